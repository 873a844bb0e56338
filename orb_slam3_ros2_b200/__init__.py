"""B200-native ORB frontend (sm_100a CUDA behind a C ABI) for the ORB-SLAM3 core driven by ORB_SLAM3_ROS2.

`orbx` binds include/orbx.h via ctypes and mirrors the reference's ORBextractor / ORBmatcher surface;
`synth` generates the deterministic synthetic inputs.  No CPU fallback exists in this package.
"""

// compat/ORBextractor.h — drop-in replacement for ORB_SLAM3/include/ORBextractor.h (reference :43-109).
//
// Same namespace, class name, constructor, operator(), accessors and the public mvImagePyramid member, so the
// reference's Frame.cc (:110-116, :299-305, :422-424, :818, :908-923), Tracking.cc (:595-601, :1587-1598) and
// KeyFrame.h (:27) compile against it unchanged.  The implementation (compat/ORBextractor.cc) forwards to the C ABI
// in include/orbx.h; every arithmetic step runs in the sm_100a kernels of liborbx.so.  There is no CPU path: if the
// CUDA library fails, operator() throws std::runtime_error (the ROS 2 node already catches std::exception around
// TrackMonocular, src/imu_mono_realsense.cpp:342).
//
// Differences a maintainer should know about (none visible to Frame/Tracking):
//   * ExtractorNode (reference :29-41) is not declared: the quadtree lives on the device (csrc/k_quadtree.cu);
//   * the protected helpers ComputePyramid / ComputeKeyPointsOctTree / DistributeOctTree / ComputeKeyPointsOld and the
//     `pattern` / `umax` members are gone for the same reason;
//   * one extra, optional constructor argument selects the CUDA device and the float-contraction mode.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <vector>
#include <opencv2/opencv.hpp>

struct orbx_extractor;  // include/orbx.h

namespace ORB_SLAM3 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // ORBextractor.cc:409-469.  `device` = CUDA ordinal (-1: current); `fpMode` 0 = STRICT (reference built with
    // -ffp-contract=off), 1 = NATIVE (the reference's own -O3 -march=native build: fmaf in the rBRIEF rotation).
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = -1,
                 int fpMode = 0);
    ~ORBextractor();
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // ORBextractor.cc:1086-1168.  Mask is ignored, as in the reference.  Returns monoIndex, or -1 for an empty image.
    int operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // Level l is a ROI at (19,19) of a freshly allocated (w+38)x(h+38) buffer on every call (reference :1176-1178),
    // so Frame copies holding older headers stay valid by reference counting.
    std::vector<cv::Mat> mvImagePyramid;

    // Set to false by callers that never read mvImagePyramid (monocular tracking) to skip the 1.3 MB/frame copy
    // back from the device.  Default true = reference behaviour.
    bool downloadPyramid = true;

    // The C-ABI handle behind this instance, for callers that move a consumer of the device-resident pyramid onto the
    // GPU too (orbx_compute_stereo_matches takes the left and right extractors' handles; Frame.cc:811-980).
    orbx_extractor* handle() const { return mHandle; }

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

private:
    orbx_extractor* mHandle;
};

}  // namespace ORB_SLAM3

#endif

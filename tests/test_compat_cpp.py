"""The C++ drop-in class (orb_slam3_ros2_b200/compat/ORBextractor.{h,cc}) — the surface Frame/Tracking compile
against (ORB_SLAM3/include/ORBextractor.h:43-109) — built against the minicv OpenCV shim and, on the GPU box, run
next to the reference's own unmodified ORBextractor.cc (oracle/_ref)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

DRIVER = os.path.join(ROOT, "tests", "cpp", "build", "compat_driver")
REF = os.path.join(ROOT, "oracle", "_ref", "liborbref_strict.so")


def _build():
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "cpp")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_compat_class_compiles_and_links():
    """Compile check of the class surface against the OpenCV-subset the reference uses; links to liborbx.so."""
    from orb_slam3_ros2_b200 import orbx
    orbx.lib()
    _build()
    assert os.path.exists(DRIVER)
    src = open(os.path.join(ROOT, "orb_slam3_ros2_b200", "compat", "ORBextractor.h")).read()
    for member in ["GetLevels", "GetScaleFactor", "GetScaleFactors", "GetInverseScaleFactors", "GetScaleSigmaSquares",
                   "GetInverseScaleSigmaSquares", "mvImagePyramid", "HARRIS_SCORE", "FAST_SCORE", "operator()"]:
        assert member in src, member


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,nframes,nfeat,lap", [(752, 480, 3, 1000, (0, 1000)), (640, 360, 2, 1250, (0, 0)),
                                                   (1241, 376, 1, 2000, (300, 600))])
def test_compat_class_matches_reference_class(tmp_path, w, h, nframes, nfeat, lap):
    from orb_slam3_ros2_b200 import synth
    if not os.path.exists(DRIVER):
        _build()
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    raw = tmp_path / "frames.raw"
    np.ascontiguousarray(synth.frames(0, nframes, w, h)).tofile(raw)
    r = subprocess.run([DRIVER, str(raw), str(w), str(h), str(nframes), str(nfeat), str(lap[0]), str(lap[1]), REF],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "compat ok" in r.stdout

// compat/ORBextractor.cc — ORB_SLAM3::ORBextractor forwarding to the C ABI (include/orbx.h -> liborbx.so).
// Replaces ORB_SLAM3/src/ORBextractor.cc in the reference's build (INTEGRATION.md shows the CMake change).
#include "ORBextractor.h"

#include <cassert>
#include <climits>
#include <stdexcept>
#include <string>

#include "orbx.h"

namespace ORB_SLAM3 {

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte record orbx writes");

namespace {
[[noreturn]] void raise(const char* what) {
    throw std::runtime_error(std::string("ORBextractor (orbx): ") + what + ": " + orbx_last_error());
}
}  // namespace

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST, int device,
                           int fpMode)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mHandle(nullptr) {
    orbx_params p;
    p.nfeatures = _nfeatures;
    p.scale_factor = _scaleFactor;
    p.nlevels = _nlevels;
    p.ini_th_fast = _iniThFAST;
    p.min_th_fast = _minThFAST;
    p.device = device;
    p.fp_mode = fpMode;
    if (orbx_create(&p, &mHandle) != ORBX_OK) raise("orbx_create");
    mvScaleFactor.resize(nlevels);
    mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    // the tables are computed once, by the library, exactly as ORBextractor.cc:414-445 does
    if (orbx_get_tables(mHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                        mvInvLevelSigma2.data(), mnFeaturesPerLevel.data()) != ORBX_OK)
        raise("orbx_get_tables");
    mvImagePyramid.resize(nlevels);
}

ORBextractor::~ORBextractor() { orbx_destroy(mHandle); }

int ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                             cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
    if (_image.empty()) return -1;  // :1090-1091

    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);  // :1094

    int n = 0, monoIndex = 0;
    // pass 1: run the pipeline, learn the keypoint count (outputs stay on the device)
    if (orbx_extract(mHandle, image.data, image.cols, image.rows, (int)image.step, vLappingArea[0], vLappingArea[1],
                     nullptr, nullptr, INT_MAX, &n, &monoIndex) != ORBX_OK)
        raise("orbx_extract");

    // pass 2: size the caller's containers as the reference does (:1108-1118) and copy straight into them
    _keypoints = std::vector<cv::KeyPoint>(n);
    if (n == 0) {
        _descriptors.release();
    } else {
        _descriptors.create(n, 32, CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        assert(descriptors.isContinuous());
        if (orbx_batch_download(mHandle, 0, reinterpret_cast<orbx_keypoint*>(_keypoints.data()), descriptors.data, n, &n,
                                &monoIndex) != ORBX_OK)
            raise("orbx_batch_download");
    }

    if (downloadPyramid) {
        for (int level = 0; level < nlevels; ++level) {
            int w = 0, h = 0;
            if (orbx_get_pyramid(mHandle, 0, level, nullptr, &w, &h) != ORBX_OK) raise("orbx_get_pyramid");
            cv::Mat temp(h + 38, w + 38, CV_8UC1);  // fresh buffer per call, like `Mat temp` at :1177
            if (orbx_get_pyramid(mHandle, 0, level, temp.data, nullptr, nullptr) != ORBX_OK) raise("orbx_get_pyramid");
            mvImagePyramid[level] = temp(cv::Rect(19, 19, w, h));
        }
    }
    return monoIndex;
}

}  // namespace ORB_SLAM3

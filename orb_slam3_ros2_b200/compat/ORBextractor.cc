// compat/ORBextractor.cc — ORB_SLAM3::ORBextractor forwarding to the C ABI (include/orbx.h -> liborbx.so).
// Replaces ORB_SLAM3/src/ORBextractor.cc in the reference's build (INTEGRATION.md shows the CMake change).
#include "ORBextractor.h"

#include <cassert>
#include <climits>
#include <cstddef>
#include <cstring>
#include <stdexcept>
#include <string>

#include "orbx.h"

namespace ORB_SLAM3 {

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte record orbx writes");
// field by field: pt (x, y), size, angle, response, octave, class_id (opencv2/core/types.hpp; ORBextractor.cc:880-890 writes them)
static_assert(offsetof(cv::KeyPoint, pt) == offsetof(orbx_keypoint, x) && offsetof(cv::KeyPoint, size) == offsetof(orbx_keypoint, size) &&
                  offsetof(cv::KeyPoint, angle) == offsetof(orbx_keypoint, angle) &&
                  offsetof(cv::KeyPoint, response) == offsetof(orbx_keypoint, response) &&
                  offsetof(cv::KeyPoint, octave) == offsetof(orbx_keypoint, octave) &&
                  offsetof(cv::KeyPoint, class_id) == offsetof(orbx_keypoint, class_id),
              "cv::KeyPoint field layout differs from orbx_keypoint");

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

    // one call, one host synchronisation: the results and (if wanted) the padded pyramid arrive in the handle's page-locked
    // staging buffer and are copied into the caller's containers here
    orbx_frame_result r;
    if (orbx_extract_frame(mHandle, image.data, image.cols, image.rows, (int)image.step, vLappingArea[0], vLappingArea[1],
                           downloadPyramid ? 1 : 0, &r) != ORBX_OK)
        raise("orbx_extract_frame");
    const int n = r.n, monoIndex = r.mono_index;

    // size the caller's containers as the reference does (:1108-1118)
    _keypoints = std::vector<cv::KeyPoint>(n);
    if (n == 0) {
        _descriptors.release();
    } else {
        std::memcpy(static_cast<void*>(_keypoints.data()), r.kps, (size_t)n * sizeof(cv::KeyPoint));
        _descriptors.create(n, 32, CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        assert(descriptors.isContinuous());
        std::memcpy(descriptors.data, r.desc, (size_t)n * 32);
    }

    if (downloadPyramid) {
        for (int level = 0; level < nlevels; ++level) {
            const orbx_frame_level& L = r.level[level];
            cv::Mat temp(L.h + 38, L.w + 38, CV_8UC1);  // fresh buffer per call, like `Mat temp` at :1177
            const uint8_t* src = r.pyramid + L.offset;
            for (int y = 0; y < L.h + 38; ++y) std::memcpy(temp.ptr(y), src + (size_t)y * L.pitch, (size_t)L.w + 38);
            mvImagePyramid[level] = temp(cv::Rect(19, 19, L.w, L.h));
        }
    }
    return monoIndex;
}

}  // namespace ORB_SLAM3

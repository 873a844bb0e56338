// ref_wrap.cpp — TEST INFRASTRUCTURE (oracle side), not product code.
//
// extern "C" façade over the reference's own, unmodified ORB_SLAM3::ORBextractor
// (ORB_SLAM3/include/ORBextractor.h:43-109, src/ORBextractor.cc) compiled against the
// minicv shim.  Built only where /root/reference exists (oracle/Makefile → oracle/_ref/).
#include <cstdint>
#include <cstring>
#include <vector>

#include "ORBextractor.h"

namespace {
// The reference keeps its per-level tables protected; a derived peek class reads them
// without touching the reference source.
struct Peek : public ORB_SLAM3::ORBextractor {
    using ORB_SLAM3::ORBextractor::ORBextractor;
    const std::vector<int>& featuresPerLevel() const { return mnFeaturesPerLevel; }
    const std::vector<int>& umaxTable() const { return umax; }
};
}  // namespace

extern "C" {

void* orbref_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    return new Peek(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}

void orbref_destroy(void* h) { delete static_cast<Peek*>(h); }

// Runs operator() (ORBextractor.cc:1086). kps: cap x 28 bytes in cv::KeyPoint layout
// (x, y, size, angle, response : f32; octave, class_id : i32); desc: cap x 32 bytes.
// Returns monoIndex (or -1 for an empty image); *n_out = number of keypoints.
int orbref_extract(void* h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, void* kps,
                   uint8_t* desc, int cap, int* n_out) {
    Peek* e = static_cast<Peek*>(h);
    cv::Mat image;
    if (img && w > 0 && hgt > 0) image = cv::Mat(hgt, w, CV_8UC1, const_cast<uint8_t*>(img), (size_t)stride);
    std::vector<cv::KeyPoint> keys;
    cv::Mat d;
    std::vector<int> lap = {lap0, lap1};
    int mono = (*e)(image, cv::Mat(), keys, d, lap);
    int n = (int)keys.size();
    if (n_out) *n_out = n;
    if (mono < 0) return mono;
    if (n > cap) n = cap;
    static_assert(sizeof(cv::KeyPoint) == 28, "KeyPoint layout");
    if (kps && n) std::memcpy(kps, keys.data(), (size_t)n * sizeof(cv::KeyPoint));
    for (int i = 0; i < n && desc; ++i) std::memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
    return mono;
}

// Padded pyramid plane of the last call: level l is a ROI at (19,19) of a (w+38)x(h+38) buffer
// (ORBextractor.cc:1176-1178).  out may be NULL to query sizes; out is (h+38) rows of (w+38) bytes.
int orbref_pyramid(void* h, int level, uint8_t* out, int* w, int* hgt) {
    Peek* e = static_cast<Peek*>(h);
    if (level < 0 || level >= (int)e->mvImagePyramid.size()) return -1;
    const cv::Mat& m = e->mvImagePyramid[level];
    if (m.empty()) return -1;
    if (w) *w = m.cols;
    if (hgt) *hgt = m.rows;
    if (out) {
        const uint8_t* base = m.data - 19 * m.step - 19;
        for (int y = 0; y < m.rows + 38; ++y) std::memcpy(out + (size_t)y * (m.cols + 38), base + (size_t)y * m.step, m.cols + 38);
    }
    return 0;
}

int orbref_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* nfeat, int* umax16) {
    Peek* e = static_cast<Peek*>(h);
    int nl = e->GetLevels();
    std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
                       d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < nl; ++i) {
        if (scale) scale[i] = a[i];
        if (inv_scale) inv_scale[i] = b[i];
        if (sigma2) sigma2[i] = c[i];
        if (inv_sigma2) inv_sigma2[i] = d[i];
        if (nfeat) nfeat[i] = e->featuresPerLevel()[i];
    }
    if (umax16) for (int i = 0; i < 16; ++i) umax16[i] = e->umaxTable()[i];
    return nl;
}

float orbref_scale_factor(void* h) { return static_cast<Peek*>(h)->GetScaleFactor(); }

}  // extern "C"

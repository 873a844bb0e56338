// minicv_capi.cpp — TEST INFRASTRUCTURE (oracle side), not product code.
// extern "C" entry points over the minicv primitives so tests can pin them to the
// cv2 4.13.0 golden vectors in tests/golden/ (tools/gen_golden.py).
#include <cstdint>
#include <cstring>
#include <vector>

#include "minicv.hpp"

extern "C" {

void minicv_resize_linear(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh) {
    cv::Mat s(sh, sw, CV_8UC1, const_cast<uint8_t*>(src)), d(dh, dw, CV_8UC1, dst);
    cv::resize(s, d, cv::Size(dw, dh), 0, 0, cv::INTER_LINEAR);
}

void minicv_border101(const uint8_t* src, int w, int h, uint8_t* dst, int b) {
    cv::Mat s(h, w, CV_8UC1, const_cast<uint8_t*>(src)), d(h + 2 * b, w + 2 * b, CV_8UC1, dst);
    cv::copyMakeBorder(s, d, b, b, b, b, cv::BORDER_REFLECT_101);
}

// returns count; out rows = (x, y, response) as int32 triplets
int minicv_fast(const uint8_t* src, int w, int h, int stride, int threshold, int nms, int32_t* out, int cap) {
    cv::Mat s(h, w, CV_8UC1, const_cast<uint8_t*>(src), (size_t)stride);
    std::vector<cv::KeyPoint> k;
    cv::FAST(s, k, threshold, nms != 0);
    int n = (int)k.size();
    for (int i = 0; i < n && i < cap; ++i) {
        out[3 * i] = (int)k[i].pt.x; out[3 * i + 1] = (int)k[i].pt.y; out[3 * i + 2] = (int)k[i].response;
    }
    return n;
}

void minicv_gauss7(const uint8_t* src, int w, int h, uint8_t* dst) {
    cv::Mat s(h, w, CV_8UC1, const_cast<uint8_t*>(src)), d(h, w, CV_8UC1, dst);
    cv::GaussianBlur(s, d, cv::Size(7, 7), 2, 2, cv::BORDER_REFLECT_101);
}

void minicv_fast_atan2(const float* y, const float* x, float* out, int n) {
    for (int i = 0; i < n; ++i) out[i] = cv::fastAtan2(y[i], x[i]);
}

// pts / out: n x 2 floats; K, P: 3 x 3 floats row-major; dist: ndist floats
void minicv_undistort_points(const float* pts, int n, const float* K, const float* dist, int ndist, const float* P, float* out) {
    cv::Mat src = cv::Mat(n, 2, CV_32F, const_cast<float*>(pts)).clone().reshape(2);
    cv::Mat Km(3, 3, CV_32F, const_cast<float*>(K)), Pm(3, 3, CV_32F, const_cast<float*>(P));
    cv::Mat Dm = ndist ? cv::Mat(ndist, 1, CV_32F, const_cast<float*>(dist)) : cv::Mat();
    cv::undistortPoints(src, src, Km, Dm, cv::Mat(), Pm);
    for (int i = 0; i < n; ++i) std::memcpy(out + 2 * i, src.ptr<float>(i), 8);
}

}  // extern "C"

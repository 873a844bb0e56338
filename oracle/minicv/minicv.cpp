// minicv.cpp — TEST INFRASTRUCTURE (oracle side), not product code.
//
// CPU restatement of the five OpenCV primitives the reference ORB path calls
// (OpenCV is an un-vendored dependency of the reference: `find_package(OpenCV 4.4)`,
// ORB_SLAM3/CMakeLists.txt:35).  Call sites in the reference:
//   resize          ORB_SLAM3/src/ORBextractor.cc:1183
//   copyMakeBorder  ORB_SLAM3/src/ORBextractor.cc:1185,1190
//   FAST            ORB_SLAM3/src/ORBextractor.cc:826,845
//   GaussianBlur    ORB_SLAM3/src/ORBextractor.cc:1133
//   fastAtan2       ORB_SLAM3/src/ORBextractor.cc:102
// Algorithms follow OpenCV 4.x (imgproc/resize.cpp, core/copy.cpp, features2d/fast.cpp +
// fast_score.cpp, imgproc/smooth.dispatch.cpp fixed-point path, core/mathfuncs_core).
// Pinned to cv2 4.13.0 by tests/golden/prim_*.npz (tools/gen_golden.py).
//
// MUST be compiled with -ffp-contract=off and without -march=native: the float code
// here emulates a prebuilt non-FMA OpenCV binary.
#include "minicv.hpp"

namespace cv {

static inline short sat_short(int v) { return (short)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); }

// ---------------------------------------------------------------------------------
// resize, INTER_LINEAR, CV_8UC1: 11-bit fixed-point coefficients, two-stage rounding.
// ---------------------------------------------------------------------------------
void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int interpolation) {
    assert(interpolation == INTER_LINEAR);
    (void)interpolation;
    Mat src = _src.getMat();
    _dst.create(dsize.height, dsize.width, CV_8UC1);
    Mat dst = _dst.getMat();
    const int sw = src.cols, sh = src.rows, dw = dsize.width, dh = dsize.height;
    const double scale_x = 1.0 / ((double)dw / sw), scale_y = 1.0 / ((double)dh / sh);

    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> alpha(2 * dw), beta(2 * dh);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cvFloor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        alpha[2 * dx] = sat_short(cvRound((1.f - fx) * 2048.f));
        alpha[2 * dx + 1] = sat_short(cvRound(fx * 2048.f));
    }
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cvFloor(fy);
        fy -= sy;
        yofs[dy] = sy;
        beta[2 * dy] = sat_short(cvRound((1.f - fy) * 2048.f));
        beta[2 * dy + 1] = sat_short(cvRound(fy * 2048.f));
    }
    auto clipy = [sh](int y) { return y < 0 ? 0 : (y >= sh ? sh - 1 : y); };
    std::vector<int> r0(dw), r1(dw);
    std::vector<uchar> out((size_t)dw * dh);
    for (int dy = 0; dy < dh; ++dy) {
        const uchar* s0 = src.ptr(clipy(yofs[dy]));
        const uchar* s1 = src.ptr(clipy(yofs[dy] + 1));
        for (int dx = 0; dx < dw; ++dx) {
            int sx = xofs[dx], sx1 = sx + 1 < sw ? sx + 1 : sw - 1;
            int a0 = alpha[2 * dx], a1 = alpha[2 * dx + 1];
            r0[dx] = s0[sx] * a0 + s0[sx1] * a1;
            r1[dx] = s1[sx] * a0 + s1[sx1] * a1;
        }
        int b0 = beta[2 * dy], b1 = beta[2 * dy + 1];
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            out[(size_t)dy * dw + dx] = (uchar)(v < 0 ? 0 : (v > 255 ? 255 : v));
        }
    }
    for (int dy = 0; dy < dh; ++dy) std::memcpy(dst.ptr(dy), &out[(size_t)dy * dw], dw);
}

// ---------------------------------------------------------------------------------
// copyMakeBorder, BORDER_REFLECT_101 (gfedcb|abcdefgh|gfedcba).  BORDER_ISOLATED only
// matters for ROI sources, whose out-of-ROI pixels it forbids reading; this shim never
// reads outside the ROI, and every reference call site passes a whole matrix or asks
// for ISOLATED (ORBextractor.cc:1185-1191), so both spellings coincide.
// ---------------------------------------------------------------------------------
static inline int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)borderType;
    Mat src = _src.getMat();
    const int w = src.cols, h = src.rows;
    Mat tmp = src.clone();  // src may be a ROI of dst (in-place use at ORBextractor.cc:1185)
    _dst.create(h + top + bottom, w + left + right, CV_8UC1);
    Mat dst = _dst.getMat();
    for (int y = 0; y < dst.rows; ++y) {
        const uchar* s = tmp.ptr(reflect101(y - top, h));
        uchar* d = dst.ptr(y);
        for (int x = 0; x < dst.cols; ++x) d[x] = s[reflect101(x - left, w)];
    }
}

// ---------------------------------------------------------------------------------
// FAST-9/16 with 3x3 non-max suppression (features2d/fast.cpp FAST_t<16>, cornerScore<16>).
// ---------------------------------------------------------------------------------
static const int kRing[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1}, {2, -2}, {1, -3},
                                 {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

// max over the 16 cyclic 9-arcs of max(min(d), min(-d)); a pixel is a corner at threshold t
// iff the result is > t, and its NMS score is result-1 (cornerScore's "-b0 - 1").
static inline int fast_arc_strength(const uchar* p, const int* ofs) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[ofs[k]];
    for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
    int best = -255;
    for (int s = 0; s < 16; ++s) {
        int mn = d[s], mx = d[s];
        for (int i = 1; i < 9; ++i) { mn = std::min(mn, d[s + i]); mx = std::max(mx, d[s + i]); }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

// Corner test at threshold t with OpenCV's antipodal-pair early reject (a 9-arc contains at
// least one point of every antipodal pair (k, k+8)), then a run-length scan over 25 samples.
static inline bool fast_is_corner(const uchar* p, const int* ofs, int t) {
    const int v = p[0], lo = v - t, hi = v + t;
    auto cls = [&](int k) { int q = p[ofs[k]]; return (q < lo ? 1 : 0) | (q > hi ? 2 : 0); };
    int d = cls(0) | cls(8);
    if (!d) return false;
    d &= cls(2) | cls(10);
    d &= cls(4) | cls(12);
    d &= cls(6) | cls(14);
    if (!d) return false;
    d &= cls(1) | cls(9);
    d &= cls(3) | cls(11);
    d &= cls(5) | cls(13);
    d &= cls(7) | cls(15);
    if (!d) return false;
    for (int bit = 1; bit <= 2; ++bit) {
        if (!(d & bit)) continue;
        int run = 0;
        for (int k = 0; k < 25; ++k) {
            if (cls(k & 15) & bit) { if (++run >= 9) return true; } else run = 0;
        }
    }
    return false;
}

void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmax) {
    Mat img = _img.getMat();
    keypoints.clear();
    threshold = std::min(std::max(threshold, 0), 255);
    const int w = img.cols, h = img.rows;
    if (w < 7 || h < 7) return;
    int ofs[16];
    for (int k = 0; k < 16; ++k) ofs[k] = kRing[k][0] + kRing[k][1] * (int)img.step;
    // arc[y*w+x] = arc strength (> threshold >= 0) for corners, 0 elsewhere; C(p) = max(arc-1, 0).
    std::vector<int> arc((size_t)w * h, 0);
    for (int y = 3; y < h - 3; ++y) {
        const uchar* row = img.ptr(y);
        for (int x = 3; x < w - 3; ++x)
            if (fast_is_corner(row + x, ofs, threshold)) arc[(size_t)y * w + x] = fast_arc_strength(row + x, ofs);
    }
    auto C = [&](int x, int y) { int a = arc[(size_t)y * w + x]; return a > 0 ? a - 1 : 0; };
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            if (arc[(size_t)y * w + x] == 0) continue;
            const int s = C(x, y);
            bool keep = true;
            if (nonmax) {
                for (int dy = -1; dy <= 1 && keep; ++dy)
                    for (int dx = -1; dx <= 1; ++dx) {
                        if (!dx && !dy) continue;
                        if (s <= C(x + dx, y + dy)) { keep = false; break; }
                    }
            }
            if (keep) keypoints.push_back(KeyPoint((float)x, (float)y, 7.f, -1, (float)s));
        }
}

// ---------------------------------------------------------------------------------
// GaussianBlur 7x7 sigma=2, CV_8UC1, BORDER_REFLECT_101: OpenCV's fixed-point path
// (8 fractional bits per pass, one final rounding).
// ---------------------------------------------------------------------------------
void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY, int borderType) {
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2.0 && (sigmaY == 2.0 || sigmaY == 0.0));
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    static const unsigned k[7] = {18, 34, 48, 56, 48, 34, 18};  // getGaussianKernel(7,2) in 0.8 fixed point
    Mat src = _src.getMat();
    const int w = src.cols, h = src.rows;
    std::vector<uint16_t> H((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uchar* s = src.ptr(y);
        for (int x = 0; x < w; ++x) {
            unsigned acc = 0;
            for (int i = 0; i < 7; ++i) acc += k[i] * s[reflect101(x + i - 3, w)];
            H[(size_t)y * w + x] = (uint16_t)acc;
        }
    }
    std::vector<uchar> out((size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 0;
            for (int j = 0; j < 7; ++j) acc += k[j] * (uint32_t)H[(size_t)reflect101(y + j - 3, h) * w + x];
            out[(size_t)y * w + x] = (uchar)((acc + 32768u) >> 16);
        }
    _dst.create(h, w, CV_8UC1);
    Mat dst = _dst.getMat();
    for (int y = 0; y < h; ++y) std::memcpy(dst.ptr(y), &out[(size_t)y * w], w);
}

// ---------------------------------------------------------------------------------
// fastAtan2 (degrees, [0,360)): scalar, non-FMA Horner form of OpenCV's atanf polynomial.
// ---------------------------------------------------------------------------------
float fastAtan2(float y, float x) {
    const float s = (float)(180.0 / CV_PI);
    const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s;
    const float p5 = 0.1555786518463281f * s, p7 = -0.04432655554792128f * s;
    float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ---------------------------------------------------------------------------------
// undistortPoints (calib3d, cvUndistortPointsInternal): called by Frame::UndistortKeyPoints / ComputeImageBounds
// (ORB_SLAM3/src/Frame.cc:766, :795) with R empty and P = mK.  Camera matrix, distortion coefficients and P are
// converted to double; the coefficient vector is (k1, k2, p1, p2[, k3[, k4, k5, k6[, s1..s4[, tauX, tauY]]]]) padded
// with zeros (tilt = identity for the 4/5-coefficient vectors of the path).  Default criteria = 5 iterations, no
// epsilon test.  Scalar double arithmetic, no FMA (baseline SSE2 build of the cv2 wheel; pinned by
// tests/golden/prim_undistort.npz).
// ---------------------------------------------------------------------------------
static double matElem(const Mat& m, int i) { return m.depth() == CV_64F ? m.at<double>(i) : (double)m.at<float>(i); }
void undistortPoints(InputArray _src, OutputArray _dst, InputArray _K, InputArray _D, InputArray _R, InputArray _P) {
    const Mat src = _src.getMat(), K = _K.getMat(), D = _D.getMat(), P = _P.getMat();
    assert(_R.empty());
    (void)_R;
    assert(src.depth() == CV_32F && src.channels() == 2 && src.cols == 1);
    const int n = src.rows;
    double k[14] = {0};
    const int nk = D.empty() ? 0 : D.rows * D.cols;
    for (int i = 0; i < nk && i < 14; ++i) k[i] = matElem(D, i);
    const double fx = matElem(K.row(0), 0), fy = matElem(K.row(1), 1), cx = matElem(K.row(0), 2), cy = matElem(K.row(1), 2);
    const double ifx = 1. / fx, ify = 1. / fy;
    double RR[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    if (!P.empty())  // RR = PP * RR with RR = I: exact
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) RR[r][c] = matElem(P.row(r), c);
    std::vector<float> out((size_t)n * 2);
    for (int i = 0; i < n; ++i) {
        const float* s = src.ptr<float>(i);
        double x = s[0], y = s[1];
        const double u = x, v = y;
        x = (x - cx) * ifx;
        y = (y - cy) * ify;
        if (nk) {
            const double x0 = x, y0 = y;
            for (int j = 0; j < 5; ++j) {
                const double r2 = x * x + y * y;
                const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
                if (icdist < 0) {  // the point is outside the valid area of the model: left as the normalised input
                    x = (u - cx) * ifx;
                    y = (v - cy) * ify;
                    break;
                }
                const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
                const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
                x = (x0 - deltaX) * icdist;
                y = (y0 - deltaY) * icdist;
            }
        }
        const double xx = RR[0][0] * x + RR[0][1] * y + RR[0][2];
        const double yy = RR[1][0] * x + RR[1][1] * y + RR[1][2];
        const double ww = 1. / (RR[2][0] * x + RR[2][1] * y + RR[2][2]);
        out[2 * i] = (float)(xx * ww);
        out[2 * i + 1] = (float)(yy * ww);
    }
    if (_dst.getMat().data != src.data || _dst.getMat().rows != n) {
        _dst.create(n, 2, CV_32F);
        _dst.getMatRef() = _dst.getMat().reshape(2);
    }
    Mat& dst = _dst.getMatRef();
    for (int i = 0; i < n; ++i) std::memcpy(dst.ptr<float>(i), &out[2 * i], 8);
}

// ---------------------------------------------------------------------------------
// KeyPointsFilter::retainBest — only reachable from the reference's dead
// ComputeKeyPointsOld (ORBextractor.cc:1049,1067; its call is commented out at :1101).
// ---------------------------------------------------------------------------------
void KeyPointsFilter::retainBest(std::vector<KeyPoint>& kps, int n) {
    if (n < 0 || (int)kps.size() <= n) return;
    if (n == 0) { kps.clear(); return; }
    auto gt = [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; };
    std::nth_element(kps.begin(), kps.begin() + n - 1, kps.end(), gt);
    float amb = kps[n - 1].response;
    auto it = std::partition(kps.begin() + n, kps.end(), [amb](const KeyPoint& k) { return k.response >= amb; });
    kps.resize(it - kps.begin());
}

}  // namespace cv

// k_pyramid.cu — K-a: the scale pyramid with its reflected border.
//
// Replaces ORBextractor::ComputePyramid (ORB_SLAM3/src/ORBextractor.cc:1170-1195): per level
// cv::resize(level l-1 -> l, INTER_LINEAR) followed by copyMakeBorder(19, BORDER_REFLECT_101) into the padded
// plane the level lives in.  cv::resize's 8-bit linear path is 11-bit fixed point with two roundings
// (SURVEY.md App. A1); coefficient tables are generated on the host exactly as OpenCV does (double -> float ->
// floor) and uploaded once per plan.
//
// One launch per level (level l reads level l-1: a true dependency), all frames of the batch in one grid.
// Every thread produces 4 horizontally adjacent bytes of the PADDED plane and stores them as one 32-bit word:
// a border pixel recomputes the interior value it mirrors, so interior and border come from one pass and the
// plane is written exactly once (algorithmic bytes: (w+38)(h+38) per level).
#include "orbx_internal.cuh"

namespace orbx {
namespace {

__device__ __forceinline__ int reflect101(int i, int n) {
    // 19 <= n always holds for levels that carry FAST cells; the loop form keeps tiny levels correct too.
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

// level 0: copy of the caller's image + border (ORBextractor.cc:1188-1191)
__global__ void __launch_bounds__(256) pyrLevel0Kernel(const uint8_t* __restrict__ img, int stride, size_t frameStride,
                                                       uint8_t* __restrict__ pyr, unsigned frameBytes, int w, int h,
                                                       int pitch) {
    const int xp = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int yp = blockIdx.y * blockDim.y + threadIdx.y;
    const int pw = w + 2 * kEdge, ph = h + 2 * kEdge;
    if (xp >= pw || yp >= ph) return;
    const uint8_t* src = img + (size_t)blockIdx.z * frameStride + (size_t)reflect101(yp - kEdge, h) * stride;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int x = min(xp + k, pw - 1);
        word |= (uint32_t)__ldg(src + reflect101(x - kEdge, w)) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(pyr + (size_t)blockIdx.z * frameBytes + (size_t)yp * pitch + xp) = word;
}

// level l >= 1 from level l-1
__global__ void __launch_bounds__(256)
    pyrResizeKernel(uint8_t* __restrict__ pyr, unsigned frameBytes, unsigned srcOff, int sw, int sh, int spitch,
                    unsigned dstOff, int dw, int dh, int dpitch, const short4* __restrict__ tabX,
                    const short4* __restrict__ tabY) {
    const int xp = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int yp = blockIdx.y * blockDim.y + threadIdx.y;
    const int pw = dw + 2 * kEdge, ph = dh + 2 * kEdge;
    if (xp >= pw || yp >= ph) return;
    uint8_t* frame = pyr + (size_t)blockIdx.z * frameBytes;
    const uint8_t* src = frame + srcOff + (size_t)kEdge * spitch + kEdge;  // ROI origin of level l-1
    const int dy = reflect101(yp - kEdge, dh);
    const short4 ty = __ldg(tabY + dy);  // (sy, b0, b1, -)
    const int sy0 = min(max((int)ty.x, 0), sh - 1), sy1 = min(max((int)ty.x + 1, 0), sh - 1);
    const uint8_t* r0p = src + (size_t)sy0 * spitch;
    const uint8_t* r1p = src + (size_t)sy1 * spitch;
    const int b0 = ty.y, b1 = ty.z;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(xp + k, pw - 1);
        const int dx = reflect101(x - kEdge, dw);
        const short4 tx = __ldg(tabX + dx);  // (sx, a0, a1, -)
        const int sx = tx.x, sx1 = min(sx + 1, sw - 1);
        const int a0 = tx.y, a1 = tx.z;
        const int r0 = (int)r0p[sx] * a0 + (int)r0p[sx1] * a1;
        const int r1 = (int)r1p[sx] * a0 + (int)r1p[sx1] * a1;
        int v = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2;
        v = min(max(v, 0), 255);
        word |= (uint32_t)v << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(frame + dstOff + (size_t)yp * dpitch + xp) = word;
}

}  // namespace

void launchPyramid(const FramePlan& P, const Buffers& B, const uint8_t* dImgs, int stride, size_t frameStride, int F,
                   cudaStream_t s) {
    const dim3 block(64, 4);
    {
        const LevelPlan& L = P.lv[0];
        const int pw = L.w + 2 * kEdge, ph = L.h + 2 * kEdge;
        dim3 grid((pw + 255) / 256, (ph + 3) / 4, F);
        pyrLevel0Kernel<<<grid, block, 0, s>>>(dImgs, stride, frameStride, B.pyr, P.frameBytes, L.w, L.h, L.pitch);
        countLaunch();
    }
    for (int l = 1; l < P.nlevels; ++l) {
        const LevelPlan& S = P.lv[l - 1];
        const LevelPlan& D = P.lv[l];
        const int pw = D.w + 2 * kEdge, ph = D.h + 2 * kEdge;
        dim3 grid((pw + 255) / 256, (ph + 3) / 4, F);
        pyrResizeKernel<<<grid, block, 0, s>>>(B.pyr, P.frameBytes, S.planeOff, S.w, S.h, S.pitch, D.planeOff, D.w, D.h,
                                               D.pitch, reinterpret_cast<const short4*>(B.tab) + D.tabX,
                                               reinterpret_cast<const short4*>(B.tab) + D.tabY);
        countLaunch();
    }
}

}  // namespace orbx

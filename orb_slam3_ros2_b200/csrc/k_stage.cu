// k_stage.cu — input staging fused ahead of the pyramid.
//
// Replaces, for the frame that is about to be extracted, System::TrackMonocular's cv::resize(im, resized, newImSize)
// (ORB_SLAM3/src/System.cc:420-425; stereo :269-270, RGB-D :345-347) followed by Tracking::GrabImage*'s
// cvtColor(RGB|BGR|RGBA|BGRA -> GRAY) (src/Tracking.cc:1569-1582) — in that order, as the reference applies them: the
// resize runs on the interleaved colour image, channel by channel, and the grey conversion on its result.
//   * resize: cv::resize INTER_LINEAR, 8-bit: the same 11-bit fixed-point two-rounding form as the pyramid
//     (an exact 2x reduction degenerates to OpenCV's area fast path bit for bit);
//   * grey: OpenCV's 8-bit RGB2Gray, (R*9798 + G*19235 + B*3735 + 2^14) >> 15 (pinned against cv2 4.13, tests/golden).
// One thread per output pixel: the raw frame is read once and the grey frame written once (HBM-bound, 4-5 B/pixel).
#include "orbx_internal.cuh"

namespace orbx {
namespace {

__global__ void __launch_bounds__(256) stageKernel(const uint8_t* __restrict__ src, int sw, int sh, int sstride, int ch, int rgb,
                                                   uint8_t* __restrict__ dst, int dw, int dh, const short4* __restrict__ tabX,
                                                   const short4* __restrict__ tabY) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= dw || y >= dh) return;
    uint32_t c[3] = {0, 0, 0};
    const int nc = ch >= 3 ? 3 : 1;
    if (tabX) {
        const short4 tx = __ldg(tabX + x), ty = __ldg(tabY + y);
        const int x0 = tx.x, x1 = min(tx.x + 1, sw - 1);
        const int y0 = min(max((int)ty.x, 0), sh - 1), y1 = min(max((int)ty.x + 1, 0), sh - 1);
        const uint8_t* r0 = src + (size_t)y0 * sstride;
        const uint8_t* r1 = src + (size_t)y1 * sstride;
        const uint32_t B0 = (uint32_t)ty.y << 16, B1 = (uint32_t)ty.z << 16;
        for (int k = 0; k < nc; ++k) {
            const uint32_t h0 = (uint32_t)((int)r0[x0 * ch + k] * tx.y + (int)r0[x1 * ch + k] * tx.z) >> 4;
            const uint32_t h1 = (uint32_t)((int)r1[x0 * ch + k] * tx.y + (int)r1[x1 * ch + k] * tx.z) >> 4;
            c[k] = (__umulhi(B0, h0) + __umulhi(B1, h1) + 2u) >> 2;
        }
    } else {
        const uint8_t* p = src + (size_t)y * sstride + (size_t)x * ch;
        for (int k = 0; k < nc; ++k) c[k] = p[k];
    }
    uint32_t g = c[0];
    if (nc == 3) {
        const uint32_t R = rgb ? c[0] : c[2], B = rgb ? c[2] : c[0];
        g = (R * 9798u + c[1] * 19235u + B * 3735u + (1u << 14)) >> 15;
    }
    dst[(size_t)y * dw + x] = (uint8_t)g;
}

}  // namespace

void launchStage(const uint8_t* dSrc, int sw, int sh, int sstride, int ch, int rgb, uint8_t* dDst, int dw, int dh,
                 const int16_t* tabX, const int16_t* tabY, cudaStream_t s) {
    stageKernel<<<dim3((dw + 255) / 256, dh), 256, 0, s>>>(dSrc, sw, sh, sstride, ch, rgb, dDst, dw, dh,
                                                         reinterpret_cast<const short4*>(tabX), reinterpret_cast<const short4*>(tabY));
    countLaunch();
}

}  // namespace orbx

// k_blur.cu — K-e: GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) of every pyramid level.
//
// Replaces `GaussianBlur(workingMat, workingMat, Size(7,7), 2, 2, BORDER_REFLECT_101)` on a clone of the
// un-padded level (ORB_SLAM3/src/ORBextractor.cc:1132-1133).  OpenCV's 8-bit path is separable fixed point
// with kernel [18,34,48,56,48,34,18]/256 per pass and ONE final rounding (V + 32768) >> 16.
//
// The clone is reflected at the true image edge; the padded plane's 19-px border already holds exactly that
// REFLECT_101 extension, so the kernel reads the padded plane and needs no boundary logic.  The level ROI starts
// at byte 19 of a 128-B aligned row; minus the 3-px apron that is byte 16, so the staged tile rows are 16-B
// aligned and are fetched with 128-bit loads.
//
// One launch covers all levels and frames: blockIdx.x enumerates (level, tile), blockIdx.y the frame.
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kTileW = 128, kTileH = 32, kInRows = kTileH + 6, kInPitch = 144;

__global__ void __launch_bounds__(256) blurKernel(const __grid_constant__ FramePlan P, const uint8_t* __restrict__ pyr,
                                                  uint8_t* __restrict__ blur) {
    __shared__ __align__(16) uint8_t sIn[kInRows * kInPitch];
    __shared__ __align__(16) uint16_t sH[kInRows * kTileW];

    int t = blockIdx.x, l = 0;
    for (; l < P.nlevels; ++l) {
        const int nt = ((P.lv[l].w + kTileW - 1) / kTileW) * ((P.lv[l].h + kTileH - 1) / kTileH);
        if (t < nt) break;
        t -= nt;
    }
    if (l >= P.nlevels) return;
    const LevelPlan& L = P.lv[l];
    const int tilesX = (L.w + kTileW - 1) / kTileW;
    const int x0 = (t % tilesX) * kTileW, y0 = (t / tilesX) * kTileH;
    const uint8_t* plane = pyr + (size_t)blockIdx.y * P.frameBytes + L.planeOff;
    const int ph = L.h + 2 * kEdge;
    const int tid = threadIdx.x;

    // stage (kTileH+6) rows x 144 bytes, padded coordinates (x0+16, y0+16)
    for (int i = tid; i < kInRows * (kInPitch / 16); i += 256) {
        const int r = i / (kInPitch / 16), c = i % (kInPitch / 16);
        const int yy = min(y0 + 16 + r, ph - 1);
        int xx = x0 + 16 + c * 16;
        if (xx + 16 > L.pitch) xx = L.pitch - 16;  // columns past the row's pitch are never used for x < w
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(plane + (size_t)yy * L.pitch + xx));
        *reinterpret_cast<uint4*>(sIn + r * kInPitch + c * 16) = v;
    }
    __syncthreads();

    // horizontal pass: 4 outputs per item from 12 staged bytes
    for (int i = tid; i < kInRows * (kTileW / 4); i += 256) {
        const int r = i / (kTileW / 4), g = i % (kTileW / 4);
        const uint32_t* p = reinterpret_cast<const uint32_t*>(sIn + r * kInPitch + g * 4);
        const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
        int b[12];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            b[k] = (w0 >> (8 * k)) & 255;
            b[4 + k] = (w1 >> (8 * k)) & 255;
            b[8 + k] = (w2 >> (8 * k)) & 255;
        }
        uint16_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            o[k] = (uint16_t)(18 * (b[k] + b[k + 6]) + 34 * (b[k + 1] + b[k + 5]) + 48 * (b[k + 2] + b[k + 4]) +
                              56 * b[k + 3]);
        uint2 pk;
        pk.x = o[0] | ((uint32_t)o[1] << 16);
        pk.y = o[2] | ((uint32_t)o[3] << 16);
        *reinterpret_cast<uint2*>(sH + r * kTileW + g * 4) = pk;
    }
    __syncthreads();

    // vertical pass: thread = 4 columns x 4 rows, sliding 10-row window
    const int tx = tid & 31, ty = tid >> 5;
    uint32_t h[10][4];
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint2 v = *reinterpret_cast<const uint2*>(sH + (ty * 4 + r) * kTileW + tx * 4);
        h[r][0] = v.x & 0xffff; h[r][1] = v.x >> 16; h[r][2] = v.y & 0xffff; h[r][3] = v.y >> 16;
    }
    uint8_t* out = blur + (size_t)blockIdx.y * P.blurBytes + L.blurOff;
    const int x = x0 + tx * 4;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int y = y0 + ty * 4 + r;
        uint32_t word = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t acc = 18u * (h[r][k] + h[r + 6][k]) + 34u * (h[r + 1][k] + h[r + 5][k]) +
                                 48u * (h[r + 2][k] + h[r + 4][k]) + 56u * h[r + 3][k];
            word |= ((acc + 32768u) >> 16) << (8 * k);
        }
        if (y < L.h && x < L.bpitch) *reinterpret_cast<uint32_t*>(out + (size_t)y * L.bpitch + x) = word;
    }
}

}  // namespace

void launchBlur(const FramePlan& P, const Buffers& B, int F, cudaStream_t s) {
    int tiles = 0;
    for (int l = 0; l < P.nlevels; ++l)
        tiles += ((P.lv[l].w + kTileW - 1) / kTileW) * ((P.lv[l].h + kTileH - 1) / kTileH);
    blurKernel<<<dim3(tiles, F), 256, 0, s>>>(P, B.pyr, B.blur);
    countLaunch();
}

}  // namespace orbx

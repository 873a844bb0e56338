// k_pyramid.cu — K-a: the scale pyramid with its reflected border.
//
// Replaces ORBextractor::ComputePyramid (ORB_SLAM3/src/ORBextractor.cc:1170-1195): per level
// cv::resize(level l-1 -> l, INTER_LINEAR) followed by copyMakeBorder(19, BORDER_REFLECT_101) into the padded
// plane the level lives in.  cv::resize's 8-bit linear path is 11-bit fixed point with two roundings
// (SURVEY.md App. A1); coefficient tables are generated on the host exactly as OpenCV does (double -> float ->
// floor) and uploaded once per plan.
//
// One launch per level (level l reads level l-1: a true dependency), all frames of the batch in one grid.  A CTA
// owns a 128 x 32 tile of the PADDED destination plane; a border pixel is the interior pixel it mirrors, so the
// same code produces interior and border and the plane is written exactly once, with 32-bit stores.
// Like OpenCV the work is separable: the source rows the tile needs are staged in shared memory (128-bit loads),
// the horizontal pass turns each of them into one row of 16-bit (a0*p0 + a1*p1) >> 4 values for the tile's 128
// columns (a thread keeps the coefficients of its 4 columns in registers), and the vertical pass blends two such
// rows per output row.  ~18 instructions per output pixel instead of ~80 for the direct per-pixel form.
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kTW = 128, kTH = 32;

__device__ __forceinline__ int reflect101(int i, int n) {
    // 19 <= n always holds for levels that carry FAST cells; the loop form keeps tiny levels correct too.
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

// level 0, generic: copy of the caller's image + border (ORBextractor.cc:1188-1191), any stride / alignment
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

// level 0, rows 16-B aligned (base, stride and frame stride multiples of 16): a thread writes 16 padded bytes with one
// 128-bit store.  Padded x = 16k maps to source x = 16(k-1) - 3, i.e. the aligned source chunk k-1 shifted by three
// bytes: one 128-bit load plus the last word of the chunk before it.  Chunks that touch the border mirror per byte.
__global__ void __launch_bounds__(256) pyrLevel0AlignedKernel(const uint8_t* __restrict__ img, int stride, size_t frameStride,
                                                              uint8_t* __restrict__ pyr, unsigned frameBytes, int w, int h,
                                                              int pitch) {
    const int k = blockIdx.x * 64 + threadIdx.x;  // 16-byte chunk of the padded row
    const int yp = blockIdx.y * 4 + threadIdx.y;
    const int pw = w + 2 * kEdge, ph = h + 2 * kEdge;
    if (k * 16 >= pw || yp >= ph) return;
    const uint8_t* src = img + (size_t)blockIdx.z * frameStride + (size_t)reflect101(yp - kEdge, h) * stride;
    const int x0 = k * 16 - kEdge;  // source x of the chunk's first byte
    uint4 out;
    if (x0 >= 1 && x0 + 19 <= w) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(src + x0 + 3));      // source bytes x0+3 .. x0+18
        const uint32_t p = __ldg(reinterpret_cast<const uint32_t*>(src + x0 - 1));  // source bytes x0-1 .. x0+2
        out.x = __byte_perm(p, a.x, 0x4321);
        out.y = __byte_perm(a.x, a.y, 0x4321);
        out.z = __byte_perm(a.y, a.z, 0x4321);
        out.w = __byte_perm(a.z, a.w, 0x4321);
    } else {
        uint32_t wd[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t word = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int x = min(k * 16 + 4 * j + b, pw - 1);
                word |= (uint32_t)__ldg(src + reflect101(x - kEdge, w)) << (8 * b);
            }
            wd[j] = word;
        }
        out = make_uint4(wd[0], wd[1], wd[2], wd[3]);
    }
    *reinterpret_cast<uint4*>(pyr + (size_t)blockIdx.z * frameBytes + (size_t)yp * pitch + k * 16) = out;
}

// level l >= 1 from level l-1, tiled and separable
__global__ void __launch_bounds__(256)
    pyrResizeKernel(uint8_t* __restrict__ pyr, unsigned frameBytes, unsigned srcOff, int sw, int sh, int spitch,
                    unsigned dstOff, int dw, int dh, int dpitch, const short4* __restrict__ tabX,
                    const short4* __restrict__ tabY, int maxRows, int SP) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ short4 sRow[kTH];  // per tile row: first / second source row (relative to S0), b0, b1
    __shared__ int sS0, sNS;
    uint8_t* sSrc = smem;                                               // [maxRows][SP] staged source rows
    uint16_t* sH = reinterpret_cast<uint16_t*>(smem + (size_t)maxRows * SP);  // [maxRows][kTW] horizontal pass

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int X0 = blockIdx.x * kTW, Y0 = blockIdx.y * kTH;
    const int pw = dw + 2 * kEdge, ph = dh + 2 * kEdge;
    uint8_t* frame = pyr + (size_t)blockIdx.z * frameBytes;

    // ---- rows of the tile -> source rows ------------------------------------------------------------------------
    if (warp == 0) {
        const int Y = min(Y0 + lane, ph - 1);
        const short4 ty = __ldg(tabY + reflect101(Y - kEdge, dh));  // (sy, b0, b1, -)
        const int sy0 = min(max((int)ty.x, 0), sh - 1), sy1 = min(max((int)ty.x + 1, 0), sh - 1);
        int lo = sy0, hi = sy1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        sRow[lane] = make_short4((short)(sy0 - lo), (short)(sy1 - lo), ty.y, ty.z);
        if (lane == 0) { sS0 = lo; sNS = hi - lo + 1; }
    }
    // ---- the 4 columns of this lane (the same in every warp) -> source columns and coefficients -----------------
    int sx[4], sx1[4], a0[4], a1[4];
    int cLo = 1 << 30, cHi = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int X = min(X0 + 4 * lane + j, pw - 1);
        const short4 tx = __ldg(tabX + reflect101(X - kEdge, dw));  // (sx, a0, a1, -)
        sx[j] = tx.x; sx1[j] = min(tx.x + 1, sw - 1);
        a0[j] = tx.y; a1[j] = tx.z;
        cLo = min(cLo, sx[j]); cHi = max(cHi, sx1[j]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cLo = min(cLo, __shfl_xor_sync(0xffffffffu, cLo, o));
        cHi = max(cHi, __shfl_xor_sync(0xffffffffu, cHi, o));
    }
    const int A0 = (kEdge + cLo) & ~15;                      // plane column of the first staged byte (16-B aligned)
    const int nC16 = ((kEdge + cHi) - A0) / 16 + 1;          // 16-B chunks per staged row
#pragma unroll
    for (int j = 0; j < 4; ++j) { sx[j] += kEdge - A0; sx1[j] += kEdge - A0; }
    __syncthreads();
    const int S0 = sS0, nS = sNS;

    // ---- stage the source rows ---------------------------------------------------------------------------------------
    {
        const uint8_t* src = frame + srcOff + (size_t)(kEdge + S0) * spitch + A0;
        const int total = nS * nC16;
        for (int i = tid; i < total; i += 256) {
            const int r = i / nC16, c = i - r * nC16;
            *reinterpret_cast<uint4*>(sSrc + r * SP + c * 16) = __ldcg(reinterpret_cast<const uint4*>(src + (size_t)r * spitch) + c);
        }
    }
    __syncthreads();

    // ---- horizontal pass: H[s][x] = (p0*a0 + p1*a1) >> 4  (<= 32640, 16 bits) ------------------------------------
    for (int s = warp; s < nS; s += 8) {
        const uint8_t* row = sSrc + s * SP;
        uint32_t hv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) hv[j] = (uint32_t)((int)row[sx[j]] * a0[j] + (int)row[sx1[j]] * a1[j]) >> 4;
        uint2 pk;
        pk.x = hv[0] | (hv[1] << 16);
        pk.y = hv[2] | (hv[3] << 16);
        *reinterpret_cast<uint2*>(sH + s * kTW + 4 * lane) = pk;
    }
    __syncthreads();

    // ---- vertical pass + store ------------------------------------------------------------------------------------
    const int X = X0 + 4 * lane;
#pragma unroll
    for (int k = 0; k < kTH / 8; ++k) {
        const int ry = warp * (kTH / 8) + k, Y = Y0 + ry;
        if (Y >= ph || X >= pw) continue;
        const short4 ri = sRow[ry];
        const uint2 u = *reinterpret_cast<const uint2*>(sH + ri.x * kTW + 4 * lane);
        const uint2 v = *reinterpret_cast<const uint2*>(sH + ri.y * kTW + 4 * lane);
        const int b0 = ri.z, b1 = ri.w;
        const int h0[4] = {(int)(u.x & 0xffff), (int)(u.x >> 16), (int)(u.y & 0xffff), (int)(u.y >> 16)};
        const int h1[4] = {(int)(v.x & 0xffff), (int)(v.x >> 16), (int)(v.y & 0xffff), (int)(v.y >> 16)};
        uint32_t word = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            // ((b0*r0>>4)>>16) + ((b1*r1>>4)>>16) + 2 >> 2 is at most 255 (b0 + b1 = 2048): no saturation needed
            const int val = (((b0 * h0[j]) >> 16) + ((b1 * h1[j]) >> 16) + 2) >> 2;
            word |= (uint32_t)val << (8 * j);
        }
        *reinterpret_cast<uint32_t*>(frame + dstOff + (size_t)Y * dpitch + X) = word;
    }
}

}  // namespace

void launchPyramid(const FramePlan& P, const Buffers& B, const uint8_t* dImgs, int stride, size_t frameStride, int F,
                   cudaStream_t s) {
    {
        const LevelPlan& L = P.lv[0];
        const int pw = L.w + 2 * kEdge, ph = L.h + 2 * kEdge;
        const dim3 block(64, 4);
        const bool aligned = ((reinterpret_cast<uintptr_t>(dImgs) | (uintptr_t)stride | (uintptr_t)frameStride) & 15) == 0;
        if (aligned) {
            dim3 grid((pw + 1023) / 1024, (ph + 3) / 4, F);
            pyrLevel0AlignedKernel<<<grid, block, 0, s>>>(dImgs, stride, frameStride, B.pyr, P.frameBytes, L.w, L.h, L.pitch);
        } else {
            dim3 grid((pw + 255) / 256, (ph + 3) / 4, F);
            pyrLevel0Kernel<<<grid, block, 0, s>>>(dImgs, stride, frameStride, B.pyr, P.frameBytes, L.w, L.h, L.pitch);
        }
        countLaunch();
    }
    for (int l = 1; l < P.nlevels; ++l) {
        const LevelPlan& S = P.lv[l - 1];
        const LevelPlan& D = P.lv[l];
        const int pw = D.w + 2 * kEdge, ph = D.h + 2 * kEdge;
        // shared-memory geometry from the level's scale: source rows / bytes a tile can need (+ slack for the 16-B alignment)
        const int maxRows = (int)((double)kTH * S.h / D.h) + 4;
        const int SP = ((int)((double)kTW * S.w / D.w) + 4 + 15 + 16 + 15) / 16 * 16;
        const size_t smem = (size_t)maxRows * (SP + kTW * 2);
        if (smem > 48 * 1024) cudaFuncSetAttribute(pyrResizeKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        dim3 grid((pw + kTW - 1) / kTW, (ph + kTH - 1) / kTH, F);
        pyrResizeKernel<<<grid, 256, smem, s>>>(B.pyr, P.frameBytes, S.planeOff, S.w, S.h, S.pitch, D.planeOff, D.w, D.h,
                                                D.pitch, reinterpret_cast<const short4*>(B.tab) + D.tabX,
                                                reinterpret_cast<const short4*>(B.tab) + D.tabY, maxRows, SP);
        countLaunch();
    }
}

}  // namespace orbx

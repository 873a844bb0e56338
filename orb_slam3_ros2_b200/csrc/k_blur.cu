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
//
// Arithmetic layout (v2): both passes are byte / half-word DOT PRODUCTS, which sm_100a executes as single instructions:
//   horizontal  H[x] = sum_i k[i] * I[x+i-3]  = dp4a(bytes x-3..x, (18,34,48,56)) + dp4a(bytes x+1..x+4, (48,34,18,0));
//               the two byte windows of each pixel are funnel shifts of three aligned words shared by 4 pixels;
//   vertical    H rows are stored as VERTICAL pairs (rows 2m, 2m+1 of the staged tile in one 32-bit word), so the 7-tap
//               column sum is four dp2a (16-bit x 8-bit, 2-way) on consecutive pairs; even and odd output rows differ
//               only in the coefficient bytes.  The rounding constant rides in the first accumulator.
// H <= 65280 fits 16 bits and V <= 65280 * 256 + 32768 fits 32 bits, exactly OpenCV's fixed-point path.
//
// v3 (from the per-line ncu profile of v2: the level walk, the tile division and the staging loop were 35 % of the
// kernel's instructions, the arithmetic a minority):
//   * the (level, tile row, tile column) of a CTA comes from a table built with the plan (one load instead of a walk over
//     the levels and an integer division);
//   * the (32+6) x 144-byte input tile is staged by TMA: ONE cp.async.bulk.tensor.3d per CTA against the level's tensor map
//     (box 144 x 38 x 1; the tile's first byte is padded column x0 + 16, always 16-byte aligned), completion on an
//     mbarrier; rows / columns past the plane are zero-filled by the copy engine and only feed outputs that are not stored;
//   * the final ">> 16" and the byte packing are one PRMT pair (the result is byte 2 of the accumulator).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "orbx_internal.cuh"
#include "tma.cuh"

namespace orbx {
namespace {

constexpr int kTileW = 128, kTileH = 32, kInRows = kTileH + 6, kInPitch = 144, kPairs = kInRows / 2;

// v2: staging with ordinary loads, level walk per CTA (kept for A/B runs: ORBX_FAST_TMA=0)
__global__ void __launch_bounds__(256) blurKernelLd(const __grid_constant__ FramePlan P, const uint8_t* __restrict__ pyr,
                                                    uint8_t* __restrict__ blur) {
    __shared__ __align__(16) uint8_t sIn[kInRows * kInPitch];
    __shared__ __align__(16) uint32_t sHp[kPairs * kTileW];  // (H[2m][x] | H[2m+1][x] << 16)

    int l = 0;
    while (l + 1 < P.nlevels && (int)blockIdx.x >= P.lv[l + 1].blurTile0) ++l;
    const LevelPlan& L = P.lv[l];
    const int t = blockIdx.x - L.blurTile0;
    const int ty = t / L.blurTilesX;
    const int x0 = (t - ty * L.blurTilesX) * kTileW, y0 = ty * kTileH;
    const uint8_t* plane = pyr + (size_t)blockIdx.y * P.frameBytes + L.planeOff;
    const int ph = L.h + 2 * kEdge;
    const int tid = threadIdx.x;

    // stage (kTileH+6) rows x 144 bytes, padded coordinates (x0+16, y0+16): staged byte 0 of a row = level column x0-3
    for (int i = tid; i < kInRows * (kInPitch / 16); i += 256) {
        const int r = i / (kInPitch / 16), c = i % (kInPitch / 16);
        const int yy = min(y0 + 16 + r, ph - 1);
        int xx = x0 + 16 + c * 16;
        if (xx + 16 > L.pitch) xx = L.pitch - 16;  // columns past the row's pitch are never used for x < w
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(plane + (size_t)yy * L.pitch + xx));
        *reinterpret_cast<uint4*>(sIn + r * kInPitch + c * 16) = v;
    }
    __syncthreads();

    // horizontal pass: item = (row pair m, 4 columns g)
    constexpr uint32_t K1 = 0x38302212u, K2 = 0x00122230u;  // (18,34,48,56), (48,34,18,0)
    for (int i = tid; i < kPairs * (kTileW / 4); i += 256) {
        const int m = i >> 5, g = i & 31;
        uint32_t h[2][4];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(sIn + (2 * m + rr) * kInPitch) + g;
            const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t a = j ? __funnelshift_r(w0, w1, 8 * j) : w0;
                const uint32_t b = j ? __funnelshift_r(w1, w2, 8 * j) : w1;
                h[rr][j] = __dp4a(a, K1, __dp4a(b, K2, 0u));
            }
        }
        uint4 pk;
        pk.x = h[0][0] | (h[1][0] << 16);
        pk.y = h[0][1] | (h[1][1] << 16);
        pk.z = h[0][2] | (h[1][2] << 16);
        pk.w = h[0][3] | (h[1][3] << 16);
        *reinterpret_cast<uint4*>(sHp + m * kTileW + g * 4) = pk;
    }
    __syncthreads();

    // vertical pass: thread = 4 columns x 4 rows; output row 4q+i taps staged rows 4q+i .. 4q+i+6 = pairs 2q .. 2q+4
    const int tx = tid & 31, q = tid >> 5;
    uint4 pr[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) pr[k] = *reinterpret_cast<const uint4*>(sHp + (2 * q + k) * kTileW + tx * 4);
    const int x = x0 + tx * 4;
    uint8_t* out = blur + (size_t)blockIdx.y * P.blurBytes + L.blurOff + (size_t)(y0 + q * 4) * L.bpitch + x;
    // coefficient byte pairs (low byte multiplies the even row of a pair)
    constexpr uint32_t Ea = 0x2212u, Eb = 0x3830u, Ec = 0x2230u, Ed = 0x0012u;   // even start: 18,34 | 48,56 | 48,34 | 18,0
    constexpr uint32_t Oa = 0x1200u, Ob = 0x3022u, Oc = 0x3038u, Od = 0x1222u;   // odd start:  0,18 | 34,48 | 56,48 | 34,18
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int y = y0 + q * 4 + i;
        const int b = i >> 1;  // first pair
        const uint32_t ca = (i & 1) ? Oa : Ea, cb = (i & 1) ? Ob : Eb, cc = (i & 1) ? Oc : Ec, cd = (i & 1) ? Od : Ed;
        uint32_t v[4];
        const uint32_t* p0 = &pr[b].x; const uint32_t* p1 = &pr[b + 1].x; const uint32_t* p2 = &pr[b + 2].x; const uint32_t* p3 = &pr[b + 3].x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t acc = __dp2a_lo(p0[k], ca, 32768u);
            acc = __dp2a_lo(p1[k], cb, acc);
            acc = __dp2a_lo(p2[k], cc, acc);
            acc = __dp2a_lo(p3[k], cd, acc);
            v[k] = acc >> 16;
        }
        const uint32_t word = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
        if (y < L.h && x < L.bpitch) *reinterpret_cast<uint32_t*>(out + i * L.bpitch) = word;
    }
}


// horizontal 7-tap sums of 4 neighbouring pixels from three aligned words WITHOUT funnel shifts: pixel x+j taps bytes (j .. j+6) of the
// 12-byte window, so each pixel is two or three dp4a against coefficient words that carry the shift (10 dp4a per 4 pixels instead of
// 8 dp4a + 6 SHF: the shifts run on the ALU pipe, the dot products on the FMA pipe)
__device__ __forceinline__ void hsum4(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t (&h)[4]) {
    h[0] = __dp4a(w0, 0x38302212u, __dp4a(w1, 0x00122230u, 0u));
    h[1] = __dp4a(w0, 0x30221200u, __dp4a(w1, 0x12223038u, 0u));
    h[2] = __dp4a(w0, 0x22120000u, __dp4a(w1, 0x22303830u, __dp4a(w2, 0x00000012u, 0u)));
    h[3] = __dp4a(w0, 0x12000000u, __dp4a(w1, 0x30383022u, __dp4a(w2, 0x00001222u, 0u)));
}

template <bool ONEWAIT, bool REMAP, bool NOSHIFT = false>
__global__ void __launch_bounds__(256, 8) blurKernel(const __grid_constant__ FramePlan P, const __grid_constant__ LevelMaps M,
                                                  const uint32_t* __restrict__ tileTab, uint8_t* __restrict__ blur, int frame0) {
    __shared__ __align__(128) uint8_t sIn[kInRows * kInPitch];
    __shared__ __align__(16) uint32_t sHp[kPairs * kTileW];  // (H[2m][x] | H[2m+1][x] << 16)
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x;
    const uint32_t te = __ldg(tileTab + blockIdx.x);  // level << 24 | tile row << 12 | tile column
    const int l = (int)(te >> 24);
    const int x0 = (int)(te & 0xfffu) * kTileW, y0 = (int)((te >> 12) & 0xfffu) * kTileH;
    const LevelPlan& L = P.lv[l];
    // thread 0 alone owns the mbarrier (init, expect, wait); the block barrier behind its wait publishes the tile
    if (tid == 0) {
        mbarInit(&bar, 1);
        mbarInitFence();
        // padded coordinates (x0 + 16, y0 + 16): staged byte 0 of a row = level column x0 - 3
        mbarExpectTx(&bar, kInRows * kInPitch);
        tmaLoadTile(sIn, &M.m[l], x0 + 16, y0 + 16, frame0 + (int)blockIdx.y, &bar);
        if (ONEWAIT) mbarWait(&bar, 0);
    }
    if (!ONEWAIT) {
        __syncthreads();
        mbarWait(&bar, 0);
    }
    // the part of the tile inside the level: nG groups of 4 columns, `rows` output rows -> nP pairs of staged rows.  Items are
    // numbered over the VALID part only, so a tile at the right / bottom edge costs its share of instructions (15 % of the
    // 128 x 32 tiles' pixels lie outside the levels at 752 x 480).
    const int nG = REMAP ? min(kTileW / 4, (L.w - x0 + 3) >> 2) : kTileW / 4;  // columns [w, bpitch) of the blurred plane are never read
    const int rows = REMAP ? min(kTileH, L.h - y0) : kTileH;
    const int nP = (rows + 7) >> 1;  // staged rows 0 .. rows + 5
    const unsigned divG = 65536u / (unsigned)nG + 1u;  // (i * divG) >> 16 == i / nG for i < 2048
    if (ONEWAIT) __syncthreads();

    // horizontal pass: item = (row pair m, 4 columns g)
    constexpr uint32_t K1 = 0x38302212u, K2 = 0x00122230u;  // (18,34,48,56), (48,34,18,0)
    for (int i = tid; i < nP * nG; i += 256) {
        const int m = REMAP ? (int)(((unsigned)i * divG) >> 16) : (i >> 5), g = i - m * nG;
        uint32_t h[2][4];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(sIn + (2 * m + rr) * kInPitch) + g;
            const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
            if (NOSHIFT) {
                hsum4(w0, w1, w2, h[rr]);
                continue;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t a = j ? __funnelshift_r(w0, w1, 8 * j) : w0;
                const uint32_t b = j ? __funnelshift_r(w1, w2, 8 * j) : w1;
                h[rr][j] = __dp4a(a, K1, __dp4a(b, K2, 0u));
            }
        }
        uint4 pk;  // low half = row 2m, high half = row 2m + 1 (H < 2^16)
        pk.x = __byte_perm(h[0][0], h[1][0], 0x5410);
        pk.y = __byte_perm(h[0][1], h[1][1], 0x5410);
        pk.z = __byte_perm(h[0][2], h[1][2], 0x5410);
        pk.w = __byte_perm(h[0][3], h[1][3], 0x5410);
        *reinterpret_cast<uint4*>(sHp + m * kTileW + g * 4) = pk;
    }
    __syncthreads();

    // vertical pass: item = (4 output rows q, 4 columns tx); output row 4q+i taps staged rows 4q+i .. 4q+i+6 = pairs 2q .. 2q+4
    // coefficient byte pairs (low byte multiplies the even row of a pair)
    constexpr uint32_t Ea = 0x2212u, Eb = 0x3830u, Ec = 0x2230u, Ed = 0x0012u;   // even start: 18,34 | 48,56 | 48,34 | 18,0
    constexpr uint32_t Oa = 0x1200u, Ob = 0x3022u, Oc = 0x3038u, Od = 0x1222u;   // odd start:  0,18 | 34,48 | 56,48 | 34,18
    const int nQ = (rows + 3) >> 2;
    for (int it = tid; it < nQ * nG; it += 256) {
        const int q = REMAP ? (int)(((unsigned)it * divG) >> 16) : (it >> 5), tx = it - q * nG;
        const int x = x0 + tx * 4, yq = y0 + q * 4;
        if (!REMAP && (x >= L.bpitch || yq >= L.h)) continue;
        uint4 pr[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) pr[k] = *reinterpret_cast<const uint4*>(sHp + (2 * q + k) * kTileW + tx * 4);
        uint8_t* out = blur + (size_t)blockIdx.y * P.blurBytes + L.blurOff + (size_t)yq * L.bpitch + x;
        const int rq = min(4, L.h - yq);
#pragma unroll
        for (int i = 0; i < 4; ++i, out += L.bpitch) {
            const int b = i >> 1;  // first pair
            const uint32_t ca = (i & 1) ? Oa : Ea, cb = (i & 1) ? Ob : Eb, cc = (i & 1) ? Oc : Ec, cd = (i & 1) ? Od : Ed;
            uint32_t v[4];
            const uint32_t* p0 = &pr[b].x; const uint32_t* p1 = &pr[b + 1].x; const uint32_t* p2 = &pr[b + 2].x; const uint32_t* p3 = &pr[b + 3].x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint32_t acc = __dp2a_lo(p0[k], ca, 32768u);
                acc = __dp2a_lo(p1[k], cb, acc);
                acc = __dp2a_lo(p2[k], cc, acc);
                v[k] = __dp2a_lo(p3[k], cd, acc);   // (V + 32768); the pixel is bits 16..23 = byte 2 (V + 32768 < 2^24)
            }
            const uint32_t word = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
            if (i < rq) *reinterpret_cast<uint32_t*>(out) = word;
        }
    }
}

// v5: persistent CTAs, the NEXT tile's TMA box is in flight while the current tile is computed (two input buffers, two
// mbarriers).  The one-tile-per-CTA kernel spends 43 % of its stall samples waiting for its only box.
__global__ void __launch_bounds__(256, 8) blurPersistKernel(const __grid_constant__ FramePlan P, const __grid_constant__ LevelMaps M,
                                                            const uint32_t* __restrict__ tileTab, uint8_t* __restrict__ blur, int frame0,
                                                            int tilesPerFrame, int total) {
    __shared__ __align__(128) uint8_t sIn[2][(kInRows * kInPitch + 127) & ~127];  // TMA destinations are 128-byte aligned
    __shared__ __align__(16) uint32_t sHp[kPairs * kTileW];  // (H[2m][x] | H[2m+1][x] << 16)
    __shared__ __align__(8) uint64_t bar[2];
    const int tid = threadIdx.x;
    const int stride = gridDim.x;
    auto issue = [&](int t, int b) {  // thread 0 only
        const int f = t / tilesPerFrame;
        const uint32_t te = __ldg(tileTab + (t - f * tilesPerFrame));
        const int l = (int)(te >> 24);
        const int x0 = (int)(te & 0xfffu) * kTileW, y0 = (int)((te >> 12) & 0xfffu) * kTileH;
        mbarExpectTx(&bar[b], kInRows * kInPitch);
        tmaLoadTile(sIn[b], &M.m[l], x0 + 16, y0 + 16, frame0 + f, &bar[b]);
    };
    int t = blockIdx.x;
    if (t >= total) return;
    if (tid == 0) {
        mbarInit(&bar[0], 1);
        mbarInit(&bar[1], 1);
        mbarInitFence();
        issue(t, 0);
    }
    constexpr uint32_t K1 = 0x38302212u, K2 = 0x00122230u;  // (18,34,48,56), (48,34,18,0)
    constexpr uint32_t Ea = 0x2212u, Eb = 0x3830u, Ec = 0x2230u, Ed = 0x0012u;   // even start: 18,34 | 48,56 | 48,34 | 18,0
    constexpr uint32_t Oa = 0x1200u, Ob = 0x3022u, Oc = 0x3038u, Od = 0x1222u;   // odd start:  0,18 | 34,48 | 56,48 | 34,18
    const int g = tid & 31, w = tid >> 5;
    for (int i = 0; t < total; ++i, t += stride) {
        const int b = i & 1;
        if (tid == 0) {
            // buffer b ^ 1 was last read by the horizontal pass of the previous iteration, which ended before that iteration's
            // second block barrier: it is free, so the next tile's box can fly while this tile is computed
            if (t + stride < total) issue(t + stride, b ^ 1);
            mbarWait(&bar[b], (uint32_t)((i >> 1) & 1));
        }
        const int f = t / tilesPerFrame;
        const uint32_t te = __ldg(tileTab + (t - f * tilesPerFrame));
        const int l = (int)(te >> 24);
        const int x0 = (int)(te & 0xfffu) * kTileW, y0 = (int)((te >> 12) & 0xfffu) * kTileH;
        const LevelPlan& L = P.lv[l];
        __syncthreads();  // the tile has landed (thread 0 saw the barrier complete); the previous vertical pass is done with sHp
        const uint8_t* in = sIn[b];
        for (int m = w; m < kPairs; m += 8) {
            uint32_t h[2][4];
#pragma unroll
            for (int rr = 0; rr < 2; ++rr) {
                const uint32_t* p = reinterpret_cast<const uint32_t*>(in + (2 * m + rr) * kInPitch) + g;
                const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t a = j ? __funnelshift_r(w0, w1, 8 * j) : w0;
                    const uint32_t bb = j ? __funnelshift_r(w1, w2, 8 * j) : w1;
                    h[rr][j] = __dp4a(a, K1, __dp4a(bb, K2, 0u));
                }
            }
            uint4 pk;
            pk.x = __byte_perm(h[0][0], h[1][0], 0x5410);
            pk.y = __byte_perm(h[0][1], h[1][1], 0x5410);
            pk.z = __byte_perm(h[0][2], h[1][2], 0x5410);
            pk.w = __byte_perm(h[0][3], h[1][3], 0x5410);
            *reinterpret_cast<uint4*>(sHp + m * kTileW + g * 4) = pk;
        }
        __syncthreads();
        const int x = x0 + g * 4, yq = y0 + w * 4;
        if (x < L.bpitch && yq < L.h) {
            uint4 pr[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) pr[k] = *reinterpret_cast<const uint4*>(sHp + (2 * w + k) * kTileW + g * 4);
            uint8_t* out = blur + (size_t)f * P.blurBytes + L.blurOff + (size_t)yq * L.bpitch + x;
            const int rq = min(4, L.h - yq);
#pragma unroll
            for (int r = 0; r < 4; ++r, out += L.bpitch) {
                const int b2 = r >> 1;  // first pair
                const uint32_t ca = (r & 1) ? Oa : Ea, cb = (r & 1) ? Ob : Eb, cc = (r & 1) ? Oc : Ec, cd = (r & 1) ? Od : Ed;
                uint32_t v[4];
                const uint32_t* p0 = &pr[b2].x; const uint32_t* p1 = &pr[b2 + 1].x; const uint32_t* p2 = &pr[b2 + 2].x; const uint32_t* p3 = &pr[b2 + 3].x;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    uint32_t acc = __dp2a_lo(p0[k], ca, 32768u);
                    acc = __dp2a_lo(p1[k], cb, acc);
                    acc = __dp2a_lo(p2[k], cc, acc);
                    v[k] = __dp2a_lo(p3[k], cd, acc);
                }
                const uint32_t word = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
                if (r < rq) *reinterpret_cast<uint32_t*>(out) = word;
            }
        }
    }
}

}  // namespace

size_t levelMapsBytes() { return sizeof(LevelMaps); }

// (level << 24 | tile row << 12 | tile column) of every 128 x 32 blur tile, in the order of the plan's blurTile0
void buildBlurTileTable(const FramePlan& P, std::vector<uint32_t>& tab) {
    tab.clear();
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelPlan& L = P.lv[l];
        const int ty = (L.h + kTileH - 1) / kTileH;
        for (int y = 0; y < ty; ++y)
            for (int x = 0; x < L.blurTilesX; ++x) tab.push_back(((uint32_t)l << 24) | ((uint32_t)y << 12) | (uint32_t)x);
    }
}

int buildBlurMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out) {
    LevelMaps* M = reinterpret_cast<LevelMaps*>(out);
    std::memset(M, 0, sizeof(LevelMaps));
    for (int l = 0; l < P.nlevels; ++l) {
        const int rc = encodeLevelMap(&M->m[l], P, l, pyrBase, capFrames, kInPitch, kInRows);
        if (rc != 0) return rc;
    }
    return 0;
}

void launchBlur(const FramePlan& P, const Buffers& B, int F, cudaStream_t s) {
    static_assert(kTileW == 128 && kTileH == 32, "the plan's blurTile0 / blurTilesX assume 128 x 32 tiles");
    // ORBX_BLUR_MODE (A/B measurements): bit 0 = one mbarrier waiter, bit 1 = items numbered over the valid part of edge tiles.
    // Measured beside the quadtree (blur stage ms / frames per s): mode 0 0.303 / 157.2 k, 1 0.305 / 156.9 k, 2 0.322 / 155.1 k,
    // 3 0.332 / 154.4 k: the 15 % fewer instructions of the edge tiles buy nothing and the extra index arithmetic costs.  Mode 4 =
    // persistent CTAs with the next tile's TMA box in flight while the current tile is computed (blurPersistKernel): 0.379 ms /
    // 153.7 k - eight one-tile CTAs per SM already overlap each other's loads, and the persistent grid leaves the co-running quadtree
    // fewer slots.  Default = one waiter, no remapping, one tile per CTA.
    static const int mode = [] { const char* e = std::getenv("ORBX_BLUR_MODE"); return e ? std::atoi(e) : 1; }();
    const LevelMaps& M = *reinterpret_cast<const LevelMaps*>(B.blurMaps);
    const dim3 grid(P.blurTiles, F);
    if (B.blurMaps && mode == 4) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const long long total = (long long)P.blurTiles * F;
        const int gridP = (int)std::min<long long>(total, (long long)sms * 8);
        blurPersistKernel<<<gridP, 256, 0, s>>>(P, M, B.blurTab, B.blur, B.frame0, P.blurTiles, (int)total);
    } else if (B.blurMaps) {
        if (mode == 5) blurKernel<true, false, true><<<grid, 256, 0, s>>>(P, M, B.blurTab, B.blur, B.frame0);
        else if (mode == 3) blurKernel<true, true><<<grid, 256, 0, s>>>(P, M, B.blurTab, B.blur, B.frame0);
        else if (mode == 2) blurKernel<false, true><<<grid, 256, 0, s>>>(P, M, B.blurTab, B.blur, B.frame0);
        else if (mode == 1) blurKernel<true, false><<<grid, 256, 0, s>>>(P, M, B.blurTab, B.blur, B.frame0);
        else blurKernel<false, false><<<grid, 256, 0, s>>>(P, M, B.blurTab, B.blur, B.frame0);
    }
    else
        blurKernelLd<<<dim3(P.blurTiles, F), 256, 0, s>>>(P, B.pyr, B.blur);
    countLaunch();
}

}  // namespace orbx

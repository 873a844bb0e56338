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
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "orbx_internal.cuh"
#include "tma.cuh"

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

// level 0, rows 16-B aligned (base, stride and frame stride multiples of 16): an item is one 16-byte chunk of a padded
// row, written with one 128-bit store.  Padded x = 16k maps to source x = 16(k-1) - 3, i.e. the aligned source chunk
// k-1 shifted by three bytes: one 128-bit load plus the last word of the chunk before it.  Items are enumerated
// interior chunks first (ph x nFast), then the few chunks per row that touch the border (mirrored per byte), so a
// warp never mixes the two paths.
__global__ void __launch_bounds__(256) pyrLevel0AlignedKernel(const uint8_t* __restrict__ img, int stride, size_t frameStride,
                                                              uint8_t* __restrict__ pyr, unsigned frameBytes, int w, int h,
                                                              int pitch, int kFirst, int nFast, int nChunks) {
    const int ph = h + 2 * kEdge, pw = w + 2 * kEdge;
    const int fastItems = ph * nFast;
    const int nbFast = (fastItems + 511) / 512;  // the first nbFast CTAs: two interior chunks per thread (both loads in flight at once)
    uint8_t* dst = pyr + (size_t)blockIdx.z * frameBytes;
    const uint8_t* frame = img + (size_t)blockIdx.z * frameStride;
    if ((int)blockIdx.x < nbFast) {
        uint4 a[2];
        uint32_t p[2];
        size_t o[2];
        bool ok[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int it = blockIdx.x * 512 + u * 256 + threadIdx.x;
            ok[u] = it < fastItems;
            const int itc = ok[u] ? it : 0;
            const int yp = itc / nFast, k = kFirst + (itc - yp * nFast);
            int y = yp - kEdge;
            y = y < 0 ? -y : (y >= h ? 2 * h - 2 - y : y);
            const uint8_t* src = frame + (size_t)y * stride + (k * 16 - kEdge);
            a[u] = __ldg(reinterpret_cast<const uint4*>(src + 3));        // source bytes x0+3 .. x0+18
            p[u] = __ldg(reinterpret_cast<const uint32_t*>(src - 1));     // source bytes x0-1 .. x0+2
            o[u] = (size_t)yp * pitch + k * 16;
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            uint4 out;
            out.x = __byte_perm(p[u], a[u].x, 0x4321);
            out.y = __byte_perm(a[u].x, a[u].y, 0x4321);
            out.z = __byte_perm(a[u].y, a[u].z, 0x4321);
            out.w = __byte_perm(a[u].z, a[u].w, 0x4321);
            if (ok[u]) *reinterpret_cast<uint4*>(dst + o[u]) = out;
        }
        return;
    }
    const int item = fastItems + ((int)blockIdx.x - nbFast) * 256 + threadIdx.x;
    const int nSlow = nChunks - nFast;
    const int sItem = item - fastItems;
    if (sItem >= ph * nSlow) return;
    const int yp = sItem / nSlow;
    int k = sItem - yp * nSlow;
    if (k >= kFirst) k += nFast;  // chunks 0..kFirst-1 on the left, kFirst+nFast.. on the right
    const uint8_t* src = frame + (size_t)reflect101(yp - kEdge, h) * stride;
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
    *reinterpret_cast<uint4*>(dst + (size_t)yp * pitch + k * 16) = make_uint4(wd[0], wd[1], wd[2], wd[3]);
}

// level l >= 1 from level l-1, tiled and separable.  A CTA owns a 128-column strip of the padded plane and walks
// kRowTiles 32-row tiles of it; the column coefficients live in registers for the whole walk.
// SPT = bytes per staged source row (compile-time so that the row loops unroll into immediate offsets).
#ifndef ORBX_PYR_ROWTILES
#define ORBX_PYR_ROWTILES 4
#endif
constexpr int kRowTiles = ORBX_PYR_ROWTILES;

// TMA = true: the source rows of a tile are staged by ONE cp.async.bulk.tensor.3d (box SPT x maxRows x 1 of the source
// level's tensor map, first byte at the 16-byte aligned plane column A0; what the box holds beyond the nS rows / nC16 chunks
// the tile needs is never read, rows and columns past the plane arrive as zeros).
template <int SPT, bool WINDOWED, bool TMA>
__global__ void __launch_bounds__(256)
    pyrResizeKernel(uint8_t* __restrict__ pyr, unsigned frameBytes, unsigned srcOff, int sw, int sh, int spitch,
                    unsigned dstOff, int dw, int dh, int dpitch, const short4* __restrict__ tabX,
                    const short4* __restrict__ tabY, int maxRows, const __grid_constant__ CUtensorMap srcMap, int frame0,
                    const uint4* __restrict__ gtab, const int* __restrict__ stripA0) {
    extern __shared__ __align__(128) uint8_t smem[];
    // per tile row: first / second source row (relative to S0), b0 << 16, b1 << 16.  TMA: two generations (tile parity), the
    // next tile's rows and its source box are set up while the current tile is computed
    __shared__ int4 sRow[2][kTH];
    __shared__ int sS0[2], sNS[2];
    __shared__ __align__(8) uint64_t bar[2];
    uint32_t barPhase[2] = {0, 0};
    if (TMA) {
        if (threadIdx.x == 0) {
            mbarInit(&bar[0], 1);
            mbarInit(&bar[1], 1);
            mbarInitFence();
        }
        // made visible to the other threads by the first __syncthreads of the tile loop
    }
    // [maxRows][SPT] staged source rows (TMA: two buffers, 128-B aligned), then [maxRows][kTW] horizontal pass
    const int srcBytes = TMA ? ((maxRows * SPT + 127) & ~127) : maxRows * SPT;
    uint8_t* sSrcBuf[2] = {smem, smem + (TMA ? srcBytes : 0)};
    uint32_t* sH = reinterpret_cast<uint32_t*>(smem + (size_t)srcBytes * (TMA ? 2 : 1));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int X0 = blockIdx.x * kTW;
    const int pw = dw + 2 * kEdge, ph = dh + 2 * kEdge;
    (void)tid;
    uint8_t* frame = pyr + (size_t)blockIdx.z * frameBytes;

    // ---- the 4 columns of this lane (the same in every warp) -> source columns and coefficients -----------------
    int o0[4] = {0, 0, 0, 0}, o1[4] = {0, 0, 0, 0}, a0[4] = {0, 0, 0, 0}, a1[4] = {0, 0, 0, 0};
    int A0 = 0, nC16 = 0, wOfs = 0, wSh = 0;
    uint32_t sel[4] = {0, 0, 0, 0}, cf[4] = {0, 0, 0, 0};
    const int X = X0 + 4 * lane;
    if (WINDOWED && gtab) {
        // host-built (buildPyrGroupTable): per group of 4 padded columns the aligned source window, the byte selectors of the
        // four (p0, p1) pairs and the packed coefficients; per strip the first staged byte.  Two 128-bit loads per lane instead
        // of four table look-ups with the border reflection and two warp reductions.
        const uint4 e0 = __ldg(gtab + 2 * (blockIdx.x * 32 + lane)), e1 = __ldg(gtab + 2 * (blockIdx.x * 32 + lane) + 1);
        A0 = __ldg(stripA0 + blockIdx.x);
        nC16 = SPT / 16;  // ordinary-load staging (ORBX_FAST_TMA=0) copies whole staged rows; the allocation has slack behind the last plane
        wOfs = (int)(e0.x & 0xffffu) - A0;
        wSh = (int)(e0.x >> 16);
        sel[0] = e0.y & 0xffu; sel[1] = (e0.y >> 8) & 0xffu; sel[2] = (e0.y >> 16) & 0xffu; sel[3] = e0.y >> 24;
        cf[0] = e0.z; cf[1] = e0.w; cf[2] = e1.x; cf[3] = e1.y;
    } else {
        int cLo = 1 << 30, cHi = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int Xc = min(X0 + 4 * lane + j, pw - 1);
            const short4 tx = __ldg(tabX + reflect101(Xc - kEdge, dw));  // (sx, a0, a1, -)
            o0[j] = tx.x; o1[j] = min(tx.x + 1, sw - 1);
            a0[j] = tx.y; a1[j] = tx.z;
            cLo = min(cLo, o0[j]); cHi = max(cHi, o1[j]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cLo = min(cLo, __shfl_xor_sync(0xffffffffu, cLo, o));
            cHi = max(cHi, __shfl_xor_sync(0xffffffffu, cHi, o));
        }
        A0 = (kEdge + cLo) & ~15;                      // plane column of the first staged byte (16-B aligned)
        nC16 = ((kEdge + cHi) - A0) / 16 + 1;          // 16-B chunks per staged row (<= 16)
#pragma unroll
        for (int j = 0; j < 4; ++j) { o0[j] += kEdge - A0; o1[j] += kEdge - A0; }
        // The 4 columns of a lane read source bytes within a window of <= 8 bytes whenever the scale step is <= 2 (always at
        // ORB-SLAM3's 1.2): then a source row costs the lane three aligned 32-bit loads and two funnel shifts, and each
        // (p0, p1) pair is one PRMT away from the two-way dot product with (a0, a1).  Wider windows keep the byte loads.
        const int wLo = min(min(o0[0], o0[1]), min(o0[2], o0[3]));
        wOfs = wLo & ~3; wSh = (wLo & 3) * 8;   // WINDOWED is chosen on the host: source step < 2 => window <= 8 bytes
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sel[j] = (uint32_t)((o0[j] - wLo) & 7) | ((uint32_t)((o1[j] - wLo) & 7) << 4);
            cf[j] = (uint32_t)a0[j] | ((uint32_t)a1[j] << 16);
        }
    }

    // rows of tile `rt` -> source rows (warp 0); lane 0 then hands the tile's source box to the copy engine (TMA)
    auto setupTile = [&](int rt, int gen) {
        const int Y0 = (blockIdx.y * kRowTiles + rt) * kTH;
        const int Y = min(Y0 + lane, ph - 1);
        const short4 ty = __ldg(tabY + reflect101(Y - kEdge, dh));  // (sy, b0, b1, -)
        const int sy0 = min(max((int)ty.x, 0), sh - 1), sy1 = min(max((int)ty.x + 1, 0), sh - 1);
        int lo = sy0, hi = sy1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        sRow[gen][lane] = make_int4((sy0 - lo) * kTW, (sy1 - lo) * kTW, (int)ty.y << 16, (int)ty.z << 16);
        if (lane == 0) {
            sS0[gen] = lo; sNS[gen] = hi - lo + 1;
            if (TMA) {
                mbarExpectTx(&bar[gen], (uint32_t)(SPT * maxRows));
                tmaLoadTile(sSrcBuf[gen], &srcMap, A0, kEdge + lo, frame0 + (int)blockIdx.z, &bar[gen]);
            }
        }
    };
    if (TMA) {
        __syncthreads();  // the mbarriers are initialised
        if (warp == 0) setupTile(0, 0);
    }

    for (int rt = 0; rt < kRowTiles; ++rt) {
        const int Y0 = (blockIdx.y * kRowTiles + rt) * kTH;
        if (Y0 >= ph) break;
        const int gen = TMA ? (rt & 1) : 0;
        __syncthreads();  // the previous tile's readers are done with sRow / sSrc / sH (TMA: of the OTHER generation too)
        if (TMA) {
            // next tile: rows + source box into the other generation while this tile is computed
            if (warp == 0 && rt + 1 < kRowTiles && Y0 + kTH < ph) setupTile(rt + 1, gen ^ 1);
            mbarWait(&bar[gen], barPhase[gen]);
            barPhase[gen] ^= 1u;
        } else {
            if (warp == 0) setupTile(rt, 0);
            __syncthreads();
        }
        const int S0 = sS0[gen], nS = sNS[gen];
        const uint8_t* sSrc = sSrcBuf[gen];

        // ---- stage the source rows (ordinary loads) -----------------------------------------------------------------
        if (!TMA) {
            // 16 threads per row, 16 rows per sweep
            const uint8_t* src = frame + srcOff + (size_t)(kEdge + S0) * spitch + A0;
            uint8_t* dstS = sSrcBuf[0];
            for (int c = tid & 15; c < nC16; c += 16)
                for (int r = tid >> 4; r < nS; r += 16)
                    *reinterpret_cast<uint4*>(dstS + r * SPT + c * 16) = __ldcg(reinterpret_cast<const uint4*>(src + (size_t)r * spitch) + c);
            __syncthreads();
        }
        (void)S0;

        // ---- horizontal pass: H[s][x] = (p0*a0 + p1*a1) >> 4  (<= 32640) ------------------------------------------
        if (WINDOWED) {
            const uint32_t* row = reinterpret_cast<const uint32_t*>(sSrc + warp * SPT + wOfs);
            uint4* hp = reinterpret_cast<uint4*>(sH + warp * kTW + 4 * lane);
            for (int s = warp; s < nS; s += 8, row += 8 * SPT / 4, hp += 8 * kTW / 4) {
                const uint32_t w0 = row[0], w1 = row[1], w2 = row[2];
                const uint32_t lo = __funnelshift_r(w0, w1, wSh), hi = __funnelshift_r(w1, w2, wSh);
                uint4 hv;
                hv.x = __dp2a_lo(cf[0], __byte_perm(lo, hi, sel[0]), 0u) >> 4;
                hv.y = __dp2a_lo(cf[1], __byte_perm(lo, hi, sel[1]), 0u) >> 4;
                hv.z = __dp2a_lo(cf[2], __byte_perm(lo, hi, sel[2]), 0u) >> 4;
                hv.w = __dp2a_lo(cf[3], __byte_perm(lo, hi, sel[3]), 0u) >> 4;
                *hp = hv;
            }
        } else {
            for (int s = warp; s < nS; s += 8) {
                const uint8_t* row = sSrc + s * SPT;
                uint4 hv;
                hv.x = (uint32_t)((int)row[o0[0]] * a0[0] + (int)row[o1[0]] * a1[0]) >> 4;
                hv.y = (uint32_t)((int)row[o0[1]] * a0[1] + (int)row[o1[1]] * a1[1]) >> 4;
                hv.z = (uint32_t)((int)row[o0[2]] * a0[2] + (int)row[o1[2]] * a1[2]) >> 4;
                hv.w = (uint32_t)((int)row[o0[3]] * a0[3] + (int)row[o1[3]] * a1[3]) >> 4;
                *reinterpret_cast<uint4*>(sH + s * kTW + 4 * lane) = hv;
            }
        }
        __syncthreads();

        // ---- vertical pass + store: ((b0*h0) >> 16) + ((b1*h1) >> 16) + 2 >> 2, at most 255 (b0 + b1 = 2048) ------------
        if (X < pw) {
            const int ry0 = warp * (kTH / 8);
            const int rowsHere = min(kTH / 8, ph - (Y0 + ry0));  // <= 0 below the plane
            uint8_t* outp = frame + dstOff + (size_t)(Y0 + ry0) * dpitch + X;
#pragma unroll
            for (int k = 0; k < kTH / 8; ++k, outp += dpitch) {
                if (k >= rowsHere) break;
                const int4 ri = sRow[gen][ry0 + k];
                const uint4 u = *reinterpret_cast<const uint4*>(sH + ri.x + 4 * lane);
                const uint4 v = *reinterpret_cast<const uint4*>(sH + ri.y + 4 * lane);
                const uint32_t B0 = (uint32_t)ri.z, B1 = (uint32_t)ri.w;  // b << 16: umulhi(b << 16, h) == (b * h) >> 16
                const uint32_t v0 = (__umulhi(B0, u.x) + __umulhi(B1, v.x) + 2u) >> 2;
                const uint32_t v1 = (__umulhi(B0, u.y) + __umulhi(B1, v.y) + 2u) >> 2;
                const uint32_t v2 = (__umulhi(B0, u.z) + __umulhi(B1, v.z) + 2u) >> 2;
                const uint32_t v3 = (__umulhi(B0, u.w) + __umulhi(B1, v.w) + 2u) >> 2;
                *reinterpret_cast<uint32_t*>(outp) = __byte_perm(__byte_perm(v0, v1, 0x0040), __byte_perm(v2, v3, 0x0040), 0x5410);
            }
        }
    }
}

}  // namespace

static int pyrMaxRows(const LevelPlan& S, const LevelPlan& D) {
    return (int)((double)kTH * S.h / D.h) + 4;  // source rows a 32-row tile can need
}

// m[l] (l >= 1) = map of the SOURCE level l-1 with the box the resize of level l stages: 192 bytes x maxRows rows
int buildPyrMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out) {
    LevelMaps* M = reinterpret_cast<LevelMaps*>(out);
    std::memset(M, 0, sizeof(LevelMaps));
    for (int l = 1; l < P.nlevels; ++l) {
        const int rows = pyrMaxRows(P.lv[l - 1], P.lv[l]);
        if (rows > 256) continue;  // falls back to ordinary loads (launchPyramid checks the same bound)
        const int rc = encodeLevelMap(&M->m[l], P, l - 1, pyrBase, capFrames, 192, rows);
        if (rc != 0) return rc;
    }
    return 0;
}

// Per level l >= 1 (windowed levels only): for every group of 4 padded columns 8 words
//   {source window byte (plane column & ~3) | funnel shift << 16, 4 byte selectors, cf0, cf1, cf2, cf3, 0, 0}
// and, behind the groups of the level, one word per 128-column strip: the plane column of the strip's first staged byte.
// offsets[l] = first word of level l's groups, stripOfs[l] = first word of its strip entries (0 = level not windowed).
void buildPyrGroupTable(const FramePlan& P, const std::vector<int16_t>& tab, std::vector<uint32_t>& out, int* offsets, int* stripOfs) {
    out.assign(8, 0u);  // word 0 stays unused so that offset 0 means "no table"
    for (int l = 0; l < kMaxLevels; ++l) offsets[l] = stripOfs[l] = 0;
    for (int l = 1; l < P.nlevels; ++l) {
        const LevelPlan& S = P.lv[l - 1];
        const LevelPlan& D = P.lv[l];
        if (!((double)S.w / D.w < 1.95)) continue;
        const int pw = D.w + 2 * kEdge;
        const int strips = (pw + kTW - 1) / kTW, groups = strips * 32;
        const int16_t* tx = tab.data() + (size_t)D.tabX * 4;
        auto refl = [](int i, int n) {
            if (n == 1) return 0;
            while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
            return i;
        };
        offsets[l] = (int)out.size();
        std::vector<int> cLo(strips, 1 << 30);
        for (int g = 0; g < groups; ++g) {
            int o0[4], o1[4], a0[4], a1[4];
            for (int j = 0; j < 4; ++j) {
                const int Xc = std::min(4 * g + j, pw - 1);
                const int16_t* e = tx + 4 * refl(Xc - kEdge, D.w);
                o0[j] = e[0]; o1[j] = std::min(e[0] + 1, S.w - 1);
                a0[j] = e[1]; a1[j] = e[2];
                cLo[g / 32] = std::min(cLo[g / 32], o0[j]);
            }
            const int wLo = kEdge + std::min(std::min(o0[0], o0[1]), std::min(o0[2], o0[3]));
            uint32_t selp = 0, cf[4];
            for (int j = 0; j < 4; ++j) {
                const uint32_t sel = (uint32_t)((kEdge + o0[j] - wLo) & 7) | ((uint32_t)((kEdge + o1[j] - wLo) & 7) << 4);
                selp |= sel << (8 * j);
                cf[j] = (uint32_t)(uint16_t)a0[j] | ((uint32_t)(uint16_t)a1[j] << 16);
            }
            out.push_back((uint32_t)(wLo & ~3) | ((uint32_t)((wLo & 3) * 8) << 16));
            out.push_back(selp);
            out.push_back(cf[0]); out.push_back(cf[1]); out.push_back(cf[2]); out.push_back(cf[3]);
            out.push_back(0u); out.push_back(0u);
        }
        stripOfs[l] = (int)out.size();
        for (int t = 0; t < strips; ++t) out.push_back((uint32_t)((kEdge + cLo[t]) & ~15));
        while (out.size() % 8) out.push_back(0u);
    }
}

template <int SPT, bool WINDOWED, bool TMA = false>
static void launchResizeT(const FramePlan& P, const Buffers& B, int l, int F, cudaStream_t s) {
    const LevelPlan& S = P.lv[l - 1];
    const LevelPlan& D = P.lv[l];
    const int pw = D.w + 2 * kEdge, ph = D.h + 2 * kEdge;
    const int maxRows = pyrMaxRows(S, D);
    size_t smem = TMA ? 2 * (size_t)((maxRows * SPT + 127) & ~127) + (size_t)maxRows * kTW * 4 : (size_t)maxRows * (SPT + kTW * 4);
    // ORBX_PYR_PAD_KB (measurements only): extra dynamic shared memory per CTA = fewer resident CTAs per SM (occupancy sensitivity)
    static const int padKb = [] { const char* e = std::getenv("ORBX_PYR_PAD_KB"); return e ? std::atoi(e) : 0; }();
    smem += (size_t)padKb * 1024;
    auto kern = pyrResizeKernel<SPT, WINDOWED, TMA>;
    // the opt-in is needed as soon as dynamic + STATIC shared memory pass 48 KB (the kernel has ~1.2 KB of static arrays: exactly
    // 48 KB of dynamic memory - 606 x 177 at scale 1.4 - used to fail the launch with "invalid argument"), hence the margin
    if (smem > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid((pw + kTW - 1) / kTW, ((ph + kTH - 1) / kTH + kRowTiles - 1) / kRowTiles, F);
    static const CUtensorMap kNoMap{};
    const CUtensorMap& map = TMA ? reinterpret_cast<const LevelMaps*>(B.pyrMaps)->m[l] : kNoMap;
    kern<<<grid, 256, smem, s>>>(B.pyr, P.frameBytes, S.planeOff, S.w, S.h, S.pitch, D.planeOff, D.w, D.h, D.pitch,
                                 reinterpret_cast<const short4*>(B.tab) + D.tabX, reinterpret_cast<const short4*>(B.tab) + D.tabY, maxRows,
                                 map, B.frame0,
                                 (WINDOWED && B.pyrGroupTab && D.gtabOff) ? reinterpret_cast<const uint4*>(B.pyrGroupTab + D.gtabOff) : nullptr,
                                 (WINDOWED && B.pyrGroupTab && D.gtabOff) ? reinterpret_cast<const int*>(B.pyrGroupTab + D.gtabStrip) : nullptr);
    countLaunch();
}

void launchPyramid(const FramePlan& P, const Buffers& B, const uint8_t* dImgs, int stride, size_t frameStride, int F,
                   cudaStream_t s, int levels, int firstLevel, int lastLevel) {
    if (levels != 1 && firstLevel <= 0) {
        const LevelPlan& L = P.lv[0];
        const int pw = L.w + 2 * kEdge, ph = L.h + 2 * kEdge;
        const bool aligned = ((reinterpret_cast<uintptr_t>(dImgs) | (uintptr_t)stride | (uintptr_t)frameStride) & 15) == 0;
        // interior chunks k: source x0 = 16k - 19 >= 1 and x0 + 19 <= w
        const int kFirst = 2, nFast = L.w / 16 - kFirst + 1, nChunks = (pw + 15) / 16;
        if (aligned && nFast > 0 && L.w >= 20 && L.h >= 20) {
            const long long fastItems = (long long)ph * nFast, slowItems = (long long)ph * (nChunks - nFast);
            dim3 grid((unsigned)((fastItems + 511) / 512 + (slowItems + 255) / 256), 1, F);
            pyrLevel0AlignedKernel<<<grid, 256, 0, s>>>(dImgs, stride, frameStride, B.pyr, P.frameBytes, L.w, L.h, L.pitch,
                                                        kFirst, nFast, nChunks);
        } else {
            dim3 grid((pw + 255) / 256, (ph + 3) / 4, F);
            pyrLevel0Kernel<<<grid, dim3(64, 4), 0, s>>>(dImgs, stride, frameStride, B.pyr, P.frameBytes, L.w, L.h, L.pitch);
        }
        countLaunch();
    }
    if (levels == 0) return;
    for (int l = std::max(1, firstLevel); l < P.nlevels && l <= lastLevel; ++l) {
        // bytes of a source row a 128-column strip can need: 128 * scale + 2, + 15 for the 16-B alignment, rounded to chunks
        const double need = 128.0 * P.lv[l - 1].w / P.lv[l].w + 2 + 15 + 16;
        // windowed horizontal pass: 4 neighbouring columns span <= 8 source bytes iff floor(3 * step) + 2 <= 7
        const bool windowed = (double)P.lv[l - 1].w / P.lv[l].w < 1.95;
        if (need <= 192 && windowed && B.pyrMaps && pyrMaxRows(P.lv[l - 1], P.lv[l]) <= 256) launchResizeT<192, true, true>(P, B, l, F, s);
        else if (need <= 192 && windowed) launchResizeT<192, true>(P, B, l, F, s);
        else if (need <= 192) launchResizeT<192, false>(P, B, l, F, s);
        else if (need <= 512) launchResizeT<512, false>(P, B, l, F, s);  // scale factors up to ~3.7
        else launchResizeT<2048, false>(P, B, l, F, s);
    }
}

}  // namespace orbx

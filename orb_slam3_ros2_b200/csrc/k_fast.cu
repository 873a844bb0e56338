// k_fast.cu — K-b: FAST-9/16 over the reference's 35-px cell grid with the iniThFAST -> minThFAST fallback.
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (ORB_SLAM3/src/ORBextractor.cc:787-872) and the
// cv::FAST(cell, kps, th, true) calls inside it (:826, :845).  Semantics reproduced:
//   * a pixel is a corner at threshold t iff some 9 of the 16 ring pixels are all > I+t or all < I-t; with
//     arc(p) = max over the sixteen 9-arcs of max(min(I-ring), min(ring-I)) that is arc(p) > t, and FAST's
//     response is arc(p) - 1 — independent of t;
//   * FAST is called on the CELL SUB-IMAGE: only pixels >= 3 px inside the cell are tested and the 3x3 strict
//     non-max suppression sees everything outside that rectangle as 0;
//   * a corner survives iff its score is strictly greater than the scores of its 8 neighbours, where a neighbour
//     that is not a corner at t scores 0 — and since such a neighbour has arc <= t < arc(p), comparing RAW arcs
//     gives the same answer, so the arc map needs no thresholding;
//   * a cell that yields nothing at iniThFAST is redone at minThFAST;
//   * output order inside a cell is row-major; cells are visited row-major (:863-868).
//
// Work decomposition (v3): ONE WARP PER CELL, eight cells per CTA, no block-level synchronisation.
//   1. the cell (plus the bytes left of it down to a 16-B boundary) is staged in shared memory with 128-bit loads;
//   2. quick reject, 16 pixels per lane (128-bit shared loads) with byte SIMD: any 9-arc contains ring[0] or ring[8] and ring[4] or ring[12],
//      so a corner needs max(|up-c|,|dn-c|) > t and max(|l-c|,|r-c|) > t.  |.| is the native VABSDIFF4; "> t" per
//      byte is one masked add whose carry lands in bit 7.  ~97 % of the pixels end here at t = 20;
//   3. the flagged 16-pixel chunks are appended, in row-major order, to a list (ballot + popc), expanded to a pixel
//      list (segments of <= 256 pixels), and the exact arc strength is computed for two pixels per lane with the packed 16-bit VIMNMX3
//      (80 three-input min/max per pair) and written to a byte arc map (0 everywhere else);
//   4. every listed pixel with arc > t is compared with its 8 neighbours in the arc map; survivors are written
//      in order (ballot + popc again) as resp<<24 | y<<12 | x.
//
// v7 staging: the narrow class (every level of a normally sized image) stages its tile with TMA — ONE
// cp.async.bulk.tensor.3d per cell, issued by one lane against a per-level tensor map of the padded pyramid planes
// (dims pitch x rows x frames, box 64 x (hCell+6) x 1, no swizzle: the box lands as dense 64-byte rows), completion on a
// per-warp mbarrier; the lanes clear the arc map while the copy is in flight.  The copy engine needs the box to START on a
// 16-byte boundary of global memory (measured: an unaligned x coordinate raises "illegal instruction",
// tools/tma_probe.cu), so the box starts at the cell's x rounded down to 16 and the tile keeps that shift
// (tile x = cell x + (gx0 & 15) <= 15 + 48 - 1 < 64: the SHIFTED layout).  The arc map stays at the unshifted 48-byte pitch
// (its rows are 8 bank phases apart and the warp's work area still fits 4 CTAs per SM).  Reads past the plane are
// zero-filled by the copy engine.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "orbx_internal.cuh"
#include "tma.cuh"
#include "fast_common.cuh"

namespace orbx {
namespace {

#ifndef ORBX_FAST_WARPS
#define ORBX_FAST_WARPS 8
#endif
constexpr int kWarpsPerCta = ORBX_FAST_WARPS;
constexpr int kPixCap = 256;      // pixel list capacity per warp (one segment of flagged pixels)

using FastMaps = LevelMaps;

// PITCH = bytes per tile row in shared memory (multiple of 16).
// SHIFTED = true : tile x = cell x + (global column & 15): rows are copied chunk for chunk (cells up to 76 px wide).
// SHIFTED = false: tile x = cell x: the staging realigns the row with funnel shifts, so a cell of <= 48 columns fits a
//                  48-byte row.  12 words per row spread consecutive rows over 8 bank phases (a 64-byte row has 2: rows y and
//                  y+2 collide, which cost the ring gathers 2.3 wavefronts per load), the quick test's chunk index becomes
//                  contiguous, and its left / right neighbour words come from the neighbouring lanes instead of shared memory.
// TMA = true (needs SHIFTED): the tile is staged by the copy engine from the per-level tensor maps in M; the arc map
// then has its own pitch APITCH = 48 and is indexed with cell x (tile x - shift).
template <int PITCH, bool SHIFTED, bool TMA>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 32 / kWarpsPerCta)
    fastCellsKernel(const __grid_constant__ FramePlan P, const __grid_constant__ FastMaps M, const uint8_t* __restrict__ pyr,
                    uint32_t* __restrict__ cand, int* __restrict__ cellCnt, int rowsMax, int itemsMax, int warpBytes,
                    const uint32_t* __restrict__ cellTab, int cellsInLaunch, int totalWork, int* __restrict__ workCounter,
                    int frame0, const unsigned long long* __restrict__ fbList, const int* __restrict__ fbCount) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int P16 = PITCH / 16;
    constexpr int APITCH = TMA ? 48 : PITCH, A16 = APITCH / 16;
    static_assert(!TMA || SHIFTED, "the TMA box starts on a 16-byte boundary: the tile keeps the shift");
    const int lane = threadIdx.x & 31, wip = threadIdx.x >> 5;
    const unsigned ltMask = (1u << lane) - 1u;
    // LIST MODE (fbList != nullptr): the work items are the (frame, cell) entries the strip kernel (k_fast_strip.cu) left
    // behind - cells with no corner at iniThFAST (entry bit 31 set: start at the minThFAST pass, ORBextractor.cc:843-846)
    // and the cells of strips whose lists overflowed (both passes); their number is only known on the device
    if (fbList) {
        totalWork = min(__ldg(fbCount), totalWork);
        if (totalWork <= 0) return;
    }
    // per-warp mbarrier of the TMA staging, behind the warps' work areas
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + (size_t)kWarpsPerCta * warpBytes) + wip;
    uint32_t barPhase = 0;
    if (TMA) {
        if (lane == 0) {
            mbarInit(bar, 1);
            mbarInitFence();
        }
        __syncwarp();
    }

    // ---- persistent warps: cells are handed out through a global counter, so a warp that drew an expensive cell (the
    // minThFAST fallback costs several times the common case) never keeps seven finished warps' shared memory idle ----
    for (;;) {
    __syncwarp();
    int work = 0;
    if (lane == 0) work = atomicAdd(workCounter, 1);
    work = __shfl_sync(0xffffffffu, work, 0);
    if (work >= totalWork) break;
    int f, pass0 = 0;
    uint32_t ce;
    if (fbList) {
        const unsigned long long e = fbList[work];
        f = (int)(e >> 32);
        ce = (uint32_t)e & 0x7fffffffu;
        pass0 = (int)((uint32_t)e >> 31);
    } else {
        f = work / cellsInLaunch;
        // (level, cell row, cell column) of the launch's k-th cell: a table built with the plan (level-major, row-major)
        ce = __ldg(cellTab + (work - f * cellsInLaunch));
    }
    const int l = (int)(ce >> 24), ci = (int)((ce >> 12) & 0xfffu), cj = (int)(ce & 0xfffu);
    const LevelPlan& L = P.lv[l];
    const int c = ci * L.nCols + cj;
    int* cntOut = cellCnt + (size_t)f * P.nCellsTotal + L.cellBase + c;
    const int maxBX = L.w - kMinBorder, maxBY = L.h - kMinBorder;
    const int iniX = kMinBorder + cj * L.wCell, iniY = kMinBorder + ci * L.hCell;
    int maxX = iniX + L.wCell + 6, maxY = iniY + L.hCell + 6;
    if (maxX > maxBX) maxX = maxBX;
    if (maxY > maxBY) maxY = maxBY;
    const int cw = maxX - iniX, ch = maxY - iniY;
    // ORBextractor.cc:813-816, :820-823 skip rules; cv::FAST returns nothing below 7x7
    if (iniY >= maxBY - 3 || iniX >= maxBX - 6 || cw < 7 || ch < 7) {
        if (lane == 0) *cntOut = 0;
        continue;
    }

    uint8_t* tile = smem + (size_t)wip * warpBytes;      // [rowsMax][PITCH] pixels; tile x = cell x + shift
    uint8_t* arcStore = tile + rowsMax * PITCH;          // [rowsMax-4][APITCH] arc strength of tile rows 2 .. ch-3
    uint8_t* arc = arcStore - 2 * APITCH;                // indexed by (tile row, tile x - ashift); 0 where not computed
    uint32_t* glist = reinterpret_cast<uint32_t*>(arcStore + (rowsMax - 4) * APITCH);  // flagged chunks: item << 16 | mask
    uint16_t* pbuf = reinterpret_cast<uint16_t*>(glist + itemsMax);  // pixel list of the current segment: y << 7 | x
    const uint32_t* tile32 = reinterpret_cast<const uint32_t*>(tile);
    const uint4* tile128 = reinterpret_cast<const uint4*>(tile);

    // ---- stage the cell with 128-bit loads; clear the arc map --------------------------------------------------
    const int gx0 = kEdge + iniX;
    const int shift = SHIFTED ? (gx0 & 15) : 0;
    if (TMA) {
        // the previous cell's reads of this warp's tile are complete (__syncwarp at the loop head): hand the tile to the
        // copy engine, clear the arc map meanwhile, then wait for the bytes
        if (lane == 0) {
            mbarExpectTx(bar, (uint32_t)(PITCH * (L.hCell + 6)));
            tmaLoadTile(tile, &M.m[l], gx0 - shift, kEdge + iniY, frame0 + f, bar);
        }
        uint4* adst = reinterpret_cast<uint4*>(arcStore);
        const int nA = (ch - 4) * A16;
        for (int i = lane; i < nA; i += 32) adst[i] = make_uint4(0, 0, 0, 0);
        mbarWait(bar, barPhase);
        barPhase ^= 1u;
    } else {
        constexpr int kRowsPerSweep = 32 / P16;           // whole rows per sweep: lane = row-in-sweep * P16 + chunk
        const int lr = lane / P16, k = lane - lr * P16;
        const bool laneOn = lr < kRowsPerSweep;
        const int s16 = gx0 & 15, wsel = s16 >> 2, bsh = (s16 & 3) * 8;
        const uint8_t* src = pyr + (size_t)f * P.frameBytes + L.planeOff + (size_t)(kEdge + iniY) * L.pitch + (gx0 - s16);
        const int nIn = (s16 + cw + 15) >> 4;  // aligned source chunks that hold pixels of the cell
        const uint4* rowp = reinterpret_cast<const uint4*>(src + (size_t)lr * L.pitch) + k;
        const size_t rowStep = (size_t)kRowsPerSweep * L.pitch / 16;
        uint4* dst = reinterpret_cast<uint4*>(tile) + lr * P16 + k;
        uint4* adst = reinterpret_cast<uint4*>(arcStore) + lr * P16 + k;
        for (int r = lr; r < ch; r += kRowsPerSweep, rowp += rowStep, dst += kRowsPerSweep * P16, adst += kRowsPerSweep * P16) {
            if (!laneOn) break;
            uint4 v;
            if (SHIFTED) {
                v = __ldg(rowp);
            } else {
                // output chunk k of a row = bytes [16k + s, 16k + s + 16) of the 16-B aligned source row, s = gx0 & 15: two
                // aligned 128-bit loads, a warp-uniform word offset and four funnel shifts
                const uint4 a = k < nIn ? __ldg(rowp) : make_uint4(0, 0, 0, 0);
                const uint4 b = (s16 != 0 && k + 1 < nIn) ? __ldg(rowp + 1) : make_uint4(0, 0, 0, 0);
                uint32_t w0, w1, w2, w3, w4;
                switch (wsel) {
                    case 0: w0 = a.x; w1 = a.y; w2 = a.z; w3 = a.w; w4 = b.x; break;
                    case 1: w0 = a.y; w1 = a.z; w2 = a.w; w3 = b.x; w4 = b.y; break;
                    case 2: w0 = a.z; w1 = a.w; w2 = b.x; w3 = b.y; w4 = b.z; break;
                    default: w0 = a.w; w1 = b.x; w2 = b.y; w3 = b.z; w4 = b.w; break;
                }
                v.x = __funnelshift_r(w0, w1, bsh); v.y = __funnelshift_r(w1, w2, bsh);
                v.z = __funnelshift_r(w2, w3, bsh); v.w = __funnelshift_r(w3, w4, bsh);
            }
            *dst = v;
            if (r < ch - 4) *adst = make_uint4(0, 0, 0, 0);
        }
    }
    __syncwarp();

    // tested pixels: tile x in [xs, xe), rows [3, ch-3); an item of the quick test is one 16-byte chunk of a row
    const int xs = shift + 3, xe = shift + cw - 3;
    const int c0 = xs >> 4, nC = ((xe - 1) >> 4) - c0 + 1, nR = ch - 6;
    const int nItems = nR * nC;
    const unsigned divC = 65536u / (unsigned)nC + 1u;   // (i * divC) >> 16 == i / nC for i < 2048, nC <= 32
    uint32_t* slot = cand + (size_t)f * P.candWords + L.candOff + (size_t)c * L.cellCap;
    const int xOut = cj * L.wCell - shift, yOut = ci * L.hCell;  // tile coords -> cell-grid coords of :865-866
    // response = arc - 1 must beat the zero score of non-corners: arc >= 2 (only matters for thresholds < 1)
    const int tIni = max(P.iniTh, 1), tMin = max(P.minTh, 1);

    int kept = 0;
    for (int pass = pass0; pass < 2; ++pass) {
        const int t = pass == 0 ? tIni : tMin;
        if (pass == 1 && (kept > 0 || tMin >= tIni)) break;  // :843-846: redo at minThFAST only if nothing was found
        // per byte: ((d & 0x7f) + K) | d has bit 7 set iff d > tq, tq = min(t, 126) (a NECESSARY condition for t > 126)
        const uint32_t K = (uint32_t)(0x7f - min(t, 126)) * 0x01010101u;

        // ---- quick reject, 16 pixels per lane: flagged chunks appended in row-major order ------------------------
        int nG = 0;
        // SHIFTED: 32 consecutive items per sweep.  Unshifted: whole rows per sweep (lane = row-in-sweep * nC + chunk), so a
        // chunk's left / right neighbour words are the neighbouring lanes' registers.
        const int rowsPerSweep = SHIFTED ? 0 : 32 / nC;
        const int itemsPerSweep = SHIFTED ? 32 : rowsPerSweep * nC;
        for (int i0 = 0; i0 < nItems; i0 += itemsPerSweep) {
            const int i = i0 + lane;
            const bool act = i < nItems && lane < itemsPerSweep;
            uint32_t m = 0;
            int cx = 0, idx = 0;
            uint4 cc = make_uint4(0, 0, 0, 0), up = cc, dn = cc;
            if (act) {
                const int r = (int)(((unsigned)i * divC) >> 16);
                cx = c0 + (i - r * nC);
                idx = (r + 3) * P16 + cx;
                cc = tile128[idx]; up = tile128[idx - 3 * P16]; dn = tile128[idx + 3 * P16];
            }
            uint32_t lw, rw;
            if (SHIFTED) {
                lw = act ? tile32[4 * idx - 1] : 0u;
                rw = act ? tile32[4 * idx + 4] : 0u;
            } else {
                // chunk 0 never needs its left word (pixels 0..2 are not tested); the last chunk of a row reads its right one
                lw = __shfl_up_sync(0xffffffffu, cc.w, 1);
                rw = __shfl_down_sync(0xffffffffu, cc.x, 1);
                if (act && cx == nC - 1) rw = tile32[4 * idx + 4];
            }
            if (act) {
                const uint32_t cw4[6] = {lw, cc.x, cc.y, cc.z, cc.w, rw};
                const uint32_t uw[4] = {up.x, up.y, up.z, up.w}, dw4[4] = {dn.x, dn.y, dn.z, dn.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t ce = cw4[k + 1];
                    const uint32_t lf = __byte_perm(cw4[k], ce, 0x4321);      // pixels x-3 .. x
                    const uint32_t rt = __byte_perm(ce, cw4[k + 2], 0x6543);  // pixels x+3 .. x+6
                    const uint32_t d0 = __vabsdiffu4(uw[k], ce), d8 = __vabsdiffu4(dw4[k], ce);
                    const uint32_t d4 = __vabsdiffu4(rt, ce), d12 = __vabsdiffu4(lf, ce);
                    const uint32_t v = (((d0 & 0x7f7f7f7fu) + K) | d0) | (((d8 & 0x7f7f7f7fu) + K) | d8);
                    const uint32_t h = (((d4 & 0x7f7f7f7fu) + K) | d4) | (((d12 & 0x7f7f7f7fu) + K) | d12);
                    const uint32_t F = v & h & 0x80808080u;
                    m |= (((F >> 7) * 0x01020408u) >> 24) << (4 * k);  // bit 4k+j = pixel 4k+j of the chunk
                }
                // drop the pixels of edge chunks that lie outside the tested columns
                const int x0 = cx << 4;
                const int lo = min(max(xs - x0, 0), 16), hi = min(max(xe - x0, 0), 16);
                m &= ((1u << hi) - 1u) & ~((1u << lo) - 1u);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, m != 0);
            if (m) glist[nG + __popc(bal & ltMask)] = ((uint32_t)i << 16) | m;
            nG += __popc(bal);
        }
        __syncwarp();

        // ---- segments of <= kPixCap flagged pixels: expand -> exact arcs (phase A); then expand -> NMS -> output (B) ----
        bool single = false;
        int nPix = 0;
        for (int phase = 0; phase < 2; ++phase) {
            int cursor = 0, nSeg = 0;
            while (cursor < nG) {
                if (!(phase == 1 && single)) {  // with a single segment the pixel list of phase A is still valid
                    nPix = 0;
                    while (cursor < nG) {
                        const int g = cursor + lane;
                        uint32_t mm = 0;
                        int y = 0, x0 = 0;
                        if (g < nG) {
                            const uint32_t e = glist[g];
                            const int i = (int)(e >> 16);
                            const int r = (int)(((unsigned)i * divC) >> 16);
                            y = r + 3;
                            x0 = (c0 + (i - r * nC)) << 4;
                            mm = e & 0xffffu;
                        }
                        const int n = __popc(mm);
                        int incl = n;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int v = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o) incl += v;
                        }
                        const bool accept = g < nG && nPix + incl <= kPixCap;  // a prefix of the lanes
                        const int nAcc = __popc(__ballot_sync(0xffffffffu, accept));
                        if (accept) {
                            // two pixels per trip: the lowest set bit goes to the front of the lane's range, the highest to
                            // its back, so the range stays sorted and the longest lane needs 8 trips instead of 16
                            int pos = nPix + incl - n, end = nPix + incl - 1;
                            const uint32_t yb = (uint32_t)(y << 7) | (uint32_t)x0;
                            while (mm) {
                                const int jl = __ffs(mm) - 1, jh = 31 - __clz(mm);
                                pbuf[pos++] = (uint16_t)(yb + jl);
                                pbuf[end--] = (uint16_t)(yb + jh);
                                mm &= mm - 1;
                                mm &= ~(1u << jh);
                            }
                        }
                        if (nAcc) nPix += __shfl_sync(0xffffffffu, incl, nAcc - 1);
                        cursor += nAcc;
                        if (nAcc < 32) break;  // list exhausted, or the segment is full
                    }
                    __syncwarp();
                } else {
                    cursor = nG;
                }
                ++nSeg;
                if (phase == 0) {
                    // exact arc strength, two pixels per lane
                    for (int b = 0; b < nPix; b += 64) {
                        const int j0 = b + lane, j1 = j0 + 32;
                        const bool v0 = j0 < nPix, v1 = j1 < nPix;
                        const int p0 = pbuf[v0 ? j0 : 0], p1 = pbuf[v1 ? j1 : (v0 ? j0 : 0)];
                        const int o0 = (p0 >> 7) * PITCH + (p0 & 127), o1 = (p1 >> 7) * PITCH + (p1 & 127);
                        const int a0 = TMA ? (p0 >> 7) * APITCH + (p0 & 127) - shift : o0;
                        const int a1 = TMA ? (p1 >> 7) * APITCH + (p1 & 127) - shift : o1;
                        constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
                        constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
                        uint32_t q[16];
#pragma unroll
                        for (int k = 0; k < 16; ++k)
                            q[k] = (uint32_t)tile[o0 + dy[k] * PITCH + dx[k]] | ((uint32_t)tile[o1 + dy[k] * PITCH + dx[k]] << 16);
                        const uint32_t cen = (uint32_t)tile[o0] | ((uint32_t)tile[o1] << 16);
                        const uint32_t a = arcStrength2(q, cen);
                        if (v0) arc[a0] = (uint8_t)(a & 0xffu);
                        if (v1) arc[a1] = (uint8_t)(a >> 16);
                    }
                    __syncwarp();
                } else {
                    // strict 3x3 non-max suppression among corners at t, ordered output
                    for (int b = 0; b < nPix; b += 32) {
                        const int j = b + lane;
                        bool keep = false;
                        int a = 0, p = 0;
                        if (j < nPix) {
                            p = pbuf[j];
                            const int o = (p >> 7) * APITCH + (p & 127) - (TMA ? shift : 0);
                            a = arc[o];
                            if (a > t) {
                                const int m0 = __vimax3_s32(arc[o - APITCH - 1], arc[o - APITCH], arc[o - APITCH + 1]);
                                const int m1 = __vimax3_s32(arc[o + APITCH - 1], arc[o + APITCH], arc[o + APITCH + 1]);
                                keep = a > __vimax3_s32(m0, m1, max((int)arc[o - 1], (int)arc[o + 1]));
                            }
                        }
                        const unsigned bal = __ballot_sync(0xffffffffu, keep);
                        if (keep) {
                            const int o = kept + __popc(bal & ltMask);
                            if (o < L.cellCap)
                                slot[o] = ((uint32_t)(a - 1) << 24) | ((uint32_t)((p >> 7) + yOut) << 12) | (uint32_t)((p & 127) + xOut);
                        }
                        kept += __popc(bal);
                    }
                }
            }
            if (phase == 0) single = nSeg == 1;
            __syncwarp();
        }
    }
    if (lane == 0) *cntOut = kept;
    }  // persistent loop
}

struct FastGeom {
    int pitch, rowsMax, itemsMax, warpBytes;
};

FastGeom geomFor(const FramePlan& P, unsigned mask, int pitch, int apitch) {
    FastGeom g{pitch, 7, 1, 0};
    for (int l = 0; l < P.nlevels; ++l) {
        if (!((mask >> l) & 1u)) continue;
        const LevelPlan& L = P.lv[l];
        const int ch = L.hCell + 6, cw = L.wCell + 6;
        g.rowsMax = ch > g.rowsMax ? ch : g.rowsMax;
        const int items = (ch - 6) * (pitch / 16);
        g.itemsMax = items > g.itemsMax ? items : g.itemsMax;
    }
    g.itemsMax = (g.itemsMax + 3) & ~3;
    g.warpBytes = g.rowsMax * pitch + (g.rowsMax - 4) * apitch + 4 * g.itemsMax + 2 * kPixCap;
    g.warpBytes = (g.warpBytes + 127) & ~127;  // TMA destinations are 128-B aligned
    return g;
}

template <int PITCH, bool SHIFTED, bool TMA>
void launchClass(const FramePlan& P, const Buffers& B, int F, unsigned mask, int cells, const uint32_t* cellTab, int* counter,
                 cudaStream_t s, int perSmCap = 0) {
    const FastGeom g = geomFor(P, mask, PITCH, TMA ? 48 : PITCH);
    const size_t smem = (size_t)g.warpBytes * kWarpsPerCta + 8 * kWarpsPerCta;  // + one mbarrier per warp
    auto kern = fastCellsKernel<PITCH, SHIFTED, TMA>;
    if (smem > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int dev = 0, sms = 148, perSm = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kern, kWarpsPerCta * 32, smem);
    const long long total = (long long)cells * F;
    const long long want = (total + kWarpsPerCta - 1) / kWarpsPerCta;
    // ORBX_FAST_PERSM (measurements only): fewer persistent CTAs per SM leave shared memory for another stream's kernels
    static const int capPerSm = [] { const char* e = std::getenv("ORBX_FAST_PERSM"); return e ? std::atoi(e) : 0; }();
    if (capPerSm > 0) perSm = std::min(perSm, capPerSm);
    if (perSmCap > 0) perSm = std::min(perSm, perSmCap);
    const int grid = (int)std::min<long long>(want, (long long)sms * std::max(perSm, 1));
    cudaMemsetAsync(counter, 0, sizeof(int), s);
    static const FastMaps kNoMaps{};
    const FastMaps& maps = TMA ? *reinterpret_cast<const FastMaps*>(B.fastMaps) : kNoMaps;
    kern<<<grid, kWarpsPerCta * 32, smem, s>>>(P, maps, B.pyr, B.cand, B.cellCnt, g.rowsMax, g.itemsMax, g.warpBytes, cellTab, cells,
                                                (int)total, counter, B.frame0, nullptr, nullptr);
    countLaunch();
}

// list mode: a fixed grid of persistent CTAs walks the device-side list (count read on the device; nothing to do = the
// CTAs leave at once)
template <int PITCH, bool SHIFTED, bool TMA>
void launchClassList(const FramePlan& P, const Buffers& B, unsigned mask, const unsigned long long* list, const int* count, int cap,
                     int* counter, cudaStream_t s) {
    const FastGeom g = geomFor(P, mask, PITCH, TMA ? 48 : PITCH);
    const size_t smem = (size_t)g.warpBytes * kWarpsPerCta + 8 * kWarpsPerCta;
    auto kern = fastCellsKernel<PITCH, SHIFTED, TMA>;
    if (smem > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = (int)std::min<long long>(((long long)cap + kWarpsPerCta - 1) / kWarpsPerCta, 2LL * sms);
    static const FastMaps kNoMaps{};
    const FastMaps& maps = TMA ? *reinterpret_cast<const FastMaps*>(B.fastMaps) : kNoMaps;
    kern<<<grid, kWarpsPerCta * 32, smem, s>>>(P, maps, B.pyr, B.cand, B.cellCnt, g.rowsMax, g.itemsMax, g.warpBytes, nullptr, 1, cap, counter,
                                                B.frame0, list, count);
    countLaunch();
}

}  // namespace

static bool wideLevel(const LevelPlan& L);
bool fastWideLevel(const LevelPlan& L) { return wideLevel(L); }

size_t fastMapsBytes() { return sizeof(FastMaps); }

// One 3-D tensor map per narrow-class level over the handle's WHOLE pyramid allocation: x = byte column of the padded
// plane (extent = pitch), y = plane row, z = frame (stride frameBytes); box = 64 x (hCell+6) x 1.  Sub-batches address
// their frames through the z coordinate (Buffers::frame0), so the maps only change when the buffer is reallocated.
int buildFastMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out) {
    FastMaps* M = reinterpret_cast<FastMaps*>(out);
    std::memset(M, 0, sizeof(FastMaps));
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelPlan& L = P.lv[l];
        if (L.nCols <= 0 || L.nRows <= 0 || wideLevel(L)) continue;
        const int rc = encodeLevelMap(&M->m[l], P, l, pyrBase, capFrames, 64, L.hCell + 6);
        if (rc != 0) return rc;
    }
    return 0;
}

// shared by every tile-staging kernel (tma.cuh)
int encodeLevelMap(CUtensorMap* out, const FramePlan& P, int l, const uint8_t* pyrBase, int capFrames, int boxW, int boxH) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn ||
            qres != cudaDriverEntryPointSuccess)
            return -1;
        encode = reinterpret_cast<EncodeFn>(fn);
    }
    const LevelPlan& L = P.lv[l];
    const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)(L.h + 2 * kEdge), (cuuint64_t)std::max(capFrames, 1)};
    const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)P.frameBytes};
    const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(pyrBase) + L.planeOff, dims, strides, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -(int)r - 100;
}

static bool wideLevel(const LevelPlan& L) { return L.wCell + 6 > 48 || L.hCell + 6 > 50; }

bool fastHasBothClasses(const FramePlan& P) {
    bool any[2] = {false, false};
    for (int l = 0; l < P.nlevels; ++l)
        if (P.lv[l].nCols > 0 && P.lv[l].nRows > 0) any[wideLevel(P.lv[l]) ? 1 : 0] = true;
    return any[0] && any[1];
}

// (level << 24 | cell row << 12 | cell column) of every FAST cell, the narrow class first, then the wide one; inside a
// class level-major and row-major, i.e. the reference's visiting order (ORBextractor.cc:805-808)
void buildFastCellTable(const FramePlan& P, std::vector<uint32_t>& tab) {
    tab.clear();
    for (int wide = 0; wide < 2; ++wide)
        for (int l = 0; l < P.nlevels; ++l) {
            const LevelPlan& L = P.lv[l];
            if (L.nCols <= 0 || L.nRows <= 0 || wideLevel(L) != (wide == 1)) continue;
            for (int i = 0; i < L.nRows; ++i)
                for (int j = 0; j < L.nCols; ++j) tab.push_back(((uint32_t)l << 24) | ((uint32_t)i << 12) | (uint32_t)j);
        }
}

// which: -1 = both classes back to back on s; 0 = only the narrow class; 1 = only the wide class; 2 = only level 0 of the narrow
// class, on at most 2 persistent CTAs per SM (it runs beside the resize of levels 1..7, whose CTAs need the rest of the shared
// memory); 3 = the narrow class without level 0.  The parts touch disjoint cells, candidate slots and work counters, so they
// may run on different streams at the same time.
void launchFast(const FramePlan& P, const Buffers& B, int F, cudaStream_t s, int which) {
    // Two launches at most: levels whose cells fit a 48-byte unshifted tile row (cell width wCell+6 <= 48) and 50 rows —
    // every level of a large image — and the few small levels with wider or taller cells, staged chunk for chunk with
    // the global misalignment kept (wCell, hCell <= 70 always, since nCols = floor(width / 35); 15 + 76 + 4 <= 96).
    int first = 0;
    for (int wide = 0; wide < 2; ++wide) {
        unsigned mask = 0;
        int cells = 0;
        for (int l = 0; l < P.nlevels; ++l) {
            const LevelPlan& L = P.lv[l];
            if (L.nCols <= 0 || L.nRows <= 0) continue;
            if (wideLevel(L) != (wide == 1)) continue;
            mask |= 1u << l;
            cells += L.nCols * L.nRows;
        }
        if (!cells) continue;
        const bool wanted = which < 0 || (wide ? which == 1 : (which == 0 || which == 2 || which == 3));
        if (!wanted) {
            first += cells;
            continue;
        }
        if (!wide) {
            // the narrow table is level-major: level 0's cells are its first entries
            int skip = 0, take = cells, cap = 0;
            unsigned m = mask;
            int* ctr = B.workCtr;
            if (which == 2 || which == 3) {
                const int c0 = (mask & 1u) ? P.lv[0].nCols * P.lv[0].nRows : 0;
                if (which == 2) { take = c0; m = mask & 1u; cap = 2; ctr = B.workCtr2; }
                else { skip = c0; take = cells - c0; m = mask & ~1u; }
            }
            if (take > 0) {
                if (B.fastMaps) launchClass<64, true, true>(P, B, F, m, take, B.fastTab + first + skip, ctr, s, cap);
                else launchClass<48, false, false>(P, B, F, m, take, B.fastTab + first + skip, ctr, s, cap);
            }
        } else {
            launchClass<96, true, false>(P, B, F, mask, cells, B.fastTab + first, B.workCtr + 1, s);
        }
        first += cells;
    }
}

// The strip kernel's leftovers: list 0 = narrow-class cells, list 1 = wide-class cells (fbList[cls * cap ..], fbCount[cls]).
// The work counters B.workCtr[0..1] must be zero (the caller clears them together with the list counts).
void launchFastFallback(const FramePlan& P, const Buffers& B, int F, const unsigned long long* fbList, const int* fbCount, int cap,
                        cudaStream_t s) {
    (void)F;
    unsigned maskN = 0, maskW = 0;
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelPlan& L = P.lv[l];
        if (L.nCols <= 0 || L.nRows <= 0) continue;
        (wideLevel(L) ? maskW : maskN) |= 1u << l;
    }
    if (maskN) {
        if (B.fastMaps) launchClassList<64, true, true>(P, B, maskN, fbList, fbCount, cap, B.workCtr, s);
        else launchClassList<48, false, false>(P, B, maskN, fbList, fbCount, cap, B.workCtr, s);
    }
    if (maskW) launchClassList<96, true, false>(P, B, maskW, fbList + cap, fbCount + 1, cap, B.workCtr + 1, s);
}

// level 0 is in the narrow class and so is at least one more level: the early level-0 launch (which = 2 / 3) applies
bool fastLevel0Early(const FramePlan& P) {
    if (P.nlevels < 2 || P.lv[0].nCols <= 0 || P.lv[0].nRows <= 0 || wideLevel(P.lv[0])) return false;
    for (int l = 1; l < P.nlevels; ++l)
        if (P.lv[l].nCols > 0 && P.lv[l].nRows > 0 && !wideLevel(P.lv[l])) return true;
    return false;
}

}  // namespace orbx

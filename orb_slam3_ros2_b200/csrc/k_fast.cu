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
//   2. quick reject, 4 pixels per lane with byte SIMD: any 9-arc contains ring[0] or ring[8] and ring[4] or ring[12],
//      so a corner needs max(|up-c|,|dn-c|) > t and max(|l-c|,|r-c|) > t.  |.| is the native VABSDIFF4; "> t" per
//      byte is one masked add whose carry lands in bit 7.  ~97 % of the pixels end here at t = 20;
//   3. the flagged 4-pixel groups are appended, in row-major order, to a list (ballot + popc), expanded to a pixel
//      list, and the exact arc strength is computed for two pixels per lane with the packed 16-bit VIMNMX3
//      (80 three-input min/max per pair) and written to a byte arc map (0 everywhere else);
//   4. every listed pixel with arc > t is compared with its 8 neighbours in the arc map; survivors are written
//      in order (ballot + popc again) as resp<<24 | y<<12 | x.
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kWarpsPerCta = 8;
constexpr int kPixCap = 512;      // pixel list capacity per warp (one segment = 128 groups of <= 4 pixels)
constexpr int kSegGroups = 128;

__device__ __forceinline__ uint32_t max3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t min3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_s16x2(a, b, c); }

// arc strength of two pixels at once: q[k] = ring pixel k of pixel A in the low half, of pixel B in the high half
__device__ __forceinline__ uint32_t arcStrength2(const uint32_t (&q)[16], uint32_t cen) {
    uint32_t mx3[16], mn3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        mx3[k] = max3(q[k], q[(k + 1) & 15], q[(k + 2) & 15]);
        mn3[k] = min3(q[k], q[(k + 1) & 15], q[(k + 2) & 15]);
    }
    uint32_t mx9[16], mn9[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        mx9[k] = max3(mx3[k], mx3[(k + 3) & 15], mx3[(k + 6) & 15]);
        mn9[k] = min3(mn3[k], mn3[(k + 3) & 15], mn3[(k + 6) & 15]);
    }
    // A = min_k max9_k (strongest "centre brighter": I - A), B = max_k min9_k ("centre darker": B - I)
    uint32_t A = min3(mx9[0], mx9[1], mx9[2]), B = max3(mn9[0], mn9[1], mn9[2]);
#pragma unroll
    for (int k = 3; k < 15; k += 2) {
        A = min3(A, mx9[k], mx9[k + 1]);
        B = max3(B, mn9[k], mn9[k + 1]);
    }
    A = __vmins2(A, mx9[15]);
    B = __vmaxs2(B, mn9[15]);
    return __vimax3_s16x2_relu(__vsub2(cen, A), __vsub2(B, cen), 0u);
}

// PITCH = bytes per tile row in shared memory (multiple of 16)
template <int PITCH>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
    fastCellsKernel(const __grid_constant__ FramePlan P, const uint8_t* __restrict__ pyr, uint32_t* __restrict__ cand,
                    int* __restrict__ cellCnt, int rowsMax, int itemsMax, int warpBytes, unsigned levelMask) {
    extern __shared__ __align__(16) uint8_t smem[];
    constexpr int P4 = PITCH / 4, P16 = PITCH / 16;
    const int lane = threadIdx.x & 31, wip = threadIdx.x >> 5;
    const unsigned ltMask = (1u << lane) - 1u;

    // ---- which cell -------------------------------------------------------------------------------------
    int c = blockIdx.x * kWarpsPerCta + wip, l = 0;
    for (; l < P.nlevels; ++l) {  // cells of the levels selected by levelMask, level-major
        if (!((levelMask >> l) & 1u)) continue;
        const int nc = P.lv[l].nCols * P.lv[l].nRows;
        if (c < nc) break;
        c -= nc;
    }
    if (l >= P.nlevels) return;
    const LevelPlan& L = P.lv[l];
    const int ci = c / L.nCols, cj = c - ci * L.nCols;
    const int f = blockIdx.y;
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
        return;
    }

    uint8_t* tile = smem + (size_t)wip * warpBytes;      // [rowsMax][PITCH] pixels; tile x = cell x + shift
    uint8_t* arc = tile + rowsMax * PITCH;               // same geometry: arc strength, 0 where not computed
    uint16_t* glist = reinterpret_cast<uint16_t*>(arc + rowsMax * PITCH);  // flagged groups: item << 4 | nibble
    uint16_t* pbuf = glist + itemsMax;                   // pixel list of the current segment: y << 7 | x (tile coords)
    const uint32_t* tile32 = reinterpret_cast<const uint32_t*>(tile);

    // ---- stage the cell with 128-bit loads; clear the arc map --------------------------------------------------
    const int gx0 = kEdge + iniX;
    const int shift = gx0 & 15;
    {
        const uint8_t* src = pyr + (size_t)f * P.frameBytes + L.planeOff + (size_t)(kEdge + iniY) * L.pitch + (gx0 - shift);
        const int n16 = ch * P16;
        for (int i = lane; i < n16; i += 32) {
            const int r = i / P16, k = i - r * P16;
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r * L.pitch) + k);
            reinterpret_cast<uint4*>(tile)[i] = v;
            reinterpret_cast<uint4*>(arc)[i] = make_uint4(0, 0, 0, 0);
        }
    }
    __syncwarp();

    // tested pixels: tile x in [xs, xe), rows [3, ch-3)
    const int xs = shift + 3, xe = shift + cw - 3;
    const int w0 = xs >> 2, nW = ((xe - 1) >> 2) - w0 + 1, nR = ch - 6;
    const int nItems = nR * nW;
    const unsigned divW = 65536u / (unsigned)nW + 1u;   // (i * divW) >> 16 == i / nW for i < 2048, nW <= 32
    uint32_t* slot = cand + (size_t)f * P.candWords + L.candOff + (size_t)c * L.cellCap;
    const int xOut = cj * L.wCell - shift, yOut = ci * L.hCell;  // tile coords -> cell-grid coords of :865-866
    // response = arc - 1 must beat the zero score of non-corners: arc >= 2 (only matters for thresholds < 1)
    const int tIni = max(P.iniTh, 1), tMin = max(P.minTh, 1);

    int kept = 0;
    for (int pass = 0; pass < 2; ++pass) {
        const int t = pass == 0 ? tIni : tMin;
        if (pass == 1 && (kept > 0 || tMin >= tIni)) break;  // :843-846: redo at minThFAST only if nothing was found
        // per byte: ((d & 0x7f) + K) | d has bit 7 set iff d > tq, tq = min(t, 126) (a NECESSARY condition for t > 126)
        const uint32_t K = (uint32_t)(0x7f - min(t, 126)) * 0x01010101u;

        // ---- quick reject: flagged 4-pixel groups appended in row-major order ----------------------------------
        int nG = 0;
        for (int i0 = 0; i0 < nItems; i0 += 32) {
            const int i = i0 + lane;
            uint32_t F = 0;
            if (i < nItems) {
                const int r = (int)(((unsigned)i * divW) >> 16);
                const int idx = (r + 3) * P4 + w0 + (i - r * nW);
                const uint32_t cc = tile32[idx];
                const uint32_t up = tile32[idx - 3 * P4], dn = tile32[idx + 3 * P4];
                const uint32_t lf = __byte_perm(tile32[idx - 1], cc, 0x4321);  // pixels x-3 .. x
                const uint32_t rt = __byte_perm(cc, tile32[idx + 1], 0x6543);  // pixels x+3 .. x+6
                const uint32_t d0 = __vabsdiffu4(up, cc), d8 = __vabsdiffu4(dn, cc);
                const uint32_t d4 = __vabsdiffu4(rt, cc), d12 = __vabsdiffu4(lf, cc);
                const uint32_t v = (((d0 & 0x7f7f7f7fu) + K) | d0) | (((d8 & 0x7f7f7f7fu) + K) | d8);
                const uint32_t h = (((d4 & 0x7f7f7f7fu) + K) | d4) | (((d12 & 0x7f7f7f7fu) + K) | d12);
                F = v & h & 0x80808080u;
            }
            const unsigned bal = __ballot_sync(0xffffffffu, F != 0);
            if (F) {
                const uint32_t nib = (((F >> 7) * 0x01020408u) >> 24) & 0xfu;  // bit j = pixel j of the group
                glist[nG + __popc(bal & ltMask)] = (uint16_t)((i << 4) | nib);
            }
            nG += __popc(bal);
        }
        __syncwarp();

        // ---- segments of <= 128 groups: expand -> exact arcs (phase A); then expand -> NMS -> output (phase B) ----
        const int nSeg = (nG + kSegGroups - 1) / kSegGroups;
        int nPix = 0;
        for (int phase = 0; phase < 2; ++phase) {
            for (int seg = 0; seg < nSeg; ++seg) {
                if (!(phase == 1 && nSeg == 1)) {  // with a single segment the pixel list of phase A is still valid
                    nPix = 0;
                    const int g1 = min(nG, (seg + 1) * kSegGroups);
                    for (int g0 = seg * kSegGroups; g0 < g1; g0 += 32) {
                        const int g = g0 + lane;
                        int nib = 0, y = 0, x0 = 0;
                        if (g < g1) {
                            const int e = glist[g];
                            const int i = e >> 4;
                            const int r = (int)(((unsigned)i * divW) >> 16);
                            y = r + 3;
                            x0 = (w0 + (i - r * nW)) << 2;
                            // drop the pixels of edge groups that lie outside the tested columns
                            const int lo = min(max(xs - x0, 0), 4), hi = min(max(xe - x0, 0), 4);
                            nib = (e & 0xf) & ((1 << hi) - 1) & ~((1 << lo) - 1);
                        }
                        const int n = __popc(nib);
                        int incl = n;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int v = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o) incl += v;
                        }
                        int pos = nPix + incl - n;
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if ((nib >> j) & 1) pbuf[pos++] = (uint16_t)((y << 7) | (x0 + j));
                        nPix += __shfl_sync(0xffffffffu, incl, 31);
                    }
                    __syncwarp();
                }
                if (phase == 0) {
                    // exact arc strength, two pixels per lane
                    for (int b = 0; b < nPix; b += 64) {
                        const int j0 = b + lane, j1 = j0 + 32;
                        const bool v0 = j0 < nPix, v1 = j1 < nPix;
                        const int p0 = pbuf[v0 ? j0 : 0], p1 = pbuf[v1 ? j1 : (v0 ? j0 : 0)];
                        const int o0 = (p0 >> 7) * PITCH + (p0 & 127), o1 = (p1 >> 7) * PITCH + (p1 & 127);
                        constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
                        constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
                        uint32_t q[16];
#pragma unroll
                        for (int k = 0; k < 16; ++k)
                            q[k] = (uint32_t)tile[o0 + dy[k] * PITCH + dx[k]] | ((uint32_t)tile[o1 + dy[k] * PITCH + dx[k]] << 16);
                        const uint32_t cen = (uint32_t)tile[o0] | ((uint32_t)tile[o1] << 16);
                        const uint32_t a = arcStrength2(q, cen);
                        if (v0) arc[o0] = (uint8_t)(a & 0xffu);
                        if (v1) arc[o1] = (uint8_t)(a >> 16);
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
                            const int o = (p >> 7) * PITCH + (p & 127);
                            a = arc[o];
                            if (a > t) {
                                const int m0 = __vimax3_s32(arc[o - PITCH - 1], arc[o - PITCH], arc[o - PITCH + 1]);
                                const int m1 = __vimax3_s32(arc[o + PITCH - 1], arc[o + PITCH], arc[o + PITCH + 1]);
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
        }
        __syncwarp();
    }
    if (lane == 0) *cntOut = kept;
}

struct FastGeom {
    int pitch, rowsMax, itemsMax, warpBytes;
};

FastGeom geomFor(const FramePlan& P, unsigned mask, int pitch) {
    FastGeom g{pitch, 7, 1, 0};
    for (int l = 0; l < P.nlevels; ++l) {
        if (!((mask >> l) & 1u)) continue;
        const LevelPlan& L = P.lv[l];
        const int ch = L.hCell + 6, cw = L.wCell + 6;
        g.rowsMax = ch > g.rowsMax ? ch : g.rowsMax;
        const int items = (ch - 6) * ((cw - 6 + 3) / 4 + 2);
        g.itemsMax = items > g.itemsMax ? items : g.itemsMax;
    }
    g.itemsMax = (g.itemsMax + 7) & ~7;
    g.warpBytes = 2 * g.rowsMax * pitch + 2 * g.itemsMax + 2 * kPixCap;
    g.warpBytes = (g.warpBytes + 15) & ~15;
    return g;
}

template <int PITCH>
void launchClass(const FramePlan& P, const Buffers& B, int F, unsigned mask, int cells, cudaStream_t s) {
    const FastGeom g = geomFor(P, mask, PITCH);
    const size_t smem = (size_t)g.warpBytes * kWarpsPerCta;
    if (smem > 48 * 1024) cudaFuncSetAttribute(fastCellsKernel<PITCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const dim3 grid((cells + kWarpsPerCta - 1) / kWarpsPerCta, F);
    fastCellsKernel<PITCH><<<grid, kWarpsPerCta * 32, smem, s>>>(P, B.pyr, B.cand, B.cellCnt, g.rowsMax, g.itemsMax, g.warpBytes, mask);
    countLaunch();
}

}  // namespace

void launchFast(const FramePlan& P, const Buffers& B, int F, cudaStream_t s) {
    // Two launches at most: levels whose cells fit the 64-byte tile row (shift <= 15, cell width wCell+6 <= 45, one
    // spare word to the right) and 50 rows — every level of a large image — and the few small levels with wider or
    // taller cells (wCell, hCell <= 70 always, since nCols = floor(width / 35); 15 + 76 + 4 <= 96).
    for (int wide = 0; wide < 2; ++wide) {
        unsigned mask = 0;
        int cells = 0;
        for (int l = 0; l < P.nlevels; ++l) {
            const LevelPlan& L = P.lv[l];
            if (L.nCols <= 0 || L.nRows <= 0) continue;
            if ((15 + L.wCell + 6 + 4 > 64 || L.hCell + 6 > 50) != (wide == 1)) continue;
            mask |= 1u << l;
            cells += L.nCols * L.nRows;
        }
        if (!cells) continue;
        if (!wide) launchClass<64>(P, B, F, mask, cells, s);
        else launchClass<96>(P, B, F, mask, cells, s);
    }
}

}  // namespace orbx

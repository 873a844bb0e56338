// k_fast.cu — K-b: FAST-9/16 over the reference's 35-px cell grid with the iniThFAST -> minThFAST fallback.
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (ORB_SLAM3/src/ORBextractor.cc:787-872) and the
// cv::FAST(cell, kps, th, true) calls inside it (:826, :845).  Semantics reproduced:
//   * a pixel is a corner at threshold t iff some 9 of the 16 ring pixels are all > I+t or all < I-t; with
//     arc(p) = max over the sixteen 9-arcs of max(min(I-ring), min(ring-I)) that is arc(p) > t, and FAST's
//     response is arc(p) - 1 — independent of t;
//   * FAST is called on the CELL SUB-IMAGE: only pixels >= 3 px inside the cell are tested and the 3x3 strict
//     non-max suppression sees everything outside that rectangle as 0;
//   * a cell that yields nothing at iniThFAST is redone at minThFAST;
//   * output order inside a cell is row-major; cells are visited row-major (:863-868).
//
// Two observations make the kernel branch-free and threshold-free until the very end:
//   (1) min over an arc of (I - ring) = I - max(ring): the arc strength needs only sliding max/min of the RAW ring
//       values, 80 three-input min/max per pixel, no subtraction and no corner pre-test;
//   (2) a corner at t survives NMS iff arc(p) > arc(q) for its 8 neighbours: neighbours that are not corners at t
//       have arc(q) <= t < arc(p), so thresholding them to 0 (what OpenCV does) changes nothing.  Hence the
//       local-maximum flag is computed once and both thresholds are just counts over it.
//
// sm_100a specifics: pixels are widened to 16 bits and processed two per thread with the native packed
// VIMNMX3.S16x2 / VIADD.16x2 instructions; the cell is kept in shared memory twice (pairs starting at even and at
// odd columns; row pitch = 2*words+1 so column 0 has an odd-aligned home) so every ring pair is ONE aligned 32-bit LDS.  One CTA per (cell, frame).
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kFT = 128;  // threads per CTA

// ring offsets (dx,dy), k = 0..15 (OpenCV features2d/fast.cpp makeOffsets, patternSize 16)
#define ORBX_RING_DX {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1}
#define ORBX_RING_DY {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3}

__device__ __forceinline__ uint32_t max3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t min3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_s16x2(a, b, c); }

// TW2 = 32-bit words per tile row in each copy (tile columns / 2)
template <int TW2>
__global__ void __launch_bounds__(kFT) fastCellsKernel(const __grid_constant__ FramePlan P,
                                                       const uint8_t* __restrict__ pyr, uint32_t* __restrict__ cand,
                                                       int* __restrict__ cellCnt, int tileRows, unsigned levelMask) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ int sWarpSum[kFT / 32];

    // ---- which cell -------------------------------------------------------------------------------------
    int c = blockIdx.x, l = 0;
    for (; l < P.nlevels; ++l) {  // blockIdx.x enumerates the cells of the levels selected by levelMask
        if (!((levelMask >> l) & 1u)) continue;
        const int nc = P.lv[l].nCols * P.lv[l].nRows;
        if (c < nc) break;
        c -= nc;
    }
    if (l >= P.nlevels) return;
    const LevelPlan& L = P.lv[l];
    const int ci = c / L.nCols, cj = c - ci * L.nCols;
    const int f = blockIdx.y, tid = threadIdx.x;
    int* cntOut = cellCnt + (size_t)f * P.nCellsTotal + L.cellBase + c;
    const int maxBX = L.w - kMinBorder, maxBY = L.h - kMinBorder;
    const int iniX = kMinBorder + cj * L.wCell, iniY = kMinBorder + ci * L.hCell;
    int maxX = iniX + L.wCell + 6, maxY = iniY + L.hCell + 6;
    if (maxX > maxBX) maxX = maxBX;
    if (maxY > maxBY) maxY = maxBY;
    const int cw = maxX - iniX, ch = maxY - iniY;
    // ORBextractor.cc:813-816, :820-823 skip rules; cv::FAST returns nothing below 7x7
    if (iniY >= maxBY - 3 || iniX >= maxBX - 6 || cw < 7 || ch < 7) {
        if (tid == 0) *cntOut = 0;
        return;
    }

    // Tile coordinates: column tc = x_cell + shift so that tile column 0 is a 4-byte aligned address in the plane.
    const int gx0 = kEdge + iniX;
    const int shift = gx0 & 3;
    uint32_t* tileE = smem;                       // [tileRows][TW2]  word i = columns (2i, 2i+1)
    uint32_t* tileO = tileE + tileRows * TW2;     // [tileRows][TW2]  word i = columns (2i-1, 2i)
    uint32_t* arcE = tileO + tileRows * TW2;      // same geometry: arc strength clamped to >= 0, 0 outside the tested rect
    uint32_t* arcO = arcE + tileRows * TW2;
    uint32_t* flags = tileE;                      // reused after the arc pass: local-maximum arcs per pair

    // ---- stage the cell, widening bytes to 16 bits, in both pair alignments ----------------------------------
    {
        const uint8_t* src = pyr + (size_t)f * P.frameBytes + L.planeOff + (size_t)(kEdge + iniY) * L.pitch + (gx0 - shift);
        const int nW = (shift + cw + 3) >> 2;     // aligned words per row
        const float invW = 1.0f / (float)nW;
        for (int i = tid; i < ch * nW; i += kFT) {
            const int r = (int)(((float)i + 0.5f) * invW), wi = i - r * nW;
            const uint32_t* g = reinterpret_cast<const uint32_t*>(src + (size_t)r * L.pitch) + wi;
            const uint32_t w0 = __ldg(g), w1 = __ldg(g + 1);
            uint2 e, o;
            e.x = __byte_perm(w0, 0, 0x4140);                        // (p0, p1)
            e.y = __byte_perm(w0, 0, 0x4342);                        // (p2, p3)
            o.x = __byte_perm(w0, 0, 0x4241);                        // (p1, p2)
            o.y = __byte_perm(w0, w1, 0x7473) & 0x00ff00ffu;         // (p3, p4)
            *reinterpret_cast<uint2*>(tileE + r * TW2 + 2 * wi) = e;
            tileO[r * TW2 + 2 * wi + 1] = o.x;
            tileO[r * TW2 + 2 * wi + 2] = o.y;
            if (wi == 0) tileO[r * TW2] = __byte_perm(w0, 0, 0x4044);  // (-, p0): column 0 lives in the high half
        }
    }
    __syncthreads();

    // ---- pass 1: arc strength of every tested pixel, two pixels per thread ---------------------------------------
    // Pair grid over the tested rectangle plus a 1-px apron (which is written as 0 for the NMS).
    const int ec0 = (2 + shift) & ~1;                               // first pair column (even)
    const int PW = ((((cw - 3 + shift) & ~1) - ec0) >> 1) + 1;      // pair columns
    const int PR = ch - 4;                                          // rows ty = 2 .. ch-3
    const float invPW = 1.0f / (float)PW;
    {
        constexpr int dx[16] = ORBX_RING_DX, dy[16] = ORBX_RING_DY;
        for (int i = tid; i < PR * PW; i += kFT) {
            const int r = (int)(((float)i + 0.5f) * invPW), pc = i - r * PW;
            const int ty = r + 2, tc = ec0 + 2 * pc;
            const int x0 = tc - shift;                               // cell x of lane 0
            const bool rowOk = ty >= 3 && ty < ch - 3;
            const bool ok0 = rowOk && x0 >= 3 && x0 < cw - 3, ok1 = rowOk && x0 + 1 >= 3 && x0 + 1 < cw - 3;
            const int b = ty * TW2 + (tc >> 1);
            uint32_t arc = 0;
            if (ok0 || ok1) {
                uint32_t q[16];
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    q[k] = (dx[k] & 1) ? tileO[b + dy[k] * TW2 + ((dx[k] + 1) >> 1)] : tileE[b + dy[k] * TW2 + (dx[k] >> 1)];
                const uint32_t cen = tileE[b];
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
                arc = __vimax3_s16x2_relu(__vsub2(cen, A), __vsub2(B, cen), 0u);
                arc &= (ok0 ? 0x0000ffffu : 0u) | (ok1 ? 0xffff0000u : 0u);
            }
            arcE[b] = arc;
            // odd-aligned copy: column tc is the high half of O word tc/2, column tc+1 the low half of O word tc/2+1
            reinterpret_cast<uint16_t*>(arcO)[2 * b + 1] = (uint16_t)(arc & 0xffffu);
            reinterpret_cast<uint16_t*>(arcO)[2 * b + 2] = (uint16_t)(arc >> 16);
        }
    }
    __syncthreads();

    // ---- pass 2: strict 3x3 local maximum (threshold independent), kept arcs into `flags` ------------------------
    const int NR = ch - 6;  // tested rows
    const int nF = NR * PW;
    for (int i = tid; i < nF; i += kFT) {
        const int r = (int)(((float)i + 0.5f) * invPW), pc = i - r * PW;
        const int b = (r + 3) * TW2 + ((ec0 >> 1) + pc);
        const uint32_t up = max3(arcO[b - TW2], arcE[b - TW2], arcO[b - TW2 + 1]);
        const uint32_t dn = max3(arcO[b + TW2], arcE[b + TW2], arcO[b + TW2 + 1]);
        const uint32_t nm = max3(up, dn, __vmaxs2(arcO[b], arcO[b + 1]));
        const uint32_t cen = arcE[b];
        flags[i] = cen & __vcmpgts2(cen, nm);  // arcs are >= 0 and invalid lanes are 0, so they never pass
    }
    __syncthreads();

    // ---- pass 3: count at both thresholds, ordered compaction (row-major), fallback to minThFAST -----------------
    const int per = (nF + kFT - 1) / kFT;
    const int i0 = min(tid * per, nF), i1 = min(i0 + per, nF);
    // response = arc - 1 must beat the zero score of non-corners: arc >= 2 (only matters for thresholds < 1)
    const int tIni = max(P.iniTh, 1), tMin = max(P.minTh, 1);
    int cnt = 0;  // (count at ini) << 16 | (count at min)
    for (int i = i0; i < i1; ++i) {
        const uint32_t v = flags[i];
        const int a0 = v & 0xffff, a1 = v >> 16;
        cnt += ((a0 > tIni) << 16) + (a0 > tMin) + ((a1 > tIni) << 16) + (a1 > tMin);
    }
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += n;
    }
    if ((tid & 31) == 31) sWarpSum[tid >> 5] = incl;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kFT / 32; ++w) {
        if (w < (tid >> 5)) base += sWarpSum[w];
        total += sWarpSum[w];
    }
    const bool useIni = (total >> 16) > 0;  // ORBextractor.cc:843-846: redo at minThFAST only if nothing was found
    const int t = useIni ? tIni : tMin;
    const int excl = base + incl - cnt;
    int o = useIni ? (excl >> 16) : (excl & 0xffff);
    uint32_t* slot = cand + (size_t)f * P.candWords + L.candOff + (size_t)c * L.cellCap;
    const int xOff = cj * L.wCell - shift + ec0, yOff = ci * L.hCell + 3;
    for (int i = i0; i < i1; ++i) {
        const uint32_t v = flags[i];
        const int a0 = v & 0xffff, a1 = v >> 16;
        if (a0 > t || a1 > t) {
            const int r = (int)(((float)i + 0.5f) * invPW), pc = i - r * PW;
            const uint32_t xy = ((uint32_t)(r + yOff) << 12) | (uint32_t)(xOff + 2 * pc);
            if (a0 > t) { if (o < L.cellCap) slot[o] = ((uint32_t)(a0 - 1) << 24) | xy; ++o; }
            if (a1 > t) { if (o < L.cellCap) slot[o] = ((uint32_t)(a1 - 1) << 24) | (xy + 1); ++o; }
        }
    }
    if (tid == 0) *cntOut = useIni ? (total >> 16) : (total & 0xffff);
}

}  // namespace

void launchFast(const FramePlan& P, const Buffers& B, int F, cudaStream_t s) {
    // Two launches at most: levels whose cells fit the narrow tile (wCell + 9 <= 48 columns: every level of a large
    // image) and the few small levels with wider cells (wCell < 70 always, since nCols = floor(width / 35)).
    for (int wide = 0; wide < 2; ++wide) {
        unsigned mask = 0;
        int cells = 0, maxH = 0;
        for (int l = 0; l < P.nlevels; ++l) {
            const LevelPlan& L = P.lv[l];
            if (L.nCols <= 0 || L.nRows <= 0) continue;
            if ((L.wCell + 9 > 48) != (wide == 1)) continue;
            mask |= 1u << l;
            cells += L.nCols * L.nRows;
            maxH = L.hCell > maxH ? L.hCell : maxH;
        }
        if (!cells) continue;
        const int tileRows = maxH + 6 + 1;
        const dim3 grid(cells, F);
        if (!wide) {
            const size_t smem = (size_t)4 * tileRows * 26 * 4;
            if (smem > 48 * 1024) cudaFuncSetAttribute(fastCellsKernel<26>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            fastCellsKernel<26><<<grid, kFT, smem, s>>>(P, B.pyr, B.cand, B.cellCnt, tileRows, mask);
        } else {
            const size_t smem = (size_t)4 * tileRows * 46 * 4;
            if (smem > 48 * 1024) cudaFuncSetAttribute(fastCellsKernel<46>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            fastCellsKernel<46><<<grid, kFT, smem, s>>>(P, B.pyr, B.cand, B.cellCnt, tileRows, mask);
        }
        countLaunch();
    }
}

}  // namespace orbx

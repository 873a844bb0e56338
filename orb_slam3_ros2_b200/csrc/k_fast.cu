// k_fast.cu — K-b: FAST-9/16 over the reference's 35-px cell grid with the iniThFAST -> minThFAST fallback.
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (ORB_SLAM3/src/ORBextractor.cc:787-872) and the
// cv::FAST(cell, kps, th, true) calls inside it (:826, :845).  Semantics reproduced:
//   * a pixel is a corner at threshold t iff some 9 of the 16 ring pixels are all > I+t or all < I-t;
//     its score is (max over the sixteen 9-arcs of min|d|) - 1 and does not depend on t;
//   * FAST is called on the CELL SUB-IMAGE: only pixels >= 3 px inside the cell are tested and the 3x3 strict
//     non-max suppression sees scores outside that rectangle (and of non-corners at t) as 0;
//   * a cell that yields nothing at iniThFAST is redone at minThFAST;
//   * output order inside a cell is row-major; cells are visited row-major (:863-868).
//
// One CTA per (cell, frame).  The cell (+3-px ring apron) is staged in shared memory, every lane classifies its
// pixels branch-free with two 16-bit ring masks (brighter / darker than the centre by more than minThFAST) and a
// shift-and trick for "9 consecutive bits on a 16-cycle"; pixels that pass are queued so the expensive arc
// strength is computed by all lanes on a dense work list.  The survivors of the NMS are compacted in row-major
// order with a block scan and written as packed words (resp<<24 | y<<12 | x) into the cell's slot.
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kThreads = 128;

// ring offsets (dx,dy), k = 0..15 (OpenCV features2d/fast.cpp makeOffsets, patternSize 16)
__constant__ int8_t cRingX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
__constant__ int8_t cRingY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

__device__ __forceinline__ bool hasArc9(uint32_t m16) {
    const uint32_t x = m16 | (m16 << 16);
    uint32_t r = x & (x >> 1);
    r &= r >> 2;
    r &= r >> 4;
    r &= x >> 8;
    return (r & 0xffffu) != 0;
}

__global__ void __launch_bounds__(kThreads) fastCellsKernel(const __grid_constant__ FramePlan P,
                                                            const uint8_t* __restrict__ pyr, uint32_t* __restrict__ cand,
                                                            int* __restrict__ cellCnt, int tilePitch, int arcPitchMax) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int sQueueN;
    __shared__ int sWarpSum[kThreads / 32];

    // ---- which cell -------------------------------------------------------------------------------------
    int c = blockIdx.x, l = 0;
    for (; l < P.nlevels; ++l) {
        const int nc = P.lv[l].nCols * P.lv[l].nRows;
        if (c < nc) break;
        c -= nc;
    }
    if (l >= P.nlevels) return;
    const LevelPlan& L = P.lv[l];
    const int ci = c / L.nCols, cj = c % L.nCols;
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
        if (threadIdx.x == 0) *cntOut = 0;
        return;
    }
    const int tw = cw - 6, th = ch - 6;  // tested rectangle
    const int npx = tw * th;
    const int arcPitch = tw + 2;  // 1-px zero apron for the NMS

    uint8_t* sTile = smem;                                            // ch x tilePitch
    uint8_t* sArc = sTile + (size_t)tilePitch * (L.hCell + 6);        // (th+2) x arcPitch
    uint16_t* sQueue = reinterpret_cast<uint16_t*>(sArc + ((arcPitchMax + 3) & ~3));  // npx entries max
    const int tid = threadIdx.x;

    // ---- stage the cell ----------------------------------------------------------------------------------
    const uint8_t* src = pyr + (size_t)f * P.frameBytes + L.planeOff + (size_t)(kEdge + iniY) * L.pitch + kEdge + iniX;
    for (int i = tid; i < cw * ch; i += kThreads) {
        const int y = i / cw, x = i - y * cw;
        sTile[y * tilePitch + x] = __ldg(src + (size_t)y * L.pitch + x);
    }
    for (int i = tid; i < (th + 2) * arcPitch; i += kThreads) sArc[i] = 0;
    if (tid == 0) sQueueN = 0;
    __syncthreads();

    int ofs[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) ofs[k] = cRingY[k] * tilePitch + cRingX[k];

    // ---- pass 1: branch-free corner test at minThFAST, queue the survivors ---------------------------------
    const int tmin = P.minTh < P.iniTh ? P.minTh : P.iniTh;
    for (int e = tid; e < npx; e += kThreads) {
        const int y = e / tw, x = e - y * tw;
        const uint8_t* p = sTile + (y + 3) * tilePitch + x + 3;
        const int v = p[0], hi = v + tmin, lo = v - tmin;
        uint32_t mb = 0, md = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const int q = p[ofs[k]];
            mb |= (uint32_t)(q > hi) << k;
            md |= (uint32_t)(q < lo) << k;
        }
        if (hasArc9(mb) || hasArc9(md)) {
            const int pos = atomicAdd(&sQueueN, 1);
            sQueue[pos] = (uint16_t)e;
        }
    }
    __syncthreads();

    // ---- pass 2: arc strength of the queued pixels ----------------------------------------------------------
    const int qn = sQueueN;
    for (int i = tid; i < qn; i += kThreads) {
        const int e = sQueue[i];
        const int y = e / tw, x = e - y * tw;
        const uint8_t* p = sTile + (y + 3) * tilePitch + x + 3;
        const int v = p[0];
        int d[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) d[k] = v - (int)p[ofs[k]];
        // min / max over every cyclic window of 9 by doubling: w2 -> w4 -> w8 -> w9
        int mn2[16], mx2[16], mn4[16], mx4[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) { mn2[k] = min(d[k], d[(k + 1) & 15]); mx2[k] = max(d[k], d[(k + 1) & 15]); }
#pragma unroll
        for (int k = 0; k < 16; ++k) { mn4[k] = min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = max(mx2[k], mx2[(k + 2) & 15]); }
        // strongest "centre brighter" arc = max_k min9(d); strongest "centre darker" arc = -min_k max9(d).
        // NOTE: written without `max(best, max(mn9, -mx9))` on purpose — ptxas 12.9 (sm_100a) fuses that into a
        // VIMNMX3 and drops the negation of the third operand (seen in SASS and in results on a B200).
        int bestMin = -255, bestMax = 255;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const int mn9 = min(min(mn4[k], mn4[(k + 4) & 15]), d[(k + 8) & 15]);
            const int mx9 = max(max(mx4[k], mx4[(k + 4) & 15]), d[(k + 8) & 15]);
            bestMin = max(bestMin, mn9);
            bestMax = min(bestMax, mx9);
        }
        const int best = (bestMin + bestMax > 0) ? bestMin : 0 - bestMax;  // max(bestMin, -bestMax) as a select
        sArc[(y + 1) * arcPitch + x + 1] = (uint8_t)best;  // > tmin >= 0 here
    }
    __syncthreads();

    // ---- pass 3: NMS at iniThFAST, then (if the cell is empty) at minThFAST; ordered compaction ---------------
    const int per = (npx + kThreads - 1) / kThreads;  // <= 64 (cells are < 76 px wide)
    const int e0 = tid * per, e1 = min(e0 + per, npx);
    uint32_t* slot = cand + (size_t)f * P.candWords + L.candOff + (size_t)c * L.cellCap;
    const int xOff = cj * L.wCell, yOff = ci * L.hCell;
    for (int pass = 0; pass < 2; ++pass) {
        const int t = pass == 0 ? P.iniTh : P.minTh;
        unsigned long long keep = 0;
        for (int e = e0; e < e1; ++e) {
            const int y = e / tw, x = e - y * tw;
            const uint8_t* a = sArc + (y + 1) * arcPitch + x + 1;
            const int ap = a[0];
            if (ap <= t) continue;
            const int s = ap - 1;
            bool k = true;
#pragma unroll
            for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
                for (int dx = -1; dx <= 1; ++dx) {
                    if (dx == 0 && dy == 0) continue;
                    const int aq = a[dy * arcPitch + dx];
                    const int cq = aq > t ? aq - 1 : 0;
                    k = k && (s > cq);
                }
            if (k) keep |= 1ull << (e - e0);
        }
        // block exclusive scan of popcounts
        const int mine = __popcll(keep);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, incl, o);
            if ((tid & 31) >= o) incl += n;
        }
        if ((tid & 31) == 31) sWarpSum[tid >> 5] = incl;
        __syncthreads();
        int base = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) {
            if (w < (tid >> 5)) base += sWarpSum[w];
            total += sWarpSum[w];
        }
        __syncthreads();
        if (total > 0 || pass == 1) {
            int o = base + incl - mine;
            while (keep) {
                const int b = __ffsll((long long)keep) - 1;
                keep &= keep - 1;
                const int e = e0 + b;
                const int y = e / tw, x = e - y * tw;
                const int resp = (int)sArc[(y + 1) * arcPitch + x + 1] - 1;
                if (o < L.cellCap) slot[o] = ((uint32_t)resp << 24) | ((uint32_t)(y + 3 + yOff) << 12) | (uint32_t)(x + 3 + xOff);
                ++o;
            }
            if (tid == 0) *cntOut = total;
            break;
        }
    }
}

}  // namespace

void launchFast(const FramePlan& P, const Buffers& B, int F, cudaStream_t s) {
    int maxW = 0, maxH = 0;
    for (int l = 0; l < P.nlevels; ++l)
        if (P.lv[l].nCols > 0 && P.lv[l].nRows > 0) {
            maxW = P.lv[l].wCell > maxW ? P.lv[l].wCell : maxW;
            maxH = P.lv[l].hCell > maxH ? P.lv[l].hCell : maxH;
        }
    if (P.nCellsTotal == 0) return;
    const int tilePitch = (maxW + 6 + 3) & ~3;
    const int arcBytes = (maxW + 2) * (maxH + 2);
    const size_t smem = (size_t)tilePitch * (maxH + 6) + ((arcBytes + 3) & ~3) + 2 * (size_t)maxW * maxH + 16;
    if (smem > 48 * 1024) cudaFuncSetAttribute(fastCellsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    fastCellsKernel<<<dim3(P.nCellsTotal, F), kThreads, smem, s>>>(P, B.pyr, B.cand, B.cellCnt, tilePitch, arcBytes);
    countLaunch();
}

}  // namespace orbx

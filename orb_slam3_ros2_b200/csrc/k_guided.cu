// k_guided.cu — the matcher-side rows next to the extractor:
//   * acceptKernel          the accept rules of ORBmatcher on top-2 results (ORBmatcher.cc:123-128, :327-329, :697-699,
//                           :848-850, :1767, Frame.cc:1149), float comparisons written as the reference writes them;
//   * gridSortKernel +      Frame::AssignFeaturesToGrid / PosInGrid (Frame.cc:385-416, :725-736) as ONE sort of
//     gridSearchKernel      (cell << 20 | index) keys: with cells numbered column-major (ix * 48 + iy) the candidates of
//                           one grid column of a GetFeaturesInArea window (Frame.cc:657-723) are a contiguous key range
//                           already in the reference's visiting order, so a warp walks a query's window column by
//                           column, 32 candidates at a time, and keeps (distance, visit position) top-2 keys;
//   * stereoRefineKernel +  the second half of Frame::ComputeStereoMatches (Frame.cc:896-980): 11 x 11 L1 correlation
//     stereoCutKernel       over +-5 px read straight from the two extractors' device-resident pyramids, parabola fit,
//                           disparity / depth, and the 1.5 * 1.4 * median cut through a histogram select.
//   * undistortKernel       Frame::UndistortKeyPoints / ComputeImageBounds (Frame.cc:747-809): cv::undistortPoints for a
//                           pinhole camera, one thread per keypoint, the 5 fixed iterations in round-to-nearest doubles
//                           without contraction (cv2 4.13.0's scalar path, bit for bit).
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kGridRows = 48, kGridCols = 64, kGridCells = kGridRows * kGridCols;  // Frame.h:44-45
constexpr unsigned kNoCell = 0xfffu;

__global__ void acceptKernel(int rule, float nnratio, const int* __restrict__ d1, const int* __restrict__ d2,
                             const long long* __restrict__ idx2, const int* __restrict__ l1, const int* __restrict__ l2, int n,
                             uint8_t* __restrict__ accept) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b1 = d1[i], b2 = d2[i];
    const float f1 = (float)b1, f2 = (float)b2;
    const float lim = __fmul_rn(nnratio, f2);
    bool ok = false;
    switch (rule) {
        case 0: {
            const bool same = (l1 ? l1[i] : 0) == (l2 ? l2[i] : 0);
            ok = b1 <= 100 && (!same || f1 <= lim);
            break;
        }
        case 1: ok = b1 <= 50 && f1 < lim; break;
        case 2: ok = b1 <= 50 && f1 < lim; break;
        case 3: ok = b1 < 50 && f1 < lim; break;
        case 4: ok = b1 <= 100; break;
        case 5: ok = b1 <= 50; break;
        case 6: ok = (!idx2 || idx2[i] >= 0) && (double)f1 < __dmul_rn((double)f2, 0.7); break;
        case 7: ok = f1 <= __fmul_rn(50.f, nnratio); break;  // int * float -> float, then (float)bestDist <= that (:523, :636)
    }
    accept[i] = ok ? 1 : 0;
}

// ---- grid: keys (cell << 20 | index) sorted by one CTA (bitonic, shared memory), then the cell boundaries -------------
// batch form (nPerFrame != NULL): CTA f sorts frame f of a [frame][cap] keypoint array into sortedOut + f * cap,
// cellStart + f * (kGridCells + 1)
__global__ void __launch_bounds__(1024) gridSortKernel(const float* __restrict__ kps, int nt, int nPow2, float minX, float minY,
                                                       float invW, float invH, uint32_t* __restrict__ sortedOut,
                                                       int* __restrict__ cellStart, const int* __restrict__ nPerFrame, int cap) {
    extern __shared__ uint32_t keys[];
    const int tid = threadIdx.x;
    if (nPerFrame) {
        const int f = blockIdx.x;
        nt = min(nPerFrame[f], cap);
        kps += (size_t)f * cap * 7;
        sortedOut += (size_t)f * cap;
        cellStart += (size_t)f * (kGridCells + 1);
    }
    for (int i = tid; i < nPow2; i += 1024) {
        uint32_t key = 0xffffffffu;
        if (i < nt) {
            // posX = round((kp.pt.x - mnMinX) * mfGridElementWidthInv)  (Frame.cc:727-728; float round, half away from 0)
            const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i * 7], minX), invW));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i * 7 + 1], minY), invH));
            const bool in = px >= 0 && px < kGridCols && py >= 0 && py < kGridRows;
            key = ((in ? (uint32_t)(px * kGridRows + py) : kNoCell) << 20) | (uint32_t)i;
        }
        keys[i] = key;
    }
    __syncthreads();
    for (int k = 2; k <= nPow2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < nPow2; i += 1024) {
                const int p = i ^ j;
                if (p > i) {
                    const uint32_t a = keys[i], b = keys[p];
                    if (((i & k) == 0) ? (a > b) : (a < b)) { keys[i] = b; keys[p] = a; }
                }
            }
            __syncthreads();
        }
    // cellStart[c] = first position whose cell >= c (cells 0..kGridCells)
    for (int i = tid; i < nt; i += 1024) {
        sortedOut[i] = keys[i];
        const int c = (int)(keys[i] >> 20);
        const int prev = i == 0 ? -1 : (int)(keys[i - 1] >> 20);
        for (int cc = prev + 1; cc <= min(c, kGridCells); ++cc) cellStart[cc] = i;
    }
    {
        const int last = nt > 0 ? (int)(keys[nt - 1] >> 20) : -1;
        for (int cc = last + 1 + tid; cc <= kGridCells; cc += 1024) cellStart[cc] = nt;
    }
}

struct GridQuery {
    float u, v, r;
    int minLevel, maxLevel;
    float ur;
};

__device__ __forceinline__ void top2Insert(unsigned long long& k1, unsigned long long& k2, unsigned long long k) {
    k2 = min(k2, max(k1, k));
    k1 = min(k1, k);
}

// one warp per query
__global__ void __launch_bounds__(256) gridSearchKernel(const float* __restrict__ kps, const uint4* __restrict__ desc,
                                                        const uint32_t* __restrict__ sorted, const int* __restrict__ cellStart,
                                                        const uint8_t* __restrict__ skip, const float* __restrict__ uRight,
                                                        float minX, float minY, float invW, float invH,
                                                        const GridQuery* __restrict__ qs, const uint4* __restrict__ descQ, int nq,
                                                        int* __restrict__ idx1, int* __restrict__ d1, int* __restrict__ idx2,
                                                        int* __restrict__ d2, int* __restrict__ lvl1, int* __restrict__ lvl2) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= nq) return;
    const GridQuery Q = qs[q];
    const uint4 a0 = descQ[2 * q], a1 = descQ[2 * q + 1];
    unsigned long long k1 = ~0ull, k2 = ~0ull;  // distance << 48 | visit position << 24 | index
    // Frame.cc:665-687
    const int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(Q.u, minX), Q.r), invW)));
    const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(Q.u, minX), Q.r), invW)));
    const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(Q.v, minY), Q.r), invH)));
    const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(Q.v, minY), Q.r), invH)));
    if (!(minCX >= kGridCols || maxCX < 0 || minCY >= kGridRows || maxCY < 0)) {
        const bool checkLevels = Q.minLevel > 0 || Q.maxLevel >= 0;
        unsigned visit = 0;
        for (int ix = minCX; ix <= maxCX; ++ix) {
            if (minCY > maxCY) break;
            const int b = cellStart[ix * kGridRows + minCY], e = cellStart[ix * kGridRows + maxCY + 1];
            for (int p0 = b; p0 < e; p0 += 32) {
                const int p = p0 + lane;
                if (p < e) {
                    const int idx = (int)(sorted[p] & 0xfffffu);
                    const float x = kps[idx * 7], y = kps[idx * 7 + 1];
                    const int oct = __float_as_int(kps[idx * 7 + 5]);
                    bool ok = true;
                    if (checkLevels) ok = !(oct < Q.minLevel) && !(Q.maxLevel >= 0 && oct > Q.maxLevel);
                    ok = ok && fabsf(__fsub_rn(x, Q.u)) < Q.r && fabsf(__fsub_rn(y, Q.v)) < Q.r;
                    if (ok && skip) ok = skip[idx] == 0;
                    if (ok && uRight) {
                        const float urt = uRight[idx];
                        if (urt > 0.f) ok = !(fabsf(__fsub_rn(Q.ur, urt)) > Q.r);
                    }
                    if (ok) {
                        const uint4 b0 = desc[2 * idx], b1 = desc[2 * idx + 1];
                        const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                                      __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                        const unsigned long long key = ((unsigned long long)d << 48) |
                                                       ((unsigned long long)(visit + (unsigned)(p - b)) << 24) | (unsigned)idx;
                        top2Insert(k1, k2, key);
                    }
                }
            }
            visit += (unsigned)(e - b);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
        top2Insert(k1, k2, o1);
        top2Insert(k1, k2, o2);
    }
    if (lane == 0) {
        const bool h1 = k1 != ~0ull, h2 = k2 != ~0ull;
        const int i1 = h1 ? (int)(k1 & 0xffffffu) : -1, i2 = h2 ? (int)(k2 & 0xffffffu) : -1;
        idx1[q] = i1; d1[q] = h1 ? (int)(k1 >> 48) : 256;
        idx2[q] = i2; d2[q] = h2 ? (int)(k2 >> 48) : 256;
        if (lvl1) lvl1[q] = h1 ? __float_as_int(kps[i1 * 7 + 5]) : -1;
        if (lvl2) lvl2[q] = h2 ? __float_as_int(kps[i2 * 7 + 5]) : -1;
    }
}

// ---- ComputeStereoMatches, Frame.cc:896-966: one warp per left keypoint ------------------------------------------------
struct StereoLevels {
    unsigned planeOff[kMaxLevels];
    int pitch[kMaxLevels];
    int cols[kMaxLevels];
    float sf[kMaxLevels], isf[kMaxLevels];
};

__global__ void __launch_bounds__(256) stereoRefineKernel(const __grid_constant__ StereoLevels LV, const uint8_t* __restrict__ pyrL,
                                                          const uint8_t* __restrict__ pyrR, const float* __restrict__ kl,
                                                          const float* __restrict__ kr, int nl, const int* __restrict__ bestIdx,
                                                          const int* __restrict__ bestDist, float mbf, float maxD,
                                                          float* __restrict__ uRight, float* __restrict__ depth,
                                                          int* __restrict__ sad, const StereoBatch SB) {
    if (SB.nl) {  // pair blockIdx.y of a batch
        const size_t o = (size_t)blockIdx.y * SB.cap;
        nl = min(SB.nl[blockIdx.y], SB.cap);
        kl += o * 7; kr += o * 7; bestIdx += o; bestDist += o; uRight += o; depth += o; sad += o;
        pyrL += (size_t)blockIdx.y * SB.pyrStrideL; pyrR += (size_t)blockIdx.y * SB.pyrStrideR;
    }
    const int iL = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (iL >= nl) return;
    float outU = -1.0f, outD = -1.0f;
    int outS = -1;
    const int bd = bestDist[iL];
    if (bd < 75) {  // thOrbDist = (TH_HIGH + TH_LOW) / 2
        const float uL = kl[iL * 7], vL = kl[iL * 7 + 1];
        const int oct = __float_as_int(kl[iL * 7 + 5]);
        const float uR0 = kr[bestIdx[iL] * 7];
        const float scaleFactor = LV.isf[oct];
        const float scaleduL = roundf(__fmul_rn(uL, scaleFactor));
        const float scaledvL = roundf(__fmul_rn(vL, scaleFactor));
        const float scaleduR0 = roundf(__fmul_rn(uR0, scaleFactor));
        const float iniu = scaleduR0, endu = __fadd_rn(scaleduR0, 11.0f);  // scaleduR0 + L - w, + L + w + 1 (small integers)
        if (!(iniu < 0 || endu >= (float)LV.cols[oct])) {
            const int pitch = LV.pitch[oct];
            // level pixel (x, y) lives at plane + (y + 19) * pitch + x + 19
            const uint8_t* pl = pyrL + LV.planeOff[oct] + (size_t)((int)scaledvL - 5 + kEdge) * pitch + ((int)scaleduL - 5 + kEdge);
            const uint8_t* pr = pyrR + LV.planeOff[oct] + (size_t)((int)scaledvL - 5 + kEdge) * pitch + ((int)scaleduR0 - 5 + kEdge);
            int acc[11];
#pragma unroll
            for (int k = 0; k < 11; ++k) acc[k] = 0;
            // lanes 0..10 own one row of the window: 11 left pixels against 21 right pixels
            if (lane < 11) {
                int a[11], b[21];
#pragma unroll
                for (int x = 0; x < 11; ++x) a[x] = pl[(size_t)lane * pitch + x];
#pragma unroll
                for (int x = 0; x < 21; ++x) b[x] = pr[(size_t)lane * pitch + x - 5];
#pragma unroll
                for (int k = 0; k < 11; ++k)
#pragma unroll
                    for (int x = 0; x < 11; ++x) acc[k] += abs(a[x] - b[x + k]);
            }
#pragma unroll
            for (int k = 0; k < 11; ++k)
#pragma unroll
                for (int o = 8; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
            // (lanes 0..15 now hold the sums of rows 0..15; rows 11..15 contributed 0)
            int best = 0x7fffffff, bestInc = 0;
#pragma unroll
            for (int k = 0; k < 11; ++k)
                if (acc[k] < best) { best = acc[k]; bestInc = k - 5; }
            if (!(bestInc == -5 || bestInc == 5)) {
                float dist1 = 0.f, dist2 = 0.f, dist3 = 0.f;
#pragma unroll
                for (int k = 1; k < 10; ++k)
                    if (k - 5 == bestInc) { dist1 = (float)acc[k - 1]; dist2 = (float)acc[k]; dist3 = (float)acc[k + 1]; }
                const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
                const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);
                if (!(deltaR < -1.f || deltaR > 1.f)) {
                    float bestuR = __fmul_rn(LV.sf[oct], __fadd_rn(__fadd_rn(scaleduR0, (float)bestInc), deltaR));
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= 0.f && disparity < maxD) {
                        if (disparity <= 0.f) {
                            disparity = 0.01f;                                  // disparity = 0.01 (double -> float)
                            bestuR = (float)__dsub_rn((double)uL, 0.01);        // bestuR = uL - 0.01
                        }
                        outD = __fdiv_rn(mbf, disparity);
                        outU = bestuR;
                        outS = best;
                    }
                }
            }
        }
    }
    if (lane == 0) { uRight[iL] = outU; depth[iL] = outD; sad[iL] = outS; }
}

// Frame.cc:969-980: median of the correlation distances (element size/2 of the sorted list) through a histogram
// (distances are <= 121 * 255 = 30855), then every match with (float)dist >= 1.5f * 1.4f * median is removed.
constexpr int kSadBins = 30856;
__global__ void __launch_bounds__(1024) stereoCutKernel(const int* __restrict__ sad, int nl, float* __restrict__ uRight,
                                                        float* __restrict__ depth, const StereoBatch SB) {
    if (SB.nl) {  // pair blockIdx.x of a batch
        const size_t o = (size_t)blockIdx.x * SB.cap;
        nl = min(SB.nl[blockIdx.x], SB.cap);
        sad += o; uRight += o; depth += o;
    }
    extern __shared__ int hist[];
    __shared__ int sPart[1024];
    __shared__ int sMedian;
    const int tid = threadIdx.x;
    for (int i = tid; i < kSadBins; i += 1024) hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < nl; i += 1024)
        if (sad[i] >= 0) atomicAdd(&hist[min(sad[i], kSadBins - 1)], 1);
    __syncthreads();
    const int per = (kSadBins + 1023) / 1024;
    const int b = tid * per, e = min(b + per, kSadBins);
    int sum = 0;
    for (int i = b; i < e; ++i) sum += hist[i];
    sPart[tid] = sum;
    __syncthreads();
    if (tid == 0) {
        int total = 0;
        for (int i = 0; i < 1024; ++i) total += sPart[i];
        int median = -1;
        if (total > 0) {
            const int rank = total / 2;  // vDistIdx[size / 2]
            int acc = 0, t = 0;
            for (; t < 1024; ++t) {
                if (acc + sPart[t] > rank) break;
                acc += sPart[t];
            }
            for (int i = t * per; i < kSadBins; ++i) {
                acc += hist[i];
                if (acc > rank) { median = i; break; }
            }
        }
        sMedian = median;
    }
    __syncthreads();
    if (sMedian < 0) return;
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)sMedian);
    for (int i = tid; i < nl; i += 1024)
        if (sad[i] >= 0 && !((float)sad[i] < thDist)) { uRight[i] = -1.0f; depth[i] = -1.0f; }
}

// ---- ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763): one warp per frame pair ------------------------------
// The loop over i1 is sequential in the reference (vMatchedDistance / vnMatches21 of earlier keypoints gate later ones), so
// one warp walks it in order; inside an i1 the window's candidates (Frame::GetFeaturesInArea order = grid columns, each a
// contiguous range of the sorted keys) are flattened and taken 32 at a time.
struct InitArgs {
    const float* kps1; const uint4* desc1; const float* kps2; const uint4* desc2;   // pair p: + p * stride keypoints
    const uint32_t* sorted2; const int* cellStart2;                                   // F2's grid: + p * stride, + p * (cells + 1)
    const int* n1Arr; const int* n2Arr; int n1, n2, stride;
    float minX, minY, invW, invH, r, nnratio;
    int checkOri, prevFromKps;
    float* prev;       // [pair][stride][2], in / out (vbPrevMatched)
    int* m12;          // [pair][stride] out (vnMatches12)
    int* matchedDist;  // scratch [pair][stride] (vMatchedDistance)
    int* m21;          // scratch [pair][stride] (vnMatches21)
    int* binOf;        // scratch [pair][stride] rotation bin of i1's push, -1 = none
    int* nMatches;     // [pair]
    // candidate lists of phase A (initCandKernel): per (pair, i1) the window's candidates in visiting order as dist << 20 | index,
    // kInitListCap entries at most; candCount > kInitListCap = the list overflowed and the walk is redone here
    uint32_t* candList; int* candCount;
};
constexpr int kInitListCap = kSearchInitListCap;

// Phase A, one warp per (pair, i1): the order-independent part of SearchForInitialization — the window query of
// Frame::GetFeaturesInArea(vbPrevMatched[i1], windowSize, level1, level1) (Frame.cc:657-723) and the Hamming distance of every
// candidate — written as a list in the reference's visiting order.  The order-dependent part (vMatchedDistance gate, best /
// second, accept, take-over) is phase B, which then only streams these lists.
__global__ void __launch_bounds__(256) initCandKernel(const InitArgs A, int pairs) {
    __shared__ int sColB[8][64], sPre[8][65];
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, wip = threadIdx.x >> 5;
    const int pair = wid / A.stride, i1 = wid - pair * A.stride;
    if (pair >= pairs) return;
    const size_t o = (size_t)pair * A.stride;
    const int n1 = A.n1Arr ? min(A.n1Arr[pair], A.stride) : A.n1;
    if (i1 >= n1) return;
    const float* k1 = A.kps1 + o * 7;
    const float* k2 = A.kps2 + o * 7;
    const int level1 = __float_as_int(k1[i1 * 7 + 5]);
    if (level1 > 0) return;
    int* colB = sColB[wip];
    int* pre = sPre[wip];
    const uint4* d1 = A.desc1 + o * 2;
    const uint4* d2 = A.desc2 + o * 2;
    const uint32_t* sorted = A.sorted2 + o;
    const int* cellStart = A.cellStart2 + (size_t)pair * (kGridCells + 1);
    const float x = A.prevFromKps ? k1[i1 * 7] : A.prev[(o + i1) * 2], y = A.prevFromKps ? k1[i1 * 7 + 1] : A.prev[(o + i1) * 2 + 1];
    const float r = A.r;
    uint32_t* list = A.candList + (o + i1) * kInitListCap;
    int count = 0;
    const int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), r), A.invW)));
    const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), r), A.invW)));
    const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), r), A.invH)));
    const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), r), A.invH)));
    if (!(minCX >= kGridCols || maxCX < 0 || minCY >= kGridRows || maxCY < 0 || minCX > maxCX || minCY > maxCY)) {
        const int ncols = maxCX - minCX + 1;
        int b0 = 0, len0 = 0, b1 = 0, len1 = 0;
        if (lane < ncols) {
            b0 = cellStart[(minCX + lane) * kGridRows + minCY];
            len0 = cellStart[(minCX + lane) * kGridRows + maxCY + 1] - b0;
        }
        if (lane + 32 < ncols) {
            b1 = cellStart[(minCX + lane + 32) * kGridRows + minCY];
            len1 = cellStart[(minCX + lane + 32) * kGridRows + maxCY + 1] - b1;
        }
        int s0 = len0, s1 = len1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t0 = __shfl_up_sync(0xffffffffu, s0, d), t1 = __shfl_up_sync(0xffffffffu, s1, d);
            if (lane >= d) { s0 += t0; s1 += t1; }
        }
        const int tot0 = __shfl_sync(0xffffffffu, s0, 31), total = tot0 + __shfl_sync(0xffffffffu, s1, 31);
        colB[lane] = b0; colB[lane + 32] = b1;
        pre[lane + 1] = s0; pre[lane + 33] = tot0 + s1;
        if (lane == 0) pre[0] = 0;
        __syncwarp();
        const uint4 a0 = d1[2 * i1], a1 = d1[2 * i1 + 1];
        for (int t0 = 0; t0 < total; t0 += 32) {
            const int t = t0 + lane;
            bool ok = false;
            uint32_t entry = 0;
            if (t < total) {
                int lo = 0, hi = ncols - 1;
                while (lo < hi) {
                    const int mid = (lo + hi + 1) >> 1;
                    if (pre[mid] <= t) lo = mid; else hi = mid - 1;
                }
                const int idx = (int)(sorted[colB[lo] + (t - pre[lo])] & 0xfffffu);
                const int oct = __float_as_int(k2[idx * 7 + 5]);
                const float x2 = k2[idx * 7], y2 = k2[idx * 7 + 1];
                ok = true;
                if (level1 >= 0) ok = !(oct < level1) && !(oct > level1);
                ok = ok && fabsf(__fsub_rn(x2, x)) < r && fabsf(__fsub_rn(y2, y)) < r;
                if (ok) {
                    const uint4 c0 = d2[2 * idx], c1 = d2[2 * idx + 1];
                    const int dist = __popc(a0.x ^ c0.x) + __popc(a0.y ^ c0.y) + __popc(a0.z ^ c0.z) + __popc(a0.w ^ c0.w) +
                                     __popc(a1.x ^ c1.x) + __popc(a1.y ^ c1.y) + __popc(a1.z ^ c1.z) + __popc(a1.w ^ c1.w);
                    entry = ((uint32_t)dist << 20) | (uint32_t)idx;
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            const int pos = count + __popc(bal & ((1u << lane) - 1u));
            if (ok && pos < kInitListCap) list[pos] = entry;
            count += __popc(bal);
        }
    }
    if (lane == 0) A.candCount[o + i1] = count;
}

__global__ void __launch_bounds__(32) searchInitKernel(const InitArgs A) {
    __shared__ int colB[64], pre[65], hist[30], keep[3];
    const int pair = blockIdx.x, lane = threadIdx.x;
    const size_t o = (size_t)pair * A.stride;
    const int n1 = A.n1Arr ? min(A.n1Arr[pair], A.stride) : A.n1, n2 = A.n2Arr ? min(A.n2Arr[pair], A.stride) : A.n2;
    const float* k1 = A.kps1 + o * 7;
    const float* k2 = A.kps2 + o * 7;
    const uint4* d1 = A.desc1 + o * 2;
    const uint4* d2 = A.desc2 + o * 2;
    const uint32_t* sorted = A.sorted2 + o;
    const int* cellStart = A.cellStart2 + (size_t)pair * (kGridCells + 1);
    float* prev = A.prev + o * 2;
    volatile int* m12 = A.m12 + o;
    volatile int* md = A.matchedDist + o;
    volatile int* m21 = A.m21 + o;
    volatile int* binOf = A.binOf + o;
    for (int i = lane; i < n2; i += 32) { md[i] = INT_MAX; m21[i] = -1; }
    for (int i = lane; i < n1; i += 32) {
        m12[i] = -1; binOf[i] = -1;
        if (A.prevFromKps) { prev[2 * i] = k1[i * 7]; prev[2 * i + 1] = k1[i * 7 + 1]; }
    }
    if (lane < 30) hist[lane] = 0;
    __syncwarp();
    int nmatches = 0;
    const float factor = 1.0f / 30.0f;  // :656
    int lvlBlock = 0, cntBlock = -1;
    for (int i1 = 0; i1 < n1; ++i1) {
        // octave and list length of 32 consecutive keypoints are fetched by the lanes at once (two dependent loads per i1 less)
        if ((i1 & 31) == 0) {
            const int j = i1 + lane;
            lvlBlock = j < n1 ? __float_as_int(k1[j * 7 + 5]) : 1;
            cntBlock = (A.candCount && j < n1 && lvlBlock <= 0) ? A.candCount[o + j] : -1;
        }
        const int level1 = __shfl_sync(0xffffffffu, lvlBlock, i1 & 31);
        const int listed = __shfl_sync(0xffffffffu, cntBlock, i1 & 31);
        if (level1 > 0) continue;  // :665-666
        unsigned long long key1 = ~0ull, key2 = ~0ull;  // distance << 48 | visit position << 24 | index
        if (listed >= 0 && listed <= kInitListCap) {
            // phase A's list (already in visiting order, distances attached): only the order-dependent gate is left
            const uint32_t* list = A.candList + (o + i1) * kInitListCap;
            for (int t = lane; t < listed; t += 32) {
                const uint32_t e = list[t];
                const int idx = (int)(e & 0xfffffu), dist = (int)(e >> 20);
                if (!(md[idx] <= dist))  // :687-688
                    top2Insert(key1, key2, ((unsigned long long)dist << 48) | ((unsigned long long)t << 24) | (unsigned)idx);
            }
        } else {
            const float x = prev[2 * i1], y = prev[2 * i1 + 1], r = A.r;
            // Frame::GetFeaturesInArea, Frame.cc:665-687
            const int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), r), A.invW)));
            const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), r), A.invW)));
            const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), r), A.invH)));
            const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), r), A.invH)));
            const bool inGrid = !(minCX >= kGridCols || maxCX < 0 || minCY >= kGridRows || maxCY < 0 || minCX > maxCX || minCY > maxCY);
            const int ncols = inGrid ? maxCX - minCX + 1 : 0;
            int b0 = 0, len0 = 0, b1 = 0, len1 = 0;
            if (lane < ncols) {
                b0 = cellStart[(minCX + lane) * kGridRows + minCY];
                len0 = cellStart[(minCX + lane) * kGridRows + maxCY + 1] - b0;
            }
            if (lane + 32 < ncols) {
                b1 = cellStart[(minCX + lane + 32) * kGridRows + minCY];
                len1 = cellStart[(minCX + lane + 32) * kGridRows + maxCY + 1] - b1;
            }
            int s0 = len0, s1 = len1;
    #pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t0 = __shfl_up_sync(0xffffffffu, s0, d), t1 = __shfl_up_sync(0xffffffffu, s1, d);
                if (lane >= d) { s0 += t0; s1 += t1; }
            }
            const int tot0 = __shfl_sync(0xffffffffu, s0, 31), total = tot0 + __shfl_sync(0xffffffffu, s1, 31);
            __syncwarp();
            colB[lane] = b0; colB[lane + 32] = b1;
            pre[lane + 1] = s0; pre[lane + 33] = tot0 + s1;
            if (lane == 0) pre[0] = 0;
            __syncwarp();
            const uint4 a0 = d1[2 * i1], a1 = d1[2 * i1 + 1];
            for (int t0 = 0; t0 < total; t0 += 32) {
                const int t = t0 + lane;
                if (t < total) {
                    int lo = 0, hi = ncols - 1;  // the column with pre[c] <= t < pre[c + 1]
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) >> 1;
                        if (pre[mid] <= t) lo = mid; else hi = mid - 1;
                    }
                    const int idx = (int)(sorted[colB[lo] + (t - pre[lo])] & 0xfffffu);
                    const int oct = __float_as_int(k2[idx * 7 + 5]);
                    const float x2 = k2[idx * 7], y2 = k2[idx * 7 + 1];
                    // minLevel = maxLevel = level1 (<= 0): bCheckLevels is true because maxLevel >= 0 or, for a negative level1, false
                    bool ok = true;
                    if (level1 >= 0) ok = !(oct < level1) && !(oct > level1);
                    ok = ok && fabsf(__fsub_rn(x2, x)) < r && fabsf(__fsub_rn(y2, y)) < r;
                    if (ok) {
                        const uint4 c0 = d2[2 * idx], c1 = d2[2 * idx + 1];
                        const int dist = __popc(a0.x ^ c0.x) + __popc(a0.y ^ c0.y) + __popc(a0.z ^ c0.z) + __popc(a0.w ^ c0.w) +
                                         __popc(a1.x ^ c1.x) + __popc(a1.y ^ c1.y) + __popc(a1.z ^ c1.z) + __popc(a1.w ^ c1.w);
                        if (!(md[idx] <= dist))  // :687-688
                            top2Insert(key1, key2, ((unsigned long long)dist << 48) | ((unsigned long long)t << 24) | (unsigned)idx);
                    }
                }
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, key1, d), o2 = __shfl_xor_sync(0xffffffffu, key2, d);
            top2Insert(key1, key2, o1);
            top2Insert(key1, key2, o2);
        }
        if (key1 == ~0ull) continue;
        const int bestDist = (int)(key1 >> 48), bestIdx2 = (int)(key1 & 0xffffffu);
        const float fBest2 = key2 == ~0ull ? 2147483648.0f : (float)(int)(key2 >> 48);  // (float)INT_MAX
        if (bestDist <= 50 && (float)bestDist < __fmul_rn(fBest2, A.nnratio)) {  // :697-699
            const int old = m21[bestIdx2];
            if (old >= 0) nmatches--;  // :701-705
            nmatches++;
            int bin = -1;
            if (A.checkOri) {
                float rot = __fsub_rn(k1[i1 * 7 + 3], k2[bestIdx2 * 7 + 3]);
                if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                bin = (int)roundf(__fmul_rn(rot, factor));
                if (bin == 30) bin = 0;
            }
            __syncwarp();
            if (lane == 0) {
                if (old >= 0) m12[old] = -1;
                m12[i1] = bestIdx2;
                m21[bestIdx2] = i1;
                md[bestIdx2] = bestDist;
                if (bin >= 0) { binOf[i1] = bin; hist[bin]++; }
            }
            __syncwarp();
        }
    }
    __syncwarp();
    if (A.checkOri) {
        if (lane == 0) {  // ComputeThreeMaxima, ORBmatcher.cc:2012-2053 (the histogram counts every push, withdrawn or not)
            int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < 30; i++) {
                const int sz = hist[i];
                if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
                else if (sz > max3) { max3 = sz; ind3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
            keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
        }
        __syncwarp();
        int dropped = 0;
        for (int i = lane; i < n1; i += 32) {
            const int b = binOf[i];
            if (b >= 0 && b != keep[0] && b != keep[1] && b != keep[2] && m12[i] >= 0) { m12[i] = -1; ++dropped; }  // :738-749
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) dropped += __shfl_xor_sync(0xffffffffu, dropped, d);
        nmatches -= dropped;
        __syncwarp();
    }
    for (int i = lane; i < n1; i += 32) {  // :756-759
        const int m = m12[i];
        if (m >= 0) { prev[2 * i] = k2[m * 7]; prev[2 * i + 1] = k2[m * 7 + 1]; }
    }
    if (lane == 0) A.nMatches[pair] = nmatches;
}

// cv::undistortPoints(src, dst, K, dist, R = I, P) as Frame::UndistortKeyPoints calls it (Frame.cc:766); every operation is
// written in the order of OpenCV's cvUndistortPointsInternal with explicit _rn intrinsics so that nvcc contracts nothing.
__device__ __forceinline__ double dm(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double da(double a, double b) { return __dadd_rn(a, b); }
__global__ void undistortKernel(const float* kpsIn, float* kpsOut, const int* __restrict__ nPerFrame,
                                int nSingle, int cap, const UndistortCam C) {
    const int f = blockIdx.y;
    const int n = nPerFrame ? nPerFrame[f] : nSingle;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* in = kpsIn + ((size_t)f * cap + i) * 7;
    float* out = kpsOut + ((size_t)f * cap + i) * 7;
    const double u = (double)in[0], v = (double)in[1];
    double x = dm(__dsub_rn(u, C.cx), C.ifx), y = dm(__dsub_rn(v, C.cy), C.ify);
    const double x0 = x, y0 = y;
    const double* k = C.k;
    for (int j = 0; j < 5; ++j) {
        const double r2 = da(dm(x, x), dm(y, y));
        const double num = da(1.0, dm(da(dm(da(dm(k[7], r2), k[6]), r2), k[5]), r2));
        const double den = da(1.0, dm(da(dm(da(dm(k[4], r2), k[1]), r2), k[0]), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) {
            x = x0;  // (u - cx) * ifx again
            y = y0;
            break;
        }
        // deltaX = 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2, left to right
        const double dX = da(da(da(dm(dm(dm(2.0, k[2]), x), y), dm(k[3], da(r2, dm(dm(2.0, x), x)))), dm(k[8], r2)), dm(dm(k[9], r2), r2));
        // deltaY = k2*(r2 + 2*y*y) + 2*k3*x*y + k10*r2 + k11*r2*r2
        const double dY = da(da(da(dm(k[2], da(r2, dm(dm(2.0, y), y))), dm(dm(dm(2.0, k[3]), x), y)), dm(k[10], r2)), dm(dm(k[11], r2), r2));
        x = dm(__dsub_rn(x0, dX), icdist);
        y = dm(__dsub_rn(y0, dY), icdist);
    }
    const double xx = da(da(dm(C.rr[0], x), dm(C.rr[1], y)), C.rr[2]);
    const double yy = da(da(dm(C.rr[3], x), dm(C.rr[4], y)), C.rr[5]);
    const double ww = __ddiv_rn(1.0, da(da(dm(C.rr[6], x), dm(C.rr[7], y)), C.rr[8]));
    out[0] = (float)dm(xx, ww);
    out[1] = (float)dm(yy, ww);
    if (out != in) {
#pragma unroll
        for (int w = 2; w < 7; ++w) out[w] = in[w];  // size, angle, response, octave, class_id carried over (:773-777)
    }
}

}  // namespace

void launchAccept(int rule, float nnratio, const int32_t* d1, const int32_t* d2, const int64_t* idx2, const int32_t* l1,
                  const int32_t* l2, int n, uint8_t* accept, cudaStream_t s) {
    if (n <= 0) return;
    acceptKernel<<<(n + 255) / 256, 256, 0, s>>>(rule, nnratio, d1, d2, reinterpret_cast<const long long*>(idx2), l1, l2, n, accept);
    countLaunch();
}

int gridSortCapacity() { return 32768; }  // keys of one CTA's shared memory (128 KB)

void launchGridSortBatch(const void* kps, const int* nPerFrame, int frames, int cap, float minX, float maxX, float minY, float maxY,
                         uint32_t* sorted, int* cellStart, cudaStream_t s) {
    if (frames <= 0) return;
    const float invW = (float)kGridCols / (maxX - minX), invH = (float)kGridRows / (maxY - minY);  // Frame.cc:251-252
    int nPow2 = 32;
    while (nPow2 < cap) nPow2 <<= 1;
    const size_t smem = (size_t)nPow2 * 4;
    if (smem > 40 * 1024) cudaFuncSetAttribute(gridSortKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    gridSortKernel<<<frames, 1024, smem, s>>>(static_cast<const float*>(kps), 0, nPow2, minX, minY, invW, invH, sorted, cellStart,
                                              nPerFrame, cap);
    countLaunch();
}

void launchGridSearch(const void* kpsT, const uint8_t* descT, int nt, float minX, float maxX, float minY, float maxY,
                      const uint8_t* skipT, const float* uRightT, const void* queries, const uint8_t* descQ, int nq,
                      uint32_t* sortedScratch, int* cellStart, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2,
                      int32_t* lvl1, int32_t* lvl2, cudaStream_t s) {
    const float invW = (float)kGridCols / (maxX - minX), invH = (float)kGridRows / (maxY - minY);  // Frame.cc:251-252
    int nPow2 = 32;
    while (nPow2 < nt) nPow2 <<= 1;
    const size_t smem = (size_t)nPow2 * 4;
    if (smem > 40 * 1024) cudaFuncSetAttribute(gridSortKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    gridSortKernel<<<1, 1024, smem, s>>>(static_cast<const float*>(kpsT), nt, nPow2, minX, minY, invW, invH, sortedScratch, cellStart,
                                         nullptr, 0);
    countLaunch();
    if (nq <= 0) return;
    gridSearchKernel<<<(nq + 7) / 8, 256, 0, s>>>(static_cast<const float*>(kpsT), reinterpret_cast<const uint4*>(descT), sortedScratch,
                                                  cellStart, skipT, uRightT, minX, minY, invW, invH,
                                                  static_cast<const GridQuery*>(queries), reinterpret_cast<const uint4*>(descQ), nq, idx1,
                                                  d1, idx2, d2, lvl1, lvl2);
    countLaunch();
}

void launchStereoRefine(const FramePlan& P, const float* sf, const float* isf, const uint8_t* pyrL, const uint8_t* pyrR,
                        const void* kl, const void* kr, int nl, const int* bestIdx, const int* bestDist, float mbf, float maxD,
                        float* uRight, float* depth, int* sad, cudaStream_t s, const StereoBatch* batch) {
    const StereoBatch SB = batch ? *batch : StereoBatch{};
    const int frames = batch ? batch->frames : 1;
    if (batch) nl = batch->cap;
    if (nl <= 0) return;
    StereoLevels LV{};
    for (int l = 0; l < P.nlevels; ++l) {
        LV.planeOff[l] = P.lv[l].planeOff; LV.pitch[l] = P.lv[l].pitch; LV.cols[l] = P.lv[l].w;
        LV.sf[l] = sf[l]; LV.isf[l] = isf[l];
    }
    stereoRefineKernel<<<dim3((nl + 7) / 8, frames), 256, 0, s>>>(LV, pyrL, pyrR, static_cast<const float*>(kl),
                                                                  static_cast<const float*>(kr), nl, bestIdx, bestDist, mbf, maxD,
                                                                  uRight, depth, sad, SB);
    countLaunch();
    const size_t smem = (size_t)kSadBins * 4;
    cudaFuncSetAttribute(stereoCutKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    stereoCutKernel<<<frames, 1024, smem, s>>>(sad, nl, uRight, depth, SB);
    countLaunch();
}

void launchUndistort(const void* kpsIn, void* kpsOut, const int* nPerFrame, int nSingle, int frames, int cap, const UndistortCam& C,
                     cudaStream_t s) {
    if (frames <= 0 || cap <= 0 || (!nPerFrame && nSingle <= 0)) return;
    const int most = nPerFrame ? cap : nSingle;
    undistortKernel<<<dim3((most + 127) / 128, frames), 128, 0, s>>>((const float*)kpsIn, (float*)kpsOut, nPerFrame, nSingle, cap, C);
    countLaunch();
}

// frames = 1: single pair (kps2 / desc2 are F2's arrays, n1 / n2 given); frames > 1: pair p = frame p -> frame p + 1 of a
// [frame][cap] batch (kps2 = kps1, counts from nPerFrame).  sorted / cellStart: scratch for F2's grids (frames x cap,
// frames x 3073 ints); prev / m12 / scratch3: pairs x cap (x 2 floats for prev; scratch3 = 3 int arrays back to back).
// candList / candCount (may be NULL: phase B then walks every window itself): pairs x cap x kInitListCap words, pairs x cap ints.
void launchSearchInit(const void* kps1, const uint8_t* desc1, int n1, const void* kps2, const uint8_t* desc2, int n2, const int* nPerFrame,
                      int frames, int cap, float minX, float maxX, float minY, float maxY, int windowSize, float nnratio, int checkOri,
                      int prevFromKps, uint32_t* sorted, int* cellStart, float* prev, int* m12, int* scratch3, int* nMatches,
                      cudaStream_t s, uint32_t* candList, int* candCount) {
    const float invW = (float)kGridCols / (maxX - minX), invH = (float)kGridRows / (maxY - minY);  // Frame.cc:251-252
    const bool batch = nPerFrame != nullptr;
    int nPow2 = 32;
    while (nPow2 < (batch ? cap : n2)) nPow2 <<= 1;
    const size_t smem = (size_t)nPow2 * 4;
    if (smem > 40 * 1024) cudaFuncSetAttribute(gridSortKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int pairs = batch ? frames - 1 : 1;
    if (pairs <= 0) return;
    InitArgs A{};
    A.kps1 = static_cast<const float*>(kps1); A.desc1 = reinterpret_cast<const uint4*>(desc1);
    A.n1Arr = batch ? nPerFrame : nullptr; A.n2Arr = batch ? nPerFrame + 1 : nullptr;
    A.n1 = n1; A.n2 = n2; A.stride = cap;
    A.minX = minX; A.minY = minY; A.invW = invW; A.invH = invH; A.r = (float)windowSize; A.nnratio = nnratio;
    A.checkOri = checkOri; A.prevFromKps = prevFromKps;
    A.prev = prev; A.m12 = m12; A.matchedDist = scratch3; A.m21 = scratch3 + (size_t)pairs * cap;
    A.binOf = scratch3 + 2 * (size_t)pairs * cap; A.nMatches = nMatches;
    if (batch) {  // grids of frames 1 .. frames-1 (sorted from frame 0 so that the indexing stays [frame])
        gridSortKernel<<<frames, 1024, smem, s>>>(static_cast<const float*>(kps1), 0, nPow2, minX, minY, invW, invH, sorted, cellStart,
                                                  nPerFrame, cap);
        A.kps2 = A.kps1 + (size_t)cap * 7; A.desc2 = A.desc1 + (size_t)cap * 2;
        A.sorted2 = sorted + cap; A.cellStart2 = cellStart + (kGridCells + 1);
    } else {
        gridSortKernel<<<1, 1024, smem, s>>>(static_cast<const float*>(kps2), n2, nPow2, minX, minY, invW, invH, sorted, cellStart, nullptr, 0);
        A.kps2 = static_cast<const float*>(kps2); A.desc2 = reinterpret_cast<const uint4*>(desc2);
        A.sorted2 = sorted; A.cellStart2 = cellStart;
    }
    A.candList = candList; A.candCount = candCount;
    if (candList && candCount) {  // phase A: one warp per (pair, i1); slots of higher-level or absent keypoints are never read
        const long long warps = (long long)pairs * cap;
        initCandKernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(A, pairs);
        countLaunch();
    }
    searchInitKernel<<<pairs, 32, 0, s>>>(A);
    countLaunch(2);
}

}  // namespace orbx

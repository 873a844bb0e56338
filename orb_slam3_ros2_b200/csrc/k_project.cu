// k_project.cu — ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), whole function
// (ORB_SLAM3/src/ORBmatcher.cc:1676-1887), the matcher Tracking::TrackWithMotionModel runs on every tracked frame
// (Tracking.cc:2876, :2906, :2914); frames with Nleft == -1 (monocular, rectified stereo, RGB-D).
//
// Per last-frame feature i that holds a map point and is no outlier: x3Dc = Tcw * x3Dw (Sophus' quaternion form, operation by
// operation: so3.hpp:358-368, se3.hpp:321-325), invzc < 0 -> skip, uv = Pinhole::project (CameraModels/Pinhole.cpp:43-49),
// image-bounds test, radius = th * mvScaleFactors[octave], level window by bForward / bBackward (:1727-1741),
// Frame::GetFeaturesInArea (Frame.cc:657-723), skip features that already hold a map point with observations (:1745-1747)
// or fail the uRight gate (:1749-1755), best Hamming distance with strict < (first minimum in visiting order), accept
// bestDist <= TH_HIGH (:1770), CurrentFrame.mvpMapPoints[bestIdx2] = pMP (the last writer wins, nmatches counts every accept),
// rotation histogram + ComputeThreeMaxima (:1864-1884, every pushed entry of a dropped bin clears its feature and decrements).
//
// The i loop is order dependent only through "already holds a map point with observations": two phases like
// SearchForInitialization.  Phase A, one warp per (pair, i), does everything that does not depend on the order - projection,
// window query over the sorted grid (contiguous key ranges per grid column, k_guided.cu), level / window / uRight filters,
// Hamming distances - and writes the candidates in visiting order as distance << 20 | index (kProjListCap per i).  Phase B,
// one warp per frame pair, walks i in order over those lists with the taken-bits in shared memory; a list that overflowed is
// re-enumerated in place.  All float operations are single rounded (__f*_rn) in the reference's order.
#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr int kGridRows = 48, kGridCols = 64, kGridCells = kGridRows * kGridCols;  // Frame.h:44-45
constexpr int kProjListCap = 64;
constexpr int kMaxCur = 32768;  // taken-bits of one pair in shared memory (= gridSortCapacity)

struct ProjArgs {
    const float* kpsLast;      // [pair][capLast] x 7 (mvKeysUn of the last frame)
    const uint8_t* lastFlags;  // bit 0 = map point and not outlier, bit 1 = Observations() > 0
    const float* world;        // [pair][capLast] x 3
    const uint4* mpDesc;       // [pair][capLast] x 32 B (MapPoint::GetDescriptor)
    const int* nLast;          // [pair]
    const float* kpsCur;       // [pair][capCur] x 7 (mvKeysUn of the current frame)
    const uint4* descCur;
    const float* uRightCur;    // may be NULL
    const uint8_t* takenCur;   // may be NULL
    const int* nCur;           // [pair]
    const float* poses;        // [pair] x 14: Tcw (qx qy qz qw tx ty tz), Tlw
    const uint32_t* sorted;    // [pair][capCur] grid keys of the current frame (gridSortKernel)
    const int* cellStart;      // [pair][kGridCells + 1]
    uint32_t* candList;        // [pair][capLast][kProjListCap]
    int* candCount;            // [pair][capLast]: -1 = i is not projected / has no window, else candidates found (may exceed the cap)
    float* proj;               // [pair][capLast] x 4: u, v, radius, ur  (phase A -> phase B re-enumeration)
    int* levels;               // [pair][capLast] x 2: minLevel, maxLevel
    uint32_t* pushes;          // [pair][capLast]: rotation bin << 20 | current feature, in push order
    int* matchCur;             // [pair][capCur] out
    int* nMatches;             // [pair] out
    int capLast, capCur;
    float fx, fy, cx, cy, minX, maxX, minY, maxY, invW, invH, mbf, mb, th;
    int mono, checkOri;
    float sf[kMaxLevels];
};

struct GridGeom {
    float minX, minY, invW, invH;
};

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ float fm(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fa(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fs(float a, float b) { return __fsub_rn(a, b); }
// Eigen's cross product (a1 b2 - a2 b1, a2 b0 - a0 b2, a0 b1 - a1 b0), every operation rounded once
__device__ __forceinline__ V3 crossq(float ax, float ay, float az, V3 b) {
    return {fs(fm(ay, b.z), fm(az, b.y)), fs(fm(az, b.x), fm(ax, b.z)), fs(fm(ax, b.y), fm(ay, b.x))};
}
// Sophus SO3 * point (so3.hpp:366-368): uv = q.vec x p; uv += uv; p + w * uv + q.vec x uv
__device__ __forceinline__ V3 rotq(float qx, float qy, float qz, float qw, V3 p) {
    V3 uv = crossq(qx, qy, qz, p);
    uv.x = fa(uv.x, uv.x); uv.y = fa(uv.y, uv.y); uv.z = fa(uv.z, uv.z);
    const V3 c = crossq(qx, qy, qz, uv);
    return {fa(fa(p.x, fm(qw, uv.x)), c.x), fa(fa(p.y, fm(qw, uv.y)), c.y), fa(fa(p.z, fm(qw, uv.z)), c.z)};
}
__device__ __forceinline__ V3 se3mul(const float* T, V3 p) {  // se3.hpp:324
    const V3 r = rotq(T[0], T[1], T[2], T[3], p);
    return {fa(r.x, T[4]), fa(r.y, T[5]), fa(r.z, T[6])};
}

// Frame::GetFeaturesInArea(u, v, radius, minLevel, maxLevel) + the order-independent filters of the candidate loop, in the
// reference's visiting order (grid column by grid column, rows inside, insertion order inside a cell), 32 candidates per trip:
// emit(ok, idx, dist, position) is called by the whole warp.  colB / pre: 64 + 65 ints of shared memory of this warp.
template <class Emit>
__device__ __forceinline__ int projEnumerate(const GridGeom A, const float* kc, const uint4* dc, const float* urc, const uint32_t* sorted,
                                             const int* cellStart, float u, float v, float radius, int minLevel, int maxLevel, float ur,
                                             const uint4 a0, const uint4 a1, int lane, int* colB, int* pre, Emit emit) {
    const int minCX = max(0, (int)floorf(fm(fs(fs(u, A.minX), radius), A.invW)));
    const int maxCX = min(kGridCols - 1, (int)ceilf(fm(fa(fs(u, A.minX), radius), A.invW)));
    const int minCY = max(0, (int)floorf(fm(fs(fs(v, A.minY), radius), A.invH)));
    const int maxCY = min(kGridRows - 1, (int)ceilf(fm(fa(fs(v, A.minY), radius), A.invH)));
    if (minCX >= kGridCols || maxCX < 0 || minCY >= kGridRows || maxCY < 0 || minCX > maxCX || minCY > maxCY) return 0;
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
    __syncwarp();
    colB[lane] = b0; colB[lane + 32] = b1;
    pre[lane + 1] = s0; pre[lane + 33] = tot0 + s1;
    if (lane == 0) pre[0] = 0;
    __syncwarp();
    const bool checkLevels = minLevel > 0 || maxLevel >= 0;
    for (int t0 = 0; t0 < total; t0 += 32) {
        const int t = t0 + lane;
        bool ok = false;
        int idx = 0, dist = 0;
        if (t < total) {
            int lo = 0, hi = ncols - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (pre[mid] <= t) lo = mid; else hi = mid - 1;
            }
            idx = (int)(sorted[colB[lo] + (t - pre[lo])] & 0xfffffu);
            const int oct = __float_as_int(kc[idx * 7 + 5]);
            const float x2 = kc[idx * 7], y2 = kc[idx * 7 + 1];
            ok = true;
            if (checkLevels) ok = !(oct < minLevel) && !(maxLevel >= 0 && oct > maxLevel);
            ok = ok && fabsf(fs(x2, u)) < radius && fabsf(fs(y2, v)) < radius;
            if (ok && urc) {  // :1749-1755 (Nleft == -1): er = fabs(ur - mvuRight[i2]) > radius -> skip
                const float urt = urc[idx];
                if (urt > 0.f) ok = !(fabsf(fs(ur, urt)) > radius);
            }
            if (ok) {
                const uint4 c0 = dc[2 * idx], c1 = dc[2 * idx + 1];
                dist = __popc(a0.x ^ c0.x) + __popc(a0.y ^ c0.y) + __popc(a0.z ^ c0.z) + __popc(a0.w ^ c0.w) +
                       __popc(a1.x ^ c1.x) + __popc(a1.y ^ c1.y) + __popc(a1.z ^ c1.z) + __popc(a1.w ^ c1.w);
            }
        }
        emit(ok, idx, dist, t);
    }
    return total;
}

// ---- phase A: one warp per (pair, i) ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) projCandKernel(const ProjArgs A, int pairs) {
    __shared__ int sColB[8][64], sPre[8][65];
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, wip = threadIdx.x >> 5;
    const int pair = wid / A.capLast, i = wid - pair * A.capLast;
    if (pair >= pairs) return;
    const int nLast = min(A.nLast[pair], A.capLast);
    if (i >= nLast) return;
    const size_t oL = (size_t)pair * A.capLast, oC = (size_t)pair * A.capCur;
    int* cnt = A.candCount + oL + i;
    if (!(A.lastFlags[oL + i] & 1)) {  // no map point, or an outlier (:1698-1701)
        if (lane == 0) *cnt = -1;
        return;
    }
    const float* T = A.poses + (size_t)pair * 14;
    // bForward / bBackward (:1686-1693): twc = Tcw.inverse().translation() = q* (-t); tlc = Tlw * twc
    const V3 twc = rotq(-T[0], -T[1], -T[2], T[3], {fm(T[4], -1.f), fm(T[5], -1.f), fm(T[6], -1.f)});
    const V3 tlc = se3mul(T + 7, twc);
    const bool bForward = tlc.z > A.mb && !A.mono, bBackward = -tlc.z > A.mb && !A.mono;
    const float* w = A.world + (oL + i) * 3;
    const V3 x3Dc = se3mul(T, {w[0], w[1], w[2]});
    const float invzc = (float)__ddiv_rn(1.0, (double)x3Dc.z);  // 1.0 / x3Dc(2): a double division, rounded to float (:1709)
    bool skip = invzc < 0;
    const float u = fa(__fdiv_rn(fm(A.fx, x3Dc.x), x3Dc.z), A.cx), v = fa(__fdiv_rn(fm(A.fy, x3Dc.y), x3Dc.z), A.cy);  // Pinhole.cpp:45-46
    skip = skip || u < A.minX || u > A.maxX || v < A.minY || v > A.maxY;
    if (skip) {
        if (lane == 0) *cnt = -1;
        return;
    }
    const int oct = __float_as_int(A.kpsLast[(oL + i) * 7 + 5]);
    const float radius = fm(A.th, A.sf[min(max(oct, 0), kMaxLevels - 1)]);
    int minLevel, maxLevel;
    if (bForward) { minLevel = oct; maxLevel = -1; }
    else if (bBackward) { minLevel = 0; maxLevel = oct; }
    else { minLevel = oct - 1; maxLevel = oct + 1; }
    const float ur = fs(u, fm(A.mbf, invzc));
    const uint4 a0 = A.mpDesc[(oL + i) * 2], a1 = A.mpDesc[(oL + i) * 2 + 1];
    uint32_t* list = A.candList + (oL + i) * kProjListCap;
    int count = 0;
    projEnumerate(GridGeom{A.minX, A.minY, A.invW, A.invH}, A.kpsCur + oC * 7, A.descCur + oC * 2, A.uRightCur ? A.uRightCur + oC : nullptr, A.sorted + oC,
                  A.cellStart + (size_t)pair * (kGridCells + 1), u, v, radius, minLevel, maxLevel, ur, a0, a1, lane, sColB[wip], sPre[wip],
                  [&](bool ok, int idx, int dist, int) {
                      const unsigned bal = __ballot_sync(0xffffffffu, ok);
                      const int pos = count + __popc(bal & ((1u << lane) - 1u));
                      if (ok && pos < kProjListCap) list[pos] = ((uint32_t)dist << 20) | (uint32_t)idx;
                      count += __popc(bal);
                  });
    if (lane == 0) {
        *cnt = count;
        float* pj = A.proj + (oL + i) * 4;
        pj[0] = u; pj[1] = v; pj[2] = radius; pj[3] = ur;
        A.levels[(oL + i) * 2] = minLevel; A.levels[(oL + i) * 2 + 1] = maxLevel;
    }
}

// ---- phase B: one warp per pair walks i in order -------------------------------------------------------------------------
__global__ void __launch_bounds__(32) projWalkKernel(const ProjArgs A) {
    __shared__ uint32_t taken[kMaxCur / 32];
    __shared__ int colB[64], pre[65], hist[30], keep[3];
    const int pair = blockIdx.x, lane = threadIdx.x;
    const size_t oL = (size_t)pair * A.capLast, oC = (size_t)pair * A.capCur;
    const int nLast = min(A.nLast[pair], A.capLast), nCur = min(min(A.nCur[pair], A.capCur), kMaxCur);
    const float* kl = A.kpsLast + oL * 7;
    const float* kc = A.kpsCur + oC * 7;
    int* match = A.matchCur + oC;
    uint32_t* pushes = A.pushes + oL;
    for (int w = lane; w < (nCur + 31) / 32; w += 32) {
        uint32_t bits = 0;
        if (A.takenCur)
            for (int b = 0; b < 32 && w * 32 + b < nCur; ++b) bits |= (A.takenCur[oC + w * 32 + b] ? 1u : 0u) << b;
        taken[w] = bits;
    }
    for (int i = lane; i < A.capCur; i += 32) match[i] = -1;
    if (lane < 30) hist[lane] = 0;
    __syncwarp();
    int nmatches = 0, nPush = 0;
    const float factor = 1.0f / 30.0f;  // 1.0f / HISTO_LENGTH (:1684)
    for (int i = 0; i < nLast; ++i) {
        const int cnt = A.candCount[oL + i];
        if (cnt <= 0) continue;  // not projected, or vIndices2.empty() (:1743-1744)
        unsigned long long best = ~0ull;  // distance << 40 | visiting position << 20 | index
        if (cnt <= kProjListCap) {
            const uint32_t* list = A.candList + (oL + i) * kProjListCap;
            for (int p0 = 0; p0 < cnt; p0 += 32) {
                const int p = p0 + lane;
                if (p < cnt) {
                    const uint32_t e = list[p];
                    const uint32_t idx = e & 0xfffffu;
                    if (!((taken[idx >> 5] >> (idx & 31)) & 1u))
                        best = min(best, ((unsigned long long)(e >> 20) << 40) | ((unsigned long long)p << 20) | idx);
                }
            }
        } else {  // the list overflowed: walk the window again, here
            const float* pj = A.proj + (oL + i) * 4;
            const uint4 a0 = A.mpDesc[(oL + i) * 2], a1 = A.mpDesc[(oL + i) * 2 + 1];
            projEnumerate(GridGeom{A.minX, A.minY, A.invW, A.invH}, kc, A.descCur + oC * 2, A.uRightCur ? A.uRightCur + oC : nullptr, A.sorted + oC,
                          A.cellStart + (size_t)pair * (kGridCells + 1), pj[0], pj[1], pj[2], A.levels[(oL + i) * 2], A.levels[(oL + i) * 2 + 1],
                          pj[3], a0, a1, lane, colB, pre, [&](bool ok, int idx, int dist, int t) {
                              if (ok && !((taken[idx >> 5] >> (idx & 31)) & 1u))
                                  best = min(best, ((unsigned long long)dist << 40) | ((unsigned long long)t << 20) | (unsigned)idx);
                          });
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (best != ~0ull && (int)(best >> 40) <= 100) {  // bestDist <= TH_HIGH (:1770)
            const int idx = (int)(best & 0xfffffu);
            ++nmatches;
            if (lane == 0) {
                match[idx] = i;  // CurrentFrame.mvpMapPoints[bestIdx2] = pMP: the last writer wins
                if (A.lastFlags[oL + i] & 2) taken[idx >> 5] |= 1u << (idx & 31);  // pMP->Observations() > 0: skipped from now on
                if (A.checkOri) {
                    float rot = fs(kl[i * 7 + 3], kc[idx * 7 + 3]);
                    if (rot < 0.0f) rot = fa(rot, 360.0f);
                    int bin = (int)roundf(fm(rot, factor));
                    if (bin == 30) bin = 0;
                    pushes[nPush] = ((uint32_t)bin << 20) | (uint32_t)idx;
                    hist[bin]++;
                }
            }
            if (A.checkOri) ++nPush;
        }
        __syncwarp();
    }
    if (A.checkOri) {
        if (lane == 0) {  // ComputeThreeMaxima, ORBmatcher.cc:2012-2053
            int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < 30; i++) {
                const int sz = hist[i];
                if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
                else if (sz > max3) { max3 = sz; ind3 = i; }
            }
            if ((float)max2 < fm(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < fm(0.1f, (float)max1)) { ind3 = -1; }
            keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
        }
        __syncwarp();
        int dropped = 0;
        for (int j = lane; j < nPush; j += 32) {  // :1873-1883: every entry of a dropped bin clears its feature and decrements
            const uint32_t e = pushes[j];
            const int b = (int)(e >> 20);
            if (b != keep[0] && b != keep[1] && b != keep[2]) {
                match[e & 0xfffffu] = -1;
                ++dropped;
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) dropped += __shfl_xor_sync(0xffffffffu, dropped, d);
        nmatches -= dropped;
    }
    if (lane == 0) A.nMatches[pair] = nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, th, bFarPoints, thFarPoints) -------------------
// (ORBmatcher.cc:43-210 + RadiusByViewingCos :215-221; Tracking::SearchLocalPoints, Tracking.cc:3410, :3433), Nleft == -1 frames.
// The map points arrive already projected (Frame::isInFrustum set mTrackProjX / Y / XR, mnTrackScaleLevel, mTrackViewCos), so phase A
// is the window query alone; phase B keeps best AND second (strict <, first in visiting order) with their octaves and applies the
// accept rule of :123-128 before F.mvpMapPoints[bestIdx] = pMP.
struct MapArgs {
    const float* track;        // [pair][capMp] x 5: mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos, mTrackDepth
    const int* level;          // mnTrackScaleLevel
    const uint8_t* flags;      // bit 0 = mbTrackInView, bit 1 = Observations() > 0, bit 2 = isBad()
    const uint4* mpDesc;
    const int* nMp;            // [pair]
    const float* kpsCur;
    const uint4* descCur;
    const float* uRightCur;
    const uint8_t* takenCur;
    const int* nCur;
    const uint32_t* sorted;
    const int* cellStart;
    uint32_t* candList;        // [pair][capMp][kProjListCap]
    int* candCount;
    float* radius;             // [pair][capMp]: r * mvScaleFactors[level]
    int* matchCur;
    int* nMatches;
    int capMp, capCur;
    float minX, minY, invW, invH, th, thFar, nnratio;
    int farPoints;
    float sf[kMaxLevels];
};

__global__ void __launch_bounds__(256) mapCandKernel(const MapArgs A, int pairs) {
    __shared__ int sColB[8][64], sPre[8][65];
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, wip = threadIdx.x >> 5;
    const int pair = wid / A.capMp, i = wid - pair * A.capMp;
    if (pair >= pairs) return;
    const int nMp = min(A.nMp[pair], A.capMp);
    if (i >= nMp) return;
    const size_t oM = (size_t)pair * A.capMp, oC = (size_t)pair * A.capCur;
    int* cnt = A.candCount + oM + i;
    const float* tr = A.track + (oM + i) * 5;
    const uint8_t fl = A.flags[oM + i];
    // :52-59: not in view, too far, bad
    if (!(fl & 1) || (A.farPoints && tr[4] > A.thFar) || (fl & 4)) {
        if (lane == 0) *cnt = -1;
        return;
    }
    const int lvl = min(max(A.level[oM + i], 0), kMaxLevels - 1);
    float r = (double)tr[3] > 0.998 ? 2.5f : 4.0f;  // RadiusByViewingCos: the float against a double literal
    if (A.th != 1.0f) r = fm(r, A.th);             // bFactor (:47, :68-69)
    const float radius = fm(r, A.sf[lvl]);
    const uint4 a0 = A.mpDesc[(oM + i) * 2], a1 = A.mpDesc[(oM + i) * 2 + 1];
    uint32_t* list = A.candList + (oM + i) * kProjListCap;
    int count = 0;
    projEnumerate(GridGeom{A.minX, A.minY, A.invW, A.invH}, A.kpsCur + oC * 7, A.descCur + oC * 2, A.uRightCur ? A.uRightCur + oC : nullptr,
                  A.sorted + oC, A.cellStart + (size_t)pair * (kGridCells + 1), tr[0], tr[1], radius, lvl - 1, lvl, tr[2], a0, a1, lane,
                  sColB[wip], sPre[wip], [&](bool ok, int idx, int dist, int) {
                      const unsigned bal = __ballot_sync(0xffffffffu, ok);
                      const int pos = count + __popc(bal & ((1u << lane) - 1u));
                      if (ok && pos < kProjListCap) list[pos] = ((uint32_t)dist << 20) | (uint32_t)idx;
                      count += __popc(bal);
                  });
    if (lane == 0) {
        *cnt = count;
        A.radius[oM + i] = radius;
    }
}

__device__ __forceinline__ void top2Key(unsigned long long& k1, unsigned long long& k2, unsigned long long k) {
    k2 = min(k2, max(k1, k));
    k1 = min(k1, k);
}

__global__ void __launch_bounds__(32) mapWalkKernel(const MapArgs A) {
    __shared__ uint32_t taken[kMaxCur / 32];
    __shared__ int colB[64], pre[65];
    const int pair = blockIdx.x, lane = threadIdx.x;
    const size_t oM = (size_t)pair * A.capMp, oC = (size_t)pair * A.capCur;
    const int nMp = min(A.nMp[pair], A.capMp), nCur = min(min(A.nCur[pair], A.capCur), kMaxCur);
    const float* kc = A.kpsCur + oC * 7;
    int* match = A.matchCur + oC;
    for (int w = lane; w < (nCur + 31) / 32; w += 32) {
        uint32_t bits = 0;
        if (A.takenCur)
            for (int b = 0; b < 32 && w * 32 + b < nCur; ++b) bits |= (A.takenCur[oC + w * 32 + b] ? 1u : 0u) << b;
        taken[w] = bits;
    }
    for (int i = lane; i < A.capCur; i += 32) match[i] = -1;
    __syncwarp();
    int nmatches = 0;
    for (int i = 0; i < nMp; ++i) {
        const int cnt = A.candCount[oM + i];
        if (cnt <= 0) continue;
        unsigned long long k1 = ~0ull, k2 = ~0ull;  // distance << 40 | visiting position << 20 | index
        if (cnt <= kProjListCap) {
            const uint32_t* list = A.candList + (oM + i) * kProjListCap;
            for (int p0 = 0; p0 < cnt; p0 += 32) {
                const int p = p0 + lane;
                if (p < cnt) {
                    const uint32_t e = list[p];
                    const uint32_t idx = e & 0xfffffu;
                    if (!((taken[idx >> 5] >> (idx & 31)) & 1u))
                        top2Key(k1, k2, ((unsigned long long)(e >> 20) << 40) | ((unsigned long long)p << 20) | idx);
                }
            }
        } else {
            const float* tr = A.track + (oM + i) * 5;
            const int lvl = min(max(A.level[oM + i], 0), kMaxLevels - 1);
            const uint4 a0 = A.mpDesc[(oM + i) * 2], a1 = A.mpDesc[(oM + i) * 2 + 1];
            projEnumerate(GridGeom{A.minX, A.minY, A.invW, A.invH}, kc, A.descCur + oC * 2, A.uRightCur ? A.uRightCur + oC : nullptr,
                          A.sorted + oC, A.cellStart + (size_t)pair * (kGridCells + 1), tr[0], tr[1], A.radius[oM + i], lvl - 1, lvl, tr[2],
                          a0, a1, lane, colB, pre, [&](bool ok, int idx, int dist, int t) {
                              if (ok && !((taken[idx >> 5] >> (idx & 31)) & 1u))
                                  top2Key(k1, k2, ((unsigned long long)dist << 40) | ((unsigned long long)t << 20) | (unsigned)idx);
                          });
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
            top2Key(k1, k2, o1);
            top2Key(k1, k2, o2);
        }
        if (k1 != ~0ull) {
            const int bestDist = (int)(k1 >> 40), bestIdx = (int)(k1 & 0xfffffu);
            const int bestDist2 = k2 != ~0ull ? (int)(k2 >> 40) : 256;
            const int bestLevel = __float_as_int(kc[bestIdx * 7 + 5]);
            const int bestLevel2 = k2 != ~0ull ? __float_as_int(kc[(int)(k2 & 0xfffffu) * 7 + 5]) : -1;
            // :123-128: bestDist <= TH_HIGH and (levels differ or (float)bestDist <= mfNNratio * bestDist2)
            const float lim = fm(A.nnratio, (float)bestDist2);
            const bool ok = bestDist <= 100 && (bestLevel != bestLevel2 || (float)bestDist <= lim);
            if (ok) {
                ++nmatches;
                if (lane == 0) {
                    match[bestIdx] = i;
                    if (A.flags[oM + i] & 2) taken[bestIdx >> 5] |= 1u << (bestIdx & 31);
                }
            }
        }
        __syncwarp();
    }
    if (lane == 0) A.nMatches[pair] = nmatches;
}

}  // namespace

size_t projScratchWordsPerLast() { return kProjListCap + 1 + 4 + 2 + 1; }  // list, count, proj, levels, pushes (32-bit words per last-frame slot)

// All pointers are device pointers; arrays are [pair][cap...].  scratch: pairs * capLast * projScratchWordsPerLast() words;
// sorted: pairs * capCur words; cellStart: pairs * 3073 ints.  The caller has run gridSortKernel's batch form over kpsCur
// (launchGridSortBatch below does).
void launchSearchByProjection(const void* kpsLast, const uint8_t* lastFlags, const float* world, const uint8_t* mpDesc, const int* nLast,
                              const void* kpsCur, const uint8_t* descCur, const float* uRightCur, const uint8_t* takenCur, const int* nCur,
                              const float* poses, int pairs, int capLast, int capCur, const ProjCam& C, uint32_t* sorted, int* cellStart,
                              uint32_t* scratch, int* matchCur, int* nMatches, cudaStream_t s) {
    if (pairs <= 0) return;
    ProjArgs A{};
    A.kpsLast = static_cast<const float*>(kpsLast); A.lastFlags = lastFlags; A.world = world;
    A.mpDesc = reinterpret_cast<const uint4*>(mpDesc); A.nLast = nLast;
    A.kpsCur = static_cast<const float*>(kpsCur); A.descCur = reinterpret_cast<const uint4*>(descCur);
    A.uRightCur = uRightCur; A.takenCur = takenCur; A.nCur = nCur; A.poses = poses;
    A.sorted = sorted; A.cellStart = cellStart;
    const size_t slots = (size_t)pairs * capLast;
    A.candList = scratch;
    A.candCount = reinterpret_cast<int*>(scratch + slots * kProjListCap);
    A.proj = reinterpret_cast<float*>(scratch + slots * (kProjListCap + 1));
    A.levels = reinterpret_cast<int*>(scratch + slots * (kProjListCap + 5));
    A.pushes = scratch + slots * (kProjListCap + 7);
    A.matchCur = matchCur; A.nMatches = nMatches;
    A.capLast = capLast; A.capCur = capCur;
    A.fx = C.fx; A.fy = C.fy; A.cx = C.cx; A.cy = C.cy;
    A.minX = C.minX; A.maxX = C.maxX; A.minY = C.minY; A.maxY = C.maxY;
    A.invW = (float)kGridCols / (C.maxX - C.minX); A.invH = (float)kGridRows / (C.maxY - C.minY);  // Frame.cc:251-252
    A.mbf = C.mbf; A.mb = C.mb; A.th = C.th; A.mono = C.mono; A.checkOri = C.checkOri;
    for (int l = 0; l < kMaxLevels; ++l) A.sf[l] = l < C.nlevels ? C.sf[l] : 1.f;
    const long long warps = (long long)pairs * capLast;
    projCandKernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(A, pairs);
    countLaunch();
    projWalkKernel<<<pairs, 32, 0, s>>>(A);
    countLaunch();
}

size_t mapScratchWordsPerPoint() { return kProjListCap + 1 + 1; }  // list, count, radius

// Tracking::SearchLocalPoints' matcher for `pairs` independent (map point set, frame) problems; device pointers, [pair][cap...].
void launchSearchByProjectionMap(const float* track, const int* level, const uint8_t* flags, const uint8_t* mpDesc, const int* nMp,
                                 const void* kpsCur, const uint8_t* descCur, const float* uRightCur, const uint8_t* takenCur, const int* nCur,
                                 int pairs, int capMp, int capCur, const ProjCam& C, float nnratio, int farPoints, float thFar,
                                 uint32_t* sorted, int* cellStart, uint32_t* scratch, int* matchCur, int* nMatches, cudaStream_t s) {
    if (pairs <= 0) return;
    MapArgs A{};
    A.track = track; A.level = level; A.flags = flags; A.mpDesc = reinterpret_cast<const uint4*>(mpDesc); A.nMp = nMp;
    A.kpsCur = static_cast<const float*>(kpsCur); A.descCur = reinterpret_cast<const uint4*>(descCur);
    A.uRightCur = uRightCur; A.takenCur = takenCur; A.nCur = nCur; A.sorted = sorted; A.cellStart = cellStart;
    const size_t slots = (size_t)pairs * capMp;
    A.candList = scratch;
    A.candCount = reinterpret_cast<int*>(scratch + slots * kProjListCap);
    A.radius = reinterpret_cast<float*>(scratch + slots * (kProjListCap + 1));
    A.matchCur = matchCur; A.nMatches = nMatches; A.capMp = capMp; A.capCur = capCur;
    A.minX = C.minX; A.minY = C.minY;
    A.invW = (float)kGridCols / (C.maxX - C.minX); A.invH = (float)kGridRows / (C.maxY - C.minY);  // Frame.cc:251-252
    A.th = C.th; A.thFar = thFar; A.nnratio = nnratio; A.farPoints = farPoints;
    for (int l = 0; l < kMaxLevels; ++l) A.sf[l] = l < C.nlevels ? C.sf[l] : 1.f;
    const long long warps = (long long)pairs * capMp;
    mapCandKernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(A, pairs);
    countLaunch();
    mapWalkKernel<<<pairs, 32, 0, s>>>(A);
    countLaunch();
}

}  // namespace orbx

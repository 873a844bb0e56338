// k_quadtree.cu — K-c: ORBextractor::DistributeOctTree rebuilt as a block-parallel, round-synchronous split.
//
// Replaces DistributeOctTree (ORB_SLAM3/src/ORBextractor.cc:555-779), ExtractorNode::DivideNode (:480-536) and
// compareNodes (:538-553); semantics checklist in SURVEY.md App. B.  One CTA per (frame, level).
//
// The reference keeps a std::list of nodes, each owning a vector of keypoints, and rebuilds it node by node.
// Here nothing is moved: every candidate carries the LIST POSITION of the node that owns it (`nid`), a round
// renumbers the whole list at once, and one sweep over the candidates re-tags them.  Facts used:
//   * points inside a node keep their original relative order (DivideNode appends in order), so "first maximum
//     response in the node" = maximum of (response, -original index): an atomicMax on a packed key;
//   * a breadth pass pushes the children of every expandable node to the FRONT one by one, so the new list is
//     [children in reverse creation order] ++ [untouched nodes in their old order]  (App. D.3);
//   * the quadrant populations of a node are counted when the node is CREATED (same sweep), so the bookkeeping
//     of the next round (how many non-empty children, where `size >= N` is first reached) needs only scans;
//   * the only inherently sequential step is std::sort on (count, UL.x) in the second phase: ties are frequent
//     and libstdc++'s introsort is not stable, so its exact move sequence is replayed by one lane (App. D).
#include <cstdlib>

#include "orbx_internal.cuh"

namespace orbx {
namespace {

// threads per CTA = template parameter kQT of the kernel: 128 for the usual 1000-2000 feature settings (the co-running blur loses
// less), 256 for large per-level targets (launchQuadtree)

struct Box {
    short ulx, urx, uly, bry;
};

__device__ __forceinline__ int ceilHalf(int extent) {
    // (int)ceil((float)extent / 2)  (ORBextractor.cc:482-483); extents are small non-negative ints
    return (extent + 1) >> 1;
}

__device__ __forceinline__ int quadrantOf(uint32_t packed, const Box& b) {
    const int x = packed & 0xfff, y = (packed >> 12) & 0xfff;
    const int mx = b.ulx + ceilHalf(b.urx - b.ulx), my = b.uly + ceilHalf(b.bry - b.uly);
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);  // n1 UL, n2 UR, n3 BL, n4 BR (:513-527)
}

__device__ __forceinline__ Box childBox(const Box& b, int q) {
    const int mx = b.ulx + ceilHalf(b.urx - b.ulx), my = b.uly + ceilHalf(b.bry - b.uly);
    Box c;
    c.ulx = (q & 1) ? mx : b.ulx;
    c.urx = (q & 1) ? b.urx : mx;
    c.uly = (q & 2) ? my : b.uly;
    c.bry = (q & 2) ? b.bry : my;
    return c;
}

// In-place exclusive scan of a[0..n) by the whole CTA; returns the total to every thread.
template <int kQT>
__device__ int blockScanExclusive(int* a, int n, int* sPartial) {
    const int tid = threadIdx.x;
    const int per = (n + kQT - 1) / kQT;
    const int b = min(tid * per, n), e = min(b + per, n);
    int sum = 0;
    for (int i = b; i < e; ++i) sum += a[i];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) sPartial[tid >> 5] = incl;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kQT / 32; ++w) {
        if (w < (tid >> 5)) base += sPartial[w];
        total += sPartial[w];
    }
    int run = base + incl - sum;
    for (int i = b; i < e; ++i) {
        const int v = a[i];
        a[i] = run;
        run += v;
    }
    __syncthreads();
    return total;
}

// ---- libstdc++ std::sort (GCC 13 bits/stl_algo.h:1792-1950, bits/stl_heap.h) replayed on packed keys -------
// key = count << 32 | UL.x << 16 | list position; the comparator sees (count, UL.x) only (compareNodes).
__device__ __forceinline__ bool lessKey(unsigned long long a, unsigned long long b) { return (a >> 16) < (b >> 16); }

__device__ void adjustHeap(unsigned long long* f, int hole, int len, unsigned long long value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (lessKey(f[child], f[child - 1])) child--;
        f[hole] = f[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        f[hole] = f[child - 1];
        hole = child - 1;
    }
    int parent = (hole - 1) / 2;
    while (hole > top && lessKey(f[parent], value)) {
        f[hole] = f[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    f[hole] = value;
}

__device__ void heapSortRange(unsigned long long* f, int len) {
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            const unsigned long long v = f[parent];
            adjustHeap(f, parent, len, v);
            if (parent == 0) break;
            parent--;
        }
    }
    int last = len;
    while (last > 1) {
        --last;
        const unsigned long long v = f[last];
        f[last] = f[0];
        adjustHeap(f, 0, last, v);
    }
}

__device__ __forceinline__ void unguardedLinearInsert(unsigned long long* a, int last) {
    const unsigned long long v = a[last];
    int next = last - 1;
    while (lessKey(v, a[next])) {
        a[last] = a[next];
        last = next;
        --next;
    }
    a[last] = v;
}

__device__ void insertionSort(unsigned long long* a, int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (lessKey(a[i], a[first])) {
            const unsigned long long v = a[i];
            for (int k = i; k > first; --k) a[k] = a[k - 1];
            a[first] = v;
        } else {
            unguardedLinearInsert(a, i);
        }
    }
}

__device__ void stdSortReplay(unsigned long long* a, int n) {
    if (n == 0) return;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;
    // __introsort_loop with an explicit stack (the two sub-ranges are disjoint, so their order is immaterial)
    int stF[64], stL[64], stD[64], sp = 0;
    stF[0] = 0; stL[0] = n; stD[0] = 2 * lg; sp = 1;
    while (sp > 0) {
        --sp;
        int first = stF[sp], last = stL[sp], depth = stD[sp];
        while (last - first > 16) {
            if (depth == 0) {
                heapSortRange(a + first, last - first);
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;
            const int ia = first + 1, ib = mid, ic = last - 1;  // __move_median_to_first
            int m;
            if (lessKey(a[ia], a[ib])) {
                if (lessKey(a[ib], a[ic])) m = ib;
                else if (lessKey(a[ia], a[ic])) m = ic;
                else m = ia;
            } else if (lessKey(a[ia], a[ic])) m = ia;
            else if (lessKey(a[ib], a[ic])) m = ic;
            else m = ib;
            { const unsigned long long t = a[first]; a[first] = a[m]; a[m] = t; }
            int lo = first + 1, hi = last;  // __unguarded_partition
            while (true) {
                while (lessKey(a[lo], a[first])) ++lo;
                --hi;
                while (lessKey(a[first], a[hi])) --hi;
                if (!(lo < hi)) break;
                const unsigned long long t = a[lo]; a[lo] = a[hi]; a[hi] = t;
                ++lo;
            }
            stF[sp] = lo; stL[sp] = last; stD[sp] = depth; ++sp;  // introsort_loop(cut, last, depth)
            last = lo;
        }
    }
    if (n > 16) {
        insertionSort(a, 0, 16);
        for (int i = 16; i != n; ++i) unguardedLinearInsert(a, i);
    } else {
        insertionSort(a, 0, n);
    }
}
struct QtShared {
    Box* boxA; Box* boxB;
    int* cntA; int* cntB;
    int* ccA; int* ccB;
    int* scan; int* visit; int* cb; int* rk;
    short* map;
    unsigned long long* keys;
    int* cellOff;
};

#ifndef ORBX_QT_UNROLL
#define ORBX_QT_UNROLL 4
#endif
constexpr int kCandUnroll = ORBX_QT_UNROLL;

template <int kQT>
__global__ void __launch_bounds__(kQT) quadtreeKernel(const __grid_constant__ FramePlan P, const uint32_t* __restrict__ cand,
                                                      const int* __restrict__ cellCnt, uint32_t* __restrict__ gath,
                                                      uint16_t* __restrict__ nidAll, LevelKp* __restrict__ lkp,
                                                      int* __restrict__ lkpCnt, int* __restrict__ errFlag, int levelMajor, int NC,
                                                      int maxCells, int lBase, const int* __restrict__ gate, int gateMode,
                                                      int gateLevels) {
    extern __shared__ __align__(16) uint8_t smemRaw[];
    __shared__ int sPartial[kQT / 32];
    __shared__ int sRed;

    // levelMajor: the grid is (frames, levels) - every frame's level 0 first, then level 1, ...: the big levels start first and the
    // small ones fill the SMs as those drain (with (levels, frames) every wave of resident CTAs ends with its own level-0 CTAs)
    const int l = (levelMajor ? blockIdx.y : blockIdx.x) + lBase, f = levelMajor ? blockIdx.x : blockIdx.y, tid = threadIdx.x;
    // early / late pass of the split pipeline (launchQuadtree): gate[0..1] = the strips' overflow counts once the levels below gateLevels
    // were done.  Early pass (1): runs those levels unless a strip overflowed (their cells then wait for the per-cell kernel); late pass (2):
    // runs them only in that case
    if (gateMode) {
        const bool over = (gate[0] | gate[1]) != 0;
        if (gateMode == 1 ? over : (l < gateLevels && !over)) return;
    }
    const LevelPlan& L = P.lv[l];
    const int N = L.nFeat;

    QtShared S;
    {
        uint8_t* p = smemRaw;
        S.keys = reinterpret_cast<unsigned long long*>(p); p += sizeof(unsigned long long) * NC;
        S.boxA = reinterpret_cast<Box*>(p); p += sizeof(Box) * NC;
        S.boxB = reinterpret_cast<Box*>(p); p += sizeof(Box) * NC;
        S.map = reinterpret_cast<short*>(p); p += sizeof(short) * 4 * NC;
        S.cntA = reinterpret_cast<int*>(p); p += sizeof(int) * NC;
        S.cntB = reinterpret_cast<int*>(p); p += sizeof(int) * NC;
        S.ccA = reinterpret_cast<int*>(p); p += sizeof(int) * 4 * NC;
        S.ccB = reinterpret_cast<int*>(p); p += sizeof(int) * 4 * NC;
        S.scan = reinterpret_cast<int*>(p); p += sizeof(int) * NC;
        S.visit = reinterpret_cast<int*>(p); p += sizeof(int) * NC;
        S.cb = reinterpret_cast<int*>(p); p += sizeof(int) * NC;
        S.rk = reinterpret_cast<int*>(p); p += sizeof(int) * NC;
        S.cellOff = reinterpret_cast<int*>(p);
    }
    int* outCnt = lkpCnt + (size_t)f * P.nlevels + l;
    LevelKp* out = lkp + (size_t)f * P.lkpTotal + L.lkpBase;

    // ---- gather the cells' candidates into one list, reference order (ORBextractor.cc:863-868) ---------------
    const int nCells = L.nCols * L.nRows;
    const int* cc = cellCnt + (size_t)f * P.nCellsTotal + L.cellBase;
    for (int c = tid; c < nCells; c += kQT) S.cellOff[c] = cc[c];
    __syncthreads();
    const int n = nCells > 0 ? blockScanExclusive<kQT>(S.cellOff, nCells, sPartial) : 0;
    const int width = L.w - 2 * kMinBorder, height = L.h - 2 * kMinBorder;
    const int nIni = (height > 0 && width > 0) ? (int)roundf((float)width / (float)height) : 0;  // :559
    if (n == 0 || nIni <= 0) {
        if (tid == 0) *outCnt = 0;
        return;
    }
    uint32_t* g = gath + (size_t)f * P.gathWords + L.gathOff;
    uint16_t* nid = nidAll + (size_t)f * P.gathWords + L.gathOff;
    {
        const uint32_t* slots = cand + (size_t)f * P.candWords + L.candOff;
        const int warp = tid >> 5, lane = tid & 31;
        // four cells per trip, the loads of all four first (a cell rarely holds more than 32 candidates: the tail loop is for those)
        constexpr int kW = kQT / 32;
        for (int c0 = warp; c0 < nCells; c0 += 4 * kW) {
            int bb[4], cnt[4];
            uint32_t v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int c = c0 + u * kW;
                bb[u] = 0; cnt[u] = 0; v[u] = 0u;
                if (c < nCells) {
                    bb[u] = S.cellOff[c];
                    cnt[u] = ((c + 1 < nCells) ? S.cellOff[c + 1] : n) - bb[u];
                    if (lane < cnt[u]) v[u] = slots[(size_t)c * L.cellCap + lane];
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (lane < cnt[u]) g[bb[u] + lane] = v[u];
                for (int k = lane + 32; k < cnt[u]; k += 32) g[bb[u] + k] = slots[(size_t)(c0 + u * kW) * L.cellCap + k];
            }
        }
    }
    // ---- roots (:559-602) -----------------------------------------------------------------------------------
    const float hX = (float)width / (float)nIni;
    for (int r = tid; r < nIni; r += kQT) {
        Box b;
        b.ulx = (short)(int)(hX * (float)r);
        b.urx = (short)(int)(hX * (float)(r + 1));
        b.uly = 0;
        b.bry = (short)height;
        S.boxB[r] = b;
        S.cntB[r] = 0;
    }
    __syncthreads();
    for (int i0 = tid; i0 < n; i0 += kCandUnroll * kQT) {
        uint32_t pts[kCandUnroll];
#pragma unroll
        for (int u = 0; u < kCandUnroll; ++u) pts[u] = i0 + u * kQT < n ? g[i0 + u * kQT] : 0u;
#pragma unroll
        for (int u = 0; u < kCandUnroll; ++u) {
            const int i = i0 + u * kQT;
            if (i >= n) continue;
            int r = (int)((float)(pts[u] & 0xfff) / hX);  // :585
            r = min(r, nIni - 1);
            nid[i] = (uint16_t)r;
            atomicAdd(&S.cntB[r], 1);
        }
    }
    __syncthreads();
    for (int r = tid; r < nIni; r += kQT) S.scan[r] = S.cntB[r] > 0;
    __syncthreads();
    int size = blockScanExclusive<kQT>(S.scan, nIni, sPartial);
    for (int r = tid; r < nIni; r += kQT)
        if (S.cntB[r] > 0) {
            const int np = S.scan[r];
            S.boxA[np] = S.boxB[r];
            S.cntA[np] = S.cntB[r];
            S.ccA[4 * np] = S.ccA[4 * np + 1] = S.ccA[4 * np + 2] = S.ccA[4 * np + 3] = 0;
            S.map[4 * r] = (short)np;
        }
    __syncthreads();
    // (the loops over the candidates take four at a time, loads first: a thread's candidates are independent, but one global round trip
    // per candidate and round is what the length of a level-0 CTA was made of)
    for (int i0 = tid; i0 < n; i0 += kCandUnroll * kQT) {
        int pp[kCandUnroll];
        uint32_t pts[kCandUnroll];
#pragma unroll
        for (int u = 0; u < kCandUnroll; ++u) {
            const int i = i0 + u * kQT;
            pp[u] = i < n ? (int)nid[i] : 0;
            pts[u] = i < n ? g[i] : 0u;
        }
#pragma unroll
        for (int u = 0; u < kCandUnroll; ++u) {
            const int i = i0 + u * kQT;
            if (i < n) {
                const int np = S.map[4 * pp[u]];
                nid[i] = (uint16_t)np;
                if (S.cntA[np] > 1) atomicAdd(&S.ccA[4 * np + quadrantOf(pts[u], S.boxA[np])], 1);
            }
        }
    }
    __syncthreads();

    // ---- rounds -------------------------------------------------------------------------------------------------
    int phase = 1, cLast = 0;
    bool finish = false;
    while (!finish) {
        const int prevSize = size;
        int K = 0;
        for (int p = tid; p < size; p += kQT) S.rk[p] = -1;
        if (phase == 1) {  // breadth pass: every node with more than one point is split (:618-681)
            for (int p = tid; p < size; p += kQT) S.scan[p] = S.cntA[p] > 1;
            __syncthreads();
            K = blockScanExclusive<kQT>(S.scan, size, sPartial);
            for (int p = tid; p < size; p += kQT)
                if (S.cntA[p] > 1) { S.visit[S.scan[p]] = p; S.rk[p] = S.scan[p]; }
            __syncthreads();
        } else {  // (:692-753) expandable nodes of the previous round, sorted, split from the back until size >= N
            for (int j = tid; j < cLast; j += kQT) S.scan[j] = S.cntA[cLast - 1 - j] > 1;
            __syncthreads();
            const int m = blockScanExclusive<kQT>(S.scan, cLast, sPartial);
            for (int j = tid; j < cLast; j += kQT) {
                const int p = cLast - 1 - j;
                if (S.cntA[p] > 1)
                    S.keys[S.scan[j]] = ((unsigned long long)(uint32_t)S.cntA[p] << 32) |
                                        ((unsigned long long)(uint16_t)S.boxA[p].ulx << 16) | (unsigned long long)p;
            }
            __syncthreads();
            if (tid == 0) { stdSortReplay(S.keys, m); sRed = m; }
            __syncthreads();
            for (int r = tid; r < m; r += kQT) {
                const int p = (int)(S.keys[m - 1 - r] & 0xffff);
                S.cb[r] = (S.ccA[4 * p] > 0) + (S.ccA[4 * p + 1] > 0) + (S.ccA[4 * p + 2] > 0) + (S.ccA[4 * p + 3] > 0);
            }
            __syncthreads();
            // first visit index at which the list has reached N (:746-747)
            {
                const int per = (m + kQT - 1) / kQT;
                // inclusive sizes need the scan; do it on a copy kept in visit[]
                for (int r = tid; r < m; r += kQT) S.visit[r] = S.cb[r];
                __syncthreads();
                blockScanExclusive<kQT>(S.visit, m, sPartial);
                for (int r = tid; r < m; r += kQT) {
                    const int sz = size + S.visit[r] + S.cb[r] - (r + 1);
                    if (sz >= N) atomicMin(&sRed, r + 1);
                }
                (void)per;
                __syncthreads();
            }
            K = sRed;
            __syncthreads();
            for (int p = tid; p < size; p += kQT) S.scan[p] = 0;
            __syncthreads();
            for (int r = tid; r < K; r += kQT) {
                const int p = (int)(S.keys[m - 1 - r] & 0xffff);
                S.visit[r] = p;
                S.rk[p] = r;
                S.scan[p] = 1;
            }
            __syncthreads();
            blockScanExclusive<kQT>(S.scan, size, sPartial);
        }
        if (K == 0) break;  // nothing expandable: size == prevSize (:685, :750)
        for (int r = tid; r < K; r += kQT) {
            const int p = S.visit[r];
            S.cb[r] = (S.ccA[4 * p] > 0) + (S.ccA[4 * p + 1] > 0) + (S.ccA[4 * p + 2] > 0) + (S.ccA[4 * p + 3] > 0);
        }
        __syncthreads();
        const int C = blockScanExclusive<kQT>(S.cb, K, sPartial);
        const int newSize = C + size - K;
        if (newSize > NC) {  // cannot happen (size <= max(N + 3, 4 * nIni)); guard the shared arrays anyway
            if (tid == 0) atomicOr(errFlag, 2);
            break;
        }
        for (int p = tid; p < size; p += kQT) {
            const int r = S.rk[p];
            if (r >= 0) {
                int j = 0;
                const Box pb = S.boxA[p];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int cq = S.ccA[4 * p + q];
                    if (cq > 0) {
                        const int np = C - 1 - (S.cb[r] + j);
                        ++j;
                        S.boxB[np] = childBox(pb, q);
                        S.cntB[np] = cq;
                        S.ccB[4 * np] = S.ccB[4 * np + 1] = S.ccB[4 * np + 2] = S.ccB[4 * np + 3] = 0;
                        S.map[4 * p + q] = (short)np;
                    }
                }
            } else {
                const int np = C + (p - S.scan[p]);
                S.boxB[np] = S.boxA[p];
                S.cntB[np] = S.cntA[p];
#pragma unroll
                for (int q = 0; q < 4; ++q) S.ccB[4 * np + q] = S.ccA[4 * p + q];
                S.map[4 * p] = (short)np;
            }
        }
        __syncthreads();
        for (int i0 = tid; i0 < n; i0 += kCandUnroll * kQT) {
            int pp[kCandUnroll];
            uint32_t pts[kCandUnroll];
#pragma unroll
            for (int u = 0; u < kCandUnroll; ++u) {
                const int i = i0 + u * kQT;
                pp[u] = i < n ? (int)nid[i] : 0;
                pts[u] = i < n ? g[i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < kCandUnroll; ++u) {
                const int i = i0 + u * kQT;
                if (i >= n) continue;
                const int p = pp[u];
                if (S.rk[p] >= 0) {
                    const uint32_t pt = pts[u];
                    const int np = S.map[4 * p + quadrantOf(pt, S.boxA[p])];
                    nid[i] = (uint16_t)np;
                    if (S.cntB[np] > 1) atomicAdd(&S.ccB[4 * np + quadrantOf(pt, S.boxB[np])], 1);
                } else {
                    nid[i] = (uint16_t)S.map[4 * p];
                }
            }
        }
        __syncthreads();
        { Box* t = S.boxA; S.boxA = S.boxB; S.boxB = t; }
        { int* t = S.cntA; S.cntA = S.cntB; S.cntB = t; }
        { int* t = S.ccA; S.ccA = S.ccB; S.ccB = t; }
        size = newSize;
        cLast = C;
        if (size >= N || size == prevSize) {
            finish = true;
        } else if (phase == 1) {
            if (tid == 0) sRed = 0;
            __syncthreads();
            int mine = 0;
            for (int p = tid; p < C; p += kQT) mine += S.cntA[p] > 1;
            if (mine) atomicAdd(&sRed, mine);
            __syncthreads();
            const int nToExpand = sRed;
            __syncthreads();
            if (size + 3 * nToExpand > N) phase = 2;  // :689
        }
    }

    // ---- best response per node, first wins ties, list order (:757-778); then :880-890 ---------------------------
    unsigned* best = reinterpret_cast<unsigned*>(S.scan);
    for (int p = tid; p < size; p += kQT) best[p] = 0;
    __syncthreads();
    for (int i0 = tid; i0 < n; i0 += kCandUnroll * kQT) {
        int pp[kCandUnroll];
        uint32_t pts[kCandUnroll];
#pragma unroll
        for (int u = 0; u < kCandUnroll; ++u) {
            const int i = i0 + u * kQT;
            pp[u] = i < n ? (int)nid[i] : 0;
            pts[u] = i < n ? g[i] : 0u;
        }
#pragma unroll
        for (int u = 0; u < kCandUnroll; ++u) {
            const int i = i0 + u * kQT;
            if (i < n) atomicMax(&best[pp[u]], (pts[u] & 0xff000000u) | (0xffffffu - (unsigned)i));
        }
    }
    __syncthreads();
    for (int p = tid; p < size; p += kQT) {
        const uint32_t pt = g[0xffffffu - (best[p] & 0xffffffu)];
        LevelKp k;
        k.x = (short)((pt & 0xfff) + kMinBorder);
        k.y = (short)(((pt >> 12) & 0xfff) + kMinBorder);
        k.resp = (short)(pt >> 24);
        k.level = (short)l;
        if (p < L.lkpCap) out[p] = k;
    }
    if (tid == 0) {
        *outCnt = min(size, L.lkpCap);
        if (size > L.lkpCap) atomicOr(errFlag, 4);
    }
}

size_t qtSmem(int NC, int maxCells) { return (size_t)NC * 88 + (size_t)(maxCells + 1) * 4 + 64; }
int qtNodeCap(const FramePlan& P) {
    int nc = 8;
    for (int l = 0; l < P.nlevels; ++l) {
        const int width = P.lv[l].w - 2 * kMinBorder, height = P.lv[l].h - 2 * kMinBorder;
        const int nIni = (height > 0 && width > 0) ? (int)roundf((float)width / (float)height) : 0;
        nc = P.lv[l].nFeat + 8 > nc ? P.lv[l].nFeat + 8 : nc;
        // the first breadth pass splits EVERY multi-point root whatever N is (:610-681): up to 4 * nIni nodes
        nc = 4 * nIni + 8 > nc ? 4 * nIni + 8 : nc;
    }
    return nc;
}
int qtMaxCells(const FramePlan& P) {
    int m = 1;
    for (int l = 0; l < P.nlevels; ++l) m = P.lv[l].nCols * P.lv[l].nRows > m ? P.lv[l].nCols * P.lv[l].nRows : m;
    return m;
}
}  // namespace

size_t quadtreeSmemBytes(const FramePlan& P) { return qtSmem(qtNodeCap(P), qtMaxCells(P)); }

template <int QT>
static void launchQuadtreeT(const FramePlan& P, const Buffers& B, int F, int NC, int maxCells, size_t smem, cudaStream_t s, int lBase,
                            int nLv, const int* gate, int gateMode, int gateLevels) {
    if (smem > 40 * 1024) cudaFuncSetAttribute(quadtreeKernel<QT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // ORBX_QT_ORDER=0: the round-1 grid (levels, frames) for A/B runs
    static const int levelMajor = [] { const char* e = std::getenv("ORBX_QT_ORDER"); return e ? std::atoi(e) : 1; }();
    const dim3 grid = levelMajor ? dim3(F, nLv) : dim3(nLv, F);
    quadtreeKernel<QT><<<grid, QT, smem, s>>>(P, B.cand, B.cellCnt, B.gath, B.nid, B.lkp, B.lkpCnt, B.errFlag, levelMajor, NC, maxCells,
                                              lBase, gate, gateMode, gateLevels);
}

void launchQuadtree(const FramePlan& P, const Buffers& B, int F, cudaStream_t s, int lBase, int nLv, const int* gate, int gateMode,
                    int gateLevels) {
    if (nLv < 0) nLv = P.nlevels - lBase;
    if (nLv <= 0) return;
    const int NC = qtNodeCap(P), maxCells = qtMaxCells(P);
    const size_t smem = qtSmem(NC, maxCells);
    // CTA size by the largest per-level target: 128 threads up to 300 nodes per level (1000 features: 141.2 -> 144.7 k frames/s
    // against 256), 256 above (2000 features, stereo pairs: 54.5 k pairs/s with 128, 55.7 k with 256; 4000 features at 1080p:
    // 28.2 k frames/s with 128, 31.0 k with 256); ORBX_QT_THREADS overrides (A/B measurements)
    static const int forced = [] { const char* e = std::getenv("ORBX_QT_THREADS"); return e ? std::atoi(e) : 0; }();
    const int qt = forced ? forced : (NC <= 300 ? 128 : 256);
    if (qt <= 128) launchQuadtreeT<128>(P, B, F, NC, maxCells, smem, s, lBase, nLv, gate, gateMode, gateLevels);
    else launchQuadtreeT<256>(P, B, F, NC, maxCells, smem, s, lBase, nLv, gate, gateMode, gateLevels);
    countLaunch();
}
}  // namespace orbx

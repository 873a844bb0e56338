// k_bow.cu — row N4: Frame::ComputeBoW (ORB_SLAM3/src/Frame.cc:738-745) =
// DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>::transform(features, BowVector&, FeatureVector&, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1196) with the ORB vocabulary's settings (TF_IDF, L1_NORM).
//
// Two kernels per batch of frames:
//   bowDescendKernel   the tree descent of transform(feature, word_id, weight, nid, levelsup) (:1217-1263): sixteen lanes
//                      per feature, one child per lane (children's descriptors are stored contiguously in visiting order,
//                      so a node's k x 32 B are one coalesced read), FORB::distance (FORB.cpp:81-100) as 8 XOR + POPC,
//                      the first-minimum rule (`d < best_d`, children in order) as a min over keys distance << 20 | position;
//   bowCollectKernel   one CTA per frame: BowVector = std::map<WordId, double> filled by addWeight in feature order
//                      (BowVector.cpp:31-43) and L1-normalised in ascending word order (:62-84); FeatureVector =
//                      std::map<NodeId, vector<feature index>> (FeatureVector.cpp:33-47).  Both maps are ONE bitonic sort
//                      of 64-bit keys (id << 32 | feature index) each: runs of equal ids are the map entries, already in
//                      the maps' iteration order, their members in insertion order.
// Doubles are bit-exact: a word's value is its weight added to itself count-1 times in sequence (what addWeight does,
// every feature of a word adds the same weight), the norm is accumulated sequentially in ascending word order by one
// thread, and the division is IEEE.
#include <algorithm>

#include "orbx_internal.cuh"

namespace orbx {
namespace {

constexpr uint64_t kSentinel = ~0ull;

__global__ void __launch_bounds__(256) bowDescendKernel(const BowVocDev V, const uint8_t* __restrict__ desc,
                                                        const int* __restrict__ nPerFrame, int cap, int F, int levelsup,
                                                        int* __restrict__ leafOut, int* __restrict__ nidOut) {
    const int lane = threadIdx.x & 31, sub = lane & 15;
    const unsigned gmask = 0xffffu << (lane & 16);
    const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    if (g >= (long long)F * cap) return;  // whole groups leave together
    const int f = (int)(g / cap), i = (int)(g - (long long)f * cap);
    if (i >= nPerFrame[f]) return;
    const uint4* q4 = reinterpret_cast<const uint4*>(desc + (size_t)g * 32);
    const uint4 qa = __ldg(q4), qb = __ldg(q4 + 1);
    const int nidLevel = V.L - levelsup;
    int cur = 0, level = 0, nid = 0;
    for (;;) {
        const int cb = __ldg(V.childBegin + cur), ce = __ldg(V.childBegin + cur + 1);
        if (cb == ce) break;  // a leaf: isLeaf() = children.empty()
        ++level;
        unsigned best = 0xffffffffu;
        for (int p = cb + sub; p < ce; p += 16) {
            const uint4* c4 = reinterpret_cast<const uint4*>(V.childDesc + (size_t)p * 32);
            const uint4 a = __ldg(c4), b = __ldg(c4 + 1);
            const int d = __popc(a.x ^ qa.x) + __popc(a.y ^ qa.y) + __popc(a.z ^ qa.z) + __popc(a.w ^ qa.w) +
                          __popc(b.x ^ qb.x) + __popc(b.y ^ qb.y) + __popc(b.z ^ qb.z) + __popc(b.w ^ qb.w);
            best = min(best, ((unsigned)d << 20) | (unsigned)(p - cb));
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(gmask, best, o));
        cur = __ldg(V.childId + cb + (int)(best & 0xfffffu));
        if (level == nidLevel) nid = cur;
    }
    if (sub == 0) {
        leafOut[g] = cur;
        nidOut[g] = nid;
    }
}

// ascending bitonic sort of n2 (a power of two) keys in shared memory
__device__ void bitonicSort(uint64_t* k, int n2) {
    for (int size = 2; size <= n2; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const uint64_t a = k[lo], b = k[hi];
                if ((a > b) == up) {
                    k[lo] = b;
                    k[hi] = a;
                }
            }
        }
    }
    __syncthreads();
}

// Runs of equal high words in the sorted keys -> entry index of every head (exclusive count of heads before it).
// Each thread owns the contiguous chunk [t * per, (t + 1) * per); returns the total number of heads.
__device__ int headIndices(const uint64_t* k, int m, int n2, int* sScan, int& myFirst, int& myBegin, int& myEnd) {
    const int per = (n2 + blockDim.x - 1) / blockDim.x;
    myBegin = min((int)threadIdx.x * per, m);
    myEnd = min(myBegin + per, m);
    int cnt = 0;
    for (int p = myBegin; p < myEnd; ++p) cnt += (p == 0 || (uint32_t)(k[p] >> 32) != (uint32_t)(k[p - 1] >> 32));
    // block-wide exclusive scan of cnt
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) sScan[w] = incl;
    __syncthreads();
    if (w == 0) {
        int v = lane < nw ? sScan[lane] : 0, s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += u;
        }
        if (lane < nw) sScan[lane] = s - v;
        if (lane == 31) sScan[32] = s;
    }
    __syncthreads();
    myFirst = sScan[w] + incl - cnt;
    const int total = sScan[32];
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(512) bowCollectKernel(const BowVocDev V, const int* __restrict__ nPerFrame, int cap, int n2,
                                                        const int* __restrict__ leaf, const int* __restrict__ nidIn,
                                                        int32_t* __restrict__ wordIds, double* __restrict__ wordVals,
                                                        int32_t* __restrict__ nWords, int32_t* __restrict__ nodeIds,
                                                        int32_t* __restrict__ nodeBegin, int32_t* __restrict__ featIdx,
                                                        int32_t* __restrict__ nNodes) {
    extern __shared__ __align__(16) uint64_t keys[];
    __shared__ int sScan[33];
    __shared__ int sValid;
    __shared__ double sNorm;
    const int f = blockIdx.x, n = min(nPerFrame[f], cap);
    const size_t base = (size_t)f * cap;
    if (threadIdx.x == 0) sValid = 0;
    __syncthreads();

    // ---- BowVector ------------------------------------------------------------------------------------------
    int local = 0;
    for (int i = threadIdx.x; i < n2; i += blockDim.x) {
        uint64_t key = kSentinel;
        if (i < n) {
            const int node = leaf[base + i];
            if (V.weight[node] > 0) {  // "not stopped" (TemplatedVocabulary.h:1157)
                key = ((uint64_t)(uint32_t)V.wordOf[node] << 32) | (uint32_t)i;
                ++local;
            }
        }
        keys[i] = key;
    }
    if (local) atomicAdd(&sValid, local);
    bitonicSort(keys, n2);
    const int m = sValid;
    int first, b, e;
    const int nu = headIndices(keys, m, n2, sScan, first, b, e);
    for (int p = b; p < e; ++p) {
        const uint32_t word = (uint32_t)(keys[p] >> 32);
        if (p != 0 && word == (uint32_t)(keys[p - 1] >> 32)) continue;
        int q = p + 1;
        while (q < m && (uint32_t)(keys[q] >> 32) == word) ++q;
        const double w = V.weight[leaf[base + (uint32_t)keys[p]]];
        double sum = w;  // addWeight: insert w, then += w for every further feature of the word, in feature order
        for (int c = p + 1; c < q; ++c) sum = __dadd_rn(sum, w);
        wordIds[base + first] = (int32_t)word;
        wordVals[base + first] = sum;
        ++first;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double norm = 0.0;  // BowVector::normalize(L1): ascending word order, one accumulator
        for (int u = 0; u < nu; ++u) norm = __dadd_rn(norm, fabs(wordVals[base + u]));
        sNorm = norm;
        nWords[f] = nu;
    }
    __syncthreads();
    const double norm = sNorm;
    if (norm > 0.0)
        for (int u = threadIdx.x; u < nu; u += blockDim.x) wordVals[base + u] = __ddiv_rn(wordVals[base + u], norm);
    __syncthreads();

    // ---- FeatureVector --------------------------------------------------------------------------------------
    for (int i = threadIdx.x; i < n2; i += blockDim.x) {
        uint64_t key = kSentinel;
        if (i < n && V.weight[leaf[base + i]] > 0) key = ((uint64_t)(uint32_t)nidIn[base + i] << 32) | (uint32_t)i;
        keys[i] = key;
    }
    bitonicSort(keys, n2);
    const int nn = headIndices(keys, m, n2, sScan, first, b, e);
    const size_t nbBase = (size_t)f * (cap + 1);
    for (int p = b; p < e; ++p) {
        featIdx[base + p] = (int32_t)(uint32_t)keys[p];
        if (p == 0 || (uint32_t)(keys[p] >> 32) != (uint32_t)(keys[p - 1] >> 32)) {
            nodeIds[base + first] = (int32_t)(uint32_t)(keys[p] >> 32);
            nodeBegin[nbBase + first] = p;
            ++first;
        }
    }
    if (threadIdx.x == 0) {
        nodeBegin[nbBase + nn] = m;
        nNodes[f] = nn;
    }
}


// ---- ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)  (ORB_SLAM3/src/ORBmatcher.cc:223-425), monocular ----------
// Features are only compared inside the same vocabulary node.  A frame feature belongs to exactly one node, so the
// reference's sequential bookkeeping (a frame feature that already received a MapPoint is skipped, :279) never crosses
// nodes: ONE WARP PER KEYFRAME NODE finds its partner in the frame's FeatureVector (binary search in the ascending node ids),
// then walks the node's keyframe features in order; the lanes take the node's frame features 32 at a time, keys
// distance << 16 | position reproduce the strict-< best / second updates (:287-296), lane 0 applies the accept rule
// (:323-327) and records the match and its rotation bin (:345-352).
struct BowSearchArgs {
    const uint4* descKF; const uint8_t* validKF; const float* angleKF;
    const int* kfNodeIds; const int* kfNodeBegin; const int* kfFeatIdx; int kfNodes;
    const uint4* descF; const float* angleF;
    const int* fNodeIds; const int* fNodeBegin; const int* fFeatIdx; int fNodes;
    float nnratio; int checkOrientation;
    int* matchF; int* binF;
    // keyframe <-> keyframe form (ORBmatcher.cc:765-905): candidates of the second side need a good MapPoint (validF == 1, :826-830)
    // and the accept rule is strict in TH_LOW (:848); validF == NULL, strictLow == 0 is the keyframe -> frame form
    const uint8_t* validF; int strictLow;
};

__global__ void __launch_bounds__(256) bowSearchKernel(const BowSearchArgs A) {
    const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (a >= A.kfNodes) return;
    const int node = A.kfNodeIds[a];
    int lo = 0, hi = A.fNodes;  // lower_bound of the keyframe's node id in the frame's ascending node ids
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (A.fNodeIds[mid] < node) lo = mid + 1; else hi = mid;
    }
    if (lo >= A.fNodes || A.fNodeIds[lo] != node) return;
    const int fbeg = A.fNodeBegin[lo], fend = A.fNodeBegin[lo + 1];
    volatile int* matchF = A.matchF;
    constexpr unsigned kNone = (256u << 16) | 0xffffu;  // bestDist = 256, no index
    for (int iKF = A.kfNodeBegin[a]; iKF < A.kfNodeBegin[a + 1]; ++iKF) {
        const int realKF = A.kfFeatIdx[iKF];
        if (A.validKF[realKF] != 1) continue;  // no MapPoint, or a bad one (:255-259)
        const uint4 q0 = A.descKF[2 * realKF], q1 = A.descKF[2 * realKF + 1];
        unsigned k1 = kNone, k2 = kNone;
        for (int iF = fbeg + lane; iF < fend; iF += 32) {
            const int realF = A.fFeatIdx[iF];
            if (matchF[realF] >= 0) continue;  // vpMapPointMatches[realIdxF] already set (:279) / vbMatched2 (:826)
            if (A.validF && A.validF[realF] != 1) continue;
            const uint4 b0 = A.descF[2 * realF], b1 = A.descF[2 * realF + 1];
            const int d = __popc(q0.x ^ b0.x) + __popc(q0.y ^ b0.y) + __popc(q0.z ^ b0.z) + __popc(q0.w ^ b0.w) +
                          __popc(q1.x ^ b1.x) + __popc(q1.y ^ b1.y) + __popc(q1.z ^ b1.z) + __popc(q1.w ^ b1.w);
            const unsigned key = ((unsigned)d << 16) | (unsigned)(iF - fbeg);
            k2 = min(k2, max(k1, key));
            k1 = min(k1, key);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned o1 = __shfl_xor_sync(0xffffffffu, k1, o), o2 = __shfl_xor_sync(0xffffffffu, k2, o);
            k2 = min(max(k1, o1), min(k2, o2));
            k1 = min(k1, o1);
        }
        const int best1 = (int)(k1 >> 16), best2 = (int)(k2 >> 16);
        if ((A.strictLow ? best1 < 50 : best1 <= 50) && (float)best1 < __fmul_rn(A.nnratio, (float)best2)) {  // TH_LOW, mfNNratio (:323-325, :848-850)
            if (lane == 0) {
                const int bestF = A.fFeatIdx[fbeg + (int)(k1 & 0xffffu)];
                matchF[bestF] = realKF;
                if (A.checkOrientation) {
                    float rot = __fsub_rn(A.angleKF[realKF], A.angleF[bestF]);
                    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                    int bin = (int)roundf(__fmul_rn(rot, 1.0f / 30));  // factor = 1.0f / HISTO_LENGTH (:234)
                    if (bin == 30) bin = 0;
                    A.binF[bestF] = bin;
                }
            }
        }
        __syncwarp();  // the match is visible to the warp's next keyframe feature
    }
}

// Rotation consistency (:403-420) and the final count: one CTA.  ComputeThreeMaxima (:2012-2053) only needs the bin sizes.
__global__ void __launch_bounds__(256) bowOrientationKernel(int* __restrict__ matchF, const int* __restrict__ binF, int nF,
                                                            int checkOrientation, int* __restrict__ nMatches) {
    __shared__ int hist[30];
    __shared__ int keep[3];
    __shared__ int count;
    if (threadIdx.x < 30) hist[threadIdx.x] = 0;
    if (threadIdx.x == 0) count = 0;
    __syncthreads();
    if (checkOrientation) {
        for (int j = threadIdx.x; j < nF; j += blockDim.x)
            if (matchF[j] >= 0) atomicAdd(&hist[binF[j]], 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < 30; i++) {
                const int s = hist[i];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
                else if (s > max3) { max3 = s; ind3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
            keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
        }
        __syncthreads();
    }
    int local = 0;
    for (int j = threadIdx.x; j < nF; j += blockDim.x) {
        if (matchF[j] < 0) continue;
        if (checkOrientation) {
            const int b = binF[j];
            if (b != keep[0] && b != keep[1] && b != keep[2]) { matchF[j] = -1; continue; }
        }
        ++local;
    }
    if (local) atomicAdd(&count, local);
    __syncthreads();
    if (threadIdx.x == 0) *nMatches = count;
}

}  // namespace

void launchBowSearch(const uint8_t* descKF, const uint8_t* validKF, const float* angleKF, const int* kfNodeIds, const int* kfNodeBegin,
                     const int* kfFeatIdx, int kfNodes, const uint8_t* descF, int nF, const float* angleF, const int* fNodeIds,
                     const int* fNodeBegin, const int* fFeatIdx, int fNodes, float nnratio, int checkOrientation, int* matchF, int* binF,
                     int* nMatches, cudaStream_t s, const uint8_t* validF, int strictLow) {
    cudaMemsetAsync(matchF, 0xff, (size_t)std::max(nF, 1) * 4, s);
    if (kfNodes > 0 && fNodes > 0 && nF > 0) {
        const BowSearchArgs A{reinterpret_cast<const uint4*>(descKF), validKF, angleKF, kfNodeIds, kfNodeBegin, kfFeatIdx, kfNodes,
                              reinterpret_cast<const uint4*>(descF), angleF, fNodeIds, fNodeBegin, fFeatIdx, fNodes, nnratio,
                              checkOrientation, matchF, binF, validF, strictLow};
        bowSearchKernel<<<(kfNodes * 32 + 255) / 256, 256, 0, s>>>(A);
        countLaunch();
    }
    bowOrientationKernel<<<1, 256, 0, s>>>(matchF, binF, nF, checkOrientation, nMatches);
    countLaunch();
}

int bowSortCapacity() { return 16384; }

void launchBowTransform(const BowVocDev& V, const uint8_t* dDesc, const int* dN, int cap, int F, int levelsup, int* leaf,
                        int* nid, int32_t* wordIds, double* wordVals, int32_t* nWords, int32_t* nodeIds, int32_t* nodeBegin,
                        int32_t* featIdx, int32_t* nNodes, cudaStream_t s) {
    const long long groups = (long long)F * cap;
    const int grid = (int)((groups * 16 + 255) / 256);
    bowDescendKernel<<<grid, 256, 0, s>>>(V, dDesc, dN, cap, F, levelsup, leaf, nid);
    int n2 = 32;
    while (n2 < cap) n2 <<= 1;
    const size_t smem = (size_t)n2 * sizeof(uint64_t);
    if (smem > 40 * 1024) cudaFuncSetAttribute(bowCollectKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    bowCollectKernel<<<F, 512, smem, s>>>(V, dN, cap, n2, leaf, nid, wordIds, wordVals, nWords, nodeIds, nodeBegin, featIdx, nNodes);
    countLaunch(2);
}

}  // namespace orbx

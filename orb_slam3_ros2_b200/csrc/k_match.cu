// k_match.cu — K-g: brute-force top-2 Hamming matching of 256-bit ORB descriptors.
//
// Replaces the inner loop of every reference matcher — ORBmatcher::DescriptorDistance
// (ORB_SLAM3/src/ORBmatcher.cc:2058-2074) inside the strict-< best/second bookkeeping (:83-120, :273-294) — and
// cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) (src/Frame.cc:1144).  Result order is (distance asc, index asc),
// which both the reference loops (candidates visited in ascending index) and knnMatch produce.
//
// Bound: the POPC pipe (8 x 32-bit POPC per pair), not HBM — the database is streamed once per query tile.
// Layout: a CTA stages a block of database rows in shared memory with 128-bit loads; every thread keeps R query
// descriptors in registers, reads the staged rows as warp-wide broadcasts and maintains two packed keys per
// query, key = distance << 23 | row-in-chunk, so that `min` reproduces the tie order:
//     k2 = min(k2, max(k1, key));  k1 = min(k1, key)
// Chunks of the database (< 2^23 rows) go to different CTAs; their per-chunk top-2 are written as 64-bit keys
// (distance << 40 | global row) and reduced by mergeKernel — the same kernel that merges the per-rank candidates
// after the NCCL all-gather of the sharded path.
#include <algorithm>
#include <cstdlib>

#include "orbx_internal.cuh"


namespace orbx {
namespace {

constexpr int kMT = 256;        // threads per CTA
constexpr int kStage = 256;     // database rows staged per iteration (8 KB)
constexpr uint32_t kNoKey32 = 0xffffffffu;
constexpr unsigned long long kNoKey64 = ~0ull;

__device__ __forceinline__ unsigned long long widenKey(uint32_t k, long long base) {
    if (k == kNoKey32) return kNoKey64;
    return ((unsigned long long)(k >> 23) << 40) | (unsigned long long)(base + (long long)(k & 0x7fffffu));
}

// 256-bit Hamming distance with 5 POPC instead of 8: three carry-save adders (sum = x^y^z, carry = maj(x,y,z), one
// LOP3 each) fold the eight XOR words into two words of weight 1 and three of weight 2.  POPC issues at 16/clk/SM,
// LOP3 at 64/clk/SM, so trading 3 POPC for 6 LOP3 lifts the POPC-pipe bound from 2 to 3.2 pairs/clk/SM.  Measured on
// B200 (cfg5): 8 POPC 545 Gpairs/s, 5 POPC 630, 4 POPC (one more adder level) 612, 5 POPC + branch-on-improvement 626.
__device__ __forceinline__ int hamming256(const uint32_t (&q)[8], const uint4& a, const uint4& b) {
    const uint32_t x0 = q[0] ^ a.x, x1 = q[1] ^ a.y, x2 = q[2] ^ a.z, x3 = q[3] ^ a.w;
    const uint32_t x4 = q[4] ^ b.x, x5 = q[5] ^ b.y, x6 = q[6] ^ b.z, x7 = q[7] ^ b.w;
    const uint32_t s0 = x0 ^ x1 ^ x2, c0 = (x0 & x1) | (x2 & (x0 | x1));
    const uint32_t s1 = x3 ^ x4 ^ x5, c1 = (x3 & x4) | (x5 & (x3 | x4));
    const uint32_t s2 = x6 ^ x7 ^ s0, c2 = (x6 & x7) | (s0 & (x6 | x7));
    return __popc(s1) + __popc(s2) + 2 * (__popc(c0) + __popc(c1) + __popc(c2));
}

// The packed key of one pair, distance << 23 | row, with everything but the logic and the counts on the FMA pipe.  ncu on the loop
// that used hamming256() + shift + or: ALU pipe 84 %, XU (POPC) 84 %, FMA 6 % - the ALU pipe carried 21 operations per pair (8 XOR,
// 6 carry-save LOP3, IADD3, OR, 3 min / max and 2 of the compiler's own loop-bound recomputation).  Here the weight-2 and weight-1
// sums stay apart and enter the key through two multiply-adds with the multipliers 2^24 and 2^23 held in registers the compiler
// cannot see through (a shift + or would go back to the ALU pipe): key = t * 2^24 + (u * 2^23 + row).  ALU operations per pair: 17.
// Measured on B200: cfg5 630 -> 666 Gpairs/s, and 702 once matchKernel may use 126 registers (launch bound 1 CTA: ptxas had settled on 64
// registers with 72 bytes of spills); whole step 164.6 -> 169.0 k frames/s, most of it from the explicit trip counts.  One more adder on the
// three weight-2 words (4 POPC + 16 LOP3): 637 Gpairs/s, 166.8 k frames/s - rejected again.
#ifndef ORBX_MATCH_FMAKEY
#define ORBX_MATCH_FMAKEY 1
#endif
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xe8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t madLo(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// (read from __constant__ memory: ptxas folds multipliers it can see - an immediate, a `mov` - back into LEA / IADD3 on the ALU pipe)
__constant__ uint32_t gKeyMul[3] = {1u << 23, 1u << 24, 1u};
struct KeyMul {
    uint32_t m23, m24, one;
    __device__ __forceinline__ KeyMul() : m23(gKeyMul[0]), m24(gKeyMul[1]), one(gKeyMul[2]) {}
};
__device__ __forceinline__ uint32_t pairKey(const uint32_t (&q)[8], const uint4& a, const uint4& b, uint32_t row, const KeyMul& M) {
#if ORBX_MATCH_FMAKEY
    // the adders are written as lop3 instructions: left to itself the compiler re-associates the XOR chains through the raw q / db words
    // (with 4 queries per thread it issued 20 LOP3 per pair instead of the 14 written here)
    const uint32_t x0 = q[0] ^ a.x, x1 = q[1] ^ a.y, x2 = q[2] ^ a.z, x3 = q[3] ^ a.w;
    const uint32_t x4 = q[4] ^ b.x, x5 = q[5] ^ b.y, x6 = q[6] ^ b.z, x7 = q[7] ^ b.w;
    const uint32_t s0 = xor3(x0, x1, x2), c0 = maj3(x0, x1, x2);
    const uint32_t s1 = xor3(x3, x4, x5), c1 = maj3(x3, x4, x5);
    const uint32_t s2 = xor3(x6, x7, s0), c2 = maj3(x6, x7, s0);
    const uint32_t u = madLo(__popc(s1), M.one, __popc(s2));                               // weight 1
    const uint32_t t = madLo(__popc(c0), M.one, madLo(__popc(c1), M.one, __popc(c2)));     // weight 2
    return madLo(t, M.m24, madLo(u, M.m23, row));
#else
    return ((uint32_t)hamming256(q, a, b) << 23) | row;
#endif
}

struct MatchProblem {
    const uint8_t* q;    // nq x 32
    const uint8_t* db;   // nd x 32
    int nq;
    long long nd;
};

// Problem z: either the single (q, db) pair, or — consecutive mode — frame z against frame z+1 of a batch.
#ifndef ORBX_MATCH_MINB
#define ORBX_MATCH_MINB 1
#endif
template <int R>
__global__ void __launch_bounds__(kMT, ORBX_MATCH_MINB) matchKernel(const uint8_t* __restrict__ q, int nq, const uint8_t* __restrict__ db,
                                                   long long nd, long long indexBase, long long chunkRows,
                                                   unsigned long long* __restrict__ part1,
                                                   unsigned long long* __restrict__ part2, const int* __restrict__ counts,
                                                   int cap) {
    __shared__ __align__(16) uint4 sDb[kStage * 2];
    const int tid = threadIdx.x;
    const int z = blockIdx.z;
    if (counts) {  // consecutive frames of a batch: q = frame z, db = frame z+1
        q += (size_t)z * cap * 32;
        db = q + (size_t)cap * 32;
        nq = counts[z];
        nd = counts[z + 1];
    }
    const int q0 = blockIdx.y * kMT * R;
    if (q0 >= nq) return;
    const long long c0 = (long long)blockIdx.x * chunkRows;
    const long long c1 = c0 + chunkRows < nd ? c0 + chunkRows : nd;

    const KeyMul M;
    constexpr int kUnroll = R >= 4 ? 2 : 4;
    uint32_t qw[R][8];
    uint32_t k1[R], k2[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int qi = q0 + r * kMT + tid;
        const uint4* p = reinterpret_cast<const uint4*>(q + (size_t)(qi < nq ? qi : 0) * 32);
        const uint4 a = qi < nq ? __ldg(p) : make_uint4(0, 0, 0, 0), b = qi < nq ? __ldg(p + 1) : make_uint4(0, 0, 0, 0);
        qw[r][0] = a.x; qw[r][1] = a.y; qw[r][2] = a.z; qw[r][3] = a.w;
        qw[r][4] = b.x; qw[r][5] = b.y; qw[r][6] = b.z; qw[r][7] = b.w;
        k1[r] = k2[r] = kNoKey32;
    }

    for (long long s0 = c0; s0 < c1; s0 += kStage) {
        const int rows = (int)(c1 - s0 < kStage ? c1 - s0 : kStage);
        __syncthreads();
        const uint4* src = reinterpret_cast<const uint4*>(db + (size_t)s0 * 32);
        for (int i = tid; i < rows * 2; i += kMT) sDb[i] = __ldg(src + i);
        __syncthreads();
        const uint32_t idx0 = (uint32_t)(s0 - c0);
        auto onePair = [&](int j) {
            const uint4 a = sDb[2 * j], b = sDb[2 * j + 1];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint32_t key = pairKey(qw[r], a, b, idx0 + (uint32_t)j, M);
                k2[r] = min(k2[r], max(k1[r], key));
                k1[r] = min(k1[r], key);
            }
        };
        // the trip counts are explicit: with `j < rows` the compiler recomputed min(c1 - s0, kStage) in every iteration
        const int rowsU = rows & ~(kUnroll - 1);
        for (int j = 0; j < rowsU; j += kUnroll) {
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) onePair(j + u);
        }
        for (int j = rowsU; j < rows; ++j) onePair(j);
    }
    const size_t nqPad = counts ? (size_t)cap : (size_t)nq;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int qi = q0 + r * kMT + tid;
        if (qi < nq) {
            const size_t o = ((size_t)z * gridDim.x + blockIdx.x) * nqPad + qi;
            part1[o] = widenKey(k1[r], indexBase + c0);
            part2[o] = widenKey(k2[r], indexBase + c0);
        }
    }
}

// Global top-2 over `nsets` candidate sets laid out [set][query]; keys are distance << 40 | global row, so the
// merged order is (distance asc, global row asc) — knnMatch's tie rule across shard boundaries.
__global__ void mergeKernel(const unsigned long long* __restrict__ k1, const unsigned long long* __restrict__ k2,
                            int nsets, int nq, size_t setStride, long long* __restrict__ idx1, int* __restrict__ d1,
                            long long* __restrict__ idx2, int* __restrict__ d2, unsigned long long* __restrict__ outK1,
                            unsigned long long* __restrict__ outK2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    unsigned long long b1 = kNoKey64, b2 = kNoKey64;
    for (int s = 0; s < nsets; ++s) {
        const unsigned long long a1 = k1[(size_t)s * setStride + i], a2 = k2[(size_t)s * setStride + i];
        const unsigned long long lo = b1 < a1 ? b1 : a1, hi = b1 < a1 ? a1 : b1;
        const unsigned long long m2 = b2 < a2 ? b2 : a2;
        b1 = lo;
        b2 = hi < m2 ? hi : m2;
    }
    if (outK1) {
        outK1[i] = b1;
        outK2[i] = b2;
        return;
    }
    const unsigned long long mask = (1ull << 40) - 1;
    idx1[i] = b1 == kNoKey64 ? -1 : (long long)(b1 & mask);
    d1[i] = b1 == kNoKey64 ? 256 : (int)(b1 >> 40);
    idx2[i] = b2 == kNoKey64 ? -1 : (long long)(b2 & mask);
    d2[i] = b2 == kNoKey64 ? 256 : (int)(b2 >> 40);
}

// Consecutive frames of a batch in ONE kernel (v2): frame z against frame z + 1, every thread keeps R queries, a CTA walks
// ALL rows of frame z + 1 (no database chunks, so no 64-bit partial keys and no merge pass) and writes the int32 results
// itself; 128-thread CTAs, 8 q-tiles x 255 pairs all resident at once.  Slots [n, cap) get the "no match" values (-1, 256).
#ifndef ORBX_MATCHC_MINB
#define ORBX_MATCHC_MINB 8
#endif
template <int R, int NT>
__global__ void __launch_bounds__(NT, R == 1 ? ORBX_MATCHC_MINB : 1) matchConsecutiveKernel(const uint8_t* __restrict__ desc, const int* __restrict__ counts, int cap,
                                                             int* __restrict__ idx1, int* __restrict__ d1, int* __restrict__ idx2,
                                                             int* __restrict__ d2) {
    __shared__ __align__(16) uint4 sDb[kStage * 2];
    const int tid = threadIdx.x;
    const int z = blockIdx.y;
    const uint8_t* q = desc + (size_t)z * cap * 32;
    const uint8_t* db = q + (size_t)cap * 32;
    const int nq = min(counts[z], cap), nd = min(counts[z + 1], cap);
    const int q0 = blockIdx.x * NT * R;

    const KeyMul M;
    constexpr int kUnroll = R >= 4 ? 2 : 4;
    uint32_t qw[R][8];
    uint32_t k1[R], k2[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int qi = q0 + r * NT + tid;
        const uint4* p = reinterpret_cast<const uint4*>(q + (size_t)(qi < nq ? qi : 0) * 32);
        const uint4 a = qi < nq ? __ldg(p) : make_uint4(0, 0, 0, 0), b = qi < nq ? __ldg(p + 1) : make_uint4(0, 0, 0, 0);
        qw[r][0] = a.x; qw[r][1] = a.y; qw[r][2] = a.z; qw[r][3] = a.w;
        qw[r][4] = b.x; qw[r][5] = b.y; qw[r][6] = b.z; qw[r][7] = b.w;
        k1[r] = k2[r] = kNoKey32;
    }
    if (q0 < nq) {
        for (int s0 = 0; s0 < nd; s0 += kStage) {
            const int rows = min(nd - s0, kStage);
            __syncthreads();
            const uint4* src = reinterpret_cast<const uint4*>(db + (size_t)s0 * 32);
            for (int i = tid; i < rows * 2; i += NT) sDb[i] = __ldg(src + i);
            __syncthreads();
            auto onePair = [&](int j) {
                const uint4 a = sDb[2 * j], b = sDb[2 * j + 1];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const uint32_t key = pairKey(qw[r], a, b, (uint32_t)(s0 + j), M);
                    k2[r] = min(k2[r], max(k1[r], key));
                    k1[r] = min(k1[r], key);
                }
            };
            const int rowsU = rows & ~(kUnroll - 1);
            for (int j = 0; j < rowsU; j += kUnroll) {
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) onePair(j + u);
            }
            for (int j = rowsU; j < rows; ++j) onePair(j);
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int qi = q0 + r * NT + tid;
        if (qi < cap) {
            const size_t o = (size_t)z * cap + qi;
            const bool v1 = qi < nq && k1[r] != kNoKey32, v2 = qi < nq && k2[r] != kNoKey32;
            idx1[o] = v1 ? (int)(k1[r] & 0x7fffffu) : -1;
            d1[o] = v1 ? (int)(k1[r] >> 23) : 256;
            idx2[o] = v2 ? (int)(k2[r] & 0x7fffffu) : -1;
            d2[o] = v2 ? (int)(k2[r] >> 23) : 256;
        }
    }
}

// consecutive mode epilogue: keys [pair][cap] -> int32 outputs [pair][cap]
__global__ void unpackPairsKernel(const unsigned long long* __restrict__ k1, const unsigned long long* __restrict__ k2,
                                  const int* __restrict__ counts, int cap, int* __restrict__ idx1, int* __restrict__ d1,
                                  int* __restrict__ idx2, int* __restrict__ d2) {
    const int z = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const size_t o = (size_t)z * cap + i;
    const bool valid = i < counts[z];
    const unsigned long long a = valid ? k1[o] : kNoKey64, b = valid ? k2[o] : kNoKey64;
    const unsigned long long mask = (1ull << 40) - 1;
    idx1[o] = a == kNoKey64 ? -1 : (int)(a & mask);
    d1[o] = a == kNoKey64 ? 256 : (int)(a >> 40);
    idx2[o] = b == kNoKey64 ? -1 : (int)(b & mask);
    d2[o] = b == kNoKey64 ? 256 : (int)(b >> 40);
}

// Frame::ComputeStereoMatches' Hamming stage (src/Frame.cc:825-894): one thread per left keypoint, right keypoints
// visited in ascending iR (the order the row buckets are filled in, :828-838).
__global__ void stereoBandedKernel(const float* __restrict__ kl, const uint4* __restrict__ dl, int nl,
                                   const float* __restrict__ kr, const uint4* __restrict__ dr, int nr,
                                   const float* __restrict__ sf, int nRows, float maxD, int* __restrict__ bestIdx,
                                   int* __restrict__ bestDist, const StereoBatch SB) {
    if (SB.nl) {  // pair blockIdx.y of a batch
        const size_t o = (size_t)blockIdx.y * SB.cap;
        nl = min(SB.nl[blockIdx.y], SB.cap); nr = min(SB.nr[blockIdx.y], SB.cap);
        kl += o * 7; kr += o * 7; dl += o * 2; dr += o * 2; bestIdx += o; bestDist += o;
    }
    // The row band, octave and column of a right keypoint do not depend on the left one: a CTA works them out once per tile of 256 right
    // keypoints (shared memory, int4 = minr, maxr, octave, uR) and every thread then walks the tile with one broadcast load and five
    // compares per candidate instead of three global loads, ceilf / floorf and a multiply (right keypoints in ascending index, strict <).
    __shared__ int4 sR[256];
    const int iL = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = iL < nl;
    int bd = 100, bi = 0;  // ORBmatcher::TH_HIGH, bestIdxR = 0 (:867-868)
    float uL = 0.f, vL = 0.f;
    int levelL = 0;
    if (live) { uL = kl[iL * 7]; vL = kl[iL * 7 + 1]; levelL = __float_as_int(kl[iL * 7 + 5]); }
    const int row = (int)vL;  // vRowIndices[vL] (:858)
    const float minU = __fsub_rn(uL, maxD), maxU = uL;
    const bool search = live && row >= 0 && row < nRows && !(maxU < 0);
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (search) { a0 = dl[2 * iL]; a1 = dl[2 * iL + 1]; }
    for (int t0 = 0; t0 < nr; t0 += 256) {
        const int nt = min(256, nr - t0);
        __syncthreads();
        for (int j = threadIdx.x; j < nt; j += blockDim.x) {
            const int iR = t0 + j;
            const float uR = kr[iR * 7], vR = kr[iR * 7 + 1];
            const int octR = __float_as_int(kr[iR * 7 + 5]);
            const float r = __fmul_rn(2.0f, sf[octR]);
            const int maxr = (int)ceilf(__fadd_rn(vR, r)), minr = (int)floorf(__fsub_rn(vR, r));  // :833-834
            sR[j] = make_int4(minr, maxr, octR, __float_as_int(uR));
        }
        __syncthreads();
        if (!search) continue;
        for (int j = 0; j < nt; ++j) {
            const int4 e = sR[j];
            if (row < e.x || row > e.y) continue;
            if (e.z < levelL - 1 || e.z > levelL + 1) continue;
            const float uR = __int_as_float(e.w);
            if (!(uR >= minU && uR <= maxU)) continue;
            const int iR = t0 + j;
            const uint4 b0 = dr[2 * iR], b1 = dr[2 * iR + 1];
            const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                          __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
            if (d < bd) { bd = d; bi = iR; }
        }
    }
    if (live) {
        bestIdx[iL] = bi;
        bestDist[iL] = bd;
    }
}


// ---- the same search over ROW LISTS (Frame.cc:820-838 builds vRowIndices the same way): a right keypoint enters the list of every image
// row of its band; a left keypoint then only meets the right keypoints of its own row (about 2 % of them on KITTI frames).  The lists
// are filled with atomics, i.e. in no particular order: the result does not depend on it, because "first strict minimum in ascending
// index" == the minimum of distance << 20 | index.  Scratch per pair: cnt / start / fill [nRows + 1] ints, the right keypoints' records
// (minr, maxr, octave, uR) and the lists (cap x bandMax ints).
struct StereoRows {
    int* cnt; int* start; int* fill;  // [pairs][nRows + 1]
    int4* rt;                          // [pairs][cap]
    int* list;                         // [pairs][cap * bandMax]
    int nRows, cap, bandMax;
};
__global__ void stereoBandKernel(const float* __restrict__ kr, int nr, const float* __restrict__ sf, const StereoBatch SB, StereoRows R, int fillPass) {
    const int pair = blockIdx.y;
    if (SB.nr) { nr = min(SB.nr[pair], SB.cap); kr += (size_t)pair * SB.cap * 7; }
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nr) return;
    int4 e;
    if (!fillPass) {
        const float uR = kr[j * 7], vR = kr[j * 7 + 1];
        const int octR = __float_as_int(kr[j * 7 + 5]);
        const float r = __fmul_rn(2.0f, sf[octR]);
        e = make_int4((int)floorf(__fsub_rn(vR, r)), (int)ceilf(__fadd_rn(vR, r)), octR, __float_as_int(uR));  // :833-834
        R.rt[(size_t)pair * R.cap + j] = e;
    } else {
        e = R.rt[(size_t)pair * R.cap + j];
    }
    const int r0 = max(e.x, 0), r1 = min(e.y, R.nRows - 1);
    int* cnt = (fillPass ? R.fill : R.cnt) + (size_t)pair * (R.nRows + 1);
    const int* start = R.start + (size_t)pair * (R.nRows + 1);
    int* list = R.list + (size_t)pair * R.cap * R.bandMax;
    for (int r = r0; r <= r1; ++r) {
        const int k = atomicAdd(&cnt[r], 1);
        if (fillPass) list[start[r] + k] = j;
    }
}
__global__ void __launch_bounds__(256) stereoRowScanKernel(StereoRows R) {
    __shared__ int sPart[256];
    const int pair = blockIdx.x, tid = threadIdx.x;
    const int* cnt = R.cnt + (size_t)pair * (R.nRows + 1);
    int* start = R.start + (size_t)pair * (R.nRows + 1);
    const int per = (R.nRows + 255) / 256, b = min(tid * per, R.nRows), e = min(b + per, R.nRows);
    int sum = 0;
    for (int r = b; r < e; ++r) sum += cnt[r];
    sPart[tid] = sum;
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int t = 0; t < 256; ++t) { const int v = sPart[t]; sPart[t] = acc; acc += v; }
        start[R.nRows] = acc;
    }
    __syncthreads();
    int acc = sPart[tid];
    for (int r = b; r < e; ++r) { start[r] = acc; acc += cnt[r]; }
}
// one WARP per left keypoint: the lanes take the candidates of its row 32 at a time (list entry -> record -> descriptor are three
// dependent global loads per candidate; one thread per keypoint walked them one after the other: 45 us for a single KITTI pair)
__global__ void __launch_bounds__(128) stereoRowMatchKernel(const float* __restrict__ kl, const uint4* __restrict__ dl, int nl,
                                                            const uint4* __restrict__ dr, float maxD, int* __restrict__ bestIdx,
                                                            int* __restrict__ bestDist, const StereoBatch SB, StereoRows R) {
    const int pair = blockIdx.y;
    if (SB.nl) {
        const size_t o = (size_t)pair * SB.cap;
        nl = min(SB.nl[pair], SB.cap);
        kl += o * 7; dl += o * 2; dr += o * 2; bestIdx += o; bestDist += o;
    }
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (iL >= nl) return;  // whole warps
    const float uL = kl[iL * 7], vL = kl[iL * 7 + 1];
    const int levelL = __float_as_int(kl[iL * 7 + 5]);
    const int row = (int)vL;  // vRowIndices[vL] (:858)
    const float minU = __fsub_rn(uL, maxD), maxU = uL;
    uint32_t best = (100u << 20);  // ORBmatcher::TH_HIGH, bestIdxR = 0 (:867-868)
    if (row >= 0 && row < R.nRows && !(maxU < 0)) {
        const uint4 a0 = dl[2 * iL], a1 = dl[2 * iL + 1];
        const int* start = R.start + (size_t)pair * (R.nRows + 1);
        const int* list = R.list + (size_t)pair * R.cap * R.bandMax;
        const int4* rt = R.rt + (size_t)pair * R.cap;
        const int tb = start[row], te = start[row + 1];
        for (int t = tb + lane; t < te; t += 32) {
            const int iR = list[t];
            const int4 e = rt[iR];
            if (e.z < levelL - 1 || e.z > levelL + 1) continue;
            const float uR = __int_as_float(e.w);
            if (!(uR >= minU && uR <= maxU)) continue;
            const uint4 b0 = dr[2 * iR], b1 = dr[2 * iR + 1];
            const uint32_t d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                               __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
            best = min(best, (d << 20) | (uint32_t)iR);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) {
        bestIdx[iL] = (int)(best & 0xfffffu);
        bestDist[iL] = (int)(best >> 20);
    }
}

}  // namespace

// How the database is cut into chunks: enough CTAs to fill the GPU a few times over, chunk < 2^23 rows.
static long long chooseChunk(long long nd, int queryTiles, int sms) {
    long long want = ((long long)sms * 4 + queryTiles - 1) / queryTiles;  // CTAs along the db axis
    if (want < 1) want = 1;
    long long rows = (nd + want - 1) / want;
    rows = ((rows + kStage - 1) / kStage) * kStage;
    if (rows < kStage) rows = kStage;
    if (rows > (1 << 23) - kStage) rows = (1 << 23) - kStage;
    return rows;
}

int matchNumChunks(int nq, int64_t nd, int sms) {
    const int tiles = (nq + kMT * 4 - 1) / (kMT * 4);
    const long long rows = chooseChunk(nd, tiles, sms);
    return (int)((nd + rows - 1) / rows);
}

// Per-chunk top-2 keys of (q x db) into dKeys1/2 laid out [chunk][nq]; returns the number of chunks via the layout
// computed by matchNumChunks (the caller sized the buffers with it).
void launchMatchTop2(const uint8_t* dq, int nq, const uint8_t* ddb, int64_t nd, int64_t base, uint64_t* dKeys1,
                     uint64_t* dKeys2, cudaStream_t s) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int tiles = (nq + kMT * 4 - 1) / (kMT * 4);
    const long long rows = chooseChunk(nd, tiles, sms);
    const int chunks = (int)((nd + rows - 1) / rows);
    if (chunks <= 0 || nq <= 0) return;
    matchKernel<4><<<dim3(chunks, tiles, 1), kMT, 0, s>>>(dq, nq, ddb, nd, base, rows,
                                                          reinterpret_cast<unsigned long long*>(dKeys1),
                                                          reinterpret_cast<unsigned long long*>(dKeys2), nullptr, 0);
    countLaunch();
}

void launchUnpackTop2Strided(const uint64_t* k1, const uint64_t* k2, int nsets, int nq, size_t setStride, int64_t* idx1,
                             int32_t* d1, int64_t* idx2, int32_t* d2, cudaStream_t s) {
    if (nq <= 0) return;
    mergeKernel<<<(nq + 127) / 128, 128, 0, s>>>(reinterpret_cast<const unsigned long long*>(k1),
                                                 reinterpret_cast<const unsigned long long*>(k2), nsets, nq, setStride,
                                                 reinterpret_cast<long long*>(idx1), d1,
                                                 reinterpret_cast<long long*>(idx2), d2, nullptr, nullptr);
    countLaunch();
}

void launchUnpackTop2(const uint64_t* k1, const uint64_t* k2, int nsets, int nq, int64_t* idx1, int32_t* d1,
                      int64_t* idx2, int32_t* d2, cudaStream_t s) {
    launchUnpackTop2Strided(k1, k2, nsets, nq, (size_t)nq, idx1, d1, idx2, d2, s);
}

// [set][nq] keys -> one candidate set (still keys): the per-rank payload of the all-gather
void launchReduceKeys(const uint64_t* k1, const uint64_t* k2, int nsets, int nq, uint64_t* outK1, uint64_t* outK2,
                      cudaStream_t s) {
    if (nq <= 0) return;
    mergeKernel<<<(nq + 127) / 128, 128, 0, s>>>(reinterpret_cast<const unsigned long long*>(k1),
                                                 reinterpret_cast<const unsigned long long*>(k2), nsets, nq, (size_t)nq,
                                                 nullptr, nullptr, nullptr, nullptr,
                                                 reinterpret_cast<unsigned long long*>(outK1),
                                                 reinterpret_cast<unsigned long long*>(outK2));
    countLaunch();
}

// frame f vs frame f+1 for f = 0..F-2; descriptors [F][cap][32], counts [F]; keys scratch [F-1][cap] x 2
void launchMatchConsecutive(const uint8_t* dDesc, const int* dN, int F, int cap, uint64_t* keys1, uint64_t* keys2,
                            int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2, cudaStream_t s) {
    if (F < 2) return;
    // ORBX_MATCH_R: queries per thread of the one-kernel form (1, 2, 4; default 1); 0 = the two-kernel form (matchKernel<1> + unpack).
    // Measured (frames/s of the whole step): R = 1 158.2 k, R = 2 145.1 k, R = 4 146.2 k, two-kernel form 156.9 k - the loop is bound by
    // the POPC (XU 84 %) and ALU (80 %) pipes, not by issue slots, so sharing the row loads between queries buys nothing and costs occupancy
    static const int rSel = [] { const char* e = std::getenv("ORBX_MATCH_R"); return e ? std::atoi(e) : 1; }();
    // ORBX_MATCH_NT: threads (= queries) per CTA of the R = 1 form (64, 96; default 128).  Measured, whole step: 166.2 k / 165.7 k / 168.0 k
    // frames/s - CTAs beyond a frame's count skip the walk, so 128-thread CTAs waste only the 17 idle slots of the eighth one
    static const int ntSel = [] { const char* e = std::getenv("ORBX_MATCH_NT"); return e ? std::atoi(e) : 128; }();
    if (rSel == 1 && (ntSel == 64 || ntSel == 96)) {
        const dim3 grid((cap + ntSel - 1) / ntSel, F - 1);
        if (ntSel == 64) matchConsecutiveKernel<1, 64><<<grid, 64, 0, s>>>(dDesc, dN, cap, idx1, d1, idx2, d2);
        else matchConsecutiveKernel<1, 96><<<grid, 96, 0, s>>>(dDesc, dN, cap, idx1, d1, idx2, d2);
        countLaunch();
        return;
    }
    if (rSel == 1 || rSel == 2 || rSel == 4) {
        constexpr int NT = 128;
        const int tiles = (cap + NT * rSel - 1) / (NT * rSel);
        const dim3 grid(tiles, F - 1);
        if (rSel == 1) matchConsecutiveKernel<1, NT><<<grid, NT, 0, s>>>(dDesc, dN, cap, idx1, d1, idx2, d2);
        else if (rSel == 2) matchConsecutiveKernel<2, NT><<<grid, NT, 0, s>>>(dDesc, dN, cap, idx1, d1, idx2, d2);
        else matchConsecutiveKernel<4, NT><<<grid, NT, 0, s>>>(dDesc, dN, cap, idx1, d1, idx2, d2);
        countLaunch();
        return;
    }
    const int tiles = (cap + kMT - 1) / kMT;
    matchKernel<1><<<dim3(1, tiles, F - 1), kMT, 0, s>>>(dDesc, 0, nullptr, 0, 0, (long long)cap,
                                                         reinterpret_cast<unsigned long long*>(keys1),
                                                         reinterpret_cast<unsigned long long*>(keys2), dN, cap);
    countLaunch();
    unpackPairsKernel<<<dim3((cap + 127) / 128, F - 1), 128, 0, s>>>(reinterpret_cast<const unsigned long long*>(keys1),
                                                                    reinterpret_cast<const unsigned long long*>(keys2),
                                                                    dN, cap, idx1, d1, idx2, d2);
    countLaunch();
}

size_t stereoRowScratchBytes(int pairs, int cap, int nRows, int bandMax) {
    return (size_t)pairs * ((size_t)3 * (nRows + 1) * 4 + (size_t)cap * 16 + (size_t)cap * bandMax * 4) + 256;
}

void launchStereoBanded(const void* kl, const uint8_t* dl, int nl, const void* kr, const uint8_t* dr, int nr,
                        const float* sf, int nRows, float maxD, int32_t* bi, int32_t* bd, cudaStream_t s,
                        const StereoBatch* batch, void* rowScratch, int bandMax) {
    const StereoBatch SB = batch ? *batch : StereoBatch{};
    if (batch) nl = batch->cap;
    if (nl <= 0) return;
    const int pairs = batch ? batch->frames : 1, cap = batch ? batch->cap : std::max(nr, 1);
    // ORBX_STEREO_ROWS=0: the all-pairs kernel (A/B runs); it is also what runs without scratch or with 2^20 keypoints or more
    static const bool rows = [] { const char* e = std::getenv("ORBX_STEREO_ROWS"); return !(e && e[0] == '0'); }();
    if (rows && rowScratch && bandMax > 0 && cap < (1 << 20) && nRows > 0) {
        StereoRows R;
        uint8_t* p = static_cast<uint8_t*>(rowScratch);
        const size_t rowInts = (size_t)pairs * (nRows + 1);
        R.cnt = reinterpret_cast<int*>(p); R.start = R.cnt + rowInts; R.fill = R.start + rowInts;
        p += ((3 * rowInts * 4 + 15) / 16) * 16;
        R.rt = reinterpret_cast<int4*>(p); p += (size_t)pairs * cap * 16;
        R.list = reinterpret_cast<int*>(p);
        R.nRows = nRows; R.cap = cap; R.bandMax = bandMax;
        cudaMemsetAsync(R.cnt, 0, 3 * rowInts * 4, s);
        if (!batch && nr <= 0) {  // no right keypoints: every list is empty
            stereoRowMatchKernel<<<dim3((nl + 3) / 4, 1), 128, 0, s>>>(static_cast<const float*>(kl), reinterpret_cast<const uint4*>(dl), nl,
                                                                         reinterpret_cast<const uint4*>(dr), maxD, bi, bd, SB, R);
            countLaunch();
            return;
        }
        const dim3 gR((cap + 255) / 256, pairs);
        stereoBandKernel<<<gR, 256, 0, s>>>(static_cast<const float*>(kr), nr, sf, SB, R, 0);
        stereoRowScanKernel<<<pairs, 256, 0, s>>>(R);
        stereoBandKernel<<<gR, 256, 0, s>>>(static_cast<const float*>(kr), nr, sf, SB, R, 1);
        stereoRowMatchKernel<<<dim3((nl + 3) / 4, pairs), 128, 0, s>>>(static_cast<const float*>(kl), reinterpret_cast<const uint4*>(dl), nl,
                                                                           reinterpret_cast<const uint4*>(dr), maxD, bi, bd, SB, R);
        countLaunch(4);
        return;
    }
    stereoBandedKernel<<<dim3((nl + 127) / 128, batch ? batch->frames : 1), 128, 0, s>>>(
        static_cast<const float*>(kl), reinterpret_cast<const uint4*>(dl), nl, static_cast<const float*>(kr),
        reinterpret_cast<const uint4*>(dr), nr, sf, nRows, maxD, bi, bd, SB);
    countLaunch();
}

}  // namespace orbx

// ---- POPC issue-rate microbenchmark: the denominator of the matcher's roofline -----------------------------------
// Every thread runs 8 independent XOR+POPC chains (the shape of one descriptor pair) with nothing else in the loop
// but the accumulation; the result is the chip's sustained 32-bit POPC rate (pairs/s = POPC/s / 8).
namespace orbx {
namespace {
__global__ void __launch_bounds__(256) popcBenchKernel(uint32_t seed, int iters, uint32_t* __restrict__ out) {
    uint32_t a[8], acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = seed * (threadIdx.x + 1u + 977u * i) + blockIdx.x; acc[i] = 0; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            acc[i] += __popc(a[i] ^ (uint32_t)it);
            a[i] += acc[i];  // keeps the chain data dependent on its own history so nothing is hoisted
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    if (s == 0x12345u) out[0] = s;  // practically never; defeats dead-code elimination
}
}  // namespace

// returns giga-POPC per second, or a negative value on failure
double benchPopcRate(cudaStream_t s) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t* d = nullptr;
    if (cudaMalloc(&d, 16) != cudaSuccess) return -1.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 1 << 15, blocks = sms * 8;
    popcBenchKernel<<<blocks, 256, 0, s>>>(12345u, 1 << 10, d);  // warm-up
    double best = -1.0;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0, s);
        popcBenchKernel<<<blocks, 256, 0, s>>>(777u + rep, iters, d);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double rate = (double)blocks * 256.0 * iters * 8.0 / (ms * 1e-3) / 1e9;
        if (rate > best) best = rate;
    }
    countLaunch(4);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    return cudaGetLastError() == cudaSuccess ? best : -1.0;
}
}  // namespace orbx

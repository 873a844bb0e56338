// k_describe.cu — K-d IC_Angle, K-f rBRIEF-256 and K-h the ordered output placement of operator().
//
// Replaces, in ORB_SLAM3/src/ORBextractor.cc:
//   IC_Angle :76-103 / computeOrientation :471-478     intensity centroid on the UN-blurred level, cv::fastAtan2
//   computeOrbDescriptor :107-146 / computeDescriptors :1077-1084   rotated 256-pair test on the blurred level
//   operator() tail :1103-1167    pt *= scale for level > 0, lapping-area front/back placement, monoIndex
//
// Float determinism: fastAtan2 is evaluated with explicitly rounded single
// operations (no FMA); cos/sin are glibc's sincosf algorithm (double range reduction + polynomials, each
// operation rounded once); the sample rotation is two rounded products and one rounded sum in STRICT mode, or
// fmaf(x, b, y*a) in NATIVE mode (what gcc -O3 -march=native emits for the reference).  cvRound = rint.
#include <algorithm>

#include <cstdlib>

#include "orbx_internal.cuh"
#include "tma.cuh"

namespace orbx {
namespace {

#include "orb_pattern.inc"  // host table kOrbPattern[256][4]



// cv::fastAtan2 (OpenCV core/mathfuncs_core, scalar path), degrees in [0,360)
__device__ __forceinline__ float fastAtan2Deg(float y, float x) {
    const float s = (float)(180.0 / 3.1415926535897932384626433832795);
    const float p1 = __fmul_rn(0.9997878412794807f, s), p3 = __fmul_rn(-0.3258083974640975f, s);
    const float p5 = __fmul_rn(0.1555786518463281f, s), p7 = __fmul_rn(-0.04432655554792128f, s);
    const float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// glibc 2.28+ sincosf (sysdeps/ieee754/flt-32/s_sincosf.[ch]), the |y| < 120 paths (y in [0, 2*pi) here)
__device__ __forceinline__ void glibcSincosf(float y, float* sinp, float* cosp) {
    const double HPI_INV = 0x1.45f306dc9c883p+23, HPI = 0x1.921fb54442d18p+0;
    const double C0 = 1.0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
    double x = (double)y;
    int n = 0;
    double s = 1.0, k = 1.0;
    if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ff)) {
        if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ff)) {
            *sinp = y;
            *cosp = 1.0f;
            return;
        }
    } else {
        const double r = __dmul_rn(x, HPI_INV);
        n = ((int)r + 0x800000) >> 24;
        x = __dsub_rn(x, __dmul_rn((double)n, HPI));
        s = (n & 3) == 1 || (n & 3) == 2 ? -1.0 : 1.0;
        k = (n & 2) ? -1.0 : 1.0;
    }
    const double xs = __dmul_rn(x, s), x2 = __dmul_rn(x, x), x3 = __dmul_rn(x2, xs), x4 = __dmul_rn(x2, x2);
    const double x5 = __dmul_rn(x3, x2), x6 = __dmul_rn(x4, x2);
    const double c2 = __dadd_rn(k * C3, __dmul_rn(x2, k * C4));
    const double s1 = __dadd_rn(S2, __dmul_rn(x2, S3));
    const double c1 = __dadd_rn(k * C0, __dmul_rn(x2, k * C1));
    const double sn = __dadd_rn(xs, __dmul_rn(x3, S1));
    const double c = __dadd_rn(c1, __dmul_rn(x4, k * C2));
    const float Sf = (float)__dadd_rn(sn, __dmul_rn(x5, s1));
    const float Cf = (float)__dadd_rn(c, __dmul_rn(x6, c2));
    if (n & 1) { *sinp = Cf; *cosp = Sf; } else { *sinp = Sf; *cosp = Cf; }
}

// ---- K-h part 1: output slot of every level keypoint (operator() :1120-1166) ---------------------------------
// The reference walks levels 0..n-1 and, inside a level, the quadtree's list order; a keypoint whose SCALED x lies
// in [lap0, lap1] is written from the back (stereoIndex--), the others from the front (monoIndex++).
__global__ void __launch_bounds__(256) slotKernel(const __grid_constant__ FramePlan P, const LevelKp* __restrict__ lkp,
                                                  const int* __restrict__ lkpCnt, LevelKp* __restrict__ sorted,
                                                  int* __restrict__ outN, int* __restrict__ outMono, int lap0, int lap1,
                                                  int* __restrict__ errFlag) {
    __shared__ int sBase[kMaxLevels + 1];
    __shared__ int sWarp[8];
    __shared__ int sCarry;
    const int f = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) {
        int acc = 0;
        for (int l = 0; l < P.nlevels; ++l) { sBase[l] = acc; acc += lkpCnt[(size_t)f * P.nlevels + l]; }
        sBase[P.nlevels] = acc;
        sCarry = 0;
    }
    __syncthreads();
    const int total = sBase[P.nlevels];
    const LevelKp* kp = lkp + (size_t)f * P.lkpTotal;
    LevelKp* sl = sorted + (size_t)f * P.outCap;  // keypoints in operator()'s output order
    for (int start = 0; start < total; start += 256) {
        const int j = start + tid;
        int inside = 0, l = 0, idx = 0;
        if (j < total) {
            while (j >= sBase[l + 1]) ++l;
            idx = P.lv[l].lkpBase + (j - sBase[l]);
            float x = (float)kp[idx].x;
            if (l != 0) x = __fmul_rn(x, P.lv[l].scale);  // :1149-1151
            inside = (x >= (float)lap0 && x <= (float)lap1) ? 1 : 0;  // :1153
        }
        int incl = inside;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if ((tid & 31) >= o) incl += v;
        }
        if ((tid & 31) == 31) sWarp[tid >> 5] = incl;
        __syncthreads();
        int base = sCarry;
        for (int w = 0; w < (tid >> 5); ++w) base += sWarp[w];
        const int insideBefore = base + incl - inside;
        if (j < total) {
            const int dst = inside ? (total - 1 - insideBefore) : (j - insideBefore);
            if (dst < P.outCap) sl[dst] = kp[idx];
        }
        __syncthreads();
        if (tid == 255) sCarry = base + incl;
        __syncthreads();
    }
    if (tid == 0) {
        outN[f] = min(total, P.outCap);
        outMono[f] = total - sCarry;
        if (total > P.outCap) atomicOr(errFlag, 8);
    }
}

// ---- K-d + K-f + K-h part 2: one warp per keypoint -----------------------------------------------------------
// IC_Angle: the 31 x 31 bounding box of the disc is read as 31 rows x 8 aligned-and-shifted 32-bit words (every row of
// the 128-B-pitched plane has the same misalignment, so one funnel shift amount serves the whole patch); each word
// goes through two byte dot products against per-(row, word) coefficient words — u (signed, 0 outside the disc) for
// m10 and the 0/1 disc mask for the row sum of m01.  248 words / 32 lanes = 8 per lane, all loads independent.
constexpr int kPatchR = 18, kPatchRows = 2 * kPatchR + 1;
// 80-byte rows (5 x 16 B, 4 used): 20 words per row spread the rows over 8 bank phases; a 64-byte pitch has only 2
constexpr int kPatchPitch = 80;
__device__ uint2 gIcCoef[248];   // x: s8 u-coefficients, y: u8 disc mask, entry = row * 8 + word
__device__ float4 gPatternF[256];

__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

#ifndef ORBX_DESC_MINB
#define ORBX_DESC_MINB 4
#endif
// the pipelined kernel: 3 CTAs per SM = 80 registers (measured, whole step: 165.4 k frames/s against 162.7 k at 4 CTAs / 64 registers and
// 164.0 k at 2 CTAs; the kernel alone is 0.255 ms either way at 3 or 4)
#ifndef ORBX_DESC_ASYNC_MINB
#define ORBX_DESC_ASYNC_MINB 3
#endif
__global__ void __launch_bounds__(256, ORBX_DESC_MINB) describeKernel(const __grid_constant__ FramePlan P, const uint8_t* __restrict__ pyr,
                                                      const uint8_t* __restrict__ blur, const LevelKp* __restrict__ sorted,
                                                      const int* __restrict__ outN, float* __restrict__ outKps,
                                                      uint8_t* __restrict__ outDesc, int F) {
    // Persistent CTAs: the pattern and the disc coefficients are copied to shared memory once, then every warp walks
    // output slots (frame, s) with a grid-wide stride, fetching the NEXT slot's record while it works on the current one.
    __shared__ float4 sPat[256];
    __shared__ uint2 sCoef[248];
    __shared__ __align__(16) uint8_t sPatch[8][kPatchRows * kPatchPitch];
    // transposed: pair 8*lane + k lives at [k][lane], so the 32 lanes of a warp read 32 consecutive float4 (the
    // [lane][k] order put every lane on the same banks: an 8-way conflict on each of the 8 loads per keypoint)
    sPat[(threadIdx.x & 7) * 32 + (threadIdx.x >> 3)] = gPatternF[threadIdx.x];
    if (threadIdx.x < 248) sCoef[threadIdx.x] = gIcCoef[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    uint8_t* patch = sPatch[threadIdx.x >> 5];
    const int total = F * P.outCap, stride = gridDim.x * 8;
    int w = blockIdx.x * 8 + (threadIdx.x >> 5);
    int nNext = 0;
    LevelKp kpNext = {0, 0, 0, 0};
    if (w < total) {
        const int f0 = w / P.outCap;
        nNext = __ldg(outN + f0);
        kpNext = sorted[w];
    }
    for (; w < total; w += stride) {
    const int f = w / P.outCap, s = w - f * P.outCap;
    const int n = nNext;
    const LevelKp kp = kpNext;
    if (w + stride < total) {  // prefetch the next slot's record
        nNext = __ldg(outN + (w + stride) / P.outCap);
        kpNext = sorted[w + stride];
    }
    if (s >= n) continue;
    const int l = kp.level;
    const LevelPlan& L = P.lv[l];
    __syncwarp();  // the previous iteration's samples are done with `patch`

    // ---- the blurred 37 x 37 patch the 512 samples can touch (|offset| <= round(13 * sqrt 2) = 18) is fetched now, as
    // 37 rows x 4 aligned 16-byte chunks, together with the orientation patch: one round trip to memory instead of two,
    // and the samples then come from shared memory.  (Sampling the plane directly costs one L1 tag lookup per distinct
    // 128-B line per request — ~25 per 32-lane byte gather — which made the L1 pipeline the kernel's bottleneck.)
    const int bx0 = (kp.x - kPatchR) & ~15;   // blurred-plane column of the patch's first byte
    uint4 pv[5];
    {
        const uint8_t* src = blur + (size_t)f * P.blurBytes + L.blurOff + (size_t)(kp.y - kPatchR) * L.bpitch + bx0;
#pragma unroll
        for (int it = 0; it < 5; ++it) {
            const int item = it * 32 + lane;  // row = item >> 2, chunk = item & 3
            pv[it] = item < kPatchRows * 4 ? __ldg(reinterpret_cast<const uint4*>(src + (size_t)((item >> 2) * L.bpitch)) + (item & 3))
                                           : make_uint4(0, 0, 0, 0);
        }
    }

    // ---- IC_Angle (ORBextractor.cc:76-103) -----------------------------------------------------------------------
    int m10 = 0, m01 = 0;
    {
        const uint8_t* plane = pyr + (size_t)f * P.frameBytes + L.planeOff;
        const int px = kp.x + kEdge - kHalfPatch;                 // plane column of u = -15
        const uint8_t* p0 = plane + (size_t)(kp.y + kEdge - kHalfPatch) * L.pitch + (px & ~3);
        const int sh = (px & 3) * 8;
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int item = it * 32 + lane;                      // row = item >> 3, word = item & 7
            const uint32_t* q = reinterpret_cast<const uint32_t*>(p0 + (size_t)((item >> 3) * L.pitch)) + (item & 7);
            const bool ok = item < 248;
            lo[it] = ok ? __ldg(q) : 0u;
            hi[it] = ok ? __ldg(q + 1) : 0u;
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int item = it * 32 + lane;
            const uint2 cf = sCoef[item < 248 ? item : 0];
            const uint32_t w = __funnelshift_r(lo[it], hi[it], sh);
            m10 = dp4a_us(w, cf.x, m10);
            m01 += ((item >> 3) - kHalfPatch) * (int)__dp4a(w, cf.y, 0u);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fastAtan2Deg((float)m01, (float)m10);

    // ---- rBRIEF: lane i computes byte i (pairs 8i .. 8i+7, LSB first) ----------------------------------------------------
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    float a, b;
    glibcSincosf(__fmul_rn(angle, factorPI), &b, &a);
#pragma unroll
    for (int it = 0; it < 5; ++it) {
        const int item = it * 32 + lane;
        if (item < kPatchRows * 4) reinterpret_cast<uint4*>(patch)[(item >> 2) * (kPatchPitch / 16) + (item & 3)] = pv[it];
    }
    __syncwarp();
    const uint8_t* center = patch + kPatchR * kPatchPitch + (kp.x - bx0);
    constexpr int step = kPatchPitch;
    const bool strict = P.fpMode == 0;
    uint32_t t0[8], t1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float4 pt = sPat[k * 32 + lane];
        int r0, c0, r1, c1;
        if (strict) {
            r0 = __float2int_rn(__fadd_rn(__fmul_rn(pt.x, b), __fmul_rn(pt.y, a)));
            c0 = __float2int_rn(__fsub_rn(__fmul_rn(pt.x, a), __fmul_rn(pt.y, b)));
            r1 = __float2int_rn(__fadd_rn(__fmul_rn(pt.z, b), __fmul_rn(pt.w, a)));
            c1 = __float2int_rn(__fsub_rn(__fmul_rn(pt.z, a), __fmul_rn(pt.w, b)));
        } else {
            r0 = __float2int_rn(__fmaf_rn(pt.x, b, __fmul_rn(pt.y, a)));
            c0 = __float2int_rn(__fmaf_rn(pt.x, a, -__fmul_rn(pt.y, b)));
            r1 = __float2int_rn(__fmaf_rn(pt.z, b, __fmul_rn(pt.w, a)));
            c1 = __float2int_rn(__fmaf_rn(pt.z, a, -__fmul_rn(pt.w, b)));
        }
        t0[k] = center[r0 * step + c0];
        t1[k] = center[r1 * step + c1];
    }
    uint32_t byte = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) byte |= (uint32_t)(t0[k] < t1[k]) << k;
    uint32_t word = byte << (8 * (lane & 3));
    word |= __shfl_xor_sync(0xffffffffu, word, 1);
    word |= __shfl_xor_sync(0xffffffffu, word, 2);
    const int half = lane & 16;
    uint4 q;
    q.x = __shfl_sync(0xffffffffu, word, half + 0);
    q.y = __shfl_sync(0xffffffffu, word, half + 4);
    q.z = __shfl_sync(0xffffffffu, word, half + 8);
    q.w = __shfl_sync(0xffffffffu, word, half + 12);
    uint8_t* d = outDesc + ((size_t)f * P.outCap + s) * 32;
    if ((lane & 15) == 0) *reinterpret_cast<uint4*>(d + (lane >> 4) * 16) = q;

    // cv::KeyPoint record (:880-890, :1149-1151)
    if (lane < 7) {
        float x = (float)kp.x, y = (float)kp.y;
        if (l != 0) { x = __fmul_rn(x, L.scale); y = __fmul_rn(y, L.scale); }
        float v;
        switch (lane) {
            case 0: v = x; break;
            case 1: v = y; break;
            case 2: v = L.patchSize; break;
            case 3: v = angle; break;
            case 4: v = (float)kp.resp; break;
            case 5: v = __int_as_float(l); break;
            default: v = __int_as_float(-1); break;
        }
        outKps[((size_t)f * P.outCap + s) * 7 + lane] = v;
    }
    }  // persistent loop
}


// ---- describeAsyncKernel (round 2, default): the same arithmetic, software-pipelined with cp.async ---------------------------
// describeKernel's first stall reason is long_scoreboard: every warp starts a keypoint with 21 global loads and waits for them.
// Here both patches of the NEXT keypoint are copied global -> shared memory by cp.async while the current one is computed: the
// orientation patch (31 rows x 3 sixteen-byte chunks of the padded plane) is requested as soon as IC_Angle has read the current
// one, the blurred 37 x 37 patch as soon as the 512 samples are taken; one buffer of each per warp, two commit groups in flight.
constexpr int kIcPitch = 48;  // 3 chunks: (px & 15 rounded down to a word) + 36 bytes <= 48
__device__ __forceinline__ void cpAsync16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cpAsyncCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cpAsyncWait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// BULK: one cp.async.bulk per patch ROW (lane r copies row r; completion on a per-warp mbarrier) instead of one cp.async per
// 16-byte chunk: 3 copy instructions per lane and keypoint instead of 8, and one address multiply each.
__device__ __forceinline__ void bulkRow(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smemAddr(dst)), "l"(src),
                 "r"(bytes), "r"(smemAddr(bar))
                 : "memory");
}

template <bool BULK>
__global__ void __launch_bounds__(256, ORBX_DESC_ASYNC_MINB) describeAsyncKernel(const __grid_constant__ FramePlan P, const uint8_t* __restrict__ pyr,
                                                           const uint8_t* __restrict__ blur, const LevelKp* __restrict__ sorted,
                                                           const int* __restrict__ outN, float* __restrict__ outKps,
                                                           uint8_t* __restrict__ outDesc, int F) {
    __shared__ float4 sPat[256];
    __shared__ uint2 sCoef[248];
    __shared__ __align__(16) uint8_t sPatch[8][kPatchRows * kPatchPitch];
    __shared__ __align__(16) uint8_t sIc[8][31 * kIcPitch];
    __shared__ __align__(8) uint64_t sBar[8][2];
    sPat[(threadIdx.x & 7) * 32 + (threadIdx.x >> 3)] = gPatternF[threadIdx.x];
    if (threadIdx.x < 248) sCoef[threadIdx.x] = gIcCoef[threadIdx.x];
    if (BULK && threadIdx.x < 16) mbarInit(&sBar[threadIdx.x >> 1][threadIdx.x & 1], 1);
    if (BULK) mbarInitFence();
    __syncthreads();
    const int lane = threadIdx.x & 31;
    uint8_t* patch = sPatch[threadIdx.x >> 5];
    uint8_t* ic = sIc[threadIdx.x >> 5];
    uint64_t* barIc = &sBar[threadIdx.x >> 5][0];
    uint64_t* barPatch = &sBar[threadIdx.x >> 5][1];
    uint32_t phIc = 0, phPatch = 0;
    const int total = F * P.outCap, stride = gridDim.x * 8;
    const int qs = stride / P.outCap, rs = stride - qs * P.outCap;

    // lane-invariant parts of the copies' addresses (shared-memory window addresses; the global side is one multiply per patch)
    const uint32_t patchDst = smemAddr(patch) + (lane >> 2) * kPatchPitch + (lane & 3) * 16;  // row (lane >> 2) + 8 it, chunk lane & 3
    const uint32_t* icRd = reinterpret_cast<const uint32_t*>(ic) + (lane >> 3) * (kIcPitch / 4) + (lane & 7);
    const uint2* coefRd = sCoef + lane;
    const int rowW = (lane >> 3) - kHalfPatch;
    auto cp16 = [](uint32_t dst, const uint8_t* src) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
    };
    auto issueIc = [&](int f, const LevelKp& kp) {
        const LevelPlan& L = P.lv[kp.level];
        const int px = kp.x + kEdge - kHalfPatch;
        const uint8_t* p0 = pyr + (size_t)f * P.frameBytes + L.planeOff + (size_t)(kp.y + kEdge - kHalfPatch) * L.pitch + (px & ~15);
        if (BULK) {
            if (lane == 0) mbarExpectTx(barIc, 31 * kIcPitch);
            if (lane < 31) bulkRow(ic + lane * kIcPitch, p0 + (size_t)(lane * L.pitch), kIcPitch, barIc);
            return;
        }
        // 93 chunks: row = item / 3, chunk = item % 3.  Three lanes share a row, so a request touches 11 lines; one row per lane
        // (trivial addresses, 31 lines per request) was measured slower: 0.269 against 0.259 ms.
#pragma unroll
        for (int it = 0; it < 3; ++it) {
            const int item = it * 32 + lane;
            const int r = (item * 43) >> 7, c = item - 3 * r;
            if (item < 93) cp16(smemAddr(ic) + r * kIcPitch + c * 16, p0 + (unsigned)(r * L.pitch + c * 16));
        }
    };
    auto issuePatch = [&](int f, const LevelKp& kp) {
        const LevelPlan& L = P.lv[kp.level];
        const int bx0 = (kp.x - kPatchR) & ~15;
        const uint8_t* src = blur + (size_t)f * P.blurBytes + L.blurOff + (size_t)(kp.y - kPatchR) * L.bpitch + bx0;
        if (BULK) {
            if (lane == 0) mbarExpectTx(barPatch, kPatchRows * 64);
            bulkRow(patch + lane * kPatchPitch, src + (size_t)(lane * L.bpitch), 64, barPatch);
            if (lane < kPatchRows - 32) bulkRow(patch + (32 + lane) * kPatchPitch, src + (size_t)((32 + lane) * L.bpitch), 64, barPatch);
            return;
        }
        // rows (lane >> 2) + 8 it, chunk lane & 3: 4 full sweeps + rows 32..36
        const uint8_t* q = src + (unsigned)((lane >> 2) * L.bpitch + (lane & 3) * 16);
        const unsigned step = 8u * (unsigned)L.bpitch;
#pragma unroll
        for (int it = 0; it < 4; ++it, q += step) cp16(patchDst + it * 8 * kPatchPitch, q);
        if (lane < 4 * (kPatchRows - 32)) cp16(patchDst + 32 * kPatchPitch, q);
    };

    int w = blockIdx.x * 8 + (threadIdx.x >> 5);
    int f = w / P.outCap, s = w - f * P.outCap;
    // slot A = w (its patches are in flight / in shared memory), slot B = w + stride (record loaded)
    LevelKp kpA = {0, 0, 0, 0}, kpB = {0, 0, 0, 0};
    int nA = 0, nB = 0;
    int fB = f + qs, sB = s + rs;
    if (sB >= P.outCap) { sB -= P.outCap; ++fB; }
    if (w < total) { nA = __ldg(outN + f); kpA = sorted[w]; }
    if (w + stride < total) { nB = __ldg(outN + fB); kpB = sorted[w + stride]; }
    if (w < total && s < nA) issueIc(f, kpA);
    if (!BULK) cpAsyncCommit();
    if (w < total && s < nA) issuePatch(f, kpA);
    if (!BULK) cpAsyncCommit();

    for (; w < total; w += stride) {
        const bool valid = s < nA, validB = w + stride < total && sB < nB;
        const LevelKp kp = kpA;
        // record of the slot after the next (fetching it at the END of the iteration instead, to shorten its live range, was
        // measured slower: 0.278 against 0.255 ms)
        int fC = fB + qs, sC = sB + rs;
        if (sC >= P.outCap) { sC -= P.outCap; ++fC; }
        LevelKp kpC = {0, 0, 0, 0};
        int nC = 0;
        if (w + 2 * stride < total) { nC = __ldg(outN + fC); kpC = sorted[w + 2 * stride]; }

        const int l = kp.level;
        const LevelPlan& L = P.lv[l];
        float angle = 0.f;
        if (valid) {
            // this slot's orientation patch has landed (its blurred patch may still be in flight)
            if (BULK) { mbarWait(barIc, phIc); phIc ^= 1; }
            else { cpAsyncWait1(); __syncwarp(); }
            // ---- IC_Angle (ORBextractor.cc:76-103), see describeKernel.  item = 32 it + lane: row 4 it + (lane >> 3), word lane & 7,
            // so every address is a lane-invariant base plus an immediate; m01 = sum (row - 15) S = rowW sum S + 4 sum it S ----
            const int px = kp.x + kEdge - kHalfPatch;
            const int sh = (px & 3) * 8;
            int m10 = 0, m01 = 0;
            const uint32_t* q = icRd + ((px & 15) >> 2);
            int sA = 0, sB4 = 0;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const uint32_t lo = q[it * kIcPitch], hi = q[it * kIcPitch + 1];  // 4 rows = 4 * kIcPitch / 4 words
                const uint2 cf = coefRd[it * 32];
                uint32_t wv = __funnelshift_r(lo, hi, sh);
                if (it == 7 && lane >= 24) wv = 0u;  // items 248..255 do not exist (what was read is other shared memory)
                m10 = dp4a_us(wv, cf.x, m10);
                const int S = (int)__dp4a(wv, cf.y, 0u);
                sA += S;
                sB4 += it * S;
            }
            m01 = rowW * sA + 4 * sB4;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                m10 += __shfl_xor_sync(0xffffffffu, m10, o);
                m01 += __shfl_xor_sync(0xffffffffu, m01, o);
            }
            angle = fastAtan2Deg((float)m01, (float)m10);
        }
        __syncwarp();  // every lane is done with `ic`
        if (validB) issueIc(fB, kpB);
        if (!BULK) cpAsyncCommit();

        if (valid) {
            // ---- rBRIEF: lane i computes byte i (pairs 8i .. 8i+7, LSB first) ----
            const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
            float a, b;
            glibcSincosf(__fmul_rn(angle, factorPI), &b, &a);
            // this slot's blurred patch has landed (the next slot's orientation patch may be in flight)
            if (BULK) { mbarWait(barPatch, phPatch); phPatch ^= 1; }
            else { cpAsyncWait1(); __syncwarp(); }
            const int bx0 = (kp.x - kPatchR) & ~15;
            const uint8_t* center = patch + kPatchR * kPatchPitch + (kp.x - bx0);
            constexpr int step = kPatchPitch;
            const bool strict = P.fpMode == 0;
            uint32_t byte = 0;
#pragma unroll
            // the sign of t0 - t1 = (t0 < t1) enters at bit 0: pair 7 first, so pair k ends at bit k.  Each pair is consumed at once: keeping
            // the 16 samples in registers until the end made ptxas spill (64 registers), 0.294 ms
            for (int k = 7; k >= 0; --k) {
                const float4 pt = sPat[k * 32 + lane];
                int r0, c0, r1, c1;
                if (strict) {
                    r0 = __float2int_rn(__fadd_rn(__fmul_rn(pt.x, b), __fmul_rn(pt.y, a)));
                    c0 = __float2int_rn(__fsub_rn(__fmul_rn(pt.x, a), __fmul_rn(pt.y, b)));
                    r1 = __float2int_rn(__fadd_rn(__fmul_rn(pt.z, b), __fmul_rn(pt.w, a)));
                    c1 = __float2int_rn(__fsub_rn(__fmul_rn(pt.z, a), __fmul_rn(pt.w, b)));
                } else {
                    r0 = __float2int_rn(__fmaf_rn(pt.x, b, __fmul_rn(pt.y, a)));
                    c0 = __float2int_rn(__fmaf_rn(pt.x, a, -__fmul_rn(pt.y, b)));
                    r1 = __float2int_rn(__fmaf_rn(pt.z, b, __fmul_rn(pt.w, a)));
                    c1 = __float2int_rn(__fmaf_rn(pt.z, a, -__fmul_rn(pt.w, b)));
                }
                const uint32_t t0 = center[r0 * step + c0], t1 = center[r1 * step + c1];
                byte = __funnelshift_l(t0 - t1, byte, 1);
            }
            uint32_t word = byte << (8 * (lane & 3));
            word |= __shfl_xor_sync(0xffffffffu, word, 1);
            word |= __shfl_xor_sync(0xffffffffu, word, 2);
            const int half = lane & 16;
            uint4 q;
            q.x = __shfl_sync(0xffffffffu, word, half + 0);
            q.y = __shfl_sync(0xffffffffu, word, half + 4);
            q.z = __shfl_sync(0xffffffffu, word, half + 8);
            q.w = __shfl_sync(0xffffffffu, word, half + 12);
            uint8_t* d = outDesc + ((size_t)f * P.outCap + s) * 32;
            if ((lane & 15) == 0) *reinterpret_cast<uint4*>(d + (lane >> 4) * 16) = q;
        }
        __syncwarp();  // every lane is done with `patch`
        if (validB) issuePatch(fB, kpB);
        if (!BULK) cpAsyncCommit();

        // cv::KeyPoint record (:880-890, :1149-1151)
        if (valid && lane < 7) {
            float x = (float)kp.x, y = (float)kp.y;
            if (l != 0) { x = __fmul_rn(x, L.scale); y = __fmul_rn(y, L.scale); }
            float v;
            switch (lane) {
                case 0: v = x; break;
                case 1: v = y; break;
                case 2: v = L.patchSize; break;
                case 3: v = angle; break;
                case 4: v = (float)kp.resp; break;
                case 5: v = __int_as_float(l); break;
                default: v = __int_as_float(-1); break;
            }
            outKps[((size_t)f * P.outCap + s) * 7 + lane] = v;
        }
        kpA = kpB; nA = nB; f = fB; s = sB;
        kpB = kpC; nB = nC; fB = fC; sB = sC;
    }
    if (!BULK) asm volatile("cp.async.wait_all;" ::: "memory");
}

}  // namespace

// per device, once per extractor (orbx_create)
cudaError_t uploadDescribeTables() {
    static const int umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};  // ORBextractor.cc:453-468
    float4 pat[256];
    for (int i = 0; i < 256; ++i)
        pat[i] = make_float4((float)kOrbPattern[i][0], (float)kOrbPattern[i][1], (float)kOrbPattern[i][2], (float)kOrbPattern[i][3]);
    cudaError_t e = cudaMemcpyToSymbol(gPatternF, pat, sizeof(pat));
    if (e != cudaSuccess) return e;
    uint2 coef[248];
    for (int r = 0; r < 31; ++r)
        for (int k = 0; k < 8; ++k) {
            uint32_t cu = 0, cm = 0;
            const int v = r - 15;
            for (int b = 0; b < 4; ++b) {
                const int u = 4 * k - 15 + b;
                const int au = u < 0 ? -u : u;
                if (au <= umax[v < 0 ? -v : v]) {  // u = 16 (last byte of word 7) is never inside
                    cu |= (uint32_t)(uint8_t)(int8_t)u << (8 * b);
                    cm |= 1u << (8 * b);
                }
            }
            coef[r * 8 + k] = make_uint2(cu, cm);
        }
    return cudaMemcpyToSymbol(gIcCoef, coef, sizeof(coef));
}

void launchDescribe(const FramePlan& P, const Buffers& B, int F, int lap0, int lap1, cudaStream_t s) {
    slotKernel<<<F, 256, 0, s>>>(P, B.lkp, B.lkpCnt, B.sorted, B.outN, B.outMono, lap0, lap1, B.errFlag);
    countLaunch();
    int dev = 0, sms = 148, perSm = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // ORBX_DESC_MODE=0: the round-1 kernel (plain loads at the start of every keypoint); default: the cp.async pipeline
    // ORBX_DESC_MODE=0: the round-1 kernel (plain loads at the start of every keypoint, 0.280 ms per 256 frames); default (1):
    // cp.async per 16-byte chunk (0.276 ms; ncu: long_scoreboard 3.2 -> 0.4 stall cycles per issue, issue slots 67.6 -> 74.7 %, but 8 %
    // more instructions for the copies' addresses); 2: one bulk copy per patch row - measured 0.476 ms (68 TMA operations of 48-64
    // bytes per keypoint: the copy engine's per-operation cost dominates), kept for A/B runs only
    static const int mode = [] { const char* e = std::getenv("ORBX_DESC_MODE"); return e ? std::atoi(e) : 1; }();
    if (mode == 2) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, describeAsyncKernel<true>, 256, 0);
    else if (mode == 1) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, describeAsyncKernel<false>, 256, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, describeKernel, 256, 0);
    const long long want = ((long long)F * P.outCap + 7) / 8;
    // ORBX_DESC_PERSM (measurements only): cap on the persistent CTAs per SM
    static const int capPerSm = [] { const char* e = std::getenv("ORBX_DESC_PERSM"); return e ? std::atoi(e) : 0; }();
    if (capPerSm > 0) perSm = std::min(perSm, capPerSm);
    const int grid = (int)std::min<long long>(want, (long long)sms * std::max(perSm, 1));
    if (mode == 2) describeAsyncKernel<true><<<grid, 256, 0, s>>>(P, B.pyr, B.blur, B.sorted, B.outN, reinterpret_cast<float*>(B.outKps), B.outDesc, F);
    else if (mode == 1) describeAsyncKernel<false><<<grid, 256, 0, s>>>(P, B.pyr, B.blur, B.sorted, B.outN, reinterpret_cast<float*>(B.outKps), B.outDesc, F);
    else describeKernel<<<grid, 256, 0, s>>>(P, B.pyr, B.blur, B.sorted, B.outN, reinterpret_cast<float*>(B.outKps), B.outDesc, F);
    countLaunch();
}

}  // namespace orbx

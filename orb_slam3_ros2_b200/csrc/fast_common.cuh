// fast_common.cuh — the exact FAST-9/16 arc strength shared by the strip kernel (k_fast_strip.cu) and the per-cell kernel
// (k_fast.cu).  cv::FAST's score: arc(p) = max over the sixteen 9-arcs of max(min(I - ring), min(ring - I));
// p is a corner at t iff arc > t and the response is arc - 1.
#pragma once
#include <cstdint>

#include <cuda_runtime.h>

namespace orbx {

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

}  // namespace orbx

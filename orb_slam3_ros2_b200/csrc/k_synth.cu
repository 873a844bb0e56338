// k_synth.cu — deterministic synthetic frames generated on the device.
//
// Not part of the reference: EuRoC / KITTI / RealSense streams are unavailable offline, so BASELINE.json's configs
// are measured on procedurally textured frames.  Integer-only, bit-identical to orb_slam3_ros2_b200/synth.py
// (tests compare the two); used so that the 10k-frame 1080p configuration never stages 20 GB from the host.
#include "orbx_internal.cuh"

namespace orbx {
namespace {

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ unsigned long long hash3(unsigned long long seed, long long a, long long b, long long c) {
    return mix64(seed ^ ((unsigned long long)a * 0x9E3779B97F4A7C15ull) ^ ((unsigned long long)b * 0xC2B2AE3D27D4EB4Full) ^
                 ((unsigned long long)c * 0x165667B19E3779F9ull));
}

__device__ int canvas(long long X, long long Y, unsigned long long seed) {
    const int cells[5] = {4, 8, 16, 32, 64}, amps[5] = {10, 25, 40, 50, 60};
    long long val = 128;
#pragma unroll
    for (int o = 0; o < 5; ++o) {
        const long long c = cells[o];
        const long long i = X / c, fx = X % c, j = Y / c, fy = Y % c;
        const long long g00 = (long long)(hash3(seed, o, i, j) & 2047) - 1024;
        const long long g10 = (long long)(hash3(seed, o, i + 1, j) & 2047) - 1024;
        const long long g01 = (long long)(hash3(seed, o, i, j + 1) & 2047) - 1024;
        const long long g11 = (long long)(hash3(seed, o, i + 1, j + 1) & 2047) - 1024;
        const long long num = g00 * (c - fx) * (c - fy) + g10 * fx * (c - fy) + g01 * (c - fx) * fy + g11 * fx * fy;
        val += (amps[o] * num) / (1024 * c * c);  // C division truncates toward zero, like synth._tdiv
    }
    const long long ti = X / 64, tj = Y / 64;
    for (int dti = -1; dti <= 0; ++dti)
        for (int dtj = -1; dtj <= 0; ++dtj)
            for (int r = 0; r < 3; ++r) {
                const unsigned long long hv = hash3(seed, 100 + r, ti + dti, tj + dtj);
                const long long ox = hv & 63, oy = (hv >> 6) & 63;
                const long long w = 4 + (long long)((hv >> 12) % 36), h = 4 + (long long)((hv >> 20) % 36);
                const long long delta = (long long)((hv >> 28) % 160) - 80;
                const bool filled = ((hv >> 40) % 10) < 7;
                const long long x0 = (ti + dti) * 64 + ox, y0 = (tj + dtj) * 64 + oy;
                const bool inside = X >= x0 && X < x0 + w && Y >= y0 && Y < y0 + h;
                const bool edge = X == x0 || X == x0 + w - 1 || Y == y0 || Y == y0 + h - 1;
                if (inside && (filled || edge)) val += delta;
            }
    return (int)(val < 0 ? 0 : (val > 255 ? 255 : val));
}

__global__ void synthKernel(uint8_t* __restrict__ out, int k0, int w, int h, unsigned long long seed, int right) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    const int k = k0 + blockIdx.z;
    const long long X = 1024 + 3ll * k + x + (right ? 16 : 0), Y = 1024 + (long long)k + y;
    out[((size_t)blockIdx.z * h + y) * w + x] = (uint8_t)canvas(X, Y, seed);
}

}  // namespace

void launchSynth(uint8_t* dOut, int k0, int F, int w, int h, uint64_t seed, int right, cudaStream_t s) {
    synthKernel<<<dim3((w + 127) / 128, h, F), 128, 0, s>>>(dOut, k0, w, h, seed, right);
    countLaunch();
}

}  // namespace orbx

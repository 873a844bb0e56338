// tma_probe.cu — stand-alone check of the FAST tile staging's TMA recipe (3-D u8 tensor map, 48 x R x 1 box, per-warp
// mbarrier), in three ways of handing the tensor map to the kernel.  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/tma_probe tools/tma_probe.cu && /tmp/tma_probe
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct Maps {
    CUtensorMap m[16];
};

__device__ __forceinline__ void loadTile(uint8_t* tile, uint64_t* bar, const CUtensorMap* map, int x, int y, int z, uint32_t bytes,
                                         uint32_t phase) {
    if ((threadIdx.x & 31) == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                smemAddr(tile)),
            "l"(map), "r"(x), "r"(y), "r"(z), "r"(smemAddr(bar))
            : "memory");
    }
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smemAddr(bar)), "r"(phase) : "memory");
}

// mode 3: plain 1-D bulk copy of 2048 bytes (no tensor map) — checks mbarrier + bulk copy on their own
__global__ void probeBulk(const uint8_t* src, uint8_t* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 4096);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(2048) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smemAddr(smem)),
                     "l"(src), "r"(2048), "r"(smemAddr(bar))
                     : "memory");
    }
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smemAddr(bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) out[i] = smem[i];
}

// mode 4: 2-D tensor map, one warp
__global__ void probe2d(const __grid_constant__ CUtensorMap one, int rows, int bw, int x, int y, uint8_t* out, int variant) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (variant == 2) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (variant == 1) {
        // CUTLASS style: whole warp 0 enters, elect.sync picks the issuing lane
        if (threadIdx.x < 32) {
            uint32_t pred = 0;
            asm volatile(
                "{\n.reg .pred P1;\n.reg .b32 r;\nelect.sync r|P1, 0xffffffff;\nselp.b32 %0, 1, 0, P1;\n}\n" : "=r"(pred));
            if (pred) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(rows * bw) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                                 smemAddr(smem)),
                             "l"(&one), "r"(x), "r"(y), "r"(smemAddr(bar))
                             : "memory");
            }
        }
    } else if (threadIdx.x == 0) {
        if (variant == 3) asm volatile("prefetch.tensormap [%0];" ::"l"(&one) : "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(rows * bw) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                         smemAddr(smem)),
                     "l"(&one), "r"(x), "r"(y), "r"(smemAddr(bar))
                     : "memory");
    }
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smemAddr(bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < rows * bw; i += blockDim.x) out[i] = smem[i];
}

template <int MODE>
__global__ void probe(const __grid_constant__ Maps M, const __grid_constant__ CUtensorMap one, const CUtensorMap* gmaps, int l, int rows,
                      int x, int y, int z, uint8_t* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int wip = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warpBytes = ((rows * 48) + 127) & ~127;
    uint8_t* tile = smem + wip * warpBytes;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 8 * warpBytes) + wip;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const CUtensorMap* map = MODE == 0 ? &M.m[l] : (MODE == 1 ? &one : gmaps + l);
    uint32_t phase = 0;
    for (int it = 0; it < 3; ++it) {
        __syncwarp();
        loadTile(tile, bar, map, x + it + wip, y + it, z, 48u * rows, phase);
        phase ^= 1;
        __syncwarp();
    }
    if (wip == 0)
        for (int i = lane; i < rows * 48; i += 32) out[i] = tile[i];
}

int main(int argc, char** argv) {
    const int only = argc > 1 ? atoi(argv[1]) : -1;
    const int pitch = 896, H = 518, F = 3, rows = 44;
    const size_t frameBytes = (size_t)pitch * H + 256;
    std::vector<uint8_t> host(frameBytes * F);
    for (size_t i = 0; i < host.size(); ++i) host[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *dout;
    cudaMalloc(&d, host.size());
    cudaMalloc(&dout, rows * 48);
    cudaMemcpy(d, host.data(), host.size(), cudaMemcpyHostToDevice);
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    EncodeFn encode = (EncodeFn)fn;
    Maps M;
    std::memset(&M, 0, sizeof(M));
    const int l = 2;
    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)H, (cuuint64_t)F};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frameBytes};
    const cuuint32_t box[3] = {48u, (cuuint32_t)rows, 1u};
    const cuuint32_t es[3] = {1, 1, 1};
    CUresult r = encode(&M.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    CUtensorMap* gm;
    cudaMalloc(&gm, sizeof(M));
    cudaMemcpy(gm, &M, sizeof(M), cudaMemcpyHostToDevice);
    const int x = 35 + 3 * 37, y = 35 + 2 * 38, z = 1;
    const int warpBytes = ((rows * 48) + 127) & ~127;
    const size_t smem = 8 * warpBytes + 64;
    if (only == 3) {
        uint8_t* o2;
        cudaMalloc(&o2, 2048);
        probeBulk<<<1, 128, 4096 + 64>>>(d + 1024, o2);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<uint8_t> got(2048);
        cudaMemcpy(got.data(), o2, 2048, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int i = 0; i < 2048; ++i) bad += got[i] != host[1024 + i];
        printf("mode 3 (1-D bulk): %s, mismatches %d\n", cudaGetErrorString(e), bad);
        return e != cudaSuccess;
    }
    if (only == 4) {
        const int bw = argc > 2 ? atoi(argv[2]) : 48;
        const int promo = argc > 3 ? atoi(argv[3]) : 2;
        CUtensorMap t2;
        const cuuint64_t d2[2] = {(cuuint64_t)pitch, (cuuint64_t)H};
        const cuuint64_t s2[1] = {(cuuint64_t)pitch};
        const cuuint32_t b2[2] = {(cuuint32_t)bw, (cuuint32_t)rows};
        const cuuint32_t e2[2] = {1, 1};
        CUresult r2 = encode(&t2, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, d2, s2, b2, e2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        const uint64_t* w = reinterpret_cast<const uint64_t*>(&t2);
        for (int i = 0; i < 16; ++i) printf("%016llx%c", (unsigned long long)w[i], i % 4 == 3 ? '\n' : ' ');
        printf("base %p\n", (void*)d);
        uint8_t* o2;
        cudaMalloc(&o2, 16384);
        const int variant = argc > 4 ? atoi(argv[4]) : 0;
        const int xx = argc > 5 ? atoi(argv[5]) : 146;
        probe2d<<<1, 128, 16384 + 64>>>(t2, rows, bw, xx, 111, o2, variant);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<uint8_t> got(16384);
        cudaMemcpy(got.data(), o2, 16384, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int rr = 0; rr < rows; ++rr)
            for (int c = 0; c < bw; ++c) bad += got[rr * bw + c] != host[(size_t)(111 + rr) * pitch + xx + c];
        printf("mode 4 (2-D, box %d x %d, promo %d): encode %d, %s, mismatches %d\n", bw, rows, promo, (int)r2, cudaGetErrorString(e), bad);
        return e != cudaSuccess;
    }
    for (int mode = 0; mode < 3; ++mode) {
        if (only >= 0 && mode != only) continue;
        cudaMemset(dout, 0, rows * 48);
        if (mode == 0) probe<0><<<2, 256, smem>>>(M, M.m[l], gm, l, rows, x, y, z, dout);
        if (mode == 1) probe<1><<<2, 256, smem>>>(M, M.m[l], gm, l, rows, x, y, z, dout);
        if (mode == 2) probe<2><<<2, 256, smem>>>(M, M.m[l], gm, l, rows, x, y, z, dout);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<uint8_t> got(rows * 48);
        cudaMemcpy(got.data(), dout, got.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int rr = 0; rr < rows; ++rr)
            for (int c = 0; c < 48; ++c)
                bad += got[rr * 48 + c] != host[z * frameBytes + (size_t)(y + 2 + rr) * pitch + x + 2 + c];
        printf("mode %d: %s, mismatches %d\n", mode, cudaGetErrorString(e), bad);
        if (e != cudaSuccess) return 1;
    }
    return 0;
}

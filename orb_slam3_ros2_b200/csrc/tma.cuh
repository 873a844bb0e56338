// tma.cuh — the TMA / mbarrier PTX used by the tile-staging kernels (k_fast.cu, k_blur.cu) and the host-side tensor-map
// encoder.  sm_100a only.  The tensor maps describe the handle's whole pyramid allocation as a 3-D u8 tensor
// (x = byte column of a padded plane, y = plane row, z = frame), one map per pyramid level and box shape.
// Measured rule (tools/tma_probe.cu): the box must START on a 16-byte boundary of global memory, i.e. x % 16 == 0 here.
#pragma once
#include <cstdint>

#include <cuda.h>
#include <cuda_runtime.h>

#include "orbx_internal.cuh"

namespace orbx {

__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
}
__device__ __forceinline__ void mbarInitFence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbarExpectTx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smemAddr(bar)), "r"(parity) : "memory");
}
// one box of a 3-D u8 tensor (x = byte column, y = row, z = frame) -> dense rows in shared memory
__device__ __forceinline__ void tmaLoadTile(void* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smemAddr(dst)),
        "l"(map), "r"(x), "r"(y), "r"(z), "r"(smemAddr(bar))
        : "memory");
}

struct LevelMaps {
    CUtensorMap m[kMaxLevels];
};

// Encodes the map of level l with box boxW x boxH x 1 (boxW a multiple of 16, <= 256; boxH <= 256).  0 on success.
int encodeLevelMap(CUtensorMap* out, const FramePlan& P, int l, const uint8_t* pyrBase, int capFrames, int boxW, int boxH);

}  // namespace orbx

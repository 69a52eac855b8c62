// TMA bulk-copy / mbarrier helpers (inline PTX, sm_90+; SASS: UBLKCP / SYNCS) and lattice coordinates.
#ifndef PSGPU_DEVICE_UTIL_CUH
#define PSGPU_DEVICE_UTIL_CUH
#include <cuda_runtime.h>
#include <stdint.h>
#include "kernels.h"

namespace psgpu {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on the mbarrier (cp.async.bulk = TMA engine)
__device__ __forceinline__ void tma_load_bulk(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra LAB_DONE;\n"
        "bra LAB_WAIT;\n"
        "LAB_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// Corner coordinate along one axis. I = MPU lattice index, i = corner index 0..7 inside the MPU.
// PSGPU_COORD_MPU_LOCAL reproduces the reference's float expressions (PS_Polygonizer.cpp:366 origin =
// lo + (float)I*side; :578-583 corner = origin + i*cellsize, each product and sum rounded separately);
// PSGPU_COORD_GLOBAL derives every corner from its global integer index so that MPUs, slabs and GPUs agree
// bit-for-bit on shared planes.
__device__ __forceinline__ float corner_coord(const GridParams& G, int axis, int I, int i) {
    if (G.coordMode == PSGPU_COORD_MPU_LOCAL) {
        const float origin = __fadd_rn(G.gridLo[axis], __fmul_rn((float)I, G.mpuSide));
        return __fadd_rn(origin, __fmul_rn((float)i, G.cellsize));
    }
    return __fmaf_rn((float)(I * PS_CELLS_PER_MPU + i), G.cellsize, G.gridLo[axis]);
}

__device__ __forceinline__ bool is_inside(const GridParams& G, float f) { return G.insideGE ? (f >= G.iso) : (f > G.iso); }

}  // namespace psgpu
#endif

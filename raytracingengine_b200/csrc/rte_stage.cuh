// rte_stage.cuh — TMA bulk copies + mbarrier helpers and the staging of a scene's filter arrays into dynamic shared memory,
// shared by the render kernels (rte_render.cu: exact FP64 path, rte_render32.cu: FP32 production path).
#pragma once
#include "rte_kernels.h"

namespace rte {

// ------------------------------------------------------------------------------ TMA / mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(phase)
            : "memory");
    }
}

constexpr uint32_t kTmaChunk = 32768;

__device__ __forceinline__ void tma_chunks(unsigned char* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    for (uint32_t off = 0; off < bytes; off += kTmaChunk)
        tma_bulk_g2s(dst + off, reinterpret_cast<const unsigned char*>(src) + off, min(kTmaChunk, bytes - off), bar);
}

// Stage the filter arrays into dynamic shared memory with TMA bulk copies.  Called by every thread of the CTA.
struct StagedScene {
    const ulonglong2 *views, *tviews;   // hot views (shared memory)
    const float *cold, *tcold;          // cold views: global memory (measured: staging them too does not pay, not even for tiny scenes)
    const float4 *sph, *tri;            // generic filter arrays: shared or global
};
__device__ __forceinline__ void stage_scene(const DevScene& sc, const SmemPlan& pl, unsigned char* dyn, uint64_t* bar, StagedScene& S) {
    unsigned char* p_tv = dyn + pl.view_bytes;
    unsigned char* p_sph = p_tv + pl.tview_bytes;
    unsigned char* p_tri = p_sph + pl.sph_bytes;
    S.views = reinterpret_cast<const ulonglong2*>(dyn);
    S.tviews = reinterpret_cast<const ulonglong2*>(p_tv);
    S.sph = pl.stage_sph ? reinterpret_cast<const float4*>(p_sph) : sc.sph32;
    S.tri = pl.stage_tri ? reinterpret_cast<const float4*>(p_tri) : sc.tri32;
    S.cold = sc.views_cold;
    S.tcold = sc.tri_views_cold;
    if (threadIdx.x == 0) mbar_init(bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t vb = pl.stream_views ? 0u : pl.view_bytes;  // streamed views are fetched by ViewStream
        mbar_expect_tx(bar, vb + pl.tview_bytes + pl.sph_bytes + pl.tri_bytes);
        tma_chunks(dyn, sc.views, vb, bar);
        tma_chunks(p_tv, sc.tri_views, pl.tview_bytes, bar);
        tma_chunks(p_sph, sc.sph32, pl.sph_bytes, bar);
        tma_chunks(p_tri, sc.tri32, pl.tri_bytes, bar);
    }
    mbar_wait(bar, 0);
}

}  // namespace rte

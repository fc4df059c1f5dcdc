// Copy-engine staging for the K1 reduce kernels: mbarrier + cp.async.bulk primitives and the host
// interface of the two ring kernels (reduce_scalar_tma.cu, reduce_wide_tma.cu).  A ring keeps
// nstage tiles (tens of KB per CTA) in flight without holding them in registers.
#pragma once
#include <cuda.h>

#include "txb_common.cuh"

namespace txb {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a byte-count mismatch must surface as a CUDA error, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin)
        if (spin > (1u << 24)) __trap();
}
// global -> shared bulk copy (16-byte aligned addresses, size a multiple of 16)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// 2-D tiled copy through a tensor map: box (rows x channels) at element coordinates (c0, r0);
// out-of-range elements arrive as zeros and still count towards the transaction bytes
__device__ __forceinline__ void tensor_g2s_2d(void* dst, const CUtensorMap* map, int c0, int r0, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(r0)
        : "memory");
}

// 8-byte asynchronous global -> shared copy (no register staging); completion by cp_async_wait_all
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

constexpr int WT_MAXSTAGE = 6;

// scalar-x ring kernel (reduce_scalar_tma.cu)
int64_t scalar_tma_blocks(int64_t nsamp, int64_t max_blocks, int order);
int launch_scalar_tma(int order, const double* x, const double* u, const double* w, int64_t nsamp,
                      const double* pivot, int64_t nb, double* ws0, double* ws1, cudaStream_t st);

// wide-x ring kernel (reduce_wide_tma.cu).  Geometry: nthreads = lanes_c * rows, a tile is
// rows * unr sample rows; `contig` tiles are whole rows (one bulk copy), otherwise channel strips
// of lanes_c * cpt channels through a tensor map.
struct WideRingPlan {
    bool ok, contig;
    int nthreads, cpt, lanes_c, unr, nstage;
    int64_t nctiles, nslabs;
    size_t smem;
};
WideRingPlan plan_wide_ring(int64_t nsamp, int64_t nchan, int64_t ldx, int64_t ldu, int order, bool has_x,
                            bool upc, bool aligned, int64_t max_slabs_x_chan);
int launch_wide_ring(const WideRingPlan& p, int order, const double* x, const double* u, const double* w,
                     int64_t nsamp, int64_t nchan, int64_t ldx, int64_t ldu, const double* pivot,
                     double* ws0, double* ws1, cudaStream_t st);

}  // namespace txb

// Shared helpers for the thermoextrap_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/txb.h"

#define TXB_MAXK 12  // highest supported u-moment order (x_is_u at order 11 -> 12)
#define TXB_NSM_FALLBACK 148

namespace txb {

void set_error(const char* fmt, ...);
int  sm_count();

// Measurement hook (txb_profile_begin / txb_profile_end in txb.h): while a tag is active,
// launches wrapped in a ProfScope of that tag are bracketed by CUDA events on their stream.
struct ProfScope {
    ProfScope(const char* tag, cudaStream_t st);
    ~ProfScope();
    cudaStream_t st_;
    int slot_;
};

#define TXB_CHECK_ARG(cond, ...)                  \
    do {                                          \
        if (!(cond)) {                            \
            txb::set_error(__VA_ARGS__);          \
            return TXB_ERR_BAD_ARG;               \
        }                                         \
    } while (0)

#define TXB_CUDA_LAUNCH_CHECK(what)                                              \
    do {                                                                         \
        cudaError_t e__ = cudaGetLastError();                                    \
        if (e__ != cudaSuccess) {                                                \
            txb::set_error("%s: CUDA error: %s", what, cudaGetErrorString(e__)); \
            return TXB_ERR_CUDA;                                                 \
        }                                                                        \
    } while (0)

// Dispatch a runtime order (1..TXB_MAXK) to a compile-time template argument.
#define TXB_ORDER_CASE(K_, ...) \
    case K_: {                  \
        constexpr int K = K_;   \
        __VA_ARGS__;            \
    } break;

#define TXB_DISPATCH_ORDER(order, ...)                                  \
    switch (order) {                                                    \
        TXB_ORDER_CASE(1, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(2, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(3, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(4, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(5, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(6, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(7, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(8, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(9, __VA_ARGS__)                                  \
        TXB_ORDER_CASE(10, __VA_ARGS__)                                 \
        TXB_ORDER_CASE(11, __VA_ARGS__)                                 \
        TXB_ORDER_CASE(12, __VA_ARGS__)                                 \
        default:                                                        \
            txb::set_error("order %d not in [1, %d]", (int)(order), TXB_MAXK); \
            return TXB_ERR_UNSUPPORTED;                                 \
    }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// Streaming (read-once) 128-bit global loads: bypass L1 allocation.
__device__ __forceinline__ double2 ldg_stream2(const double* p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];"
                 : "=d"(v.x), "=d"(v.y)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ double ldg_stream1(const double* p) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}

// L2 cache-policy helpers: keep gather tables resident, let streams pass through.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ double2 ldg_hint2(const void* p, uint64_t pol) {
    double2 v;
    asm volatile("ld.global.nc.L2::cache_hint.v2.f64 {%0,%1}, [%2], %3;"
                 : "=d"(v.x), "=d"(v.y)
                 : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ double2 ldg_stream_hint2(const void* p, uint64_t pol) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0,%1}, [%2], %3;"
                 : "=d"(v.x), "=d"(v.y)
                 : "l"(p), "l"(pol));
    return v;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace txb

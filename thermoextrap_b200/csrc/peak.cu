// Micro-benchmark: fp64 FMA peak of the device (roofline denominator for K2;
// MEASURED_PEAKS.json carries no fp64 figure).
#include "txb_common.cuh"

namespace txb {
__global__ void __launch_bounds__(256) k_fp64_peak(int iters, double seed, double* out) {
    double a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed + i * 1e-3 + threadIdx.x * 1e-6;
    const double m = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fma(a[i], m, c);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == 123.456) out[0] = s;  // never true; keeps the chain alive
}
}  // namespace txb

// Enqueues the benchmark kernel; *flops receives the fp64 flop count of the launch.
extern "C" int txb_fp64_peak_launch(int iters, int blocks_per_sm, double* scratch, double* flops,
                                    void* stream) {
    TXB_CHECK_ARG(iters > 0 && blocks_per_sm > 0 && scratch && flops, "txb_fp64_peak_launch: bad args");
    const int nb = txb::sm_count() * blocks_per_sm;
    txb::k_fp64_peak<<<nb, 256, 0, (cudaStream_t)stream>>>(iters, 1.0, scratch);
    TXB_CUDA_LAUNCH_CHECK("k_fp64_peak");
    *flops = 2.0 * 16.0 * (double)iters * 256.0 * (double)nb;
    return TXB_OK;
}

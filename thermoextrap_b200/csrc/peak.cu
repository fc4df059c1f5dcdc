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
// Same pipe, register-tiled contraction pattern: acc[i][c] += f[i] * p[c] with a TR x TC tile of
// independent accumulators and three distinct register operands per DFMA (what k_resample and
// k_bucket_consume issue), instead of one register operand plus two constants.
template <int TR, int TC>
__global__ void __launch_bounds__(256) k_fp64_tile_peak(int iters, double seed, double* out) {
    double acc[TR][TC], f[TR], p[TC];
#pragma unroll
    for (int i = 0; i < TR; ++i) f[i] = seed + i * 1e-3 + threadIdx.x * 1e-6;
#pragma unroll
    for (int c = 0; c < TC; ++c) p[c] = 1e-9 * (c + 1) + threadIdx.x * 1e-12;
#pragma unroll
    for (int i = 0; i < TR; ++i)
#pragma unroll
        for (int c = 0; c < TC; ++c) acc[i][c] = i + c;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < TR; ++i)
#pragma unroll
            for (int c = 0; c < TC; ++c) acc[i][c] = fma(f[i], p[c], acc[i][c]);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < TR; ++i)
#pragma unroll
        for (int c = 0; c < TC; ++c) s += acc[i][c];
    if (s == 123.456) out[0] = s;
}
// fp64 tensor-core forms of the same contraction (mma.sync DMMA), NACC independent accumulator fragments
__global__ void __launch_bounds__(256) k_dmma884_peak(int iters, double seed, double* out) {
    constexpr int NACC = 8;
    double c0[NACC], c1[NACC];
    const double a = seed + threadIdx.x * 1e-6, b = 1e-9 * (threadIdx.x + 1);
#pragma unroll
    for (int i = 0; i < NACC; ++i) c0[i] = i, c1[i] = -i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c0[i] + c1[i];
    if (s == 123.456) out[0] = s;
}
__global__ void __launch_bounds__(256) k_dmma16816_peak(int iters, double seed, double* out) {
    constexpr int NACC = 4;
    double c[NACC][4], a[8], b[4];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = seed + threadIdx.x * 1e-6 + k * 1e-3;
#pragma unroll
    for (int k = 0; k < 4; ++k) b[k] = 1e-9 * (threadIdx.x + 1 + k);
#pragma unroll
    for (int i = 0; i < NACC; ++i)
#pragma unroll
        for (int k = 0; k < 4; ++k) c[i][k] = i + k;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile(
                "mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, "
                "{%12,%13,%14,%15}, {%0,%1,%2,%3};"
                : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                  "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (s == 123.456) out[0] = s;
}
// int8 tensor-core form reachable with mma.sync (IMMA m16n8k32, u8 x s8 -> s32): the rate an exact
// integer-sliced (Ozaki-style) version of the bootstrap contraction could issue at without tcgen05.
__global__ void __launch_bounds__(256) k_imma16832_peak(int iters, double seed, double* out) {
    constexpr int NACC = 8;
    int c[NACC][4];
    unsigned a[4], b[2];
#pragma unroll
    for (int k = 0; k < 4; ++k) a[k] = 0x01020304u * (threadIdx.x + k + 1) + (unsigned)seed;
    b[0] = 0x7f01fe02u + threadIdx.x, b[1] = 0x03fd04fcu ^ threadIdx.x;
#pragma unroll
    for (int i = 0; i < NACC; ++i)
#pragma unroll
        for (int k = 0; k < 4; ++k) c[i][k] = i + k;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile(
                "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    int s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (s == 123456789) out[0] = s;
}
// Shared-memory atomic throughput with returned values on spread addresses: the slot reservation of
// k_bucket_scatter (one atomicAdd on tail[bucket] per draw, ~1200-1500 buckets, result needed).
__global__ void __launch_bounds__(512) k_smem_atomic_peak(int iters, unsigned mask, unsigned* out) {
    extern __shared__ unsigned bins[];
    for (unsigned i = threadIdx.x; i <= mask; i += 512) bins[i] = 0u;
    __syncthreads();
    unsigned s = threadIdx.x * 2654435761u + blockIdx.x * 40503u, acc = 0u;
    for (int it = 0; it < iters; ++it) {
        unsigned slot[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s = s * 1664525u + 1013904223u;
            slot[k] = atomicAdd(bins + ((s >> 10) & mask), 1u);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) acc += slot[k];
    }
    if (acc == 0xdeadbeefu) out[0] = acc;
}
}  // namespace txb

// *ops = shared-memory atomics issued (lanes); nbins must be a power of two <= 8192
extern "C" int txb_smem_atomic_peak_launch(int iters, int blocks_per_sm, int nbins, double* scratch, double* ops,
                                           void* stream) {
    TXB_CHECK_ARG(iters > 0 && blocks_per_sm > 0 && scratch && ops && nbins >= 32 && nbins <= 8192 &&
                      (nbins & (nbins - 1)) == 0, "txb_smem_atomic_peak_launch: bad args");
    const int nb = txb::sm_count() * blocks_per_sm;
    txb::k_smem_atomic_peak<<<nb, 512, nbins * sizeof(unsigned), (cudaStream_t)stream>>>(
        iters, (unsigned)nbins - 1u, reinterpret_cast<unsigned*>(scratch));
    TXB_CUDA_LAUNCH_CHECK("k_smem_atomic_peak");
    *ops = 8.0 * (double)iters * 512.0 * (double)nb;
    return TXB_OK;
}

// tile = 0: 4 x 14 accumulators (k_resample at order 6, 8 warps per SM at blocks_per_sm = 1);
// tile = 1: 2 x 14; tile = 2 / 3: the tensor-core forms (DMMA m8n8k4 / m16n8k16).  *flops as below.
extern "C" int txb_fp64_tile_peak_launch(int iters, int blocks_per_sm, int tile, double* scratch, double* flops,
                                         void* stream) {
    TXB_CHECK_ARG(iters > 0 && blocks_per_sm > 0 && scratch && flops, "txb_fp64_tile_peak_launch: bad args");
    const int nb = txb::sm_count() * blocks_per_sm;
    double per_thread = 0.0;  // flops per thread and iteration
    if (tile == 0) {
        txb::k_fp64_tile_peak<4, 14><<<nb, 256, 0, (cudaStream_t)stream>>>(iters, 1.0, scratch);
        per_thread = 2.0 * 56.0;
    } else if (tile == 1) {
        txb::k_fp64_tile_peak<2, 14><<<nb, 256, 0, (cudaStream_t)stream>>>(iters, 1.0, scratch);
        per_thread = 2.0 * 28.0;
    } else if (tile == 2) {  // DMMA m8n8k4: 512 flops per warp instruction, 8 per iteration
        txb::k_dmma884_peak<<<nb, 256, 0, (cudaStream_t)stream>>>(iters, 1.0, scratch);
        per_thread = 8.0 * 512.0 / 32.0;
    } else if (tile == 3) {  // DMMA m16n8k16: 4096 flops per warp instruction, 4 per iteration
        txb::k_dmma16816_peak<<<nb, 256, 0, (cudaStream_t)stream>>>(iters, 1.0, scratch);
        per_thread = 4.0 * 4096.0 / 32.0;
    } else {                 // IMMA m16n8k32 (u8 x s8): 8192 integer ops per warp instruction, 8 per iteration
        txb::k_imma16832_peak<<<nb, 256, 0, (cudaStream_t)stream>>>(iters, 1.0, scratch);
        per_thread = 8.0 * 8192.0 / 32.0;
    }
    TXB_CUDA_LAUNCH_CHECK("k_fp64_tile_peak");
    *flops = per_thread * (double)iters * 256.0 * (double)nb;
    return TXB_OK;
}

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

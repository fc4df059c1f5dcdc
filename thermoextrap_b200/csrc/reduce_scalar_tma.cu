// Scalar-x (or u-only) shifted-power sums, samples staged through a shared-memory ring filled by
// cp.async.bulk.  Arithmetic and partial layout are those of k_reduce_scalar (reduce.cu); the ring
// keeps ~100 KB per CTA in flight so the pass runs at the copy bandwidth of HBM instead of at what
// register-staged loads can keep outstanding.
#include <stdlib.h>

#include "reduce_tma.cuh"

namespace txb {

constexpr int ST_TILE = 2048;  // samples per tile and array: 16 KB, 4 double2 per thread

template <int K, bool HAS_X, bool HAS_W>
__global__ void __launch_bounds__(256, 2)
    k_reduce_scalar_tma(const double* __restrict__ x, const double* __restrict__ u,
                        const double* __restrict__ w, int64_t nsamp, const double* __restrict__ pivot,
                        int nstage, double* __restrict__ ws0, double* __restrict__ ws1) {
    extern __shared__ __align__(128) double ring[];
    __shared__ uint64_t full[WT_MAXSTAGE];
    constexpr int NARR = 1 + (HAS_X ? 1 : 0) + (HAS_W ? 1 : 0);
    constexpr uint32_t ARR_BYTES = ST_TILE * sizeof(double);
    const int tid = threadIdx.x;

    const double cx = HAS_X ? pivot[0] : 0.0;
    const double cu = pivot[HAS_X ? 1 : 0];
    double s0[K + 1], s1[K + 1];
#pragma unroll
    for (int j = 0; j <= K; ++j) s0[j] = 0.0, s1[j] = 0.0;
    long long cnt = 0;
    auto body = [&](double xv, double uv, double wv) {
        const double du = uv - cu;
        double p = HAS_W ? wv : 1.0;
        const double dx = HAS_X ? xv - cx : 0.0;
        if (HAS_W) s0[0] += p; else ++cnt;
        if (HAS_X) s1[0] = HAS_W ? fma(p, dx, s1[0]) : s1[0] + dx;
#pragma unroll
        for (int j = 1; j <= K; ++j) {
            p *= du;
            s0[j] += p;
            if (HAS_X) s1[j] = fma(p, dx, s1[j]);
        }
    };

    // full tiles, grid-strided: neighbouring CTAs stream neighbouring DRAM pages
    const int64_t ntile = nsamp / ST_TILE;
    const int64_t mine = (int64_t)blockIdx.x < ntile ? (ntile - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto issue = [&](int64_t i) {  // thread 0
        const int s = (int)(i % nstage);
        const int64_t off = ((int64_t)blockIdx.x + i * gridDim.x) * ST_TILE;
        double* dst = ring + (size_t)s * NARR * ST_TILE;
        mbar_arrive_expect_tx(&full[s], ARR_BYTES * NARR);
        bulk_g2s(dst, u + off, ARR_BYTES, &full[s]);
        if (HAS_X) bulk_g2s(dst + ST_TILE, x + off, ARR_BYTES, &full[s]);
        if (HAS_W) bulk_g2s(dst + (HAS_X ? 2 : 1) * ST_TILE, w + off, ARR_BYTES, &full[s]);
    };
    if (tid == 0) {
        for (int s = 0; s < nstage; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int64_t i = 0; i < mine && i < nstage; ++i) issue(i);
    }
    __syncthreads();
    for (int64_t i = 0; i < mine; ++i) {
        const int s = (int)(i % nstage);
        mbar_wait(&full[s], (uint32_t)((i / nstage) & 1));
        const double2* us = reinterpret_cast<const double2*>(ring + (size_t)s * NARR * ST_TILE);
        const double2* xs = us + ST_TILE / 2;
        const double2* wsm = us + (HAS_X ? 2 : 1) * (ST_TILE / 2);
#pragma unroll
        for (int q = 0; q < ST_TILE / 512; ++q) {
            const double2 uv = us[q * 256 + tid];
            double2 xv = make_double2(0, 0), wv = make_double2(1, 1);
            if (HAS_X) xv = xs[q * 256 + tid];
            if (HAS_W) wv = wsm[q * 256 + tid];
            body(xv.x, uv.x, wv.x);
            body(xv.y, uv.y, wv.y);
        }
        __syncthreads();  // every reader is done with slot s
        if (tid == 0 && i + nstage < mine) issue(i + nstage);
    }
    // ragged tail (< one tile): plain loads
    for (int64_t i = ntile * ST_TILE + (int64_t)blockIdx.x * 256 + tid; i < nsamp; i += (int64_t)gridDim.x * 256)
        body(HAS_X ? x[i] : 0.0, u[i], HAS_W ? w[i] : 1.0);
    if (!HAS_W) s0[0] = (double)cnt;

    __shared__ double sm[8][2 * (K + 1)];
    const int lane = tid & 31, wrp = tid >> 5;
#pragma unroll
    for (int j = 0; j <= K; ++j) {
        const double a = warp_sum(s0[j]);
        const double b = HAS_X ? warp_sum(s1[j]) : 0.0;
        if (lane == 0) sm[wrp][j] = a, sm[wrp][K + 1 + j] = b;
    }
    __syncthreads();
    if (tid < 2 * (K + 1)) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[k][tid];
        if (tid <= K)
            ws0[(int64_t)blockIdx.x * (K + 1) + tid] = t;
        else if (HAS_X)
            ws1[(int64_t)blockIdx.x * (K + 1) + tid - (K + 1)] = t;
    }
}

template <int K, bool HAS_X, bool HAS_W>
static int launch_k(const double* x, const double* u, const double* w, int64_t nsamp, const double* pivot,
                    int64_t nb, double* ws0, double* ws1, cudaStream_t st) {
    constexpr int NARR = 1 + (HAS_X ? 1 : 0) + (HAS_W ? 1 : 0);
    auto kern = k_reduce_scalar_tma<K, HAS_X, HAS_W>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10);
        TXB_CUDA_LAUNCH_CHECK("k_reduce_scalar_tma attribute");
        attr_set = true;
    }
    const size_t stage = (size_t)NARR * ST_TILE * sizeof(double);
    int ns = (int)((100u << 10) / stage);
    if (ns > WT_MAXSTAGE) ns = WT_MAXSTAGE;
    kern<<<(unsigned)nb, 256, stage * ns, st>>>(x, u, w, nsamp, pivot, ns, ws0, ws1);
    TXB_CUDA_LAUNCH_CHECK("k_reduce_scalar_tma");
    return TXB_OK;
}

// Blocks the ring kernel would use (0: not applicable -> the register-staged kernel runs).
// Measured (1e8 samples): the register-staged kernel already streams 6.3-7.1 TB/s up to order 6;
// the ring wins from order 8 (+2..11 %), where accumulators leave too few registers for loads.
// TXB_SCALAR_TMA=0/1 forces the choice (A/B measurements).
int64_t scalar_tma_blocks(int64_t nsamp, int64_t max_blocks, int order) {
    const char* e = getenv("TXB_SCALAR_TMA");
    if (e ? atoi(e) == 0 : order < 8) return 0;
    int64_t nb = (int64_t)sm_count() * 2;
    if (nb > max_blocks) nb = max_blocks;
    if (nsamp / ST_TILE < 4 * nb) return 0;  // short inputs: the pipeline never fills
    return nb;
}

int launch_scalar_tma(int order, const double* x, const double* u, const double* w, int64_t nsamp,
                      const double* pivot, int64_t nb, double* ws0, double* ws1, cudaStream_t st) {
    int rc = TXB_OK;
    TXB_DISPATCH_ORDER(order, {
        if (x) {
            if (w) rc = launch_k<K, true, true>(x, u, w, nsamp, pivot, nb, ws0, ws1, st);
            else rc = launch_k<K, true, false>(x, u, w, nsamp, pivot, nb, ws0, ws1, st);
        } else {
            if (w) rc = launch_k<K, false, true>(x, u, w, nsamp, pivot, nb, ws0, ws1, st);
            else rc = launch_k<K, false, false>(x, u, w, nsamp, pivot, nb, ws0, ws1, st);
        }
    });
    return rc;
}

}  // namespace txb

// K1 for PLANAR channels: x_0[nsamp] .. x_{C-1}[nsamp] given as SEPARATE arrays next to u[nsamp].
//
// The volume path (reference volume.py:92-134, 137-209) needs <x>, <W>, <x W> and <dxdq> from three
// per-sample streams (x, dxdq, W = beta * virial).  Reducing (x, W) and (dxdq, W) as two K1 passes
// reads W twice (32 B / sample); here every stream is read once (24 B / sample):
//   S0[j]    = sum_s du^j               j = 0..K     (du = u - cu)
//   S1[c][j] = sum_s (x_c - cx_c) du^j
// Tiles of every stream go through a shared-memory ring filled by cp.async.bulk (the copy-engine
// pipeline of reduce_scalar_tma.cu), so the pass runs at the copy bandwidth of HBM.
#include "reduce_tma.cuh"

namespace txb {

int launch_sum_parts(const double* ws, int64_t nparts, int64_t n, double* out, cudaStream_t st);

constexpr int PL_TILE = 2048;   // samples per tile and stream: 16 KB
constexpr int PL_MAXCH = 4;

struct PlanarArgs {
    const double* x[PL_MAXCH];
    const double* u;
    const double* pivot;   // [C] cx, then cu
    int64_t nsamp;
    int nstage;
    double* ws0;           // [nblocks][K+1]
    double* ws1;           // [nblocks][C][K+1]
};

template <int K, int C>
__global__ void __launch_bounds__(256, 2) k_reduce_planar(const PlanarArgs a) {
    extern __shared__ __align__(128) double ring[];
    __shared__ uint64_t full[WT_MAXSTAGE];
    constexpr int NARR = 1 + C;
    constexpr uint32_t ARR_BYTES = PL_TILE * sizeof(double);
    const int tid = threadIdx.x;
    const int nstage = a.nstage;

    double cx[C];
#pragma unroll
    for (int c = 0; c < C; ++c) cx[c] = a.pivot[c];
    const double cu = a.pivot[C];
    double s0[K + 1], s1[C][K + 1];
#pragma unroll
    for (int j = 0; j <= K; ++j) {
        s0[j] = 0.0;
#pragma unroll
        for (int c = 0; c < C; ++c) s1[c][j] = 0.0;
    }
    long long cnt = 0;
    auto body = [&](const double (&xv)[C], double uv) {
        const double du = uv - cu;
        double dx[C];
#pragma unroll
        for (int c = 0; c < C; ++c) {
            dx[c] = xv[c] - cx[c];
            s1[c][0] += dx[c];
        }
        ++cnt;
        double p = 1.0;
#pragma unroll
        for (int j = 1; j <= K; ++j) {
            p *= du;
            s0[j] += p;
#pragma unroll
            for (int c = 0; c < C; ++c) s1[c][j] = fma(p, dx[c], s1[c][j]);
        }
    };

    const int64_t ntile = a.nsamp / PL_TILE;
    const int64_t mine = (int64_t)blockIdx.x < ntile ? (ntile - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto issue = [&](int64_t i) {  // thread 0
        const int s = (int)(i % nstage);
        const int64_t off = ((int64_t)blockIdx.x + i * gridDim.x) * PL_TILE;
        double* dst = ring + (size_t)s * NARR * PL_TILE;
        mbar_arrive_expect_tx(&full[s], ARR_BYTES * NARR);
        bulk_g2s(dst, a.u + off, ARR_BYTES, &full[s]);
#pragma unroll
        for (int c = 0; c < C; ++c) bulk_g2s(dst + (size_t)(1 + c) * PL_TILE, a.x[c] + off, ARR_BYTES, &full[s]);
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
        const double2* us = reinterpret_cast<const double2*>(ring + (size_t)s * NARR * PL_TILE);
#pragma unroll
        for (int q = 0; q < PL_TILE / 512; ++q) {
            const double2 uv = us[q * 256 + tid];
            double xa[C], xb[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const double2 v = us[(size_t)(1 + c) * (PL_TILE / 2) + q * 256 + tid];
                xa[c] = v.x, xb[c] = v.y;
            }
            body(xa, uv.x);
            body(xb, uv.y);
        }
        __syncthreads();  // every reader is done with slot s
        if (tid == 0 && i + nstage < mine) issue(i + nstage);
    }
    // ragged tail (< one tile): plain loads
    for (int64_t i = ntile * PL_TILE + (int64_t)blockIdx.x * 256 + tid; i < a.nsamp; i += (int64_t)gridDim.x * 256) {
        double xv[C];
#pragma unroll
        for (int c = 0; c < C; ++c) xv[c] = a.x[c][i];
        body(xv, a.u[i]);
    }
    s0[0] = (double)cnt;

    constexpr int NV = (1 + C) * (K + 1);
    __shared__ double sm[8][NV];
    const int lane = tid & 31, wrp = tid >> 5;
#pragma unroll
    for (int j = 0; j <= K; ++j) {
        const double t = warp_sum(s0[j]);
        if (lane == 0) sm[wrp][j] = t;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const double v = warp_sum(s1[c][j]);
            if (lane == 0) sm[wrp][(1 + c) * (K + 1) + j] = v;
        }
    }
    __syncthreads();
    for (int v = tid; v < NV; v += 256) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[k][v];
        if (v <= K)
            a.ws0[(int64_t)blockIdx.x * (K + 1) + v] = t;
        else
            a.ws1[(int64_t)blockIdx.x * C * (K + 1) + v - (K + 1)] = t;
    }
}

template <int K, int C>
static int launch_planar(PlanarArgs a, int64_t nb, cudaStream_t st) {
    auto kern = k_reduce_planar<K, C>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10);
        TXB_CUDA_LAUNCH_CHECK("k_reduce_planar attribute");
        attr_set = true;
    }
    const size_t stage = (size_t)(1 + C) * PL_TILE * sizeof(double);
    int ns = (int)((100u << 10) / stage);
    if (ns > WT_MAXSTAGE) ns = WT_MAXSTAGE;
    if (ns < 2) ns = 2;
    a.nstage = ns;
    {
        ProfScope prof("reduce", st);
        ProfScope ring("reduce_planar", st);
        kern<<<(unsigned)nb, 256, stage * ns, st>>>(a);
    }
    TXB_CUDA_LAUNCH_CHECK("k_reduce_planar");
    return TXB_OK;
}

static int64_t planar_blocks() { return (int64_t)sm_count() * 2; }

}  // namespace txb

using namespace txb;

extern "C" size_t txb_reduce_planar_workspace_bytes(int nplanes, int order) {
    return align_up((size_t)planar_blocks() * (size_t)(1 + nplanes) * (order + 1) * sizeof(double) + 256, 256);
}

extern "C" int txb_reduce_sums_planar_f64(const double* const* x_planes, int nplanes, const double* u,
                                          int64_t nsamp, int order, const double* pivot, double* S0, double* S1,
                                          void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(x_planes && u && pivot && S0 && S1 && workspace, "txb_reduce_sums_planar_f64: null pointer");
    TXB_CHECK_ARG(nplanes >= 1 && nplanes <= PL_MAXCH, "txb_reduce_sums_planar_f64: 1..%d planes, got %d", PL_MAXCH,
                  nplanes);
    TXB_CHECK_ARG(nsamp > 0, "txb_reduce_sums_planar_f64: nsamp must be > 0");
    if (ws_bytes < txb_reduce_planar_workspace_bytes(nplanes, order)) {
        set_error("txb_reduce_sums_planar_f64: workspace too small (%zu bytes)", ws_bytes);
        return TXB_ERR_WORKSPACE;
    }
    PlanarArgs a{};
    for (int c = 0; c < nplanes; ++c) {
        TXB_CHECK_ARG(x_planes[c] != nullptr && (reinterpret_cast<uintptr_t>(x_planes[c]) & 15) == 0,
                      "txb_reduce_sums_planar_f64: plane %d is null or not 16-byte aligned", c);
        a.x[c] = x_planes[c];
    }
    TXB_CHECK_ARG((reinterpret_cast<uintptr_t>(u) & 15) == 0, "txb_reduce_sums_planar_f64: u must be 16-byte aligned");
    a.u = u;
    a.pivot = pivot;
    a.nsamp = nsamp;
    const int64_t nb = planar_blocks();
    a.ws0 = reinterpret_cast<double*>(workspace);
    a.ws1 = a.ws0 + nb * (order + 1);
    cudaStream_t st = (cudaStream_t)stream;
    int rc = TXB_OK;
    TXB_DISPATCH_ORDER(order, {
        switch (nplanes) {
            case 1: rc = launch_planar<K, 1>(a, nb, st); break;
            case 2: rc = launch_planar<K, 2>(a, nb, st); break;
            case 3: rc = launch_planar<K, 3>(a, nb, st); break;
            default: rc = launch_planar<K, 4>(a, nb, st); break;
        }
    });
    if (rc != TXB_OK) return rc;
    rc = launch_sum_parts(a.ws0, nb, order + 1, S0, st);
    if (rc == TXB_OK) rc = launch_sum_parts(a.ws1, nb, (int64_t)nplanes * (order + 1), S1, st);
    return rc;
}

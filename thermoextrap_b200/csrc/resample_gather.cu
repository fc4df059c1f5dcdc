// K2 (index form): bootstrap straight from the sampler's INDEX table for scalar x.
//
//   S[r, :] = sum_j P[idx[r, j], :]
//
// For scalar x the frequency-table detour (K6 atomics + contraction) costs more
// than it saves: one random 16-byte gather per draw (L2-resident pairs) beats
// one random atomic per draw.  The index table (int64, the reference sampler's
// own, /root/reference/src/thermoextrap/data.py:420-421, adaptive_interp.py:543-546)
// is streamed exactly once with coalesced 128-bit loads; (dx, du[, w]) pairs
// are pre-shifted and packed so that a draw is a single LDG.128.
#include "txb_common.cuh"

namespace txb {

int launch_sum_parts(const double* ws, int64_t nparts, int64_t n, double* out, cudaStream_t st);

// packed[s] = (x - cx, u - cu)   and, when weighted, wpk[s] = w
__global__ void __launch_bounds__(256)
    k_pack_shifted(const double* __restrict__ x, int64_t ldx, const double* __restrict__ u,
                   int64_t nsamp, const double* __restrict__ pivot, double2* __restrict__ packed) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nsamp) return;
    const bool has_x = x != nullptr;
    const double dx = has_x ? x[i * ldx] - pivot[0] : 0.0;
    const double du = u[i] - pivot[has_x ? 1 : 0];
    packed[i] = make_double2(dx, du);
}

int launch_pack_shifted(const double* x, int64_t ldx, const double* u, int64_t nsamp,
                        const double* pivot, double2* packed, cudaStream_t st) {
    k_pack_shifted<<<(unsigned)((nsamp + 255) / 256), 256, 0, st>>>(x, ldx, u, nsamp, pivot, packed);
    TXB_CUDA_LAUNCH_CHECK("k_pack_shifted");
    return TXB_OK;
}

template <int K, bool HAS_X, bool HAS_W>
__global__ void __launch_bounds__(256, 2)
    k_resample_gather(const double2* __restrict__ packed, const double* __restrict__ w,
                      const int64_t* __restrict__ idx, int64_t ldi, int64_t ndraw, int64_t ndat,
                      int64_t nrep, int64_t slab_len, double* __restrict__ ws0,
                      double* __restrict__ ws1, int* __restrict__ status) {
    const int64_t rep = blockIdx.y;
    const int64_t j0 = (int64_t)blockIdx.x * slab_len;
    const int64_t j1 = min(ndraw, j0 + slab_len);
    const int64_t* row = idx + rep * ldi;

    double s0[K + 1], s1[K + 1];
#pragma unroll
    for (int j = 0; j <= K; ++j) s0[j] = 0.0, s1[j] = 0.0;
    long long cnt = 0;
    int bad = 0;

    auto body = [&](int64_t i) {
        if ((uint64_t)i >= (uint64_t)ndat) {
            bad = 1;
            return;
        }
        const double2 v = __ldg(packed + i);
        double p = HAS_W ? __ldg(w + i) : 1.0;
        if (HAS_W) s0[0] += p; else ++cnt;
        if (HAS_X) s1[0] = HAS_W ? fma(p, v.x, s1[0]) : s1[0] + v.x;
#pragma unroll
        for (int j = 1; j <= K; ++j) {
            p *= v.y;
            s0[j] += p;
            if (HAS_X) s1[j] = fma(p, v.x, s1[j]);
        }
    };

    const uint64_t pol_keep = l2_policy_evict_last();
    const uint64_t pol_stream = l2_policy_evict_first();
    constexpr int UNR = 4;  // x2 indices per 128-bit load -> 8 gathers in flight per thread
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(row + j0) & 15) == 0);
    int64_t j = j0;
    if (vec_ok) {
        const int64_t npair = (j1 - j0) / 2;
        const longlong2* row2 = reinterpret_cast<const longlong2*>(row + j0);
        const int64_t nfull = npair / (UNR * 256);  // block-uniform trip count
        for (int64_t m = 0; m < nfull; ++m) {
            const int64_t q = m * (UNR * 256) + threadIdx.x;
            longlong2 ii[UNR];
#pragma unroll
            for (int t = 0; t < UNR; ++t) {
                const double2 raw = ldg_stream_hint2(row2 + q + t * 256, pol_stream);
                ii[t].x = __double_as_longlong(raw.x);
                ii[t].y = __double_as_longlong(raw.y);
            }
            // clamp first so that all gathers can be issued back to back
            double2 v[2 * UNR];
            double wv[2 * UNR];
#pragma unroll
            for (int t = 0; t < UNR; ++t) {
                int64_t a = ii[t].x, b = ii[t].y;
                if ((uint64_t)a >= (uint64_t)ndat) a = 0, bad = 1;
                if ((uint64_t)b >= (uint64_t)ndat) b = 0, bad = 1;
                v[2 * t] = ldg_hint2(packed + a, pol_keep);
                v[2 * t + 1] = ldg_hint2(packed + b, pol_keep);
                if (HAS_W) wv[2 * t] = __ldg(w + a), wv[2 * t + 1] = __ldg(w + b);
            }
#pragma unroll
            for (int t = 0; t < 2 * UNR; ++t) {
                double p = HAS_W ? wv[t] : 1.0;
                if (HAS_W) s0[0] += p;
                if (HAS_X) s1[0] = HAS_W ? fma(p, v[t].x, s1[0]) : s1[0] + v[t].x;
#pragma unroll
                for (int jj = 1; jj <= K; ++jj) {
                    p *= v[t].y;
                    s0[jj] += p;
                    if (HAS_X) s1[jj] = fma(p, v[t].x, s1[jj]);
                }
            }
            cnt += 2 * UNR;
        }
        j = j0 + 2 * nfull * (UNR * 256);
    }
    for (int64_t jj = j + threadIdx.x; jj < j1; jj += 256) body(row[jj]);

    if (!HAS_W) s0[0] = (double)cnt;
    if (bad && status) atomicOr(status, 1);

    __shared__ double sm[8][2 * (K + 1)];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
#pragma unroll
    for (int jj = 0; jj <= K; ++jj) {
        const double a = warp_sum(s0[jj]);
        const double b = HAS_X ? warp_sum(s1[jj]) : 0.0;
        if (lane == 0) sm[wrp][jj] = a, sm[wrp][K + 1 + jj] = b;
    }
    __syncthreads();
    if (threadIdx.x < 2 * (K + 1)) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[k][threadIdx.x];
        const int64_t o = ((int64_t)blockIdx.x * nrep + rep) * (K + 1);
        if (threadIdx.x <= K)
            ws0[o + threadIdx.x] = t;
        else if (HAS_X)
            ws1[o + threadIdx.x - (K + 1)] = t;
    }
}

static int64_t gather_slabs(int64_t nrep, int64_t ndraw) {
    const int64_t target = (int64_t)sm_count() * 4;
    int64_t ns = (target + nrep - 1) / nrep;
    const int64_t maxs = (ndraw + 4095) / 4096;
    if (ns > maxs) ns = maxs;
    return ns < 1 ? 1 : ns;
}

}  // namespace txb

using namespace txb;

extern "C" size_t txb_resample_indices_workspace_bytes(int64_t ndat, int order, int64_t nrep,
                                                       int64_t ndraw) {
    const int64_t ns = gather_slabs(nrep, ndraw);
    const size_t dbl = (size_t)(2 * ndat) + (size_t)(2 * ns * nrep * (order + 2)) +
                       (size_t)(2 * nrep * (order + 2)) + 8;
    return align_up(dbl * sizeof(double) + 512, 256);
}

extern "C" int txb_resample_indices_sums_f64(const double* x, const double* u, const double* w,
                                             int64_t ndat, int64_t ldx, int order,
                                             const double* pivot, const int64_t* indices,
                                             int64_t nrep, int64_t ndraw, int64_t ldi, double* S0,
                                             double* S1, int* status, void* workspace,
                                             size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && pivot && indices && S0 && workspace, "txb_resample_indices_sums_f64: null pointer");
    TXB_CHECK_ARG(ndat > 0 && nrep > 0 && nrep <= 65535 && ndraw > 0 && ldi >= ndraw,
                  "txb_resample_indices_sums_f64: bad sizes (nrep must be <= 65535 per call)");
    TXB_CHECK_ARG(!x || (S1 && ldx >= 1), "txb_resample_indices_sums_f64: S1/ldx missing for x");
    const int64_t ns = gather_slabs(nrep, ndraw);
    const size_t need = ((size_t)(2 * ndat) + (size_t)(2 * ns * nrep * (order + 1))) * sizeof(double) + 256;
    if (ws_bytes < need) {
        set_error("txb_resample_indices_sums_f64: workspace too small: %zu < %zu", ws_bytes, need);
        return TXB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    // carve: packed (16-byte aligned) | ws0 | ws1
    uintptr_t base = (reinterpret_cast<uintptr_t>(workspace) + 15) & ~(uintptr_t)15;
    double2* packed = reinterpret_cast<double2*>(base);
    double* ws0 = reinterpret_cast<double*>(packed + ndat);
    double* ws1 = ws0 + ns * nrep * (order + 1);
    k_pack_shifted<<<(unsigned)((ndat + 255) / 256), 256, 0, st>>>(x, x ? ldx : 0, u, ndat, pivot, packed);
    TXB_CUDA_LAUNCH_CHECK("k_pack_shifted");
    if (status && cudaMemsetAsync(status, 0, sizeof(int), st) != cudaSuccess) {
        set_error("cudaMemsetAsync(status) failed");
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    int64_t slab_len = (ndraw + ns - 1) / ns;
    slab_len = (slab_len + 1) & ~(int64_t)1;  // keep slab starts 16-byte aligned
    const dim3 grid((unsigned)((ndraw + slab_len - 1) / slab_len), (unsigned)nrep);
    const int64_t nslab = grid.x;
    TXB_DISPATCH_ORDER(order, {
        if (x) {
            if (w) k_resample_gather<K, true, true><<<grid, 256, 0, st>>>(packed, w, indices, ldi, ndraw, ndat, nrep, slab_len, ws0, ws1, status);
            else   k_resample_gather<K, true, false><<<grid, 256, 0, st>>>(packed, w, indices, ldi, ndraw, ndat, nrep, slab_len, ws0, ws1, status);
        } else {
            if (w) k_resample_gather<K, false, true><<<grid, 256, 0, st>>>(packed, w, indices, ldi, ndraw, ndat, nrep, slab_len, ws0, ws1, status);
            else   k_resample_gather<K, false, false><<<grid, 256, 0, st>>>(packed, w, indices, ldi, ndraw, ndat, nrep, slab_len, ws0, ws1, status);
        }
    });
    TXB_CUDA_LAUNCH_CHECK("k_resample_gather");
    int rc = launch_sum_parts(ws0, nslab, nrep * (order + 1), S0, st);
    if (rc == TXB_OK && x) rc = launch_sum_parts(ws1, nslab, nrep * (order + 1), S1, st);
    return rc;
}

extern "C" int txb_resample_indices_f64(const double* x, const double* u, const double* w,
                                        int64_t ndat, int64_t ldx, int order, int mode,
                                        const int64_t* indices, int64_t nrep, int64_t ndraw,
                                        int64_t ldi, double* out, int* status, void* workspace,
                                        size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && out && workspace && indices, "txb_resample_indices_f64: null pointer");
    TXB_CHECK_ARG(order >= 1, "txb_resample_indices_f64: order=%d must be >= 1", order);
    TXB_CHECK_ARG(mode == TXB_MODE_COMOMENTS ? x != nullptr : (mode == TXB_MODE_X_IS_U && x == nullptr),
                  "txb_resample_indices_f64: mode COMOMENTS needs x, mode X_IS_U needs x == NULL");
    if (ws_bytes < txb_resample_indices_workspace_bytes(ndat, order, nrep, ndraw)) {
        set_error("txb_resample_indices_f64: workspace too small");
        return TXB_ERR_WORKSPACE;
    }
    const int ks = mode == TXB_MODE_X_IS_U ? order + 1 : order;
    double* ws = (double*)workspace;
    double* pivot = ws;  ws += 4;
    double* S0 = ws;     ws += nrep * (order + 2);
    double* S1 = ws;     ws += nrep * (order + 2);
    const size_t used = (size_t)((char*)ws - (char*)workspace);
    int rc = txb_pivot_f64(x, u, ndat, 1, x ? ldx : 1, 0, pivot, stream);
    if (rc != TXB_OK) return rc;
    rc = txb_resample_indices_sums_f64(x, u, w, ndat, ldx, ks, pivot, indices, nrep, ndraw, ldi, S0,
                                       S1, status, ws, ws_bytes - used, stream);
    if (rc != TXB_OK) return rc;
    return txb_finalize_f64(S0, S1, pivot, nrep, 1, 1, order, mode, out, stream);
}

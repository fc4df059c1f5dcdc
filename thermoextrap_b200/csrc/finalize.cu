// Small O(k^2)-per-element kernels on the (2, order+1) comoment tail:
//   finalize      pivot sums -> cmomy layout (binomial re-centring)
//   merge         Chan/Pebay merge of pre-reduced blocks (K3/K4)
//   convert       central <-> raw, .cmom()                (K5)
// All are latency/launch bound (report us, not roofline %).
#include "txb_common.cuh"
#include <mutex>

namespace txb {

__constant__ double c_binom[TXB_MAXK + 2][TXB_MAXK + 2];
// __constant__ memory is per device (and per context): readiness is tracked per device, under a lock
// (a C-ABI caller may drive several GPUs from one process, or call from several threads)
static std::mutex g_binom_mu;
static bool g_binom_ready[64] = {};

static int ensure_binom() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("ensure_binom: cudaGetDevice failed or device index %d out of range", dev);
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(g_binom_mu);
    if (g_binom_ready[dev]) return TXB_OK;
    static double h[TXB_MAXK + 2][TXB_MAXK + 2];
    for (int n = 0; n < TXB_MAXK + 2; ++n)
        for (int k = 0; k < TXB_MAXK + 2; ++k)
            h[n][k] = (k == 0 || k == n) ? 1.0 : (k > n ? 0.0 : h[n - 1][k - 1] + h[n - 1][k]);
    if (cudaMemcpyToSymbol(c_binom, h, sizeof(h)) != cudaSuccess) {
        set_error("cudaMemcpyToSymbol(c_binom) failed");
        return TXB_ERR_CUDA;
    }
    g_binom_ready[dev] = true;
    return TXB_OK;
}

// thread per (rep, chan).  S0[rep][nu][ks+1], S1[rep][nchan][k+1].
__global__ void __launch_bounds__(128)
    k_finalize(const double* __restrict__ S0, const double* __restrict__ S1,
               const double* __restrict__ pivot, int64_t nrep, int64_t nchan, int64_t nu, int k,
               int mode, double* __restrict__ out) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nrep * nchan) return;
    const int64_t rep = idx / nchan, c = idx % nchan;
    const int ks = mode == TXB_MODE_X_IS_U ? k + 1 : k;
    const double* s0 = S0 + (rep * nu + (nu == 1 ? 0 : c)) * (ks + 1);
    const double cu = pivot[(mode == TXB_MODE_X_IS_U ? 0 : nchan) + (nu == 1 ? 0 : c)];
    double* o = out + idx * 2 * (k + 1);

    const double W = s0[0];
    if (!(W != 0.0)) {
        for (int j = 0; j < 2 * (k + 1); ++j) o[j] = 0.0;
        return;
    }
    double a0[TXB_MAXK + 2], m0[TXB_MAXK + 2], pw[TXB_MAXK + 2];
    for (int j = 0; j <= ks; ++j) a0[j] = s0[j] / W;
    const double du = a0[1];
    pw[0] = 1.0;
    for (int j = 1; j <= ks; ++j) pw[j] = pw[j - 1] * (-du);
    for (int q = 0; q <= ks; ++q) {
        double t = 0.0;
        for (int j = 0; j <= q; ++j) t += c_binom[q][j] * pw[q - j] * a0[j];
        m0[q] = t;
    }
    if (mode == TXB_MODE_X_IS_U) {
        const double mean = cu + du;
        o[0] = W;
        o[1] = mean;
        for (int q = 2; q <= k; ++q) o[q] = m0[q];
        o[k + 1] = mean;
        for (int q = 1; q <= k; ++q) o[k + 1 + q] = m0[q + 1];
        return;
    }
    const double* s1 = S1 + idx * (k + 1);
    const double cx = pivot[c];
    const double dx = s1[0] / W;
    o[0] = W;
    o[1] = cu + du;
    for (int q = 2; q <= k; ++q) o[q] = m0[q];
    o[k + 1] = cx + dx;
    for (int q = 1; q <= k; ++q) {
        double t = 0.0;
        for (int j = 0; j <= q; ++j) t += c_binom[q][j] * pw[q - j] * (s1[j] / W - dx * a0[j]);
        o[k + 1 + q] = t;
    }
}

// acc <- acc (+) blk * f   (Chan/Pebay; in-place, highest order first)
__device__ void merge_into(double* acc, const double* blk, double f, int k) {
    const double wb = blk[0] * f;
    if (!(wb != 0.0)) return;
    const double wa = acc[0];
    const double w = wa + wb;
    const double fa = wa / w, fb = wb / w;
    const double dx = blk[k + 1] - acc[k + 1];
    const double du = blk[1] - acc[1];
    double pa[TXB_MAXK + 1], pb[TXB_MAXK + 1];
    pa[0] = pb[0] = 1.0;
    for (int j = 1; j <= k; ++j) pa[j] = pa[j - 1] * (-fb * du), pb[j] = pb[j - 1] * (fa * du);
    auto A = [&](int p, int q) -> double {
        return (p + q == 0) ? 1.0 : (p + q == 1 ? 0.0 : acc[p * (k + 1) + q]);
    };
    auto B = [&](int p, int q) -> double {
        return (p + q == 0) ? 1.0 : (p + q == 1 ? 0.0 : blk[p * (k + 1) + q]);
    };
    for (int p = 1; p >= 0; --p) {
        for (int q = k; q >= 0; --q) {
            if (p + q < 2) continue;
            double t = 0.0;
            for (int j = 0; j <= q; ++j) {
                const double cb = c_binom[q][j];
                t += cb * (fa * A(p, q - j) * pa[j] + fb * B(p, q - j) * pb[j]);
                if (p == 1)
                    t += cb * (fa * A(0, q - j) * (-fb * dx) * pa[j] + fb * B(0, q - j) * (fa * dx) * pb[j]);
            }
            acc[p * (k + 1) + q] = t;
        }
    }
    acc[0] = w;
    acc[1] += fb * du;
    acc[k + 1] += fb * dx;
}

// thread per (rep, val): sequential merge over blocks.  freq == nullptr -> all ones.
__global__ void __launch_bounds__(128)
    k_merge_blocks(const double* __restrict__ data, const int64_t* __restrict__ freq, int64_t nrep,
                   int64_t nblock, int64_t nval, int k, double* __restrict__ out) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nrep * nval) return;
    const int64_t rep = idx / nval, v = idx % nval;
    const int nt = 2 * (k + 1);
    double acc[2 * (TXB_MAXK + 1)], blk[2 * (TXB_MAXK + 1)];
    for (int j = 0; j < nt; ++j) acc[j] = 0.0;
    for (int64_t b = 0; b < nblock; ++b) {
        const double f = freq ? (double)freq[rep * nblock + b] : 1.0;
        if (f == 0.0) continue;
        const double* src = data + (b * nval + v) * nt;
        for (int j = 0; j < nt; ++j) blk[j] = src[j];
        merge_into(acc, blk, f, k);
    }
    double* o = out + idx * nt;
    for (int j = 0; j < nt; ++j) o[j] = acc[j];
}

// kind 0: central -> raw ; 1: raw -> central ; 2: central -> cmom
__global__ void __launch_bounds__(128)
    k_convert(const double* __restrict__ in, double* __restrict__ out, int64_t nelem, int k, int kind) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nelem) return;
    const int nt = 2 * (k + 1);
    const double* a = in + idx * nt;
    double* o = out + idx * nt;
    if (kind == 2) {
        for (int j = 0; j < nt; ++j) o[j] = a[j];
        o[0] = 1.0;
        o[1] = 0.0;
        o[k + 1] = 0.0;
        return;
    }
    const double xa = a[k + 1], ua = a[1];
    const double sgn = kind == 0 ? 1.0 : -1.0;
    double pu[TXB_MAXK + 1];
    pu[0] = 1.0;
    for (int j = 1; j <= k; ++j) pu[j] = pu[j - 1] * (sgn * ua);
    auto M = [&](int p, int q) -> double {
        if (p + q == 0) return 1.0;
        if (kind == 0 && p + q == 1) return 0.0;  // central input: first moments are zero
        return a[p * (k + 1) + q];
    };
    for (int p = 0; p < 2; ++p)
        for (int q = 0; q <= k; ++q) {
            double t = 0.0;
            for (int j = 0; j <= q; ++j) {
                t += c_binom[q][j] * M(p, j) * pu[q - j];
                if (p == 1) t += c_binom[q][j] * M(0, j) * (sgn * xa) * pu[q - j];
            }
            o[p * (k + 1) + q] = t;
        }
    if (kind == 0) {
        o[0] = 1.0;
    } else {
        o[0] = a[0];
        o[1] = ua;
        o[k + 1] = xa;
    }
}

}  // namespace txb

using namespace txb;

extern "C" int txb_finalize_f64(const double* S0, const double* S1, const double* pivot,
                                int64_t nrep, int64_t nchan, int64_t nu, int order, int mode,
                                double* out, void* stream) {
    TXB_CHECK_ARG(S0 && pivot && out, "txb_finalize_f64: null pointer");
    TXB_CHECK_ARG(mode == TXB_MODE_X_IS_U || S1, "txb_finalize_f64: S1 null in comoment mode");
    TXB_CHECK_ARG(order >= 1 && order + (mode == TXB_MODE_X_IS_U) <= TXB_MAXK,
                  "txb_finalize_f64: order %d unsupported", order);
    TXB_CHECK_ARG(nu == 1 || nu == nchan, "txb_finalize_f64: nu must be 1 or nchan");
    if (nrep * nchan <= 0) return TXB_OK;
    int rc = ensure_binom();
    if (rc != TXB_OK) return rc;
    const int64_t n = nrep * nchan;
    k_finalize<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(S0, S1, pivot, nrep, nchan,
                                                                              nu, order, mode, out);
    TXB_CUDA_LAUNCH_CHECK("k_finalize");
    return TXB_OK;
}

extern "C" int txb_reduce_data_f64(const double* data, int64_t nblock, int64_t nval, int order,
                                   double* out, void* stream) {
    TXB_CHECK_ARG(data && out, "txb_reduce_data_f64: null pointer");
    TXB_CHECK_ARG(order >= 1 && order <= TXB_MAXK, "txb_reduce_data_f64: order %d unsupported", order);
    if (nval <= 0) return TXB_OK;
    int rc = ensure_binom();
    if (rc != TXB_OK) return rc;
    k_merge_blocks<<<(unsigned)((nval + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        data, nullptr, 1, nblock, nval, order, out);
    TXB_CUDA_LAUNCH_CHECK("k_merge_blocks");
    return TXB_OK;
}

extern "C" int txb_resample_data_f64(const double* data, const int64_t* freq, int64_t nrep,
                                     int64_t nblock, int64_t nval, int order, double* out,
                                     void* stream) {
    TXB_CHECK_ARG(data && freq && out, "txb_resample_data_f64: null pointer");
    TXB_CHECK_ARG(order >= 1 && order <= TXB_MAXK, "txb_resample_data_f64: order %d unsupported", order);
    const int64_t n = nrep * nval;
    if (n <= 0) return TXB_OK;
    int rc = ensure_binom();
    if (rc != TXB_OK) return rc;
    k_merge_blocks<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(data, freq, nrep, nblock,
                                                                                  nval, order, out);
    TXB_CUDA_LAUNCH_CHECK("k_merge_blocks");
    return TXB_OK;
}

extern "C" int txb_convert_f64(const double* in, double* out, int64_t nelem, int order, int kind,
                               void* stream) {
    TXB_CHECK_ARG(in && out, "txb_convert_f64: null pointer");
    TXB_CHECK_ARG(order >= 1 && order <= TXB_MAXK, "txb_convert_f64: order %d unsupported", order);
    TXB_CHECK_ARG(kind >= 0 && kind <= 2, "txb_convert_f64: kind %d unknown", kind);
    if (nelem <= 0) return TXB_OK;
    int rc = ensure_binom();
    if (rc != TXB_OK) return rc;
    k_convert<<<(unsigned)((nelem + 127) / 128), 128, 0, (cudaStream_t)stream>>>(in, out, nelem, order, kind);
    TXB_CUDA_LAUNCH_CHECK("k_convert");
    return TXB_OK;
}

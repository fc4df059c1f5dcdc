// K7 device-resident: comoment records -> beta derivatives -> (-log) -> / i! -> Taylor sum, one launch.
//
// The reference evaluates lambdified sympy expressions on xarray objects and sums the series with numpy
// (/root/reference/src/thermoextrap/models.py:371 funcs[i](*args), :373-377 minus_log and / i!, :535-565
// Taylor sum, :909-932 InterpModel).  Here the four numeric beta families are GENERATED device functions
// (beta_derivs_gen.cuh, built by gen_beta_derivs.py from the reference's recurrences, beta.py:32-266) and the
// kernel reads the bootstrap kernel's output [nelem, 2, K+1] where it lies in HBM: nothing but the final
// [nalpha, nelem] predictions (or the [order+1, nelem] derivatives) ever crosses PCIe.
//   k_beta_fused     one thread per (replicate x channel) record
//   k_combine        out[a, e] (+)= sum_t W[a, t] D[t, e]      (InterpModel: W = alpha powers x inverse
//                    Hermite matrix, D = one state's normalised derivatives)
//   k_mean_var       mean / variance over the replicate axis of [A, R, B]  (stack.py:157-182 to_mean_var)
#include "beta_derivs_gen.cuh"
#include "txb_common.cuh"

namespace txb {

constexpr int BF_MODE_SUM = 0, BF_MODE_CUMSUM = 1, BF_MODE_TERMS = 2, BF_MODE_DERIVS = 3;
constexpr int BF_FLAG_MINUS_LOG = 1, BF_FLAG_NORM = 2;

struct BetaFusedArgs {
    const double* comom;   // [nelem, 2, kdim], central comoments in cmomy's layout (data.py:1233-1236)
    int64_t nelem;
    int kdim;              // K + 1 of the stored records (>= order + 1)
    int flags, mode;
    const double* dalpha;  // [nalpha] (device)
    int nalpha;
    double* out;
    // lnPi families (4 central, 5 raw; lnpi.py:42-116): d_0 = lnPi0, d_1 = mu.N - <u>, d_n = -d^(n-1)<u>/dbeta^(n-1)
    const double* lnpi0;   // [ninner]
    const double* mudotn;  // [ninner]; element e uses index e % ninner (leading dims such as rep broadcast)
    int64_t ninner;
};

template <int FAM, int ORDER>
__global__ void __launch_bounds__(128) k_beta_fused(const BetaFusedArgs a) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= a.nelem) return;
    const double* rec = a.comom + e * 2 * a.kdim;
    // BF / BO: the beta family and order actually evaluated.  The lnPi families are the u_ave families one order
    // lower, shifted by one and offset by lnPi0 and mu.N.
    constexpr int BF = FAM < 4 ? FAM : (FAM == 4 ? 1 : 3);
    constexpr int BO = FAM < 4 ? ORDER : (ORDER > 0 ? ORDER - 1 : 0);
    // central moments of the record.  x_ave families (0, 2): <du^n> from row 0, <dx du^n> from row 1.
    // u_ave families (1, 3) are used with x_is_u data (x == u, beta.py:633-646): <du^n> = <dx du^(n-1)>, which
    // reaches n = BO + 1 -- the moment the BO-th derivative of <u> needs.
    constexpr int NU = BO + 2;
    double r0[NU], r1[NU];
    double ub, xb;
    if (BF == 0 || BF == 2) {
#pragma unroll
        for (int n = 0; n <= BO; ++n) {
            r0[n] = rec[n];
            r1[n] = rec[a.kdim + n];
        }
        r0[BO + 1] = r1[BO + 1] = 0.0;
        ub = r0[1], xb = r1[0];
    } else {
#pragma unroll
        for (int n = 1; n <= BO + 1; ++n) r0[n] = rec[a.kdim + n - 1];
#pragma unroll
        for (int n = 0; n < NU; ++n) r1[n] = 0.0;
        r0[0] = 1.0;
        ub = r0[1], xb = r0[1];
    }
    if (BF >= 2) {
        // central -> raw (what DataCentralMomentsBase.rmom() hands the raw families, data.py:845-848):
        //   <u^n> = sum_k C(n,k) <du^k> ubar^(n-k),  <x u^n> = sum_k C(n,k) ubar^(n-k) (xbar <du^k> + <dx du^k>)
        double cdu[NU], cdx[NU], up[NU];
#pragma unroll
        for (int k = 0; k < NU; ++k) {
            cdu[k] = k == 0 ? 1.0 : k == 1 ? 0.0 : r0[k];
            cdx[k] = k == 0 ? 0.0 : r1[k];
        }
        up[0] = 1.0;
#pragma unroll
        for (int k = 1; k < NU; ++k) up[k] = up[k - 1] * ub;
#pragma unroll
        for (int n = 0; n < NU; ++n) {
            double s0 = 0.0, s1 = 0.0, binom = 1.0;
#pragma unroll
            for (int k = 0; k <= n; ++k) {
                s0 += binom * cdu[k] * up[n - k];
                s1 += binom * up[n - k] * (xb * cdu[k] + cdx[k]);
                binom = binom * (double)(n - k) / (double)(k + 1);
            }
            r0[n] = s0;
            r1[n] = s1;
        }
    }
    double db[BO + 1];
    if (BF == 0) beta_derivs_x_ave_central<BO>(r0, r1, db);
    if (BF == 1) beta_derivs_u_ave_central<BO>(r0, r1, db);
    if (BF == 2) beta_derivs_x_ave_raw<BO>(r0, r1, db);
    if (BF == 3) beta_derivs_u_ave_raw<BO>(r0, r1, db);
    double d[ORDER + 1];
    if (FAM < 4) {
#pragma unroll
        for (int n = 0; n <= ORDER; ++n) d[n] = db[n < BO + 1 ? n : BO];
    } else {
        const int64_t ei = e % a.ninner;
        d[0] = a.lnpi0[ei];
        if (ORDER >= 1) d[1 <= ORDER ? 1 : 0] = a.mudotn[ei] - db[0];
#pragma unroll
        for (int n = 2; n <= ORDER; ++n) d[n] = -db[n - 1];
    }

    if (a.flags & BF_FLAG_MINUS_LOG) {
        // y = -log f (models.py:261-281 SymMinusLog): with g = log f,  f g' = f'  differentiated n - 1 times gives
        //   g^(n) = ( f^(n) - sum_{k=1}^{n-1} C(n-1, k) f^(k) g^(n-k) ) / f
        double g[ORDER + 1];
        const double f = d[0];
        g[0] = log(f);
#pragma unroll
        for (int n = 1; n <= ORDER; ++n) {
            double s = d[n], binom = 1.0;   // C(n-1, 0)
#pragma unroll
            for (int k = 1; k <= n - 1; ++k) {
                binom = binom * (double)(n - k) / (double)k;   // C(n-1, k)
                s -= binom * d[k] * g[n - k];
            }
            g[n] = s / f;
        }
#pragma unroll
        for (int n = 0; n <= ORDER; ++n) d[n] = -g[n];
    }
    if ((a.flags & BF_FLAG_NORM) || a.mode != BF_MODE_DERIVS) {
        double fact = 1.0;
#pragma unroll
        for (int n = 1; n <= ORDER; ++n) {
            fact *= (double)n;
            d[n] = d[n] / fact;
        }
    }
    if (a.mode == BF_MODE_DERIVS) {
#pragma unroll
        for (int n = 0; n <= ORDER; ++n) a.out[(int64_t)n * a.nelem + e] = d[n];
        return;
    }
    for (int al = 0; al < a.nalpha; ++al) {
        const double da = a.dalpha[al];
        double pw = 1.0, acc = 0.0;
#pragma unroll
        for (int p = 0; p <= ORDER; ++p) {
            const double term = pw * d[p];
            acc += term;
            if (a.mode == BF_MODE_CUMSUM) a.out[((int64_t)al * (ORDER + 1) + p) * a.nelem + e] = acc;
            if (a.mode == BF_MODE_TERMS) a.out[((int64_t)al * (ORDER + 1) + p) * a.nelem + e] = term;
            pw *= da;
        }
        if (a.mode == BF_MODE_SUM) a.out[(int64_t)al * a.nelem + e] = acc;
    }
}

template <int FAM>
int launch_beta_fused(const BetaFusedArgs& a, int order, cudaStream_t st) {
    const unsigned grid = (unsigned)((a.nelem + 127) / 128);
    switch (order) {
#define TXB_BF_CASE(O) case O: k_beta_fused<FAM, O><<<grid, 128, 0, st>>>(a); break;
        TXB_BF_CASE(0) TXB_BF_CASE(1) TXB_BF_CASE(2) TXB_BF_CASE(3) TXB_BF_CASE(4) TXB_BF_CASE(5)
        TXB_BF_CASE(6) TXB_BF_CASE(7) TXB_BF_CASE(8) TXB_BF_CASE(9) TXB_BF_CASE(10)
#undef TXB_BF_CASE
        default: return TXB_ERR_UNSUPPORTED;
    }
    return TXB_OK;
}

// out[a, e] = (accumulate ? out[a, e] : 0) + sum_t W[a, t] D[t, e]
__global__ void __launch_bounds__(256)
    k_combine(const double* __restrict__ W, int nalpha, int nt, const double* __restrict__ D, int64_t nelem,
              int accumulate, double* __restrict__ out) {
    extern __shared__ double sW[];   // [nalpha_tile = gridDim.y rows at a time][nt]
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int a0 = blockIdx.y * 8; a0 < nalpha; a0 += gridDim.y * 8) {
        const int na = min(8, nalpha - a0);
        __syncthreads();
        for (int i = threadIdx.x; i < na * nt; i += blockDim.x) sW[i] = W[(int64_t)a0 * nt + i];
        __syncthreads();
        if (e < nelem) {
            double acc[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[q] = 0.0;
            for (int t = 0; t < nt; ++t) {
                const double v = D[(int64_t)t * nelem + e];
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q < na) acc[q] += sW[q * nt + t] * v;
            }
#pragma unroll
            for (int q = 0; q < 8; ++q)
                if (q < na) {
                    double* o = out + (int64_t)(a0 + q) * nelem + e;
                    *o = accumulate ? *o + acc[q] : acc[q];
                }
        }
    }
}

// in[A, R, B] -> mean[A, B], var[A, B] over R (ddof delta degrees of freedom).  block = 32 (b) x 8 (r slices);
// per-thread Welford, slices merged pairwise (Chan): no cancellation, one pass over the data.
__global__ void __launch_bounds__(256)
    k_mean_var(const double* __restrict__ in, int64_t A, int64_t R, int64_t B, int ddof,
               double* __restrict__ mean, double* __restrict__ var) {
    __shared__ double s_n[8][32], s_m[8][32], s_q[8][32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t b = (int64_t)blockIdx.x * 32 + tx;
    const int64_t a = blockIdx.y;
    double n = 0.0, m = 0.0, q = 0.0;
    if (b < B) {
        const double* p = in + (a * R) * B + b;
        for (int64_t r = ty; r < R; r += 8) {
            const double v = p[r * B];
            n += 1.0;
            const double dlt = v - m;
            m += dlt / n;
            q += dlt * (v - m);
        }
    }
    s_n[ty][tx] = n, s_m[ty][tx] = m, s_q[ty][tx] = q;
    __syncthreads();
    if (ty == 0 && b < B) {
        for (int k = 1; k < 8; ++k) {
            const double n2 = s_n[k][tx], m2 = s_m[k][tx], q2 = s_q[k][tx];
            if (n2 == 0.0) continue;
            const double nt = n + n2, dlt = m2 - m;
            m += dlt * (n2 / nt);
            q += q2 + dlt * dlt * (n * n2 / nt);
            n = nt;
        }
        mean[a * B + b] = m;
        var[a * B + b] = q / (n - (double)ddof);
    }
}

}  // namespace txb

using namespace txb;

extern "C" int txb_beta_predict_f64(const double* comom, int64_t nelem, int kdim, int family, int order,
                                    int flags, int mode, const double* dalpha, int nalpha, double* out,
                                    void* stream) {
    TXB_CHECK_ARG(comom && out, "txb_beta_predict_f64: null pointer");
    TXB_CHECK_ARG(family >= 0 && family < BETA_GEN_NFAMILY, "txb_beta_predict_f64: family %d outside [0, %d)", family,
                  BETA_GEN_NFAMILY);
    TXB_CHECK_ARG(mode >= BF_MODE_SUM && mode <= BF_MODE_DERIVS, "txb_beta_predict_f64: bad mode %d", mode);
    TXB_CHECK_ARG(order >= 0 && kdim >= order + 1, "txb_beta_predict_f64: records of %d moments cannot give order %d",
                  kdim, order);
    TXB_CHECK_ARG(mode == BF_MODE_DERIVS || (dalpha && nalpha >= 0), "txb_beta_predict_f64: dalpha is NULL");
    if (order > BETA_GEN_MAX_ORDER) {
        set_error("txb_beta_predict_f64: order %d above the generated maximum %d (use txb_eval_exprs_f64)", order,
                  BETA_GEN_MAX_ORDER);
        return TXB_ERR_UNSUPPORTED;
    }
    if (nelem == 0 || (mode != BF_MODE_DERIVS && nalpha == 0)) return TXB_OK;
    BetaFusedArgs a{comom, nelem, kdim, flags, mode, dalpha, nalpha, out, nullptr, nullptr, 1};
    cudaStream_t st = (cudaStream_t)stream;
    int rc = TXB_OK;
    switch (family) {
        case 0: rc = launch_beta_fused<0>(a, order, st); break;
        case 1: rc = launch_beta_fused<1>(a, order, st); break;
        case 2: rc = launch_beta_fused<2>(a, order, st); break;
        default: rc = launch_beta_fused<3>(a, order, st); break;
    }
    if (rc != TXB_OK) return rc;
    TXB_CUDA_LAUNCH_CHECK("k_beta_fused");
    return TXB_OK;
}

extern "C" int txb_lnpi_predict_f64(const double* comom, int64_t nelem, int kdim, int central, int order, int flags,
                                    int mode, const double* dalpha, int nalpha, const double* lnpi0,
                                    const double* mudotn, int64_t ninner, double* out, void* stream) {
    TXB_CHECK_ARG(comom && out && lnpi0 && mudotn, "txb_lnpi_predict_f64: null pointer");
    TXB_CHECK_ARG(mode >= BF_MODE_SUM && mode <= BF_MODE_DERIVS, "txb_lnpi_predict_f64: bad mode %d", mode);
    TXB_CHECK_ARG(order >= 0 && kdim >= order && kdim >= 1,
                  "txb_lnpi_predict_f64: x_is_u records of %d moments cannot give lnPi order %d", kdim, order);
    TXB_CHECK_ARG(ninner >= 1 && nelem % ninner == 0, "txb_lnpi_predict_f64: nelem=%lld is not a multiple of ninner=%lld",
                  (long long)nelem, (long long)ninner);
    TXB_CHECK_ARG(mode == BF_MODE_DERIVS || (dalpha && nalpha >= 0), "txb_lnpi_predict_f64: dalpha is NULL");
    if (order > BETA_GEN_MAX_ORDER) {
        set_error("txb_lnpi_predict_f64: order %d above the generated maximum %d (use txb_eval_exprs_f64)", order,
                  BETA_GEN_MAX_ORDER);
        return TXB_ERR_UNSUPPORTED;
    }
    if (nelem == 0 || (mode != BF_MODE_DERIVS && nalpha == 0)) return TXB_OK;
    TXB_CHECK_ARG((flags & BF_FLAG_MINUS_LOG) == 0, "txb_lnpi_predict_f64: the lnPi family has no -log form");
    BetaFusedArgs a{comom, nelem, kdim, flags & BF_FLAG_NORM, mode, dalpha, nalpha, out, lnpi0, mudotn, ninner};
    cudaStream_t st = (cudaStream_t)stream;
    const int rc = central ? launch_beta_fused<4>(a, order, st) : launch_beta_fused<5>(a, order, st);
    if (rc != TXB_OK) return rc;
    TXB_CUDA_LAUNCH_CHECK("k_beta_fused<lnPi>");
    return TXB_OK;
}

extern "C" int txb_combine_f64(const double* W, int nalpha, int nt, const double* D, int64_t nelem, int accumulate,
                               double* out, void* stream) {
    TXB_CHECK_ARG(W && D && out, "txb_combine_f64: null pointer");
    TXB_CHECK_ARG(nalpha >= 0 && nt >= 1 && nt <= 512 && nelem >= 0, "txb_combine_f64: bad sizes (nt <= 512)");
    if (nalpha == 0 || nelem == 0) return TXB_OK;
    const int gy = (nalpha + 7) / 8;
    const dim3 grid((unsigned)((nelem + 255) / 256), (unsigned)(gy < 65535 ? gy : 65535));
    k_combine<<<grid, 256, (size_t)8 * nt * sizeof(double), (cudaStream_t)stream>>>(W, nalpha, nt, D, nelem,
                                                                                   accumulate, out);
    TXB_CUDA_LAUNCH_CHECK("k_combine");
    return TXB_OK;
}

extern "C" int txb_mean_var_f64(const double* in, int64_t na, int64_t nrep, int64_t nb, int ddof, double* mean,
                                double* var, void* stream) {
    TXB_CHECK_ARG(in && mean && var, "txb_mean_var_f64: null pointer");
    TXB_CHECK_ARG(na >= 0 && nrep >= 1 && nb >= 0 && na <= 65535, "txb_mean_var_f64: bad sizes (na <= 65535)");
    TXB_CHECK_ARG(ddof >= 0 && ddof < nrep, "txb_mean_var_f64: ddof=%d must be in [0, nrep)", ddof);
    if (na == 0 || nb == 0) return TXB_OK;
    const dim3 grid((unsigned)((nb + 31) / 32), (unsigned)na);
    k_mean_var<<<grid, 256, 0, (cudaStream_t)stream>>>(in, na, nrep, nb, ddof, mean, var);
    TXB_CUDA_LAUNCH_CHECK("k_mean_var");
    return TXB_OK;
}

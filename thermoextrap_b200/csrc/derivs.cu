// K7: batched evaluation of the sympy-derived derivative expressions.
//
// The reference evaluates  out[i] = funcs[i](*args)  with lambdified sympy code on
// xarray objects (/root/reference/src/thermoextrap/models.py:241-243, 371), i.e. a
// handful of numpy kernels per order.  Every expression the reference builds
// (beta.py:32-266 recurrences, lnpi.py:42-116, volume.py:63-78, the -log
// transform models.py:261-281) expands to
//     sum_t  coef_t * prod_f var[f]^pow_f        (integer powers, possibly < 0)
// plus at most a  coef * log(var)  term, so one table-driven kernel evaluates all
// orders for every (replicate x state point x channel) element in a single launch.
// Terms are summed in sympy's own term order.  Taylor / interpolation sums
// (models.py:535-565, 929-932) reuse the same kernel shape: see k_poly_predict.
#include "txb_common.cuh"

namespace txb {

// vars[nvar][nelem]; expression i owns terms [term_ptr[i], term_ptr[i+1]);
// term t owns factors [fac_ptr[t], fac_ptr[t+1]) with (fac_var, fac_pow);
// fac_pow == 127 marks log(var).
__global__ void __launch_bounds__(256)
    k_eval_exprs(const double* __restrict__ vars, int64_t nelem, int nexpr,
                 const int* __restrict__ term_ptr, const double* __restrict__ coef,
                 const int* __restrict__ fac_ptr, const int* __restrict__ fac_var,
                 const int* __restrict__ fac_pow, double* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nelem) return;
    for (int i = 0; i < nexpr; ++i) {
        double acc = 0.0;
        for (int t = term_ptr[i]; t < term_ptr[i + 1]; ++t) {
            double term = coef[t];
            for (int f = fac_ptr[t]; f < fac_ptr[t + 1]; ++f) {
                const double v = vars[(int64_t)fac_var[f] * nelem + e];
                const int p = fac_pow[f];
                if (p == 127) {
                    term *= log(v);
                } else {
                    double pw = 1.0;   // p == 0 -> 1 (sympy never emits it, the ABI accepts it)
                    const int ap = p < 0 ? -p : p;
                    for (int q = 0; q < ap; ++q) pw *= v;
                    term = p < 0 ? term / pw : term * pw;
                }
            }
            acc += term;
        }
        out[(int64_t)i * nelem + e] = acc;
    }
}

// out[a][e] (mode 0: sum, mode 1: cumsum -> out[a][p][e], mode 2: no sum -> out[a][p][e])
//   = sum_p coefs[p][e] * dalpha[a]^p          (ExtrapModel.predict, InterpModel.predict)
__global__ void __launch_bounds__(256)
    k_poly_predict(const double* __restrict__ coefs, int np1, int64_t nelem,
                   const double* __restrict__ dalpha, int nalpha, int mode, double* __restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nelem) return;
    for (int a = blockIdx.y; a < nalpha; a += gridDim.y) {   // any nalpha: grid.y is capped at 65535
    const double d = dalpha[a];
    double pw = 1.0, acc = 0.0;
    for (int p = 0; p < np1; ++p) {
        const double term = pw * coefs[(int64_t)p * nelem + e];
        acc += term;
        if (mode == 1) out[((int64_t)a * np1 + p) * nelem + e] = acc;
        if (mode == 2) out[((int64_t)a * np1 + p) * nelem + e] = term;
        pw *= d;
    }
    if (mode == 0) out[(int64_t)a * nelem + e] = acc;
    }
}

}  // namespace txb

using namespace txb;

extern "C" int txb_eval_exprs_f64(const double* vars, int64_t nelem, int nexpr, const int* term_ptr,
                                  const double* coef, const int* fac_ptr, const int* fac_var,
                                  const int* fac_pow, double* out, void* stream) {
    TXB_CHECK_ARG(vars && term_ptr && coef && fac_ptr && fac_var && fac_pow && out,
                  "txb_eval_exprs_f64: null pointer");
    TXB_CHECK_ARG(nexpr >= 0 && nelem >= 0, "txb_eval_exprs_f64: negative size");
    if (nelem == 0 || nexpr == 0) return TXB_OK;
    k_eval_exprs<<<(unsigned)((nelem + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        vars, nelem, nexpr, term_ptr, coef, fac_ptr, fac_var, fac_pow, out);
    TXB_CUDA_LAUNCH_CHECK("k_eval_exprs");
    return TXB_OK;
}

extern "C" int txb_poly_predict_f64(const double* coefs, int np1, int64_t nelem, const double* dalpha,
                                    int nalpha, int mode, double* out, void* stream) {
    TXB_CHECK_ARG(coefs && dalpha && out, "txb_poly_predict_f64: null pointer");
    TXB_CHECK_ARG(np1 >= 1 && nalpha >= 0 && mode >= 0 && mode <= 2,
                  "txb_poly_predict_f64: bad sizes/mode");
    if (nelem == 0 || nalpha == 0) return TXB_OK;
    const dim3 grid((unsigned)((nelem + 255) / 256), (unsigned)(nalpha < 65535 ? nalpha : 65535));
    k_poly_predict<<<grid, 256, 0, (cudaStream_t)stream>>>(coefs, np1, nelem, dalpha, nalpha, mode, out);
    TXB_CUDA_LAUNCH_CHECK("k_poly_predict");
    return TXB_OK;
}

// PerturbModel.predict (free-energy perturbation / exponential reweighting) as one streaming pass.
//
//   out[a, c] = sum_s e_a(s) x[s, c] / sum_s e_a(s),   e_a(s) = exp(-dalpha_a u_s - max_s(-dalpha_a u_s))
//
// replaces the numpy / xarray expression of /root/reference/src/thermoextrap/models.py:1009-1046
// (legacy/reweight.py:89-101): dalpha_uv = -dalpha * uv; expvals = exp(dalpha_uv - max); num =
// dot(expvals, xv) / N; den = mean(expvals); num / den.  Same data and streaming shape as K1:
// every sample row is read once per tile of PT_NA alphas, HBM-bound for wide x.
//   * max_s(-dalpha u_s) is -dalpha * min(u) or -dalpha * max(u): one min/max pass over u.
//   * lane <-> sample for the exponentials (one exp per sample and alpha), then the 32 weights of a
//     warp are broadcast by shuffle while lanes walk the channels of each row (coalesced).
#include "txb_common.cuh"

namespace txb {

constexpr int PT_THREADS = 256;
constexpr int PT_NW = PT_THREADS / 32;
constexpr int PT_NA = 4;   // alphas per CTA
constexpr int PT_QC = 4;   // 32-channel chunks per CTA (widest form; narrower x uses QC = 1 or 2 with deeper row unrolling)

struct PtArgs {
    const double* x;       // [nsamp, ldx]
    const double* u;       // [nsamp]
    const double* dalpha;  // [nalpha]
    const double* umm;     // {min u, max u}
    double* part;          // [nslab][nalpha][nchan + 1]  (last entry: sum of the weights)
    int64_t nsamp, nchan, ldx, slab_len;
    int nalpha;
};

__global__ void __launch_bounds__(256) k_minmax_part(const double* __restrict__ u, int64_t n, double* __restrict__ part) {
    double lo = INFINITY, hi = -INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
        const double v = ldg_stream1(u + i);
        lo = fmin(lo, v);
        hi = fmax(hi, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    __shared__ double s[2][8];
    if ((threadIdx.x & 31) == 0) s[0][threadIdx.x >> 5] = lo, s[1][threadIdx.x >> 5] = hi;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) lo = fmin(lo, s[0][k]), hi = fmax(hi, s[1][k]);
        part[2 * blockIdx.x] = lo;
        part[2 * blockIdx.x + 1] = hi;
    }
}

__global__ void k_minmax_final(const double* __restrict__ part, int nparts, double* __restrict__ umm) {
    double lo = INFINITY, hi = -INFINITY;
    for (int i = threadIdx.x; i < nparts; i += 32) lo = fmin(lo, part[2 * i]), hi = fmax(hi, part[2 * i + 1]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (threadIdx.x == 0) umm[0] = lo, umm[1] = hi;
}

// QC = 32-channel chunks per CTA, UNR = rows whose loads are in flight per warp.  ncu on the first form (QC = 4,
// UNR = 4 for every width): long-scoreboard stalls at 16 warps per SM -- the row loop is a chain of UNR-row latency
// rounds, so x of <= 64 channels (half the loads masked off) had 8 loads per lane in flight.  Narrow x now spends the
// registers on rows instead of masked chunks: QC * UNR = 16 loads per lane in every form.
template <bool SCALAR, int QC_, int UNR>
__global__ void __launch_bounds__(PT_THREADS) k_perturb(const PtArgs a) {
    constexpr int QC = SCALAR ? 1 : QC_;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int a0 = blockIdx.y * PT_NA;
    const int64_t c0 = (int64_t)blockIdx.z * (QC * 32);
    const int64_t s_begin = (int64_t)blockIdx.x * a.slab_len;
    const int64_t s_end = min(a.nsamp, s_begin + a.slab_len);
    const double umin = a.umm[0], umax = a.umm[1];
    double nda[PT_NA], m[PT_NA];
#pragma unroll
    for (int i = 0; i < PT_NA; ++i) {
        nda[i] = a0 + i < a.nalpha ? -1.0 * a.dalpha[a0 + i] : 0.0;
        m[i] = fmax(nda[i] * umin, nda[i] * umax);
    }
    double den[PT_NA], num[PT_NA][QC];
#pragma unroll
    for (int i = 0; i < PT_NA; ++i) {
        den[i] = 0.0;
#pragma unroll
        for (int q = 0; q < QC; ++q) num[i][q] = 0.0;
    }
    for (int64_t s0 = s_begin + (int64_t)wrp * 32; s0 < s_end; s0 += PT_NW * 32) {
        const int64_t s = s0 + lane;
        const bool ok = s < s_end;
        const double uu = ok ? ldg_stream1(a.u + s) : 0.0;
        double e[PT_NA];
#pragma unroll
        for (int i = 0; i < PT_NA; ++i) {
            e[i] = ok ? exp(nda[i] * uu - m[i]) : 0.0;
            den[i] += e[i];
        }
        if (SCALAR) {
            const double xx = ok ? ldg_stream1(a.x + s * a.ldx) : 0.0;
#pragma unroll
            for (int i = 0; i < PT_NA; ++i) num[i][0] = fma(e[i], xx, num[i][0]);
        } else {
            const int nrow = (int)min((int64_t)32, s_end - s0);
            for (int j0 = 0; j0 < nrow; j0 += UNR) {
                double xv[UNR][QC];
#pragma unroll
                for (int jj = 0; jj < UNR; ++jj) {   // all loads of UNR rows first (rows past the slab end: zeros)
                    const double* row = a.x + (s0 + j0 + jj) * a.ldx + c0 + lane;
#pragma unroll
                    for (int q = 0; q < QC; ++q)
                        xv[jj][q] = (j0 + jj < nrow && c0 + 32 * q + lane < a.nchan) ? ldg_stream1(row + 32 * q) : 0.0;
                }
#pragma unroll
                for (int jj = 0; jj < UNR; ++jj) {
#pragma unroll
                    for (int i = 0; i < PT_NA; ++i) {
                        const double ej = __shfl_sync(0xffffffffu, e[i], (j0 + jj) & 31);
#pragma unroll
                        for (int q = 0; q < QC; ++q) num[i][q] = fma(ej, xv[jj][q], num[i][q]);
                    }
                }
            }
        }
    }
    // ---- block reduction in a fixed order, then this slab's partial
    __shared__ double red[PT_NW][PT_NA][QC * 32 + 1];
#pragma unroll
    for (int i = 0; i < PT_NA; ++i) {
        const double d = warp_sum(den[i]);
        if (SCALAR) {
            const double v = warp_sum(num[i][0]);
            if (lane == 0) red[wrp][i][0] = v;
        } else {
#pragma unroll
            for (int q = 0; q < QC; ++q) red[wrp][i][32 * q + lane] = num[i][q];
        }
        if (lane == 0) red[wrp][i][QC * 32] = d;
    }
    __syncthreads();
    const int ncol = SCALAR ? 1 : QC * 32;
    for (int e = threadIdx.x; e < PT_NA * (ncol + 1); e += PT_THREADS) {
        const int i = e / (ncol + 1), col = e % (ncol + 1);
        if (a0 + i >= a.nalpha) continue;
        const int src = col == ncol ? QC * 32 : col;
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < PT_NW; ++w) v += red[w][i][src];
        double* dst = a.part + ((int64_t)blockIdx.x * a.nalpha + a0 + i) * (a.nchan + 1);
        if (col == ncol) {
            if (blockIdx.z == 0) dst[a.nchan] = v;
        } else if (c0 + col < a.nchan) {
            dst[c0 + col] = v;
        }
    }
}

__global__ void __launch_bounds__(256) k_perturb_final(const double* __restrict__ part, int64_t nslab, int nalpha,
                                                       int64_t nchan, double* __restrict__ out) {
    // one warp per output element: lanes stride over the slabs (fixed order -> deterministic), then a shuffle sum
    const int64_t e = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (e >= (int64_t)nalpha * nchan) return;
    const int lane = threadIdx.x & 31;
    const int64_t ai = e / nchan, c = e % nchan;
    double num = 0.0, den = 0.0;
    for (int64_t s = lane; s < nslab; s += 32) {
        const double* p = part + (s * nalpha + ai) * (nchan + 1);
        num += p[c];
        den += p[nchan];
    }
    num = warp_sum(num);
    den = warp_sum(den);
    if (lane == 0) out[e] = num / den;
}

// 32-channel chunks per CTA for x of `nchan` channels
static int pt_qc(int64_t nchan) { return nchan <= 32 ? 1 : nchan <= 64 ? 2 : PT_QC; }

static int64_t pt_nslab(int64_t nsamp, int64_t nchan, int nalpha) {
    const int64_t cw = pt_qc(nchan) * 32;
    const int64_t tiles = ((nalpha + PT_NA - 1) / PT_NA) * (nchan == 1 ? 1 : (nchan + cw - 1) / cw);
    int64_t n = (4 * (int64_t)sm_count() + tiles - 1) / tiles;
    const int64_t most = (nsamp + PT_THREADS - 1) / PT_THREADS;
    if (n > most) n = most;
    return n < 1 ? 1 : n;
}

}  // namespace txb

using namespace txb;

extern "C" size_t txb_perturb_workspace_bytes(int64_t nsamp, int64_t nchan, int nalpha) {
    if (nchan < 1) nchan = 1;
    const size_t parts = (size_t)pt_nslab(nsamp, nchan, nalpha) * (size_t)nalpha * (size_t)(nchan + 1);
    return align_up((parts + 2 * 1024 + 8) * sizeof(double), 256);
}

extern "C" int txb_perturb_predict_f64(const double* x, const double* u, int64_t nsamp, int64_t nchan,
                                       int64_t ldx, const double* dalpha, int nalpha, double* out,
                                       void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(x && u && dalpha && out && workspace, "txb_perturb_predict_f64: null pointer");
    TXB_CHECK_ARG(nsamp >= 1 && nchan >= 1 && ldx >= nchan && nalpha >= 1,
                  "txb_perturb_predict_f64: bad sizes nsamp=%lld nchan=%lld ldx=%lld nalpha=%d", (long long)nsamp,
                  (long long)nchan, (long long)ldx, nalpha);
    if (ws_bytes < txb_perturb_workspace_bytes(nsamp, nchan, nalpha)) {
        set_error("txb_perturb_predict_f64: workspace too small");
        return TXB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    double* ws = reinterpret_cast<double*>(workspace);
    double* mm_part = ws;          // [1024][2]
    double* umm = ws + 2 * 1024;   // [2]
    double* part = umm + 8;
    int nmm = 4 * sm_count();
    if (nmm > 1024) nmm = 1024;
    if ((int64_t)nmm * 256 > nsamp) nmm = (int)((nsamp + 255) / 256);
    k_minmax_part<<<nmm, 256, 0, st>>>(u, nsamp, mm_part);
    TXB_CUDA_LAUNCH_CHECK("k_minmax_part");
    k_minmax_final<<<1, 32, 0, st>>>(mm_part, nmm, umm);
    TXB_CUDA_LAUNCH_CHECK("k_minmax_final");
    PtArgs a{};
    a.x = x; a.u = u; a.dalpha = dalpha; a.umm = umm; a.part = part;
    a.nsamp = nsamp; a.nchan = nchan; a.ldx = ldx; a.nalpha = nalpha;
    const int64_t nslab = pt_nslab(nsamp, nchan, nalpha);
    a.slab_len = (nsamp + nslab - 1) / nslab;
    const int64_t nslab_used = (nsamp + a.slab_len - 1) / a.slab_len;
    const unsigned atiles = (unsigned)((nalpha + PT_NA - 1) / PT_NA);
    {
        ProfScope prof("perturb", st);
        if (nchan == 1) {
            k_perturb<true, 1, 1><<<dim3((unsigned)nslab_used, atiles, 1), PT_THREADS, 0, st>>>(a);
        } else {
            const int qc = pt_qc(nchan);
            const unsigned ctiles = (unsigned)((nchan + qc * 32 - 1) / (qc * 32));
            const dim3 grid((unsigned)nslab_used, atiles, ctiles);
            if (qc == 1)      k_perturb<false, 1, 16><<<grid, PT_THREADS, 0, st>>>(a);
            else if (qc == 2) k_perturb<false, 2, 8><<<grid, PT_THREADS, 0, st>>>(a);
            else              k_perturb<false, PT_QC, 4><<<grid, PT_THREADS, 0, st>>>(a);
        }
    }
    TXB_CUDA_LAUNCH_CHECK("k_perturb");
    const int64_t nout = (int64_t)nalpha * nchan;
    k_perturb_final<<<(unsigned)((nout + 7) / 8), 256, 0, st>>>(part, nslab_used, nalpha, nchan, out);
    TXB_CUDA_LAUNCH_CHECK("k_perturb_final");
    return TXB_OK;
}

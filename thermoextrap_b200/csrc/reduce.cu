// K1 reduce_vals: single-pass fp64 shifted-power accumulation (HBM-bound).
//
// Replaces cmomy's per-sample Welford push (O(k^2)/sample, sequential) reached
// from /root/reference/src/thermoextrap/data.py:1632-1640 (and :485-489,
// :528-532, :1183-1203) with pivoted power sums
//     S0[j] = sum_s w (u-cu)^j        S1[c][j] = sum_s w (x_c-cx_c) (u-cu)^j
// (k+1 FMA per sample-channel), additive across threads / blocks / GPUs, and
// a final O(k^2) binomial re-centring per channel (finalize.cu).
#include <stdarg.h>
#include <string.h>

#include <stdlib.h>

#include "txb_common.cuh"
#include "reduce_tma.cuh"

namespace txb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---------------------------------------------------------------- profiling hook
static char g_prof_tag[64] = "";
static const int PROF_MAX = 1024;
static cudaEvent_t g_prof_ev[2 * PROF_MAX];
static int g_prof_made = 0, g_prof_used = 0;

ProfScope::ProfScope(const char* tag, cudaStream_t st) : st_(st), slot_(-1) {
    // a scope "family.member" matches the active tag "family" as well as "family.member"
    const size_t n = strlen(g_prof_tag);
    const bool hit = n != 0 && strncmp(tag, g_prof_tag, n) == 0 && (tag[n] == 0 || tag[n] == '.');
    if (!hit || g_prof_used >= PROF_MAX) return;
    while (g_prof_made <= g_prof_used) {
        if (cudaEventCreate(&g_prof_ev[2 * g_prof_made]) != cudaSuccess ||
            cudaEventCreate(&g_prof_ev[2 * g_prof_made + 1]) != cudaSuccess) {
            cudaGetLastError();
            return;
        }
        ++g_prof_made;
    }
    slot_ = g_prof_used++;
    cudaEventRecord(g_prof_ev[2 * slot_], st_);
}
ProfScope::~ProfScope() {
    if (slot_ >= 0) cudaEventRecord(g_prof_ev[2 * slot_ + 1], st_);
}

int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = TXB_NSM_FALLBACK;
    }
    return n;
}

// ---------------------------------------------------------------- pivot
// Column means over <= 256 strided rows.  block = 32 columns x 8 row phases,
// 32 independent loads per thread (one latency round).
__global__ void __launch_bounds__(256) k_colmean_sample(const double* __restrict__ a,
                                                        int64_t nsamp, int64_t ncol, int64_t ld,
                                                        double* __restrict__ out) {
    __shared__ double sm[8][33];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int64_t c = (int64_t)blockIdx.x * 32 + lane;
    const int nrows = nsamp < 256 ? (int)nsamp : 256;
    double acc = 0.0;
    if (c < ncol) {
        double v[32];
#pragma unroll
        for (int q = 0; q < 32; ++q) {
            const int i = wrp + 8 * q;
            const int64_t r = ((int64_t)i * nsamp) / nrows;
            v[q] = i < nrows ? a[r * ld + c] : 0.0;
        }
#pragma unroll
        for (int q = 0; q < 32; ++q) acc += v[q];
    }
    sm[wrp][lane] = acc;
    __syncthreads();
    if (wrp == 0 && c < ncol) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[k][lane];
        out[c] = t / (double)nrows;
    }
}

// ---------------------------------------------------------------- scalar x
// x[nsamp], u[nsamp] (or u only): one fused pass, 128-bit loads, 4 vectors in
// flight per thread.  Partials: ws0[block][K+1], ws1[block][K+1].
template <int K, bool HAS_X, bool HAS_W, int VEC>
__global__ void __launch_bounds__(256)
    k_reduce_scalar(const double* __restrict__ x, const double* __restrict__ u,
                    const double* __restrict__ w, int64_t nsamp, const double* __restrict__ pivot,
                    double* __restrict__ ws0, double* __restrict__ ws1) {
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

    const int64_t nvec = nsamp / VEC;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    constexpr int UNR = 4;
    if (VEC == 2) {
        // block-contiguous tiles (UNR * 256 double2 = 16 KB per array), grid-strided: neighbouring
        // blocks stream neighbouring DRAM pages
        constexpr int TILE = UNR * 256;
        const int64_t ntile = nvec / TILE;
        for (int64_t t = blockIdx.x; t < ntile; t += gridDim.x) {
            const int64_t base = t * TILE + threadIdx.x;
            double2 xv[UNR], uv[UNR], wv[UNR];
#pragma unroll
            for (int q = 0; q < UNR; ++q) {
                uv[q] = ldg_stream2(u + 2 * (base + q * 256));
                if (HAS_X) xv[q] = ldg_stream2(x + 2 * (base + q * 256));
                if (HAS_W) wv[q] = ldg_stream2(w + 2 * (base + q * 256));
            }
#pragma unroll
            for (int q = 0; q < UNR; ++q) {
                body(HAS_X ? xv[q].x : 0.0, uv[q].x, HAS_W ? wv[q].x : 1.0);
                body(HAS_X ? xv[q].y : 0.0, uv[q].y, HAS_W ? wv[q].y : 1.0);
            }
        }
        for (i = ntile * TILE + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
            const double2 uv = ldg_stream2(u + 2 * i);
            double2 xv = make_double2(0, 0), wv = make_double2(1, 1);
            if (HAS_X) xv = ldg_stream2(x + 2 * i);
            if (HAS_W) wv = ldg_stream2(w + 2 * i);
            body(xv.x, uv.x, wv.x);
            body(xv.y, uv.y, wv.y);
        }
        if ((nsamp & 1) && blockIdx.x == 0 && threadIdx.x == 0)
            body(HAS_X ? x[nsamp - 1] : 0.0, u[nsamp - 1], HAS_W ? w[nsamp - 1] : 1.0);
    } else {
        for (; i + (UNR - 1) * stride < nvec; i += UNR * stride) {
            double xv[UNR], uv[UNR], wv[UNR];
#pragma unroll
            for (int q = 0; q < UNR; ++q) {
                uv[q] = ldg_stream1(u + i + q * stride);
                xv[q] = HAS_X ? ldg_stream1(x + i + q * stride) : 0.0;
                wv[q] = HAS_W ? ldg_stream1(w + i + q * stride) : 1.0;
            }
#pragma unroll
            for (int q = 0; q < UNR; ++q) body(xv[q], uv[q], wv[q]);
        }
        for (; i < nvec; i += stride)
            body(HAS_X ? x[i] : 0.0, u[i], HAS_W ? w[i] : 1.0);
    }
    if (!HAS_W) s0[0] = (double)cnt;

    // block merge: warp shuffles, then one smem hop across the 8 warps
    __shared__ double sm[8][2 * (K + 1)];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j <= K; ++j) {
        const double a = warp_sum(s0[j]);
        const double b = HAS_X ? warp_sum(s1[j]) : 0.0;
        if (lane == 0) sm[wrp][j] = a, sm[wrp][K + 1 + j] = b;
    }
    __syncthreads();
    if (threadIdx.x < 2 * (K + 1)) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += sm[k][threadIdx.x];
        if (threadIdx.x <= K)
            ws0[(int64_t)blockIdx.x * (K + 1) + threadIdx.x] = t;
        else if (HAS_X)
            ws1[(int64_t)blockIdx.x * (K + 1) + threadIdx.x - (K + 1)] = t;
    }
}

// ---------------------------------------------------------------- wide x
// x[nsamp, ldx] with channels fastest.  Lanes run along channels (coalesced
// 128-bit loads, CPT channels per thread), the remaining threads of the block
// take different rows; grid = (channel tiles, sample slabs).
//   HAS_X && !UPC : S1[c][j]            (u shared: u[nsamp]; S0 comes from the
//                                        scalar kernel over u alone)
//   !HAS_X && UPC : S0[c][j]            (x_is_u wide, u[nsamp, ldu])
//   HAS_X && UPC  : both                (one energy per channel)
// Partials: ws[slab][chan][K+1].
template <int K, int CPT, bool HAS_X, bool UPC>
__global__ void __launch_bounds__(256)
    k_reduce_wide(const double* __restrict__ x, const double* __restrict__ u,
                  const double* __restrict__ w, int64_t nsamp, int64_t nchan, int64_t ldx,
                  int64_t ldu, const double* __restrict__ pivot, int lanes_c,
                  double* __restrict__ ws0, double* __restrict__ ws1) {
    static_assert(HAS_X || UPC, "u-only shared-u sums use the scalar kernel");
    const int tid = threadIdx.x;
    const int lane_c = tid & (lanes_c - 1);
    const int rowslot = tid / lanes_c;
    const int rows = 256 / lanes_c;
    // channel of element cc of this thread: pairs are interleaved across lanes so that one
    // LDG.128 instruction reads 32 B * lanes_c contiguous bytes per row (full sectors)
    const int64_t tile_base = (int64_t)blockIdx.x * lanes_c * CPT;
    auto chan_of = [&](int cc) -> int64_t {
        return CPT == 1 ? tile_base + lane_c
                        : tile_base + (int64_t)(cc >> 1) * 2 * lanes_c + 2 * lane_c + (cc & 1);
    };
    bool act[CPT];
    bool active = false;
#pragma unroll
    for (int cc = 0; cc < CPT; ++cc) act[cc] = chan_of(cc) < nchan, active = active || act[cc];
    const bool has_w = (w != nullptr);

    const int64_t rows_per_slab = (nsamp + gridDim.y - 1) / gridDim.y;
    const int64_t r_begin = (int64_t)blockIdx.y * rows_per_slab;
    const int64_t r_end = min(nsamp, r_begin + rows_per_slab);

    double cx[CPT], cu[CPT];
#pragma unroll
    for (int cc = 0; cc < CPT; ++cc) {
        cx[cc] = (HAS_X && act[cc]) ? pivot[chan_of(cc)] : 0.0;
        cu[cc] = UPC ? (act[cc] ? pivot[(HAS_X ? nchan : 0) + chan_of(cc)] : 0.0) : pivot[nchan];
    }
    double s0[UPC ? CPT : 1][K + 1], s1[HAS_X ? CPT : 1][K + 1];
#pragma unroll
    for (int j = 0; j <= K; ++j) {
#pragma unroll
        for (int cc = 0; cc < (UPC ? CPT : 1); ++cc) s0[cc][j] = 0.0;
#pragma unroll
        for (int cc = 0; cc < (HAS_X ? CPT : 1); ++cc) s1[cc][j] = 0.0;
    }

    // p = row base pointer
    auto load_vec = [&](const double* p, double (&v)[CPT]) {
        if (CPT == 1) {
            v[0] = ldg_stream1(p + chan_of(0));
        } else {
#pragma unroll
            for (int h = 0; h < CPT / 2; ++h) {
                double2 t = make_double2(0.0, 0.0);
                if (act[2 * h]) t = ldg_stream2(p + chan_of(2 * h));
                v[2 * h] = t.x;
                v[2 * h + 1] = t.y;
            }
        }
    };
    auto body = [&](const double (&xv)[CPT], const double (&uv)[CPT], double wv) {
        if (UPC) {
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) {
                const double du = uv[cc] - cu[cc];
                const double dx = HAS_X ? xv[cc] - cx[cc] : 0.0;
                double p = wv;
                s0[cc][0] += p;
                if (HAS_X) s1[cc][0] = fma(p, dx, s1[cc][0]);
#pragma unroll
                for (int j = 1; j <= K; ++j) {
                    p *= du;
                    s0[cc][j] += p;
                    if (HAS_X) s1[cc][j] = fma(p, dx, s1[cc][j]);
                }
            }
        } else {
            const double du = uv[0] - cu[0];
            double dx[CPT];
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) dx[cc] = xv[cc] - cx[cc];
            double p = wv;
#pragma unroll
            for (int cc = 0; cc < CPT; ++cc) s1[cc][0] = fma(p, dx[cc], s1[cc][0]);
#pragma unroll
            for (int j = 1; j <= K; ++j) {
                p *= du;
#pragma unroll
                for (int cc = 0; cc < CPT; ++cc) s1[cc][j] = fma(p, dx[cc], s1[cc][j]);
            }
        }
    };

    if (active) {
        constexpr int UNR = 4;
        int64_t r = r_begin + rowslot;
        for (; r + (int64_t)(UNR - 1) * rows < r_end; r += (int64_t)UNR * rows) {
            double xv[UNR][CPT], uv[UNR][CPT], wv[UNR];
#pragma unroll
            for (int q = 0; q < UNR; ++q) {
                const int64_t rr = r + (int64_t)q * rows;
                if (HAS_X) load_vec(x + rr * ldx, xv[q]);
                if (UPC) load_vec(u + rr * ldu, uv[q]);
                else uv[q][0] = __ldg(u + rr);
                wv[q] = has_w ? __ldg(w + rr) : 1.0;
            }
#pragma unroll
            for (int q = 0; q < UNR; ++q) body(xv[q], uv[q], wv[q]);
        }
        for (; r < r_end; r += rows) {
            double xv[CPT], uv[CPT];
            if (HAS_X) load_vec(x + r * ldx, xv);
            if (UPC) load_vec(u + r * ldu, uv);
            else uv[0] = __ldg(u + r);
            body(xv, uv, has_w ? __ldg(w + r) : 1.0);
        }
    }

    // block merge over the row slots that share a channel lane
    __shared__ double sm[256];
    auto block_merge = [&](double v, double* dst, bool ok) {
        sm[tid] = v;
        __syncthreads();
        if (tid < lanes_c && ok) {
            double t = 0.0;
            for (int rs = 0; rs < rows; ++rs) t += sm[rs * lanes_c + tid];
            *dst = t;
        }
        __syncthreads();
    };
#pragma unroll
    for (int cc = 0; cc < CPT; ++cc) {
        const int64_t base = ((int64_t)blockIdx.y * nchan + chan_of(cc)) * (K + 1);
#pragma unroll
        for (int j = 0; j <= K; ++j) {
            if (UPC) block_merge(s0[cc][j], ws0 + base + j, act[cc]);
            if (HAS_X) block_merge(s1[cc][j], ws1 + base + j, act[cc]);
        }
    }
}

// out[i] = sum_p ws[p][i]  (fixed association -> deterministic).
// block = out_l outputs x (256 / out_l) part phases; phases merged through smem.
__global__ void __launch_bounds__(256) k_sum_parts(const double* __restrict__ ws, int64_t nparts,
                                                   int64_t n, int out_l, double* __restrict__ out) {
    __shared__ double sm[256];
    const int tid = threadIdx.x;
    const int lo = tid & (out_l - 1);
    const int ph = tid / out_l;
    const int nph = 256 / out_l;
    const int64_t i = (int64_t)blockIdx.x * out_l + lo;
    double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
    if (i < n) {
        int64_t p = ph;
        for (; p + 3 * nph < nparts; p += 4 * nph) {
            t0 += ws[p * n + i];
            t1 += ws[(p + nph) * n + i];
            t2 += ws[(p + 2 * nph) * n + i];
            t3 += ws[(p + 3 * nph) * n + i];
        }
        for (; p < nparts; p += nph) t0 += ws[p * n + i];
    }
    sm[tid] = (t0 + t1) + (t2 + t3);
    __syncthreads();
    if (ph == 0 && i < n) {
        double t = 0.0;
        for (int k = 0; k < nph; ++k) t += sm[k * out_l + lo];
        out[i] = t;
    }
}

int launch_sum_parts(const double* ws, int64_t nparts, int64_t n, double* out, cudaStream_t st) {
    if (n <= 0) return TXB_OK;
    // few outputs / many parts -> more part phases per block
    int out_l = 32;
    while (out_l > 1 && (out_l / 2 >= n || (n <= 4096 && nparts >= 64 && out_l > 8))) out_l >>= 1;
    k_sum_parts<<<(unsigned)((n + out_l - 1) / out_l), 256, 0, st>>>(ws, nparts, n, out_l, out);
    TXB_CUDA_LAUNCH_CHECK("k_sum_parts");
    return TXB_OK;
}

// ---------------------------------------------------------------- planning
struct WidePlan {
    int cpt, lanes_c;
    int64_t nctiles, nslabs;
};

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static WidePlan plan_wide(int64_t nsamp, int64_t nchan, int cpt) {
    WidePlan p;
    p.cpt = cpt;
    const int64_t lanes_needed = (nchan + cpt - 1) / cpt;
    int lc = 1;
    while (lc < 32 && lc < lanes_needed) lc <<= 1;
    p.lanes_c = lc;
    p.nctiles = (lanes_needed + lc - 1) / lc;
    const int rows = 256 / lc;
    const int64_t target = (int64_t)sm_count() * 8;
    int64_t nslabs = (target + p.nctiles - 1) / p.nctiles;
    const int64_t max_slabs = (nsamp + (int64_t)rows * 4 - 1) / ((int64_t)rows * 4);
    if (nslabs > max_slabs) nslabs = max_slabs;
    if (nslabs < 1) nslabs = 1;
    p.nslabs = nslabs;
    return p;
}

static int64_t scalar_blocks(int64_t nsamp) {
    int64_t nb = (nsamp / 2 + 256 * 4 - 1) / (256 * 4);
    const int64_t cap = (int64_t)sm_count() * 8;
    if (nb > cap) nb = cap;
    if (nb < 1) nb = 1;
    return nb;
}

static int64_t max_wide_parts_x_chan(int64_t nsamp, int64_t nchan) {
    int64_t m = 0;
    for (int cpt = 1; cpt <= 4; cpt *= 2) {
        const WidePlan p = plan_wide(nsamp, nchan, cpt);
        if (p.nslabs * nchan > m) m = p.nslabs * nchan;
    }
    // the ring kernel (reduce_wide_tma.cu) runs one slab per resident CTA: 2 per SM at 256 threads
    // (<= 128 channels), 1 per SM at 512 threads
    const int64_t ring = (int64_t)sm_count() * (nchan <= 128 ? 2 : 1) * nchan;
    return m > ring ? m : ring;
}

// u-only / scalar-x sums -> S0[K+1] (and S1[K+1])
template <int K>
static int launch_scalar(const double* x, const double* u, const double* w, int64_t nsamp,
                         const double* pivot, double* S0, double* S1, double* ws, cudaStream_t st) {
    const bool vec = aligned16(u) && (!x || aligned16(x)) && (!w || aligned16(w));
    const int64_t nb_ring = vec ? scalar_tma_blocks(nsamp, scalar_blocks(nsamp), K) : 0;
    const int64_t nb = nb_ring ? nb_ring : scalar_blocks(nsamp);
    double* ws0 = ws;
    double* ws1 = ws + nb * (K + 1);
    if (nb_ring) {
        ProfScope prof("reduce", st);
        ProfScope ring("reduce_ring", st);
        const int rc = launch_scalar_tma(K, x, u, w, nsamp, pivot, nb, ws0, ws1, st);
        if (rc != TXB_OK) return rc;
    } else {
    ProfScope prof("reduce", st);
#define TXB_LAUNCH_SC(HX, HW, V) \
    k_reduce_scalar<K, HX, HW, V><<<(unsigned)nb, 256, 0, st>>>(x, u, w, nsamp, pivot, ws0, ws1)
    if (x) {
        if (w) { if (vec) TXB_LAUNCH_SC(true, true, 2); else TXB_LAUNCH_SC(true, true, 1); }
        else   { if (vec) TXB_LAUNCH_SC(true, false, 2); else TXB_LAUNCH_SC(true, false, 1); }
    } else {
        if (w) { if (vec) TXB_LAUNCH_SC(false, true, 2); else TXB_LAUNCH_SC(false, true, 1); }
        else   { if (vec) TXB_LAUNCH_SC(false, false, 2); else TXB_LAUNCH_SC(false, false, 1); }
    }
#undef TXB_LAUNCH_SC
    }
    TXB_CUDA_LAUNCH_CHECK("k_reduce_scalar");
    int rc = launch_sum_parts(ws0, nb, K + 1, S0, st);
    if (rc == TXB_OK && x) rc = launch_sum_parts(ws1, nb, K + 1, S1, st);
    return rc;
}

template <int K, bool HAS_X, bool UPC>
static int launch_wide(const double* x, const double* u, const double* w, int64_t nsamp,
                       int64_t nchan, int64_t ldx, int64_t ldu, const double* pivot, double* S0,
                       double* S1, double* ws, cudaStream_t st) {
    int cpt = 1;
    const bool xa = !HAS_X || (aligned16(x) && ldx % 2 == 0);
    const bool ua = !UPC || (aligned16(u) && ldu % 2 == 0);
    if (xa && ua && nchan % 2 == 0) cpt = 2;
    if (cpt == 2 && nchan % 4 == 0 && !(HAS_X && UPC)) cpt = 4;
    // copy-engine ring wherever it applies (reduce_wide_tma.cu), else register-staged loads
    const WideRingPlan t = plan_wide_ring(nsamp, nchan, ldx, ldu, K, HAS_X, UPC, xa && ua,
                                          max_wide_parts_x_chan(nsamp, nchan));
    const WidePlan p = plan_wide(nsamp, nchan, cpt);
    const int64_t nslabs = t.ok ? t.nslabs : p.nslabs;
    double* ws0 = ws;
    double* ws1 = ws + (UPC ? nslabs * nchan * (K + 1) : 0);
    const dim3 grid((unsigned)p.nctiles, (unsigned)p.nslabs);
    {
    ProfScope prof("reduce", st);
    if (t.ok) {
        ProfScope ring("reduce_ring", st);
        const int rc = launch_wide_ring(t, K, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
        if (rc != TXB_OK) return rc;
    } else if (cpt == 4)
        k_reduce_wide<K, (HAS_X && UPC) ? 2 : 4, HAS_X, UPC>
            <<<grid, 256, 0, st>>>(x, u, w, nsamp, nchan, ldx, ldu, pivot, p.lanes_c, ws0, ws1);
    else if (cpt == 2)
        k_reduce_wide<K, 2, HAS_X, UPC>
            <<<grid, 256, 0, st>>>(x, u, w, nsamp, nchan, ldx, ldu, pivot, p.lanes_c, ws0, ws1);
    else
        k_reduce_wide<K, 1, HAS_X, UPC>
            <<<grid, 256, 0, st>>>(x, u, w, nsamp, nchan, ldx, ldu, pivot, p.lanes_c, ws0, ws1);
    }
    TXB_CUDA_LAUNCH_CHECK("k_reduce_wide");
    int rc = TXB_OK;
    if (UPC) rc = launch_sum_parts(ws0, nslabs, nchan * (K + 1), S0, st);
    if (rc == TXB_OK && HAS_X) rc = launch_sum_parts(ws1, nslabs, nchan * (K + 1), S1, st);
    return rc;
}

}  // namespace txb

// ================================================================ C ABI
using namespace txb;

extern "C" int txb_version(void) { return 100; }
extern "C" const char* txb_last_error_string(void) { return txb::g_err; }
extern "C" int txb_max_order(void) { return TXB_MAXK; }

extern "C" int txb_pivot_f64(const double* x, const double* u, int64_t nsamp, int64_t nchan,
                             int64_t ldx, int64_t ldu, double* pivot, void* stream) {
    TXB_CHECK_ARG(u && pivot && nsamp > 0, "txb_pivot_f64: null pointer or nsamp <= 0");
    TXB_CHECK_ARG(!x || (nchan >= 1 && ldx >= nchan), "txb_pivot_f64: bad nchan/ldx");
    TXB_CHECK_ARG(ldu == 0 || ldu >= nchan, "txb_pivot_f64: bad ldu");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nx = x ? nchan : 0;
    if (x) k_colmean_sample<<<(unsigned)((nchan + 31) / 32), 256, 0, st>>>(x, nsamp, nchan, ldx, pivot);
    const int64_t nu = ldu == 0 ? 1 : nchan;
    k_colmean_sample<<<(unsigned)((nu + 31) / 32), 256, 0, st>>>(u, nsamp, nu, ldu == 0 ? 1 : ldu,
                                                                 pivot + nx);
    TXB_CUDA_LAUNCH_CHECK("k_colmean_sample");
    return TXB_OK;
}

static size_t reduce_partials_doubles(int64_t nsamp, int64_t nchan, int order) {
    const int64_t kp = order + 1;
    const int64_t sc = 2 * scalar_blocks(nsamp) * kp;
    const int64_t wd = 2 * max_wide_parts_x_chan(nsamp, nchan < 1 ? 1 : nchan) * kp;
    return (size_t)(sc > wd ? sc : wd);
}

extern "C" size_t txb_reduce_workspace_bytes(int64_t nsamp, int64_t nchan, int order) {
    if (nchan < 1) nchan = 1;
    const size_t sums = (size_t)(2 * nchan * (order + 2)) + (size_t)(2 * nchan + 2);
    return align_up((reduce_partials_doubles(nsamp, nchan, order + 1) + sums) * sizeof(double) + 256, 256);
}

extern "C" int txb_reduce_sums_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                                   int64_t nchan, int64_t ldx, int64_t ldu, int order,
                                   const double* pivot, double* S0, double* S1, void* workspace,
                                   size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && pivot && S0 && workspace, "txb_reduce_sums_f64: null pointer");
    TXB_CHECK_ARG(nsamp > 0, "txb_reduce_sums_f64: nsamp=%lld must be > 0", (long long)nsamp);
    TXB_CHECK_ARG(!x || S1, "txb_reduce_sums_f64: S1 is null but x is given");
    if (!x && ldu == 0) nchan = 1;
    TXB_CHECK_ARG(nchan >= 1 && (!x || ldx >= nchan) && (ldu == 0 || ldu >= nchan),
                  "txb_reduce_sums_f64: bad nchan=%lld ldx=%lld ldu=%lld", (long long)nchan,
                  (long long)ldx, (long long)ldu);
    if (ws_bytes < reduce_partials_doubles(nsamp, nchan, order) * sizeof(double)) {
        set_error("txb_reduce_sums_f64: workspace too small (%zu bytes)", ws_bytes);
        return TXB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    double* ws = (double*)workspace;
    const bool upc = (ldu != 0);
    int rc = TXB_OK;
    TXB_DISPATCH_ORDER(order, {
        if (!upc && (!x || nchan == 1)) {
            // scalar x (or u only): one fused pass
            rc = launch_scalar<K>(x, u, w, nsamp, pivot, S0, S1, ws, st);
        } else if (!upc) {
            // wide x, shared u: S0 from u alone (1/nchan of the traffic), S1 from the wide pass
            rc = launch_scalar<K>(nullptr, u, w, nsamp, pivot + nchan, S0, nullptr, ws, st);
            if (rc == TXB_OK)
                rc = launch_wide<K, true, false>(x, u, w, nsamp, nchan, ldx, 0, pivot, S0, S1, ws, st);
        } else if (!x) {
            rc = launch_wide<K, false, true>(nullptr, u, w, nsamp, nchan, 0, ldu, pivot, S0, S1, ws, st);
        } else {
            rc = launch_wide<K, true, true>(x, u, w, nsamp, nchan, ldx, ldu, pivot, S0, S1, ws, st);
        }
    });
    return rc;
}

extern "C" int txb_reduce_vals_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                                   int64_t nchan, int64_t ldx, int64_t ldu, int order, int mode,
                                   double* out, void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && out && workspace, "txb_reduce_vals_f64: null pointer");
    TXB_CHECK_ARG(order >= 1, "txb_reduce_vals_f64: order=%d must be >= 1", order);
    TXB_CHECK_ARG(mode == TXB_MODE_COMOMENTS || mode == TXB_MODE_X_IS_U, "bad mode %d", mode);
    TXB_CHECK_ARG(mode == TXB_MODE_COMOMENTS ? x != nullptr : x == nullptr,
                  "txb_reduce_vals_f64: mode COMOMENTS needs x, mode X_IS_U needs x == NULL");
    if (!x && ldu == 0) nchan = 1;
    TXB_CHECK_ARG(nchan >= 1, "txb_reduce_vals_f64: nchan must be >= 1");
    if (ws_bytes < txb_reduce_workspace_bytes(nsamp, nchan, order)) {
        set_error("txb_reduce_vals_f64: workspace too small: %zu < %zu", ws_bytes,
                  txb_reduce_workspace_bytes(nsamp, nchan, order));
        return TXB_ERR_WORKSPACE;
    }
    const int ks = mode == TXB_MODE_X_IS_U ? order + 1 : order;  // order of the sums
    const int64_t nu = ldu == 0 ? 1 : nchan;
    double* ws = (double*)workspace;
    double* pivot = ws;                 ws += 2 * nchan + 2;
    double* S0 = ws;                    ws += nchan * (order + 2);
    double* S1 = ws;                    ws += nchan * (order + 2);
    const size_t used = (size_t)((char*)ws - (char*)workspace);
    int rc = txb_pivot_f64(x, u, nsamp, nchan, ldx, ldu, pivot, stream);
    if (rc != TXB_OK) return rc;
    rc = txb_reduce_sums_f64(x, u, w, nsamp, nchan, ldx, ldu, ks, pivot, S0, S1, ws, ws_bytes - used,
                             stream);
    if (rc != TXB_OK) return rc;
    return txb_finalize_f64(S0, S1, pivot, 1, nchan, nu, order, mode, out, stream);
}

// ---------------------------------------------------------------- measurement hook (C ABI)
extern "C" int txb_profile_begin(const char* tag) {
    TXB_CHECK_ARG(tag && strlen(tag) < sizeof(txb::g_prof_tag), "txb_profile_begin: bad tag");
    strcpy(txb::g_prof_tag, tag);
    txb::g_prof_used = 0;
    return TXB_OK;
}

extern "C" int txb_profile_end(double* total_ms, int* nlaunch) {
    txb::g_prof_tag[0] = 0;
    double tot = 0.0;
    for (int i = 0; i < txb::g_prof_used; ++i) {
        float ms = 0.f;
        if (cudaEventSynchronize(txb::g_prof_ev[2 * i + 1]) != cudaSuccess ||
            cudaEventElapsedTime(&ms, txb::g_prof_ev[2 * i], txb::g_prof_ev[2 * i + 1]) != cudaSuccess) {
            txb::set_error("txb_profile_end: %s", cudaGetErrorString(cudaGetLastError()));
            return TXB_ERR_CUDA;
        }
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (nlaunch) *nlaunch = txb::g_prof_used;
    txb::g_prof_used = 0;
    return TXB_OK;
}

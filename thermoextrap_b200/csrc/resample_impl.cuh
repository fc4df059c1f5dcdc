// Host-side planning + launch of k_resample for one frequency element type.
#pragma once
#include "resample_mma_scalar.cuh"
#include "resample_mma2.cuh"
#include <cstdlib>

namespace txb {

int launch_sum_parts(const double* ws, int64_t nparts, int64_t n, double* out, cudaStream_t st);

struct RsPlan {
    int64_t nctile, nreptile, nslab, slab_len;
};

inline RsPlan rs_plan(int64_t nsamp, int64_t nchan, int64_t nrep, int rt, int st, int ct) {
    RsPlan p;
    p.nctile = ct > 0 ? (nchan + ct - 1) / ct : 1;
    p.nreptile = (nrep + rt - 1) / rt;
    const int64_t target = (int64_t)sm_count() * 2;
    int64_t nslab = target / (p.nctile * p.nreptile);
    if (nslab < 1) nslab = 1;
    int64_t slab_len = (nsamp + nslab - 1) / nslab;
    slab_len = (slab_len + st - 1) / st * st;
    p.slab_len = slab_len;
    p.nslab = (nsamp + slab_len - 1) / slab_len;
    return p;
}

// upper bound on the number of slabs any configuration will use
inline int64_t rs_max_slabs(int64_t nrep, int64_t nchan_tiles_min) {
    const int64_t target = (int64_t)sm_count() * 2;
    int64_t nreptile = (nrep + 127) / 128;
    int64_t n = target / (nreptile * (nchan_tiles_min < 1 ? 1 : nchan_tiles_min));
    return n < 1 ? 1 : n;
}

template <int K, int MODE, int NCG, typename FT>
int rs_launch(RsArgs a, double* S0, double* S1, double* ws, cudaStream_t st) {
    using C = RsCfg<K, MODE, NCG, FT>;
    const RsPlan p = rs_plan(a.nsamp, a.nchan, a.nrep, C::RT, C::ST, MODE == RS_WIDE ? C::CT : 0);
    const bool upc = a.ldu != 0;
    const int64_t nu = upc ? a.nchan : 1;
    const int64_t n0 = a.nrep * nu * (K + 1);
    const int64_t n1 = a.nrep * a.nchan * (K + 1);
    const bool want0 = (MODE == RS_SCALAR) || upc;
    const bool want1 = a.x != nullptr;
    a.ws0 = ws;
    a.ws1 = ws + (want0 ? p.nslab * n0 : 0);
    a.slab_len = p.slab_len;
    const int es = FreqElem<FT>::ES;
    a.aligned = ((reinterpret_cast<uintptr_t>(a.freq) & 15) == 0) && ((a.ldf * es) % 16 == 0);
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(k_resample<K, MODE, NCG, FT>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM) != cudaSuccess) {
            set_error("k_resample: cannot opt in to %zu bytes of shared memory", (size_t)C::SMEM);
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        attr_done = true;
    }
    const dim3 grid((unsigned)p.nctile, (unsigned)p.nreptile, (unsigned)p.nslab);
    {
        ProfScope prof("resample", st);
        k_resample<K, MODE, NCG, FT><<<grid, C::NT, C::SMEM, st>>>(a);
    }
    TXB_CUDA_LAUNCH_CHECK("k_resample");
    int rc = TXB_OK;
    if (want0) rc = launch_sum_parts(a.ws0, p.nslab, n0, S0, st);
    if (rc == TXB_OK && want1) rc = launch_sum_parts(a.ws1, p.nslab, n1, S1, st);
    return rc;
}

// wide x on the fp64 tensor cores.  TXB_RESAMPLE_MMA = 0: DFMA tile; 1: k_resample_mma (round 1);
// 2: k_resample_mma2 (operand generation overlapped with the DMMAs), 8 warps; 3: the same with 16
// warps -- measured at 1e8 x 64, order 6, uint8 table, 256 replicates (profiles/r02_exp_mma_forms.json):
// form 1 31.4, form 2 31.2, form 3 32.1 TFLOP/s; 4 (default): form 3 with uint8 entries converted through a
// shared-memory table instead of the 2^52 DADD: 32.6 TFLOP/s
extern int g_rs_mma_version;   // resample.cu; txb_set_resample_kernel overrides the environment
inline int rs_mma_version() {
    if (g_rs_mma_version < 0) {
        const char* e = getenv("TXB_RESAMPLE_MMA");
        g_rs_mma_version = (e && e[0] >= '0' && e[0] <= '4') ? e[0] - '0' : 4;
    }
    return g_rs_mma_version;
}
inline bool rs_use_mma() { return rs_mma_version() != 0; }

template <int K, typename FT>
int rs_launch_mma(RsArgs a, double* S0, double* S1, double* ws, cudaStream_t st) {
    using C = RmCfg<K, FT>;
    const RsPlan p = rs_plan(a.nsamp, a.nchan, a.nrep, C::RT, C::ST, C::CT);
    const bool upc = a.ldu != 0;
    const int64_t n = a.nrep * a.nchan * (K + 1);
    a.ws0 = ws;
    a.ws1 = ws;
    a.slab_len = p.slab_len;
    const int es = FreqElem<FT>::ES;
    a.aligned = ((reinterpret_cast<uintptr_t>(a.freq) & 15) == 0) && ((a.ldf * es) % 16 == 0);
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(k_resample_mma<K, FT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)C::SMEM) != cudaSuccess) {
            set_error("k_resample_mma: cannot opt in to %zu bytes of shared memory", (size_t)C::SMEM);
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        attr_done = true;
    }
    const dim3 grid((unsigned)p.nctile, (unsigned)p.nreptile, (unsigned)p.nslab);
    {
        ProfScope prof("resample.wide", st);
        k_resample_mma<K, FT><<<grid, C::NT, C::SMEM, st>>>(a);
    }
    TXB_CUDA_LAUNCH_CHECK("k_resample_mma");
    return launch_sum_parts(ws, p.nslab, n, upc ? S0 : S1, st);
}

// second form (operand generation overlapped with the DMMAs); WM = 4: 8 warps, WM = 8: 16 warps
template <int K, typename FT, int WM, bool LUT = false>
int rs_launch_mma2(RsArgs a, double* S0, double* S1, double* ws, cudaStream_t st) {
    using C = Rm2Cfg<K, FT, WM>;
    const RsPlan p = rs_plan(a.nsamp, a.nchan, a.nrep, C::RT, C::ST, C::CT);
    const bool upc = a.ldu != 0;
    const int64_t n = a.nrep * a.nchan * (K + 1);
    a.ws0 = ws;
    a.ws1 = ws;
    a.slab_len = p.slab_len;
    const int es = FreqElem<FT>::ES;
    a.aligned = ((reinterpret_cast<uintptr_t>(a.freq) & 15) == 0) && ((a.ldf * es) % 16 == 0);
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(k_resample_mma2<K, FT, WM, LUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)C::SMEM) != cudaSuccess) {
            set_error("k_resample_mma2: cannot opt in to %zu bytes of shared memory", (size_t)C::SMEM);
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        attr_done = true;
    }
    const dim3 grid((unsigned)p.nctile, (unsigned)p.nreptile, (unsigned)p.nslab);
    {
        ProfScope prof("resample.wide", st);
        k_resample_mma2<K, FT, WM, LUT><<<grid, C::NT, C::SMEM, st>>>(a);
    }
    TXB_CUDA_LAUNCH_CHECK("k_resample_mma2");
    return launch_sum_parts(ws, p.nslab, n, upc ? S0 : S1, st);
}

// wide dispatch for > 4 channels
template <int K, typename FT>
int rs_launch_wide_mma(const RsArgs& a, double* S0, double* S1, double* ws, cudaStream_t st) {
    const int v = rs_mma_version();
    if (v == 4) {   // 16 warps + shared-memory conversion table (uint8 tables; other dtypes: form 3)
        if constexpr (sizeof(FT) == 1) return rs_launch_mma2<K, FT, 8, true>(a, S0, S1, ws, st);
        else return rs_launch_mma2<K, FT, 8>(a, S0, S1, ws, st);
    }
    if (v == 3) return rs_launch_mma2<K, FT, 8>(a, S0, S1, ws, st);
    if (v == 2) return rs_launch_mma2<K, FT, 4>(a, S0, S1, ws, st);
    if (v == 1) return rs_launch_mma<K, FT>(a, S0, S1, ws, st);
    return rs_launch<K, RS_WIDE, 8, FT>(a, S0, S1, ws, st);
}

// scalar x on DMMA: always for 4- and 8-byte tables (the kernel then waits on the table, not on the FMA
// pipe), for narrow tables only where padding the columns to whole 8-wide fragments wastes <= 15 %
template <int K, bool HAS_X, typename FT>
inline bool rs_use_mma_scalar() {
    using C = RmsCfg<K, HAS_X, FT>;
    return rs_use_mma() && (C::ES >= 4 || (C::NCOLP - C::NCOL) * 100 <= 15 * C::NCOLP);
}

template <int K, bool HAS_X, typename FT>
int rs_launch_mma_scalar(RsArgs a, double* S0, double* S1, double* ws, cudaStream_t st) {
    using C = RmsCfg<K, HAS_X, FT>;
    const RsPlan p = rs_plan(a.nsamp, 1, a.nrep, C::RT, C::ST, 0);
    const int64_t n0 = a.nrep * (K + 1);
    a.ws0 = ws;
    a.ws1 = ws + p.nslab * n0;
    a.slab_len = p.slab_len;
    const int es = FreqElem<FT>::ES;
    a.aligned = ((reinterpret_cast<uintptr_t>(a.freq) & 15) == 0) && ((a.ldf * es) % 16 == 0);
    static bool attr_done = false;
    if (!attr_done) {
        if (cudaFuncSetAttribute(k_resample_mma_scalar<K, HAS_X, FT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)C::SMEM) != cudaSuccess) {
            set_error("k_resample_mma_scalar: cannot opt in to %zu bytes of shared memory", (size_t)C::SMEM);
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        attr_done = true;
    }
    const dim3 grid(1u, (unsigned)p.nreptile, (unsigned)p.nslab);
    {
        ProfScope prof("resample.scalar", st);
        k_resample_mma_scalar<K, HAS_X, FT><<<grid, C::NT, C::SMEM, st>>>(a);
    }
    TXB_CUDA_LAUNCH_CHECK("k_resample_mma_scalar");
    int rc = launch_sum_parts(a.ws0, p.nslab, n0, S0, st);
    if (rc == TXB_OK && HAS_X) rc = launch_sum_parts(a.ws1, p.nslab, n0, S1, st);
    return rc;
}

// scalar-mode dispatch: DMMA where it pays, else the DFMA tile
template <int K, typename FT>
int rs_launch_scalar(RsArgs a, double* S0, double* S1, double* ws, cudaStream_t st) {
    if (a.x != nullptr) {
        if (rs_use_mma_scalar<K, true, FT>()) return rs_launch_mma_scalar<K, true, FT>(a, S0, S1, ws, st);
    } else {
        if (rs_use_mma_scalar<K, false, FT>()) return rs_launch_mma_scalar<K, false, FT>(a, S0, S1, ws, st);
    }
    return rs_launch<K, RS_SCALAR, 1, FT>(a, S0, S1, ws, st);
}

template <typename FT>
int resample_sums_ft(const double* x, const double* u, const double* w, int64_t nsamp, int64_t nchan,
                     int64_t ldx, int64_t ldu, int order, const double* pivot, const void* freq,
                     int64_t nrep, int64_t ldf, double* S0, double* S1, double* ws, cudaStream_t st);

#define TXB_RESAMPLE_FT_IMPL(FT)                                                                     \
    template <>                                                                                      \
    int resample_sums_ft<FT>(const double* x, const double* u, const double* w, int64_t nsamp,       \
                             int64_t nchan, int64_t ldx, int64_t ldu, int order,                     \
                             const double* pivot, const void* freq, int64_t nrep, int64_t ldf,       \
                             double* S0, double* S1, double* ws, cudaStream_t st) {                  \
        RsArgs a;                                                                                    \
        a.x = x; a.u = u; a.w = w; a.freq = freq; a.pivot = pivot; a.ws0 = a.ws1 = nullptr;          \
        a.nsamp = nsamp; a.nchan = nchan; a.ldx = ldx; a.ldu = ldu; a.nrep = nrep; a.ldf = ldf;      \
        a.slab_len = 0; a.aligned = 0;                                                               \
        int rc = TXB_OK;                                                                             \
        TXB_DISPATCH_ORDER(order, {                                                                  \
            if (ldu == 0 && (!x || nchan == 1)) {                                                    \
                a.nchan = 1;                                                                         \
                rc = rs_launch_scalar<K, FT>(a, S0, S1, ws, st);                                     \
            } else if (ldu == 0) {                                                                   \
                RsArgs b = a;                                                                        \
                b.x = nullptr; b.nchan = 1; b.ldx = 0; b.pivot = pivot + nchan;                      \
                rc = rs_launch_scalar<K, FT>(b, S0, nullptr, ws, st);                                \
                if (rc == TXB_OK)                                                                    \
                    rc = nchan <= 4 ? rs_launch<K, RS_WIDE, 2, FT>(a, S0, S1, ws, st)                \
                                    : rs_launch_wide_mma<K, FT>(a, S0, S1, ws, st);                  \
            } else {                                                                                 \
                rc = nchan <= 4 ? rs_launch<K, RS_WIDE, 2, FT>(a, S0, S1, ws, st)                    \
                                : rs_launch_wide_mma<K, FT>(a, S0, S1, ws, st);                      \
            }                                                                                        \
        });                                                                                          \
        return rc;                                                                                   \
    }

}  // namespace txb

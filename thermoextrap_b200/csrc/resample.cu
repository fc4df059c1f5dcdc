// C ABI for K2 (resample_vals) and K6 (indices -> freq).
#include "resample_impl.cuh"

namespace txb {

int g_rs_mma_version = -1;

// ---------------------------------------------------------------- K6
// freq[r, idx[r, j]] += 1.  Integer work, bit-exact.  Narrow tables are
// incremented through a 32-bit atomicAdd on the containing word; the old
// value tells us when a lane was already saturated (status |= 2).
template <typename FT>
__global__ void __launch_bounds__(256)
    k_indices_to_freq(const int64_t* __restrict__ idx, int64_t nrep, int64_t nsamp, int64_t ndat,
                      FT* __restrict__ freq, int64_t ldf, int64_t slab_len, int* __restrict__ status) {
    const int64_t r = blockIdx.y;
    const int64_t j0 = (int64_t)blockIdx.x * slab_len;
    const int64_t j1 = min(nsamp, j0 + slab_len);
    const int64_t* row = idx + r * nsamp;
    int bad = 0;
    auto bump = [&](int64_t i) {
        if ((uint64_t)i >= (uint64_t)ndat) {
            bad |= 1;
            return;
        }
        const int64_t off = r * ldf + i;
        if (sizeof(FT) == 8) {
            atomicAdd(reinterpret_cast<unsigned long long*>(freq) + off, 1ull);
        } else if (sizeof(FT) == 4) {
            const unsigned old = atomicAdd(reinterpret_cast<unsigned*>(freq) + off, 1u);
            if (old == 0x7fffffffu) bad |= 2;
        } else {
            constexpr int BITS = 8 * (int)sizeof(FT);
            constexpr unsigned MASK = (1u << BITS) - 1u;
            const int64_t boff = off * (int64_t)sizeof(FT);
            unsigned* word = reinterpret_cast<unsigned*>(reinterpret_cast<unsigned char*>(freq) + (boff & ~3ll));
            const int sh = (int)(boff & 3) * 8;
            const unsigned old = atomicAdd(word, 1u << sh);
            if (((old >> sh) & MASK) == MASK) bad |= 2;
        }
    };
    constexpr int UNR = 8;
    int64_t j = j0 + threadIdx.x;
    for (; j + (UNR - 1) * 256 < j1; j += UNR * 256) {
        int64_t ii[UNR];
#pragma unroll
        for (int t = 0; t < UNR; ++t) ii[t] = __ldcs(row + j + t * 256);
#pragma unroll
        for (int t = 0; t < UNR; ++t) bump(ii[t]);
    }
    for (; j < j1; j += 256) bump(row[j]);
    if (bad && status) atomicOr(status, bad);
}

}  // namespace txb

using namespace txb;

extern "C" int txb_indices_to_freq(const int64_t* indices, int64_t nrep, int64_t nsamp, int64_t ndat,
                                   void* freq, int freq_dtype, int64_t ldf, int* status,
                                   void* stream) {
    TXB_CHECK_ARG(indices && freq, "txb_indices_to_freq: null pointer");
    TXB_CHECK_ARG(nrep >= 0 && nsamp >= 0 && ndat > 0 && ldf >= ndat,
                  "txb_indices_to_freq: bad sizes nrep=%lld nsamp=%lld ndat=%lld ldf=%lld",
                  (long long)nrep, (long long)nsamp, (long long)ndat, (long long)ldf);
    static const int es_of[4] = {8, 4, 2, 1};
    TXB_CHECK_ARG(freq_dtype >= 0 && freq_dtype <= 3, "txb_indices_to_freq: bad freq_dtype %d", freq_dtype);
    const int es = es_of[freq_dtype];
    TXB_CHECK_ARG(es >= 4 || ((reinterpret_cast<uintptr_t>(freq) & 3) == 0),
                  "txb_indices_to_freq: narrow table must be 4-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    // rows of a narrow table share 32-bit words only if ldf*es % 4 != 0; the word
    // atomics stay correct either way as long as the allocation is padded to 4 bytes.
    if (cudaMemsetAsync(freq, 0, (size_t)nrep * ldf * es, st) != cudaSuccess) {
        set_error("txb_indices_to_freq: cudaMemsetAsync failed");
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    if (status && cudaMemsetAsync(status, 0, sizeof(int), st) != cudaSuccess) {
        set_error("txb_indices_to_freq: cudaMemsetAsync(status) failed");
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    if (nrep * nsamp == 0) return TXB_OK;
    TXB_CHECK_ARG(nrep <= 65535, "txb_indices_to_freq: nrep must be <= 65535 per call");
    // many slabs per row: rows are processed (almost) one at a time, so the row being
    // incremented stays L2-resident
    int64_t ns = (int64_t)sm_count() * 4;
    const int64_t maxs = (nsamp + 2047) / 2048;
    if (ns > maxs) ns = maxs;
    if (ns < 1) ns = 1;
    const int64_t slab_len = (nsamp + ns - 1) / ns;
    const dim3 grid((unsigned)((nsamp + slab_len - 1) / slab_len), (unsigned)nrep);
    switch (freq_dtype) {
        case TXB_I64: k_indices_to_freq<int64_t><<<grid, 256, 0, st>>>(indices, nrep, nsamp, ndat, (int64_t*)freq, ldf, slab_len, status); break;
        case TXB_I32: k_indices_to_freq<int32_t><<<grid, 256, 0, st>>>(indices, nrep, nsamp, ndat, (int32_t*)freq, ldf, slab_len, status); break;
        case TXB_U16: k_indices_to_freq<uint16_t><<<grid, 256, 0, st>>>(indices, nrep, nsamp, ndat, (uint16_t*)freq, ldf, slab_len, status); break;
        default:      k_indices_to_freq<uint8_t><<<grid, 256, 0, st>>>(indices, nrep, nsamp, ndat, (uint8_t*)freq, ldf, slab_len, status); break;
    }
    TXB_CUDA_LAUNCH_CHECK("k_indices_to_freq");
    return TXB_OK;
}

extern "C" int txb_set_resample_kernel(int version) {
    const int prev = rs_mma_version();
    if (version >= 0 && version <= 4) g_rs_mma_version = version;
    return prev;
}

static size_t rs_partials_doubles(int64_t nsamp, int64_t nchan, int order, int64_t nrep) {
    (void)nsamp;
    const int64_t kp = order + 1;
    const int64_t ch = nchan < 1 ? 1 : nchan;
    // scalar pass: nslab * nrep * 2 (K+1);  wide pass: nslab * nrep * nchan * (K+1)
    const int64_t sc = rs_max_slabs(nrep, 1) * nrep * 2 * kp;
    const int64_t wd = rs_max_slabs(nrep, (ch + 15) / 16) * nrep * ch * kp;
    return (size_t)(sc > wd ? sc : wd);
}

extern "C" size_t txb_resample_workspace_bytes(int64_t nsamp, int64_t nchan, int order, int64_t nrep) {
    const int64_t ch = nchan < 1 ? 1 : nchan;
    const size_t sums = (size_t)(2 * nrep * ch * (order + 2)) + (size_t)(2 * ch + 2);
    return align_up((rs_partials_doubles(nsamp, nchan, order + 1, nrep) + sums) * sizeof(double) + 256, 256);
}

extern "C" int txb_resample_sums_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                                     int64_t nchan, int64_t ldx, int64_t ldu, int order,
                                     const double* pivot, const void* freq, int freq_dtype,
                                     int64_t nrep, int64_t ldf, double* S0, double* S1,
                                     void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && pivot && freq && S0 && workspace, "txb_resample_sums_f64: null pointer");
    TXB_CHECK_ARG(nsamp > 0 && nrep > 0, "txb_resample_sums_f64: nsamp and nrep must be > 0");
    TXB_CHECK_ARG(nsamp < (1ll << 31), "txb_resample_sums_f64: nsamp must be < 2^31");
    TXB_CHECK_ARG(ldf >= nsamp, "txb_resample_sums_f64: ldf=%lld < nsamp=%lld", (long long)ldf, (long long)nsamp);
    TXB_CHECK_ARG(!x || S1, "txb_resample_sums_f64: S1 is null but x is given");
    if (!x && ldu == 0) nchan = 1;
    TXB_CHECK_ARG(nchan >= 1 && (!x || ldx >= nchan) && (ldu == 0 || ldu >= nchan),
                  "txb_resample_sums_f64: bad nchan/ldx/ldu");
    if (x && ldu != 0) {
        set_error("txb_resample_sums_f64: per-channel u together with x is not supported");
        return TXB_ERR_UNSUPPORTED;
    }
    if (ws_bytes < rs_partials_doubles(nsamp, nchan, order, nrep) * sizeof(double)) {
        set_error("txb_resample_sums_f64: workspace too small (%zu bytes)", ws_bytes);
        return TXB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    double* ws = (double*)workspace;
    switch (freq_dtype) {
        case TXB_I64: return resample_sums_ft<int64_t>(x, u, w, nsamp, nchan, ldx, ldu, order, pivot, freq, nrep, ldf, S0, S1, ws, st);
        case TXB_I32: return resample_sums_ft<int32_t>(x, u, w, nsamp, nchan, ldx, ldu, order, pivot, freq, nrep, ldf, S0, S1, ws, st);
        case TXB_U16: return resample_sums_ft<uint16_t>(x, u, w, nsamp, nchan, ldx, ldu, order, pivot, freq, nrep, ldf, S0, S1, ws, st);
        case TXB_U8:  return resample_sums_ft<uint8_t>(x, u, w, nsamp, nchan, ldx, ldu, order, pivot, freq, nrep, ldf, S0, S1, ws, st);
        default:
            set_error("txb_resample_sums_f64: bad freq_dtype %d", freq_dtype);
            return TXB_ERR_BAD_ARG;
    }
}

extern "C" int txb_resample_vals_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                                     int64_t nchan, int64_t ldx, int64_t ldu, int order, int mode,
                                     const void* freq, int freq_dtype, int64_t nrep, int64_t ldf,
                                     double* out, void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && out && workspace && freq, "txb_resample_vals_f64: null pointer");
    TXB_CHECK_ARG(order >= 1, "txb_resample_vals_f64: order=%d must be >= 1", order);
    TXB_CHECK_ARG(mode == TXB_MODE_COMOMENTS || mode == TXB_MODE_X_IS_U, "bad mode %d", mode);
    TXB_CHECK_ARG(mode == TXB_MODE_COMOMENTS ? x != nullptr : x == nullptr,
                  "txb_resample_vals_f64: mode COMOMENTS needs x, mode X_IS_U needs x == NULL");
    if (!x && ldu == 0) nchan = 1;
    TXB_CHECK_ARG(nchan >= 1 && nrep >= 1, "txb_resample_vals_f64: nchan, nrep must be >= 1");
    if (ws_bytes < txb_resample_workspace_bytes(nsamp, nchan, order, nrep)) {
        set_error("txb_resample_vals_f64: workspace too small: %zu < %zu", ws_bytes,
                  txb_resample_workspace_bytes(nsamp, nchan, order, nrep));
        return TXB_ERR_WORKSPACE;
    }
    const int ks = mode == TXB_MODE_X_IS_U ? order + 1 : order;
    const int64_t nu = ldu == 0 ? 1 : nchan;
    double* ws = (double*)workspace;
    double* pivot = ws;  ws += 2 * nchan + 2;
    double* S0 = ws;     ws += nrep * nchan * (order + 2);
    double* S1 = ws;     ws += nrep * nchan * (order + 2);
    const size_t used = (size_t)((char*)ws - (char*)workspace);
    int rc = txb_pivot_f64(x, u, nsamp, nchan, ldx, ldu, pivot, stream);
    if (rc != TXB_OK) return rc;
    rc = txb_resample_sums_f64(x, u, w, nsamp, nchan, ldx, ldu, ks, pivot, freq, freq_dtype, nrep, ldf,
                               S0, S1, ws, ws_bytes - used, stream);
    if (rc != TXB_OK) return rc;
    return txb_finalize_f64(S0, S1, pivot, nrep, nchan, nu, order, mode, out, stream);
}

// K2 resample_vals as an fp64 contraction with fused operand generation.
//
//   S[r, col] = sum_s F[r, s] * P[s, col]
//
// replaces cmomy.resample_vals' nrep*nsamp weighted pushes
// (/root/reference/src/thermoextrap/data.py:1354-1366, 1803-1810).
//   F  the sampler's frequency table, consumed in its own integer dtype:
//      cp.async (LDGSTS) 16-byte chunks -> smem, 3-stage pipeline; converted to
//      fp64 in registers with the 2^52 magic-number trick (1 DADD).
//   P  shifted-power products, never materialised in HBM: regenerated per
//      sample tile into smem from x, u, w.
//        scalar mode   cols = [ w du^j | w dx du^j ],  j = 0..K
//        wide mode     cols (c, j) = w dx_c du^j   (or w du_c^j for x_is_u)
//   micro-kernel: lane <-> replicate (TR reps per lane), warp <-> (column
//      group, replicate half, sample phase); each thread keeps a TR x TC
//      accumulator tile in registers, A operands come from conflict-free
//      LDS.128, B operands are warp-broadcast LDS.128.  8 warps per SM; the
//      16-warp split (NT = 512, TR / 2, RH = 2) measured 10 % slower.  A DFMA
//      tile with three register operands cannot pass ~27 TFLOP/s on a B200
//      (peak.cu); the wide-x case therefore runs on DMMA (resample_mma.cuh).
// grid = (column tiles, replicate tiles, sample slabs); slab partials are
// summed in a fixed order by k_sum_parts (deterministic, no atomics).
#pragma once
#include "txb_common.cuh"

namespace txb {

template <typename FT> struct FreqElem;
template <> struct FreqElem<int64_t>  { static constexpr int ES = 8; static constexpr int ID = TXB_I64; };
template <> struct FreqElem<int32_t>  { static constexpr int ES = 4; static constexpr int ID = TXB_I32; };
template <> struct FreqElem<uint16_t> { static constexpr int ES = 2; static constexpr int ID = TXB_U16; };
template <> struct FreqElem<uint8_t>  { static constexpr int ES = 1; static constexpr int ID = TXB_U8; };

__device__ __forceinline__ double u32_to_f64(uint32_t v) {
    // (double)v for v < 2^32 with one DADD: 2^52 + v is exact.
    return __hiloint2double(0x43300000, (int)v) - 4503599627370496.0;
}

// v-th element of a 16-byte chunk, as double (counts are < 2^31)
template <int ES>
__device__ __forceinline__ double chunk_elem(const uint4& q, int v) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    if (ES == 8) return u32_to_f64(w[2 * v]);
    if (ES == 4) return u32_to_f64(w[v]);
    if (ES == 2) return u32_to_f64((w[v >> 1] >> (16 * (v & 1))) & 0xffffu);
    return u32_to_f64((w[v >> 2] >> (8 * (v & 3))) & 0xffu);
}

__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem, int src_bytes) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

enum { RS_SCALAR = 0, RS_WIDE = 1 };

// Compile-time tile configuration.
template <int K, int MODE, int NCG_, typename FT>
struct RsCfg {
    static constexpr int ES = FreqElem<FT>::ES;
    static constexpr int VS = 16 / ES;                  // samples per 16-byte chunk
    static constexpr int TC = 2 * (K + 1);              // columns per thread
    static constexpr int NT = 256;                      // threads per CTA
    static constexpr int TR = (TC <= 18) ? 4 : 2;       // replicates per lane
    static constexpr int RH = 1;                        // warps along replicates (2 with NT = 512, TR / 2)
    static constexpr int RT = 32 * TR * RH;             // replicate tile
    static constexpr int NCG = NCG_;                    // column groups (warps along columns)
    static constexpr int KW = NT / 32 / (NCG * RH);     // sample phases (warps along samples)
    static constexpr int NCOL = NCG * TC;               // columns per CTA
    static constexpr int CT = (MODE == RS_WIDE) ? 2 * NCG : 1;   // channels per CTA
    static constexpr int ST = (MODE == RS_WIDE) ? 32 : 256 / ES; // samples per stage
    static constexpr int NG = ST / VS;                  // 16-byte groups per row per stage
    static constexpr int PITCH = ST * ES + 16;          // bytes; PITCH/16 odd -> LDS.128 conflict-free
    static constexpr int NSTAGE = 3;
    static constexpr int A_BYTES = RT * PITCH;
    static constexpr int P_DBL = ST * NCOL;             // doubles per P buffer
    static constexpr int RED_PITCH = NCOL + 1;          // doubles (odd -> conflict-free)
    static constexpr size_t SMEM_PIPE = (size_t)NSTAGE * A_BYTES + 2 * (size_t)P_DBL * 8;
    static constexpr size_t SMEM_RED = (size_t)RT * RED_PITCH * 8;
    static constexpr size_t SMEM = SMEM_PIPE > SMEM_RED ? SMEM_PIPE : SMEM_RED;
    static_assert((PITCH / 16) % 2 == 1, "row pitch must be an odd number of 16B units");
};

struct RsArgs {
    const double* x;      // [nsamp, ldx] or null
    const double* u;      // [nsamp] (ldu == 0) or [nsamp, ldu]
    const double* w;      // [nsamp] or null
    const void* freq;     // [nrep, ldf]
    const double* pivot;  // cx[nchan] (if x), cu[nu]
    double* ws0;          // [nslab][nrep][nu][K+1]     (scalar mode; wide x_is_u mode)
    double* ws1;          // [nslab][nrep][nchan][K+1]
    int64_t nsamp, nchan, ldx, ldu, nrep, ldf;
    int64_t slab_len;     // samples per slab (multiple of ST)
    int aligned;          // rows of freq are 16-byte aligned
};

template <int K, int MODE, int NCG, typename FT>
__global__ void __launch_bounds__(256, 1) k_resample(const RsArgs a) {
    using C = RsCfg<K, MODE, NCG, FT>;
    constexpr int ES = C::ES, VS = C::VS, TC = C::TC, TR = C::TR, RT = C::RT, KW = C::KW;
    constexpr int NCOL = C::NCOL, CT = C::CT, ST = C::ST, NG = C::NG, PITCH = C::PITCH;
    constexpr int NSTAGE = C::NSTAGE, NT = C::NT, RH = C::RH;

    extern __shared__ __align__(16) unsigned char smem[];
    unsigned char* sA = smem;
    double* sP = reinterpret_cast<double*>(smem + (size_t)NSTAGE * C::A_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int cg = wrp % NCG;          // column group
    const int rh = (wrp / NCG) % RH;   // replicate half
    const int kw = wrp / (NCG * RH);   // sample phase
    const int64_t c_base = (int64_t)blockIdx.x * CT;   // first channel of this CTA (wide)
    const int64_t r_base = (int64_t)blockIdx.y * RT;
    const int64_t s_begin = (int64_t)blockIdx.z * a.slab_len;
    const int64_t s_end = min(a.nsamp, s_begin + a.slab_len);
    const int nstage = (int)((s_end - s_begin + ST - 1) / ST);
    const bool has_x = a.x != nullptr, has_w = a.w != nullptr;
    const bool upc = a.ldu != 0;  // wide x_is_u: per-channel u, no x

    // ---- operand A: cp.async the F tile of stage `it` into buffer it % NSTAGE
    const unsigned char* fbase = reinterpret_cast<const unsigned char*>(a.freq);
    auto issue_A = [&](int it) {
        if (it < nstage) {
            unsigned char* dst = sA + (size_t)(it % NSTAGE) * C::A_BYTES;
            const int64_t s0 = s_begin + (int64_t)it * ST;
            if (a.aligned) {
                for (int e = tid; e < RT * NG; e += NT) {
                    const int r = e / NG, g = e % NG;
                    int64_t rr = r_base + r;
                    if (rr >= a.nrep) rr = a.nrep - 1;
                    const int64_t s = s0 + (int64_t)g * VS;
                    int64_t left = (a.nsamp - s) * ES;
                    const int nb = left >= 16 ? 16 : (left > 0 ? (int)left : 0);
                    const unsigned char* src = fbase + (rr * a.ldf + (nb > 0 ? s : 0)) * ES;
                    cp_async_16(dst + (size_t)r * PITCH + g * 16, src, nb);
                }
            } else {
                const FT* f = reinterpret_cast<const FT*>(a.freq);
                for (int e = tid; e < RT * ST; e += NT) {
                    const int r = e / ST, sl = e % ST;
                    int64_t rr = r_base + r;
                    if (rr >= a.nrep) rr = a.nrep - 1;
                    const int64_t s = s0 + sl;
                    FT v = s < a.nsamp ? f[rr * a.ldf + s] : (FT)0;
                    *reinterpret_cast<FT*>(dst + (size_t)r * PITCH + sl * ES) = v;
                }
            }
        }
        cp_async_commit();
    };

    // ---- operand B: registers prefetched one stage ahead, expanded into sP
    constexpr int NPRE = (MODE == RS_WIDE) ? (ST * CT + NT - 1) / NT : 1;
    double px[NPRE], pu[NPRE], pwt[NPRE];
    auto prefetch_P = [&](int it) {
        const int64_t s0 = s_begin + (int64_t)it * ST;
#pragma unroll
        for (int q = 0; q < NPRE; ++q) {
            px[q] = 0.0, pu[q] = 0.0, pwt[q] = 0.0;
            const int e = tid + NT * q;
            if (MODE == RS_SCALAR) {
                const int64_t s = s0 + e;
                if (e < ST && s < s_end && it < nstage) {
                    pu[q] = __ldg(a.u + s);
                    px[q] = has_x ? __ldg(a.x + s * (a.ldx > 0 ? a.ldx : 1)) : 0.0;
                    pwt[q] = has_w ? __ldg(a.w + s) : 1.0;
                }
            } else {
                const int sl = e / CT, cl = e % CT;
                const int64_t s = s0 + sl, c = c_base + cl;
                if (e < ST * CT && s < s_end && c < a.nchan && it < nstage) {
                    pu[q] = upc ? __ldg(a.u + s * a.ldu + c) : __ldg(a.u + s);
                    px[q] = has_x ? __ldg(a.x + s * a.ldx + c) : 0.0;
                    pwt[q] = has_w ? __ldg(a.w + s) : 1.0;
                }
            }
        }
    };
    auto write_P = [&](int it) {
        double* dst = sP + (size_t)(it & 1) * C::P_DBL;
#pragma unroll
        for (int q = 0; q < NPRE; ++q) {
            const int e = tid + NT * q;
            if (MODE == RS_SCALAR) {
                if (e < ST) {
                    const double du = pu[q] - a.pivot[has_x ? 1 : 0];
                    const double dx = has_x ? px[q] - a.pivot[0] : 0.0;
                    double p = pwt[q];  // 0 for out-of-range samples
                    double* row = dst + (size_t)e * NCOL;
#pragma unroll
                    for (int j = 0; j <= K; ++j) {
                        row[j] = p;
                        row[K + 1 + j] = p * dx;
                        p *= du;
                    }
                }
            } else {
                if (e < ST * CT) {
                    const int sl = e / CT, cl = e % CT;
                    const int64_t c = c_base + cl;
                    const bool ok = c < a.nchan;
                    double du, lead;
                    if (upc) {   // cols (c, j) = w du_c^j
                        du = pu[q] - (ok ? a.pivot[c] : 0.0);
                        lead = pwt[q];
                    } else {     // cols (c, j) = w dx_c du^j
                        du = pu[q] - a.pivot[a.nchan];
                        lead = pwt[q] * (px[q] - (ok ? a.pivot[c] : 0.0));
                    }
                    double p = lead;
                    double* row = dst + (size_t)sl * NCOL + cl * (K + 1);
#pragma unroll
                    for (int j = 0; j <= K; ++j) {
                        row[j] = p;
                        p *= du;
                    }
                }
            }
        }
    };

    double acc[TR][TC];
#pragma unroll
    for (int i = 0; i < TR; ++i)
#pragma unroll
        for (int c = 0; c < TC; ++c) acc[i][c] = 0.0;

    // ---- pipeline
#pragma unroll
    for (int p = 0; p < NSTAGE - 1; ++p) issue_A(p);
    prefetch_P(0);

    for (int it = 0; it < nstage; ++it) {
        write_P(it);
        prefetch_P(it + 1);
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();
        issue_A(it + NSTAGE - 1);

        const unsigned char* bufA = sA + (size_t)(it % NSTAGE) * C::A_BYTES;
        const double* bufP = sP + (size_t)(it & 1) * C::P_DBL + cg * TC;
#pragma unroll 1
        for (int g = kw; g < NG; g += KW) {
            uint4 raw[TR];
#pragma unroll
            for (int i = 0; i < TR; ++i)
                raw[i] = *reinterpret_cast<const uint4*>(bufA + (size_t)((rh * TR + i) * 32 + lane) * PITCH + g * 16);
#pragma unroll
            for (int v = 0; v < VS; ++v) {
                const double* prow = bufP + (size_t)(g * VS + v) * NCOL;
                double pv[TC];
#pragma unroll
                for (int c = 0; c < TC; c += 2) {
                    const double2 t = *reinterpret_cast<const double2*>(prow + c);
                    pv[c] = t.x, pv[c + 1] = t.y;
                }
#pragma unroll
                for (int i = 0; i < TR; ++i) {
                    const double f = chunk_elem<ES>(raw[i], v);
#pragma unroll
                    for (int c = 0; c < TC; ++c) acc[i][c] = fma(f, pv[c], acc[i][c]);
                }
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();

    // ---- merge the KW sample phases through smem, then write the slab partial
    double* red = reinterpret_cast<double*>(smem);
    for (int ph = 0; ph < KW; ++ph) {
        if (kw == ph) {
#pragma unroll
            for (int i = 0; i < TR; ++i)
#pragma unroll
                for (int c = 0; c < TC; ++c) {
                    double* d = red + (size_t)((rh * TR + i) * 32 + lane) * C::RED_PITCH + cg * TC + c;
                    *d = (ph == 0) ? acc[i][c] : *d + acc[i][c];
                }
        }
        __syncthreads();
    }
    const int64_t slab = blockIdx.z;
    for (int e = tid; e < RT * NCOL; e += NT) {
        const int r = e / NCOL, col = e % NCOL;
        const int64_t rr = r_base + r;
        if (rr >= a.nrep) continue;
        const double v = red[(size_t)r * C::RED_PITCH + col];
        if (MODE == RS_SCALAR) {
            if (col <= K)
                a.ws0[(slab * a.nrep + rr) * (K + 1) + col] = v;
            else if (has_x)
                a.ws1[(slab * a.nrep + rr) * (K + 1) + col - (K + 1)] = v;
        } else {
            const int cl = col / (K + 1), j = col % (K + 1);
            const int64_t c = c_base + cl;
            if (c < a.nchan) {
                double* dstp = upc ? a.ws0 : a.ws1;
                dstp[((slab * a.nrep + rr) * a.nchan + c) * (K + 1) + j] = v;
            }
        }
    }
}

}  // namespace txb

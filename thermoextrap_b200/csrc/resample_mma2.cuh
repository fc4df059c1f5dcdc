// K2 wide-x contraction on the fp64 tensor cores, second form (round 2).
//
// Same contraction and fragment layout as k_resample_mma (resample_mma.cuh); three changes, each aimed at a
// bubble the first form's ncu capture showed (DMMA pipe 83 % active, one block-wide barrier per stage with
// the P tile written between the barrier and the first DMMA):
//   1. the P tile of stage it + 1 is written WHILE the DMMAs of stage it issue (one 16-sample piece after
//      every 16-sample group of DMMAs, into the other P buffer), so the tensor pipe no longer drains while
//      all warps regenerate operands; one barrier per stage remains and only orders buffer reuse;
//   2. the samples of a 16-sample group are taken in the order (4 fk + i) instead of (4 i + fk): summation
//      order over K is free, and a lane then needs FOUR CONSECUTIVE table entries per group -- one 32-bit
//      shared-memory load per fragment row instead of four byte loads (P pitch is odd so the B-fragment
//      loads of this order stay conflict-free);
//   3. WM (warps along replicates) is a template parameter: 4 (8 warps, 4 x (K+1) fragments each) or 8 (16
//      warps, 2 x (K+1) fragments each, <= 128 registers) -- ncu on the first form shows the warps stalled on
//      `wait` and `math_pipe_throttle` in equal parts with only two warps per scheduler to cover each other.
// (Folding the u-moment row S0 in as an extra fragment of channel tile 0 was tried and removed: the warp-uniform
//  branch around its mma.sync put a WARPSYNC into every k-step and cost 13 %; the separate scalar pass is 2 %.)
#pragma once
#include "resample_mma.cuh"

#ifndef TXB_MMA2_ST96
#define TXB_MMA2_ST96 1
#endif

namespace txb {

template <int K, typename FT, int WM_>
struct Rm2Cfg {
    static constexpr int ES = FreqElem<FT>::ES;
    static constexpr int VS = 16 / ES;
    static constexpr int WM = WM_, WN = 2;
    static constexpr int NT = WM * WN * 32;
    static constexpr int RT = (K + 1 <= 9) ? 128 : 64;    // replicate tile, as k_resample_mma
    static constexpr int MT = RT / (WM * 8);              // 8-replicate fragments per warp
    static constexpr int NTL = K + 1;
    static constexpr int CT = 16;
    static constexpr int NCOL = CT * (K + 1);
    static constexpr int PP = NCOL + 5;                   // odd pitch (NCOL is even): see header note 2
    static constexpr size_t SMEM64 = (size_t)3 * RT * (64 * ES + 16) + 2 * (size_t)64 * PP * 8;
    static constexpr size_t SMEM96 = (size_t)3 * RT * (96 * ES + 16) + 2 * (size_t)96 * PP * 8;
    // samples per stage = samples per block-wide barrier: as many as shared memory holds (227 KB per CTA minus the
    // 2 KB conversion table and the runtime's 1 KB)
    static constexpr int ST = (TXB_MMA2_ST96 && ES == 1 && SMEM96 <= 222 * 1024 && (96 * CT) % NT == 0)
                                  ? 96 : SMEM64 <= 200 * 1024 ? 64 : 32;
    static constexpr int NGRP = ST / 16;                  // 16-sample groups per stage
    static constexpr int NG = ST / VS;
    static constexpr int PITCH = ST * ES + 16;
    static constexpr int NSTAGE = 3;
    static constexpr int A_BYTES = RT * PITCH;
    static constexpr int P_DBL = ST * PP;
    static constexpr size_t SMEM = (size_t)NSTAGE * A_BYTES + 2 * (size_t)P_DBL * 8;
    static_assert(NCOL == WN * NTL * 8, "column tile must split into 8-wide fragments");
    static constexpr int NPIECE = ST * CT / NT;           // P elements per thread and stage
    static constexpr int PROWS = NT / CT;                 // samples covered by one piece
    static constexpr int GPP = NGRP / NPIECE;             // 16-sample groups per piece
    static_assert(MT >= 1 && NPIECE >= 1 && NPIECE * GPP == NGRP, "tile / thread geometry");
    static_assert((PP & 1) == 1, "P pitch must be odd");
};

// four consecutive table entries of one fragment row, as loaded from shared memory
template <int ES> struct APack;
template <> struct APack<1> { uint32_t w; };
template <> struct APack<2> { uint2 w; };
template <> struct APack<4> { uint4 w; };
template <> struct APack<8> { uint4 w0, w1; };

template <int ES>
__device__ __forceinline__ APack<ES> apack_load(const unsigned char* p);
template <> __device__ __forceinline__ APack<1> apack_load<1>(const unsigned char* p) {
    return APack<1>{*reinterpret_cast<const uint32_t*>(p)};
}
template <> __device__ __forceinline__ APack<2> apack_load<2>(const unsigned char* p) {
    return APack<2>{*reinterpret_cast<const uint2*>(p)};
}
template <> __device__ __forceinline__ APack<4> apack_load<4>(const unsigned char* p) {
    return APack<4>{*reinterpret_cast<const uint4*>(p)};
}
template <> __device__ __forceinline__ APack<8> apack_load<8>(const unsigned char* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    return APack<8>{q[0], q[1]};
}
template <int I> __device__ __forceinline__ uint32_t apack_get(const APack<1>& a) { return (a.w >> (8 * I)) & 0xffu; }
template <int I> __device__ __forceinline__ uint32_t apack_get(const APack<2>& a) {
    return ((I < 2 ? a.w.x : a.w.y) >> (16 * (I & 1))) & 0xffffu;
}
template <int I> __device__ __forceinline__ uint32_t apack_get(const APack<4>& a) {
    return I == 0 ? a.w.x : I == 1 ? a.w.y : I == 2 ? a.w.z : a.w.w;
}
template <int I> __device__ __forceinline__ uint32_t apack_get(const APack<8>& a) {   // low words (counts < 2^31)
    return I == 0 ? a.w0.x : I == 1 ? a.w0.z : I == 2 ? a.w1.x : a.w1.z;
}

// LUT: table entries become fp64 through a 256-entry shared-memory table instead of the 2^52 magic DADD (uint8 tables
// only).  ncu on the DADD form: the hottest stalled instructions of the kernel are those DADDs, queued behind the
// DMMAs in the same fp64 pipe with the warp's next fragments waiting on them.
template <int K, typename FT, int WM, bool LUT = false>
__global__ void __launch_bounds__(WM * 64, 1) k_resample_mma2(const RsArgs a) {
    using C = Rm2Cfg<K, FT, WM>;
    static_assert(!LUT || sizeof(FT) == 1, "the conversion table covers uint8 entries");
    constexpr int ES = C::ES, VS = C::VS, NT = C::NT, MT = C::MT, NTL = C::NTL, RT = C::RT, CT = C::CT;
    constexpr int PP = C::PP, ST = C::ST, NG = C::NG, PITCH = C::PITCH, NSTAGE = C::NSTAGE, NGRP = C::NGRP;
    constexpr int NPIECE = C::NPIECE, PROWS = C::PROWS, GPP = C::GPP;

    extern __shared__ __align__(16) unsigned char smem[];
    unsigned char* sA = smem;
    double* sP = reinterpret_cast<double*>(smem + (size_t)NSTAGE * C::A_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int wm = wrp % C::WM, wn = wrp / C::WM;
    const int fg = lane >> 2, fk = lane & 3;
    const int64_t c_base = (int64_t)blockIdx.x * CT;
    const int64_t r_base = (int64_t)blockIdx.y * RT;
    const int64_t s_begin = (int64_t)blockIdx.z * a.slab_len;
    const int64_t s_end = min(a.nsamp, s_begin + a.slab_len);
    const int nstage = (int)((s_end - s_begin + ST - 1) / ST);
    const bool has_x = a.x != nullptr, has_w = a.w != nullptr;
    const bool upc = a.ldu != 0;
    __shared__ double s_cvt[LUT ? 256 : 1];
    if (LUT) {
        for (int i = tid; i < 256; i += NT) s_cvt[i] = (double)i;   // visible after the prologue's barrier
    }

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

    // P operands: this thread regenerates channel cl of samples (tid / 16 + PROWS q), q = 0 .. NPIECE-1
    const int cl = tid % CT, sl0 = tid / CT;
    const int64_t my_c = c_base + cl;
    const bool c_ok = my_c < a.nchan;
    const double piv_c = c_ok ? a.pivot[my_c] : 0.0;          // cx[c] (comoments) or cu[c] (x_is_u wide)
    const double piv_u = upc ? 0.0 : a.pivot[a.nchan];
    double px[NPIECE], pu[NPIECE], pwt[NPIECE];
    auto prefetch_P = [&](int it) {
        const int64_t s0 = s_begin + (int64_t)it * ST;
#pragma unroll
        for (int q = 0; q < NPIECE; ++q) {
            // rows past the slab end carry weight 0 (every column of the row is then 0); channels past nchan
            // carry dx = 0 (x := its pivot)
            px[q] = piv_c, pu[q] = 0.0, pwt[q] = 0.0;
            const int64_t s = s0 + sl0 + PROWS * q;
            if (s < s_end && it < nstage) {
                pwt[q] = has_w ? __ldg(a.w + s) : 1.0;
                if (upc) {
                    if (c_ok) pu[q] = __ldg(a.u + s * a.ldu + my_c);
                } else {
                    pu[q] = __ldg(a.u + s);
                    if (c_ok && has_x) px[q] = __ldg(a.x + s * a.ldx + my_c);
                }
            }
        }
    };
    auto write_piece = [&](int it, int q) {
        double* row = sP + (size_t)(it & 1) * C::P_DBL + (size_t)(sl0 + PROWS * q) * PP;
        double du, lead;
        if (upc) {   // cols (c, j) = w du_c^j
            du = pu[q] - piv_c;
            lead = c_ok ? pwt[q] : 0.0;
        } else {     // cols (c, j) = w dx_c du^j
            du = pu[q] - piv_u;
            lead = pwt[q] * (px[q] - piv_c);
        }
        // (powers by squaring -- depth log2 K, 5 more multiplies -- measured 0.8 % slower: the extra fp64 ops cost more
        // in the shared pipe than the shorter dependency chain saves)
        double p = lead;
#pragma unroll
        for (int j = 0; j <= K; ++j) {
            row[cl * (K + 1) + j] = p;
            p *= du;
        }
    };

    double acc[MT][NTL][2];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int n = 0; n < NTL; ++n) acc[m][n][0] = 0.0, acc[m][n][1] = 0.0;

    // ---- prologue: A stages 0 and 1 in flight, P(0) written, P(1) operands in registers
#pragma unroll
    for (int p = 0; p < NSTAGE - 1; ++p) issue_A(p);
    prefetch_P(0);
#pragma unroll
    for (int q = 0; q < NPIECE; ++q) write_piece(0, q);
    prefetch_P(1);
    cp_async_wait<NSTAGE - 2>();
    __syncthreads();

    const int a_row0 = wm * MT * 8 + fg;
    const int b_col0 = wn * NTL * 8 + fg;
    for (int it = 0; it < nstage; ++it) {
        // buffer (it + 2) % 3 was read in stage it - 1; every warp has passed that stage's barrier
        issue_A(it + NSTAGE - 1);
        const unsigned char* bufA = sA + (size_t)(it % NSTAGE) * C::A_BYTES + (size_t)a_row0 * PITCH + 4 * fk * ES;
        const double* bufP = sP + (size_t)(it & 1) * C::P_DBL + (size_t)(4 * fk) * PP;

        APack<ES> aw[MT];
        double bf[NTL];
        auto fetch_a = [&](int g) {
#pragma unroll
            for (int m = 0; m < MT; ++m) aw[m] = apack_load<ES>(bufA + (size_t)(m * 8) * PITCH + g * 16 * ES);
        };
        auto fetch_b = [&](int g, int i) {
            const double* rowp = bufP + (size_t)(g * 16 + i) * PP;
#pragma unroll
            for (int n = 0; n < NTL; ++n) bf[n] = rowp[b_col0 + n * 8];
        };
        fetch_a(0);
        fetch_b(0, 0);
#pragma unroll
        for (int g = 0; g < NGRP; ++g) {
            APack<ES> ac[MT];
#pragma unroll
            for (int m = 0; m < MT; ++m) ac[m] = aw[m];
            if (g + 1 < NGRP) fetch_a(g + 1);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                double af[MT], bc[NTL];
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    const uint32_t v = i == 0 ? apack_get<0>(ac[m]) : i == 1 ? apack_get<1>(ac[m])
                                     : i == 2 ? apack_get<2>(ac[m]) : apack_get<3>(ac[m]);
                    af[m] = LUT ? s_cvt[v] : u32_to_f64(v);
                }
#pragma unroll
                for (int n = 0; n < NTL; ++n) bc[n] = bf[n];
                if (i + 1 < 4) fetch_b(g, i + 1);
                else if (g + 1 < NGRP) fetch_b(g + 1, 0);
#pragma unroll
                for (int m = 0; m < MT; ++m)
#pragma unroll
                    for (int n = 0; n < NTL; ++n) dmma884(acc[m][n][0], acc[m][n][1], af[m], bc[n]);
            }
            // operands of the NEXT stage, into the other P buffer (last read in stage it - 1)
            if ((g + 1) % GPP == 0) write_piece(it + 1, g / GPP);
        }
        prefetch_P(it + 2);
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();
    }
    cp_async_wait<0>();

    // ---- fragments straight to the slab partial
    const int64_t slab = blockIdx.z;
    double* dstp = upc ? a.ws0 : a.ws1;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const int64_t rr = r_base + a_row0 + m * 8;
        if (rr >= a.nrep) continue;
#pragma unroll
        for (int n = 0; n < NTL; ++n) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int col = (wn * NTL + n) * 8 + 2 * fk + h;
                const int ch = col / (K + 1), j = col % (K + 1);
                const int64_t c = c_base + ch;
                if (c < a.nchan) dstp[((slab * a.nrep + rr) * a.nchan + c) * (K + 1) + j] = acc[m][n][h];
            }
        }
    }
}

}  // namespace txb

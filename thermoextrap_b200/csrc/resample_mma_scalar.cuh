// K2 scalar-x contraction on the fp64 tensor cores (DMMA.8x8x4): the table-driven bootstrap for the
// reference sampler's own frequency table.
//
//   S[r, col] = sum_s F[r, s] * P[s, col],   cols = [ w du^j | w dx du^j ], j = 0..K   (second half only with x)
//
// Same pipeline as k_resample<RS_SCALAR> (resample.cuh) and the same slab partials; the micro-kernel is
// the DMMA one of resample_mma.cuh.  What it buys: with an int64 table (cmomy's freq[nrep, ndat], consumed
// unchanged) the DFMA tile needed 72 FMAs per 8-byte table element per thread-tile and ran at 48 % of the HBM
// roofline at order 8 (profiles/r01_configs_c4_c5_ksweep_perturb.jsonl); here a warp instruction carries 256
// FMAs and the kernel waits on the table instead.  Columns are padded to whole 8-wide fragments (NCOLP), so for
// narrow tables it only replaces the DFMA tile where the padding is small (see rs_use_mma_scalar).
//   8 warps along the replicates, each MT row fragments x NTL column fragments.
#pragma once
#include "resample_mma.cuh"

namespace txb {

template <int K, bool HAS_X, typename FT>
struct RmsCfg {
    static constexpr int ES = FreqElem<FT>::ES;
    static constexpr int VS = 16 / ES;
    static constexpr int NT = 256;
    // wide table elements: one row fragment per warp (RT = 64) so that three CTAs fit on an SM (73 KB of
    // shared memory each) and the per-stage barrier of one overlaps the loads / DMMAs of the others
    static constexpr int MT = ES >= 4 ? 1 : 2;
    static constexpr int RT = (NT / 32) * MT * 8;                 // replicates per CTA (64 or 128)
    static constexpr int NCOL = HAS_X ? 2 * (K + 1) : (K + 1);    // useful columns
    static constexpr int NTL = (NCOL + 7) / 8;                    // 8-column fragments
    static constexpr int NCOLP = NTL * 8;
    static constexpr int PP = (NCOLP + 15) / 16 * 16 + 4;         // P pitch (doubles), == 4 mod 16: B loads conflict-free
    static constexpr int ST = ES == 1 ? 128 : 256 / ES;           // samples per stage (128, 128, 64, 32)
    static constexpr int NG = ST / VS;
    static constexpr int PITCH = ST * ES + (ES == 8 ? 32 : 16);   // F row pitch (bytes): A loads conflict-free
    static constexpr int NSTAGE = 3;
    static constexpr int A_BYTES = RT * PITCH;
    static constexpr int P_DBL = ST * PP;
    static constexpr size_t SMEM = (size_t)NSTAGE * A_BYTES + 2 * (size_t)P_DBL * 8;
};

template <int K, bool HAS_X, typename FT>
__global__ void __launch_bounds__(256, 3) k_resample_mma_scalar(const RsArgs a) {
    using C = RmsCfg<K, HAS_X, FT>;
    constexpr int ES = C::ES, VS = C::VS, NT = C::NT, MT = C::MT, NTL = C::NTL, RT = C::RT;
    constexpr int PP = C::PP, ST = C::ST, NG = C::NG, PITCH = C::PITCH, NSTAGE = C::NSTAGE;

    extern __shared__ __align__(16) unsigned char smem[];
    unsigned char* sA = smem;
    double* sP = reinterpret_cast<double*>(smem + (size_t)NSTAGE * C::A_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int fg = lane >> 2, fk = lane & 3;
    const int64_t r_base = (int64_t)blockIdx.y * RT;
    const int64_t s_begin = (int64_t)blockIdx.z * a.slab_len;
    const int64_t s_end = min(a.nsamp, s_begin + a.slab_len);
    const int nstage = (int)((s_end - s_begin + ST - 1) / ST);
    const bool has_w = a.w != nullptr;
    // uint8 entries become fp64 through a shared-memory table, not the 2^52 DADD (which queues behind the DMMAs in the
    // fp64 pipe: resample_mma2.cuh); visible after the first block-wide barrier of the stage loop
    constexpr bool LUT = sizeof(FT) == 1;
    __shared__ double s_cvt[LUT ? 256 : 1];
    if (LUT) {
        for (int i = tid; i < 256; i += NT) s_cvt[i] = (double)i;
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

    // ---- operand B: one sample per thread (threads >= ST idle), prefetched one stage ahead
    double px = 0.0, pu = 0.0, pwt = 0.0;
    auto prefetch_P = [&](int it) {
        const int64_t s = s_begin + (int64_t)it * ST + tid;
        px = 0.0, pu = 0.0, pwt = 0.0;
        if (tid < ST && s < s_end && it < nstage) {
            pu = __ldg(a.u + s);
            px = HAS_X ? __ldg(a.x + s * (a.ldx > 0 ? a.ldx : 1)) : 0.0;
            pwt = has_w ? __ldg(a.w + s) : 1.0;
        }
    };
    auto write_P = [&](int it) {
        if (tid < ST) {
            double* row = sP + (size_t)(it & 1) * C::P_DBL + (size_t)tid * PP;
            const double du = pu - a.pivot[HAS_X ? 1 : 0];
            const double dx = HAS_X ? px - a.pivot[0] : 0.0;
            double p = pwt;  // 0 for out-of-range samples
#pragma unroll
            for (int j = 0; j <= K; ++j) {
                row[j] = p;
                if (HAS_X) row[K + 1 + j] = p * dx;
                p *= du;
            }
#pragma unroll
            for (int j = C::NCOL; j < C::NCOLP; ++j) row[j] = 0.0;
        }
    };

    double acc[MT][NTL][2];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int n = 0; n < NTL; ++n) acc[m][n][0] = 0.0, acc[m][n][1] = 0.0;

#pragma unroll
    for (int p = 0; p < NSTAGE - 1; ++p) issue_A(p);
    prefetch_P(0);

    const int a_row0 = wrp * MT * 8 + fg;
    for (int it = 0; it < nstage; ++it) {
        write_P(it);
        prefetch_P(it + 1);
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();
        issue_A(it + NSTAGE - 1);

        const unsigned char* bufA = sA + (size_t)(it % NSTAGE) * C::A_BYTES + (size_t)a_row0 * PITCH;
        const double* bufP = sP + (size_t)(it & 1) * C::P_DBL + fg;
#pragma unroll 4
        for (int t = 0; t < ST / 4; ++t) {
            const int sl = 4 * t + fk;
            double af[MT], bf[NTL];
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const FT v = *reinterpret_cast<const FT*>(bufA + (size_t)(m * 8) * PITCH + sl * ES);
                af[m] = LUT ? s_cvt[(uint32_t)v & 0xffu] : u32_to_f64((uint32_t)v);  // counts are < 2^31
            }
#pragma unroll
            for (int n = 0; n < NTL; ++n) bf[n] = bufP[(size_t)sl * PP + n * 8];
#pragma unroll
            for (int m = 0; m < MT; ++m)
#pragma unroll
                for (int n = 0; n < NTL; ++n) dmma884(acc[m][n][0], acc[m][n][1], af[m], bf[n]);
        }
    }
    cp_async_wait<0>();

    const int64_t slab = blockIdx.z;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const int64_t rr = r_base + a_row0 + m * 8;
        if (rr >= a.nrep) continue;
#pragma unroll
        for (int n = 0; n < NTL; ++n) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int col = n * 8 + 2 * fk + h;
                if (col <= K)
                    a.ws0[(slab * a.nrep + rr) * (K + 1) + col] = acc[m][n][h];
                else if (HAS_X && col < 2 * (K + 1))
                    a.ws1[(slab * a.nrep + rr) * (K + 1) + col - (K + 1)] = acc[m][n][h];
            }
        }
    }
}

}  // namespace txb

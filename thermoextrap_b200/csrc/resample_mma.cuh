// K2 wide-x contraction on the fp64 tensor cores (DMMA.8x8x4 via mma.sync.m8n8k4).
//
//   S[r, (c, j)] = sum_s F[r, s] * P[s, (c, j)],   P[s, (c, j)] = w_s dx_{s,c} du_s^j  (or w_s du_{s,c}^j)
//
// Same operands, pipeline and slab partials as k_resample (resample.cuh); only the micro-kernel
// differs.  Why: a register-tiled DFMA contraction needs three 64-bit register operands per FMA and
// tops out at 26.6-28.3 TFLOP/s on a B200 (txb_fp64_tile_peak_launch, profiles/r01_fp64_peaks.json),
// the same loop issued as DMMA reaches the pipe's 36.5-37.0 TFLOP/s with 8 warps per SM because one
// warp instruction carries 256 FMAs for 4 operand registers.
//   M <-> replicates (F, converted to fp64 in registers: one DADD each), N <-> columns, K <-> samples.
//   CTA tile: RT replicates x 16 channels x (K+1) powers; 8 warps as 4 (replicates) x 2 (columns);
//   a warp owns MT x (K+1) accumulator fragments (8 x 8 each, 2 doubles per thread).
//   fragment layout (PTX ISA, mma.m8n8k4 .f64): g = lane >> 2, k = lane & 3:
//     A[g][k]   B[k][g]   C[g][2k], C[g][2k+1]
#pragma once
#include "resample.cuh"

namespace txb {

template <int K, typename FT>
struct RmCfg {
    static constexpr int ES = FreqElem<FT>::ES;
    static constexpr int VS = 16 / ES;
    static constexpr int NT = 256;
    static constexpr int WM = 4, WN = 2;                  // warps along replicates / columns
    static constexpr int MT = (K + 1 <= 9) ? 4 : 2;       // 8-replicate fragments per warp
    static constexpr int NTL = K + 1;                     // 8-column fragments per warp
    static constexpr int RT = WM * MT * 8;                // replicate tile (128 or 64, as k_resample)
    static constexpr int CT = 16;                         // channels per CTA
    static constexpr int NCOL = CT * (K + 1);             // = WN * NTL * 8
    static constexpr int PP = NCOL + 4;                   // P row pitch (doubles): B-fragment loads conflict-free
    // samples per stage: 64 (half the block-wide syncs of 32) while three F stages + two P buffers fit
    static constexpr size_t SMEM64 = (size_t)3 * RT * (64 * ES + 16) + 2 * (size_t)64 * PP * 8;
    static constexpr int ST = SMEM64 <= 160 * 1024 ? 64 : 32;
    static constexpr int NG = ST / VS;
    static constexpr int PITCH = ST * ES + 16;            // F row pitch (bytes)
    static constexpr int NSTAGE = 3;
    static constexpr int A_BYTES = RT * PITCH;
    static constexpr int P_DBL = ST * PP;
    static constexpr size_t SMEM = (size_t)NSTAGE * A_BYTES + 2 * (size_t)P_DBL * 8;
    static_assert(NCOL == WN * NTL * 8, "column tile must split into 8-wide fragments");
};

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

template <int K, typename FT>
__global__ void __launch_bounds__(256, 1) k_resample_mma(const RsArgs a) {
    using C = RmCfg<K, FT>;
    constexpr int ES = C::ES, VS = C::VS, NT = C::NT, MT = C::MT, NTL = C::NTL, RT = C::RT, CT = C::CT;
    constexpr int PP = C::PP, ST = C::ST, NG = C::NG, PITCH = C::PITCH, NSTAGE = C::NSTAGE;

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

    constexpr int NPRE = (ST * CT + NT - 1) / NT;
    double px[NPRE], pu[NPRE], pwt[NPRE];
    auto prefetch_P = [&](int it) {
        const int64_t s0 = s_begin + (int64_t)it * ST;
#pragma unroll
        for (int q = 0; q < NPRE; ++q) {
            px[q] = 0.0, pu[q] = 0.0, pwt[q] = 0.0;
            const int e = tid + NT * q;
            const int sl = e / CT, cl = e % CT;
            const int64_t s = s0 + sl, c = c_base + cl;
            if (e < ST * CT && s < s_end && c < a.nchan && it < nstage) {
                pu[q] = upc ? __ldg(a.u + s * a.ldu + c) : __ldg(a.u + s);
                px[q] = has_x ? __ldg(a.x + s * a.ldx + c) : 0.0;
                pwt[q] = has_w ? __ldg(a.w + s) : 1.0;
            }
        }
    };
    auto write_P = [&](int it) {
        double* dst = sP + (size_t)(it & 1) * C::P_DBL;
#pragma unroll
        for (int q = 0; q < NPRE; ++q) {
            const int e = tid + NT * q;
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
                double* row = dst + (size_t)sl * PP + cl * (K + 1);
#pragma unroll
                for (int j = 0; j <= K; ++j) {
                    row[j] = p;
                    p *= du;
                }
            }
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

    // this thread's A element of fragment m sits in row (wm * MT + m) * 8 + fg, its B element of
    // fragment n in column (wn * NTL + n) * 8 + fg; both at sample (4 t + fk) of the stage
    const int a_row0 = wm * MT * 8 + fg;
    const int b_col0 = wn * NTL * 8 + fg;
    for (int it = 0; it < nstage; ++it) {
        write_P(it);
        prefetch_P(it + 1);
        cp_async_wait<NSTAGE - 2>();
        __syncthreads();
        issue_A(it + NSTAGE - 1);

        const unsigned char* bufA = sA + (size_t)(it % NSTAGE) * C::A_BYTES + (size_t)a_row0 * PITCH;
        const double* bufP = sP + (size_t)(it & 1) * C::P_DBL + b_col0;
        // fragments of step t+1 are fetched (LDS) before the DMMAs of step t are issued
        FT av[MT];
        double bf[NTL];
        auto fetch = [&](int t) {
            const int sl = 4 * t + fk;
#pragma unroll
            for (int m = 0; m < MT; ++m) av[m] = *reinterpret_cast<const FT*>(bufA + (size_t)(m * 8) * PITCH + sl * ES);
#pragma unroll
            for (int n = 0; n < NTL; ++n) bf[n] = bufP[(size_t)sl * PP + n * 8];
        };
        fetch(0);
#pragma unroll 4
        for (int t = 0; t < ST / 4; ++t) {
            double af[MT], bc[NTL];
#pragma unroll
            for (int m = 0; m < MT; ++m) af[m] = u32_to_f64((uint32_t)av[m]);
#pragma unroll
            for (int n = 0; n < NTL; ++n) bc[n] = bf[n];
            if (t + 1 < ST / 4) fetch(t + 1);
#pragma unroll
            for (int m = 0; m < MT; ++m)
#pragma unroll
                for (int n = 0; n < NTL; ++n) dmma884(acc[m][n][0], acc[m][n][1], af[m], bc[n]);
        }
    }
    cp_async_wait<0>();

    // ---- fragments straight to the slab partial (one sample phase: nothing to merge)
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
                const int cl = col / (K + 1), j = col % (K + 1);
                const int64_t c = c_base + cl;
                if (c < a.nchan) dstp[((slab * a.nrep + rr) * a.nchan + c) * (K + 1) + j] = acc[m][n][h];
            }
        }
    }
}

}  // namespace txb

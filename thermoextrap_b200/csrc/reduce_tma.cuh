// Wide-x shifted-power sums with rows staged through a shared-memory ring that the copy engine
// fills (cp.async.bulk + mbarrier transaction counts).  Same arithmetic, thread->channel map and
// partial layout as k_reduce_wide; what changes is where the memory-level parallelism lives: the
// ring keeps nstage tiles (tens of KB per CTA) in flight without holding them in registers.
#pragma once
#include <cuda.h>

#include "txb_common.cuh"

namespace txb {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a byte-count mismatch must surface as a CUDA error, never as a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin)
        if (spin > (1u << 24)) __trap();
}
// global -> shared bulk copy (16-byte aligned addresses, size a multiple of 16)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// 2-D tiled copy through a tensor map: box (rows x channels) at element coordinates (c0, r0);
// out-of-range elements arrive as zeros and still count towards the transaction bytes
__device__ __forceinline__ void tensor_g2s_2d(void* dst, const CUtensorMap* map, int c0, int r0, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(r0)
        : "memory");
}

constexpr int WT_UNR = 4;       // rows per thread per tile
constexpr int WT_MAXSTAGE = 4;

template <int K, int CPT, bool HAS_X, bool UPC>
__global__ void __launch_bounds__(256, 2)
    k_reduce_wide_tma(const double* __restrict__ x, const double* __restrict__ u,
                      const double* __restrict__ w, int64_t nsamp, int64_t nchan, int64_t ldx,
                      int64_t ldu, const double* __restrict__ pivot, int lanes_c, int nstage,
                      double* __restrict__ ws0, double* __restrict__ ws1, int use_map,
                      const __grid_constant__ CUtensorMap map_x,
                      const __grid_constant__ CUtensorMap map_u) {
    static_assert(HAS_X || UPC, "u-only shared-u sums use the scalar kernel");
    static_assert(CPT % 2 == 0, "16-byte shared loads");
    extern __shared__ __align__(128) double ring[];
    __shared__ uint64_t full[WT_MAXSTAGE];
    constexpr int NARR = (HAS_X ? 1 : 0) + (UPC ? 1 : 0);

    const int tid = threadIdx.x;
    const int lane_c = tid & (lanes_c - 1);
    const int rowslot = tid / lanes_c;
    const int rows = 256 / lanes_c;
    const int R = rows * WT_UNR;  // rows per tile
    const int W = lanes_c * CPT;  // channels per tile
    const int64_t tile_base = (int64_t)blockIdx.x * W;
    const int w_eff = (int)min((int64_t)W, nchan - tile_base);
    auto chan_of = [&](int cc) -> int64_t {
        return tile_base + (int64_t)(cc >> 1) * 2 * lanes_c + 2 * lane_c + (cc & 1);
    };
    bool act[CPT];
#pragma unroll
    for (int cc = 0; cc < CPT; ++cc) act[cc] = chan_of(cc) < nchan;
    const bool has_w = (w != nullptr);

    const int64_t rows_per_slab = (nsamp + gridDim.y - 1) / gridDim.y;
    const int64_t r_begin = (int64_t)blockIdx.y * rows_per_slab;
    const int64_t r_end = min(nsamp, r_begin + rows_per_slab);
    const int64_t ntiles = r_end > r_begin ? (r_end - r_begin + R - 1) / R : 0;

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

    const size_t arr_doubles = (size_t)R * W;
    const size_t stage_doubles = arr_doubles * NARR;
    const uint32_t row_bytes = (uint32_t)w_eff * 8u;
    // whole tile is one contiguous run of memory -> one copy per array
    const bool one_x = !HAS_X || (w_eff == W && ldx == W);
    const bool one_u = !UPC || (w_eff == W && ldu == W);

    // warp 0: start the copies of tile t into its ring slot
    auto issue = [&](int64_t t) {
        const int lane = tid & 31;
        const int s = (int)(t % nstage);
        const int64_t r0 = r_begin + t * R;
        const int nr = (int)min((int64_t)R, r_end - r0);
        double* dst = ring + (size_t)s * stage_doubles;
        if (use_map) {
            // strided rows: one tiled copy per array; the box is always R x W (zero filled past the edges)
            if (lane == 0) {
                mbar_arrive_expect_tx(&full[s], (uint32_t)(stage_doubles * sizeof(double)));
                if (HAS_X) tensor_g2s_2d(dst, &map_x, (int)tile_base, (int)r0, &full[s]);
                if (UPC) tensor_g2s_2d(dst + (HAS_X ? arr_doubles : 0), &map_u, (int)tile_base, (int)r0, &full[s]);
            }
            return;
        }
        if (lane == 0) mbar_arrive_expect_tx(&full[s], (uint32_t)nr * row_bytes * NARR);
        __syncwarp();
        if (HAS_X) {
            if (one_x) {
                if (lane == 0) bulk_g2s(dst, x + r0 * ldx, (uint32_t)nr * row_bytes, &full[s]);
            } else {
                for (int i = lane; i < nr; i += 32)
                    bulk_g2s(dst + (size_t)i * W, x + (r0 + i) * ldx + tile_base, row_bytes, &full[s]);
            }
        }
        if (UPC) {
            double* du_ = dst + (HAS_X ? arr_doubles : 0);
            if (one_u) {
                if (lane == 0) bulk_g2s(du_, u + r0 * ldu, (uint32_t)nr * row_bytes, &full[s]);
            } else {
                for (int i = lane; i < nr; i += 32)
                    bulk_g2s(du_ + (size_t)i * W, u + (r0 + i) * ldu + tile_base, row_bytes, &full[s]);
            }
        }
    };

    if (tid == 0) {
        for (int s = 0; s < nstage; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid < 32)
        for (int64_t t = 0; t < ntiles && t < nstage; ++t) issue(t);

    // per-row scalars (shared u, weights) come straight from global, one tile ahead
    double un[WT_UNR], wn[WT_UNR];
    auto fetch_row_scalars = [&](int64_t t) {
#pragma unroll
        for (int q = 0; q < WT_UNR; ++q) {
            const int64_t rr = r_begin + t * R + rowslot + (int64_t)q * rows;
            const bool ok = t < ntiles && rr < r_end;
            un[q] = (!UPC && ok) ? __ldg(u + rr) : 0.0;
            wn[q] = (has_w && ok) ? __ldg(w + rr) : 1.0;
        }
    };
    fetch_row_scalars(0);

    for (int64_t t = 0; t < ntiles; ++t) {
        const int s = (int)(t % nstage);
        const uint32_t parity = (uint32_t)((t / nstage) & 1);
        const int nr = (int)min((int64_t)R, r_end - (r_begin + t * R));
        double uc[WT_UNR], wc[WT_UNR];
#pragma unroll
        for (int q = 0; q < WT_UNR; ++q) uc[q] = un[q], wc[q] = wn[q];
        fetch_row_scalars(t + 1);
        mbar_wait(&full[s], parity);
        const double* xs = ring + (size_t)s * stage_doubles;
        const double* us = xs + (HAS_X ? arr_doubles : 0);
#pragma unroll
        for (int q = 0; q < WT_UNR; ++q) {
            const int i = rowslot + q * rows;
            if (i < nr) {
                double xv[CPT], uv[CPT];
#pragma unroll
                for (int h = 0; h < CPT / 2; ++h) {
                    const size_t off = (size_t)i * W + (size_t)h * 2 * lanes_c + 2 * lane_c;
                    if (HAS_X) {
                        const double2 tx = *reinterpret_cast<const double2*>(xs + off);
                        xv[2 * h] = tx.x, xv[2 * h + 1] = tx.y;
                    }
                    if (UPC) {
                        const double2 tu = *reinterpret_cast<const double2*>(us + off);
                        uv[2 * h] = tu.x, uv[2 * h + 1] = tu.y;
                    }
                }
                const double wv = wc[q];
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
                    const double du = uc[q] - cu[0];
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
            }
        }
        __syncthreads();  // every reader is done with slot s
        if (tid < 32 && t + nstage < ntiles) issue(t + nstage);
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

}  // namespace txb

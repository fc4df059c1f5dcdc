// Wide-x shifted-power sums with rows staged through a shared-memory ring that the copy engine
// fills (cp.async.bulk / cp.async.bulk.tensor + mbarrier transaction counts).  Same arithmetic,
// thread->channel map and partial layout (ws[slab][chan][K+1]) as k_reduce_wide in reduce.cu.
//
// Two tile shapes:
//   contiguous : the CTA covers every channel of a row (lanes_c * CPT >= nchan), so a tile of R rows
//                is one run of R * ld doubles and arrives with ONE bulk copy; shared-memory row
//                pitch = ld.  256 threads for <= 128 channels, 512 threads (one CTA per SM) for
//                rows up to 2048 channels -- long runs keep DRAM pages open, which 1 KB channel
//                strips do not (ncu: 54 % vs 76 % DRAM throughput).
//   tensor map : channel strips of lanes_c * CPT channels for strided / very wide rows (opt-in).
#include <stdlib.h>

#include <algorithm>

#include "reduce_tma.cuh"

namespace txb {

constexpr int WT_UNR = 4;  // max rows per thread per tile

template <int K, int CPT, bool HAS_X, bool UPC, int NT>
__global__ void __launch_bounds__(NT, NT == 256 ? 2 : 1)
    k_reduce_wide_tma(const double* __restrict__ x, const double* __restrict__ u,
                      const double* __restrict__ w, int64_t nsamp, int64_t nchan, int64_t ldx,
                      int64_t ldu, const double* __restrict__ pivot, int lanes_c, int unr, int nstage,
                      double* __restrict__ ws0, double* __restrict__ ws1, int use_map,
                      const __grid_constant__ CUtensorMap map_x,
                      const __grid_constant__ CUtensorMap map_u) {
    static_assert(HAS_X || UPC, "u-only shared-u sums use the scalar kernel");
    static_assert(CPT % 2 == 0, "16-byte shared loads");
    extern __shared__ __align__(128) double ring[];
    __shared__ uint64_t full[WT_MAXSTAGE];

    const int tid = threadIdx.x;
    // lanes_c need not be a power of two: threads past rows * lanes_c idle (rowslot = R)
    const int rows = NT / lanes_c;
    const int R = rows * unr;     // rows per tile
    const bool spare = tid >= rows * lanes_c;
    const int lane_c = tid % lanes_c;
    const int rowslot = spare ? R : tid / lanes_c;
    const int W = lanes_c * CPT;  // channels the CTA covers
    const int64_t tile_base = (int64_t)blockIdx.x * W;
    auto chan_of = [&](int cc) -> int64_t {
        return tile_base + (int64_t)(cc >> 1) * 2 * lanes_c + 2 * lane_c + (cc & 1);
    };
    bool act[CPT];
#pragma unroll
    for (int cc = 0; cc < CPT; ++cc) act[cc] = !spare && chan_of(cc) < nchan;
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

    // shared-memory row pitch (doubles) and array sizes of one stage
    const int pitch_x = use_map ? W : (int)ldx;
    const int pitch_u = use_map ? W : (int)ldu;
    const size_t arr_x = HAS_X ? (((size_t)R * pitch_x + 15) & ~(size_t)15) : 0;
    const size_t arr_u = UPC ? (((size_t)R * pitch_u + 15) & ~(size_t)15) : 0;
    const size_t stage_doubles = arr_x + arr_u;

    // warp 0, lane 0: start the copies of tile t into its ring slot
    auto issue = [&](int64_t t) {
        const int s = (int)(t % nstage);
        const int64_t r0 = r_begin + t * R;
        const int nr = (int)min((int64_t)R, r_end - r0);
        double* dst = ring + (size_t)s * stage_doubles;
        if (use_map) {
            // the box is always R x W: elements past the edges arrive as zeros and count as bytes
            mbar_arrive_expect_tx(&full[s], (uint32_t)(((size_t)R * W * sizeof(double)) * ((HAS_X ? 1 : 0) + (UPC ? 1 : 0))));
            if (HAS_X) tensor_g2s_2d(dst, &map_x, (int)tile_base, (int)r0, &full[s]);
            if (UPC) tensor_g2s_2d(dst + arr_x, &map_u, (int)tile_base, (int)r0, &full[s]);
        } else {
            // whole rows: one run per array, without the padding after the last row
            const uint32_t bx = HAS_X ? (uint32_t)(((int64_t)(nr - 1) * ldx + nchan) * sizeof(double)) : 0u;
            const uint32_t bu = UPC ? (uint32_t)(((int64_t)(nr - 1) * ldu + nchan) * sizeof(double)) : 0u;
            mbar_arrive_expect_tx(&full[s], bx + bu);
            if (HAS_X) bulk_g2s(dst, x + r0 * ldx, bx, &full[s]);
            if (UPC) bulk_g2s(dst + arr_x, u + r0 * ldu, bu, &full[s]);
        }
    };

    if (tid == 0) {
        for (int s = 0; s < nstage; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        for (int64_t t = 0; t < ntiles && t < nstage; ++t) issue(t);
    }
    __syncthreads();

    // per-row scalars (shared u, weights): 8-byte cp.async into a double-buffered side array, one
    // tile ahead -- no registers held across the tile, visible after the tile's closing barrier
    double* rowsc = ring + (size_t)nstage * stage_doubles + W;  // [2][2][R]: (buffer, u|w, row)
    auto fetch_row_scalars = [&](int64_t t) {
        if (t >= ntiles) return;
        const int64_t r0 = r_begin + t * R;
        double* dst = rowsc + (size_t)(t & 1) * 2 * R;
        for (int k = tid; k < R && r0 + k < r_end; k += NT) {
            if (!UPC) cp_async8(dst + k, u + r0 + k);
            if (has_w) cp_async8(dst + R + k, w + r0 + k);
        }
    };
    fetch_row_scalars(0);
    cp_async_wait_all();
    __syncthreads();

    for (int64_t t = 0; t < ntiles; ++t) {
        const int s = (int)(t % nstage);
        const uint32_t parity = (uint32_t)((t / nstage) & 1);
        const int nr = (int)min((int64_t)R, r_end - (r_begin + t * R));
        const double* usc = rowsc + (size_t)(t & 1) * 2 * R;
        fetch_row_scalars(t + 1);
        mbar_wait(&full[s], parity);
        const double* xs = ring + (size_t)s * stage_doubles;
        const double* us = xs + arr_x;
#pragma unroll
        for (int q = 0; q < WT_UNR; ++q) {
            const int i = rowslot + q * rows;
            if (i < nr) {  // nr <= rows * unr, so this also bounds q by unr
                double xv[CPT], uv[CPT];
                // channels past nchan read whatever follows in the ring (the host pads it by one
                // tile width); their sums are never written
#pragma unroll
                for (int h = 0; h < CPT / 2; ++h) {
                    const int coff = h * 2 * lanes_c + 2 * lane_c;
                    if (HAS_X) {
                        const double2 tx = *reinterpret_cast<const double2*>(xs + (size_t)i * pitch_x + coff);
                        xv[2 * h] = tx.x, xv[2 * h + 1] = tx.y;
                    }
                    if (UPC) {
                        const double2 tu = *reinterpret_cast<const double2*>(us + (size_t)i * pitch_u + coff);
                        uv[2 * h] = tu.x, uv[2 * h + 1] = tu.y;
                    }
                }
                const double wv = has_w ? usc[R + i] : 1.0;
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
                    const double du = usc[i] - cu[0];
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
        cp_async_wait_all();
        __syncthreads();  // every reader is done with slot s; next tile's row scalars have landed
        if (tid == 0 && t + nstage < ntiles) issue(t + nstage);
    }

    // block merge over the row slots that share a channel lane
    __shared__ double sm[NT];
    auto block_merge = [&](double v, double* dst, bool ok) {
        if (rows == 1) {
            if (ok) *dst = v;
            return;
        }
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

// ---------------------------------------------------------------- host side
// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*TensorMapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                      CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                      CUtensorMapFloatOOBfill);
static TensorMapEncodeFn tensor_map_encoder() {
    static TensorMapEncodeFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (TensorMapEncodeFn)p;
        cudaGetLastError();
    }
    return fn;
}

// [nsamp, nchan] fp64 matrix with row pitch ld, boxes of R rows x W channels
static bool make_tensor_map(CUtensorMap* m, const double* a, int64_t nsamp, int64_t nchan, int64_t ld,
                            int R, int W) {
    const cuuint64_t dims[2] = {(cuuint64_t)nchan, (cuuint64_t)nsamp};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(double)};
    const cuuint32_t box[2] = {(cuuint32_t)W, (cuuint32_t)R};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = tensor_map_encoder()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(a), dims,
                                            strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) set_error("cuTensorMapEncodeTiled failed (%d)", (int)r);
    return r == CUDA_SUCCESS;
}

// TXB_WIDE_TMA: 0 = register-staged kernel only; 1 (default) = ring kernel wherever a CTA can cover
// whole rows; 2 = also tensor-map channel strips for what is left (measured no faster than the
// register-staged kernel: 1 KB strips do not keep DRAM pages open); 3 = as 2 but never 512 threads.
static int ring_mode() {
    const char* e = getenv("TXB_WIDE_TMA");
    return e ? atoi(e) : 1;
}

static size_t round16(size_t n) { return (n + 15) & ~(size_t)15; }

WideRingPlan plan_wide_ring(int64_t nsamp, int64_t nchan, int64_t ldx, int64_t ldu, int order, bool has_x,
                            bool upc, bool aligned, int64_t max_slabs_x_chan) {
    WideRingPlan p{};
    const int mode = ring_mode();
    if (!mode || !aligned || nchan % 2 != 0) return p;
    const int narr = (has_x ? 1 : 0) + (upc ? 1 : 0);
    const int64_t ldmax = std::max<int64_t>(has_x ? ldx : 0, upc ? ldu : 0);
    const bool packed = ldmax <= nchan + nchan / 4;  // whole-row runs carry the row padding along
    // four channels per thread where the accumulators fit in 128 registers (order < 10)
    const int cpt4 = (nchan % 4 == 0 && narr == 1 && order < 10) ? 4 : 2;
    const int lanes = (int)((nchan + cpt4 - 1) / cpt4);
    if (packed && lanes <= 32) {
        p.contig = true, p.nthreads = 256, p.cpt = cpt4, p.lanes_c = lanes, p.nctiles = 1;
    } else if (packed && mode != 3 && lanes <= 512) {
        p.contig = true, p.nthreads = 512, p.cpt = cpt4, p.lanes_c = lanes, p.nctiles = 1;
    } else if (mode >= 2 && nsamp < ((int64_t)1 << 31) && tensor_map_encoder()) {
        p.contig = false, p.nthreads = 256, p.cpt = cpt4, p.lanes_c = 32;
        p.nctiles = (lanes + 31) / 32;
    } else {
        return p;
    }
    const int rows = p.nthreads / p.lanes_c;
    const size_t pad = (size_t)p.lanes_c * p.cpt * sizeof(double);  // over-read slack after the last slot
    const size_t budget = (p.nthreads == 256 ? (110u << 10) : (216u << 10)) - pad - 4 * (size_t)rows * WT_UNR * sizeof(double);
    size_t stage = 0;
    for (p.unr = WT_UNR; p.unr >= 1; --p.unr) {
        const size_t R = (size_t)rows * p.unr;
        if (!p.contig && R > 256) continue;  // tensor-map box limit
        stage = p.contig ? (has_x ? round16(R * ldx) : 0) + (upc ? round16(R * ldu) : 0)
                         : R * p.lanes_c * p.cpt * narr;
        stage *= sizeof(double);
        if (2 * stage <= budget) break;
    }
    if (p.unr < 1) return p;
    p.nstage = (int)std::min<size_t>(WT_MAXSTAGE, budget / stage);
    p.smem = stage * p.nstage + pad + 4 * (size_t)rows * p.unr * sizeof(double);  // + row scalars [2][2][R]
    const int64_t R = (int64_t)rows * p.unr;
    int64_t nslabs = ((int64_t)sm_count() * (p.nthreads == 256 ? 2 : 1) + p.nctiles - 1) / p.nctiles;
    nslabs = std::min(nslabs, nsamp / (R * 2));  // at least two tiles per CTA
    nslabs = std::min(nslabs, max_slabs_x_chan / nchan);
    if (nslabs * p.nctiles * 2 < sm_count()) return p;  // short inputs: finer slabs of the register kernel win
    p.nslabs = nslabs;
    p.ok = true;
    return p;
}

template <int K, int CPT, bool HAS_X, bool UPC, int NT>
static int launch_k(const WideRingPlan& p, const double* x, const double* u, const double* w, int64_t nsamp,
                    int64_t nchan, int64_t ldx, int64_t ldu, const double* pivot, double* ws0, double* ws1,
                    cudaStream_t st) {
    auto kern = k_reduce_wide_tma<K, CPT, HAS_X, UPC, NT>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 << 10);
        TXB_CUDA_LAUNCH_CHECK("k_reduce_wide_tma attribute");
        attr_set = true;
    }
    CUtensorMap mx{}, mu{};
    if (!p.contig) {
        const int R = (NT / p.lanes_c) * p.unr, W = p.lanes_c * CPT;
        if (HAS_X && !make_tensor_map(&mx, x, nsamp, nchan, ldx, R, W)) return TXB_ERR_CUDA;
        if (UPC && !make_tensor_map(&mu, u, nsamp, nchan, ldu, R, W)) return TXB_ERR_CUDA;
    }
    const dim3 grid((unsigned)p.nctiles, (unsigned)p.nslabs);
    kern<<<grid, NT, p.smem, st>>>(x, u, w, nsamp, nchan, ldx, ldu, pivot, p.lanes_c, p.unr, p.nstage, ws0, ws1,
                                   p.contig ? 0 : 1, mx, mu);
    TXB_CUDA_LAUNCH_CHECK("k_reduce_wide_tma");
    return TXB_OK;
}

template <int K, bool HAS_X, bool UPC>
static int launch_mode(const WideRingPlan& p, const double* x, const double* u, const double* w, int64_t nsamp,
                       int64_t nchan, int64_t ldx, int64_t ldu, const double* pivot, double* ws0, double* ws1,
                       cudaStream_t st) {
    if (p.cpt == 4) {
        if constexpr (!(HAS_X && UPC)) {
            if (p.nthreads == 512)
                return launch_k<K, 4, HAS_X, UPC, 512>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
            return launch_k<K, 4, HAS_X, UPC, 256>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
        }
    }
    if (p.nthreads == 512)
        return launch_k<K, 2, HAS_X, UPC, 512>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
    return launch_k<K, 2, HAS_X, UPC, 256>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
}

int launch_wide_ring(const WideRingPlan& p, int order, const double* x, const double* u, const double* w,
                     int64_t nsamp, int64_t nchan, int64_t ldx, int64_t ldu, const double* pivot,
                     double* ws0, double* ws1, cudaStream_t st) {
    const bool has_x = x != nullptr, upc = ldu != 0;
    int rc = TXB_OK;
    TXB_DISPATCH_ORDER(order, {
        if (has_x && !upc) rc = launch_mode<K, true, false>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
        else if (!has_x) rc = launch_mode<K, false, true>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
        else rc = launch_mode<K, true, true>(p, x, u, w, nsamp, nchan, ldx, ldu, pivot, ws0, ws1, st);
    });
    return rc;
}

}  // namespace txb

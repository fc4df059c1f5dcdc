// K6 on the device: the reference sampler's integer stream, bit-exact, never leaving HBM.
//
// The reference draws bootstrap indices with numpy:
//     rng.choice(ndat, (nrep, nsamp), replace=True)  ==  rng.integers(0, ndat, (nrep, nsamp))
// (/root/reference/src/thermoextrap/adaptive_interp.py:543-544, tests/test_data.py:98-100,
//  legacy/extrap.py:133-136; cmomy.factory_sampler does the same for {"nrep": ...} specs,
//  data.py:1782-1789).  numpy's Generator(PCG64) produces that stream as
//     next64 : s = s * M + inc (mod 2^128);  x = hi64(s) ^ lo64(s);  rotr64(x, s >> 122)
//     next32 : low half of a next64 first, then its high half (has_uint32 / uinteger buffer)
//     draw   : m = next32 * ndat;  rejected iff lo32(m) < 2^32 mod ndat;  value = hi32(m)
// so raw 32-bit draw #p depends only on jump(s0, p / 2), acceptance only on that draw, and the
// output is the stream compaction of the accepted raw draws.
//
// Device scheme (integer work; every thread block owns one contiguous range of LCG steps):
//   pass A  k_sampler<MODE_COUNT>  : accepted draws per block
//   scan    k_sampler_scan         : exclusive prefix -> first output position of each block
//   pass B  k_sampler<MODE_...>    : regenerate the range and, per accepted draw at output
//                                    position p (replicate p / nsamp):
//             FILL   indices[p] = value                         (sampler.indices)
//             HIST   freq[p / nsamp, value] += 1                (sampler.freq, any width)
//             GATHER S[p / nsamp, :] += powers(x[value], u[value])   (K2 fused, scalar x)
// Regenerating the stream costs ~15 integer instructions per draw, far below the price of
// storing and re-reading an int64 table (8 B/draw; 80 GB at BASELINE config 2).
#include "txb_common.cuh"

namespace txb {

int launch_pack_shifted(const double* x, int64_t ldx, const double* u, int64_t nsamp,
                        const double* pivot, double2* packed, cudaStream_t st);
int launch_sum_parts(const double* ws, int64_t nparts, int64_t n, double* out, cudaStream_t st);

constexpr uint64_t PCG_MULT_HI = 0x2360ED051FC65DA4ull;
constexpr uint64_t PCG_MULT_LO = 0x4385DF649FCCF645ull;
constexpr int SMP_T = 8;          // LCG steps per thread per iteration (16 raw draws)
constexpr int SMP_THREADS = 256;
constexpr int SMP_NPOW = 48;      // jump table covers step offsets < 2^48

struct Affine {  // s -> a * s + c  (mod 2^128)
    uint64_t a_hi, a_lo, c_hi, c_lo;
};

struct SamplerArgs {
    uint64_t s_hi, s_lo, inc_hi, inc_lo;
    uint32_t pre_valid, pre_u32;  // numpy's has_uint32 / uinteger
    uint32_t n, thresh;           // ndat, 2^32 mod ndat
    int64_t count;                // accepted draws wanted (nrep * nsamp)
    int64_t nsamp;                // draws per replicate
    int64_t steps_per_block;      // multiple of 512 * SMP_T (both thread geometries tile it)
    Affine iter_jump;             // (SMP_THREADS - 1) * SMP_T steps
    Affine iter_jump2;            // (BK_THREADS - 1) * SMP_T steps (512-thread kernels)
    Affine iter_jump3;            // (1024 - 1) * SMP_T / 2 steps (1024-thread partition kernel)
    Affine pow2[SMP_NPOW];        // 2^i steps
};

enum { MODE_COUNT = 0, MODE_FILL = 1, MODE_HIST = 2, MODE_GATHER = 3, MODE_LOCATE = 4 };

__device__ __forceinline__ void affine_apply(uint64_t& hi, uint64_t& lo, const Affine& f) {
    const uint64_t plo = f.a_lo * lo;
    const uint64_t phi = __umul64hi(f.a_lo, lo) + f.a_lo * hi + f.a_hi * lo;
    const uint64_t rlo = plo + f.c_lo;
    hi = phi + f.c_hi + (rlo < plo ? 1ull : 0ull);
    lo = rlo;
}

__device__ __forceinline__ uint64_t pcg_step_out(uint64_t& hi, uint64_t& lo, uint64_t inc_hi,
                                                 uint64_t inc_lo) {
    const uint64_t plo = PCG_MULT_LO * lo;
    const uint64_t phi = __umul64hi(PCG_MULT_LO, lo) + PCG_MULT_LO * hi + PCG_MULT_HI * lo;
    const uint64_t rlo = plo + inc_lo;
    hi = phi + inc_hi + (rlo < plo ? 1ull : 0ull);
    lo = rlo;
    const uint64_t x = hi ^ lo;
    const unsigned rot = (unsigned)(hi >> 58);
    return (x >> rot) | (x << ((64u - rot) & 63u));
}

__device__ __forceinline__ uint64_t pcg_out(uint64_t hi, uint64_t lo) {
    const uint64_t x = hi ^ lo;
    const unsigned rot = (unsigned)(hi >> 58);
    return (x >> rot) | (x << ((64u - rot) & 63u));
}

// exclusive block scan of one int per thread (256 threads); returns the block total in `total`
__device__ __forceinline__ int block_scan_excl(int v, int* sm_warp, int& total) {
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();  // previous users of sm_warp are done
    if (lane == 31) sm_warp[wrp] = inc;
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < SMP_THREADS / 32; ++k) {
        const int w = sm_warp[k];
        if (k < wrp) base += w;
        tot += w;
    }
    total = tot;
    return base + inc - v;
}

struct EmitArgs {
    const int64_t* block_base;  // pass B: first output position of each block
    int64_t* block_count;       // pass A: accepted draws per block
    int64_t* raw_used;          // device scalar: raw 32-bit draws consumed by `count` outputs
    int* status;                // 1 = index out of range, 2 = narrow-table overflow
    int64_t* indices;           // FILL
    void* freq;                 // HIST
    int64_t ldf;
    const double2* packed;      // GATHER: (x - cx, u - cu) pairs
    const double* w;
    double* S0;                 // [nrep, K+1]  (zeroed by the host wrapper)
    double* S1;
    int64_t* iter_base;         // per-iteration output positions (512 * SMP_T steps each)
    int* iter_count;
    int64_t pos_lo, pos_hi;     // only outputs with pos_lo <= position < pos_hi are consumed (rank window)
    int64_t rep_lo;             // first replicate of the window: output rows are relative to it
    int block_offset;           // COUNT: first block of this launch (ranks count disjoint block ranges)
};

template <typename FT>
__device__ __forceinline__ void freq_bump(FT* freq, int64_t off, int& bad) {
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
        // no return value -> RED.E.ADD (fire and forget); a saturated lane carries into its neighbour
        // (or out of the word) and is caught afterwards by k_freq_rowsum_check
        atomicAdd(word, 1u << sh);
        (void)MASK;
    }
}

// Narrow-table overflow check.  Every replicate row receives exactly `expect` increments; a
// byte / half-word lane that wraps loses 2^BITS from the row's lane sum (its carry adds at most
// 1 elsewhere), so the sum of the lanes equals `expect` iff no lane overflowed.
template <int ES>
__global__ void __launch_bounds__(1024) k_freq_rowsum_check(const unsigned char* __restrict__ freq, int64_t row_bytes,
                                                            long long expect, int* __restrict__ status,
                                                            long long* __restrict__ sums = nullptr,
                                                            int64_t len_bytes = -1) {
    // row_bytes = pitch; len_bytes = leading bytes of every row that are summed (default: the whole pitch)
    const unsigned char* row = freq + (int64_t)blockIdx.x * row_bytes;
    const int64_t len = len_bytes < 0 ? row_bytes : len_bytes;
    unsigned long long sum = 0;
    auto lanes = [](unsigned w) -> unsigned {
        return ES == 1 ? __dp4a(w, 0x01010101u, 0u) : (w & 0xffffu) + (w >> 16);
    };
    if ((reinterpret_cast<uintptr_t>(row) & 15) == 0) {
        const uint4* p = reinterpret_cast<const uint4*>(row);
        const int64_t n = len / 16;
        unsigned acc = 0;  // <= expect < 2^31 even before any wrap
        for (int64_t i = threadIdx.x; i < n; i += 1024) {
            const uint4 q = __ldcs(p + i);
            acc += lanes(q.x) + lanes(q.y) + lanes(q.z) + lanes(q.w);
        }
        for (int64_t i = n * 16 + (int64_t)threadIdx.x * ES; i < len; i += 1024 * ES)   // tail elements
            acc += ES == 1 ? (unsigned)row[i] : (unsigned)*reinterpret_cast<const unsigned short*>(row + i);
        sum = acc;
    } else {
        const unsigned* p = reinterpret_cast<const unsigned*>(row);
        const int64_t n = len / 4;
        unsigned acc = 0;
        for (int64_t i = threadIdx.x; i < n; i += 1024) acc += lanes(__ldcs(p + i));
        for (int64_t i = n * 4 + (int64_t)threadIdx.x * ES; i < len; i += 1024 * ES)
            acc += ES == 1 ? (unsigned)row[i] : (unsigned)*reinterpret_cast<const unsigned short*>(row + i);
        sum = acc;
    }
    __shared__ unsigned long long sm[32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x < 32) {
        sum = sm[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        // column-range tables (one rank's sample shard) cannot be checked locally: the caller adds the ranks' sums
        if (threadIdx.x == 0 && sums) sums[blockIdx.x] = (long long)sum;
        else if (threadIdx.x == 0 && sum != (unsigned long long)expect) atomicOr(status, 2);
    }
}

template <int K, bool HAS_X>
struct Acc {
    double s0[K + 1], s1[K + 1];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int j = 0; j <= K; ++j) s0[j] = 0.0, s1[j] = 0.0;
    }
    // p0 = weight of the draw (0 for an empty slot)
    __device__ __forceinline__ void push(double2 v, double p0) {
        double p = p0;
        s0[0] += p;
        if (HAS_X) s1[0] = fma(p, v.x, s1[0]);
#pragma unroll
        for (int j = 1; j <= K; ++j) {
            p *= v.y;
            s0[j] += p;
            if (HAS_X) s1[j] = fma(p, v.x, s1[j]);
        }
    }
    __device__ __forceinline__ void flush_thread(double* S0, double* S1, int64_t rep) {
#pragma unroll
        for (int j = 0; j <= K; ++j) {
            atomicAdd(S0 + rep * (K + 1) + j, s0[j]);
            if (HAS_X) atomicAdd(S1 + rep * (K + 1) + j, s1[j]);
        }
    }
    // all threads of the block must call this
    __device__ __forceinline__ void flush_block(double (*sm)[2 * (K + 1)], double* S0, double* S1,
                                                int64_t rep) {
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
        __syncthreads();
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
            for (int k = 0; k < SMP_THREADS / 32; ++k) t += sm[k][threadIdx.x];
            if (threadIdx.x <= K)
                atomicAdd(S0 + rep * (K + 1) + threadIdx.x, t);
            else if (HAS_X)
                atomicAdd(S1 + rep * (K + 1) + threadIdx.x - (K + 1), t);
        }
    }
};

template <int MODE, int K, bool HAS_X, bool HAS_W, typename FT>
__global__ void __launch_bounds__(SMP_THREADS, 2)
    k_sampler(const __grid_constant__ SamplerArgs a, const EmitArgs e) {
    constexpr int T = SMP_T;
    constexpr int NSLOT = 2 * T;
    constexpr int STAGE = SMP_THREADS * NSLOT + 1;  // + the buffered 32-bit half
    __shared__ int sm_warp[SMP_THREADS / 32];
    __shared__ uint32_t sm_val[MODE == MODE_COUNT ? 1 : STAGE];
    __shared__ double sm_red[MODE == MODE_GATHER ? SMP_THREADS / 32 : 1][2 * (K + 1)];

    const int bid = blockIdx.x + (MODE == MODE_COUNT ? e.block_offset : 0);
    int64_t P = 0;
    if (MODE != MODE_COUNT) {
        P = e.block_base[bid];
        if (P >= a.count) return;
        const int64_t P_end = P + e.block_count[bid];
        // LOCATE: only the block that emits output #count-1 walks its range (to report raw_used)
        if (MODE == MODE_LOCATE && P_end < a.count) return;
        if (MODE != MODE_LOCATE && (P >= e.pos_hi || P_end <= e.pos_lo)) return;  // outside the window
    }
    const int64_t q_block = (int64_t)bid * a.steps_per_block;
    uint64_t hi = a.s_hi, lo = a.s_lo;
    {
        uint64_t q = (uint64_t)q_block + (uint64_t)threadIdx.x * T;
        for (int i = 0; q != 0 && i < SMP_NPOW; ++i, q >>= 1)
            if (q & 1) affine_apply(hi, lo, a.pow2[i]);
    }
    const int64_t niter = a.steps_per_block / (SMP_THREADS * T);
    const uint32_t n = a.n, thresh = a.thresh;
    long long cnt_total = 0;
    int bad = 0;

    // GATHER state
    Acc<K, HAS_X> acc;
    acc.zero();
    bool dirty = false;
    int64_t cur_rep = 0, next_boundary = 0;
    if (MODE == MODE_HIST || MODE == MODE_GATHER) {
        cur_rep = P / a.nsamp;
        next_boundary = (cur_rep + 1) * a.nsamp;
    }

    for (int64_t it = 0; it < niter; ++it) {
        uint32_t val[NSLOT];
        uint32_t mask = 0;
#pragma unroll
        for (int k = 0; k < T; ++k) {
            const uint64_t o = pcg_step_out(hi, lo, a.inc_hi, a.inc_lo);
            const uint64_t m0 = (uint64_t)(uint32_t)o * n;
            const uint64_t m1 = (o >> 32) * (uint64_t)n;
            val[2 * k] = (uint32_t)(m0 >> 32);
            val[2 * k + 1] = (uint32_t)(m1 >> 32);
            if ((uint32_t)m0 >= thresh) mask |= 1u << (2 * k);
            if ((uint32_t)m1 >= thresh) mask |= 1u << (2 * k + 1);
        }
        const bool pre = it == 0 && bid == 0 && threadIdx.x == 0 && a.pre_valid != 0;
        uint32_t pre_val = 0;
        bool pre_acc = false;
        if (pre) {
            const uint64_t m = (uint64_t)a.pre_u32 * n;
            pre_val = (uint32_t)(m >> 32);
            pre_acc = (uint32_t)m >= thresh;
        }
        const int c = __popc(mask) + (pre_acc ? 1 : 0);
        if (MODE == MODE_COUNT) {
            cnt_total += c;
            affine_apply(hi, lo, a.iter_jump);
            continue;
        }

        int total;
        int off = block_scan_excl(c, sm_warp, total);  // (syncs: previous readers of sm_val done)
        {
            // raw 32-bit draw number of slot 0 of this thread (0-based, the buffered half is #0)
            const int64_t q0 = q_block + (it * SMP_THREADS + threadIdx.x) * T;
            const int64_t raw0 = (int64_t)a.pre_valid + 2 * q0;
            if (pre_acc) {
                if (P + off == a.count - 1) *e.raw_used = 1;
                sm_val[off++] = pre_val;
            }
#pragma unroll
            for (int k = 0; k < NSLOT; ++k) {
                if (mask & (1u << k)) {
                    if (P + off == a.count - 1) *e.raw_used = raw0 + k + 1;
                    sm_val[off++] = val[k];
                }
            }
        }
        __syncthreads();
        // entries [nlow, nvalid) of this iteration lie inside [pos_lo, min(pos_hi, count))
        const int64_t limit = (e.pos_hi < a.count ? e.pos_hi : a.count) - P;
        const int nvalid = limit <= 0 ? 0 : ((int64_t)total < limit ? total : (int)limit);
        const int64_t below = e.pos_lo - P;
        const int nlow = below <= 0 ? 0 : (below < (int64_t)nvalid ? (int)below : nvalid);

        if (MODE == MODE_FILL) {
            for (int i = nlow + threadIdx.x; i < nvalid; i += SMP_THREADS)
                e.indices[P + i - e.pos_lo] = (int64_t)sm_val[i];
        } else if (MODE == MODE_HIST) {
            FT* freq = reinterpret_cast<FT*>(e.freq);
            if (P + nvalid <= next_boundary) {
                const int64_t row = (cur_rep - e.rep_lo) * e.ldf;
                for (int i = nlow + threadIdx.x; i < nvalid; i += SMP_THREADS) freq_bump<FT>(freq, row + sm_val[i], bad);
            } else {
                for (int i = nlow + threadIdx.x; i < nvalid; i += SMP_THREADS)
                    freq_bump<FT>(freq, ((P + i) / a.nsamp - e.rep_lo) * e.ldf + sm_val[i], bad);
            }
        } else if (MODE == MODE_GATHER) {
            if (P + nvalid <= next_boundary) {
                // fast path: every draw of this iteration belongs to replicate cur_rep
                constexpr int G = 8;
                for (int i0 = nlow + threadIdx.x; i0 < nvalid; i0 += G * SMP_THREADS) {
                    double2 v[G];
                    double pw[G];
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        const int i = i0 + g * SMP_THREADS;
                        const bool ok = i < nvalid;
                        const uint32_t s = ok ? sm_val[i] : 0u;
                        v[g] = ok ? __ldg(e.packed + s) : make_double2(0.0, 0.0);
                        pw[g] = ok ? (HAS_W ? __ldg(e.w + s) : 1.0) : 0.0;
                    }
#pragma unroll
                    for (int g = 0; g < G; ++g) acc.push(v[g], pw[g]);
                }
                dirty = dirty || nvalid > nlow;
            } else {
                // replicate boundary inside this iteration: thread-local tagged accumulation
                if (dirty) {
                    acc.flush_block(sm_red, e.S0, e.S1, cur_rep - e.rep_lo);
                    acc.zero();
                    dirty = false;
                }
                const int per = (nvalid - nlow + SMP_THREADS - 1) / SMP_THREADS;
                const int i_lo = nlow + threadIdx.x * per;
                const int i_hi = min(nvalid, i_lo + per);
                int64_t tag = -1;
                for (int i = i_lo; i < i_hi; ++i) {
                    const int64_t rep = (P + i) / a.nsamp;
                    if (rep != tag) {
                        if (tag >= 0) {
                            acc.flush_thread(e.S0, e.S1, tag - e.rep_lo);
                            acc.zero();
                        }
                        tag = rep;
                    }
                    const uint32_t s = sm_val[i];
                    acc.push(__ldg(e.packed + s), HAS_W ? __ldg(e.w + s) : 1.0);
                }
                if (tag >= 0) {
                    acc.flush_thread(e.S0, e.S1, tag - e.rep_lo);
                    acc.zero();
                }
            }
        }

        P += total;
        if (P >= a.count || (MODE != MODE_LOCATE && P >= e.pos_hi)) break;
        if (MODE == MODE_HIST || MODE == MODE_GATHER) {
            if (P >= next_boundary) {
                if (MODE == MODE_GATHER && dirty) {
                    acc.flush_block(sm_red, e.S0, e.S1, cur_rep - e.rep_lo);
                    acc.zero();
                    dirty = false;
                }
                while (P >= next_boundary) ++cur_rep, next_boundary += a.nsamp;
            }
        }
        affine_apply(hi, lo, a.iter_jump);
    }

    if (MODE == MODE_COUNT) {
        __shared__ long long sm_cnt[SMP_THREADS / 32];
        long long v = cnt_total;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) sm_cnt[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            long long t = 0;
            for (int k = 0; k < SMP_THREADS / 32; ++k) t += sm_cnt[k];
            e.block_count[bid] = t;
        }
    }
    if (MODE == MODE_GATHER && dirty) acc.flush_block(sm_red, e.S0, e.S1, cur_rep - e.rep_lo);
    if (MODE == MODE_HIST && bad && e.status) atomicOr(e.status, bad);
}

// pass A (512 threads): accepted draws per iteration (512 * SMP_T LCG steps) and per block
__global__ void __launch_bounds__(512, 2)
    k_sampler_count(const __grid_constant__ SamplerArgs a, int block_offset, int64_t* __restrict__ block_count,
                    int* __restrict__ iter_count) {
    constexpr int T = SMP_T;
    constexpr int NT = 512;
    __shared__ int sm_c[2][NT / 32];
    const int bid = blockIdx.x + block_offset;
    const int64_t q_block = (int64_t)bid * a.steps_per_block;
    uint64_t hi = a.s_hi, lo = a.s_lo;
    {
        uint64_t q = (uint64_t)q_block + (uint64_t)threadIdx.x * T;
        for (int i = 0; q != 0 && i < SMP_NPOW; ++i, q >>= 1)
            if (q & 1) affine_apply(hi, lo, a.pow2[i]);
    }
    const int64_t niter = a.steps_per_block / (NT * T);
    const uint32_t n = a.n, thresh = a.thresh;
    long long total = 0;
    for (int64_t it = 0; it < niter; ++it) {
        int c = 0;
#pragma unroll
        for (int k = 0; k < T; ++k) {
            const uint64_t o = pcg_step_out(hi, lo, a.inc_hi, a.inc_lo);
            c += (uint32_t)((uint64_t)(uint32_t)o * n) >= thresh ? 1 : 0;
            c += (uint32_t)((o >> 32) * (uint64_t)n) >= thresh ? 1 : 0;
        }
        if (it == 0 && bid == 0 && threadIdx.x == 0 && a.pre_valid != 0)
            c += (uint32_t)((uint64_t)a.pre_u32 * n) >= thresh ? 1 : 0;
        c = __reduce_add_sync(0xffffffffu, c);
        int* buf = sm_c[it & 1];
        if ((threadIdx.x & 31) == 0) buf[threadIdx.x >> 5] = c;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
#pragma unroll
            for (int k = 0; k < NT / 32; ++k) t += buf[k];
            iter_count[(int64_t)bid * niter + it] = t;
            total += t;
        }
        affine_apply(hi, lo, a.iter_jump2);
    }
    if (threadIdx.x == 0) block_count[bid] = total;
}

// iter_base[j] = output position of the first draw of iteration j (one warp per block)
__global__ void k_sampler_iter_scan(const int64_t* __restrict__ block_base, const int* __restrict__ iter_count,
                                    int64_t* __restrict__ iter_base, int nblocks, int64_t niter) {
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= nblocks) return;
    int64_t run = block_base[b];
    for (int64_t i0 = 0; i0 < niter; i0 += 32) {
        const int64_t i = i0 + lane;
        const int v = i < niter ? iter_count[(int64_t)b * niter + i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (i < niter) iter_base[(int64_t)b * niter + i] = run + inc - v;
        run += __shfl_sync(0xffffffffu, inc, 31);
    }
}

// exclusive scan of the per-block counts (a few thousand entries); flags a short raw range
__global__ void k_sampler_scan(const int64_t* __restrict__ cnt, int64_t* __restrict__ base, int nblocks,
                               int64_t count, int* __restrict__ status) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int64_t run = 0;
    for (int b = 0; b < nblocks; ++b) {
        const int64_t c = cnt[b];
        base[b] = run;
        run += c;
    }
    if (run < count && status) atomicOr(status, 4);
}

__global__ void k_status_or(int* status, int bits) { atomicOr(status, bits); }

}  // namespace txb

#include "sampler_bucket.cuh"

// ------------------------------------------------------------------ host side
using namespace txb;
typedef unsigned __int128 u128;

namespace {

constexpr size_t SMP_WS_HEAD = 64 * 1024;  // block counts + bases + scalars

inline u128 mk128(uint64_t hi, uint64_t lo) { return ((u128)hi << 64) | lo; }
inline Affine mk_affine(u128 a, u128 c) {
    return Affine{(uint64_t)(a >> 64), (uint64_t)a, (uint64_t)(c >> 64), (uint64_t)c};
}

// threads of the partition kernel: 512 unless TXB_SCATTER_THREADS=1024 / txb_set_scatter_threads(1024).  Measured
// (profiles/r02_exp_mma_forms.json): 512 threads 100.5 ms, 1024 threads 102.5 ms per 128-replicate window of 1e8 draws;
// ncu (profiles/r02_target_kernels_ncu_full.txt): the kernel is instruction-bound (ALU 35 % + IMAD 17 % of the pipes,
// 44 % issue-active), so twice the warps with 18 % more instructions does not pay.
int g_scatter_threads = -1;
inline int scatter_threads() {
    if (g_scatter_threads < 0) {
        const char* e = getenv("TXB_SCATTER_THREADS");
        g_scatter_threads = (e && atoi(e) == 1024) ? 1024 : 512;
    }
    return g_scatter_threads;
}

// per-iteration drain of the partition kernel: the compacted per-warp flush unless TXB_SCATTER_DRAIN=0 (the
// per-bucket walk of round 1).  Measured (profiles/r02_exp_drain.json): 91.1 -> 84.8 ms per 256 replicates of 1e8
// draws, config 2's partition 32.9 -> 30.2 ms; same lists, bit-identical tables.
inline int scatter_compact_drain() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("TXB_SCATTER_DRAIN");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v;
}

struct Plan {
    SamplerArgs a;
    int nblocks;
    int64_t ipb;        // iterations (512 * SMP_T steps) per block
    size_t head_bytes;  // scalars | block_count | block_base | iter_base | iter_count
};

// (a, c) of `delta` LCG steps
void lcg_jump(u128 inc, uint64_t delta, u128& a_out, u128& c_out) {
    u128 acc_a = 1, acc_c = 0, cur_a = mk128(PCG_MULT_HI, PCG_MULT_LO), cur_c = inc;
    while (delta) {
        if (delta & 1) {
            acc_a = acc_a * cur_a;
            acc_c = acc_c * cur_a + cur_c;
        }
        cur_c = (cur_a + 1) * cur_c;
        cur_a = cur_a * cur_a;
        delta >>= 1;
    }
    a_out = acc_a;
    c_out = acc_c;
}

int make_plan(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t nsamp, Plan& p) {
    TXB_CHECK_ARG(rng != nullptr, "sampler: rng state is NULL");
    // ndat == 1 consumes no draws in numpy (rng == 0 shortcut); the Python layer handles it
    TXB_CHECK_ARG(ndat >= 2 && ndat <= 0xFFFFFFFFll,
                  "sampler: ndat=%lld outside [2, 2^32-1] (numpy's 32-bit Lemire path)", (long long)ndat);
    TXB_CHECK_ARG(count >= 1 && nsamp >= 1, "sampler: count=%lld nsamp=%lld must be >= 1",
                  (long long)count, (long long)nsamp);
    SamplerArgs& a = p.a;
    a.s_hi = rng->state_hi, a.s_lo = rng->state_lo, a.inc_hi = rng->inc_hi, a.inc_lo = rng->inc_lo;
    a.pre_valid = rng->has_uint32 ? 1u : 0u;
    a.pre_u32 = rng->uinteger;
    a.n = (uint32_t)ndat;
    a.thresh = (uint32_t)((0x100000000ull - (uint64_t)ndat) % (uint64_t)ndat);
    a.count = count;
    a.nsamp = nsamp;
    // raw draws needed: mean count / (1 - p) plus a > 12 sigma margin
    const double prej = (double)a.thresh / 4294967296.0;
    const double want = ((double)count + 12.0 * sqrt((double)count) + 4096.0) / (1.0 - prej);
    TXB_CHECK_ARG(want < 1.0e14, "sampler: %lld draws is beyond the jump table", (long long)count);
    const int64_t raw = (int64_t)want + 1;
    const int64_t steps = (raw + 1) / 2;
    const int64_t per_iter = (int64_t)BK_THREADS * SMP_T;
    const int64_t iters = (steps + per_iter - 1) / per_iter;
    int64_t nb = (int64_t)sm_count() * 8;
    if (nb > iters) nb = iters;
    const int64_t ipb = (iters + nb - 1) / nb;
    nb = (iters + ipb - 1) / ipb;
    p.nblocks = (int)nb;
    a.steps_per_block = ipb * per_iter;
    const u128 inc = mk128(rng->inc_hi, rng->inc_lo);
    u128 ja, jc;
    lcg_jump(inc, (uint64_t)(SMP_THREADS - 1) * SMP_T, ja, jc);
    a.iter_jump = mk_affine(ja, jc);
    lcg_jump(inc, (uint64_t)(BK_THREADS - 1) * SMP_T, ja, jc);
    a.iter_jump2 = mk_affine(ja, jc);
    lcg_jump(inc, (uint64_t)(2 * BK_THREADS - 1) * (SMP_T / 2), ja, jc);
    a.iter_jump3 = mk_affine(ja, jc);
    u128 ca = mk128(PCG_MULT_HI, PCG_MULT_LO), cc = inc;
    for (int i = 0; i < SMP_NPOW; ++i) {
        a.pow2[i] = mk_affine(ca, cc);
        cc = (ca + 1) * cc;
        ca = ca * ca;
    }
    TXB_CHECK_ARG((size_t)p.nblocks * 16 + 64 <= SMP_WS_HEAD, "sampler: too many blocks");
    p.ipb = ipb;
    p.head_bytes = align_up(SMP_WS_HEAD + (size_t)(nb * ipb) * 12 + 64, 256);
    return TXB_OK;
}

struct Head {
    int64_t* block_count;
    int64_t* block_base;
    int64_t* raw_used;
    int64_t* iter_base;
    int* iter_count;
};

Head carve_head(void* ws, const Plan& pl) {
    int64_t* p = reinterpret_cast<int64_t*>(ws);
    int64_t* ib = reinterpret_cast<int64_t*>(reinterpret_cast<char*>(ws) + SMP_WS_HEAD);
    return Head{p + 8, p + 8 + pl.nblocks, p, ib, reinterpret_cast<int*>(ib + (int64_t)pl.nblocks * pl.ipb)};
}

// Workspace head (first txb_sampler_workspace_bytes(...) bytes):
//   int64 [0] raw_used scratch, [1..7] reserved, [8 .. 8+nblocks) block_count, [8+nblocks .. 8+2 nblocks) block_base
//   at byte 65536: int64 iter_base[nblocks * ipb], then int32 iter_count[nblocks * ipb]
void bind_head(const Plan& p, EmitArgs& e, void* ws, int64_t* raw_used, int* status) {
    Head h = carve_head(ws, p);
    e.block_count = h.block_count;
    e.block_base = h.block_base;
    e.iter_base = h.iter_base;
    e.iter_count = h.iter_count;
    e.raw_used = raw_used ? raw_used : h.raw_used;
    e.status = status;
    e.pos_lo = 0;
    e.pos_hi = p.a.count;
    e.rep_lo = 0;
    e.block_offset = 0;
}

// pass A over blocks [block_lo, block_hi); the other entries of block_count are zeroed so that
// ranks counting disjoint ranges can combine their heads with a plain integer SUM
int run_count(const Plan& p, EmitArgs& e, int block_lo, int block_hi, cudaStream_t st) {
    if (cudaMemsetAsync(e.block_count, 0, sizeof(int64_t) * p.nblocks, st) != cudaSuccess ||
        cudaMemsetAsync(e.iter_count, 0, sizeof(int) * (size_t)p.nblocks * p.ipb, st) != cudaSuccess) {
        set_error("sampler: cudaMemsetAsync failed");
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    if (block_hi <= block_lo) return TXB_OK;
    {
        ProfScope prof("sampler_count", st);
        k_sampler_count<<<block_hi - block_lo, 512, 0, st>>>(p.a, block_lo, e.block_count, e.iter_count);
    }
    TXB_CUDA_LAUNCH_CHECK("k_sampler_count");
    return TXB_OK;
}

// scan of block_count -> block_base, then find the raw position of output #count-1
int run_scan(const Plan& p, EmitArgs& e, cudaStream_t st) {
    if (cudaMemsetAsync(e.raw_used, 0, sizeof(int64_t), st) != cudaSuccess ||
        (e.status && cudaMemsetAsync(e.status, 0, sizeof(int), st) != cudaSuccess)) {
        set_error("sampler: cudaMemsetAsync failed");
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    k_sampler_scan<<<1, 32, 0, st>>>(e.block_count, const_cast<int64_t*>(e.block_base), p.nblocks, p.a.count,
                                     e.status);
    TXB_CUDA_LAUNCH_CHECK("k_sampler_scan");
    k_sampler_iter_scan<<<(p.nblocks + 7) / 8, 256, 0, st>>>(e.block_base, e.iter_count, e.iter_base, p.nblocks, p.ipb);
    TXB_CUDA_LAUNCH_CHECK("k_sampler_iter_scan");
    k_sampler_locate<<<1, BK_THREADS, 0, st>>>(p.a, e.iter_base, (int64_t)p.nblocks * p.ipb, e.raw_used);
    TXB_CUDA_LAUNCH_CHECK("k_sampler_locate");
    return TXB_OK;
}

// everything an emit call needs before its pass B
int begin_emit(const Plan& p, EmitArgs& e, void* ws, int64_t* raw_used, int* status, int prepared,
               cudaStream_t st) {
    bind_head(p, e, ws, raw_used, status);
    if (prepared) {
        if (status && cudaMemsetAsync(status, 0, sizeof(int), st) != cudaSuccess) {
            set_error("sampler: cudaMemsetAsync failed");
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        return TXB_OK;
    }
    int rc = run_count(p, e, 0, p.nblocks, st);
    if (rc != TXB_OK) return rc;
    return run_scan(p, e, st);
}

int check_head(const char* what, void* workspace, size_t ws_bytes, const Plan& p) {
    TXB_CHECK_ARG(workspace != nullptr, "%s: null workspace", what);
    if (ws_bytes < p.head_bytes) {
        set_error("%s: workspace too small", what);
        return TXB_ERR_WORKSPACE;
    }
    return TXB_OK;
}

}  // namespace

static size_t head_bytes_for(int64_t ndat, int64_t count) {
    const txb_pcg64 dummy = {0, 0, 0, 1, 0, 0};
    Plan p;
    return make_plan(&dummy, ndat, count, count, p) == TXB_OK ? p.head_bytes : 0;
}

extern "C" size_t txb_sampler_workspace_bytes(int64_t ndat, int64_t count) { return head_bytes_for(ndat, count); }

extern "C" int txb_sampler_layout(int64_t ndat, int64_t count, int64_t* out) {
    const txb_pcg64 dummy = {0, 0, 0, 1, 0, 0};
    Plan p;
    int rc = make_plan(&dummy, ndat, count, count, p);
    if (rc != TXB_OK) return rc;
    TXB_CHECK_ARG(out != nullptr, "txb_sampler_layout: null pointer");
    out[0] = p.nblocks;                      // entries of block_count (int64, byte offset 64)
    out[1] = (int64_t)p.nblocks * p.ipb;     // entries of iter_count (int32)
    out[2] = (int64_t)SMP_WS_HEAD + 8 * out[1];  // byte offset of iter_count
    out[3] = (int64_t)p.head_bytes;
    return TXB_OK;
}

extern "C" int txb_sampler_nblocks(const txb_pcg64* rng, int64_t ndat, int64_t count) {
    Plan p;
    return make_plan(rng, ndat, count, count, p) == TXB_OK ? p.nblocks : -1;
}

extern "C" int txb_sampler_count(const txb_pcg64* rng, int64_t ndat, int64_t count, int block_lo,
                                 int block_hi, void* workspace, size_t ws_bytes, void* stream) {
    Plan p;
    int rc = make_plan(rng, ndat, count, count, p);
    if (rc != TXB_OK) return rc;
    rc = check_head("txb_sampler_count", workspace, ws_bytes, p);
    if (rc != TXB_OK) return rc;
    if (block_hi < 0 || block_hi > p.nblocks) block_hi = p.nblocks;
    TXB_CHECK_ARG(block_lo >= 0 && block_lo <= block_hi, "txb_sampler_count: bad block range");
    EmitArgs e{};
    bind_head(p, e, workspace, nullptr, nullptr);
    return run_count(p, e, block_lo, block_hi, (cudaStream_t)stream);
}

extern "C" int txb_sampler_scan(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t* raw_used,
                                int* status, void* workspace, size_t ws_bytes, void* stream) {
    Plan p;
    int rc = make_plan(rng, ndat, count, count, p);
    if (rc != TXB_OK) return rc;
    rc = check_head("txb_sampler_scan", workspace, ws_bytes, p);
    if (rc != TXB_OK) return rc;
    EmitArgs e{};
    bind_head(p, e, workspace, raw_used, status);
    return run_scan(p, e, (cudaStream_t)stream);
}

extern "C" int txb_sampler_prepare(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t* raw_used,
                                   int* status, void* workspace, size_t ws_bytes, void* stream) {
    Plan p;
    int rc = make_plan(rng, ndat, count, count, p);
    if (rc != TXB_OK) return rc;
    rc = check_head("txb_sampler_prepare", workspace, ws_bytes, p);
    if (rc != TXB_OK) return rc;
    EmitArgs e{};
    return begin_emit(p, e, workspace, raw_used, status, 0, (cudaStream_t)stream);
}
extern "C" int txb_pcg64_advance(txb_pcg64* rng, int64_t raw_used) {
    TXB_CHECK_ARG(rng && raw_used >= 0, "txb_pcg64_advance: bad argument");
    int64_t r = raw_used;
    if (r == 0) return TXB_OK;
    if (rng->has_uint32) {
        rng->has_uint32 = 0;  // numpy leaves the stale `uinteger` in place
        --r;
        if (r == 0) return TXB_OK;
    }
    const uint64_t steps = (uint64_t)(r + 1) / 2;
    u128 ja, jc;
    lcg_jump(mk128(rng->inc_hi, rng->inc_lo), steps, ja, jc);
    const u128 s = ja * mk128(rng->state_hi, rng->state_lo) + jc;
    rng->state_hi = (uint64_t)(s >> 64);
    rng->state_lo = (uint64_t)s;
    {  // the high half of the last 64-bit output is buffered (still pending when r is odd)
        const uint64_t x = rng->state_hi ^ rng->state_lo;
        const unsigned rot = (unsigned)(rng->state_hi >> 58);
        const uint64_t o = (x >> rot) | (x << ((64u - rot) & 63u));
        rng->has_uint32 = (uint32_t)(r & 1);
        rng->uinteger = (uint32_t)(o >> 32);
    }
    return TXB_OK;
}

extern "C" int txb_sampler_indices(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t pos_lo,
                                   int64_t pos_hi, int64_t* indices, int64_t* raw_used, int* status,
                                   int prepared, void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(indices != nullptr, "txb_sampler_indices: null pointer");
    TXB_CHECK_ARG(0 <= pos_lo && pos_lo < pos_hi && pos_hi <= count,
                  "txb_sampler_indices: bad window [%lld, %lld) of %lld", (long long)pos_lo, (long long)pos_hi,
                  (long long)count);
    Plan p;
    int rc = make_plan(rng, ndat, count, count, p);
    if (rc != TXB_OK) return rc;
    rc = check_head("txb_sampler_indices", workspace, ws_bytes, p);
    if (rc != TXB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    EmitArgs e{};
    rc = begin_emit(p, e, workspace, raw_used, status, prepared, st);
    if (rc != TXB_OK) return rc;
    e.indices = indices;
    e.pos_lo = pos_lo;
    e.pos_hi = pos_hi;
    k_sampler<MODE_FILL, 1, false, false, int64_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e);
    TXB_CUDA_LAUNCH_CHECK("k_sampler<fill>");
    return TXB_OK;
}

namespace {

// geometry of the bucketed path for one call
struct BucketPlan {
    bool ok;
    int nb, cap;
    double m;              // expected entries of one (replicate, bucket) list
    size_t fixed_bytes;    // everything except the per-replicate part
};

// segments per replicate, draws per segment and list capacity for a window of `nwin` replicates
struct SegPlan {
    int nseg, cap_lines;
    int64_t seg_len;
    size_t per_rep_bytes;  // line counts + partials + lists for one replicate
};

SegPlan seg_plan(const BucketPlan& b, int64_t nsamp, int ks, int64_t nwin) {
    SegPlan s{};
    // one CTA per SM at a time (the rings fill shared memory): choose the number of segments per replicate so
    // that the CTA count fills whole waves.  ncu of round 1's rule (>= 4 CTAs per SM overall) showed 640 CTAs
    // for a 128-replicate window on 148 SMs: 4.3 waves run as 5, 14 % of the kernel idle.  Candidates between
    // ~8 and ~40 waves (startup -- binary search, 48-step jump, ring clear -- stays < 1 % of a CTA), most
    // efficient first, fewer segments on ties.
    const int64_t sms = sm_count();
    const int64_t max_seg = (int64_t)(b.m / (8.0 * BK_LINE));     // keep lists mostly full lines
    // histogram consumer (ks < 0): 8 .. 40 waves.  Comoment consumer: every extra segment is one more list (and
    // one more padded last line) per (replicate, bucket) for k_bucket_consume -- measured +1.7 ms per 5 segments
    // at config 2 -- so stay within [1x, 2x] of round 1's rule (>= 4 CTAs per SM overall)
    int64_t lo_seg = ((ks < 0 ? 8 : 4) * sms + nwin - 1) / nwin, hi_seg = ((ks < 0 ? 40 : 8) * sms + nwin - 1) / nwin;
    if (lo_seg < 1) lo_seg = 1;
    if (hi_seg > max_seg) hi_seg = max_seg;
    if (hi_seg > 255) hi_seg = 255;   // k_bucket_hist keeps the segment sizes in a 256-entry shared array
    if (lo_seg > hi_seg) lo_seg = hi_seg;
    if (lo_seg < 1) lo_seg = hi_seg = 1;
    int64_t nseg = lo_seg;
    double best = -1.0;
    for (int64_t c = lo_seg; c <= hi_seg; ++c) {
        const int64_t ctas = c * nwin, waves = (ctas + sms - 1) / sms;
        const double eff = (double)ctas / (double)(waves * sms);
        if (eff > best + 1e-9) best = eff, nseg = c;
    }
    s.nseg = (int)nseg;
    s.seg_len = (nsamp + nseg - 1) / nseg;
    const double ms = b.m * (double)s.seg_len / (double)nsamp;
    s.cap_lines = (int)((ms + 10.0 * sqrt(ms) + 32.0) / BK_LINE) + 3;
    s.per_rep_bytes = (size_t)b.nb * (sizeof(unsigned) * s.nseg + 2 * sizeof(double) * (ks + 1) +
                                      (size_t)s.nseg * s.cap_lines * BK_LINE * 2) + 64 + 16;   // + a row-sum accumulator
    return s;
}

BucketPlan bucket_plan(const Plan& p, int64_t ndat, int64_t nsamp, int64_t nrep_out, int order) {
    BucketPlan b{};
    b.nb = (int)((ndat + BK_SIZE - 1) / BK_SIZE);
    b.m = ndat <= BK_SIZE ? (double)nsamp : (double)nsamp * BK_SIZE / (double)ndat;
    // ring: up to 31 pending entries + one iteration's arrivals (Poisson, mean 8192 / nb).  The fast
    // append wraps (and flags the call for a redo on the gather path) when arrivals > cap - pending:
    // choose the power of two that makes this a < 1e-12 event per bucket and iteration.
    const double arrivals = (double)(BK_THREADS * 2 * SMP_T) / b.nb;
    b.cap = BK_MINCAP;
    for (;; b.cap *= 2) {
        double term = exp(-arrivals), cdf = 0.0;  // P(X <= cap - 31)
        for (int k = 0; k <= b.cap - 31; ++k) {
            cdf += term;
            term *= arrivals / (k + 1);
        }
        if (1.0 - cdf < 1e-12 || b.cap >= 4096) break;
    }
    // worth it only for tables too big for L2-resident gathers, with mostly full lines, and when the
    // rings fit in shared memory (a window of bucket counts around 800-1170 does not: gather path)
    b.ok = ndat >= (int64_t)4 << 20 && b.nb < BK_MAXNB && b.m >= 8.0 * BK_LINE &&
           (size_t)(b.nb + 1) * (b.cap * 2 + 12) <= (size_t)BK_MAXNB * (BK_MINCAP * 2 + 12);
    b.fixed_bytes = p.head_bytes + 256 + sizeof(double2) * (size_t)ndat +
                    sizeof(double) * (8 + 2 * (size_t)nrep_out * (order + 2));
    return b;
}

// geometry of the bucketed histogram (txb_sampler_freq): 65520-sample buckets
BucketPlan hist_bucket_plan(const Plan& p, int64_t ndat, int64_t nsamp) {
    BucketPlan b{};
    b.nb = (int)((ndat + BK_HSIZE - 1) / BK_HSIZE);
    b.m = (double)nsamp * BK_HSIZE / (double)ndat;
    const double arrivals = (double)(BK_THREADS * 2 * SMP_T) / b.nb;
    b.cap = BK_MINCAP;
    for (;; b.cap *= 2) {  // same ring sizing rule as bucket_plan
        double term = exp(-arrivals), cdf = 0.0;
        for (int k = 0; k <= b.cap - 31; ++k) {
            cdf += term;
            term *= arrivals / (k + 1);
        }
        if (1.0 - cdf < 1e-12 || b.cap >= 4096) break;
    }
    b.ok = b.nb >= 8 && b.nb < BK_MAXNB && b.m >= 8.0 * BK_LINE && b.cap < 4096 &&
           (size_t)(b.nb + 1) * (b.cap * 2 + 12) <= (size_t)BK_MAXNB * (BK_MINCAP * 2 + 12);
    b.fixed_bytes = p.head_bytes + 256;
    return b;
}

template <int K>
int launch_consume(const ConsumeArgs& c, bool has_x, bool has_w, int gy, cudaStream_t st) {
    const size_t smem = sizeof(double2) * BK_SIZE + (has_w ? sizeof(double) * BK_SIZE : 0);
    const dim3 grid((unsigned)c.nb, (unsigned)gy);
#define TXB_CONSUME(HX, HW)                                                                                  \
    {                                                                                                        \
        static bool done = false;                                                                            \
        if (!done) {                                                                                         \
            if (cudaFuncSetAttribute(k_bucket_consume<K, HX, HW>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                     (int)(sizeof(double2) * BK_SIZE + sizeof(double) * BK_SIZE)) != cudaSuccess) { \
                set_error("k_bucket_consume: shared memory opt-in failed");                                  \
                cudaGetLastError();                                                                          \
                return TXB_ERR_CUDA;                                                                         \
            }                                                                                                \
            done = true;                                                                                     \
        }                                                                                                    \
        k_bucket_consume<K, HX, HW><<<grid, BK_THREADS, smem, st>>>(c);                                      \
    }
    if (has_x) {
        if (has_w) TXB_CONSUME(true, true) else TXB_CONSUME(true, false)
    } else {
        if (has_w) TXB_CONSUME(false, true) else TXB_CONSUME(false, false)
    }
#undef TXB_CONSUME
    TXB_CUDA_LAUNCH_CHECK("k_bucket_consume");
    return TXB_OK;
}

}  // namespace

extern "C" int txb_set_scatter_threads(int nthreads) {
    const int prev = scatter_threads();
    if (nthreads == 512 || nthreads == 1024) g_scatter_threads = nthreads;
    return prev;
}

extern "C" size_t txb_sampler_freq_workspace_bytes(int64_t ndat, int64_t nrep, int64_t nsamp, int64_t nrep_out) {
    const txb_pcg64 dummy = {0, 0, 0, 1, 0, 0};
    Plan p;
    if (make_plan(&dummy, ndat, nrep * nsamp, nsamp, p) != TXB_OK) return 0;
    const BucketPlan b = hist_bucket_plan(p, ndat, nsamp);
    if (!b.ok) return align_up(p.head_bytes + 256, 256);
    const SegPlan sp = seg_plan(b, nsamp, -1, nrep_out);
    return align_up(align_up(p.head_bytes + 256, 256) + sp.per_rep_bytes * (size_t)nrep_out + 8192, 256);
}

// Shared body of txb_sampler_freq (all columns) and txb_sampler_freq_cols (columns [col_lo, col_hi) only:
// the frequency columns of one rank's sample shard; `freq` then holds col_hi - col_lo columns per row and
// `rowsum` receives the row sums, which the caller adds over ranks to redo the overflow check).
static int sampler_freq_impl(const txb_pcg64* rng, int64_t ndat, int64_t nrep, int64_t nsamp,
                             int64_t rep_lo, int64_t rep_hi, int64_t col_lo, int64_t col_hi, void* freq,
                             int freq_dtype, int64_t ldf, int64_t* raw_used, long long* rowsum, int64_t sum_cols,
                             int* status, int prepared, void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(freq != nullptr, "txb_sampler_freq: null pointer");
    const bool partial = !(col_lo == 0 && col_hi == ndat);
    const int64_t ncol = col_hi - col_lo;
    TXB_CHECK_ARG(0 <= col_lo && col_lo < col_hi && col_hi <= ndat && col_lo % BK_HSIZE == 0,
                  "txb_sampler_freq: column range [%lld, %lld) must lie in [0, ndat) and start at a multiple of %d",
                  (long long)col_lo, (long long)col_hi, BK_HSIZE);
    TXB_CHECK_ARG(nrep >= 1 && ldf >= ncol && 0 <= rep_lo && rep_lo < rep_hi && rep_hi <= nrep,
                  "txb_sampler_freq: bad sizes nrep=%lld ldf=%lld columns=%lld window [%lld, %lld)", (long long)nrep,
                  (long long)ldf, (long long)ncol, (long long)rep_lo, (long long)rep_hi);
    TXB_CHECK_ARG(freq_dtype >= 0 && freq_dtype <= 3, "txb_sampler_freq: bad freq_dtype %d", freq_dtype);
    static const size_t esz[4] = {8, 4, 2, 1};
    TXB_CHECK_ARG((reinterpret_cast<uintptr_t>(freq) & 3) == 0 && ((ldf * esz[freq_dtype]) & 3) == 0,
                  "txb_sampler_freq: narrow table rows must be 4-byte aligned");
    TXB_CHECK_ARG(esz[freq_dtype] >= 4 || nsamp < (1ll << 31),
                  "txb_sampler_freq: nsamp must be < 2^31 for a narrow table (row-sum overflow check)");
    Plan p;
    int rc = make_plan(rng, ndat, nrep * nsamp, nsamp, p);
    if (rc != TXB_OK) return rc;
    rc = check_head("txb_sampler_freq", workspace, ws_bytes, p);
    if (rc != TXB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the atomic form increments a zeroed table; the bucketed form stores every lane of every row itself
    // (k_bucket_hist also covers the row padding when it is short), so it skips this pass over the table
    auto zero_table = [&]() -> int {
        if (cudaMemsetAsync(freq, 0, (size_t)(rep_hi - rep_lo) * ldf * esz[freq_dtype], st) != cudaSuccess) {
            set_error("txb_sampler_freq: cudaMemsetAsync failed");
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        return TXB_OK;
    };
    EmitArgs e{};
    rc = begin_emit(p, e, workspace, raw_used, status, prepared & TXB_SAMPLER_PREPARED, st);
    if (rc != TXB_OK) return rc;
    e.freq = freq;
    e.ldf = ldf;
    e.rep_lo = rep_lo;
    e.pos_lo = rep_lo * nsamp;
    e.pos_hi = rep_hi * nsamp;
    // ---- bucketed histogram (narrow tables, scratch beyond the head): partition the draws by sample
    // bucket in shared memory, then count every (bucket, replicate) in a shared-memory histogram
    bool bucketed = false, rowsum_folded = false;
    if (esz[freq_dtype] < 4 && !(prepared & TXB_SAMPLER_NO_BUCKETS) && nsamp < (1ll << 31)) {
        const BucketPlan bp = hist_bucket_plan(p, ndat, nsamp);
        const size_t used = align_up(p.head_bytes + 256, 256);
        const size_t avail = ws_bytes > used + 4096 ? ws_bytes - used - 4096 : 0;
        const int64_t nwin = rep_hi - rep_lo;
        int64_t wrep = 0;
        SegPlan sp{};
        // buckets kept by this call (all of them unless a column range was given); lists are sized for those
        const int b_lo = (int)(col_lo / BK_HSIZE);
        const int nbl = (int)((col_hi + BK_HSIZE - 1) / BK_HSIZE) - b_lo;
        BucketPlan bl = bp;
        bl.nb = nbl;
        if (bp.ok) {
            wrep = nwin;
            sp = seg_plan(bl, nsamp, -1, wrep);
            while (wrep > 0 && sp.per_rep_bytes * (size_t)wrep > avail) {
                const int64_t fit = (int64_t)(avail / sp.per_rep_bytes);
                wrep = fit < wrep ? fit : wrep - 1;
                if (wrep > 0) sp = seg_plan(bl, nsamp, -1, wrep);
            }
        }
        if (wrep > 0) {
            // row padding [ncol, ldf) lies inside the last bucket's slice?  else clear the table as before
            if (ldf > (int64_t)nbl * BK_HSIZE && (rc = zero_table()) != TXB_OK) return rc;
            char* dyn = reinterpret_cast<char*>(workspace) + used;
            unsigned* nlines = reinterpret_cast<unsigned*>(dyn);
            uint16_t* lists = reinterpret_cast<uint16_t*>(dyn + align_up(sizeof(unsigned) * (size_t)wrep * nbl * sp.nseg, 64));
            const size_t hist_smem = (size_t)BK_HSIZE * esz[freq_dtype];
            static bool attr_done = false;
            if (!attr_done) {
                if (cudaFuncSetAttribute(k_bucket_scatter<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         BK_MAXNB * (BK_MINCAP * 2 + 12)) != cudaSuccess ||
                    cudaFuncSetAttribute(k_bucket_scatter<1, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         BK_MAXNB * (BK_MINCAP * 2 + 12)) != cudaSuccess ||
                    cudaFuncSetAttribute(k_bucket_hist<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         BK_HSIZE) != cudaSuccess ||
                    cudaFuncSetAttribute(k_bucket_hist<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         BK_HSIZE * 2) != cudaSuccess) {
                    set_error("txb_sampler_freq: shared memory opt-in failed");
                    cudaGetLastError();
                    return TXB_ERR_CUDA;
                }
                attr_done = true;
            }
            const size_t sc_smem = (size_t)(nbl + 1) * (bp.cap * 2 + 12);  // + the ring for rejected / foreign draws
            // whole-table calls fold the overflow check into the histogram's store loop (no second pass over
            // the table); column-range calls keep k_freq_rowsum_check, which also reports the partial row sums
            const size_t lists_bytes = (size_t)wrep * nbl * sp.nseg * sp.cap_lines * BK_LINE * 2;
            unsigned long long* rowacc = (!partial && status)
                ? reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(lists) + align_up(lists_bytes, 16))
                : nullptr;
            rowsum_folded = rowacc != nullptr;
            for (int64_t r0 = rep_lo; r0 < rep_hi; r0 += wrep) {
                const int64_t r1 = r0 + wrep < rep_hi ? r0 + wrep : rep_hi;
                const int64_t nw = r1 - r0;
                BucketArgs ba{};
                ba.iter_base = e.iter_base;
                ba.iter_count = e.iter_count;
                ba.niter = (int64_t)p.nblocks * p.ipb;
                ba.status = status;
                ba.lists = lists;
                ba.nlines = nlines;
                ba.nb = nbl;
                ba.b_lo = b_lo;
                ba.compact_drain = scatter_compact_drain();
                ba.cap = bp.cap;
                ba.cap_lines = sp.cap_lines;
                ba.nseg = sp.nseg;
                ba.seg_len = sp.seg_len;
                ba.rep_lo = r0;
                {
                    ProfScope prof("bucket_scatter", st);
                    if (scatter_threads() == 1024)
                        k_bucket_scatter<1, 1024><<<dim3((unsigned)sp.nseg, (unsigned)nw), 1024, sc_smem, st>>>(p.a, ba);
                    else
                        k_bucket_scatter<1><<<dim3((unsigned)sp.nseg, (unsigned)nw), BK_THREADS, sc_smem, st>>>(p.a, ba);
                }
                TXB_CUDA_LAUNCH_CHECK("k_bucket_scatter<1>");
                HistArgs h{};
                h.lists = lists;
                h.nlines = nlines;
                h.nb = nbl;
                h.cap_lines = sp.cap_lines;
                h.nseg = sp.nseg;
                h.freq = reinterpret_cast<unsigned char*>(freq) + (size_t)(r0 - rep_lo) * ldf * esz[freq_dtype];
                h.ldf = ldf;
                h.ndat = ncol;   // columns are relative to col_lo (a multiple of the bucket size)
                h.rowacc = rowacc;
                if (rowacc && cudaMemsetAsync(rowacc, 0, sizeof(unsigned long long) * (size_t)nw, st) != cudaSuccess) {
                    set_error("txb_sampler_freq: cudaMemsetAsync(rowacc) failed");
                    cudaGetLastError();
                    return TXB_ERR_CUDA;
                }
                {
                    ProfScope prof("bucket_hist", st);
                    const dim3 grid((unsigned)nbl, (unsigned)nw);
                    if (freq_dtype == TXB_U16) k_bucket_hist<uint16_t><<<grid, BK_THREADS, hist_smem, st>>>(h);
                    else                       k_bucket_hist<uint8_t><<<grid, BK_THREADS, hist_smem, st>>>(h);
                }
                TXB_CUDA_LAUNCH_CHECK("k_bucket_hist");
                if (rowacc) {
                    k_rowacc_verify<<<(unsigned)((nw + 255) / 256), 256, 0, st>>>(rowacc, nw, (long long)nsamp, status);
                    TXB_CUDA_LAUNCH_CHECK("k_rowacc_verify");
                }
            }
            bucketed = true;
        }
    }
    if (!bucketed && partial) {
        set_error("txb_sampler_freq_cols: a column range needs the bucketed histogram (uint8 / uint16 table, >= %d "
                  "samples, scratch beyond the sampler head); build the full table and slice it instead", 8 * BK_HSIZE);
        return TXB_ERR_UNSUPPORTED;
    }
    if (!bucketed && (rc = zero_table()) != TXB_OK) return rc;
    if (!bucketed) switch (freq_dtype) {
        case TXB_I64: k_sampler<MODE_HIST, 1, false, false, int64_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e); break;
        case TXB_I32: k_sampler<MODE_HIST, 1, false, false, int32_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e); break;
        case TXB_U16: k_sampler<MODE_HIST, 1, false, false, uint16_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e); break;
        default:      k_sampler<MODE_HIST, 1, false, false, uint8_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e); break;
    }
    TXB_CUDA_LAUNCH_CHECK("k_sampler<hist>");
    if (status && esz[freq_dtype] < 4 && !rowsum_folded) {  // narrow lanes are bumped without a return value: verify the row sums
        const unsigned char* f = reinterpret_cast<const unsigned char*>(freq);
        const int64_t row_bytes = ldf * (int64_t)esz[freq_dtype];
        const unsigned rows = (unsigned)(rep_hi - rep_lo);
        const int64_t len = rowsum ? sum_cols * (int64_t)esz[freq_dtype] : -1;
        if (freq_dtype == TXB_U16) k_freq_rowsum_check<2><<<rows, 1024, 0, st>>>(f, row_bytes, nsamp, status, rowsum, len);
        else                       k_freq_rowsum_check<1><<<rows, 1024, 0, st>>>(f, row_bytes, nsamp, status, rowsum, len);
        TXB_CUDA_LAUNCH_CHECK("k_freq_rowsum_check");
    }
    return TXB_OK;
}

extern "C" int txb_sampler_freq(const txb_pcg64* rng, int64_t ndat, int64_t nrep, int64_t nsamp,
                                int64_t rep_lo, int64_t rep_hi, void* freq, int freq_dtype, int64_t ldf,
                                int64_t* raw_used, int* status, int prepared, void* workspace,
                                size_t ws_bytes, void* stream) {
    return sampler_freq_impl(rng, ndat, nrep, nsamp, rep_lo, rep_hi, 0, ndat, freq, freq_dtype, ldf, raw_used,
                             nullptr, 0, status, prepared, workspace, ws_bytes, stream);
}

extern "C" int txb_sampler_freq_cols(const txb_pcg64* rng, int64_t ndat, int64_t nrep, int64_t nsamp,
                                     int64_t rep_lo, int64_t rep_hi, int64_t col_lo, int64_t col_hi, void* freq,
                                     int freq_dtype, int64_t ldf, int64_t* rowsum, int64_t sum_cols, int* status,
                                     int prepared, void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(rowsum != nullptr && status != nullptr, "txb_sampler_freq_cols: rowsum and status are required");
    TXB_CHECK_ARG(sum_cols >= 0 && sum_cols <= col_hi - col_lo, "txb_sampler_freq_cols: sum_cols out of range");
    TXB_CHECK_ARG(freq_dtype == TXB_U8 || freq_dtype == TXB_U16, "txb_sampler_freq_cols: uint8 / uint16 tables only");
    return sampler_freq_impl(rng, ndat, nrep, nsamp, rep_lo, rep_hi, col_lo, col_hi, freq, freq_dtype, ldf, nullptr,
                             reinterpret_cast<long long*>(rowsum), sum_cols, status, prepared, workspace, ws_bytes,
                             stream);
}

extern "C" size_t txb_sampler_resample_workspace_bytes(int64_t ndat, int order, int64_t nrep, int64_t nsamp,
                                                       int64_t nrep_out) {
    const size_t head = head_bytes_for(ndat, nrep * nsamp);
    if (head == 0) return 0;
    const size_t dbl = (size_t)(2 * ndat) + (size_t)(2 * nrep_out * (order + 2)) + 16;
    return align_up(head + dbl * sizeof(double) + 512, 256);
}

extern "C" size_t txb_sampler_resample_fast_workspace_bytes(int64_t ndat, int order, int mode, int64_t nrep,
                                                            int64_t nsamp, int64_t nrep_out) {
    const txb_pcg64 dummy = {0, 0, 0, 1, 0, 0};
    Plan p;
    if (make_plan(&dummy, ndat, nrep * nsamp, nsamp, p) != TXB_OK) return 0;
    const int ks = mode == TXB_MODE_X_IS_U ? order + 1 : order;
    const BucketPlan b = bucket_plan(p, ndat, nsamp, nrep_out, order);
    if (!b.ok) return txb_sampler_resample_workspace_bytes(ndat, order, nrep, nsamp, nrep_out);
    const SegPlan sp = seg_plan(b, nsamp, ks, nrep_out);
    return align_up(b.fixed_bytes + sp.per_rep_bytes * (size_t)nrep_out + 4096, 256);
}

extern "C" int txb_sampler_resample_f64(const txb_pcg64* rng, const double* x, const double* u,
                                        const double* w, int64_t ndat, int64_t ldx, int order, int mode,
                                        int64_t nrep, int64_t nsamp, int64_t rep_lo, int64_t rep_hi,
                                        double* out, int64_t* raw_used, int* status, int flags,
                                        void* workspace, size_t ws_bytes, void* stream) {
    TXB_CHECK_ARG(u && out && workspace, "txb_sampler_resample_f64: null pointer");
    TXB_CHECK_ARG(order >= 1 && nrep >= 1 && 0 <= rep_lo && rep_lo < rep_hi && rep_hi <= nrep,
                  "txb_sampler_resample_f64: order=%d nrep=%lld window [%lld, %lld)", order, (long long)nrep,
                  (long long)rep_lo, (long long)rep_hi);
    TXB_CHECK_ARG(mode == TXB_MODE_COMOMENTS ? x != nullptr : (mode == TXB_MODE_X_IS_U && x == nullptr),
                  "txb_sampler_resample_f64: mode COMOMENTS needs x, mode X_IS_U needs x == NULL");
    const int prepared = flags & TXB_SAMPLER_PREPARED;
    const int64_t nwin = rep_hi - rep_lo;
    Plan p;
    int rc = make_plan(rng, ndat, nrep * nsamp, nsamp, p);
    if (rc != TXB_OK) return rc;
    if (ws_bytes < txb_sampler_resample_workspace_bytes(ndat, order, nrep, nsamp, nwin)) {
        set_error("txb_sampler_resample_f64: workspace too small");
        return TXB_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int ks = mode == TXB_MODE_X_IS_U ? order + 1 : order;
    TXB_CHECK_ARG(ks <= TXB_MAXK, "txb_sampler_resample_f64: order %d not supported", order);
    char* base = reinterpret_cast<char*>(workspace) + p.head_bytes;
    base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(base) + 15) & ~(uintptr_t)15);
    double2* packed = reinterpret_cast<double2*>(base);
    double* pivot = reinterpret_cast<double*>(packed + ndat);
    double* S0 = pivot + 4;
    double* S1 = S0 + nwin * (order + 2);
    char* dyn = reinterpret_cast<char*>(S1 + nwin * (order + 2));
    dyn = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(dyn) + 63) & ~(uintptr_t)63);
    rc = txb_pivot_f64(x, u, ndat, 1, x ? ldx : 1, 0, pivot, stream);
    if (rc != TXB_OK) return rc;
    rc = launch_pack_shifted(x, x ? ldx : 0, u, ndat, pivot, packed, st);
    if (rc != TXB_OK) return rc;
    EmitArgs e{};
    rc = begin_emit(p, e, workspace, raw_used, status, prepared, st);
    if (rc != TXB_OK) return rc;

    // ---- bucketed path when the table is too big for L2-resident gathers and scratch allows it
    const BucketPlan bp = bucket_plan(p, ndat, nsamp, nwin, order);
    const size_t used = (size_t)(dyn - reinterpret_cast<char*>(workspace));
    const size_t avail = ws_bytes > used + 4096 ? ws_bytes - used - 4096 : 0;
    int64_t wrep = 0;
    SegPlan sp{};
    if (bp.ok && !(flags & TXB_SAMPLER_NO_BUCKETS)) {
        wrep = nwin;
        sp = seg_plan(bp, nsamp, ks, wrep);
        while (wrep > 0 && sp.per_rep_bytes * (size_t)wrep > avail) {  // fewer replicates per window
            const int64_t fit = (int64_t)(avail / sp.per_rep_bytes);
            wrep = fit < wrep ? fit : wrep - 1;
            if (wrep > 0) sp = seg_plan(bp, nsamp, ks, wrep);
        }
        if (wrep < 32 && wrep < nwin) wrep = 0;  // too little scratch to be worth it
    }
    if (wrep > 0) {
        unsigned* nlines = reinterpret_cast<unsigned*>(dyn);
        double* ws0 = reinterpret_cast<double*>(dyn + align_up(sizeof(unsigned) * (size_t)wrep * bp.nb * sp.nseg, 64));
        double* ws1 = ws0 + (size_t)bp.nb * wrep * (ks + 1);
        uint16_t* lists = reinterpret_cast<uint16_t*>(ws1 + (size_t)bp.nb * wrep * (ks + 1));
        lists = reinterpret_cast<uint16_t*>((reinterpret_cast<uintptr_t>(lists) + 63) & ~(uintptr_t)63);
        static bool attr_done = false;
        if (!attr_done) {
            if (cudaFuncSetAttribute(k_bucket_scatter<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     BK_MAXNB * (BK_MINCAP * 2 + 12)) != cudaSuccess ||
                cudaFuncSetAttribute(k_bucket_scatter<0, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     BK_MAXNB * (BK_MINCAP * 2 + 12)) != cudaSuccess) {
                set_error("k_bucket_scatter: shared memory opt-in failed");
                cudaGetLastError();
                return TXB_ERR_CUDA;
            }
            attr_done = true;
        }
        const size_t sc_smem = (size_t)(bp.nb + 1) * (bp.cap * 2 + 12);  // + the ring for rejected draws
        for (int64_t r0 = rep_lo; r0 < rep_hi; r0 += wrep) {
            const int64_t r1 = r0 + wrep < rep_hi ? r0 + wrep : rep_hi;
            const int64_t nw = r1 - r0;
            BucketArgs ba{};
            ba.iter_base = e.iter_base;
            ba.iter_count = e.iter_count;
            ba.niter = (int64_t)p.nblocks * p.ipb;
            ba.status = status;
            ba.lists = lists;
            ba.nlines = nlines;
            ba.nb = bp.nb;
            ba.cap = bp.cap;
            ba.cap_lines = sp.cap_lines;
            ba.nseg = sp.nseg;
            ba.seg_len = sp.seg_len;
            ba.rep_lo = r0;
            ba.compact_drain = scatter_compact_drain();
            {
                ProfScope prof("bucket_scatter", st);
                if (scatter_threads() == 1024)
                    k_bucket_scatter<0, 1024><<<dim3((unsigned)sp.nseg, (unsigned)nw), 1024, sc_smem, st>>>(p.a, ba);
                else
                    k_bucket_scatter<0><<<dim3((unsigned)sp.nseg, (unsigned)nw), BK_THREADS, sc_smem, st>>>(p.a, ba);
            }
            TXB_CUDA_LAUNCH_CHECK("k_bucket_scatter");
            ConsumeArgs c{};
            c.packed = packed;
            c.w = w;
            c.ndat = ndat;
            c.lists = lists;
            c.nlines = nlines;
            c.nb = bp.nb;
            c.cap_lines = sp.cap_lines;
            c.nseg = sp.nseg;
            c.nwin = nw;
            c.ws0 = ws0;
            c.ws1 = ws1;
            int gy = (2 * sm_count() + bp.nb - 1) / bp.nb;
            const int gy_max = (int)((nw + BK_THREADS / 32 - 1) / (BK_THREADS / 32));
            if (gy > gy_max) gy = gy_max;
            if (gy < 1) gy = 1;
            {
                ProfScope prof("bucket_consume", st);
                TXB_DISPATCH_ORDER(ks, { rc = launch_consume<K>(c, x != nullptr, w != nullptr, gy, st); });
            }
            if (rc != TXB_OK) return rc;
            rc = launch_sum_parts(ws0, bp.nb, nw * (ks + 1), S0 + (r0 - rep_lo) * (ks + 1), st);
            if (rc == TXB_OK && x) rc = launch_sum_parts(ws1, bp.nb, nw * (ks + 1), S1 + (r0 - rep_lo) * (ks + 1), st);
            if (rc != TXB_OK) return rc;
        }
        rc = txb_finalize_f64(S0, S1, pivot, nwin, 1, 1, order, mode, out, stream);
        if (rc == TXB_OK && status) {  // tell the caller which path ran (bit 32)
            k_status_or<<<1, 1, 0, st>>>(status, 32);
            TXB_CUDA_LAUNCH_CHECK("k_status_or");
        }
        return rc;
    }

    // ---- gather path
    if (cudaMemsetAsync(S0, 0, sizeof(double) * 2 * nwin * (order + 2), st) != cudaSuccess) {
        set_error("txb_sampler_resample_f64: cudaMemsetAsync failed");
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    e.packed = packed;
    e.w = w;
    e.S0 = S0;
    e.S1 = S1;
    e.rep_lo = rep_lo;
    e.pos_lo = rep_lo * nsamp;
    e.pos_hi = rep_hi * nsamp;
    {
    ProfScope prof("sampler_gather", st);
    TXB_DISPATCH_ORDER(ks, {
        if (x) {
            if (w) k_sampler<MODE_GATHER, K, true, true, int64_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e);
            else   k_sampler<MODE_GATHER, K, true, false, int64_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e);
        } else {
            if (w) k_sampler<MODE_GATHER, K, false, true, int64_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e);
            else   k_sampler<MODE_GATHER, K, false, false, int64_t><<<p.nblocks, SMP_THREADS, 0, st>>>(p.a, e);
        }
    });
    }
    TXB_CUDA_LAUNCH_CHECK("k_sampler<gather>");
    return txb_finalize_f64(S0, S1, pivot, nwin, 1, 1, order, mode, out, stream);
}

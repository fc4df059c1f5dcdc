// K6 + K2 for scalar x without random HBM gathers: "bucketed" bootstrap.  Included by sampler.cu.
//
// The fused gather kernel (k_sampler<GATHER>) is bound by random 64-byte DRAM bursts: 60 B of
// traffic per 16-byte draw on a table that cannot stay in L2 (profiles/r01_gather_ncu_full.txt).
// Here the draws are first partitioned by sample bucket (8192 consecutive samples) with software
// write-combining through shared memory, then every bucket is consumed from a shared-memory copy
// of its samples:
//
//   k_bucket_scatter  regenerates the sampler's stream (same blocks / offsets as every other
//                     pass-B kernel), appends each draw's 13-bit local index to a 64-entry ring
//                     per bucket in shared memory, and writes full 32-entry lines (64 B) to the
//                     (replicate, bucket) list in HBM.  Rings are drained (padded lines) at every
//                     replicate boundary and at the end of the block.
//   k_bucket_consume  one CTA per bucket: stages the bucket's (dx, du[, w]) in shared memory, then
//                     each warp walks the lists of its replicates with coalesced 16-byte loads and
//                     gathers from shared memory; per-(replicate, bucket) power sums are reduced
//                     with shuffles and written as partials (summed by k_sum_parts).
//
// HBM traffic: 2 bytes per draw written + read, streaming, instead of a random 64-byte burst.
#pragma once

namespace txb {

constexpr int BK_LOG = 13;
constexpr int BK_SIZE = 1 << BK_LOG;     // samples per bucket
constexpr int BK_MINCAP = 64;            // smallest ring (entries per bucket, uint16); rings are a power of two
constexpr int BK_LINE = 32;              // entries per flushed line (64 bytes)
constexpr int BK_MAXNB = 1536;           // buckets a producer CTA can hold rings for
constexpr int BK_THREADS = 512;
constexpr uint16_t BK_EMPTY = 0xFFFFu;
// bucket geometry.  0: 8192-sample buckets (the comoment consumer stages a bucket's samples in shared
// memory).  1: 65520-sample buckets for the histogram consumer (16-bit local index with 0xFFFF left
// free for padding; a multiple of 16 so every bucket's slice of a frequency row stays 16-byte aligned).
constexpr int BK_HSIZE = 65520;
template <int GEOM>
__device__ __forceinline__ uint32_t bk_bucket(uint32_t v) {
    return GEOM ? v / (uint32_t)BK_HSIZE : v >> BK_LOG;
}
template <int GEOM>
__device__ __forceinline__ uint16_t bk_local(uint32_t v, uint32_t b) {
    return (uint16_t)(GEOM ? v - b * (uint32_t)BK_HSIZE : v & (BK_SIZE - 1));
}

struct BucketArgs {
    const int64_t* iter_base;   // [niter] output position of the first draw of each iteration
    const int* iter_count;      // [niter]
    int64_t niter;
    int* status;          // |= 8 when a list overflows its capacity
    uint16_t* lists;      // [nwin][nb][nseg][cap_lines][BK_LINE]
    unsigned* nlines;     // [nwin][nb][nseg] lines written
    int nb;               // buckets
    int cap;              // ring entries per bucket (power of two >= BK_MINCAP)
    int cap_lines;        // per segment list
    int nseg;
    int64_t seg_len;      // draws per segment
    int64_t rep_lo;       // first replicate of the window
    int b_lo;             // first bucket kept (draws of other buckets go to the sink ring): a rank's column range
    int compact_drain;    // 1: per-warp compacted flush of full lines (default); 0: every thread walks its buckets
};

// exclusive block scan for NTH threads
template <int NTH>
__device__ __forceinline__ int bk_block_scan(int v, int* sm_warp, int& total) {
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();
    if (lane == 31) sm_warp[wrp] = inc;
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < NTH / 32; ++k) {
        const int w = sm_warp[k];
        if (k < wrp) base += w;
        tot += w;
    }
    total = tot;
    return base + inc - v;
}

// NTH = 512 (8 LCG steps per thread and iteration) or 1024 (4 steps): the same 4096-step iterations either
// way, so the position table of the counting pass serves both.  The 1024-thread form halves the registers
// a thread needs for its draws (<= 64, the cap of a 1024-thread CTA) and doubles the warps that hide the
// shared-memory atomic latency (ncu on the 512-thread form: 25 % warps active, 38 % issue-active).
template <int GEOM, int NTH = BK_THREADS>
__global__ void __launch_bounds__(NTH, 1)
    k_bucket_scatter(const __grid_constant__ SamplerArgs a, const BucketArgs e) {
    constexpr int T = SMP_T * BK_THREADS / NTH;
    constexpr int NSLOT = 2 * T;
    extern __shared__ __align__(16) unsigned char bk_smem[];
    const int nb = e.nb;
    const unsigned cap = (unsigned)e.cap;
    // ring nb is a sink: rejected raw draws are "appended" there so that the fast path is branch-free
    // (it is never drained; its tail just wraps)
    const int nbp = nb + 1;
    uint16_t* rings = reinterpret_cast<uint16_t*>(bk_smem);                         // [nb + 1][cap]
    unsigned* tail = reinterpret_cast<unsigned*>(bk_smem + (size_t)nbp * cap * 2);  // [nb + 1]
    unsigned* head = tail + nbp;                                                    // [nb + 1]
    unsigned* lines = head + nbp;                                                   // [nb + 1] lines written
    __shared__ int sm_warp[NTH / 32];
    __shared__ long long sm_j0;

    const int seg = blockIdx.x;
    const int64_t r = blockIdx.y;  // window-relative replicate
    const int64_t rep_begin = (e.rep_lo + r) * a.nsamp;
    const int64_t A = rep_begin + (int64_t)seg * e.seg_len;
    int64_t B = A + e.seg_len;
    if (B > rep_begin + a.nsamp) B = rep_begin + a.nsamp;
    const int64_t list0 = (r * nb) * e.nseg + seg;  // + b * nseg

    for (int b = threadIdx.x; b < nbp; b += NTH) tail[b] = 0u, head[b] = 0u, lines[b] = 0u;
    if (threadIdx.x == 0) {
        // last iteration whose first output position is <= A
        int64_t lo = 0, hi = e.niter - 1;
        while (lo < hi) {
            const int64_t mid = (lo + hi + 1) >> 1;
            if (e.iter_base[mid] <= A) lo = mid; else hi = mid - 1;
        }
        sm_j0 = lo;
    }
    __syncthreads();
    if (A >= B) {  // empty segment
        for (int b = threadIdx.x; b < nb; b += NTH) e.nlines[list0 + (int64_t)b * e.nseg] = 0u;
        return;
    }
    const int64_t j0 = sm_j0;
    uint64_t hi = a.s_hi, lo = a.s_lo;
    {
        uint64_t q = ((uint64_t)j0 * NTH + threadIdx.x) * T;
        for (int i = 0; q != 0 && i < SMP_NPOW; ++i, q >>= 1)
            if (q & 1) affine_apply(hi, lo, a.pow2[i]);
    }
    const uint32_t n = a.n, thresh = a.thresh;

    auto write_line = [&](int b, unsigned h, int cnt) {
        const unsigned line = lines[b];
        lines[b] = line + 1;
        if (line >= (unsigned)e.cap_lines) {
            if (e.status) atomicOr(e.status, 8);
            return;
        }
        uint4* dst = reinterpret_cast<uint4*>(e.lists + ((list0 + (int64_t)b * e.nseg) * e.cap_lines + line) * BK_LINE);
        const uint16_t* src = rings + (size_t)b * cap + (h & (cap - 1));  // h is a multiple of BK_LINE
        if (cnt == BK_LINE) {
            const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll
            for (int q = 0; q < 4; ++q) __stcs(dst + q, s4[q]);
        } else {
            uint16_t tmp[BK_LINE];
#pragma unroll
            for (int q = 0; q < BK_LINE; ++q) tmp[q] = q < cnt ? src[q] : BK_EMPTY;
            const uint4* t4 = reinterpret_cast<const uint4*>(tmp);
#pragma unroll
            for (int q = 0; q < 4; ++q) __stcs(dst + q, t4[q]);
        }
    };
    // every bucket ring is drained by one fixed thread: full lines always, the remainder when `all`
    auto drain = [&](bool all) {
        for (int b = threadIdx.x; b < nb; b += NTH) {
            unsigned t = tail[b], h = head[b];
            if (t - h > cap) {
                // more arrivals than the ring holds (> 6 sigma).  The checked append writes those as
                // one-entry lines; the fast append has wrapped over unread entries: flag it, the
                // caller redoes the call on the gather path.
                if (e.status) atomicOr(e.status, 16);
                t = h + cap;
            }
            while (t - h >= (unsigned)BK_LINE) {
                write_line(b, h, BK_LINE);
                h += BK_LINE;
            }
            if (all) {
                if (t != h) write_line(b, h, (int)(t - h));
                t = h = 0u;
            }
            tail[b] = t;
            head[b] = h;
        }
    };
    // The per-iteration drain, compacted.  ncu (instructions executed per source region, round 2): the walk above
    // costs 210 of the 856 warp-instructions of an iteration although only one bucket in six has a full line --
    // the flush code runs predicated for every bucket of every thread.  Here a warp first lists the buckets that
    // do have a full line (ballot + popc into a per-warp shared list), then its lanes flush one listed bucket each.
    constexpr int DR_K = (BK_MAXNB + NTH - 1) / NTH;      // buckets per thread
    __shared__ uint16_t sm_wl[NTH / 32][DR_K * 32];
    auto drain_lines = [&]() {
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
        int nfl = 0;
#pragma unroll
        for (int k = 0; k < DR_K; ++k) {
            const int b = threadIdx.x + k * NTH;
            bool need = false;
            if (b < nb) {
                const unsigned t = tail[b], h = head[b];
                need = t - h >= (unsigned)BK_LINE;
                if (t - h > cap) {   // ring overrun (> 6 sigma): flag, clamp (as in drain())
                    if (e.status) atomicOr(e.status, 16);
                    tail[b] = h + cap;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, need);
            if (need) sm_wl[wrp][nfl + __popc(m & ((1u << lane) - 1u))] = (uint16_t)b;
            nfl += __popc(m);
        }
        __syncwarp();
        for (int i = lane; i < nfl; i += 32) {
            const int b = sm_wl[wrp][i];
            const unsigned t = tail[b];
            unsigned h = head[b];
            do {
                write_line(b, h, BK_LINE);
                h += BK_LINE;
            } while (t - h >= (unsigned)BK_LINE);
            head[b] = h;
        }
    };
    auto append = [&](uint32_t v) {
        const uint32_t rb = bk_bucket<GEOM>(v) - (uint32_t)e.b_lo;
        if (rb >= (uint32_t)nb) return;  // outside this call's column range
        const int b = (int)rb;
        const uint16_t loc = bk_local<GEOM>(v, rb + (uint32_t)e.b_lo);
        const unsigned slot = atomicAdd(tail + b, 1u);
        if (slot - head[b] < cap) {
            rings[(size_t)b * cap + (slot & (cap - 1))] = loc;
        } else {  // ring full (> 6 sigma): a one-entry line, slot taken with a shared-memory atomic
            const unsigned line = atomicAdd(lines + b, 1u);
            if (line >= (unsigned)e.cap_lines) {
                if (e.status) atomicOr(e.status, 8);
            } else {
                uint16_t* dst = e.lists + ((list0 + (int64_t)b * e.nseg) * e.cap_lines + line) * BK_LINE;
                dst[0] = loc;
                for (int q = 1; q < BK_LINE; ++q) dst[q] = BK_EMPTY;
            }
        }
    };

    // The draws of iteration j + 1 are generated (pure ALU: 8 dependent 128-bit LCG steps) right after
    // this thread's share of the drain of iteration j, i.e. between the two block-wide barriers, so
    // that warps still waiting on the drain's shared-memory loads overlap with warps already issuing
    // integer work.
    uint32_t val[NSLOT];
    uint32_t mask = 0;
    auto generate = [&]() {
        mask = 0;
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
    };
    generate();
    // position / size of the next iteration are fetched one iteration ahead (global loads off the critical path)
    int64_t P_next = e.iter_base[j0];
    int total_next = e.iter_count[j0];
    for (int64_t j = j0; j < e.niter; ++j) {
        const int64_t P = P_next;
        if (P >= B) break;
        const int total = total_next;
        if (j + 1 < e.niter) {
            P_next = e.iter_base[j + 1];
            total_next = e.iter_count[j + 1];
        }
        const bool pre = j == 0 && threadIdx.x == 0 && a.pre_valid != 0;
        uint32_t pre_val = 0;
        bool pre_acc = false;
        if (pre) {
            const uint64_t m = (uint64_t)a.pre_u32 * n;
            pre_val = (uint32_t)(m >> 32);
            pre_acc = (uint32_t)m >= thresh;
        }
        if (P >= A && P + total <= B) {  // block-uniform: the whole iteration belongs to this segment
            if (pre_acc) append(pre_val);
            // all slot reservations first (independent shared-memory atomics), then the stores
            unsigned slot[NSLOT];
            uint32_t bk[NSLOT];
#pragma unroll
            for (int k = 0; k < NSLOT; ++k) {
                // bucket relative to the kept range; below b_lo it wraps to a huge value -> sink, like a rejected draw
                const uint32_t rb = bk_bucket<GEOM>(val[k]) - (uint32_t)e.b_lo;
                bk[k] = ((mask & (1u << k)) && rb < (uint32_t)nb) ? rb : (uint32_t)nb;
                slot[k] = atomicAdd(tail + bk[k], 1u);
            }
#pragma unroll
            for (int k = 0; k < NSLOT; ++k)
                rings[bk[k] * cap + (slot[k] & (cap - 1))] = bk_local<GEOM>(val[k], bk[k] + (uint32_t)e.b_lo);
        } else {
            int tot;
            const int c = __popc(mask) + (pre_acc ? 1 : 0);
            int64_t pos = P + bk_block_scan<NTH>(c, sm_warp, tot);
            if (pre_acc) {
                if (pos >= A && pos < B) append(pre_val);
                ++pos;
            }
#pragma unroll
            for (int k = 0; k < NSLOT; ++k) {
                if (mask & (1u << k)) {
                    if (pos >= A && pos < B) append(val[k]);
                    ++pos;
                }
            }
        }
        __syncthreads();
        if (e.compact_drain) drain_lines(); else drain(false);
        affine_apply(hi, lo, NTH == BK_THREADS ? a.iter_jump2 : a.iter_jump3);
        generate();  // iteration j + 1 (one unused set at the end of the block's range)
        __syncthreads();
    }
    drain(true);
    for (int b = threadIdx.x; b < nb; b += NTH) {
        const unsigned nl = lines[b];
        e.nlines[list0 + (int64_t)b * e.nseg] = nl < (unsigned)e.cap_lines ? nl : (unsigned)e.cap_lines;
    }
}

// Raw position of output #count-1 (what numpy's generator has consumed when the table is complete), from the
// per-iteration position table: binary-search the iteration that emits it, regenerate that ONE iteration (4096 LCG
// steps) and scan.  Replaces the MODE_LOCATE pass of k_sampler, where the single block holding the last output
// walked its whole range serially (30 ms on the critical path of every 1e11-draw plan, ncu launch list r2).
__global__ void __launch_bounds__(BK_THREADS, 1)
    k_sampler_locate(const __grid_constant__ SamplerArgs a, const int64_t* __restrict__ iter_base, int64_t niter,
                     int64_t* __restrict__ raw_used) {
    constexpr int T = SMP_T, NSLOT = 2 * T;
    __shared__ int sm_warp[BK_THREADS / 32];
    __shared__ long long sm_j;
    const int64_t target = a.count - 1;
    if (threadIdx.x == 0) {
        int64_t lo = 0, hi = niter - 1;
        while (lo < hi) {   // last iteration whose first output position is <= target
            const int64_t mid = (lo + hi + 1) >> 1;
            if (iter_base[mid] <= target) lo = mid; else hi = mid - 1;
        }
        sm_j = lo;
    }
    __syncthreads();
    const int64_t j = sm_j;
    const int64_t P = iter_base[j];
    const int64_t q0 = (j * BK_THREADS + threadIdx.x) * T;   // first LCG step of this thread
    uint64_t hi = a.s_hi, lo = a.s_lo;
    {
        uint64_t q = (uint64_t)q0;
        for (int i = 0; q != 0 && i < SMP_NPOW; ++i, q >>= 1)
            if (q & 1) affine_apply(hi, lo, a.pow2[i]);
    }
    const uint32_t n = a.n, thresh = a.thresh;
    uint32_t mask = 0;
#pragma unroll
    for (int k = 0; k < T; ++k) {
        const uint64_t o = pcg_step_out(hi, lo, a.inc_hi, a.inc_lo);
        if ((uint32_t)((uint64_t)(uint32_t)o * n) >= thresh) mask |= 1u << (2 * k);
        if ((uint32_t)((o >> 32) * (uint64_t)n) >= thresh) mask |= 1u << (2 * k + 1);
    }
    const bool pre_acc = j == 0 && threadIdx.x == 0 && a.pre_valid != 0 &&
                         (uint32_t)((uint64_t)a.pre_u32 * n) >= thresh;
    int total;
    int off = bk_block_scan<BK_THREADS>(__popc(mask) + (pre_acc ? 1 : 0), sm_warp, total);
    // raw 32-bit draw number of slot 0 of this thread (0-based, the buffered half is #0)
    const int64_t raw0 = (int64_t)a.pre_valid + 2 * q0;
    if (pre_acc) {
        if (P + off == target) *raw_used = 1;
        ++off;
    }
#pragma unroll
    for (int k = 0; k < NSLOT; ++k) {
        if (mask & (1u << k)) {
            if (P + off == target) *raw_used = raw0 + k + 1;
            ++off;
        }
    }
}

struct ConsumeArgs {
    const double2* packed;   // [ndat] (x - cx, u - cu)
    const double* w;         // [ndat] or NULL
    int64_t ndat;
    const uint16_t* lists;
    const unsigned* nlines;
    int nb, cap_lines, nseg;
    int64_t nwin;            // replicates in this window
    double* ws0;             // [nb][nwin][K+1] partial sums per bucket
    double* ws1;
};

// one CTA per bucket; warps take replicates round-robin
template <int K, bool HAS_X, bool HAS_W>
__global__ void __launch_bounds__(BK_THREADS, 1) k_bucket_consume(const ConsumeArgs c) {
    extern __shared__ __align__(16) unsigned char bk_smem[];
    double2* tile = reinterpret_cast<double2*>(bk_smem);                                  // [BK_SIZE]
    double* wt = reinterpret_cast<double*>(bk_smem + sizeof(double2) * BK_SIZE);          // [BK_SIZE] if HAS_W
    const int b = blockIdx.x;
    const int64_t s0 = (int64_t)b * BK_SIZE;
    for (int i = threadIdx.x; i < BK_SIZE; i += BK_THREADS) {
        const int64_t s = s0 + i;
        tile[i] = s < c.ndat ? c.packed[s] : make_double2(0.0, 0.0);
        if (HAS_W) wt[i] = s < c.ndat ? c.w[s] : 0.0;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    constexpr int NW = BK_THREADS / 32;
    for (int64_t r = blockIdx.y * NW + wrp; r < c.nwin; r += (int64_t)gridDim.y * NW) {
        Acc<K, HAS_X> acc;
        acc.zero();
        for (int seg = 0; seg < c.nseg; ++seg) {
            const int64_t list = (r * c.nb + b) * c.nseg + seg;
            const uint4* L = reinterpret_cast<const uint4*>(c.lists + list * c.cap_lines * BK_LINE);
            const int nq = (int)c.nlines[list] * 4;  // 16-byte chunks (8 entries each)
            uint4 q = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
            if (lane < nq) q = __ldcs(L + lane);
            for (int i = lane; i < nq; i += 32) {
                const uint4 cur = q;
                if (i + 32 < nq) q = __ldcs(L + i + 32);  // prefetch the next chunk
                const uint32_t wds[4] = {cur.x, cur.y, cur.z, cur.w};
                double2 v[8];
                double pw[8];
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const uint32_t loc = (wds[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                    const bool ok = loc != BK_EMPTY;
                    const uint32_t li = ok ? loc : 0u;
                    v[t] = tile[li];
                    pw[t] = ok ? (HAS_W ? wt[li] : 1.0) : 0.0;
                }
#pragma unroll
                for (int t = 0; t < 8; ++t) acc.push(v[t], pw[t]);
            }
        }
        const int64_t o = ((int64_t)b * c.nwin + r) * (K + 1);
#pragma unroll
        for (int j = 0; j <= K; ++j) {
            const double a0 = warp_sum(acc.s0[j]);
            if (lane == 0) c.ws0[o + j] = a0;
            if (HAS_X) {
                const double a1 = warp_sum(acc.s1[j]);
                if (lane == 0) c.ws1[o + j] = a1;
            }
        }
    }
}

struct HistArgs {
    const uint16_t* lists;
    const unsigned* nlines;
    int nb, cap_lines, nseg;
    void* freq;        // row 0 = first replicate of this window
    int64_t ldf, ndat;
    unsigned long long* rowacc;   // [replicates of the window] sum of the stored lanes (overflow check), or NULL
};

// K6 frequency table from the bucket lists: one CTA per (bucket, replicate) counts the bucket's draws in
// a shared-memory histogram (shared-memory atomics; global RED.ADD on a 10-100 MB row measured only
// 2.2e10 increments/s) and writes its 65520-entry slice of the replicate's row with coalesced stores.
// A lane that wraps carries into its neighbour exactly as in the global-atomic form, so the row-sum
// check (k_freq_rowsum_check) covers both.
template <typename FT>
__global__ void __launch_bounds__(BK_THREADS) k_bucket_hist(const HistArgs c) {
    extern __shared__ __align__(16) unsigned char bk_smem[];
    constexpr int ES = (int)sizeof(FT);
    constexpr int NWORD = BK_HSIZE * ES / 4;
    unsigned* h = reinterpret_cast<unsigned*>(bk_smem);
    const int b = blockIdx.x;
    const int64_t r = blockIdx.y;
    // the lists of the (replicate, bucket)'s segments are walked as ONE flattened sequence of 16-byte chunks, so
    // that all 512 threads stay busy however the draws are split into segments (37 segments of ~220 chunks at the
    // target shape): s_off[sg] = first flattened chunk of segment sg (exclusive scan by warp 0)
    __shared__ int s_off[257];
    for (int i = threadIdx.x; i < NWORD; i += BK_THREADS) h[i] = 0u;
    const int64_t list0 = (r * c.nb + b) * c.nseg;
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int v[8], sum = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int sg = lane * 8 + k;
            v[k] = sg < c.nseg ? (int)c.nlines[list0 + sg] * 4 : 0;
            sum += v[k];
        }
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        int run = inc - sum;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s_off[lane * 8 + k] = run;
            run += v[k];
        }
        if (lane == 31) s_off[256] = run;
    }
    __syncthreads();
    const int total = s_off[256];
    // U independent 16-byte loads in flight per thread (ncu on the one-load-at-a-time loop: 58 % of the stall
    // samples on the list load's scoreboard)
    constexpr int U = 4;
    int seg = 0, f = threadIdx.x;
    auto next_chunk = [&]() -> const uint4* {
        if (f >= total) return nullptr;
        while (f >= s_off[seg + 1]) ++seg;
        const uint4* p = reinterpret_cast<const uint4*>(c.lists + (list0 + seg) * c.cap_lines * BK_LINE) + (f - s_off[seg]);
        f += BK_THREADS;
        return p;
    };
    for (;;) {
        const uint4* p[U];
#pragma unroll
        for (int u = 0; u < U; ++u) p[u] = next_chunk();
        if (p[0] == nullptr) break;
        uint4 q[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            q[u] = p[u] ? __ldcs(p[u]) : make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t wds[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const uint32_t loc = (wds[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                if (loc != BK_EMPTY) {
                    const uint32_t byte = loc * ES;
                    atomicAdd(h + (byte >> 2), 1u << (8 * (byte & 3)));
                }
            }
        }
    }
    __syncthreads();
    const int64_t s0 = (int64_t)b * BK_HSIZE;
    const int64_t left = c.ndat - s0;
    // the last bucket also stores the (zero) lanes of the row padding [ndat, ldf) when they fit in its slice,
    // so the caller does not have to clear the table first; rows are whole words
    const int64_t cover = (b == c.nb - 1 ? c.ldf : c.ndat) - s0;
    const int n_here = cover < BK_HSIZE ? (int)cover : BK_HSIZE;
    (void)left;
    const int nword = (n_here * ES + 3) / 4;
    unsigned char* dst = reinterpret_cast<unsigned char*>(c.freq) + (r * c.ldf + s0) * ES;
    // sum of the lanes as stored: a lane that wrapped loses 2^BITS from it, so the row total equals the number
    // of draws iff no lane overflowed (the test k_freq_rowsum_check makes in a second pass over the table)
    auto lanes = [](unsigned w) -> unsigned {
        return ES == 1 ? __dp4a(w, 0x01010101u, 0u) : (w & 0xffffu) + (w >> 16);
    };
    unsigned acc = 0;
    if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        const int nq = nword / 4;
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        const uint4* h4 = reinterpret_cast<const uint4*>(h);
        for (int i = threadIdx.x; i < nq; i += BK_THREADS) {
            const uint4 q = h4[i];
            d4[i] = q;
            acc += lanes(q.x) + lanes(q.y) + lanes(q.z) + lanes(q.w);
        }
        unsigned* d1 = reinterpret_cast<unsigned*>(dst);
        for (int i = nq * 4 + threadIdx.x; i < nword; i += BK_THREADS) {
            d1[i] = h[i];
            acc += lanes(h[i]);
        }
    } else {
        unsigned* d1 = reinterpret_cast<unsigned*>(dst);
        for (int i = threadIdx.x; i < nword; i += BK_THREADS) {
            d1[i] = h[i];
            acc += lanes(h[i]);
        }
    }
    if (c.rowacc) {
        acc = __reduce_add_sync(0xffffffffu, acc);
        if ((threadIdx.x & 31) == 0 && acc) atomicAdd(c.rowacc + r, (unsigned long long)acc);
    }
}

// rowacc[r] (accumulated by k_bucket_hist) must equal the draws of a replicate: else a narrow lane overflowed
__global__ void k_rowacc_verify(const unsigned long long* __restrict__ rowacc, int64_t nrows, long long expect,
                                int* __restrict__ status) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < nrows && rowacc[r] != (unsigned long long)expect) atomicOr(status, 2);
}

}  // namespace txb

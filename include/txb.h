/* txb.h -- C ABI of libtxb.so: B200 (sm_100a) kernels for thermoextrap's hot path.
 *
 * This is the drop-in boundary.  The reference (usnistgov/thermoextrap, pure
 * Python) reaches its numeric backend, the third-party package `cmomy`
 * (numba, CPU), through the call sites cited on each entry point below
 * (paths relative to /root/reference/src/thermoextrap).  A maintainer binds
 * these symbols with ctypes; no torch / numpy types
 * appear in any signature.
 *
 * Conventions
 *  - every function returns an int status (TXB_OK == 0); on failure
 *    txb_last_error_string() describes the problem.  No exceptions cross
 *    the boundary.
 *  - all data pointers are DEVICE pointers (fp64 unless stated), caller
 *    owned; inputs are const, outputs fully overwritten; the library never
 *    allocates device memory: scratch comes from the caller's `workspace`
 *    (size from the matching *_workspace_bytes query).
 *  - work is enqueued on `stream` (a cudaStream_t passed as void*; NULL =
 *    legacy default stream); no hidden synchronisation.
 *  - samples ("rec") are rows: x[nsamp, ldx] row-major, channel ("val")
 *    fastest, nchan <= ldx.  u is u[nsamp] when ldu == 0, else u[nsamp, ldu]
 *    (one energy per channel).  w is w[nsamp] or NULL.
 *  - comoment layout (data.py:1233-1236): out[..., 2, order+1] with
 *    [0,0]=sum w, [1,0]=<x>, [0,1]=<u>, else <dx^i du^j>.
 *
 * Intermediate "pivot sums" (additive across thread blocks, GPUs and streamed
 * chunks that share one pivot):
 *    pivot = { cx[nchan], cu[nu] }                nu = 1 (ldu==0) or nchan
 *    S0[nrep][nu   ][K+1] = sum_s f w (u-cu)^j
 *    S1[nrep][nchan][K+1] = sum_s f w (x-cx) (u-cu)^j
 * txb_finalize_f64 re-centres them into the cmomy layout.
 */
#ifndef TXB_H
#define TXB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    TXB_OK = 0,
    TXB_ERR_BAD_ARG = 1,
    TXB_ERR_UNSUPPORTED = 2,
    TXB_ERR_CUDA = 3,
    TXB_ERR_WORKSPACE = 4,
    TXB_ERR_OVERFLOW = 5
};

/* element type of a frequency table */
enum { TXB_I64 = 0, TXB_I32 = 1, TXB_U16 = 2, TXB_U8 = 3 };

/* finalize modes */
enum {
    TXB_MODE_COMOMENTS = 0, /* S0 (order K) + S1 (order K) -> [nchan,2,K+1]           */
    TXB_MODE_X_IS_U = 1     /* S0 only, order K+1 -> moments_to_comoments(mom=(1,K))  */
};

int txb_version(void);
const char* txb_last_error_string(void);
/* highest supported `order` (u-moment order of the accumulated sums) */
int txb_max_order(void);

/* ---- pivot (cheap strided pre-pass, <= 1024 rows) ------------------------
 * pivot[0:nchan] ~ mean x per channel (omitted when x == NULL),
 * pivot[nchan : nchan+nu] ~ mean u.  Any value near the mean works; sums are
 * exact for ANY pivot up to rounding, the pivot only controls conditioning. */
int txb_pivot_f64(const double* x, const double* u, int64_t nsamp, int64_t nchan, int64_t ldx,
                  int64_t ldu, double* pivot, void* stream);

/* ---- K1: reduce_vals ------------------------------------------------------
 * replaces cmomy.wrap_reduce_vals(x, u, weight=, mom=(1,order), dim=rec)
 * data.py:485-489, 528-532, 1183-1203, 1632-1640; lnpi.py:278-282.          */
size_t txb_reduce_workspace_bytes(int64_t nsamp, int64_t nchan, int order);
int txb_reduce_sums_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                        int64_t nchan, int64_t ldx, int64_t ldu, int order, const double* pivot,
                        double* S0, double* S1, void* workspace, size_t ws_bytes, void* stream);
/* x == NULL -> u-only sums (S1 untouched).  */

int txb_finalize_f64(const double* S0, const double* S1, const double* pivot, int64_t nrep,
                     int64_t nchan, int64_t nu, int order, int mode, double* out, void* stream);

/* pivot + sums + finalize in one call; `pivot_scratch` = nchan + nu doubles,
 * carved from the workspace.  mode TXB_MODE_X_IS_U: x must be NULL and the
 * sums run to order+1 (data.py:1182-1191).                                    */
int txb_reduce_vals_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                        int64_t nchan, int64_t ldx, int64_t ldu, int order, int mode, double* out,
                        void* workspace, size_t ws_bytes, void* stream);

/* txb_reduce_vals_static_f64: txb_reduce_vals_f64 WITHOUT a workspace argument and without any allocation.
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch is a file-scope
 * `static __device__` array of TXB_STATIC_SCRATCH_BYTES (16 MiB) inside the library, held by every GPU that
 * loads the library until the process exits; shapes whose txb_reduce_workspace_bytes() exceeds it are rejected
 * with TXB_ERR_WORKSPACE (16 MiB covers scalar x at any order and 64 channels at order 6; 2000 channels do not
 * fit).  Same kernels, same results as txb_reduce_vals_f64.  All calls on a device share the array (it keeps the
 * last call's contents; nothing needs clearing, every word read is written earlier in the same call): calls that
 * OVERLAP on the GPU -- two streams, two host threads -- race; serialise them per device.               */
enum { TXB_STATIC_SCRATCH_BYTES = 16 << 20 };
int txb_reduce_vals_static_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                               int64_t nchan, int64_t ldx, int64_t ldu, int order, int mode, double* out,
                               void* stream);

/* ---- K1, planar channels: x_0[nsamp] .. x_{C-1}[nsamp] as SEPARATE arrays (1 <= C <= 4) next to u[nsamp].
 * The volume path (reference volume.py:92-134: VolumeDataCallback.dxdq mean; :137-209 factory_extrapmodel's
 * DataValues.from_vals raw moments) needs <x>, <W>, <xW>, <dxdq> from three per-sample streams; one pass over
 * all of them (24 B / sample at C = 2) instead of two K1 passes that both read W.  Additive pivot sums about
 * pivot = [cx_0 .. cx_{C-1}, cu]:  S0[order+1] = sum du^j,  S1[C][order+1] = sum (x_c - cx_c) du^j; feed them
 * to txb_finalize_f64 (nchan = C, nu = 1).  Arrays must be 16-byte aligned (cp.async.bulk staging).          */
size_t txb_reduce_planar_workspace_bytes(int nplanes, int order);
int txb_reduce_sums_planar_f64(const double* const* x_planes, int nplanes, const double* u, int64_t nsamp,
                               int order, const double* pivot, double* S0, double* S1, void* workspace,
                               size_t ws_bytes, void* stream);

/* ---- K6: indices -> freq -------------------------------------------------
 * replaces cmomy's indices_to_freq inside factory_sampler
 * (data.py:420-421, 1028-1037, 1344-1352, 1782-1789).  Integer, bit-exact.
 * indices[nrep, nsamp] int64 in [0, ndat);  freq[nrep, ldf] of `freq_dtype`
 * is zero-filled then incremented.  *status (device int, may be NULL) is set
 * non-zero on out-of-range indices (1) or narrow-type overflow (2).          */
int txb_indices_to_freq(const int64_t* indices, int64_t nrep, int64_t nsamp, int64_t ndat,
                        void* freq, int freq_dtype, int64_t ldf, int* status, void* stream);

/* ---- K2: resample_vals ----------------------------------------------------
 * replaces cmomy.wrap_resample_vals(x, u, weight=, sampler=, mom=(1,order))
 * data.py:1354-1366, 1803-1810.  freq[nrep, ldf] is the sampler's table
 * (int64 unchanged, or a narrowed copy).  Counts must be < 2^31.
 * The contraction S[r, col] = sum_s freq[r, s] * P[s, col] runs on the fp64 tensor
 * cores (DMMA.8x8x4) for more than 4 channels and for scalar x with 4/8-byte tables
 * (or little column padding), else on a register-tiled DFMA kernel; environment
 * TXB_RESAMPLE_MMA=0 forces the DFMA kernels (A/B measurements).  Rows that are
 * 16-byte aligned (freq pointer and ldf * element size) take the cp.async path. */
size_t txb_resample_workspace_bytes(int64_t nsamp, int64_t nchan, int order, int64_t nrep);
int txb_resample_sums_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                          int64_t nchan, int64_t ldx, int64_t ldu, int order, const double* pivot,
                          const void* freq, int freq_dtype, int64_t nrep, int64_t ldf, double* S0,
                          double* S1, void* workspace, size_t ws_bytes, void* stream);
int txb_resample_vals_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                          int64_t nchan, int64_t ldx, int64_t ldu, int order, int mode,
                          const void* freq, int freq_dtype, int64_t nrep, int64_t ldf, double* out,
                          void* workspace, size_t ws_bytes, void* stream);

/* ---- K2 (index form): bootstrap straight from the sampler's index table ------
 * scalar x only (x[ndat] with stride ldx, or x == NULL for x_is_u): one random
 * 16-byte gather per draw instead of one atomic per draw + a contraction.
 * indices[nrep, ldi] int64, ndraw draws per replicate, values in [0, ndat)
 * (data.py:420-431 gather semantics == data.py:1803-1810 freq semantics,
 * pinned by tests/test_data.py:94-112).  *status (device int, may be NULL) is
 * set to 1 when an index is out of range.  nrep <= 65535 per call.             */
size_t txb_resample_indices_workspace_bytes(int64_t ndat, int order, int64_t nrep, int64_t ndraw);
int txb_resample_indices_sums_f64(const double* x, const double* u, const double* w, int64_t ndat,
                                  int64_t ldx, int order, const double* pivot,
                                  const int64_t* indices, int64_t nrep, int64_t ndraw, int64_t ldi,
                                  double* S0, double* S1, int* status, void* workspace,
                                  size_t ws_bytes, void* stream);
int txb_resample_indices_f64(const double* x, const double* u, const double* w, int64_t ndat,
                             int64_t ldx, int order, int mode, const int64_t* indices,
                             int64_t nrep, int64_t ndraw, int64_t ldi, double* out, int* status,
                             void* workspace, size_t ws_bytes, void* stream);

/* ---- K6 on the device: the reference sampler's stream, bit-exact ----------------
 * Reproduces numpy Generator(PCG64).integers(0, ndat, size) == .choice(ndat, size,
 * replace=True), the call cmomy.factory_sampler / thermoextrap make for {"nrep": n}
 * specs (adaptive_interp.py:543-544, tests/test_data.py:98-100, data.py:1782-1789),
 * for ndat <= 2^32 - 1, continuing from the generator state *rng
 * (bit_generator.state: state, inc, has_uint32, uinteger).  rng is a HOST struct.
 * *raw_used (device int64, may be NULL) receives the number of raw 32-bit draws the
 * outputs consumed; txb_pcg64_advance (host only, no GPU work) moves *rng past them
 * so the caller's numpy Generator continues exactly where numpy would have.
 * *status (device int, may be NULL): 2 = narrow table overflow, 4 = internal raw-range
 * estimate too short (never observed; > 12 sigma).
 * txb_sampler_prepare runs the counting pass once (and reports raw_used) into the first
 * txb_sampler_workspace_bytes(ndat, count) bytes of `workspace`; a later emit call given the SAME
 * (rng, ndat, count) and that workspace with prepared != 0 skips the pass.
 *   txb_sampler_indices : indices[count] int64              (IndexSampler.indices)
 *   txb_sampler_freq    : freq[nrep, ldf] of freq_dtype     (IndexSampler.freq)
 *   txb_sampler_resample_f64 : K2 fused with the generator for scalar x -- neither
 *       table is materialised; out[nrep, 2, order+1] as txb_resample_indices_f64.  */
typedef struct txb_pcg64 {
    uint64_t state_hi, state_lo, inc_hi, inc_lo;
    uint32_t has_uint32, uinteger;
} txb_pcg64;
/* head of every sampler workspace: block counts / offsets and the per-iteration position table */
size_t txb_sampler_workspace_bytes(int64_t ndat, int64_t count);
/* out[0] = entries of block_count (int64 at byte 64), out[1] = entries of iter_count (int32),
 * out[2] = byte offset of iter_count, out[3] = txb_sampler_workspace_bytes(ndat, count)      */
int txb_sampler_layout(int64_t ndat, int64_t count, int64_t* out);
int txb_pcg64_advance(txb_pcg64* rng, int64_t raw_used);
int txb_sampler_prepare(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t* raw_used,
                        int* status, void* workspace, size_t ws_bytes, void* stream);
/* Multi-GPU form of txb_sampler_prepare: the counting pass is split by thread block.  Rank g runs
 * txb_sampler_count on blocks [block_lo, block_hi) of txb_sampler_nblocks(...) (entries outside
 * the range are zeroed), the ranks SUM the two count arrays of the head (block_count, int64, and
 * iter_count, int32: see txb_sampler_layout; one small all-reduce each), and txb_sampler_scan
 * finishes the preparation.
 * block_hi < 0 means "to the end".                                                        */
int txb_sampler_nblocks(const txb_pcg64* rng, int64_t ndat, int64_t count);
int txb_sampler_count(const txb_pcg64* rng, int64_t ndat, int64_t count, int block_lo, int block_hi,
                      void* workspace, size_t ws_bytes, void* stream);
int txb_sampler_scan(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t* raw_used,
                     int* status, void* workspace, size_t ws_bytes, void* stream);
/* Emit calls take a window so that ranks can own slices of ONE numpy table:
 * positions [pos_lo, pos_hi) of the flattened table / replicates [rep_lo, rep_hi); outputs hold
 * only the window (indices[pos_hi - pos_lo], freq[rep_hi - rep_lo, ldf], out[rep_hi - rep_lo, ...]). */
int txb_sampler_indices(const txb_pcg64* rng, int64_t ndat, int64_t count, int64_t pos_lo,
                        int64_t pos_hi, int64_t* indices, int64_t* raw_used, int* status,
                        int prepared, void* workspace, size_t ws_bytes, void* stream);
/* txb_sampler_freq `prepared` is a flag word (TXB_SAMPLER_PREPARED | TXB_SAMPLER_NO_BUCKETS, below).
 * With only the head as workspace every draw is a global RED.ADD on the table (2.2e10 draws/s on a
 * B200).  Given txb_sampler_freq_workspace_bytes (or part of it: replicates go in windows) and a
 * uint8 / uint16 table, draws are partitioned by 65520-sample bucket through shared memory and every
 * (bucket, replicate) slice is counted in a shared-memory histogram and stored coalesced
 * (sampler_bucket.cuh).  *status |= 8 / 16 as for txb_sampler_resample_f64 (rerun with NO_BUCKETS);
 * |= 2 when a narrow lane overflowed (checked through the row sums: every row must sum to nsamp). */
size_t txb_sampler_freq_workspace_bytes(int64_t ndat, int64_t nrep, int64_t nsamp, int64_t nrep_out);
int txb_sampler_freq(const txb_pcg64* rng, int64_t ndat, int64_t nrep, int64_t nsamp,
                     int64_t rep_lo, int64_t rep_hi, void* freq, int freq_dtype, int64_t ldf,
                     int64_t* raw_used, int* status, int prepared, void* workspace, size_t ws_bytes,
                     void* stream);
/* Column-range form for SAMPLE-SHARDED bootstraps (reference semantics data.py:1751-1813
 * with the samples of one state split over ranks): only draws whose index lies in [col_lo, col_hi) are counted;
 * freq[rep_hi - rep_lo, ldf >= col_hi - col_lo] holds those columns (column c of the table = sample col_lo + c).
 * col_lo must be a multiple of TXB_FREQ_BUCKET; uint8 / uint16 tables on the bucketed path only (else
 * TXB_ERR_UNSUPPORTED: build the full table and slice it).  A column range cannot verify "every row sums to nsamp"
 * on its own, so rowsum[rep_hi - rep_lo] (int64, device) receives the sums of the first sum_cols columns of every
 * row: ranks whose ranges overlap (a range starts at the bucket boundary below the rank's first sample) each pass
 * the width up to the next rank's start, the caller adds the sums over the ranks (one integer all_reduce) and
 * compares with nsamp -- a wrapped lane still loses 2^bits from that total.                                        */
enum { TXB_FREQ_BUCKET = 65520 };
int txb_sampler_freq_cols(const txb_pcg64* rng, int64_t ndat, int64_t nrep, int64_t nsamp,
                          int64_t rep_lo, int64_t rep_hi, int64_t col_lo, int64_t col_hi, void* freq,
                          int freq_dtype, int64_t ldf, int64_t* rowsum, int64_t sum_cols, int* status,
                          int prepared, void* workspace, size_t ws_bytes, void* stream);
/* txb_sampler_resample_f64 `flags`: TXB_SAMPLER_PREPARED = the workspace head holds the result of
 * txb_sampler_prepare (or count + scan) for the same (rng, ndat, nrep * nsamp);
 * TXB_SAMPLER_NO_BUCKETS = always use the gather kernel.  With the minimum workspace
 * (txb_sampler_resample_workspace_bytes) the bootstrap gathers (x, u) pairs from HBM; given
 * txb_sampler_resample_fast_workspace_bytes (or as much of it as the caller can spare -- replicates
 * are processed in as many windows as needed) draws are first partitioned by sample bucket and
 * consumed from shared memory (sampler_bucket.cuh), which is ~2.5x faster on tables beyond L2.
 * *status |= 8 if a bucket list overflowed its > 10 sigma capacity, |= 16 if a shared-memory ring
 * wrapped (< 1e-12 per bucket and iteration): rerun with NO_BUCKETS.  |= 32 (informational) when
 * the bucketed path ran.                                                                   */
enum { TXB_SAMPLER_PREPARED = 1, TXB_SAMPLER_NO_BUCKETS = 2 };
size_t txb_sampler_resample_workspace_bytes(int64_t ndat, int order, int64_t nrep, int64_t nsamp,
                                            int64_t nrep_out);
size_t txb_sampler_resample_fast_workspace_bytes(int64_t ndat, int order, int mode, int64_t nrep,
                                                 int64_t nsamp, int64_t nrep_out);
int txb_sampler_resample_f64(const txb_pcg64* rng, const double* x, const double* u,
                             const double* w, int64_t ndat, int64_t ldx, int order, int mode,
                             int64_t nrep, int64_t nsamp, int64_t rep_lo, int64_t rep_hi, double* out,
                             int64_t* raw_used, int* status, int flags, void* workspace,
                             size_t ws_bytes, void* stream);

/* ---- K3 / K4: merge pre-reduced blocks (Chan/Pebay) -----------------------
 * data[nblock, nval, 2, order+1] (cmomy layout).
 * txb_reduce_data_f64   : CentralMomentsData.reduce(dim)          data.py:996
 * txb_resample_data_f64 : .resample_and_reduce(sampler)           data.py:1048-1052
 *                         freq[nrep, nblock] int64                            */
int txb_reduce_data_f64(const double* data, int64_t nblock, int64_t nval, int order, double* out,
                        void* stream);
int txb_resample_data_f64(const double* data, const int64_t* freq, int64_t nrep, int64_t nblock,
                          int64_t nval, int order, double* out, void* stream);

/* ---- K5: conversions on the (2, order+1) tail -------------------------------
 * kind 0: central -> raw   (.rmom(),  data.py:845-847)  out[0,0] = 1
 * kind 1: raw -> central   (convert.moments_type(to="central"), data.py:1108-1115)
 * kind 2: central -> cmom  (.cmom(),  data.py:850-852)  [0,0]=1,[1,0]=[0,1]=0
 * nelem = number of (2, order+1) tails.                                       */
int txb_convert_f64(const double* in, double* out, int64_t nelem, int order, int kind,
                    void* stream);

/* ---- K7: derivative expressions and Taylor / polynomial prediction -----------
 * txb_eval_exprs_f64 replaces  [funcs[i](*args) for i in range(order+1)]
 * (models.py:241-243, 371; expressions from beta.py:32-266, lnpi.py:42-116,
 * volume.py:63-78, -log transform models.py:261-281).
 *   vars[nvar][nelem] fp64; expression i = sum over terms t in
 *   [term_ptr[i], term_ptr[i+1]) of coef[t] * prod over factors f in
 *   [fac_ptr[t], fac_ptr[t+1]) of vars[fac_var[f]] ^ fac_pow[f]
 *   (integer power, may be negative; 127 = natural log).  out[nexpr][nelem].
 *   All table pointers are DEVICE pointers (int32 / fp64).
 * txb_poly_predict_f64 replaces the Taylor sum of ExtrapModel.predict
 * (models.py:535-565) and InterpModel.predict (models.py:929-932):
 *   coefs[np1][nelem], dalpha[nalpha] (device) ->
 *   mode 0: out[nalpha][nelem] = sum_p coefs[p] dalpha^p
 *   mode 1: out[nalpha][np1][nelem] running sum (cumsum); mode 2: individual terms. */
int txb_eval_exprs_f64(const double* vars, int64_t nelem, int nexpr, const int* term_ptr,
                       const double* coef, const int* fac_ptr, const int* fac_var,
                       const int* fac_pow, double* out, void* stream);
int txb_poly_predict_f64(const double* coefs, int np1, int64_t nelem, const double* dalpha,
                         int nalpha, int mode, double* out, void* stream);

/* ---- PerturbModel.predict (models.py:1009-1046; legacy/reweight.py:89-101) -----
 * Exponential reweighting from one state: for every dalpha[a] = alpha_a - alpha0 (DEVICE array)
 *   out[a, c] = sum_s e_a(s) x[s, c] / sum_s e_a(s),  e_a(s) = exp(-dalpha_a u_s - max_s(-dalpha_a u_s)),
 * x[nsamp, ldx] (first nchan columns), u[nsamp], out[nalpha, nchan].  One streaming pass over the
 * samples per tile of 4 alphas (plus a min/max pass over u). */
size_t txb_perturb_workspace_bytes(int64_t nsamp, int64_t nchan, int nalpha);
int txb_perturb_predict_f64(const double* x, const double* u, int64_t nsamp, int64_t nchan,
                            int64_t ldx, const double* dalpha, int nalpha, double* out,
                            void* workspace, size_t ws_bytes, void* stream);

/* ---- K7 device-resident: comoment records -> beta derivatives -> Taylor prediction, one launch ----------
 * Replaces, for the four numeric beta families, the chain  data.derivs_args (data.py:945-962) ->
 * funcs[i](*args) (models.py:371) -> minus_log (:373-374) -> / i! (:376-377) -> Taylor sum (:535-565)
 * with GENERATED device functions (csrc/beta_derivs_gen.cuh, built by csrc/gen_beta_derivs.py from the
 * recurrences of beta.py:32-266) reading the bootstrap kernels' output where it lies.
 *   comom   [nelem, 2, kdim] central comoment records in cmomy's layout ([0,1] = <u>, [0,n] = <du^n>,
 *           [1,0] = <x>, [1,n] = <dx du^n>), kdim >= order + 1
 *   family  0 x_ave central, 1 u_ave central, 2 x_ave raw, 3 u_ave raw (raw: converted in registers).  The u_ave
 *           families expect x_is_u records (x == u, beta.py:633-646): <du^n> is read as <dx du^(n-1)>
 *   flags   1 = -log transform (SymMinusLog), 2 = divide by i! (mode 3 only; modes 0-2 always do)
 *   mode    0 sum -> out[nalpha, nelem]; 1 cumsum / 2 terms -> out[nalpha, order+1, nelem];
 *           3 derivatives only -> out[order+1, nelem] (dalpha unused)
 *   dalpha  [nalpha] device doubles (alpha - alpha0).  order <= 10 (TXB_ERR_UNSUPPORTED above).     */
int txb_beta_predict_f64(const double* comom, int64_t nelem, int kdim, int family, int order, int flags,
                         int mode, const double* dalpha, int nalpha, double* out, void* stream);
/* The same launch for the lnPi family (lnpi.py:42-116: d_0 = lnPi0, d_1 = mu.N - <u>, d_n = -d^(n-1)<u>/dbeta^(n-1),
 * i.e. the u_ave family one order lower, shifted): comom = x_is_u records [nelem, 2, kdim] with kdim >= order;
 * central != 0 selects the central form; lnpi0 / mudotn [ninner] device doubles, element e reads index e % ninner
 * (leading dims such as rep broadcast).  flags: 2 = divide by i! in mode 3; 1 (-log) is rejected (the reference has
 * none for lnPi); modes as above.                                                                            */
int txb_lnpi_predict_f64(const double* comom, int64_t nelem, int kdim, int central, int order, int flags, int mode,
                         const double* dalpha, int nalpha, const double* lnpi0, const double* mudotn,
                         int64_t ninner, double* out, void* stream);
/* out[a, e] = (accumulate ? out[a, e] : 0) + sum_t W[a, t] D[t, e]: InterpModel.predict (models.py:909-932)
 * with W = alpha powers x inverse Hermite matrix folded on the host ([nalpha, nt], nt <= 512) and D = one
 * state's normalised derivatives [nt, nelem], both on the device.                                   */
int txb_combine_f64(const double* W, int nalpha, int nt, const double* D, int64_t nelem, int accumulate,
                    double* out, void* stream);
/* mean[a, b], var[a, b] over r of in[na, nrep, nb] (ddof delta degrees of freedom; xarray's default is 0):
 * stack.to_mean_var (stack.py:157-182) over the replicate axis of derivative arrays, on the device.   */
int txb_mean_var_f64(const double* in, int64_t na, int64_t nrep, int64_t nb, int ddof, double* mean,
                     double* var, void* stream);

/* ---- measurement helper: fp64 FMA peak micro-benchmark (bench.py roofline) ---
 * scratch: >= 1 device double; *flops (HOST pointer) = flop count of the launch. */
int txb_fp64_peak_launch(int iters, int blocks_per_sm, double* scratch, double* flops, void* stream);
/* the same pipe driven by a register-tiled contraction (TR x 14 accumulators, three register operands
 * per DFMA; tile 0: TR = 4, tile 1: TR = 2): the ceiling a k_resample-style micro-kernel can reach. */
int txb_fp64_tile_peak_launch(int iters, int blocks_per_sm, int tile, double* scratch, double* flops,
                              void* stream);
/* shared-memory atomicAdd (value returned) on `nbins` spread counters, 512-thread CTAs: the ceiling of the
 * slot reservation in k_bucket_scatter (one per draw).  *ops (HOST) = atomics issued by the launch. */
int txb_smem_atomic_peak_launch(int iters, int blocks_per_sm, int nbins, double* scratch, double* ops,
                                void* stream);

/* Which wide-x (> 4 channels) contraction kernel txb_resample_* launches: 0 = register-tiled DFMA,
 * 1 = k_resample_mma (DMMA, round 1), 2 = k_resample_mma2 (operand generation overlapped with the DMMAs,
 * 4-sample table loads; 8 warps), 3 = k_resample_mma2 with 16 warps, 4 (the default) = form 3 with uint8 table
 * entries converted through a shared-memory table instead of a DADD (other table dtypes: as form 3).
 * Same results within rounding; exists so tests and bench.py can A/B the forms in one process (also
 * TXB_RESAMPLE_MMA in the environment).  Returns the previous setting; other values only query.  No
 * counterpart in the reference (data.py:1803-1810 has one cmomy.resample_vals).                     */
int txb_set_resample_kernel(int version);

/* Threads per CTA of the draw-partition kernel (k_bucket_scatter) behind txb_sampler_resample_f64 and
 * txb_sampler_freq: 512 (default; 8 LCG steps per thread and iteration) or 1024 (4 steps; measured 2 % slower).  Identical
 * output (the lists hold the same draws); exists to A/B the forms in one process (also TXB_SCATTER_THREADS).
 * Returns the previous setting; other values only query.  No counterpart in the reference.             */
int txb_set_scatter_threads(int nthreads);

/* Kernel timing for bench.py's roofline: between txb_profile_begin(tag) and
 * txb_profile_end, every launch of the kernel family `tag` ("reduce", "resample",
 * "sampler_count", "sampler_gather"; "resample.wide" / "resample.scalar" = only the DMMA contraction of wide x /
 * of the scalar (u-moment) columns, both also counted under "resample"; "reduce_ring" = only the copy-engine-staged reduce
 * kernels, which is how the tests check which kernel ran) is bracketed by CUDA events on its own stream;
 * txb_profile_end synchronises on them and returns the summed duration and the number of
 * launches.  No effect on results; not thread safe (measurement only).            */
int txb_profile_begin(const char* tag);
int txb_profile_end(double* total_ms, int* nlaunch);

#ifdef __cplusplus
}
#endif
#endif /* TXB_H */

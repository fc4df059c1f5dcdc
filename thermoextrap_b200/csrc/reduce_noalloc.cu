// K1 (reduce_vals) for callers that cannot allocate: txb_reduce_vals_static_f64.
//
// No counterpart in the reference.  Some callers can not call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it) and have no
// workspace to hand in.  Same kernels and arithmetic as txb_reduce_vals_f64 (reduce.cu); the only difference is
// where the scratch lives: a file-scope `static __device__` array of TXB_STATIC_SCRATCH_BYTES that every GPU
// which loads the library holds until the process exits.  The call allocates nothing (no cudaMalloc / cudaFree /
// allocating library routine); inputs and outputs are still the caller's.
//
// The array's name in HOST code is a host-side stand-in, not a device address: handing `g_static_scratch` itself
// to a kernel would fault (or silently touch host memory where the GPU can reach it).  The device address is
// fetched once per device, with that device current, through cudaGetSymbolAddress below.
//
// Sharing: every call on a device uses the same array and finds the previous call's contents in it.  Nothing here
// needs a cleared scratch -- each word a kernel of this path reads (pivot, block partials, summed S0 / S1) is
// written earlier in the SAME call (tests/test_zz_no_alloc.py runs two differently shaped calls back to back) --
// but two calls that overlap on the GPU (different streams, or two host threads on one device) race on it:
// callers must serialise calls to this entry point per device.
#include <mutex>

#include "txb_common.cuh"

namespace txb {

static __device__ __align__(256) unsigned char g_static_scratch[TXB_STATIC_SCRATCH_BYTES];

static std::mutex g_static_mu;
static void* g_static_ptr[64] = {};   // device address of g_static_scratch, per device ordinal

// Device address of the scratch array on the CURRENT device (looked up once per device).
static int static_scratch_ptr(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("txb_reduce_vals_static_f64: cudaGetDevice failed or device ordinal %d out of range: %s", dev,
                  cudaGetErrorString(e));
        cudaGetLastError();
        return TXB_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);   // the symbol, resolved for the current device
        if (e != cudaSuccess || !p) {
            set_error("txb_reduce_vals_static_f64: cudaGetSymbolAddress failed on device %d: %s", dev,
                      cudaGetErrorString(e));
            cudaGetLastError();
            return TXB_ERR_CUDA;
        }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return TXB_OK;
}

}  // namespace txb

using namespace txb;

extern "C" int txb_reduce_vals_static_f64(const double* x, const double* u, const double* w, int64_t nsamp,
                                          int64_t nchan, int64_t ldx, int64_t ldu, int order, int mode,
                                          double* out, void* stream) {
    TXB_CHECK_ARG(u && out, "txb_reduce_vals_static_f64: null pointer");
    TXB_CHECK_ARG(nsamp > 0 && order >= 1, "txb_reduce_vals_static_f64: nsamp=%lld, order=%d must be >= 1",
                  (long long)nsamp, order);
    const size_t need = txb_reduce_workspace_bytes(nsamp, (!x && ldu == 0) ? 1 : nchan, order);
    if (need > (size_t)TXB_STATIC_SCRATCH_BYTES) {
        set_error("txb_reduce_vals_static_f64: this shape (nsamp=%lld, nchan=%lld, order=%d) needs %zu bytes of scratch, "
                  "the built-in array holds %d; use txb_reduce_vals_f64 with a caller-provided workspace",
                  (long long)nsamp, (long long)nchan, order, need, (int)TXB_STATIC_SCRATCH_BYTES);
        return TXB_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    const int rc = static_scratch_ptr(&scratch);
    if (rc != TXB_OK) return rc;
    return txb_reduce_vals_f64(x, u, w, nsamp, nchan, ldx, ldu, order, mode, out, scratch,
                               (size_t)TXB_STATIC_SCRATCH_BYTES, stream);
}

"""txb_reduce_vals_static_f64 -- the entry point that allocates nothing (scratch = a static __device__ array inside
libtxb.so; no counterpart in the reference).

Called through the C ABI (ctypes) and compared with the existing txb_reduce_vals_f64 on the same inputs, on two
differently shaped calls in a row (the second finds the first one's contents in the shared scratch), within the
tolerance of the other GPU tests: ``|a - b| <= 1e-10 * max(|b|, sigma_x^i sigma_u^j)``.  Collected last on purpose
(a failure here must not hide other tests); passes alone in a fresh process and after the rest.
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _pair(lib, _lib, torch, x, u, w, order, mode):
    """(static-scratch result, workspace result) of one reduce through the raw C ABI."""
    dev = torch.device("cuda", torch.cuda.current_device())
    ud = torch.as_tensor(u, device=dev)
    xd = None if x is None else torch.as_tensor(x, device=dev)
    wd = None if w is None else torch.as_tensor(w, device=dev)
    nsamp = ud.shape[0]
    nchan = 1 if xd is None or xd.ndim == 1 else xd.shape[1]
    ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    out_s = torch.full((nchan, 2, order + 1), float("nan"), dtype=torch.float64, device=dev)
    out_w = torch.full((nchan, 2, order + 1), float("nan"), dtype=torch.float64, device=dev)
    nb = lib.txb_reduce_workspace_bytes(nsamp, nchan, order)
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    st = lib.txb_reduce_vals_static_f64(ptr(xd), ptr(ud), ptr(wd), nsamp, nchan, nchan, 0, order, mode,
                                        out_s.data_ptr(), None)
    _lib.check(st, "txb_reduce_vals_static_f64")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before   # nothing went through torch's allocator either
    st = lib.txb_reduce_vals_f64(ptr(xd), ptr(ud), ptr(wd), nsamp, nchan, nchan, 0, order, mode,
                                 out_w.data_ptr(), ws.data_ptr(), ws.numel(), None)
    _lib.check(st, "txb_reduce_vals_f64")
    torch.cuda.synchronize()
    return out_s.cpu().numpy(), out_w.cpu().numpy()


def test_static_scratch_reduce_matches_workspace_reduce_twice():
    import torch

    from oracle import comoments as oc
    from thermoextrap_b200 import _lib

    lib = _lib.load()
    rng = np.random.default_rng(0)
    # call 1: wide x (64 channels, order 6: 9.3 MiB of the 16 MiB scratch)
    n1 = 200_003
    u1 = rng.normal(175.0, 5.3, n1)
    x1 = 1e-3 * u1[:, None] + rng.normal(0.0, 0.02, (n1, 64))
    # call 2, straight after: scalar x, weighted, another order -- a different carve-up of the same scratch
    n2 = 1_000_001
    u2 = rng.normal(-3.0, 2.0, n2)
    x2 = 0.5 * u2 + rng.normal(0.0, 1.0, n2)
    w2 = rng.random(n2) + 0.5
    for x, u, w, order in ((x1, u1, None, 6), (x2, u2, w2, 4), (x1, u1, None, 6)):
        got, ref = _pair(lib, _lib, torch, x, u, w, order, _lib.MODE_COMOMENTS)
        assert np.isfinite(got).all()
        scale = oc.tolerance_scale(x, u, order)
        scale = scale.reshape(ref.shape) if np.ndim(scale) else scale
        err = np.abs(got - ref) / np.maximum(np.abs(ref), scale)
        assert err.max() <= TOL, f"static-scratch reduce differs from the workspace reduce: {err.max():.3e}"
        # and both are right (oracle: longdouble two-pass)
        exact = oc.comoments_two_pass(x, u, order, w=w).reshape(ref.shape)
        assert (np.abs(got - exact) / np.maximum(np.abs(exact), scale)).max() <= TOL
    # x_is_u mode (x == NULL)
    got, ref = _pair(lib, _lib, torch, None, u2, None, 5, _lib.MODE_X_IS_U)
    sc = oc.tolerance_scale(u2, u2, 5)
    assert (np.abs(got - ref) / np.maximum(np.abs(ref), sc)).max() <= TOL


def test_static_scratch_rejects_oversize_shapes():
    import torch

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    n, c, order = 4096, 2000, 4   # needs ~27 MiB of partials: beyond the 16 MiB array
    assert lib.txb_reduce_workspace_bytes(n, c, order) > 16 << 20
    x = torch.zeros((n, c), dtype=torch.float64, device="cuda")
    u = torch.zeros(n, dtype=torch.float64, device="cuda")
    out = torch.zeros((c, 2, order + 1), dtype=torch.float64, device="cuda")
    st = lib.txb_reduce_vals_static_f64(x.data_ptr(), u.data_ptr(), None, n, c, c, 0, order, _lib.MODE_COMOMENTS,
                                        out.data_ptr(), None)
    assert st == 4  # TXB_ERR_WORKSPACE
    msg = lib.txb_last_error_string().decode()
    assert "built-in array" in msg and "txb_reduce_vals_f64" in msg

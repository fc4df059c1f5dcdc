"""Parity of the CUDA kernels (through the C ABI) against the CPU oracle.

Tolerance: ``|a - b| <= 1e-10 * max(|b|, sigma_x^i sigma_u^j)``
for fp64 comoments; integer tables bit-exact.
"""

import numpy as np
import pytest

from oracle import comoments as oc

pytestmark = pytest.mark.gpu

TOL = 1e-10


@pytest.fixture(scope="module")
def K():
    import torch

    from thermoextrap_b200 import kernels

    assert torch.cuda.is_available()
    return kernels


def relerr(a, ref, scale):
    return float((np.abs(a - ref) / np.maximum(np.abs(ref), scale)).max())


def make(n, nv, seed=0, offset=175.0):
    rng = np.random.default_rng(seed)
    u = rng.normal(offset, 5.3, n)
    if nv == 0:
        x = 1e-3 * u + rng.normal(0.0, 0.02, n)
    else:
        a = rng.uniform(0.5, 2.0, nv)
        x = 1e-3 * a * u[:, None] + rng.normal(0.0, 0.02, (n, nv))
    return x, u, rng


# ------------------------------------------------------------------ K1
@pytest.mark.parametrize("order", [1, 2, 5, 6, 8, 10])
@pytest.mark.parametrize("n,nv", [(100, 5), (4097, 0), (100_000, 0), (3000, 64), (513, 7), (300, 2000), (257, 2)])
def test_reduce_vals(K, n, nv, order):
    x, u, _ = make(n, nv)
    out = K.reduce_vals(x, u, order).cpu().numpy()
    ref = oc.comoments_two_pass(x, u, order)
    assert out.shape == ref.shape
    assert relerr(out, ref, oc.tolerance_scale(x, u, order)) < TOL


@pytest.mark.parametrize("n,nv", [(1000, 0), (1001, 6), (777, 3)])
def test_reduce_vals_weighted(K, n, nv):
    x, u, rng = make(n, nv, seed=3)
    w = rng.random(n)
    w[::7] = 0.0
    out = K.reduce_vals(x, u, 6, w=w).cpu().numpy()
    ref = oc.comoments_two_pass(x, u, 6, w=w)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 6)) < TOL


def test_reduce_vals_unaligned_views(K):
    import torch

    x, u, _ = make(10_001, 0, seed=5)
    xd = torch.from_numpy(np.concatenate([[0.0], x])).cuda()[1:]
    ud = torch.from_numpy(np.concatenate([[0.0], u])).cuda()[1:]
    assert xd.data_ptr() % 16 == 8
    out = K.reduce_vals(xd, ud, 7).cpu().numpy()
    ref = oc.comoments_two_pass(x, u, 7)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 7)) < TOL


def test_reduce_vals_fixture_uniform(K):
    """The reference's FixtureData shape (tests/conftest.py:15-28): U(0,1), n=100, nv=5, order 5."""
    rng = np.random.default_rng(0)
    u = rng.random(100)
    x = rng.random((100, 5))
    out = K.reduce_vals(x, u, 5).cpu().numpy()
    ref = oc.comoments_two_pass(x, u, 5)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 5)) < TOL
    # tests/test_data.py:7-38: raw moments == np.average(x * u**i)
    raw = K.convert(out, K.CONVERT_CENTRAL_TO_RAW).cpu().numpy()
    for i in range(6):
        np.testing.assert_allclose(raw[:, 0, i], np.average(u**i), rtol=1e-12)
        np.testing.assert_allclose(raw[:, 1, i], np.average(x * u[:, None] ** i, axis=0), rtol=1e-12)


# ------------------------------------------------------------------ K1, copy-engine ring kernels
def ring_launches(fn):
    """Run fn; return (result, number of cp.async.bulk-staged reduce kernels it launched)."""
    import ctypes as C

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    lib.txb_profile_begin(b"reduce_ring")
    out = fn()
    ms, nl = C.c_double(), C.c_int()
    lib.txb_profile_end(C.byref(ms), C.byref(nl))
    return out, nl.value


# shapes large enough that the ring path is taken (reduce_wide_tma.cu:plan_wide_ring): 256-thread whole-row
# tiles (<= 128 channels), 512-thread tiles (<= 2048 channels, lane counts that are not powers of two),
# two channels per thread at order >= 10, ragged last tiles
@pytest.mark.parametrize("n,nv,order", [(20_001, 64, 6), (3001, 2000, 4), (6000, 130, 6), (40_003, 16, 8),
                                        (9000, 500, 12), (330_000, 2, 3), (5000, 1022, 5), (12_345, 36, 10)])
def test_reduce_vals_wide_ring(K, n, nv, order):
    x, u, _ = make(n, nv, seed=11)
    out, nring = ring_launches(lambda: K.reduce_vals(x, u, order))
    assert nring == 1
    ref = oc.comoments_two_pass(x, u, order)
    assert relerr(out.cpu().numpy(), ref, oc.tolerance_scale(x, u, order)) < TOL


def test_reduce_vals_wide_ring_weighted_strided_upc(K):
    import torch

    x, u, rng = make(20_000, 64, seed=12)
    w = rng.random(20_000)
    w[::5] = 0.0
    ref = oc.comoments_two_pass(x, u, 6, w=w)
    scale = oc.tolerance_scale(x, u, 6)
    out, nring = ring_launches(lambda: K.reduce_vals(x, u, 6, w=w))
    assert nring == 1 and relerr(out.cpu().numpy(), ref, scale) < TOL
    # rows with padding, straight through the C ABI (the Python layer always packs): ld = 72 is streamed
    # by whole-row copies that carry the padding along, ld = 128 stays on the register-staged kernel
    from thermoextrap_b200 import _lib

    lib = _lib.load()
    ud, wd = torch.from_numpy(u).cuda(), torch.from_numpy(w).cuda()
    for ld, expect in [(72, 1), (128, 0)]:
        buf = torch.full((20_000, ld), float("nan"), dtype=torch.float64, device="cuda")
        buf[:, :64] = torch.from_numpy(x).cuda()
        outd = torch.empty((64, 2, 7), dtype=torch.float64, device="cuda")
        nb = lib.txb_reduce_workspace_bytes(20_000, 64, 6)
        ws = torch.empty(nb, dtype=torch.uint8, device="cuda")

        def call():
            _lib.check(lib.txb_reduce_vals_f64(buf.data_ptr(), ud.data_ptr(), wd.data_ptr(), 20_000, 64, ld, 0, 6, 0,
                                               outd.data_ptr(), ws.data_ptr(), nb, None), "txb_reduce_vals_f64")
            torch.cuda.synchronize()
            return outd

        out, nring = ring_launches(call)
        assert nring == expect and relerr(out.cpu().numpy(), ref, scale) < TOL
    # one energy per channel (both arrays wide) and x_is_u wide
    uu = u[:, None] + rng.normal(0, 1.0, x.shape)
    out, nring = ring_launches(lambda: K.reduce_vals(x, uu, 5))
    assert nring == 1
    assert relerr(out.cpu().numpy(), oc.comoments_two_pass(x, uu, 5), oc.tolerance_scale(x, uu, 5)) < TOL
    uw = np.ascontiguousarray(np.tile(uu[:8000], (1, 4))[:, :200]) + rng.normal(0, 0.5, (8000, 200))
    out, nring = ring_launches(lambda: K.reduce_vals(None, uw, 4, x_is_u=True))
    assert nring == 1
    refu = oc.comoments_two_pass(uw, uw, 4)
    assert relerr(out.cpu().numpy(), refu, oc.tolerance_scale(uw, uw, 4).reshape(refu.shape)) < TOL


@pytest.mark.parametrize("order,weighted", [(8, False), (9, True), (11, False)])
def test_reduce_vals_scalar_ring(K, order, weighted):
    n = 2_500_003  # ragged: 1220 full tiles of 2048 + 1443 samples
    x, u, rng = make(n, 0, seed=13)
    w = rng.random(n) if weighted else None
    out, nring = ring_launches(lambda: K.reduce_vals(x, u, order, w=w))
    assert nring == 1
    ref = oc.comoments_two_pass(x, u, order, w=w)
    assert relerr(out.cpu().numpy(), ref, oc.tolerance_scale(x, u, order)) < TOL
    out, nring = ring_launches(lambda: K.reduce_vals(None, u, order, w=w, x_is_u=True))
    assert nring == 1
    refu = oc.comoments_two_pass(u, u, order, w=w).reshape(out.shape)
    assert relerr(out.cpu().numpy(), refu, oc.tolerance_scale(u, u, order).reshape(out.shape)) < TOL


@pytest.mark.parametrize("n,nv", [(5000, 0), (2000, 40), (999, 3)])
@pytest.mark.parametrize("order", [1, 4, 9])
def test_reduce_vals_x_is_u(K, n, nv, order):
    _, u, rng = make(n, 0, seed=1)
    if nv:
        u = u[:, None] * rng.uniform(0.5, 1.5, nv) + rng.normal(0, 1.0, (n, nv))
    out = K.reduce_vals(None, u, order, x_is_u=True).cpu().numpy()
    ref = oc.comoments_two_pass(u, u, order)
    if nv == 0:
        ref = ref.reshape(out.shape)
    assert relerr(out, ref, oc.tolerance_scale(u, u, order).reshape(out.shape)) < TOL


def test_reduce_vals_per_channel_u(K):
    x, u, rng = make(1500, 6, seed=2)
    uu = u[:, None] + rng.normal(0, 1.0, x.shape)
    out = K.reduce_vals(x, uu, 5).cpu().numpy()
    ref = oc.comoments_two_pass(x, uu, 5)
    assert relerr(out, ref, oc.tolerance_scale(x, uu, 5)) < TOL


def test_reduce_vals_errors(K):
    x, u, _ = make(10, 0)
    with pytest.raises(ValueError):
        K.reduce_vals(x, u, 0)
    with pytest.raises(ValueError):
        K.reduce_vals(x, u, 40)
    with pytest.raises(ValueError):
        K.reduce_vals(x[:5], u, 3)


# ------------------------------------------------------------------ K6
@pytest.mark.parametrize("dtype", ["int64", "int32", "uint16", "uint8"])
@pytest.mark.parametrize("nrep,nsamp,ndat", [(10, 100, 100), (3, 1001, 777), (130, 4096, 4096)])
def test_indices_to_freq(K, dtype, nrep, nsamp, ndat):
    import torch

    rng = np.random.default_rng(4)
    idx = rng.choice(ndat, (nrep, nsamp), replace=True)
    freq = K.indices_to_freq(idx, ndat, dtype=getattr(torch, dtype))
    ref = oc.indices_to_freq(idx, ndat)
    got = freq.cpu().numpy().astype(np.int64)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)
    assert np.array_equal(got.sum(axis=1), np.full(nrep, nsamp))


def test_indices_to_freq_overflow_and_range(K):
    import torch

    idx = np.zeros((1, 300), dtype=np.int64)
    with pytest.raises(OverflowError):
        K.indices_to_freq(idx, 10, dtype=torch.uint8)
    assert int(K.indices_to_freq(idx, 10, dtype=torch.uint16)[0, 0]) == 300
    with pytest.raises(ValueError):
        K.indices_to_freq(np.full((1, 4), 10), 10)


# ------------------------------------------------------------------ K2
@pytest.mark.parametrize("dtype", ["int64", "int32", "uint16", "uint8"])
@pytest.mark.parametrize("n,nv,nrep,order", [(100, 5, 10, 5), (1000, 0, 7, 8), (4099, 0, 130, 3), (333, 20, 5, 6), (500, 3, 4, 10), (2048, 64, 3, 6)])
def test_resample_vals(K, dtype, n, nv, nrep, order):
    import torch

    x, u, rng = make(n, nv, seed=7)
    idx = rng.choice(n, (nrep, n), replace=True)
    freq = K.indices_to_freq(idx, n, dtype=getattr(torch, dtype))
    out = K.resample_vals(x, u, freq, order).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, order, idx)
    assert out.shape == ref.shape
    assert relerr(out, ref, oc.tolerance_scale(x, u, order)) < TOL


def test_resample_vals_host_int64_table_unaligned(K):
    """The reference sampler's own int64 table, odd ndat (rows not 16-byte aligned)."""
    x, u, rng = make(1001, 0, seed=8)
    idx = rng.choice(1001, (9, 1001), replace=True)
    freq = oc.indices_to_freq(idx, 1001)
    out = K.resample_vals(x, u, freq, 6).cpu().numpy()
    ref = oc.resample_vals_freq(x, u, 6, freq)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 6)) < TOL


def test_resample_vals_weighted_and_x_is_u(K):
    x, u, rng = make(2000, 4, seed=9)
    w = rng.random(2000)
    idx = rng.choice(2000, (6, 2000), replace=True)
    freq = oc.indices_to_freq(idx, 2000)
    out = K.resample_vals(x, u, freq, 5, w=w).cpu().numpy()
    ref = oc.resample_vals_freq(x, u, 5, freq, w=w)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 5)) < TOL
    # x_is_u scalar and wide
    out = K.resample_vals(None, u, freq, 4, x_is_u=True).cpu().numpy()
    ref = oc.resample_vals_freq(u, u, 4, freq).reshape(out.shape)
    assert relerr(out, ref, oc.tolerance_scale(u, u, 4).reshape(out.shape[1:])) < TOL
    uu = u[:, None] + rng.normal(0, 1.0, (2000, 6))
    out = K.resample_vals(None, uu, freq, 4, x_is_u=True).cpu().numpy()
    ref = oc.resample_vals_freq(uu, uu, 4, freq)
    assert relerr(out, ref, oc.tolerance_scale(uu, uu, 4)) < TOL


@pytest.mark.parametrize("dtype", ["uint8", "int64"])
@pytest.mark.parametrize("n,nv,nrep,order,kind", [
    (1500, 17, 150, 1, "x"), (1500, 17, 150, 8, "x"), (1203, 33, 70, 9, "xw"), (900, 16, 9, 12, "x"),
    (1500, 20, 140, 5, "upc"), (777, 9, 66, 10, "upc"),
])
def test_resample_vals_wide_tensor_core_tiles(K, dtype, n, nv, nrep, order, kind):
    """Wide x (> 4 channels) runs on the DMMA contraction (k_resample_mma): replicate tiles of 128 / 64
    with ragged edges, channel tiles of 16 with ragged edges, weights, and the per-channel-u form."""
    import torch

    x, u, rng = make(n, nv, seed=12)
    idx = rng.choice(n, (nrep, n), replace=True)
    freq = K.indices_to_freq(idx, n, dtype=getattr(torch, dtype))
    if kind == "upc":
        uu = u[:, None] + x
        out = K.resample_vals(None, uu, freq, order, x_is_u=True).cpu().numpy()
        sub = [0, nrep // 2, nrep - 1]
        ref = oc.resample_vals_gather(uu, uu, order, idx[sub])
        assert relerr(out[sub], ref, oc.tolerance_scale(uu, uu, order)) < TOL
        return
    w = rng.random(n) if kind == "xw" else None
    out = K.resample_vals(x, u, freq, order, w=w).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, order, idx, w=w)
    assert out.shape == ref.shape
    assert relerr(out, ref, oc.tolerance_scale(x, u, order)) < TOL


@pytest.mark.parametrize("dtype", ["int64", "int32", "uint16", "uint8"])
@pytest.mark.parametrize("n,nrep,order,kind", [
    (1001, 130, 1, "x"), (4099, 300, 6, "x"), (2500, 9, 7, "xw"), (777, 129, 12, "x"), (3000, 70, 3, "u"), (1500, 140, 11, "uw"),
])
def test_resample_vals_scalar_tensor_core_tiles(K, dtype, n, nrep, order, kind):
    """Scalar x from a frequency table (the reference sampler's own int64 table included) on the DMMA
    contraction (k_resample_mma_scalar) or the DFMA tile, whichever the dispatch picks: ragged replicate tiles,
    rows that are not 16-byte multiples, weights, x_is_u."""
    import torch

    x, u, rng = make(n, 0, seed=13)
    idx = rng.choice(n, (nrep, n), replace=True)
    freq = K.indices_to_freq(idx, n, dtype=getattr(torch, dtype))
    w = rng.random(n) if kind.endswith("w") else None
    sub = [0, 1, nrep // 2, nrep - 2, nrep - 1]
    if kind[0] == "u":
        out = K.resample_vals(None, u, freq, order, w=w, x_is_u=True).cpu().numpy()
        ref = oc.resample_vals_gather(u, u, order, idx[sub], w=w).reshape(len(sub), *out.shape[1:])
        scale = oc.tolerance_scale(u, u, order).reshape(out.shape[1:])
    else:
        out = K.resample_vals(x, u, freq, order, w=w).cpu().numpy()
        ref = oc.resample_vals_gather(x, u, order, idx[sub], w=w)
        scale = oc.tolerance_scale(x, u, order)
    assert out.shape[0] == nrep
    assert relerr(out[sub], ref, scale) < TOL


def test_resample_identity_freq_equals_reduce(K):
    """freq == 1 for every sample reproduces reduce_vals (size-independent property)."""
    x, u, _ = make(50_000, 8, seed=10)
    freq = np.ones((2, 50_000), dtype=np.int64)
    a = K.resample_vals(x, u, freq, 6).cpu().numpy()
    b = K.reduce_vals(x, u, 6).cpu().numpy()
    assert relerr(a[0], b, oc.tolerance_scale(x, u, 6)) < TOL
    assert np.array_equal(a[0], a[1])


@pytest.mark.parametrize("n,ndraw,nrep,order", [(1000, 1000, 7, 8), (4099, 4099, 13, 3), (100_000, 100_000, 5, 6), (501, 1203, 3, 5)])
def test_resample_indices_scalar(K, n, ndraw, nrep, order):
    x, u, rng = make(n, 0, seed=13)
    idx = rng.choice(n, (nrep, ndraw), replace=True)
    out = K.resample_indices(x, u, idx, order).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, order, idx)
    assert out.shape == ref.shape
    scale = oc.tolerance_scale(x, u, order)
    assert relerr(out, ref, scale) < TOL
    # same answer through the frequency-table path (tests/test_data.py:94-112)
    if ndraw == n:
        viaf = K.resample_vals(x, u, K.indices_to_freq(idx, n), order).cpu().numpy()
        assert relerr(viaf, ref, scale) < TOL


def test_resample_indices_weighted_x_is_u_and_errors(K):
    x, u, rng = make(3000, 0, seed=14)
    w = rng.random(3000)
    idx = rng.choice(3000, (4, 3000), replace=True)
    out = K.resample_indices(x, u, idx, 6, w=w).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, 6, idx, w=w)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 6)) < TOL
    out = K.resample_indices(None, u, idx, 5, x_is_u=True).cpu().numpy()
    ref = oc.resample_vals_gather(u, u, 5, idx)
    assert relerr(out, ref, oc.tolerance_scale(u, u, 5)) < TOL
    bad = idx.copy()
    bad[2, 17] = 3000
    with pytest.raises(ValueError):
        K.resample_indices(x, u, bad, 6)
    # odd row pitch -> unaligned rows take the scalar-load path
    idx2 = rng.choice(3000, (3, 2999), replace=True)
    out = K.resample_indices(x, u, idx2, 4).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, 4, idx2)
    assert relerr(out, ref, oc.tolerance_scale(x, u, 4)) < TOL


# ------------------------------------------------------------------ K3/K4/K5
def test_reduce_and_resample_data(K):
    x, u, rng = make(1000, 3, seed=11)
    blocks = np.stack([oc.comoments_two_pass(x[i * 100 : (i + 1) * 100], u[i * 100 : (i + 1) * 100], 6) for i in range(10)])
    out = K.reduce_data(blocks).cpu().numpy()
    ref = oc.comoments_two_pass(x, u, 6)
    scale = oc.tolerance_scale(x, u, 6)
    assert relerr(out, ref, scale) < TOL
    freq = oc.indices_to_freq(rng.choice(10, (20, 10)), 10)
    out = K.resample_data(blocks, freq).cpu().numpy()
    ref = oc.resample_blocks(blocks, freq)
    assert relerr(out, ref, scale) < TOL


def test_convert_roundtrip(K):
    x, u, _ = make(500, 4, seed=12, offset=3.0)
    c = oc.comoments_two_pass(x, u, 7)
    raw = K.convert(c, K.CONVERT_CENTRAL_TO_RAW).cpu().numpy()
    np.testing.assert_allclose(raw, oc.central_to_raw(c), rtol=1e-12)
    raw[..., 0, 0] = c[..., 0, 0]
    back = K.convert(raw, K.CONVERT_RAW_TO_CENTRAL).cpu().numpy()
    assert relerr(back, c, oc.tolerance_scale(x, u, 7)) < 1e-9  # raw moments at <u>=3 lose digits
    cm = K.convert(c, K.CONVERT_CENTRAL_TO_CMOM).cpu().numpy()
    assert np.all(cm[..., 0, 0] == 1) and np.all(cm[..., 0, 1] == 0) and np.all(cm[..., 1, 0] == 0)
    assert np.array_equal(cm[..., 0, 2:], c[..., 0, 2:])

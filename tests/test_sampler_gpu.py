"""Device sampler (K6): bit-exact against live numpy -- the reference's own sampler.

The reference draws ``rng.choice(ndat, (nrep, nsamp), replace=True)`` on a numpy
``Generator(PCG64)`` (adaptive_interp.py:543-544, tests/test_data.py:98-100); integer
tables must be identical, and the generator must be left in the state numpy would
leave it in (so later draws agree too).  The fused bootstrap is compared with the
oracle's gather-then-reduce on the numpy table (tests/test_data.py:94-112 semantics),
tolerance ``1e-10 * max(|ref|, sigma_x^i sigma_u^j)``.
"""

import numpy as np
import pytest

from oracle import comoments as oc

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def K():
    from thermoextrap_b200 import kernels

    return kernels


def _same_state(a, b):
    sa, sb = a.bit_generator.state, b.bit_generator.state
    return (sa["state"] == sb["state"] and sa["has_uint32"] == sb["has_uint32"]
            and (not sa["has_uint32"] or sa["uinteger"] == sb["uinteger"]))


CASES = [
    (2, 5, 1),  # ndat == 1: numpy returns zeros and consumes nothing
    (3, 100, 100),
    (10, 1000, 1000),
    (7, 4096, 4096),
    (5, 4097, 5000),
    (64, 100_000, 100_000),
    (2, 3_000_000, 3_000_000),
    (3, 1_000_003, 3_000_000_000),  # 30 % rejection
    (2, 2_000_000, 2_147_483_649),  # ~50 % rejection
    (4, 65_537, 100_000_000),       # 2.2 % rejection
]


@pytest.mark.parametrize("nrep,nsamp,ndat", CASES)
@pytest.mark.parametrize("odd_start", [False, True])
def test_indices_bit_exact(K, nrep, nsamp, ndat, odd_start):
    a, b = np.random.default_rng(11), np.random.default_rng(11)
    if odd_start:  # leave a buffered 32-bit half in the generator
        assert a.integers(0, 10, 3).tolist() == b.integers(0, 10, 3).tolist()
        assert a.bit_generator.state["has_uint32"] == 1
    ref = a.choice(ndat, size=(nrep, nsamp), replace=True)
    out = K.sampler_indices(b, nrep, nsamp, ndat).cpu().numpy()
    assert out.dtype == np.int64
    assert np.array_equal(out, ref)
    assert _same_state(a, b)
    # and the streams keep agreeing afterwards
    assert np.array_equal(a.integers(0, 1000, 5), b.integers(0, 1000, 5))


@pytest.mark.parametrize("dtype_name", ["uint8", "uint16", "int32", "int64"])
@pytest.mark.parametrize("nrep,nsamp,ndat", [(3, 100, 100), (9, 5000, 4097), (33, 100_000, 100_000), (2, 1_000_000, 999_999)])
def test_freq_bit_exact(K, nrep, nsamp, ndat, dtype_name):
    import torch

    a, b = np.random.default_rng(3), np.random.default_rng(3)
    idx = a.choice(ndat, size=(nrep, nsamp), replace=True)
    ref = oc.indices_to_freq(idx, ndat)
    out = K.sampler_freq(b, nrep, nsamp, ndat, dtype=getattr(torch, dtype_name)).cpu().numpy()
    assert np.array_equal(out.astype(np.int64), ref)
    assert out.sum(axis=1).tolist() == [nsamp] * nrep
    assert _same_state(a, b)


def test_freq_overflow_is_loud(K):
    import torch

    with pytest.raises(OverflowError):
        K.sampler_freq(np.random.default_rng(0), 1, 5000, 3, dtype=torch.uint8)


@pytest.mark.parametrize("order", [1, 6, 8])
@pytest.mark.parametrize("nrep,n", [(5, 100), (16, 4096), (7, 10_000), (3, 300_000)])
def test_fused_resample_matches_numpy_table(K, nrep, n, order):
    rng = np.random.default_rng(1)
    u = rng.normal(175.0, 5.3, n)
    x = 1e-3 * u + rng.normal(0.0, 0.02, n)
    a, b = np.random.default_rng(21), np.random.default_rng(21)
    idx = a.choice(n, size=(nrep, n), replace=True)
    ref = oc.resample_vals_gather(x, u, order, idx)
    out = K.sampler_resample(b, x, u, nrep, order).cpu().numpy()
    scale = oc.tolerance_scale(x, u, order)
    assert out.shape == ref.shape
    assert float((np.abs(out - ref) / np.maximum(np.abs(ref), scale)).max()) < TOL
    assert _same_state(a, b)


def test_fused_resample_weighted_and_x_is_u(K):
    n, nrep, order = 5000, 6, 5
    rng = np.random.default_rng(2)
    u = rng.normal(-3.0, 2.0, n)
    x = 0.5 * u + rng.normal(0.0, 1.0, n)
    w = rng.random(n)
    a, b, c = (np.random.default_rng(8) for _ in range(3))
    idx = a.choice(n, size=(nrep, n), replace=True)
    scale = oc.tolerance_scale(x, u, order)
    out = K.sampler_resample(b, x, u, nrep, order, w=w).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, order, idx, w=w)
    assert float((np.abs(out - ref) / np.maximum(np.abs(ref), scale)).max()) < TOL
    out = K.sampler_resample(c, None, u, nrep, order, x_is_u=True).cpu().numpy()
    ref = K.resample_indices(None, u, idx, order, x_is_u=True).cpu().numpy()
    assert np.allclose(out, ref, rtol=1e-11, atol=0)


def test_full_size_properties(K):
    """BASELINE config-2 width (ndat = 1e7), a slice of its replicates: row sums and parity of two routes."""
    import torch

    n, nrep, order = 10_000_000, 8, 8
    g = torch.Generator(device="cuda").manual_seed(0)
    u = torch.randn(n, dtype=torch.float64, device="cuda", generator=g) * 5.3 + 175.0
    x = 1e-3 * u + 0.02 * torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
    a, b, c = (np.random.default_rng(4) for _ in range(3))
    fused = K.sampler_resample(a, x, u, nrep, order)
    idx = K.sampler_indices(b, nrep, n)
    assert np.array_equal(idx[0, :1000].cpu().numpy(), c.choice(n, size=(nrep, n))[0, :1000])
    via_table = K.resample_indices(x, u, idx, order)
    assert torch.allclose(fused, via_table, rtol=1e-9, atol=0)
    assert fused[:, 0, 0].tolist() == [float(n)] * nrep
    freq = K.indices_to_freq(idx, n, dtype=torch.uint8)
    via_freq = K.resample_vals(x, u, freq, order)
    assert torch.allclose(fused, via_freq, rtol=1e-9, atol=0)


# ------------------------------------------------------------------ windows / split counting (multi-GPU form)
def test_replicate_windows_tile_the_table(K):
    import torch

    nrep, n, order = 11, 50_000, 6
    rng = np.random.default_rng(3)
    u = rng.normal(0.0, 1.0, n)
    x = u + rng.normal(0.0, 1.0, n)
    ref_idx = np.random.default_rng(9).choice(n, (nrep, n))
    plan = K.SamplerPlan(np.random.default_rng(9), nrep, n)
    whole = K.sampler_resample(plan, x, u, order=order)
    cuts = [(0, 4), (4, 5), (5, 11)]
    parts = [K.sampler_resample(plan, x, u, order=order, reps=c) for c in cuts]
    # the gather path flushes replicate sums with fp64 atomics (order varies run to run): compare on the
    # comoment scale, not with atol=0 (odd central moments are ~0)
    scale = oc.tolerance_scale(x, u, order)
    got, want = torch.cat(parts).cpu().numpy(), whole.cpu().numpy()
    assert float((np.abs(got - want) / np.maximum(np.abs(want), scale)).max()) < 1e-12
    for a, b in cuts:
        assert np.array_equal(K.sampler_indices(plan, reps=(a, b)).cpu().numpy(), ref_idx[a:b])
        f = K.sampler_freq(plan, dtype=torch.int32, reps=(a, b)).cpu().numpy()
        assert np.array_equal(f.astype(np.int64), oc.indices_to_freq(ref_idx[a:b], n))
    with pytest.raises(ValueError):
        K.sampler_indices(plan, reps=(5, 12))


@pytest.mark.parametrize("world", [2, 3, 8])
def test_split_count_equals_prepare(K, world):
    """Ranks counting disjoint block ranges + integer SUM == the single-GPU preparation."""
    import ctypes as C

    import torch

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    nrep, n = 5, 3_000_000
    plan = K.SamplerPlan(np.random.default_rng(4), nrep, n)
    state = K.pcg64_state(np.random.default_rng(4))
    nb = lib.txb_sampler_nblocks(C.byref(state), n, nrep * n)
    assert nb > 0
    lay = (C.c_int64 * 4)()
    assert lib.txb_sampler_layout(n, nrep * n, lay) == 0
    assert lay[0] == nb and lay[3] == lib.txb_sampler_workspace_bytes(n, nrep * n) == plan.head.numel()
    heads = []
    for r in range(world):
        h = torch.zeros(lay[3], dtype=torch.uint8, device="cuda")
        lo, hi = nb * r // world, nb * (r + 1) // world
        assert lib.txb_sampler_count(C.byref(state), n, nrep * n, lo, hi, h.data_ptr(), h.numel(), None) == 0
        heads.append(h)
    total = heads[0].clone()
    for h in heads[1:]:
        total.view(torch.int64)[8 : 8 + nb] += h.view(torch.int64)[8 : 8 + nb]
        total[lay[2] : lay[2] + 4 * lay[1]].view(torch.int32).add_(h[lay[2] : lay[2] + 4 * lay[1]].view(torch.int32))
    aux = torch.zeros(2, dtype=torch.int64, device="cuda")
    assert lib.txb_sampler_scan(C.byref(state), n, nrep * n, aux.data_ptr(), aux[1:].data_ptr(), total.data_ptr(),
                                total.numel(), None) == 0
    assert torch.equal(total[64:], plan.head[64:])
    assert int(aux[0]) == plan.raw_used
    it = plan.head[lay[2] : lay[2] + 4 * lay[1]].view(torch.int32)
    assert int(it.sum()) >= nrep * n  # the counted raw range covers the requested draws


def test_sums_are_additive_across_shards(K):
    import torch

    x, u, _ = (lambda r: (r.normal(1.0, 2.0, (9000, 4)), r.normal(175.0, 5.3, 9000), None))(np.random.default_rng(0))
    order = 6
    pivot, _, _, _ = K.reduce_sums(x[:1024], u[:1024], 1)
    tot0 = tot1 = None
    for lo, hi in [(0, 1000), (1000, 1001), (1001, 9000)]:
        _, s0, s1, vs = K.reduce_sums(x[lo:hi], u[lo:hi], order, pivot=pivot)
        tot0 = s0 if tot0 is None else tot0 + s0
        tot1 = s1 if tot1 is None else tot1 + s1
    out = K.finalize_sums(pivot, tot0, tot1, order, val_shape=vs).cpu().numpy()
    ref = oc.comoments_two_pass(x, u, order)
    assert float((np.abs(out - ref) / np.maximum(np.abs(ref), oc.tolerance_scale(x, u, order))).max()) < TOL
    assert torch.allclose(torch.from_numpy(out), K.reduce_vals(x, u, order).cpu(), rtol=1e-11, atol=1e-13)


# ------------------------------------------------------------------ bucketed (shared-memory) bootstrap
@pytest.mark.parametrize("n,nrep,order", [(4_200_000, 3, 8), (5_000_003, 2, 6), (6_300_000, 2, 2)])
def test_bucketed_equals_gather_and_oracle(K, n, nrep, order):
    import torch

    rng = np.random.default_rng(5)
    u = rng.normal(175.0, 5.3, n)
    x = 1e-3 * u + rng.normal(0.0, 0.02, n)
    w = rng.random(n)
    idx = np.random.default_rng(31).choice(n, (nrep, n))
    scale = oc.tolerance_scale(x, u, order)
    for wt in (None, w):
        a = K.sampler_resample(np.random.default_rng(31), x, u, nrep, order, w=wt, buckets=True)
        b = K.sampler_resample(np.random.default_rng(31), x, u, nrep, order, w=wt, buckets=False)
        assert torch.allclose(a, b, rtol=1e-10, atol=1e-300)
        if wt is None:
            ref = oc.resample_vals_gather(x, u, order, idx[:1], w=wt)
            assert float((np.abs(a[:1].cpu().numpy() - ref) / np.maximum(np.abs(ref), scale)).max()) < TOL
    a = K.sampler_resample(np.random.default_rng(31), None, u, nrep, order, x_is_u=True, buckets=True)
    b = K.sampler_resample(np.random.default_rng(31), None, u, nrep, order, x_is_u=True, buckets=False)
    assert torch.allclose(a, b, rtol=1e-10, atol=1e-300)


def test_bucketed_windows_when_scratch_is_small(K):
    import torch

    n, nrep, order = 4_200_000, 70, 4
    rng = np.random.default_rng(6)
    u = rng.normal(0.0, 1.0, n)
    x = u + rng.normal(0.0, 1.0, n)
    full = K.sampler_resample(np.random.default_rng(2), x, u, nrep, order, buckets=False)
    # scratch for ~33 of the 70 replicates: the library must loop over replicate windows
    per_rep = 513 * (4 + 2 * 8 * (order + 1) + 300 * 64)
    part = K.sampler_resample(np.random.default_rng(2), x, u, nrep, order, scratch_bytes=33 * per_rep)
    assert torch.allclose(part, full, rtol=1e-10, atol=1e-300)
    # and replicate windows requested by the caller
    plan = K.SamplerPlan(np.random.default_rng(2), nrep, n)
    mid = K.sampler_resample(plan, x, u, order=order, reps=(17, 60))
    assert torch.allclose(mid, full[17:60], rtol=1e-10, atol=1e-300)


@pytest.mark.parametrize("n,nchan,nrep,order,weighted", [(3001, 5, 300, 3, False), (4096, 64, 130, 6, True)])
def test_wide_bootstrap_in_replicate_windows(K, n, nchan, nrep, order, weighted):
    """Wide x, lazy sampler: the frequency table is built and contracted one replicate window at a
    time (data.py:1803-1810 semantics at sizes where the reference's dense table cannot exist).
    Windows of 128 replicates must give what the numpy table gives (tests/test_data.py:94-112)."""
    rng = np.random.default_rng(4)
    u = rng.normal(10.0, 2.0, n)
    x = rng.normal(0.0, 1.0, (n, nchan)) + 0.3 * u[:, None]
    w = rng.random(n) if weighted else None
    a, b, c = (np.random.default_rng(33) for _ in range(3))
    idx = a.choice(n, size=(nrep, n), replace=True)
    timing = {}
    out = K.sampler_resample_wide(b, x, u, nrep, order, w=w, table_bytes=1, timing=timing).cpu().numpy()
    assert timing["windows"] == -(-nrep // 128) and _same_state(a, b)
    one = K.sampler_resample_wide(c, x, u, nrep, order, w=w).cpu().numpy()  # one window
    assert out.shape == one.shape == (nrep, nchan, 2, order + 1)
    scale = oc.tolerance_scale(x, u, order)
    assert float((np.abs(out - one) / np.maximum(np.abs(one), scale)).max()) < 1e-12  # slab split differs only
    for r in (0, 127, 128, nrep - 1):
        ref = oc.resample_vals_gather(x, u, order, idx[r : r + 1], w=w)[0]
        assert float((np.abs(out[r] - ref) / np.maximum(np.abs(ref), scale)).max()) < TOL


def test_wide_bootstrap_through_the_api_uses_windows(K, monkeypatch):
    """``wrap_resample_vals`` with ``sampler={"nrep", "rng"}`` and wide x takes the windowed path when the
    table is over the budget, and agrees with the one-table path."""
    import thermoextrap_b200.backend as B
    from thermoextrap_b200.core.xr import DataArray

    n, nchan, nrep, order = 2000, 7, 200, 4
    rng = np.random.default_rng(5)
    u = rng.normal(1.0, 1.0, n)
    x = rng.normal(0.0, 1.0, (n, nchan)) + u[:, None]
    xv, uv = DataArray(x, dims=("rec", "val")), DataArray(u, dims=("rec",))
    calls = []
    real = K.sampler_resample_wide
    monkeypatch.setattr(K, "sampler_resample_wide", lambda *a, **k: calls.append(1) or real(*a, **k))
    full = B.wrap_resample_vals(xv, uv, mom=(1, order), sampler={"nrep": nrep, "rng": np.random.default_rng(9)},
                                dim="rec").obj.values
    assert not calls
    monkeypatch.setattr(B, "WIDE_TABLE_BYTES", 1000)
    win = B.wrap_resample_vals(xv, uv, mom=(1, order), sampler={"nrep": nrep, "rng": np.random.default_rng(9)},
                               dim="rec").obj.values
    assert calls and win.shape == full.shape == (nrep, nchan, 2, order + 1)
    scale = oc.tolerance_scale(x, u, order)
    assert float((np.abs(win - full) / np.maximum(np.abs(full), scale)).max()) < 1e-12


@pytest.mark.parametrize("nrep,nsamp,ndat,dtype_name,scratch", [
    (5, 2_000_000, 2_000_000, "uint8", None),
    (5, 2_000_003, 2_000_003, "uint8", 14_000_000),   # rows not 16-byte multiples; two replicates per window
    (3, 1_500_000, 2_000_000, "uint16", None),        # nsamp != ndat
    (2, 9_000_000, 9_000_000, "uint8", None),         # 138 buckets
])
def test_bucketed_histogram_bit_exact(K, nrep, nsamp, ndat, dtype_name, scratch):
    """K6 frequency table through the shared-memory path (bucket partition + per-bucket histogram):
    identical integers to ``indices_to_freq(rng.choice(...))`` and to the global-atomic form."""
    import torch

    dt = getattr(torch, dtype_name)
    a, b, c = (np.random.default_rng(17) for _ in range(3))
    ref = oc.indices_to_freq(a.choice(ndat, size=(nrep, nsamp), replace=True), ndat)
    lib = __import__("thermoextrap_b200._lib", fromlist=["load"]).load()
    assert lib.txb_sampler_freq_workspace_bytes(ndat, nrep, nsamp, nrep) > lib.txb_sampler_workspace_bytes(ndat, nrep * nsamp)
    lib.txb_profile_begin(b"bucket_hist")
    out = K.sampler_freq(b, nrep, nsamp, ndat, dtype=dt, scratch_bytes=scratch)
    import ctypes as C

    ms, nl = C.c_double(), C.c_int()
    lib.txb_profile_end(C.byref(ms), C.byref(nl))
    assert nl.value >= (1 if scratch is None else 2), "the bucketed histogram kernel did not run"
    assert np.array_equal(out.cpu().numpy().astype(np.int64), ref)
    assert _same_state(a, b)
    slow = K.sampler_freq(c, nrep, nsamp, ndat, dtype=dt, buckets=False)
    assert torch.equal(out, slow)


def test_bucketed_histogram_overflow_is_loud(K):
    """A uint8 lane that wraps inside the shared-memory histogram is caught by the row-sum check."""
    import torch

    ndat, nsamp = 600_000, 120_000_000  # mean count 200: tens of samples exceed 255
    with pytest.raises(OverflowError):
        K.sampler_freq(np.random.default_rng(3), 1, nsamp, ndat, dtype=torch.uint8)
    with pytest.raises(OverflowError):
        K.sampler_freq(np.random.default_rng(3), 1, nsamp, ndat, dtype=torch.uint8, buckets=False)
    f16 = K.sampler_freq(np.random.default_rng(3), 1, nsamp, ndat, dtype=torch.uint16)
    ref = np.bincount(np.random.default_rng(3).choice(ndat, size=nsamp, replace=True), minlength=ndat)
    assert ref.max() > 255 and np.array_equal(f16[0].cpu().numpy().astype(np.int64), ref)


def test_deferred_status_gives_the_same_bootstrap(K):
    """``status_out`` skips the per-call status read-back (bench steady state); ``check_deferred`` verifies later."""
    n, nrep, order = 4_500_000, 3, 4   # large enough for the bucketed path
    rng = np.random.default_rng(6)
    u = rng.normal(0.0, 1.0, n)
    x = u + rng.normal(0.0, 1.0, n)
    a = K.sampler_resample(np.random.default_rng(2), x, u, nrep, order, sync_rng=False)
    pending = []
    b = K.sampler_resample(np.random.default_rng(2), x, u, nrep, order, sync_rng=False, status_out=pending)
    assert len(pending) == 1
    K.check_deferred(pending)
    assert not pending
    # same draws, same lists; the order of entries inside a list depends on the order of the shared-memory
    # atomics, so the sums agree to rounding, not bit for bit
    scale = oc.tolerance_scale(x, u, order)
    a, b = a.cpu().numpy(), b.cpu().numpy()
    assert float((np.abs(a - b) / np.maximum(np.abs(a), scale)).max()) < 1e-12

"""Chunked upload on the copy stream (backend.to_device / kernels.upload_async): every consumer is ordered after
the copy on its own stream, and the wide bootstrap's first window follows the chunks as they land.  The copy stream
is held back artificially so that the consumers really find the upload in flight."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture
def small_async(monkeypatch):
    import thermoextrap_b200.backend as backend
    from thermoextrap_b200 import kernels as K

    monkeypatch.setattr(K, "ASYNC_UPLOAD_MIN_BYTES", 1 << 20)
    monkeypatch.setattr(K, "UPLOAD_CHUNKS", 4)
    monkeypatch.setattr(K, "UPLOAD_TEST_DELAY_CYCLES", 2_000_000_000)   # ~1 s
    monkeypatch.setattr(backend, "ASYNC_UPLOAD", True)
    backend.clear_device_cache()
    yield backend, K
    torch.cuda.synchronize()
    backend.clear_device_cache()


def _samples(n, c, seed=0):
    rng = np.random.default_rng(seed)
    u = rng.normal(size=n) * 1.5 + 3.0
    x = rng.normal(size=(n, c)) + 0.3 * u[:, None] + np.arange(c)[None]
    return x, u


def test_consumers_wait_for_an_upload_in_flight(small_async):
    backend, K = small_async
    x, u = _samples(50_000, 8)
    xd = backend.to_device(x)
    up = K.pending_upload(xd)
    assert up is not None and len(up.bounds) == 4 and up.bounds[0] == (0, 16384) and up.bounds[-1][1] == 50_000
    ud = backend.to_device(u)          # 400 KB: the plain copy on the compute stream
    assert K.pending_upload(ud) is None
    out = K.reduce_vals(xd, ud, 3).cpu().numpy()      # queued behind the last chunk, no host wait
    from oracle import comoments as oc

    ref = oc.comoments_two_pass(x, u, 3)
    scale = oc.tolerance_scale(x, u, 3)
    assert np.all(np.abs(out - ref) <= 1e-10 * np.maximum(np.abs(ref), scale))
    assert torch.equal(xd.cpu(), torch.from_numpy(x))
    assert K.pending_upload(xd) is None     # landed: forgotten


@pytest.mark.parametrize(("order", "max_follow"), [(2, 1), (6, 2), (6, 8)])
def test_first_windows_follow_the_chunks(small_async, monkeypatch, order, max_follow):
    """from_vals -> model -> resample -> predict with the samples still uploading: K1 is deferred, the first replicate
    windows (1, 2, all 3 of them) are contracted chunk by chunk; same predictions as with the one-copy upload."""
    import thermoextrap_b200 as xt

    backend, K = small_async
    n, c, nrep = 20_000, 16, 300
    x, u = _samples(n, c, seed=order)
    monkeypatch.setattr(backend, "WIDE_TABLE_BYTES", 128 * n)   # three windows of <= 128 replicates
    monkeypatch.setattr(K, "FOLLOW_MAX_WINDOWS", max_follow)
    xv, uv = xt.DataArray(x, dims=("rec", "val")), xt.DataArray(u, dims=("rec",))

    def run():
        backend.clear_device_cache()
        data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
        model = xt.beta.factory_extrapmodel(beta=0.4, data=data)
        boot = model.resample(sampler={"nrep": nrep, "rng": np.random.default_rng(7)})
        return model.predict([0.3, 0.5]).values, boot.predict([0.3, 0.5]).values, boot.data.dxduave.obj.values

    before = K.follow_windows
    monkeypatch.setattr(backend, "ASYNC_UPLOAD", False)
    p0, b0, m0 = run()            # also warms the symbolic caches, so that the next run reaches the bootstrap
    assert K.follow_windows == before   # well inside the artificial delay of the copy stream
    monkeypatch.setattr(backend, "ASYNC_UPLOAD", True)
    p1, b1, m1 = run()
    assert K.follow_windows == before + min(max_follow, 3)
    np.testing.assert_allclose(p1, p0, rtol=1e-13, atol=0)
    # a followed window differs only in the order its partial sums were added
    assert m1.shape == m0.shape
    from oracle import comoments as oc

    scale = oc.tolerance_scale(x, u, order)
    assert np.all(np.abs(m1 - m0) <= 1e-12 * np.maximum(np.abs(m0), scale[None]))
    np.testing.assert_allclose(b1, b0, rtol=1e-9, atol=0)
    assert m1[:, :, 0, 0].min() == m1[:, :, 0, 0].max() == float(n)


def test_integer_tables_take_the_plain_copy(small_async):
    backend, K = small_async
    idx = np.random.default_rng(0).integers(0, 1000, size=(300, 1000))   # 2.4 MB of int64
    t = backend.to_device(idx)
    assert K.pending_upload(t) is None and torch.equal(t.cpu(), torch.from_numpy(idx))


def test_following_windows_with_trailing_value_dims(small_async, monkeypatch):
    """x[rec, val, deriv] (the xalpha layout): the followed windows flatten the trailing dims like the resident path."""
    backend, K = small_async
    n, nrep, order = 12_000, 200, 3
    rng = np.random.default_rng(5)
    u = rng.normal(size=n) + 2.0
    x = rng.normal(size=(n, 6, 3)) + 0.2 * u[:, None, None]
    plan = lambda: K.SamplerPlan(np.random.default_rng(3), nrep, n)   # noqa: E731
    first = plan()                                 # built (and its kernels loaded) before the clock of the delay starts
    ud = backend.to_device(u)
    xd = backend.to_device(x)                      # 1.7 MB: chunked, held back by the artificial delay
    assert K.pending_upload(xd) is not None
    before = K.follow_windows
    got = K.sampler_resample_wide(first, xd, ud, order=order, table_bytes=128 * n)
    assert K.follow_windows == before + 2 and got.shape == (nrep, 6, 3, 2, order + 1)
    torch.cuda.synchronize()
    ref = K.sampler_resample_wide(plan(), xd, ud, order=order, table_bytes=128 * n)   # resident now
    assert K.follow_windows == before + 2
    from oracle import comoments as oc

    scale = oc.tolerance_scale(x, u, order)
    g, r = got.cpu().numpy(), ref.cpu().numpy()
    assert np.all(np.abs(g - r) <= 1e-12 * np.maximum(np.abs(r), scale[None]))

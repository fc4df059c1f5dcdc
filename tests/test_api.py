"""Host-side mirror of the reference API, exercised the way the reference's own tests exercise it.

Runs against the real CUDA kernels on a GPU box and against the oracle-backed stand-in
(tests/fake_kernels.py) elsewhere -- either way the host logic (dims, samplers, data classes,
factories, models) is the thing under test.  Each test names the reference test it mirrors
(paths relative to /root/reference/tests).
"""

import json
from pathlib import Path

import numpy as np
import pytest

import thermoextrap_b200 as xtrap
from thermoextrap_b200 import DataArray

GOLD = Path(__file__).parent / "golden"


@pytest.fixture(autouse=True)
def _kernels(backend_kernels):
    return backend_kernels


class Fixture:
    """FixtureData(n=100, nv=5, order=5, seed=0)  (conftest.py:15-28)."""

    def __init__(self):
        rng = np.random.default_rng(0)
        self.n, self.nv, self.order = 100, 5, 5
        self.u = rng.random(self.n)
        self.x = rng.random((self.n, self.nv))
        self.ub = rng.random(self.n)
        self.xb = rng.random((self.n, self.nv))
        self.uv = xtrap.xrwrap_uv(self.u)
        self.xv = xtrap.xrwrap_xv(self.x)
        self.beta0 = 0.5
        self.betas = [0.3, 0.4]

    def vals(self, central=True):
        return xtrap.DataCentralMomentsVals.from_vals(xv=self.xv, uv=self.uv, order=self.order, central=central)

    def cmom(self, central=True):
        return xtrap.DataCentralMoments.from_vals(xv=self.xv, uv=self.uv, order=self.order, central=central, dim="rec")

    def values(self, central=False):
        return xtrap.factory_data_values(uv=self.u, xv=self.x, order=self.order, central=central)


@pytest.fixture(scope="module")
def fx():
    return Fixture()


@pytest.fixture(scope="module")
def legacy():
    return json.loads((GOLD / "legacy_derivs.json").read_text())


def test_fixture_matches_golden_generator(fx, legacy):
    assert fx.u.sum() == pytest.approx(legacy["check_u_sum"], rel=1e-14)
    assert fx.xb.sum() == pytest.approx(legacy["check_xb_sum"], rel=1e-14)


# ---------------------------------------------------------------- test_data.py:7-48
@pytest.mark.parametrize("maker", ["vals", "cmom"])
@pytest.mark.parametrize("central", [True, False])
def test_raw_moments_match_reference_legacy(fx, legacy, maker, central):
    data = getattr(fx, maker)(central)
    u = data.u.values
    np.testing.assert_allclose(u, np.broadcast_to(legacy["u_raw"], u.shape), rtol=1e-10)
    xu = data.xu.transpose("umom", ...).values
    np.testing.assert_allclose(xu, np.asarray(legacy["xu_raw"]), rtol=1e-10)


def test_values_classes_agree(fx, legacy):
    for central in (False, True):
        d = fx.values(central)
        args = d.derivs_args
        assert len(args) >= 2
    u = fx.values(False).u.values
    np.testing.assert_allclose(u, np.broadcast_to(legacy["u_raw"], u.shape), rtol=1e-10)


# ---------------------------------------------------------------- test_beta.py:17-47
@pytest.mark.parametrize("maker,central", [("vals", True), ("vals", False), ("cmom", True), ("cmom", False), ("values", False), ("values", True)])
def test_beta_derivs_match_reference_legacy(fx, legacy, maker, central):
    data = getattr(fx, maker)(central)
    model = xtrap.beta.factory_extrapmodel(beta=fx.beta0, data=data)
    d = model.derivs(norm=False).transpose("order", ...).values
    ref = np.asarray(legacy["derivs"])
    scale = np.abs(ref).max(axis=1, keepdims=True)
    assert np.all(np.abs(d - ref) <= 1e-10 * np.maximum(np.abs(ref), scale))


# ---------------------------------------------------------------- test_data.py:94-112
def test_resample_freq_equals_gather(fx):
    rng = np.random.default_rng(0)
    nrep = 10
    idx = rng.choice(fx.n, (nrep, fx.n), replace=True)
    a = fx.vals().resample(sampler={"indices": idx})
    assert a.dxduave.obj.dims[0] == "rep"
    from oracle import comoments as oc

    ref = oc.resample_vals_gather(fx.x, fx.u, fx.order, idx)
    scale = oc.tolerance_scale(fx.x, fx.u, fx.order)  # sigma_x^i sigma_u^j: odd central moments are ~ 0
    got = a.dxduave.obj.values
    assert np.all(np.abs(got - ref) <= 1e-10 * np.maximum(np.abs(ref), scale[None]))
    # a freq table gives the same thing
    freq = oc.indices_to_freq(idx, fx.n)
    b = fx.vals().resample(sampler={"freq": freq})
    np.testing.assert_allclose(b.dxduave.obj.values, a.dxduave.obj.values, rtol=1e-11, atol=1e-15)


def test_resample_random_spec_equals_numpy_table(fx):
    """{"nrep": n, "rng": g} must bootstrap with exactly the table numpy would have drawn, and leave g where numpy would."""
    g1, g2 = np.random.default_rng(42), np.random.default_rng(42)
    idx = g1.choice(fx.n, (7, fx.n), replace=True)
    a = fx.vals().resample(sampler={"nrep": 7, "rng": g2})
    b = fx.vals().resample(sampler={"indices": idx})
    np.testing.assert_allclose(a.dxduave.obj.values, b.dxduave.obj.values, rtol=1e-11, atol=1e-15)
    assert g1.integers(0, 1 << 30, 4).tolist() == g2.integers(0, 1 << 30, 4).tolist()
    # scalar x goes through the fused sampler kernel on the GPU
    xs = DataArray(fx.x[:, 0].copy(), dims=("rec",))
    d = xtrap.DataCentralMomentsVals.from_vals(xv=xs, uv=fx.uv, order=3, central=True)
    g3 = np.random.default_rng(42)
    c = d.resample(sampler={"nrep": 7, "rng": g3})
    np.testing.assert_allclose(c.dxduave.obj.values[..., :4], a.dxduave.obj.values[:, 0, :, :4], rtol=1e-10, atol=1e-14)
    s = xtrap.backend.factory_sampler({"nrep": 3, "rng": np.random.default_rng(5)}, ndat=50)
    assert np.array_equal(s.indices, np.random.default_rng(5).choice(50, (3, 50)))
    assert np.array_equal(s.freq.sum(axis=1), [50] * 3)


def test_deferred_reduce_launches_on_first_use(fx, monkeypatch):
    """K1 of ``from_vals`` may be deferred (samples still uploading: backend.wrap_reduce_vals): building the data
    object, the model and a bootstrap must not launch it; asking for values does, once, with the same result."""
    import thermoextrap_b200.backend as backend

    eager = fx.vals()
    ref_pred = xtrap.beta.factory_extrapmodel(beta=fx.beta0, data=eager).predict(fx.betas).values
    calls = []
    real = backend.K.reduce_vals

    def counted(*a, **k):
        calls.append(1)
        return real(*a, **k)

    monkeypatch.setattr(backend.K, "reduce_vals", counted)
    monkeypatch.setattr(backend, "DEFER_REDUCE", True)
    data = fx.vals()
    assert not calls and data.dxduave.dims == eager.dxduave.dims and data.dxduave.mom == eager.dxduave.mom
    assert dict(data.dxduave.sizes) == dict(eager.dxduave.sizes)
    model = xtrap.beta.factory_extrapmodel(beta=fx.beta0, data=data)
    boot = model.resample(sampler={"nrep": 4, "rng": np.random.default_rng(1)})
    assert not calls, "the bootstrap needs the samples, not their comoments"
    assert boot.predict(fx.betas).sizes["rep"] == 4
    np.testing.assert_allclose(model.predict(fx.betas).values, ref_pred, rtol=1e-13, atol=0)
    assert len(calls) == 1
    np.testing.assert_array_equal(data.dxduave.obj.values, eager.dxduave.obj.values)
    assert len(calls) == 1 and data.dxduave.device_records() is not None
    # a deferred record pickles as its values
    import pickle

    lazy = fx.vals().dxduave
    assert lazy._obj is None
    back = pickle.loads(pickle.dumps(lazy))
    np.testing.assert_array_equal(back.obj.values, eager.dxduave.obj.values)
    assert back.mom_dims == eager.dxduave.mom_dims and back.device_records() is None


def test_pinned_empty_is_a_plain_numpy_array():
    import thermoextrap_b200.backend as backend

    a = backend.pinned_empty((7, 3))
    assert isinstance(a, np.ndarray) and a.shape == (7, 3) and a.dtype == np.float64 and a.flags.c_contiguous
    a[:] = 2.5
    assert float(a.sum()) == 52.5
    assert backend.pinned_empty(5, np.int64).shape == (5,)


def test_block_resample_and_reduce(fx):
    """DataCentralMoments over pre-reduced blocks (data.py:974-1055; notebook cells 39-41)."""
    xx = DataArray(fx.x[:, 0].reshape(10, 10), dims=("rec", "block"))
    uu = DataArray(fx.u.reshape(10, 10), dims=("rec", "block"))
    blocks = xtrap.DataCentralMoments.from_vals(xv=xx, uv=uu, order=3, dim="block", central=True)
    assert blocks.dxduave.obj.shape == (10, 2, 4)
    total = blocks.reduce(dim="rec")
    direct = xtrap.DataCentralMoments.from_vals(
        xv=DataArray(fx.x[:, 0].copy(), dims=("rec",)), uv=fx.uv, order=3, dim="rec", central=True)
    np.testing.assert_allclose(total.dxduave.obj.values, direct.dxduave.obj.values, rtol=1e-11, atol=1e-15)
    boot = blocks.resample(sampler={"nrep": 4, "rng": np.random.default_rng(1)})
    assert boot.dxduave.obj.shape == (4, 2, 4)
    with pytest.raises(ValueError):
        direct.resample(sampler={"nrep": 4})

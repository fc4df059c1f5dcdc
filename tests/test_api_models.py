"""Models / factories of the reference API (known answers, golden vectors, analytic ideal gas).

Mirrors /root/reference/tests: test_beta.py:77-128 (ideal gas vs analytic), :431-452 (exact interpolation
polynomial), test_lnPi.py:122-159 (golden vectors, subset committed under tests/golden), test_volume.py:56-74.
Runs on the CUDA kernels on a GPU box, on the oracle stand-in elsewhere.
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


def test_extrapmodel_idealgas_vs_analytic():
    """BASELINE config 1 at the reference test's size: order 3 here, 1e5 samples, nrep=100."""
    ref_beta, max_order, test_betas = 5.0, 3, np.array([4.9, 5.1])
    rng = np.random.default_rng(0)
    xdata, udata = xtrap.idealgas.generate_data((100_000, 1), ref_beta, 1.0, rng=rng)
    dat = xtrap.DataCentralMomentsVals.from_vals(
        order=max_order, xv=DataArray(xdata, dims=("rec",)), uv=DataArray(udata, dims=("rec",)), central=True)
    ex = xtrap.beta.factory_extrapmodel(ref_beta, dat, xalpha=False)
    ex_res = ex.resample(sampler={"nrep": 100, "rng": rng})
    for o in range(max_order + 1):
        true_extrap, true_derivs = xtrap.idealgas.x_beta_extrap(o, ref_beta, test_betas, 1.0)
        test_derivs = ex.derivs(order=o, norm=False).values
        test_extrap = ex.predict(test_betas, order=o).values
        derr = 2.0 * ex_res.derivs(order=o, norm=False).std("rep").values[-1]
        eerr = 2.0 * np.max(ex_res.predict(test_betas, order=o).std("rep").values)
        np.testing.assert_allclose(true_derivs[-1], test_derivs[-1], rtol=0.0, atol=derr * 5)
        np.testing.assert_allclose(true_extrap, test_extrap, rtol=0.0, atol=eerr * 2)


def test_interpmodel_polynomial_exact():
    xdat2 = DataArray(np.array([0.5, 1.5]), dims=("rec",))
    for i in range(3):
        xdat1 = DataArray(((-1.0) ** (i + 1)) * xdat2.values, dims=("rec",))
        udat1 = DataArray((i + 1) * np.array([-2.0, 2.0]), dims=("rec",))
        udat2 = DataArray((i + 1) * np.array([2.0, -2.0]), dims=("rec",))
        dat1 = xtrap.DataCentralMomentsVals.from_vals(order=1, xv=xdat1, uv=udat1, central=True)
        dat2 = xtrap.DataCentralMomentsVals.from_vals(order=1, xv=xdat2, uv=udat2, central=True)
        ex1 = xtrap.beta.factory_extrapmodel(-1.0, dat1, xalpha=False)
        ex2 = xtrap.beta.factory_extrapmodel(1.0, dat2, xalpha=False)
        interp = xtrap.InterpModel([ex1, ex2])
        check = np.zeros(4)
        check[i + 1] = 1.0
        np.testing.assert_allclose(interp.coefs().values, check, rtol=0, atol=1e-14)
        np.testing.assert_allclose(interp.predict([0.0, 0.5]).values, [0.5 ** (i + 1) * a for a in (0.0, 1.0)], atol=1e-14)


# ---------------------------------------------------------------- lnPi golden vectors
@pytest.fixture(scope="module")
def lnpi_gold():
    d = json.loads((GOLD / "lnpi_sample.json").read_text())

    def prep(rec):
        lnpi = np.asarray(rec["lnPi"], dtype=float)
        lnpi = lnpi - lnpi[0]
        energy = np.asarray(rec["energy"], dtype=float)
        energy = np.concatenate((np.ones((len(lnpi), 1)), energy), axis=-1)
        return {"lnpi": DataArray(lnpi, dims=("n",)), "energy": DataArray(energy, dims=("n", "umom")),
                "mu": DataArray(np.atleast_1d(rec["mu"]), dims=("comp",)), "beta": 1.0 / rec["temp"],
                "order": rec["order"]}

    return d["stride"], prep(d["ref"]), [prep(s) for s in d["samples"]]


@pytest.mark.parametrize("central", [False, True])
def test_lnpi_golden_u_and_lnpi(lnpi_gold, central):
    stride, ref, samples = lnpi_gold
    betas = np.unique([s["beta"] for s in samples])
    data_u = xtrap.DataCentralMoments.from_ave_raw(u=ref["energy"], xu=None, x_is_u=True, central=central, meta=None)
    em_u = xtrap.beta.factory_extrapmodel(beta=ref["beta"], data=data_u, name="u_ave")
    out_u = em_u.predict(betas, cumsum=True)
    # macrostate coordinate of the strided subset: N = stride * i
    ncoords = DataArray(np.arange(ref["lnpi"].shape[0], dtype=float)[None, :] * stride, dims=("comp", "n"))
    meta = xtrap.lnpi.lnPiDataCallback(ref["lnpi"], ref["mu"], dims_n=["n"], dims_comp="comp", ncoords=ncoords)
    em_l = xtrap.lnpi.factory_extrapmodel_lnPi(beta=ref["beta"], data=data_u.new_like(meta=meta))
    out_l = em_l.predict(betas, cumsum=True)
    for s in samples:
        ib = int(np.argmin(np.abs(betas - s["beta"])))
        b = out_u.transpose("beta", "order", ...).values[ib, s["order"]]
        np.testing.assert_allclose(s["energy"].values[:, 1], b, rtol=1e-5)
        lp = out_l.transpose("beta", "order", ...).values[ib, s["order"]]
        np.testing.assert_allclose(s["lnpi"].values, lp - lp[0], rtol=1e-7, atol=1e-9)


# ---------------------------------------------------------------- volume (test_volume.py:56-74)
def test_volume_models_agree():
    rng = np.random.default_rng(0)
    u = rng.random(100)
    x = rng.random((100, 5))
    volume, volumes = 1.0, [0.1, 0.5, 1.5, 2.0]
    xem_ig = xtrap.volume_idealgas.factory_extrapmodel(order=1, volume=volume, uv=u, xv=x)
    uv, xv = xtrap.xrwrap_uv(u), xtrap.xrwrap_xv(x)
    xem = xtrap.volume.factory_extrapmodel(uv=uv, xv=xv, order=1, dxdqv=xv, volume=volume, ndim=1)
    a, b = xem_ig.predict(volumes).values, xem.predict(volumes).values
    np.testing.assert_allclose(a, b, rtol=1e-12)
    # closed form (volume_idealgas.py:44-55): d0 = <x>, d1 = (<xW> - <x><W>)/V + <x>/V
    d0 = x.mean(0)
    d1 = ((x * u[:, None]).mean(0) - d0 * u.mean()) / volume + d0 / volume
    ref = d0[None, :] + d1[None, :] * (np.asarray(volumes)[:, None] - volume)
    np.testing.assert_allclose(a, ref, rtol=1e-10)
    with pytest.raises(ValueError):
        xtrap.volume.factory_extrapmodel(uv=uv, xv=xv, order=2, dxdqv=xv, volume=volume, ndim=1)
    boot = xem.resample(sampler={"nrep": 6, "rng": np.random.default_rng(1)})
    assert boot.predict(volumes).values.shape == (4, 6, 5)

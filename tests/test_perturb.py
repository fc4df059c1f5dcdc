"""PerturbModel.predict (models.py:1009-1046; reference test tests/test_beta.py:154-162).

CPU: the oracle restatement against closed forms, and the host class through the oracle-backed
stand-in.  GPU: ``txb_perturb_predict_f64`` against the oracle, tolerance ``1e-10 * max(|ref|, sigma_x)``.
"""

import numpy as np
import pytest

from oracle import perturb as op


def _samples(n, nchan=None, seed=0):
    rng = np.random.default_rng(seed)
    u = rng.normal(3.0, 1.5, n)
    shape = (n,) if nchan is None else (n, nchan)
    x = rng.normal(0.0, 1.0, shape) + (0.7 * u if nchan is None else 0.7 * u[:, None] * np.linspace(0.5, 2.0, nchan))
    return x, u


def _golden():
    import json
    from pathlib import Path

    g = json.loads((Path(__file__).parent / "golden" / "perturb.json").read_text())
    rng = np.random.default_rng(g["seed"])  # the reference's FixtureData draw order (tests/conftest.py:15-28)
    u = rng.random(g["n"])
    x = rng.random((g["n"], g["nv"]))
    assert abs(u.sum() - g["check_u_sum"]) < 1e-12 and abs(x.sum() - g["check_x_sum"]) < 1e-10
    return g, x, u


def test_oracle_matches_reference_legacy_golden():
    """tests/golden/perturb.json holds ``legacy.PerturbModel.predict`` run from /root/reference (make_golden.py)."""
    g, x, u = _golden()
    out = op.perturb_predict(x, u, np.asarray(g["betas"]) - g["beta0"])
    assert np.allclose(out, np.asarray(g["predict"]), rtol=1e-13, atol=0)


@pytest.mark.gpu
def test_kernel_matches_reference_legacy_golden():
    from thermoextrap_b200 import kernels as K

    g, x, u = _golden()
    out = K.perturb_predict(x, u, np.asarray(g["betas"]) - g["beta0"]).cpu().numpy()
    assert np.allclose(out, np.asarray(g["predict"]), rtol=1e-10, atol=0)


def test_oracle_dalpha_zero_is_the_mean():
    x, u = _samples(1000, 3)
    assert np.allclose(op.perturb_predict(x, u, [0.0])[0], x.mean(axis=0), rtol=1e-13)


def test_oracle_matches_gaussian_closed_form():
    """For u ~ N(mu, s^2) and x = u the reweighted mean is mu - dalpha * s^2."""
    rng = np.random.default_rng(1)
    mu, s, n = 2.0, 0.8, 2_000_000
    u = rng.normal(mu, s, n)
    for da in (-0.3, 0.2):
        got = op.perturb_predict(u, u, [da])[0]
        assert abs(got - (mu - da * s * s)) < 6 * s / np.sqrt(n) * np.exp(da * da * s * s)


def test_model_predict_host_logic(backend_kernels):
    import thermoextrap_b200 as xt

    x, u = _samples(500, 4, seed=2)
    xv, uv = xt.DataArray(x, dims=("rec", "val")), xt.DataArray(u, dims=("rec",))
    pm = xt.beta.factory_perturbmodel(0.5, uv=uv, xv=xv)
    out = pm.predict([0.3, 0.7])
    assert out.dims == ("beta", "val") and out.shape == (2, 4)
    ref = op.perturb_predict(x, u, np.array([0.3, 0.7]) - 0.5)
    assert np.allclose(out.values, ref, rtol=1e-12)
    one = pm.predict(0.3)
    assert one.dims == ("val",) and np.allclose(one.values, ref[0], rtol=1e-12)
    # resampled data: uv[rep, rec], xv[rep, rec, val] -> (beta, rep, val), each replicate reweighted on its own
    idx = np.random.default_rng(3).choice(500, size=(6, 500))
    boot = pm.resample(sampler={"indices": idx}).predict([0.3, 0.7])
    assert boot.dims == ("beta", "rep", "val") and boot.shape == (2, 6, 4)
    ref2 = op.perturb_predict(x[idx[2]], u[idx[2]], np.array([0.3, 0.7]) - 0.5)
    assert np.allclose(boot.values[:, 2], ref2, rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("n,nchan,nalpha", [(1, None, 1), (37, None, 3), (100_000, None, 5), (257, 1, 2), (5000, 7, 4),
                                            (3000, 130, 9), (20_000, 2000, 2), (1_000_003, 64, 6)])
def test_kernel_matches_oracle(n, nchan, nalpha):
    from thermoextrap_b200 import kernels as K

    x, u = _samples(n, nchan, seed=n % 97)
    dalpha = np.linspace(-0.4, 0.5, nalpha)
    ref = op.perturb_predict(x, u, dalpha)
    out = K.perturb_predict(x, u, dalpha).cpu().numpy()
    assert out.shape == ref.shape
    scale = np.maximum(np.abs(ref), x.std(axis=0) if n > 1 else 1.0)
    assert float((np.abs(out - ref) / scale).max()) < 1e-10


@pytest.mark.gpu
def test_kernel_large_shift_is_stable():
    """|dalpha * u| ~ 1e4: the max shift keeps every exponent <= 0 (models.py:1040-1041)."""
    from thermoextrap_b200 import kernels as K

    x, u = _samples(50_000, 3, seed=5)
    u = 1000.0 * u
    for da in (-3.0, 4.0):
        ref = op.perturb_predict(x, u, [da])
        out = K.perturb_predict(x, u, [da]).cpu().numpy()
        assert np.isfinite(out).all() and np.allclose(out, ref, rtol=1e-9, atol=1e-12)


@pytest.mark.gpu
def test_model_predict_on_gpu():
    import thermoextrap_b200 as xt

    x, u = _samples(20_000, 4, seed=8)
    xv, uv = xt.DataArray(x, dims=("rec", "val")), xt.DataArray(u, dims=("rec",))
    out = xt.beta.factory_perturbmodel(0.5, uv=uv, xv=xv).predict([0.3, 0.7])
    assert np.allclose(out.values, op.perturb_predict(x, u, np.array([0.3, 0.7]) - 0.5), rtol=1e-10)

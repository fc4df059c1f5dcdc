"""K7 device-resident path through the public API: generated device functions (txb_beta_predict_f64), the
InterpModel combine, and stack.to_mean_var.  On the CPU these run over the oracle-backed stand-in kernels (host
logic only); ``tests/test_api_gpu.py`` re-runs them on the CUDA kernels.

Tolerance 1e-10 relative to ``max(|ref|, sum_p |term_p|)`` (derivative / Taylor sums cancel).
"""

import math

import numpy as np
import pytest

import thermoextrap_b200 as xtrap
from thermoextrap_b200 import models
from thermoextrap_b200.core.xr import DataArray

TOL = 1e-10


@pytest.fixture(autouse=True)
def _kernels(backend_kernels):
    return backend_kernels


@pytest.fixture
def calls(backend_kernels, monkeypatch):
    """Count launches of the fused kernel entry."""
    n = {"beta_predict": 0, "eval_exprs": 0}
    for name in n:
        real = getattr(backend_kernels, name)

        def wrapped(*a, _real=real, _name=name, **k):
            n[_name] += 1
            return _real(*a, **k)

        monkeypatch.setattr(backend_kernels, name, wrapped)
    return n


def _samples(seed=5, n=600, nv=3):
    rng = np.random.default_rng(seed)
    u = rng.normal(2.0, 0.6, n)
    x = 4.0 + 0.3 * u[:, None] * np.linspace(0.5, 1.5, nv)[None, :] + rng.normal(0.0, 0.2, (n, nv))
    return DataArray(x, dims=("rec", "val")), DataArray(u, dims=("rec",))


def _scaled_close(a, b, terms_scale=None):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    assert a.shape == b.shape, (a.shape, b.shape)
    s = np.abs(b) if terms_scale is None else np.maximum(np.abs(b), terms_scale)
    err = np.abs(a - b) / np.maximum(s, 1e-300)
    assert err.max() <= TOL, f"max scaled error {err.max():.3e}"


def _term_scales(records, family, order):
    """``S[i, elem] = sum_t |coef_t| prod |var|^p`` of the i-th derivative polynomial: what fp64 rounding is relative
    to when the terms cancel (raw families: <x u^n> ~ 1e2 cancel down to derivatives ~ 1e-2).  Polynomials from the
    build-time generator, moment magnitudes from the oracle's reading of the records."""
    import importlib.util
    from pathlib import Path

    from oracle import derivs as od

    path = Path(__file__).resolve().parent.parent / "thermoextrap_b200" / "csrc" / "gen_beta_derivs.py"
    spec = importlib.util.spec_from_file_location("gen_beta_derivs", path)
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    rec = np.asarray(records, dtype=float)
    lead = rec.shape[:-2]
    rec = rec.reshape(-1, 2, rec.shape[-1])
    ub, xb, du, dxdu = od.central_rows(rec, x_is_u=family in (1, 3))
    ub, xb, du = np.abs(ub).astype(float), np.abs(xb).astype(float), np.abs(du).astype(float)
    nmax = du.shape[1]
    if family < 2:
        rows = {0: du.copy(), 1: None if dxdu is None else np.abs(dxdu).astype(float)}
        rows[0][:, 1] = ub
        if rows[1] is not None:
            rows[1][:, 0] = xb
    else:   # |raw moments| <= sum_k C(n,k) |du_k| |ub|^(n-k) (+ the x row)
        dx = np.zeros_like(du) if dxdu is None else np.abs(dxdu).astype(float)
        r0 = np.stack([sum(math.comb(n, k) * du[:, k] * ub ** (n - k) for k in range(n + 1)) for n in range(nmax)], 1)
        r1 = np.stack([sum(math.comb(n, k) * ub ** (n - k) * (xb * du[:, k] + dx[:, k]) for k in range(n + 1))
                       for n in range(nmax)], 1)
        rows = {0: r0, 1: r1}
    out = np.zeros((order + 1, rec.shape[0]))
    for i, poly in enumerate(gen.family_polys(family, order)):
        for mono, c in poly.items():
            term = np.full(rec.shape[0], float(abs(c)))
            for (row, n), p in mono:
                term = term * rows[row][:, n] ** p
            out[i] += term
    return out.reshape(order + 1, *lead)


def _both(fn, monkeypatch):
    """Result of ``fn()`` on the fused path and on the expression-table path."""
    monkeypatch.setattr(models, "FUSED_K7", True)
    fast = fn()
    monkeypatch.setattr(models, "FUSED_K7", False)
    slow = fn()
    monkeypatch.setattr(models, "FUSED_K7", True)
    return fast, slow


@pytest.mark.parametrize("central", [True, False])
@pytest.mark.parametrize("minus_log", [False, True])
@pytest.mark.parametrize("mode", ["sum", "cumsum", "no_sum"])
def test_extrapmodel_fused_equals_table_path(central, minus_log, mode, monkeypatch, calls):
    """ExtrapModel.predict / derivs / coefs of x_ave: one fused launch from the device-resident comoments equals the
    expression-table route (models.py:371-377, 535-565), original and resampled data, every summation mode."""
    xv, uv = _samples()
    order, betas = 5, [0.8, 0.95, 1.1]
    kw = {"cumsum": mode == "cumsum", "no_sum": mode == "no_sum", "minus_log": minus_log}

    def run():
        data = xtrap.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=central)
        m = xtrap.beta.factory_extrapmodel(beta=1.0, data=data)
        b = m.resample(sampler={"nrep": 7, "rng": np.random.default_rng(3)})
        return (m.predict(betas, **kw), b.predict(betas, **kw), b.derivs(minus_log=minus_log), b.coefs(minus_log=minus_log),
                m.predict(0.9, **kw))

    before = calls["beta_predict"]
    fast, slow = _both(run, monkeypatch)
    assert calls["beta_predict"] - before >= 4, "the fused kernel entry was not used"
    for f, s in zip(fast, slow):
        assert f.dims == s.dims
        assert set(f.coords) == set(s.coords)
        for k in s.coords:
            assert np.array_equal(np.asarray(f.coords[k].values), np.asarray(s.coords[k].values)), k
    # tolerance scale: |coef| x |moment| products of every derivative polynomial (they cancel), Taylor-weighted for
    # the predictions; the -log transform divides by <x> and mixes the orders: bound it by the largest of them
    monkeypatch.setattr(models, "FUSED_K7", False)
    data = xtrap.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=central)
    m = xtrap.beta.factory_extrapmodel(beta=1.0, data=data)
    b = m.resample(sampler={"nrep": 7, "rng": np.random.default_rng(3)})
    fam = 0 if central else 2
    S_m, S_b = _term_scales(m.data.values.values, fam, order), _term_scales(b.data.values.values, fam, order)
    if minus_log:
        amp = 2.0**order / np.abs(np.asarray(m.data.values.values)[..., 1, 0]).min()
        S_m, S_b = np.maximum.accumulate(S_m, 0) * amp, np.maximum.accumulate(S_b, 0) * amp
    fact = np.array([math.factorial(i) for i in range(order + 1)], dtype=float)

    def pred_scale(S, dbs):
        w = np.array([[abs(db) ** p / fact[p] for p in range(order + 1)] for db in dbs])    # [alpha, order]
        terms = w.reshape(*w.shape, *([1] * (S.ndim - 1))) * S[None]
        return terms.sum(1) if mode == "sum" else (np.cumsum(terms, 1) if mode == "cumsum" else terms)

    dbs = [bb - 1.0 for bb in betas]
    _scaled_close(fast[0].values, slow[0].values, pred_scale(S_m, dbs))
    _scaled_close(fast[1].values, slow[1].values, pred_scale(S_b, dbs))
    _scaled_close(fast[2].values, slow[2].values, S_b)
    _scaled_close(fast[3].values, slow[3].values, S_b / fact.reshape(-1, *([1] * (S_b.ndim - 1))))
    _scaled_close(fast[4].values, slow[4].values, pred_scale(S_m, [0.9 - 1.0])[0])


def test_u_ave_fused_on_x_is_u_data(monkeypatch, calls):
    """u_ave families read <du^n> as <dx du^(n-1)> of x_is_u records, up to n = order + 1."""
    _, uv = _samples()
    for central in (True, False):
        def run():
            data = xtrap.DataCentralMomentsVals.from_vals(xv=None, uv=uv, order=4, central=central, x_is_u=True)
            m = xtrap.beta.factory_extrapmodel(beta=1.0, data=data, name="u_ave")
            b = m.resample(sampler={"nrep": 5, "rng": np.random.default_rng(1)})
            return m.predict([0.9, 1.2]), b.predict([0.9, 1.2]), b.derivs()

        before = calls["beta_predict"]
        fast, slow = _both(run, monkeypatch)
        assert calls["beta_predict"] > before
        for f, s in zip(fast, slow):
            assert f.dims == s.dims
            _scaled_close(f.values, s.values, terms_scale=1e2 * np.abs(s.values).max() if not central else None)


def test_fused_path_is_skipped_when_it_does_not_apply(monkeypatch, calls):
    """post_func / n-th moment families keep the table path; comoments built on the host are uploaded (they are small)
    and take the fused launch like the kernel-produced ones."""
    xv, uv = _samples()
    data = xtrap.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=3, central=True)
    before = calls["beta_predict"]
    m = xtrap.beta.factory_extrapmodel(beta=1.0, data=data, post_func="minus_log")
    assert m.derivatives.family is None
    m.predict([0.9])
    m2 = xtrap.beta.factory_extrapmodel(beta=1.0, data=data, name="dxdun_ave", n=2, order=1)
    m2.predict([0.9])
    assert calls["beta_predict"] == before
    host = xtrap.DataCentralMoments.from_data(data.values, central=True)   # built from a host array: no device copy
    m3 = xtrap.beta.factory_extrapmodel(beta=1.0, data=host)
    assert m3.derivatives.family == 0
    fast = m3.predict([0.9, 1.2]).values
    assert calls["beta_predict"] == before + 1
    monkeypatch.setattr(models, "FUSED_K7", False)
    slow = xtrap.beta.factory_extrapmodel(beta=1.0, data=host).predict([0.9, 1.2]).values
    _scaled_close(fast, slow)


@pytest.mark.parametrize("central", [True, False])
def test_lnpi_fused_equals_table_path(central, monkeypatch, backend_kernels):
    """The lnPi family (lnpi.py:42-116) on the fused launch: d_0 = lnPi0, d_1 = mu.N - <u>, d_n = -d^(n-1)<u>:
    from samples (uv[rec, n], x_is_u) and from pre-averaged moments (from_ave_raw, the reference's golden path),
    model order = data order + 1."""
    n_dim, order = 6, 3
    rng = np.random.default_rng(4)
    nvals = np.arange(1, n_dim + 1)
    uv = rng.normal(-4.8 * nvals, 0.9 * np.sqrt(nvals), (900, n_dim))
    lnpi0 = DataArray(rng.normal(size=n_dim), dims=("n",), coords={"n": nvals})
    mu = DataArray(np.array([-2.63]), dims=("comp",))
    n_calls = {"k": 0}
    real = backend_kernels.lnpi_predict

    def counted(*a, **k):
        n_calls["k"] += 1
        return real(*a, **k)

    monkeypatch.setattr(backend_kernels, "lnpi_predict", counted)

    def run():
        meta = xtrap.lnpi.lnPiDataCallback(lnpi0, mu, dims_n=["n"], dims_comp="comp")
        d1 = xtrap.DataCentralMomentsVals.from_vals(xv=None, uv=DataArray(uv, dims=("rec", "n")), order=order, central=central,
                                                    x_is_u=True, meta=meta)
        m1 = xtrap.lnpi.factory_extrapmodel_lnPi(beta=1.0, data=d1)
        raw_u = np.stack([(uv**j).mean(0) for j in range(order + 2)], axis=-1)          # <u^j>, j = 0 .. order + 1
        d2 = xtrap.DataCentralMoments.from_ave_raw(u=DataArray(raw_u, dims=("n", "umom")), xu=None, x_is_u=True,
                                                   central=central, meta=meta)
        m2 = xtrap.lnpi.factory_extrapmodel_lnPi(beta=1.0, data=d2)
        return (m1.predict([0.9, 1.1]), m1.derivs(), m2.predict([0.9, 1.1], cumsum=True), m2.coefs(),
                m1.predict(1.05, order=2))

    fast, slow = _both(run, monkeypatch)
    assert n_calls["k"] >= 5, "the fused lnPi entry was not used"
    # tolerance scale per order: the u_ave polynomial's |coef| |moment| products one order lower (raw form: <u^j> of
    # macrostate n reaches |4.8 n|^j and cancels), plus |lnPi0| and |mu.N| for orders 0 and 1
    monkeypatch.setattr(models, "FUSED_K7", False)
    meta = xtrap.lnpi.lnPiDataCallback(lnpi0, mu, dims_n=["n"], dims_comp="comp")
    d1 = xtrap.DataCentralMomentsVals.from_vals(xv=None, uv=DataArray(uv, dims=("rec", "n")), order=order, central=central,
                                                x_is_u=True, meta=meta)
    L = order + 1
    Su = _term_scales(d1.values.values, 1 if central else 3, L - 1)                 # [L, n]
    Sl = np.concatenate([np.abs(lnpi0.values)[None], (np.abs(meta.mudotN.values) + Su[0])[None], Su[1:]], axis=0)
    fact = np.array([math.factorial(i) for i in range(L + 1)], dtype=float)

    def pred_scale(dbs, upto, cumsum=False):
        w = np.array([[abs(db) ** p / fact[p] for p in range(upto + 1)] for db in dbs])
        terms = w[:, :, None] * Sl[None, : upto + 1]
        return np.cumsum(terms, 1) if cumsum else terms.sum(1)

    for f, s in zip(fast, slow):
        assert f.dims == s.dims and f.shape == s.shape
    _scaled_close(fast[0].values, slow[0].values, pred_scale([-0.1, 0.1], L))
    _scaled_close(fast[1].values, slow[1].values, Sl)
    _scaled_close(fast[2].values, slow[2].values, pred_scale([-0.1, 0.1], L, cumsum=True))
    _scaled_close(fast[3].values, slow[3].values, Sl / fact[:, None])
    _scaled_close(fast[4].values, slow[4].values, pred_scale([0.05], 2)[0])


def test_interpmodel_fused_equals_table_path(monkeypatch, calls):
    """InterpModel over 3 states with rep x val dims: derivatives, mat_inv . derivs and the polynomial on the device."""
    betas0 = [0.7, 1.0, 1.4]

    def run():
        states = []
        for i, b0 in enumerate(betas0):
            xv, uv = _samples(seed=20 + i)
            data = xtrap.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=3, central=True)
            states.append(xtrap.beta.factory_extrapmodel(beta=b0, data=data))
        im = xtrap.InterpModel(states)
        rs = im.resample(sampler={"nrep": 6, "rng": np.random.default_rng(8)})
        return im.predict([0.8, 1.2]), rs.predict([0.8, 1.2]), rs.coefs(), rs.predict(0.9)

    before = calls["beta_predict"]
    fast, slow = _both(run, monkeypatch)
    assert calls["beta_predict"] - before >= 9

    # tolerance scale: the 12 x 12 Hermite inverse has entries ~ 1e5 that cancel in mat_inv . derivs; rounding is
    # relative to  sum_p |alpha|^p sum_(state, order) |mat_inv[p, s, o]| |derivs[s][o]|  (models.py:893-912, 929-932)
    def scales(im, alphas):
        order = im.order
        minv = np.abs(im._mat_inv(order))                                                  # [p, s, o]
        d = np.stack([np.abs(np.asarray(m.derivs(order, norm=False).values)) for m in im.states])   # [s, o, *lead]
        coef = np.tensordot(minv, d, axes=([1, 2], [0, 1]))                                # [p, *lead]
        pw = np.abs(np.asarray(alphas, dtype=float))[:, None] ** np.arange(coef.shape[0])[None, :]
        return np.tensordot(pw, coef, axes=(1, 0)), coef

    monkeypatch.setattr(models, "FUSED_K7", False)
    states = []
    for i, b0 in enumerate(betas0):
        xv, uv = _samples(seed=20 + i)
        data = xtrap.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=3, central=True)
        states.append(xtrap.beta.factory_extrapmodel(beta=b0, data=data))
    im = xtrap.InterpModel(states)
    rs = im.resample(sampler={"nrep": 6, "rng": np.random.default_rng(8)})
    for f, s in zip(fast, slow):
        assert f.dims == s.dims and f.shape == s.shape
    _scaled_close(fast[0].values, slow[0].values, scales(im, [0.8, 1.2])[0])
    _scaled_close(fast[1].values, slow[1].values, scales(rs, [0.8, 1.2])[0])
    _scaled_close(fast[2].values, slow[2].values, scales(rs, [1.0])[1])
    _scaled_close(fast[3].values, slow[3].values, scales(rs, [0.9])[0][0])
    # and the property that pins both: the interpolant reproduces every state's own value at its alpha0
    for path in (True, False):
        monkeypatch.setattr(models, "FUSED_K7", path)
        at_nodes = im.predict(betas0).values
        own = np.stack([np.asarray(m.derivs(0).values)[0] for m in im.states])
        _scaled_close(at_nodes, own, scales(im, betas0)[0])


@pytest.mark.parametrize("ddof", [0, 1])
def test_to_mean_var_matches_numpy(ddof):
    """stack.to_mean_var (stack.py:157-182): mean / var over ``rep`` stacked along ``stats``."""
    rng = np.random.default_rng(2)
    vals = rng.normal(3.0, 1.0, (5, 700, 9)) * np.linspace(1, 1e3, 9)
    da = DataArray(vals, dims=("order", "rep", "val"), coords={"val": np.arange(9) * 2.0})
    out = xtrap.stack.to_mean_var(da, "rep", ddof=ddof)
    assert out.dims == ("stats", "order", "val")
    assert list(out.coords["stats"].values) == ["mean", "var"]
    assert np.array_equal(np.asarray(out.coords["val"].values), np.arange(9) * 2.0)
    ld = vals.astype(np.longdouble)
    _scaled_close(out.values[0], ld.mean(axis=1).astype(float))
    _scaled_close(out.values[1], ld.var(axis=1, ddof=ddof).astype(float))
    last = xtrap.stack.to_mean_var(DataArray(vals, dims=("a", "b", "rep")), "rep")   # reduced axis last
    _scaled_close(last.values[1], ld.var(axis=2).astype(float))
    with pytest.raises(ValueError):
        xtrap.stack.to_mean_var(da, "nope")


def test_states_derivs_concat():
    states = []
    for i, b0 in enumerate([0.7, 1.0]):
        xv, uv = _samples(seed=30 + i)
        data = xtrap.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=3, central=True)
        states.append(xtrap.beta.factory_extrapmodel(beta=b0, data=data))
    sc = xtrap.StateCollection(states).resample(sampler={"nrep": 4, "rng": np.random.default_rng(0)})
    out = xtrap.stack.states_derivs_concat(sc)
    assert out.dims == ("beta", "order", "rep", "val") and out.shape == (2, 4, 4, 3)
    assert np.array_equal(np.asarray(out.coords["beta"].values), [0.7, 1.0])
    ref = np.stack([np.asarray(s.derivs(norm=False).values) for s in sc])
    assert np.array_equal(out.values, ref)
    mv = xtrap.stack.to_mean_var(out, "rep")
    assert mv.dims == ("stats", "beta", "order", "val")
    fact = math.factorial(3)
    normed = xtrap.stack.states_derivs_concat(sc, norm=True)
    _scaled_close(normed.values[:, 3], ref[:, 3] / fact)

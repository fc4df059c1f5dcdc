"""Round-2 kernels and paths on the real device, against the oracle (``-m gpu``).

Tolerance: integers bit-exact; fp64 ``|a - b| <= 1e-10 * max(|b|, sigma_x^i sigma_u^j)`` (north_star's rtol
with the SURVEY.md 8c scale: odd central moments are ~ 0, plain rtol is ill-posed there).
"""

import math

import numpy as np
import pytest

from oracle import comoments as oc

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def K():
    from thermoextrap_b200 import kernels

    return kernels


def close(got, ref, scale, tol=TOL):
    got, ref = np.asarray(got), np.asarray(ref)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    err = np.abs(got - ref) / np.maximum(np.abs(ref), scale)
    assert err.max() <= tol, f"max scaled error {err.max():.3e}"


# ------------------------------------------------------------------ planar K1 (volume path, config 5)
@pytest.mark.parametrize("n,nplanes,order", [(5000, 1, 1), (100_003, 2, 1), (1_000_000, 2, 3), (300_001, 4, 6), (777, 3, 2)])
def test_reduce_planar_matches_oracle(K, n, nplanes, order):
    """txb_reduce_sums_planar_f64 -> txb_finalize_f64 == two-pass longdouble comoments of the stacked planes."""
    rng = np.random.default_rng(n)
    u = rng.normal(0.3, 1.1, n)
    planes = [rng.uniform(0, 1, n) + 0.1 * (c + 1) * u for c in range(nplanes)]
    pivot, S0, S1 = K.reduce_sums_planar(planes, u, order)
    out = K.finalize_sums(pivot, S0, S1, order, val_shape=(nplanes,)).cpu().numpy()
    x = np.stack(planes, axis=1)
    ref = oc.comoments_two_pass(x, u, order)
    close(out, ref, oc.tolerance_scale(x, u, order))
    # same thing as the interleaved wide kernel
    wide = K.reduce_vals(x, u, order).cpu().numpy()
    close(out, wide, oc.tolerance_scale(x, u, order), tol=1e-11)


def test_volume_stream_equals_resident_model_with_drifting_chunks(K):
    """volume.factory_extrapmodel_stream (double-buffered H2D ring, per-chunk pivots, Chan/Pebay merge) ==
    volume.factory_extrapmodel on the concatenated samples, with chunk means that drift by many sigma
    (the case a first-chunk pivot would lose digits on) and ragged chunk sizes."""
    import thermoextrap_b200 as xt

    rng = np.random.default_rng(3)
    sizes = [40_000, 65_536, 1, 30_001, 123_457]
    xs, qs, ws = [], [], []
    for i, m in enumerate(sizes):
        w = rng.normal(50.0 * i, 1.0, m)          # mean drifts by 50 sigma per chunk
        x = rng.uniform(0, 1, m) + 0.1 * w
        xs.append(x), qs.append(0.5 * x + 0.01 * rng.normal(size=m)), ws.append(w)
    x, q, w = np.concatenate(xs), np.concatenate(qs), np.concatenate(ws)
    volume, vols = 2.0, [1.5, 2.0, 2.6]
    ref_model = xt.volume.factory_extrapmodel(volume=volume, uv=xt.xrwrap_uv(w), xv=xt.DataArray(x, dims=("rec",)),
                                              dxdqv=xt.DataArray(q, dims=("rec",)), ndim=3, order=1)
    stream_model = xt.volume.factory_extrapmodel_stream(volume, zip(xs, qs, ws), ndim=3)
    a, b = stream_model.predict(vols).values, ref_model.predict(vols).values
    # closed form (volume.py:73-77) in longdouble as the oracle
    xl, ql, wl = (np.asarray(v, dtype=np.longdouble) for v in (x, q, w))
    d0 = xl.mean()
    d1 = (-d0 * wl.mean() + (xl * wl).mean() + ql.mean()) / (volume * 3)
    oracle = np.array([float(d0 + d1 * (v - volume)) for v in vols])
    scale = float(abs(d0) + (abs(d0 * wl.mean()) + abs((xl * wl).mean()) + abs(ql.mean())) / (volume * 3))
    close(a, oracle, scale)
    close(b, oracle, scale)
    with pytest.raises((NotImplementedError, ValueError)):   # nothing to resample: no samples, no rec dim
        stream_model.resample(sampler={"nrep": 3})
    # value dims go through the wide kernels
    xs2 = [np.stack([v, 2 * v], axis=1) for v in xs]
    qs2 = [np.stack([v, -v], axis=1) for v in qs]
    m2 = xt.volume.factory_extrapmodel_stream(volume, zip(xs2, qs2, ws), ndim=3)
    p2 = m2.predict(vols).values
    assert p2.shape == (3, 2)
    close(p2[:, 0], oracle, scale)


# ------------------------------------------------------------------ column-range frequency table (sample sharding)
@pytest.mark.parametrize("ndat,nrep,cuts", [(700_000, 5, [0, 131_040, 400_000, 700_000]), (1_000_003, 3, [0, 500_001, 1_000_003])])
def test_sampler_freq_cols_bit_exact(K, ndat, nrep, cuts):
    """txb_sampler_freq_cols: the columns of one row range of numpy's table, bit-exact, and row sums that add up
    to nsamp over a partition of the rows (the overflow check of the sample-sharded bootstrap)."""
    import torch

    g = np.random.default_rng(9)
    plan = K.SamplerPlan(g, nrep, ndat)
    ref = oc.indices_to_freq(np.random.default_rng(9).choice(ndat, (nrep, ndat)), ndat)
    total = torch.zeros(nrep, dtype=torch.int64, device="cuda")
    for r in range(len(cuts) - 1):
        row_lo, row_hi = cuts[r], cuts[r + 1]
        col_lo = row_lo // K.FREQ_BUCKET * K.FREQ_BUCKET
        nxt = ndat if r == len(cuts) - 2 else cuts[r + 1] // K.FREQ_BUCKET * K.FREQ_BUCKET
        f, rs = K.sampler_freq_cols(plan, col_lo, row_hi, sum_cols=max(0, min(row_hi, nxt) - col_lo))
        assert np.array_equal(f.cpu().numpy().astype(np.int64), ref[:, col_lo:row_hi])
        total += rs
    assert total.tolist() == [ndat] * nrep


def test_sample_sharded_sums_equal_whole(K):
    """Pivot sums of row shards (each with the frequency columns of its own rows) add up to the sums of the
    whole sample set: the single-process form of parallel.bootstrap_sample_sharded."""
    import torch

    rng = np.random.default_rng(4)
    n, nv, order, nrep = 900_000, 8, 4, 6
    u = rng.normal(175.0, 5.3, n)
    x = 1e-3 * u[:, None] + rng.normal(0.0, 0.02, (n, nv))
    plan = K.SamplerPlan(np.random.default_rng(21), nrep, n)
    idx = np.random.default_rng(21).choice(n, (nrep, n))
    pivot, _, _, _ = K.reduce_sums(x[:1024], u[:1024], 1)
    cuts = [0, 300_000, 533_333, 900_000]
    tot0 = tot1 = rows = None
    for r in range(3):
        nxt = None if r == 2 else cuts[r + 1]
        S0, S1, val_shape, rs = K.sampler_resample_sums_rows(plan, x[cuts[r]:cuts[r + 1]], u[cuts[r]:cuts[r + 1]], order,
                                                             cuts[r], pivot, next_row_lo=nxt)
        tot0 = S0 if tot0 is None else tot0 + S0
        tot1 = S1 if tot1 is None else tot1 + S1
        rows = rs if rows is None else rows + rs
    assert rows.tolist() == [n] * nrep
    out = K.finalize_sums(pivot, tot0, tot1, order, val_shape=val_shape).cpu().numpy()
    ref = oc.resample_vals_gather(x, u, order, idx)
    close(out, ref, oc.tolerance_scale(x, u, order)[None])
    whole = K.sampler_resample_wide(plan, x, u, order=order).cpu().numpy()
    close(out, whole, oc.tolerance_scale(x, u, order)[None], tol=1e-11)
    del torch


# ------------------------------------------------------------------ C2's real width against the oracle
@pytest.mark.parametrize("variant", ["plain", "weighted", "x_is_u"])
def test_bucketed_bootstrap_at_config2_width_matches_oracle(K, variant):
    """The headline scalar path (k_bucket_scatter + k_bucket_consume) at BASELINE config 2's own width -- 1e7 samples,
    order 8 -- for 3 replicates of numpy's table, against the longdouble gather-then-reduce oracle."""
    n, order, nrep = 10_000_000, 8, 3
    rng = np.random.default_rng(0)
    u = rng.normal(175.0, 5.3, n)
    x = 1e-3 * u + rng.normal(0.0, 0.02, n)
    w = rng.random(n) + 0.5 if variant == "weighted" else None
    xis = variant == "x_is_u"
    idx = np.random.default_rng(5).choice(n, (nrep, n))
    status: list = []
    out = K.sampler_resample(np.random.default_rng(5), None if xis else x, u, nrep, order, w=w, x_is_u=xis,
                             sync_rng=False, status_out=status)
    flag = int(status[0].item())
    assert flag & 32 and not flag & 24, f"bucketed path did not run cleanly (status {flag})"
    ref = oc.resample_vals_gather(u if xis else x, u, order, idx, w=w)
    close(out.cpu().numpy().reshape(ref.shape), ref, oc.tolerance_scale(u if xis else x, u, order))


# ------------------------------------------------------------------ ADVICE r1 fixes in the C ABI
def test_poly_predict_many_alphas_and_pow_zero(K):
    rng = np.random.default_rng(1)
    coefs = rng.normal(size=(4, 3))
    d = np.linspace(-0.3, 0.3, 70_000)   # > 65535: the grid's y dimension is strided
    out = K.poly_predict(coefs, d).cpu().numpy()
    ref = sum(coefs[p][None, :] * d[:, None] ** p for p in range(4))
    np.testing.assert_allclose(out, ref, rtol=1e-12, atol=1e-15)
    # factor power 0 evaluates to 1 (not to the variable)
    v = rng.normal(size=(2, 5))
    got = K.eval_exprs(v, [0, 1], [2.5], [0, 2], [0, 1], [0, 2]).cpu().numpy()
    np.testing.assert_allclose(got[0], 2.5 * v[1] ** 2, rtol=1e-14)


def test_binomials_per_device(K):
    """finalize on a second device in the same process must not run with unset binomials."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    from thermoextrap_b200 import _lib

    lib = _lib.load()
    rng = np.random.default_rng(0)
    data = rng.normal(size=(3, 2, 2, 4))
    data[..., 0, 0] = 10.0
    outs = []
    for d in (0, 1):
        with torch.cuda.device(d):
            t = torch.as_tensor(data, device=f"cuda:{d}")
            o = torch.empty((2, 2, 4), dtype=torch.float64, device=f"cuda:{d}")
            _lib.check(lib.txb_reduce_data_f64(t.data_ptr(), 3, 2, 3, o.data_ptr(), None), "txb_reduce_data_f64")
            torch.cuda.synchronize(d)
            outs.append(o.cpu().numpy())
    np.testing.assert_array_equal(outs[0], outs[1])


# ------------------------------------------------------------------ wide-x contraction kernels, every form
@pytest.mark.parametrize("version", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("dtype", ["uint8", "uint16", "int32", "int64"])
@pytest.mark.parametrize("n,nv,nrep,order,kind", [
    (1500, 17, 150, 1, "x"), (2111, 64, 130, 6, "x"), (1203, 33, 70, 7, "xw"), (1500, 5, 129, 6, "x"),
    (900, 16, 9, 8, "x"), (640, 70, 20, 10, "xw"), (1500, 20, 140, 5, "upc"), (777, 9, 66, 7, "upc"),
])
def test_wide_contraction_kernel_forms_match_oracle(K, version, dtype, n, nv, nrep, order, kind):
    """``txb_set_resample_kernel``: the DFMA tile (0), k_resample_mma (1) and k_resample_mma2 with 8 warps (2) and
    16 warps (3) all meet the gather oracle: ragged replicate / channel / sample tiles, weights, both replicate
    tile sizes (orders <= 8 / > 8), the per-channel-u form, every table dtype."""
    import torch

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    rng = np.random.default_rng(21)
    u = rng.normal(175.0, 5.3, n)
    x = u[:, None] * np.linspace(0.5e-3, 2e-3, nv)[None, :] + rng.normal(0.0, 0.02, (n, nv))
    idx = rng.choice(n, (nrep, n), replace=True)
    freq = K.indices_to_freq(idx, n, dtype=getattr(torch, dtype))
    sub = sorted({0, 1, nrep // 2, nrep - 2, nrep - 1})
    prev = lib.txb_set_resample_kernel(version)
    try:
        assert lib.txb_set_resample_kernel(-1) == version
        if kind == "upc":
            uu = u[:, None] + 50.0 * x
            out = K.resample_vals(None, uu, freq, order, x_is_u=True).cpu().numpy()
            ref = oc.resample_vals_gather(uu, uu, order, idx[sub])
            close(out[sub], ref, oc.tolerance_scale(uu, uu, order))
            return
        w = rng.random(n) if kind == "xw" else None
        out = K.resample_vals(x, u, freq, order, w=w).cpu().numpy()
        ref = oc.resample_vals_gather(x, u, order, idx[sub], w=w)
        close(out[sub], ref, oc.tolerance_scale(x, u, order))
    finally:
        lib.txb_set_resample_kernel(prev)


# ------------------------------------------------------------------ partition kernel: 512- and 1024-thread forms
def test_scatter_thread_forms_agree(K):
    """``txb_set_scatter_threads``: the 1024-thread partition kernel (4 LCG steps per thread) and the 512-thread
    form (8 steps, the default) walk the same 4096-step iterations of numpy's stream: frequency tables bit-identical
    (and equal to numpy's own table), scalar bootstrap sums equal to rounding (list order differs)."""
    import torch

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    prev = lib.txb_set_scatter_threads(-1)
    try:
        ndat, nrep = 2_000_003, 4
        ref = oc.indices_to_freq(np.random.default_rng(17).choice(ndat, size=(nrep, ndat), replace=True), ndat)
        n2, order = 4_200_000, 6
        rng = np.random.default_rng(3)
        u = rng.normal(175.0, 5.3, n2)
        x = 1e-3 * u + rng.normal(0.0, 0.02, n2)
        got = {}
        for nth in (512, 1024):
            lib.txb_set_scatter_threads(nth)
            assert lib.txb_set_scatter_threads(-1) == nth
            f = K.sampler_freq(np.random.default_rng(17), nrep, ndat, ndat, dtype=torch.uint8)
            assert np.array_equal(f.cpu().numpy().astype(np.int64), ref), f"{nth}-thread table differs from numpy's"
            status: list = []
            b = K.sampler_resample(np.random.default_rng(5), x, u, 3, order, sync_rng=False, status_out=status)
            flag = int(status[0].item())
            assert flag & 32 and not flag & 24, f"bucketed path did not run cleanly (status {flag})"
            got[nth] = b.cpu().numpy()
        idx = np.random.default_rng(5).choice(n2, (3, n2))
        ref_b = oc.resample_vals_gather(x, u, order, idx)
        for nth in (512, 1024):
            close(got[nth].reshape(ref_b.shape), ref_b, oc.tolerance_scale(x, u, order))
    finally:
        lib.txb_set_scatter_threads(prev)


# ------------------------------------------------------------------ K7 generated device functions vs the oracle
def _records(seed, nelem, k, x_is_u=False):
    """Central comoment records [nelem, 2, k+1] of genuinely different sample sets (one per element)."""
    rng = np.random.default_rng(seed)
    out = np.empty((nelem, 2, k + 1))
    for e in range(nelem):
        u = rng.normal(1.0 + 0.1 * e, 0.4, 500)
        x = u if x_is_u else 3.0 + 0.5 * u + 0.2 * u * u + rng.normal(0.0, 0.1, 500)
        out[e] = oc.comoments_two_pass(x, u, k)[0] if np.ndim(x) > 1 else oc.comoments_two_pass(x, u, k).reshape(2, k + 1)
    return out


@pytest.mark.parametrize("family", [0, 1, 2, 3])
@pytest.mark.parametrize("order", [0, 1, 3, 6, 8])
def test_beta_fused_derivs_and_predict_match_oracle(K, family, order):
    """txb_beta_predict_f64 (generated device functions) against oracle/derivs.py, which lambdifies the REFERENCE's
    expression strings (tests/golden/exprs.json): derivatives, normalised coefficients, and the three Taylor modes.
    Scale: sum over the polynomial's |coef| |moment| products (the terms cancel)."""
    from oracle import derivs as od
    from test_fused_k7 import _term_scales

    rec = _records(40 + family, 37, order + 1, x_is_u=family in (1, 3))   # records carry one moment more than needed
    S = _term_scales(rec, family, order)
    ref = od.beta_derivs(rec, order, family)
    got = K.beta_predict(rec, order, family, mode=K.PREDICT_DERIVS).cpu().numpy()
    close(got, ref, S)
    fact = np.array([math.factorial(i) for i in range(order + 1)], dtype=float)[:, None]
    close(K.beta_predict(rec, order, family, mode=K.PREDICT_DERIVS, norm=True).cpu().numpy(), ref / fact, S / fact)
    dal = np.array([-0.3, 0.05, 0.4])
    w = np.abs(dal)[:, None] ** np.arange(order + 1)[None, :] / fact[:, 0][None, :]
    terms = w[:, :, None] * S[None]
    for mode, sc in ((K.PREDICT_SUM, terms.sum(1)), (K.PREDICT_CUMSUM, np.cumsum(terms, 1)), (K.PREDICT_TERMS, terms)):
        close(K.beta_predict(rec, order, family, dal, mode=mode).cpu().numpy(),
              od.beta_predict(rec, order, family, dal, mode=mode), sc)


@pytest.mark.parametrize("family", [0, 2])
@pytest.mark.parametrize("order", [1, 3, 5])
def test_beta_fused_minus_log_matches_reference_expressions(K, family, order):
    """The in-kernel -log recurrence against the reference's SymMinusLog expressions (golden, orders <= 5)."""
    from oracle import derivs as od
    from test_fused_k7 import _term_scales

    rec = _records(50 + family, 29, order)
    ref = od.beta_derivs(rec, order, family, minus_log=True)
    got = K.beta_predict(rec, order, family, minus_log=True, mode=K.PREDICT_DERIVS).cpu().numpy()
    S = np.maximum.accumulate(_term_scales(rec, family, order), 0) * 2.0**order / np.abs(rec[:, 1, 0]).min()
    close(got, ref, S)


def test_beta_fused_orders_9_10_match_the_expression_table_kernel(K):
    """Orders beyond the golden file: the generated functions against the table interpreter fed with the package's
    sympy expressions (themselves golden-pinned to order 8 and recurrence-checked beyond, tests/test_beta_codegen.py)."""
    import thermoextrap_b200 as xtrap
    from thermoextrap_b200 import models
    from thermoextrap_b200.core.xr import DataArray

    rng = np.random.default_rng(9)
    u = rng.normal(1.0, 0.3, 4000)
    x = 2.0 + 0.4 * u[:, None] + rng.normal(0.0, 0.05, (4000, 3))
    for central in (True, False):
        data = xtrap.DataCentralMomentsVals.from_vals(xv=DataArray(x, dims=("rec", "val")), uv=DataArray(u, dims=("rec",)),
                                                      order=10, central=central)
        m = xtrap.beta.factory_extrapmodel(beta=1.0, data=data)
        models.FUSED_K7 = True
        fast = m.derivs().values
        try:
            models.FUSED_K7 = False
            slow = xtrap.beta.factory_extrapmodel(beta=1.0, data=data).derivs().values
        finally:
            models.FUSED_K7 = True
        from test_fused_k7 import _term_scales

        close(fast, slow, _term_scales(data.values.values, 0 if central else 2, 10))


def test_combine_and_mean_var_kernels(K):
    rng = np.random.default_rng(4)
    W, D = rng.normal(size=(19, 35)), rng.normal(size=(35, 3001))
    ref = W.astype(np.longdouble) @ D.astype(np.longdouble)
    sc = np.abs(W) @ np.abs(D)
    out = K.combine(W, D)
    close(out.cpu().numpy(), ref.astype(float), sc)
    out = K.combine(W, D, out=out)
    close(out.cpu().numpy(), (2 * ref).astype(float), 2 * sc)
    arr = rng.normal(5.0, 2.0, (7, 10_000, 64)) * np.logspace(-3, 3, 64)
    for ddof in (0, 1):
        mean, var = K.mean_var(arr, ddof=ddof)
        ld = arr.astype(np.longdouble)
        close(mean.cpu().numpy(), ld.mean(1).astype(float), np.abs(arr).mean(1))
        close(var.cpu().numpy(), ld.var(1, ddof=ddof).astype(float), 0.0)
    one = rng.normal(size=(2, 1, 3))
    mean, var = K.mean_var(one)
    assert np.array_equal(mean.cpu().numpy(), one[:, 0]) and not var.cpu().numpy().any()


def test_wide_bootstrap_redoes_an_overflowed_window(K):
    """kernels.sampler_resample_wide issues its uint8 windows without reading the status back; a window whose row-sum
    check (folded into k_bucket_hist) reports a count above 255 is redone with an int32 table.  ndat = 6e5 samples
    drawn 1.2e8 times per replicate (mean count 200: tens of samples exceed 255) -- the result must equal the
    contraction of the int32 table built with global atomics (itself bit-exact against numpy,
    tests/test_sampler_gpu.py::test_bucketed_histogram_overflow_is_loud)."""
    import torch

    ndat, nsamp, nrep, nchan, order = 600_000, 120_000_000, 2, 5, 3
    rng = np.random.default_rng(6)
    u = rng.normal(175.0, 5.3, ndat)
    x = u[:, None] * np.linspace(0.5e-3, 2e-3, nchan)[None, :] + rng.normal(0.0, 0.02, (ndat, nchan))
    got = K.sampler_resample_wide(np.random.default_rng(3), x, u, nrep, order, nsamp=nsamp).cpu().numpy()
    f32 = K.sampler_freq(np.random.default_rng(3), nrep, nsamp, ndat, dtype=torch.int32, buckets=False)
    assert int(f32.max()) > 255
    ref = K.resample_vals(x, u, f32, order).cpu().numpy()
    assert got.shape == ref.shape == (nrep, nchan, 2, order + 1)
    assert (got[:, :, 0, 0] == float(nsamp)).all()
    close(got, ref, oc.tolerance_scale(x, u, order), tol=1e-12)


@pytest.mark.parametrize("central", [True, False])
@pytest.mark.parametrize("order", [0, 1, 2, 5, 6])
def test_lnpi_fused_matches_reference_expressions(K, central, order):
    """txb_lnpi_predict_f64 against the reference's lnPi expression strings (tests/golden/exprs.json, lnpi.py:42-116)
    evaluated by oracle/derivs.py: x_is_u records, lnPi0 / mu.N broadcast over a leading replicate axis."""
    from oracle import derivs as od
    from test_fused_k7 import _term_scales

    nrep, nn = 5, 11
    rec = _records(60 + order, nrep * nn, max(order, 1), x_is_u=True).reshape(nrep, nn, 2, max(order, 1) + 1)
    rng = np.random.default_rng(3)
    lnpi0, mudotn = rng.normal(size=nn), -2.63 * np.arange(nn)
    ref = od.lnpi_derivs(rec, order, central, lnpi0, mudotn)
    got = K.lnpi_predict(rec, order, central, lnpi0, mudotn, mode=K.PREDICT_DERIVS).cpu().numpy()
    S = np.zeros_like(ref)
    S[0] = np.tile(np.abs(lnpi0), nrep)
    if order >= 1:
        Su = _term_scales(rec.reshape(-1, 2, rec.shape[-1]), 1 if central else 3, order - 1)
        S[1] = np.tile(np.abs(mudotn), nrep) + Su[0]
        S[2:] = Su[1:]
    close(got, ref, S)
    dal = np.array([-0.2, 0.3])
    fact = np.array([math.factorial(i) for i in range(order + 1)], dtype=float)
    terms = (np.abs(dal)[:, None] ** np.arange(order + 1)[None, :] / fact[None, :])[:, :, None] * S[None]
    for mode, sc in ((K.PREDICT_SUM, terms.sum(1)), (K.PREDICT_CUMSUM, np.cumsum(terms, 1)), (K.PREDICT_TERMS, terms)):
        close(K.lnpi_predict(rec, order, central, lnpi0, mudotn, dal, mode=mode).cpu().numpy(),
              od.lnpi_predict(rec, order, central, lnpi0, mudotn, dal, mode=mode), sc)


@pytest.mark.parametrize("central", [True, False])
def test_lnpi_fused_orders_beyond_golden_follow_u_ave(K, central):
    """lnPi orders 7..10 (the golden strings stop at 6): d_n = -d^(n-1)<u>/dbeta^(n-1) exactly as the u_ave family
    the same kernel evaluates (itself pinned to the reference's strings to order 8 and to the package's sympy classes
    to order 10), d_1 = mu.N - <u>, d_0 = lnPi0."""
    order, nn = 10, 7
    rec = _records(70, nn, order, x_is_u=True)
    rng = np.random.default_rng(5)
    lnpi0, mudotn = rng.normal(size=nn), -2.63 * np.arange(nn)
    got = K.lnpi_predict(rec, order, central, lnpi0, mudotn, mode=K.PREDICT_DERIVS).cpu().numpy()
    ud = K.beta_predict(rec, order - 1, 1 if central else 3, mode=K.PREDICT_DERIVS).cpu().numpy()
    assert np.array_equal(got[0], lnpi0)
    assert np.array_equal(got[1], mudotn - ud[0])
    assert np.array_equal(got[2:], -ud[1:])

"""The numba port of cmomy's push algorithm (bench.py's CPU baseline) against the exact oracle."""

import numpy as np
import pytest

from oracle import cmomy_port as cp
from oracle import comoments as oc


@pytest.mark.parametrize("order", [1, 3, 6, 8])
def test_reduce_matches_two_pass(order):
    rng = np.random.default_rng(0)
    n = 2000
    u = rng.normal(175.0, 5.3, n)
    x = 1e-3 * u + rng.normal(0.0, 0.02, n)
    w = rng.random(n)
    scale = oc.tolerance_scale(x, u, order)
    for wt in (None, w):
        out = cp.reduce_vals(x, u, order, w=wt)
        ref = oc.comoments_two_pass(x, u, order, w=wt)
        assert np.all(np.abs(out - ref) <= 1e-9 * np.maximum(np.abs(ref), scale))
    xw = np.stack([x, 2 * x + 1], axis=1)
    out = cp.reduce_vals(xw, u, order)
    ref = oc.comoments_two_pass(xw, u, order)
    assert np.all(np.abs(out - ref) <= 1e-9 * np.maximum(np.abs(ref), oc.tolerance_scale(xw, u, order)))


def test_resample_matches_gather():
    rng = np.random.default_rng(1)
    n, nrep, order = 500, 7, 5
    u = rng.normal(0.0, 1.0, n)
    x = u + rng.normal(0.0, 1.0, n)
    idx = rng.choice(n, size=(nrep, n), replace=True)
    freq = cp.indices_to_freq(idx, n)
    assert np.array_equal(freq, oc.indices_to_freq(idx, n))
    out = cp.resample_vals(x, u, freq, order)
    ref = oc.resample_vals_gather(x, u, order, idx)
    assert np.all(np.abs(out - ref) <= 1e-9 * np.maximum(np.abs(ref), oc.tolerance_scale(x, u, order)))
    out2 = cp.bootstrap_reference_path(x, u, order, nrep, np.random.default_rng(5))
    ref2 = oc.resample_vals_gather(x, u, order, np.random.default_rng(5).choice(n, size=(nrep, n)))
    assert np.all(np.abs(out2 - ref2) <= 1e-9 * np.maximum(np.abs(ref2), oc.tolerance_scale(x, u, order)))


def test_resample_wide_matches_gather():
    """Channels x[nsamp, nchan]: one push loop per (replicate, channel), as cmomy's gufunc broadcasts."""
    rng = np.random.default_rng(2)
    n, nv, nrep, order = 300, 3, 5, 4
    u = rng.normal(0.0, 1.0, n)
    x = u[:, None] + rng.normal(0.0, 1.0, (n, nv))
    idx = rng.choice(n, size=(nrep, n), replace=True)
    out = cp.resample_vals(x, u, cp.indices_to_freq(idx, n), order)
    ref = oc.resample_vals_gather(x, u, order, idx)
    assert out.shape == ref.shape == (nrep, nv, 2, order + 1)
    assert np.all(np.abs(out - ref) <= 1e-9 * np.maximum(np.abs(ref), oc.tolerance_scale(x, u, order)[None]))

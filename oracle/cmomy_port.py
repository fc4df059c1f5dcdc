"""CPU port of the reference's numeric backend for the hot path.  TEST / BASELINE INFRASTRUCTURE ONLY.

parity: pinned against ``oracle.comoments.comoments_two_pass`` (itself pinned to the
reference's own test oracles, see ``oracle/__init__.py``) in ``tests/test_cmomy_port.py``.

The reference (``/root/reference/src/thermoextrap/data.py:1632-1640, 1803-1810``) hands the
arithmetic to the third-party package ``cmomy`` 0.24.0 (``uv.lock:454-455``; numba
``guvectorize`` loops), which is neither vendored under ``/root/reference`` nor installable
here.  This module restates cmomy's published algorithm with the same tool (numba) so that
``bench.py`` can time "what the reference does on the host cores":

* ``push_val``      one-pass weighted update of a ``(2, k+1)`` central-comoment accumulator with
                    one sample (Welford / Pebay single-point merge), O(k^2) per sample;
* ``reduce_vals``   ``for s: push_val(x[s], u[s], w[s])``                       (cmomy K1);
* ``indices_to_freq`` histogram of the sampler's index table                    (cmomy K6);
* ``resample_vals`` ``for r: for s: if freq[r, s]: push_val(x[s], u[s], w[s] * freq[r, s], out[r])``
                    (cmomy K2) -- parallel over replicates with ``prange``, which is *more*
                    host parallelism than the reference gets for scalar ``x`` (cmomy's parallel
                    target only spans broadcast dimensions).

Layout (``data.py:1233-1236``): ``out[0,0]=sum w, out[1,0]=<x>, out[0,1]=<u>``, else
``<dx^i du^j>``.
"""

from __future__ import annotations

import numpy as np
from numba import njit, prange

_BINOM = np.zeros((16, 16))
for _n in range(16):
    _BINOM[_n, 0] = 1.0
    for _k in range(1, _n + 1):
        _BINOM[_n, _k] = _BINOM[_n - 1, _k - 1] + _BINOM[_n - 1, _k]


@njit(cache=True, inline="always")
def _get(data, i, j):
    # central view of the accumulator: [0,0] -> 1, [1,0] and [0,1] -> 0
    if i + j == 0:
        return 1.0
    if i + j == 1:
        return 0.0
    return data[i, j]


@njit(cache=True)
def push_val(x0, x1, w, data, binom):
    """Merge one sample ``(x0, x1)`` of weight ``w`` into ``data[2, k+1]`` in place."""
    if w == 0.0:
        return
    k = data.shape[1] - 1
    data[0, 0] += w
    alpha = w / data[0, 0]
    one_alpha = 1.0 - alpha
    d0 = x0 - data[1, 0]
    d1 = x1 - data[0, 1]
    data[1, 0] += alpha * d0
    data[0, 1] += alpha * d1
    m0 = -alpha * d0  # old mean - new mean
    m1 = -alpha * d1
    p0 = one_alpha * d0  # new point - new mean
    p1 = one_alpha * d1
    # descending order: every right-hand side still holds the old value
    for a0 in range(1, -1, -1):
        a1_min = 1 if a0 == 1 else 2
        for a1 in range(k, a1_min - 1, -1):
            tmp = 0.0
            m0b = 1.0
            for b0 in range(a0 + 1):
                m1b = 1.0
                for b1 in range(a1 + 1):
                    old = _get(data, a0 - b0, a1 - b1)
                    if old != 0.0:
                        tmp += binom[a0, b0] * binom[a1, b1] * m0b * m1b * old
                    m1b *= m1
                m0b *= m0
            data[a0, a1] = one_alpha * tmp + alpha * (p0**a0) * (p1**a1)


@njit(cache=True)
def _reduce_one(x, u, w, has_w, order, binom):
    data = np.zeros((2, order + 1))
    for s in range(u.shape[0]):
        push_val(x[s], u[s], w[s] if has_w else 1.0, data, binom)
    return data


def reduce_vals(x, u, order, w=None):
    """Comoments of scalar ``x[nsamp]`` (or channels ``x[nsamp, nv]``) with ``u[nsamp]``."""
    x = np.asarray(x, dtype=np.float64)
    u = np.ascontiguousarray(u, dtype=np.float64)
    has_w = w is not None
    wv = np.ascontiguousarray(w, dtype=np.float64) if has_w else np.zeros(1)
    if x.ndim == 1:
        return _reduce_one(np.ascontiguousarray(x), u, wv, has_w, order, _BINOM)
    return _reduce_wide(np.ascontiguousarray(x.T), u, wv, has_w, order, _BINOM)


@njit(cache=True, parallel=True)
def _reduce_wide(xt, u, w, has_w, order, binom):
    out = np.zeros((xt.shape[0], 2, order + 1))
    for c in prange(xt.shape[0]):
        out[c] = _reduce_one(xt[c], u, w, has_w, order, binom)
    return out


@njit(cache=True, parallel=True)
def indices_to_freq(indices, ndat):
    nrep = indices.shape[0]
    freq = np.zeros((nrep, ndat), dtype=np.int64)
    for r in prange(nrep):
        for j in range(indices.shape[1]):
            freq[r, indices[r, j]] += 1
    return freq


@njit(cache=True, parallel=True)
def _resample(x, u, w, has_w, freq, order, binom):
    nrep = freq.shape[0]
    out = np.zeros((nrep, 2, order + 1))
    for r in prange(nrep):
        data = out[r]
        for s in range(u.shape[0]):
            f = freq[r, s]
            if f != 0:
                push_val(x[s], u[s], (w[s] if has_w else 1.0) * f, data, binom)
    return out


@njit(cache=True, parallel=True)
def _resample_wide(xt, u, w, has_w, freq, order, binom):
    # cmomy's gufunc broadcasts over (rep, channel): one independent push loop per pair
    nrep, nchan = freq.shape[0], xt.shape[0]
    out = np.zeros((nrep, nchan, 2, order + 1))
    for rc in prange(nrep * nchan):
        r, c = rc // nchan, rc % nchan
        data = out[r, c]
        x = xt[c]
        for s in range(u.shape[0]):
            f = freq[r, s]
            if f != 0:
                push_val(x[s], u[s], (w[s] if has_w else 1.0) * f, data, binom)
    return out


def resample_vals(x, u, freq, order, w=None):
    """``out[rep] = comoments`` of the samples reweighted by ``freq[rep]`` (scalar ``x[nsamp]``), or
    ``out[rep, chan]`` for channels ``x[nsamp, nchan]``."""
    has_w = w is not None
    wv = np.ascontiguousarray(w, dtype=np.float64) if has_w else np.zeros(1)
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 2:
        return _resample_wide(
            np.ascontiguousarray(x.T), np.ascontiguousarray(u, dtype=np.float64), wv, has_w,
            np.ascontiguousarray(freq, dtype=np.int64), order, _BINOM,
        )
    return _resample(
        np.ascontiguousarray(x, dtype=np.float64), np.ascontiguousarray(u, dtype=np.float64), wv, has_w,
        np.ascontiguousarray(freq, dtype=np.int64), order, _BINOM,
    )


def bootstrap_reference_path(x, u, order, nrep, rng):
    """The reference's whole bootstrap for ``{"nrep": nrep, "rng": rng}``: table, histogram, pushes."""
    n = u.shape[0]
    idx = rng.choice(n, size=(nrep, n), replace=True)
    freq = indices_to_freq(idx, n)
    return resample_vals(x, u, freq, order)

"""Oracle for weighted central comoments ``mom=(1, k)`` (TEST INFRASTRUCTURE ONLY).

Layout (``/root/reference/src/thermoextrap/data.py:1233-1236``)::

    data[..., 0, 0] = sum of weights
    data[..., 1, 0] = <x>
    data[..., 0, 1] = <u>
    data[..., i, j] = <(x-<x>)**i (u-<u>)**j>      otherwise, i in {0,1}, j<=k

Three restatements are provided:

``comoments_two_pass``   definitionally exact two-pass sums in ``np.longdouble``
                         (ground truth at 1e-10; what ``tests/test_data.py:7-38``
                         pins via ``np.average(x*u**i)`` after central->raw).
``push_vals``            cmomy 0.24's published algorithm: one-pass weighted
                         push of single samples = the singleton case of the
                         Chan/Pebay merge (call sites ``data.py:1632-1640``).
``merge_blocks``         Chan/Pebay merge of pre-reduced blocks
                         (``resample_and_reduce`` ``data.py:1048-1052``,
                         ``reduce`` ``data.py:996``).
"""

from __future__ import annotations

from math import comb

import numpy as np

LD = np.longdouble


def _prep(x, u, w):
    x = np.asarray(x, dtype=np.float64)
    u = np.asarray(u, dtype=np.float64)
    nsamp = x.shape[0]
    val_shape = x.shape[1:]
    xf = x.reshape(nsamp, -1)
    if u.ndim == 1:
        uf = u.reshape(nsamp, 1)
    else:
        uf = np.broadcast_to(u, x.shape).reshape(nsamp, -1)
    if w is None:
        wf = np.ones((nsamp, 1))
    else:
        w = np.asarray(w, dtype=np.float64)
        wf = w.reshape(nsamp, 1) if w.ndim == 1 else np.broadcast_to(w, x.shape).reshape(nsamp, -1)
    return xf, uf, wf, val_shape


def comoments_two_pass(x, u, order, w=None, dtype=np.float64):
    """Exact (longdouble, two-pass) comoments of ``x[nsamp, *val]`` with ``u``.

    Returns ``out[*val, 2, order+1]`` in the cmomy layout.
    """
    xf, uf, wf, val_shape = _prep(x, u, w)
    xl, ul, wl = xf.astype(LD), uf.astype(LD), wf.astype(LD)
    nchan = xl.shape[1]
    wsum = np.broadcast_to(wl.sum(axis=0), (nchan,)) if wl.shape[1] == 1 else wl.sum(axis=0)
    xave = (wl * xl).sum(axis=0) / wsum
    uave = np.broadcast_to((wl * ul).sum(axis=0) / (wl.sum(axis=0)), (nchan,))
    dx = xl - xave
    du = ul - (uave if ul.shape[1] == nchan else uave[:1])
    out = np.zeros((nchan, 2, order + 1), dtype=LD)
    p = np.ones_like(du)
    for j in range(order + 1):
        out[:, 0, j] = (wl * p).sum(axis=0) / wsum
        out[:, 1, j] = (wl * dx * p).sum(axis=0) / wsum
        p = p * du
    out[:, 0, 0] = wsum
    out[:, 1, 0] = xave
    out[:, 0, 1] = uave
    return out.astype(dtype).reshape(*val_shape, 2, order + 1)


def merge_pair(a, b):
    """Chan/Pebay merge of two comoment arrays ``[..., 2, k+1]`` (cmomy layout).

    m_pq = sum_{i<=p, j<=q} C(p,i) C(q,j) [ fa * A_{p-i,q-j} (-fb dx)^i (-fb du)^j
                                           + fb * B_{p-i,q-j} ( fa dx)^i ( fa du)^j ]
    with ``fa = Wa/W``, ``fb = Wb/W``, ``d* = mean_b - mean_a`` and the
    convention ``M_00 = 1, M_10 = M_01 = 0`` inside the sum.
    """
    a = np.asarray(a)
    b = np.asarray(b)
    wa = a[..., 0, 0]
    wb = b[..., 0, 0]
    w = wa + wb
    with np.errstate(invalid="ignore", divide="ignore"):
        fa = np.where(w != 0, wa / w, 0.0)
        fb = np.where(w != 0, wb / w, 0.0)
    dx = b[..., 1, 0] - a[..., 1, 0]
    du = b[..., 0, 1] - a[..., 0, 1]
    k = a.shape[-1] - 1

    def cm(arr, i, j):
        if i == 0 and j == 0:
            return np.ones_like(w)
        if i + j == 1:
            return np.zeros_like(w)
        return arr[..., i, j]

    out = np.zeros(np.broadcast(a, b).shape, dtype=np.result_type(a, b))
    for p in range(2):
        for q in range(k + 1):
            if p + q < 2:
                continue
            acc = np.zeros_like(w)
            for i in range(p + 1):
                for j in range(q + 1):
                    c = comb(p, i) * comb(q, j)
                    ta = fa * cm(a, p - i, q - j) * (-fb * dx) ** i * (-fb * du) ** j
                    tb = fb * cm(b, p - i, q - j) * (fa * dx) ** i * (fa * du) ** j
                    acc = acc + c * (ta + tb)
            out[..., p, q] = acc
    out[..., 0, 0] = w
    out[..., 1, 0] = a[..., 1, 0] + fb * dx
    out[..., 0, 1] = a[..., 0, 1] + fb * du
    return out


def merge_blocks(data, freq=None, axis=0, dtype=np.float64):
    """Sequential Chan/Pebay merge of blocks along ``axis``.

    ``freq`` (optional, ``[nblock]``) scales each block's weight
    (``push_data_scale`` semantics of ``resample_and_reduce``).
    """
    data = np.moveaxis(np.asarray(data, dtype=LD), axis, 0)
    acc = np.zeros(data.shape[1:], dtype=LD)
    for ib in range(data.shape[0]):
        blk = data[ib].copy()
        if freq is not None:
            if freq[ib] == 0:
                continue
            blk[..., 0, 0] = blk[..., 0, 0] * freq[ib]
        acc = merge_pair(acc, blk)
    return acc.astype(dtype)


def resample_blocks(data, freq, axis=0, dtype=np.float64):
    """``out[rep] = merge_blocks(data, freq[rep])``  (K3, ``data.py:1048-1052``)."""
    return np.stack([merge_blocks(data, f, axis=axis, dtype=dtype) for f in np.asarray(freq)])


def push_vals(x, u, order, w=None, dtype=np.float64):
    """One-pass per-sample push (cmomy's algorithm): singleton-block merges in fp64."""
    xf, uf, wf, val_shape = _prep(x, u, w)
    nsamp, nchan = xf.shape
    acc = np.zeros((nchan, 2, order + 1), dtype=dtype)
    blk = np.zeros_like(acc)
    for s in range(nsamp):
        if wf.shape[1] == 1 and wf[s, 0] == 0:
            continue
        blk[...] = 0.0
        blk[:, 0, 0] = wf[s]
        blk[:, 1, 0] = xf[s]
        blk[:, 0, 1] = uf[s]
        acc = merge_pair(acc, blk)
    return acc.reshape(*val_shape, 2, order + 1)


def resample_vals_gather(x, u, order, indices, w=None, dtype=np.float64):
    """Bootstrap oracle, gather form (``data.py:404-449`` semantics):
    ``out[rep] = comoments(x[idx[rep]], u[idx[rep]])``."""
    x = np.asarray(x)
    u = np.asarray(u)
    out = []
    for idx in np.asarray(indices):
        wi = None if w is None else np.asarray(w)[idx]
        out.append(comoments_two_pass(x[idx], u[idx], order, w=wi, dtype=dtype))
    return np.stack(out)


def resample_vals_freq(x, u, order, freq, w=None, dtype=np.float64):
    """Bootstrap oracle, freq form (``data.py:1803-1810``): weights ``w_s * freq[rep, s]``."""
    x = np.asarray(x)
    nsamp = x.shape[0]
    w0 = np.ones(nsamp) if w is None else np.asarray(w, dtype=np.float64)
    out = []
    for f in np.asarray(freq):
        out.append(comoments_two_pass(x, u, order, w=w0 * f, dtype=dtype))
    return np.stack(out)


def indices_to_freq(indices, ndat=None):
    """``freq[r, i] = #{j : indices[r, j] == i}`` (int64, exact)."""
    indices = np.asarray(indices)
    ndat = indices.shape[1] if ndat is None else ndat
    freq = np.zeros((indices.shape[0], ndat), dtype=np.int64)
    for r in range(indices.shape[0]):
        freq[r] = np.bincount(indices[r], minlength=ndat)
    return freq


# ---------------------------------------------------------------- conversions
def central_to_raw(data, dtype=np.float64):
    """cmomy ``.rmom()``: raw ``<x^i u^j>`` with ``[0,0]=1`` (notebook cells 16-22)."""
    d = np.asarray(data, dtype=LD)
    k = d.shape[-1] - 1
    xa, ua = d[..., 1, 0], d[..., 0, 1]

    def cm(i, j):
        if i == 0 and j == 0:
            return np.ones_like(xa)
        if i + j == 1:
            return np.zeros_like(xa)
        return d[..., i, j]

    out = np.zeros_like(d)
    for p in range(2):
        for q in range(k + 1):
            acc = np.zeros_like(xa)
            for i in range(p + 1):
                for j in range(q + 1):
                    acc = acc + comb(p, i) * comb(q, j) * cm(i, j) * xa ** (p - i) * ua ** (q - j)
            out[..., p, q] = acc
    return out.astype(dtype)


def raw_to_central(raw, dtype=np.float64):
    """``convert.moments_type(raw, to="central")`` (``data.py:1108-1115``).

    ``raw[...,0,0]`` is the weight and is preserved.
    """
    r = np.asarray(raw, dtype=LD)
    k = r.shape[-1] - 1
    xa, ua = r[..., 1, 0], r[..., 0, 1]

    def rm(i, j):
        if i == 0 and j == 0:
            return np.ones_like(xa)
        return r[..., i, j]

    out = np.zeros_like(r)
    for p in range(2):
        for q in range(k + 1):
            acc = np.zeros_like(xa)
            for i in range(p + 1):
                for j in range(q + 1):
                    acc = acc + comb(p, i) * comb(q, j) * rm(i, j) * (-xa) ** (p - i) * (-ua) ** (q - j)
            out[..., p, q] = acc
    out[..., 0, 0] = r[..., 0, 0]
    out[..., 1, 0] = xa
    out[..., 0, 1] = ua
    return out.astype(dtype)


def raw_to_central_1d(raw, dtype=np.float64):
    """1-D version (``x_is_u`` branch, ``data.py:1108-1111``): ``raw[..., j] = <u^j>``."""
    r = np.asarray(raw, dtype=LD)
    k = r.shape[-1] - 1
    ua = r[..., 1]
    out = np.zeros_like(r)
    for q in range(k + 1):
        acc = np.zeros_like(ua)
        for j in range(q + 1):
            rj = np.ones_like(ua) if j == 0 else r[..., j]
            acc = acc + comb(q, j) * rj * (-ua) ** (q - j)
        out[..., q] = acc
    out[..., 0] = r[..., 0]
    out[..., 1] = ua
    return out.astype(dtype)


def moments_to_comoments(m, order):
    """``moments_to_comoments(mom=(1, order))`` (``data.py:1191,1266-1268``):
    ``out[0, j] = m[j]``, ``out[1, j] = m[j+1]``; in the central layout
    ``out[1, 0]`` is ``<u>`` (``x == u``)."""
    m = np.asarray(m)
    if order < 0:
        order = m.shape[-1] - 2
    out = np.zeros(m.shape[:-1] + (2, order + 1), dtype=m.dtype)
    out[..., 0, :] = m[..., : order + 1]
    out[..., 1, :] = m[..., 1 : order + 2]
    return out


def tolerance_scale(x, u, order):
    """Per-entry scale ``sigma_x^i sigma_u^j`` for the 1e-10 comparison."""
    xf, uf, _, val_shape = _prep(x, u, None)
    sx = xf.std(axis=0)
    su = np.broadcast_to(uf.std(axis=0), sx.shape)
    out = np.zeros((xf.shape[1], 2, order + 1))
    for i in range(2):
        for j in range(order + 1):
            out[:, i, j] = sx**i * su**j
    return out.reshape(*val_shape, 2, order + 1)

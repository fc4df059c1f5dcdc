"""TEST-ONLY stand-in for ``thermoextrap_b200.kernels`` backed by the CPU oracle.

Lets the ``-m "not gpu"`` suite exercise the HOST logic (dims, samplers, data classes,
models, factories) in a container without a GPU.  The product never imports this module;
on a GPU box the same tests run against the real CUDA kernels.
"""

from __future__ import annotations

import numpy as np
import torch

from oracle import comoments as oc

CONVERT_CENTRAL_TO_RAW, CONVERT_RAW_TO_CENTRAL, CONVERT_CENTRAL_TO_CMOM = 0, 1, 2
launch_count = 0


def device():
    return torch.device("cpu")


def pending_upload(t):
    """No copy stream on the CPU: nothing is ever in flight."""
    return None


def wait_upload(t):
    return None


def _np(t):
    if t is None:
        return None
    return t.numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def reduce_vals(x, u, order, w=None, x_is_u=False):
    u = _np(u)
    x = u if x_is_u else _np(x)
    if order < 1 or order > 12:
        raise ValueError("order")
    out = oc.comoments_two_pass(x, u, order, w=_np(w))
    return _t(out)


def indices_to_freq(indices, ndat=None, dtype=torch.int64):
    f = oc.indices_to_freq(_np(indices), ndat)
    info = torch.iinfo(dtype)
    if f.max() > info.max:
        raise OverflowError
    return _t(f).to(dtype)


def resample_vals(x, u, freq, order, w=None, x_is_u=False):
    u = _np(u)
    x = u if x_is_u else _np(x)
    return _t(oc.resample_vals_freq(x, u, order, _np(freq).astype(np.int64), w=_np(w)))


def resample_indices(x, u, indices, order, w=None, x_is_u=False):
    u = _np(u)
    x = u if x_is_u else _np(x)
    return _t(oc.resample_vals_gather(x, u, order, _np(indices), w=_np(w)))


def reduce_data(data):
    return _t(oc.merge_blocks(_np(data)))


def resample_data(data, freq):
    return _t(oc.resample_blocks(_np(data), _np(freq)))


def convert(data, kind):
    d = _np(data)
    if kind == CONVERT_CENTRAL_TO_RAW:
        return _t(oc.central_to_raw(d))
    if kind == CONVERT_RAW_TO_CENTRAL:
        return _t(oc.raw_to_central(d))
    out = d.copy()
    out[..., 0, 0] = 1.0
    out[..., 0, 1] = 0.0
    out[..., 1, 0] = 0.0
    return _t(out)


PREDICT_SUM, PREDICT_CUMSUM, PREDICT_TERMS = 0, 1, 2


def eval_exprs(vars_, term_ptr, coef, fac_ptr, fac_var, fac_pow):
    v = _np(vars_).astype(np.float64)
    nexpr = len(term_ptr) - 1
    out = np.zeros((nexpr, v.shape[1]))
    for i in range(nexpr):
        for t in range(term_ptr[i], term_ptr[i + 1]):
            term = np.full(v.shape[1], float(coef[t]))
            for f in range(fac_ptr[t], fac_ptr[t + 1]):
                p = int(fac_pow[f])
                x = v[int(fac_var[f])]
                term = term * (np.log(x) if p == 127 else x ** float(p) if p < 0 else x**p)
            out[i] += term
    return _t(out)


def poly_predict(coefs, dalpha, mode=0):
    c = _np(coefs)
    d = np.atleast_1d(np.asarray(dalpha, dtype=np.float64))
    p = np.arange(c.shape[0])
    terms = d[:, None, None] ** p[None, :, None] * c[None]
    if mode == 0:
        return _t(terms.sum(axis=1))
    if mode == 1:
        return _t(np.cumsum(terms, axis=1))
    return _t(terms)


class SamplerPlan:
    """Stand-in for ``kernels.SamplerPlan``: live numpy draws the table (it IS the reference sampler)."""

    def __init__(self, rng, nrep, nsamp, ndat=None, group=None):
        self.nrep, self.nsamp = int(nrep), int(nsamp)
        self.ndat = self.nsamp if ndat is None else int(ndat)
        self._idx = rng.choice(self.ndat, size=(self.nrep, self.nsamp), replace=True)


def _win(plan, reps):
    return slice(None) if reps is None else slice(int(reps[0]), int(reps[1]))


def sampler_indices(src, nrep=None, nsamp=None, ndat=None, reps=None):
    plan = src if isinstance(src, SamplerPlan) else SamplerPlan(src, nrep, nsamp, ndat)
    return _t(plan._idx[_win(plan, reps)])


def sampler_freq(src, nrep=None, nsamp=None, ndat=None, dtype=torch.uint8, reps=None):
    plan = src if isinstance(src, SamplerPlan) else SamplerPlan(src, nrep, nsamp, ndat)
    return indices_to_freq(plan._idx[_win(plan, reps)], plan.ndat, dtype=dtype)


def sampler_resample(src, x, u, nrep=None, order=1, w=None, x_is_u=False, nsamp=None, sync_rng=True, reps=None,
                     buckets=True):
    n = _np(u).shape[0]
    plan = src if isinstance(src, SamplerPlan) else SamplerPlan(src, nrep, n if nsamp is None else nsamp, n)
    return resample_indices(x, u, plan._idx[_win(plan, reps)], order, w=w, x_is_u=x_is_u)


def sampler_resample_wide(src, x, u, nrep=None, order=1, w=None, x_is_u=False, nsamp=None, table_bytes=None,
                          timing=None, reps=None):
    n = _np(u).shape[0]
    plan = src if isinstance(src, SamplerPlan) else SamplerPlan(src, nrep, n if nsamp is None else nsamp, n)
    uu = _np(u)
    xx = uu if x_is_u else _np(x)
    return _t(oc.resample_vals_gather(xx, uu, order, plan._idx[_win(plan, reps)], w=_np(w)))


def reduce_sums(x, u, order, w=None, pivot=None):
    u = _np(u).astype(np.float64)
    ww = np.ones(u.shape[0]) if w is None else _np(w)
    if x is None:  # x_is_u: powers of u (per channel when u is 2-D)
        u2 = u.reshape(u.shape[0], -1)
        if pivot is None:
            pivot = _t(u2[:256].mean(0))
        du = u2 - _np(pivot)[None, :]
        pw = du[:, :, None] ** np.arange(order + 1)[None, None, :] * ww[:, None, None]
        return pivot, _t(pw.sum(0)), None, u.shape[1:]
    x = _np(x).astype(np.float64)
    val_shape = x.shape[1:]
    x2 = x.reshape(x.shape[0], -1)
    if pivot is None:
        pivot = _t(np.concatenate([x2[:256].mean(0), [u[:256].mean()]]))
    pv = _np(pivot)
    dx, du = x2 - pv[:-1], u - pv[-1]
    pw = du[:, None] ** np.arange(order + 1)[None, :] * ww[:, None]
    return pivot, _t(pw.sum(0)[None, :]), _t(dx.T @ pw), val_shape


def finalize_sums(pivot, S0, S1, order, x_is_u=False, val_shape=()):
    pv, s0, s1 = _np(pivot), _np(S0), (None if S1 is None else _np(S1))
    lead = s0.shape[:-2]
    s0 = s0.reshape(-1, *s0.shape[-2:])
    outs = []
    for r in range(s0.shape[0]):
        if x_is_u:
            nu = s0.shape[1]
            m = s0[r] / s0[r, :, :1]                     # raw moments about the pivot, orders 0..order+1
            raw = np.zeros((nu, 2, order + 2))
            raw[:, 0, :] = m
            raw[:, 0, 0] = s0[r, :, 0]
            cen = oc.raw_to_central(raw)[:, 0, :]        # central moments of u (about its mean), [nu, order+2]
            out = np.zeros((nu, 2, order + 1))
            out[:, 0, :] = cen[:, : order + 1]
            out[:, 1, :] = cen[:, 1:]
            mean = cen[:, 1] + pv[:nu]
            out[:, 0, 1] = mean
            out[:, 1, 0] = mean
        else:
            s1r = s1.reshape(-1, *s1.shape[-2:])[r]
            nchan = s1r.shape[0]
            raw = np.zeros((nchan, 2, order + 1))
            raw[:, 0, :] = s0[r, 0] / s0[r, 0, 0]
            raw[:, 1, :] = s1r / s0[r, 0, 0]
            raw[:, 0, 0] = s0[r, 0, 0]
            out = oc.raw_to_central(raw)
            out[:, 1, 0] += pv[:-1]
            out[:, 0, 1] += pv[-1]
        outs.append(out)
    res = np.stack(outs).reshape(*lead, *val_shape, 2, order + 1)
    return _t(res)


FREQ_BUCKET = 65520


def sampler_resample_sums_rows(plan, x_local, u_local, order, row_lo, pivot, w=None, table_bytes=None,
                               next_row_lo=None):
    x, u = _np(x_local).astype(np.float64), _np(u_local).astype(np.float64)
    n_local = u.shape[0]
    val_shape = x.shape[1:]
    x2 = x.reshape(n_local, -1)
    idx = plan._idx
    pv = _np(pivot)
    dx, du = x2 - pv[:-1], u - pv[-1]
    ww = np.ones_like(u) if w is None else _np(w)
    pw = du[:, None] ** np.arange(order + 1)[None, :] * ww[:, None]
    nrep = idx.shape[0]
    S0 = np.zeros((nrep, 1, order + 1))
    S1 = np.zeros((nrep, x2.shape[1], order + 1))
    rowsum = np.zeros(nrep, dtype=np.int64)
    for r in range(nrep):
        sel = idx[r][(idx[r] >= row_lo) & (idx[r] < row_lo + n_local)] - row_lo
        f = np.bincount(sel, minlength=n_local).astype(np.float64)
        S0[r, 0] = f @ pw
        S1[r] = (dx * f[:, None]).T @ pw
        # the real kernel counts table columns from the bucket boundary below row_lo up to the next rank's table
        col_lo = row_lo // FREQ_BUCKET * FREQ_BUCKET
        nxt = plan.ndat if next_row_lo is None else next_row_lo // FREQ_BUCKET * FREQ_BUCKET
        rowsum[r] = int(((idx[r] >= col_lo) & (idx[r] < min(nxt, row_lo + n_local))).sum())
    return _t(S0), _t(S1), val_shape, _t(rowsum)


def free_table_bytes() -> int:
    return 1 << 62


def perturb_predict(x, u, dalpha):
    from oracle import perturb as op

    return _t(op.perturb_predict(_np(x), _np(u), dalpha))


# ------------------------------------------------------------------ K7 device-resident stand-ins
BETA_FAMILIES = {("x_ave", True): 0, ("u_ave", True): 1, ("x_ave", False): 2, ("u_ave", False): 3}
BETA_MAX_ORDER = 8   # what the oracle's golden expressions cover
PREDICT_DERIVS = 3


def beta_predict(comom, order, family, dalpha=None, minus_log=False, mode=0, norm=False):
    from oracle import derivs as od

    c = _np(comom)
    if minus_log and order > 5:   # beyond the golden -log expressions: restate the recurrence (host-logic tests only)
        import math

        d = od.beta_derivs(c, order, family).astype(np.longdouble)
        g = np.empty_like(d)
        g[0] = np.log(d[0])
        for n in range(1, order + 1):
            s = d[n].copy()
            for k in range(1, n):
                s -= math.comb(n - 1, k) * d[k] * g[n - k]
            g[n] = s / d[0]
        der = (-g).astype(np.float64)
        fact = np.array([math.factorial(i) for i in range(order + 1)], dtype=np.float64)[:, None]
        if mode == PREDICT_DERIVS:
            return _t(der / fact if norm else der)
        return poly_predict(_t(der / fact), dalpha, mode)
    if mode == PREDICT_DERIVS:
        return _t(od.beta_derivs(c, order, family, minus_log=minus_log, norm=norm))
    return _t(od.beta_predict(c, order, family, dalpha, minus_log=minus_log, mode=mode))


LNPI_FAMILIES = {True: 4, False: 5}


def lnpi_predict(comom, order, central, lnpi0, mudotn, dalpha=None, mode=0, norm=False):
    from oracle import derivs as od

    c = _np(comom)
    if mode == PREDICT_DERIVS:
        return _t(od.lnpi_derivs(c, order, central, _np(lnpi0), _np(mudotn), norm=norm))
    return _t(od.lnpi_predict(c, order, central, _np(lnpi0), _np(mudotn), dalpha, mode=mode))


def combine(W, D, out=None):
    res = np.asarray(_np(W), dtype=np.float64) @ _np(D)
    if out is not None:
        out += _t(res)
        return out
    return _t(res)


def mean_var(arr, ddof=0):
    a = _np(arr)
    return _t(a.mean(axis=1)), _t(a.var(axis=1, ddof=ddof))

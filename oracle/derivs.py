"""TEST INFRASTRUCTURE (oracle): CPU restatement of K7 for the four numeric beta families.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.

What the reference does (/root/reference/src/thermoextrap): ``Derivatives.derivs`` evaluates the lambdified
sympy expressions ``funcs[i](*data.derivs_args)`` (models.py:371), applies the ``-log`` transform (models.py:261-281,
373-374), divides by ``i!`` (:376-377); ``ExtrapModel.predict`` sums the Taylor series (:535-565).

The expressions evaluated here are the strings in ``tests/golden/exprs.json`` -- produced by the REFERENCE's own
sympy classes (beta.py:32-266) by ``tests/golden/make_golden.py`` -- lambdified with numpy; nothing is taken from
the product's generator.  Orders above the golden file's (8; 5 for ``-log``) raise.
"""

from __future__ import annotations

import json
import math
from functools import lru_cache
from pathlib import Path

import numpy as np

GOLDEN = Path(__file__).resolve().parent.parent / "tests" / "golden" / "exprs.json"
KEYS = {
    0: "x_ave|central=True,post_func=None,xalpha=False",
    1: "u_ave|central=True,post_func=None",
    2: "x_ave|central=False,post_func=None,xalpha=False",
    3: "u_ave|central=False,post_func=None",
}


@lru_cache(None)
def _funcs(key: str):
    import sympy as sp

    rec = json.loads(GOLDEN.read_text())[key]
    names = rec.get("args") or ["X"]
    scalars = {"x1", "lnPi0", "mudotN"}
    loc = {n: (sp.Symbol(n) if n in scalars or (n == "u" and "du" in names) else sp.IndexedBase(n)) for n in names}
    args = [loc[n] for n in names]
    return [sp.lambdify(args, sp.sympify(e, locals=loc), "numpy") for e in rec["exprs"]]


def central_rows(comom: np.ndarray, x_is_u: bool):
    """``(ubar, xbar, du[n], dxdu[n])`` of records ``[nelem, 2, K+1]``; ``x_is_u``: ``<du^n> = <dx du^(n-1)>``."""
    c = np.asarray(comom, dtype=np.longdouble)
    k1 = c.shape[-1]
    if x_is_u:
        du = np.concatenate([np.ones_like(c[:, 1, :1]), c[:, 1, :]], axis=1)   # n = 0 .. K+1, [1] = <u>
        ub = du[:, 1].copy()
        du[:, 1] = 0.0
        return ub, ub, du, None
    du = c[:, 0, :].copy()
    ub = du[:, 1].copy()
    du[:, 0], du[:, 1] = 1.0, 0.0
    dxdu = c[:, 1, :].copy()
    xb = dxdu[:, 0].copy()
    dxdu[:, 0] = 0.0
    assert du.shape[1] == k1
    return ub, xb, du, dxdu


def beta_derivs(comom, order: int, family: int, minus_log: bool = False, norm: bool = False) -> np.ndarray:
    """``[order+1, nelem]`` derivatives of family 0..3 from central comoment records ``[..., 2, K+1]``."""
    c = np.asarray(comom, dtype=np.float64)
    c = c.reshape(-1, 2, c.shape[-1])
    funcs = _funcs(KEYS[family])
    if order + 1 > len(funcs):
        raise NotImplementedError(f"oracle holds orders <= {len(funcs) - 1}")
    ub, xb, du, dxdu = central_rows(c, x_is_u=family in (1, 3))
    nel = c.shape[0]
    if family < 2:
        duT = [du[:, n] for n in range(du.shape[1])]
        if family == 0:
            dxT = [dxdu[:, n] for n in range(dxdu.shape[1])]
            args = (xb, duT, dxT)
        else:
            args = (ub, duT)
    else:
        nmax = du.shape[1]
        u_raw, xu_raw = [], []
        for n in range(nmax):
            u_raw.append(sum(math.comb(n, k) * du[:, k] * ub ** (n - k) for k in range(n + 1)))
            if family == 2:
                xu_raw.append(sum(math.comb(n, k) * ub ** (n - k) * (xb * du[:, k] + dxdu[:, k]) for k in range(n + 1)))
        args = (u_raw, xu_raw) if family == 2 else (u_raw,)
    out = np.empty((order + 1, nel), dtype=np.longdouble)
    for i in range(order + 1):
        out[i] = np.broadcast_to(funcs[i](*args), (nel,))
    if minus_log:
        ml = _funcs("SymMinusLog")
        if order + 1 > len(ml):
            raise NotImplementedError(f"oracle holds -log orders <= {len(ml) - 1}")
        X = [out[i] for i in range(order + 1)]
        out = np.stack([np.broadcast_to(ml[i](X), (nel,)) for i in range(order + 1)])
    if norm:
        out = out / np.array([math.factorial(i) for i in range(order + 1)], dtype=np.longdouble)[:, None]
    return out.astype(np.float64)


def lnpi_derivs(comom, order: int, central: bool, lnpi0, mudotn, norm: bool = False) -> np.ndarray:
    """``[order+1, nelem]`` lnPi derivatives (golden key ``lnPi|central=...``, lnpi.py:42-116) from x_is_u records
    ``[..., 2, K+1]``; ``lnpi0`` / ``mudotn`` (``ninner`` values) broadcast over the leading elements."""
    c = np.asarray(comom, dtype=np.float64)
    c = c.reshape(-1, 2, c.shape[-1])
    funcs = _funcs(f"lnPi|central={bool(central)}")
    if order + 1 > len(funcs):
        raise NotImplementedError(f"oracle holds orders <= {len(funcs) - 1}")
    ub, _, du, _ = central_rows(c, x_is_u=True)
    nel = c.shape[0]
    l0 = np.resize(np.asarray(lnpi0, dtype=np.longdouble).reshape(-1), nel)
    mn = np.resize(np.asarray(mudotn, dtype=np.longdouble).reshape(-1), nel)
    if central:
        args = (ub, [du[:, n] for n in range(du.shape[1])], l0, mn)
    else:
        u_raw = [sum(math.comb(n, k) * du[:, k] * ub ** (n - k) for k in range(n + 1)) for n in range(du.shape[1])]
        args = (u_raw, l0, mn)
    out = np.stack([np.broadcast_to(funcs[i](*args), (nel,)) for i in range(order + 1)]).astype(np.longdouble)
    if norm:
        out = out / np.array([math.factorial(i) for i in range(order + 1)], dtype=np.longdouble)[:, None]
    return out.astype(np.float64)


def lnpi_predict(comom, order, central, lnpi0, mudotn, dalpha, mode=0) -> np.ndarray:
    coefs = lnpi_derivs(comom, order, central, lnpi0, mudotn, norm=True).astype(np.longdouble)
    d = np.atleast_1d(np.asarray(dalpha, dtype=np.longdouble))
    terms = d[:, None, None] ** np.arange(order + 1)[None, :, None] * coefs[None]
    if mode == 0:
        return terms.sum(axis=1).astype(np.float64)
    return (np.cumsum(terms, axis=1) if mode == 1 else terms).astype(np.float64)


def beta_predict(comom, order, family, dalpha, minus_log=False, mode=0) -> np.ndarray:
    """Taylor prediction (mode 0 sum ``[nalpha, nelem]``, 1 cumsum / 2 terms ``[nalpha, order+1, nelem]``)."""
    coefs = beta_derivs(comom, order, family, minus_log=minus_log, norm=True).astype(np.longdouble)
    d = np.atleast_1d(np.asarray(dalpha, dtype=np.longdouble))
    terms = d[:, None, None] ** np.arange(order + 1)[None, :, None] * coefs[None]
    if mode == 0:
        return terms.sum(axis=1).astype(np.float64)
    return (np.cumsum(terms, axis=1) if mode == 1 else terms).astype(np.float64)


def scale_of_terms(comom, order, family, dalpha, minus_log=False) -> np.ndarray:
    """``sum_p |term_p|`` per (alpha, element): the SURVEY 8c tolerance scale for derivative / predict sums."""
    return np.abs(beta_predict(comom, order, family, dalpha, minus_log=minus_log, mode=2)).sum(axis=1)

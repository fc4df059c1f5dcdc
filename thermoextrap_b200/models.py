"""Extrapolation / interpolation models (mirror of ``thermoextrap.models``).

Symbolic layer (``SymFuncBase``, ``SymDerivBase``, ``SymSubs``, ``Lambdify``, ``SymMinusLog``:
reference models.py:60-287) builds the same sympy expressions as the reference; the numeric
layer compiles them into monomial tables evaluated on the GPU (``ExprProgram`` ->
``txb_eval_exprs_f64``) instead of lambdified numpy, and sums Taylor / interpolation series
with ``txb_poly_predict_f64``.
"""

from __future__ import annotations

import math
from collections.abc import Mapping, Sequence
from functools import lru_cache
from typing import Any

import attrs
import numpy as np
import sympy as sp
from attrs import converters as attc
from attrs import field
from attrs import validators as attv

from . import backend as cm
from .core import cached
from .core.attrs_utils import MyAttrsMixin, convert_mapping_or_none_to_dict
from .core.sputils import get_default_indexed, get_default_symbol
from .core.xr import DataArray, concat, dot, is_dataarray
from .core.xrutils import xrwrap_alpha
from .data import AbstractData

__all__ = [
    "Derivatives",
    "ExtrapModel",
    "ExtrapWeightedModel",
    "InterpModel",
    "InterpModelPiecewise",
    "MBARModel",
    "PerturbModel",
    "StateCollection",
]


################################################################################
# symbolic layer
################################################################################
class SymFuncBase(sp.Function):
    """Base class for user-defined sympy functions with hand-written derivative rules.

    Subclasses implement ``deriv_args`` (symbols the lambdified function takes), ``fdiff``
    (the recurrence) and ``eval`` (``beta='None'`` turns the function into an indexed symbol).
    """

    @classmethod
    def deriv_args(cls):
        msg = "must specify in subclass"
        raise NotImplementedError(msg)

    def fdiff(self, argindex=1):
        msg = "must specify in subclass"
        raise NotImplementedError(msg)

    @classmethod
    def eval(cls, beta):
        msg = "must specify in subclass"
        raise NotImplementedError(msg)


class SymDerivBase:
    """``self[n]`` = n-th derivative of ``func`` w.r.t. ``self.beta`` (recursive, cached)."""

    def __init__(self, func, args=None, expand=True, post_func=None) -> None:
        if args is None:
            args = func.deriv_args()
        self._func_orig = func
        self._post_func = post_func
        if post_func is not None:
            if isinstance(post_func, str):
                if post_func == "minus_log":
                    post_func = lambda f: -sp.log(f)  # noqa: E731
                elif post_func.startswith("pow_"):
                    i = int(post_func.split("_")[-1])
                    post_func = lambda f: pow(f, i)  # noqa: E731
                else:
                    msg = "post_func must be callable or in {minus_log, pow_1, pow_2, ...}"
                    raise ValueError(msg)
            func = post_func(func)
        self.func = func
        self.args = args
        self.expand = expand
        self._cache: dict = {}

    @cached.meth
    def __getitem__(self, order):
        if order == 0:
            return self.func
        out = self[order - 1].diff(self.beta, 1)
        return out.expand() if self.expand else out


@attrs.define
class SymSubs:
    """Apply substitutions to the expressions of a :class:`SymDerivBase`."""

    funcs: Any = field()
    subs: Sequence | None = field(default=None)
    subs_final: Sequence | None = field(default=None)
    subs_all: Mapping | None = field(default=None)
    recursive: bool = field(default=True)
    simplify: bool = field(default=False)
    expand: bool = field(default=True)
    _cache: dict = field(init=False, repr=False, factory=dict)

    @cached.meth
    def __getitem__(self, order):
        func = self.funcs[order]
        if self.subs is not None:
            if self.recursive:
                for o in range(order, -1, -1):
                    func = func.subs(self.subs[o])
            else:
                func = func.subs(self.subs[order])
        if self.subs_final is not None:
            func = func.subs(self.subs_final[order])
        if self.subs_all is not None:
            func = func.subs(self.subs_all)
        if self.simplify:
            func = func.simplify()
        if self.expand:
            func = func.expand()
        return func


@attrs.define
class Lambdify:
    """``self[n]`` = ``sympy.lambdify(args, exprs[n])`` (host callable; kept for API parity --
    the models themselves evaluate through :class:`ExprProgram` on the device)."""

    exprs: Any = field()
    args: Sequence | None = field(default=None)
    lambdify_kws: Mapping | None = field(kw_only=True, default=None, converter=attc.default_if_none(factory=dict))
    _cache: dict = field(init=False, repr=False, factory=dict)

    @cached.meth
    def __getitem__(self, order):
        return sp.lambdify(self.args, self.exprs[order], **self.lambdify_kws)

    @classmethod
    def from_u_xu(cls, exprs, **lambdify_kws):
        u, xu = get_default_indexed("u", "xu")
        return cls(exprs=exprs, args=(u, xu), lambdify_kws=lambdify_kws)

    @classmethod
    def from_du_dxdu(cls, exprs, xalpha=False, **lambdify_kws):
        x1 = get_default_indexed("x1") if xalpha else get_default_symbol("x1")
        du, dxdu = get_default_indexed("du", "dxdu")
        return cls(exprs=exprs, args=(x1, du, dxdu), lambdify_kws=lambdify_kws)


class SymMinusLog:
    """Derivatives of ``Y = -log(X)`` in terms of the derivatives ``X[i]`` (Faa di Bruno)."""

    X, dX = get_default_indexed("X", "dX")

    def __init__(self) -> None:
        self._cache: dict = {}

    @cached.meth
    def __getitem__(self, order):
        if order == 0:
            return -sp.log(self.X[0])
        expr = 0
        for k in range(1, order + 1):
            expr += sp.factorial(k - 1) * (-1 / self.X[0]) ** k * sp.bell(order, k, self.dX)
        subs = {self.dX[j]: self.X[j + 1] for j in range(order + 1)}
        return expr.subs(subs).expand().simplify()


@lru_cache(5)
def factory_minus_log():
    s = SymMinusLog()
    return Lambdify(s, (s.X,))


################################################################################
# numeric layer: expressions -> monomial tables -> GPU
################################################################################
LOG_POW = 127  # factor power code for log(var) (include/txb.h)


class ExprProgram:
    """A list of sympy expressions compiled for ``txb_eval_exprs_f64``.

    Every expression must expand to ``sum_t coef_t * prod_f var_f**p_f`` with integer powers
    (negative allowed) where ``var`` is a Symbol or an Indexed entry of one of ``args``; a
    factor may also be ``log(var)``.  Terms keep sympy's ``as_ordered_terms`` order, the
    order ``lambdify`` would evaluate them in.
    """

    def __init__(self, exprs: Sequence, args: Sequence) -> None:
        self.args = list(args)
        self.vars: list = []  # sympy atoms, position = variable index
        index: dict = {}
        self.term_ptr = [0]
        self.coef: list[float] = []
        self.fac_ptr = [0]
        self.fac_var: list[int] = []
        self.fac_pow: list[int] = []

        def var_id(atom):
            if atom not in index:
                if not self._is_var(atom):
                    msg = f"cannot compile {atom!r}: not a Symbol or Indexed of {self.args}"
                    raise NotImplementedError(msg)
                index[atom] = len(self.vars)
                self.vars.append(atom)
            return index[atom]

        for expr in exprs:
            expr = sp.expand(sp.sympify(expr))
            for term in expr.as_ordered_terms():
                coef, factors = term.as_coeff_mul()
                if not coef.is_number:
                    msg = f"non-numeric coefficient in {term}"
                    raise NotImplementedError(msg)
                nf = 0
                for f in factors:
                    if f.is_number:  # e.g. Float factors, which as_coeff_mul leaves in place
                        coef = coef * f
                        continue
                    base, exp = f.as_base_exp()
                    if isinstance(base, sp.log):
                        if exp != 1:
                            msg = f"power of a logarithm in {term}"
                            raise NotImplementedError(msg)
                        self.fac_var.append(var_id(base.args[0]))
                        self.fac_pow.append(LOG_POW)
                    else:
                        if not (exp.is_Integer and abs(int(exp)) < LOG_POW):
                            msg = f"non-integer power {exp} in {term}"
                            raise NotImplementedError(msg)
                        self.fac_var.append(var_id(base))
                        self.fac_pow.append(int(exp))
                    nf += 1
                self.coef.append(float(coef))
                self.fac_ptr.append(self.fac_ptr[-1] + nf)
            self.term_ptr.append(len(self.coef))
        self.nexpr = len(self.term_ptr) - 1

    def _is_var(self, atom) -> bool:
        if isinstance(atom, sp.Indexed):
            return atom.base in self.args
        return isinstance(atom, sp.Symbol) and atom in self.args

    def _value_of(self, atom, values):
        if isinstance(atom, sp.Indexed):
            obj = values[self.args.index(atom.base)]
            idx = tuple(int(i) for i in atom.indices)
            return obj[idx[0]] if len(idx) == 1 else obj[idx]
        return values[self.args.index(atom)]

    def __call__(self, *values) -> list[DataArray]:
        """Evaluate all expressions; returns one DataArray per expression."""
        if len(values) != len(self.args):
            msg = f"expected {len(self.args)} arguments, got {len(values)}"
            raise ValueError(msg)
        vals = [self._value_of(a, values) for a in self.vars]
        dims: list = []
        sizes: dict = {}
        coords: dict = {}
        for v in vals:
            if is_dataarray(v):
                for d in v.dims:
                    if d not in dims:
                        dims.append(d)
                        sizes[d] = v.sizes[d]
                    elif sizes[d] != v.sizes[d]:
                        msg = f"size mismatch along {d!r}"
                        raise ValueError(msg)
                for k, c in v.coords.items():
                    if k in v.dims and k not in coords:
                        coords[k] = np.asarray(c.values)
        shape = tuple(sizes[d] for d in dims)
        nelem = int(np.prod(shape, dtype=np.int64)) if shape else 1
        mat = np.empty((max(len(vals), 1), nelem))
        for i, v in enumerate(vals):
            if is_dataarray(v):
                order = [d for d in dims if d in v.dims]
                a = np.asarray(v.transpose(*order).values, dtype=np.float64)
                a = a.reshape([sizes[d] if d in v.dims else 1 for d in dims])
                mat[i] = np.broadcast_to(a, shape).reshape(-1)
            else:
                mat[i] = float(v)
        out = cm._host(
            cm.K.eval_exprs(mat, self.term_ptr, self.coef, self.fac_ptr, self.fac_var, self.fac_pow)
        )
        return [DataArray(out[i].reshape(shape), dims=dims, coords=coords or None) for i in range(self.nexpr)]


@lru_cache(10)
def _minus_log_program(order: int) -> ExprProgram:
    s = SymMinusLog()
    return ExprProgram([s[i] for i in range(order + 1)], (s.X,))


@attrs.define
class Derivatives(MyAttrsMixin):
    """Derivative functions up to some order.

    ``funcs[i](*args)`` is the i-th derivative.  When ``exprs`` and ``args`` (sympy) are
    available -- every factory in this package provides them -- all orders are evaluated
    in one GPU launch; plain callables supplied by the user are called as they are.
    """

    funcs: Any = field()
    exprs: Any = field(kw_only=True, default=None)
    args: Sequence | None = field(kw_only=True, default=None)
    #: index of the generated device function family (``kernels.BETA_FAMILIES``) when these are the plain
    #: beta derivatives of ``x_ave`` / ``u_ave``: models then evaluate them on the device in one fused launch
    family: int | None = field(kw_only=True, default=None)
    _cache: dict = field(init=False, repr=False, factory=dict)

    @cached.meth
    def _program(self, order: int) -> ExprProgram:
        return ExprProgram([self.exprs[i] for i in range(order + 1)], self.args)

    @staticmethod
    def _apply_minus_log(X, order):
        return _minus_log_program(order)(X)

    def derivs(self, data=None, order=None, args=None, minus_log=False, order_dim="order", concat_kws=None, norm=False):
        """Derivatives for orders ``0..order``; concatenated along ``order_dim`` unless it is None."""
        if data is not None:
            args = data.derivs_args
            if order is None:
                order = data.order
        if args is None:
            msg = "must specify args or data"
            raise ValueError(msg)
        if order is None:
            msg = "must specify order or data"
            raise ValueError(msg)
        out = None
        if self.exprs is not None and self.args is not None:
            try:
                out = self._program(order)(*args)
            except NotImplementedError:
                # not a polynomial / log expression in Symbols and Indexed entries (a user's
                # Derivatives.from_sympy may hold anything): evaluate the lambdified functions
                # as the reference does (models.py:371)
                out = None
        if out is None:
            out = [self.funcs[i](*args) for i in range(order + 1)]
        if minus_log:
            out = self._apply_minus_log(X=out, order=order)
        if norm:
            out = [x / math.factorial(i) for i, x in enumerate(out)]
        if order_dim is not None:
            out = concat(out, dim=order_dim, **(concat_kws or {}))
        return out

    def coefs(self, data=None, args=None, order=None, minus_log=False, order_dim="order"):
        """Taylor coefficients: ``derivs(..., norm=True)``."""
        return self.derivs(data=data, args=args, order=order, minus_log=minus_log, order_dim=order_dim, norm=True)

    @classmethod
    def from_sympy(cls, exprs, args, family=None):
        return cls(funcs=Lambdify(exprs, args=args), exprs=exprs, args=args, family=family)


@lru_cache(10)
def taylor_series_norm(order, order_dim="order"):
    """``taylor_series_coefficients = derivs * taylor_series_norm``."""
    out = np.array([1 / math.factorial(i) for i in range(order + 1)])
    return DataArray(out, dims=order_dim) if order_dim is not None else out


################################################################################
# models
################################################################################
#: evaluate the numeric beta families with the generated device functions (``txb_beta_predict_f64``) straight
#: from the comoment records the kernels left on the device; False forces the expression-table path
FUSED_K7 = True


class _FusedPlan:
    """What the fused K7 launch needs: the comoment records on the device, the dims / coords of their leading axes,
    the generated family, and (lnPi) the two offset arrays."""

    __slots__ = ("rec", "lead_dims", "lead_shape", "coords", "family", "lnpi0", "mudotn")

    def __init__(self, rec, lead_dims, lead_shape, coords, family, lnpi0=None, mudotn=None) -> None:
        self.rec, self.lead_dims, self.lead_shape, self.coords = rec, lead_dims, lead_shape, coords
        self.family, self.lnpi0, self.mudotn = family, lnpi0, mudotn

    def launch(self, order, dalpha=None, minus_log=False, mode=None, norm=False):
        K = cm.K
        mode = K.PREDICT_SUM if mode is None else mode
        if self.family >= 4:
            return K.lnpi_predict(self.rec, order, self.family == 4, self.lnpi0, self.mudotn, dalpha, mode=mode, norm=norm)
        return K.beta_predict(self.rec, order, self.family, dalpha, minus_log=bool(minus_log), mode=mode, norm=norm)


def _fused_records(model, order, minus_log=False):
    """A :class:`_FusedPlan` when ``model`` qualifies for the fused K7 kernels: plain beta derivatives of ``x_ave`` /
    ``u_ave`` or the lnPi family (``derivatives.family``), comoment data (left on the device by a kernel of this
    package, else uploaded: the records are small), no ``deriv`` dimension, the default ``meta`` callback (lnPi: its
    own callback), order within the generated range.  None otherwise: the expression-table path runs."""
    from .data import DataCallback, DataCentralMomentsBase

    fam = getattr(model.derivatives, "family", None)
    data = model.data
    if not FUSED_K7 or fam is None or not isinstance(data, DataCentralMomentsBase):
        return None
    lnpi = fam >= 4
    if order > cm.K.BETA_MAX_ORDER or data.xalpha or (lnpi and minus_log):
        return None
    if order > (data.order + 1 if lnpi else data.order):
        return None
    if bool(data.central) != (fam in (0, 1, 4)) or bool(data.x_is_u) != (fam in (1, 3, 4, 5)):
        return None
    if data.dxduave.mom_ndim != 2 or tuple(data.dxduave.mom_dims) != (data.xmom_dim, data.umom_dim):
        return None
    obj = data.dxduave.obj
    lead_dims = tuple(obj.dims[:-2])
    lead_shape = tuple(obj.shape[:-2])
    lnpi0 = mudotn = None
    if lnpi:
        meta = data.meta
        if type(meta).__name__ != "lnPiDataCallback":
            return None
        # lnPi0 and mu.N broadcast against each other must span a SUFFIX of the records' leading dims
        l0, mn = meta.lnPi0_ave, meta.mudotN
        inner = tuple(d for d in lead_dims if d in l0.dims or d in mn.dims)
        if set(l0.dims) | set(mn.dims) != set(inner) or inner != lead_dims[len(lead_dims) - len(inner):] or not inner:
            return None
        shape = tuple(obj.sizes[d] for d in inner)

        def spread(a):
            v = np.asarray(a.transpose(*[d for d in inner if d in a.dims]).values, dtype=np.float64)
            v = v.reshape([obj.sizes[d] if d in a.dims else 1 for d in inner])
            return np.ascontiguousarray(np.broadcast_to(v, shape)).reshape(-1)

        if any(a.sizes[d] != obj.sizes[d] for a in (l0, mn) for d in a.dims):
            return None
        lnpi0, mudotn = spread(l0), spread(mn)
    elif type(data.meta) is not DataCallback:
        return None
    rec = data.dxduave.device_records()
    if rec is None:
        rec = cm.to_device(np.ascontiguousarray(np.asarray(obj.values, dtype=np.float64)))
    coords = {k: np.asarray(c.values) for k, c in obj.coords.items() if k in lead_dims}
    return _FusedPlan(rec, lead_dims, lead_shape, coords, fam, lnpi0, mudotn)


def _predict_series(coefs: DataArray, order_dim: str, dvals: np.ndarray, mode: int) -> np.ndarray:
    """``sum_p coefs[p] * d**p`` on the device; returns ``[nd, (np1,) *rest]``."""
    c = coefs.transpose(order_dim, ...)
    rest_shape = c.shape[1:]
    flat = np.ascontiguousarray(np.asarray(c.values, dtype=np.float64).reshape(c.shape[0], -1))
    out = cm._host(cm.K.poly_predict(flat, dvals, mode))
    lead = (len(dvals),) if mode == cm.K.PREDICT_SUM else (len(dvals), c.shape[0])
    return out.reshape(*lead, *rest_shape), c.dims[1:]


@attrs.define
class ExtrapModel(MyAttrsMixin):
    """Taylor-series extrapolation about ``alpha0``."""

    alpha0: float = field(converter=float)
    data: AbstractData = field(validator=attv.instance_of(AbstractData))
    derivatives: Derivatives = field(validator=attv.instance_of(Derivatives))
    order: int | None = field(default=attrs.Factory(lambda self: self.data.order, takes_self=True))
    minus_log: bool | None = field(kw_only=True, default=False, converter=attc.default_if_none(False))
    alpha_name: str = field(kw_only=True, default="alpha", converter=str)
    _cache: dict = field(init=False, repr=False, factory=dict)

    @cached.meth
    def _derivs(self, order, order_dim, minus_log):
        fused = _fused_records(self, order, minus_log) if order_dim is not None else None
        if fused is not None:
            out = cm._host(fused.launch(order, minus_log=minus_log, mode=cm.K.PREDICT_DERIVS))
            return DataArray(out.reshape(order + 1, *fused.lead_shape), dims=(order_dim, *fused.lead_dims),
                             coords=fused.coords or None)
        return self.derivatives.derivs(
            data=self.data, order=order, norm=False, minus_log=minus_log, order_dim=order_dim
        )

    def derivs(self, order=None, order_dim="order", minus_log=None, norm=False):
        if minus_log is None:
            minus_log = self.minus_log
        if order is None:
            order = self.order
        out = self._derivs(order=order, order_dim=order_dim, minus_log=minus_log)
        if norm:
            return out * taylor_series_norm(order, order_dim)
        return out

    def coefs(self, order=None, order_dim="order", minus_log=None):
        return self.derivs(order=order, order_dim=order_dim, minus_log=minus_log, norm=True)

    def __call__(self, *args, **kwargs):
        return self.predict(*args, **kwargs)

    def predict(self, alpha, order=None, order_dim="order", cumsum=False, no_sum=False, minus_log=None,
                alpha_name=None, dalpha_coords="dalpha", alpha0_coords=True):
        """Evaluate the Taylor series at ``alpha`` (sum, ``cumsum`` over orders, or ``no_sum``)."""
        if order is None:
            order = self.order
        if alpha_name is None:
            alpha_name = self.alpha_name
        if minus_log is None:
            minus_log = self.minus_log
        alpha = xrwrap_alpha(alpha, name=alpha_name)
        dalpha = alpha - self.alpha0
        if alpha.ndim > 1:
            msg = "alpha must be a scalar or 1-D"
            raise ValueError(msg)
        mode = cm.K.PREDICT_TERMS if no_sum else (cm.K.PREDICT_CUMSUM if cumsum else cm.K.PREDICT_SUM)
        dvals = np.atleast_1d(np.asarray(dalpha.values, dtype=np.float64))
        fused = _fused_records(self, order, minus_log)
        if fused is not None:
            # records -> derivatives -> -log -> / i! -> Taylor sum in ONE launch; only the predictions come back
            rest_dims = fused.lead_dims
            vals = cm._host(fused.launch(order, dvals, minus_log=minus_log, mode=mode))
            lead = (len(dvals),) if mode == cm.K.PREDICT_SUM else (len(dvals), order + 1)
            vals = vals.reshape(*lead, *fused.lead_shape)
            coords = {alpha_name: np.atleast_1d(np.asarray(alpha.values)), **fused.coords}
        else:
            coefs = self.coefs(order=order, order_dim=order_dim, minus_log=minus_log)
            vals, rest_dims = _predict_series(coefs, order_dim, dvals, mode)
            coords = {alpha_name: np.atleast_1d(np.asarray(alpha.values))}
            for k, c in coefs.coords.items():
                if k in rest_dims:
                    coords[k] = np.asarray(c.values)
        dims = [alpha_name] + ([order_dim] if mode != cm.K.PREDICT_SUM else []) + list(rest_dims)
        out = DataArray(vals, dims=dims, coords=coords)
        if dalpha_coords is not None:
            out = out.assign_coords({dalpha_coords: (alpha_name, dvals)})
        if alpha.ndim == 0:
            out = out.isel({alpha_name: 0})
        if alpha0_coords:
            if not isinstance(alpha0_coords, str):
                alpha0_coords = alpha_name + "0"
            out = out.assign_coords({alpha0_coords: self.alpha0})
        return out

    def resample(self, sampler, **kws):
        """New model on bootstrap-resampled data."""
        return self.new_like(
            order=self.order, alpha0=self.alpha0, derivatives=self.derivatives,
            data=self.data.resample(sampler=sampler, **kws), minus_log=self.minus_log, alpha_name=self.alpha_name,
        )


@attrs.define
class StateCollection(MyAttrsMixin):
    """Sequence of models (states)."""

    states: Sequence = field()
    kws: dict = field(kw_only=True, converter=convert_mapping_or_none_to_dict, default=None)
    _cache: dict = field(init=False, repr=False, factory=dict)

    def __call__(self, *args, **kwargs):
        return self.predict(*args, **kwargs)

    def __len__(self) -> int:
        return len(self.states)

    def __getitem__(self, idx):
        return self.states[idx]

    @property
    def alpha_name(self):
        return getattr(self[0], "alpha_name", "alpha")

    def resample(self, sampler, **kws):
        """Resample every state.  One sampler spec (mapping / array / IndexSampler) is applied to
        all states -- a mapping such as ``{"nrep": n}`` makes each state draw its own indices."""
        if isinstance(sampler, (np.ndarray, DataArray, cm.IndexSampler, Mapping)):
            sampler = [sampler] * len(self)
        elif len(sampler) != len(self):
            msg = f"{len(sampler)=} must equal {len(self)=}"
            raise ValueError(msg)
        return type(self)(
            states=tuple(state.resample(sampler=s, **kws) for state, s in zip(self.states, sampler)), **self.kws
        )

    def map(self, func, *args, **kwargs):
        if isinstance(func, str):
            return [getattr(s, func)(*args, **kwargs) for s in self]
        return [func(s, *args, **kwargs) for s in self]

    def map_concat(self, func, concat_dim=None, concat_kws=None, *args, **kwargs):
        out = self.map(func, *args, **kwargs)
        if is_dataarray(out[0]):
            if concat_dim is None:
                concat_dim = DataArray(np.asarray(self.alpha0), dims=self.alpha_name, name=self.alpha_name)
            out = concat(out, dim=concat_dim, **(concat_kws or {}))
        return out

    def append(self, states, sort=True, key=None, **kws):
        new_states = list(self.states) + list(states)
        if sort:
            new_states = sorted(new_states, key=key or (lambda x: x.alpha0), **kws)
        return type(self)(new_states, **self.kws)

    @property
    def order(self):
        return min(m.order for m in self)

    @property
    def alpha0(self):
        return [m.alpha0 for m in self]

    def _check_alpha(self, alpha, bounded=False) -> None:
        if bounded:
            try:
                seq = iter(alpha)
            except TypeError:
                seq = [alpha]
            lb, ub = self[0].alpha0, self[-1].alpha0
            for a in seq:
                if a < lb or a > ub:
                    msg = f"{a} outside of bounds [{lb}, {ub}]"
                    raise ValueError(msg)


def xr_weights_minkowski(deltas, m=20, dim="state"):
    deltas_m = deltas**m
    return 1.0 - deltas_m / deltas_m.sum(dim)


class PiecewiseMixin:
    """Pick the two states used to predict at a given ``alpha``."""

    def _indices_between_alpha(self, alpha):
        idx = int(np.digitize(alpha, self.alpha0, right=False)) - 1
        idx = min(max(idx, 0), len(self) - 2)
        return [idx, idx + 1]

    def _indices_nearest_alpha(self, alpha):
        return np.argsort(np.abs(np.array(self.alpha0) - alpha))[:2]

    def _indices_alpha(self, alpha, method):
        if method is None or method == "between":
            return self._indices_between_alpha(alpha)
        if method == "nearest":
            return self._indices_nearest_alpha(alpha)
        msg = f"unknown method {method}"
        raise ValueError(msg)

    def _states_alpha(self, alpha, method):
        return [self[i] for i in self._indices_alpha(alpha, method)]


@attrs.define
class ExtrapWeightedModel(StateCollection, PiecewiseMixin):
    """Blend the extrapolations from two states with Minkowski-distance weights.

    ``states`` must be sorted by ``alpha0``.
    """

    def predict(self, alpha, order=None, order_dim="order", cumsum=False, minus_log=None, alpha_name=None,
                method=None, bounded=False):
        """``method``: None / 'between' (bracketing states) or 'nearest' (two closest states)."""
        self._check_alpha(alpha, bounded)
        if order is None:
            order = self.order
        if alpha_name is None:
            alpha_name = self.alpha_name
        if len(self) == 2:
            states = self.states
        else:
            if np.array(alpha).ndim > 0:
                return concat(
                    (
                        self.predict(alpha=a, order=order, order_dim=order_dim, cumsum=cumsum, minus_log=minus_log,
                                     alpha_name=alpha_name, method=method)
                        for a in alpha
                    ),
                    dim=alpha_name,
                )
            states = self._states_alpha(alpha, method)
        out = concat(
            [
                m.predict(alpha, order=order, order_dim=order_dim, cumsum=cumsum, minus_log=minus_log,
                          alpha_name=alpha_name)
                for m in states
            ],
            dim="state",
        )
        w = xr_weights_minkowski(abs(out["dalpha"]))
        return (out * w).sum("state") / w.sum("state")


@attrs.define
class InterpModel(StateCollection):
    """Polynomial through the derivatives of all states (generalised Hermite interpolation)."""

    @cached.meth
    def _mat_inv(self, order) -> np.ndarray:
        """Inverse of the generalised Hermite matrix, ``[npoly, nstate, order+1]`` (models.py:877-893)."""
        nstate = len(self)
        npoly = nstate * (order + 1)
        from scipy.special import factorial as sp_factorial

        power = np.arange(npoly)
        fact = sp_factorial(power)
        # row (state i, derivative j): d^j/dalpha^j of alpha**p at alpha_i = alpha_i**(p-j) * p! / (p-j)!
        # (same operation order as the reference, models.py:877-891, so that exact
        # known-answer tests on the coefficients hold bit for bit)
        mat = np.zeros((npoly, npoly))
        for i, m in enumerate(self.states):
            a = m.alpha0
            for j in range(order + 1):
                ok = power >= j
                pj = np.where(ok, power - j, 0)
                mat[i * (order + 1) + j] = np.where(ok, (a**pj) * fact / sp_factorial(pj), 0.0)
        return np.linalg.inv(mat).reshape(npoly, nstate, order + 1)

    def _fused_coefs(self, order, minus_log):
        """Polynomial coefficients ``[npoly, nelem]`` ON THE DEVICE when every state qualifies for the fused K7
        kernels (same leading dims): per state one launch for the derivatives of all records, one for its share of
        ``mat_inv . derivs``.  Returns ``(coefs_dev, lead_dims, lead_shape, coords)`` or None."""
        mls = [m.minus_log if minus_log is None else minus_log for m in self.states]
        fused = [_fused_records(m, order, ml) for m, ml in zip(self.states, mls)]
        if any(f is None for f in fused) or len({(f.lead_dims, f.lead_shape) for f in fused}) != 1:
            return None
        mat_inv = self._mat_inv(order)
        coefs_dev = None
        for s, (f, ml) in enumerate(zip(fused, mls)):
            d = f.launch(order, minus_log=ml, mode=cm.K.PREDICT_DERIVS)
            coefs_dev = cm.K.combine(np.ascontiguousarray(mat_inv[:, s, :]), d, out=coefs_dev)
        return coefs_dev, fused[0].lead_dims, fused[0].lead_shape, fused[0].coords

    @cached.meth
    def coefs(self, order=None, order_dim="porder", minus_log=None):
        if order is None:
            order = self.order
        fused = self._fused_coefs(order, minus_log)
        if fused is not None:
            coefs_dev, lead_dims, lead_shape, coords = fused
            vals = cm._host(coefs_dev).reshape(-1, *lead_shape)
            return DataArray(vals, dims=(order_dim, *lead_dims), coords=coords or None)
        mat_inv = DataArray(self._mat_inv(order), dims=[order_dim, "state", "order"])
        derivs = concat(
            [m.derivs(order, norm=False, minus_log=minus_log, order_dim="order") for m in self.states], dim="state"
        )
        return dot(mat_inv, derivs, dims=["state", "order"])

    def predict(self, alpha, order=None, order_dim="porder", minus_log=None, alpha_name=None):
        if order is None:
            order = self.order
        if alpha_name is None:
            alpha_name = self.alpha_name
        alpha = xrwrap_alpha(alpha, name=alpha_name)
        avals = np.atleast_1d(np.asarray(alpha.values, dtype=np.float64))
        fused = self._fused_coefs(order, minus_log)
        if fused is not None:
            # derivatives, mat_inv . derivs and the polynomial all stay on the device: one D2H of [alpha, *lead]
            coefs_dev, rest_dims, lead_shape, fcoords = fused
            vals = cm._host(cm.K.poly_predict(coefs_dev, avals, cm.K.PREDICT_SUM)).reshape(len(avals), *lead_shape)
            coords = {alpha_name: avals, **fcoords}
        else:
            coefs = self.coefs(order=order, order_dim=order_dim, minus_log=minus_log)
            vals, rest_dims = _predict_series(coefs, order_dim, avals, cm.K.PREDICT_SUM)
            coords = {alpha_name: avals}
            for k, c in coefs.coords.items():
                if k in rest_dims:
                    coords[k] = np.asarray(c.values)
        out = DataArray(vals, dims=[alpha_name, *rest_dims], coords=coords)
        return out.isel({alpha_name: 0}) if alpha.ndim == 0 else out


class InterpModelPiecewise(StateCollection, PiecewiseMixin):
    """Two-state :class:`InterpModel` chosen per ``alpha``."""

    @cached.meth
    def single_interpmodel(self, *state_indices):
        state0, state1 = (self[i] for i in state_indices)
        return InterpModel([state0, state1])

    def predict(self, alpha, order=None, order_dim="porder", minus_log=None, alpha_name=None, method=None,
                bounded=False):
        self._check_alpha(alpha, bounded)
        if alpha_name is None:
            alpha_name = self.alpha_name
        kws = dict(order=order, order_dim=order_dim, minus_log=minus_log, alpha_name=alpha_name)
        if len(self) == 2:
            return self.single_interpmodel(0, 1).predict(alpha=alpha, **kws)
        try:
            seq = iter(alpha)
        except TypeError:
            seq = [alpha]
        out = []
        for a in seq:
            model = self.single_interpmodel(*(int(i) for i in self._indices_alpha(alpha=a, method=method)))
            out.append(model.predict(alpha=a, **kws))
        return out[0] if len(out) == 1 else concat(out, dim=alpha_name)


@attrs.define
class PerturbModel(MyAttrsMixin):
    """Free-energy perturbation (exponential reweighting) from one state."""

    alpha0: float = field(converter=float)
    data: AbstractData = field(validator=attv.instance_of(AbstractData))
    alpha_name: str | None = field(default="alpha", converter=attc.default_if_none("alpha"))

    def predict(self, alpha, alpha_name=None):
        if alpha_name is None:
            alpha_name = self.alpha_name
        alpha = xrwrap_alpha(alpha, name=alpha_name)
        uv, xv = self.data.uv, self.data.xv
        rec_dim = self.data.rec_dim
        dalpha = np.atleast_1d(np.asarray(alpha.values, dtype=np.float64)) - self.alpha0
        scalar = alpha.ndim == 0
        if rec_dim not in uv.dims or rec_dim not in xv.dims:
            msg = f"uv and xv need dimension {rec_dim!r}"
            raise ValueError(msg)
        # uv[(rep,) rec], xv[(rep,) rec, *val]: one streaming reweighting kernel per replicate
        # (``txb_perturb_predict_f64``); the result has dims (alpha, [rep,] *val) like the reference's
        # broadcast ``num / den`` (models.py:1037-1046).
        lead = [d for d in uv.dims if d != rec_dim]
        if any(d not in xv.dims for d in lead):
            msg = f"dims of uv {uv.dims} beyond {rec_dim!r} must be dims of xv {xv.dims}"
            raise ValueError(msg)
        val_dims = [d for d in xv.dims if d != rec_dim and d not in lead]
        xt = xv.transpose(*lead, rec_dim, *val_dims)
        ut = uv.transpose(*lead, rec_dim)
        xa, ua = np.asarray(xt.values, dtype=np.float64), np.asarray(ut.values, dtype=np.float64)
        lead_shape = ua.shape[:-1]
        nrec, val_shape = ua.shape[-1], xa.shape[len(lead_shape) + 1 :]
        xa = xa.reshape(-1, nrec, *val_shape)
        ua = ua.reshape(-1, nrec)
        parts = []
        for r in range(ua.shape[0]):
            xd, ud = cm.to_device(np.ascontiguousarray(xa[r])), cm.to_device(np.ascontiguousarray(ua[r]))
            parts.append(cm._host(cm.K.perturb_predict(xd, ud, dalpha)))
        out = np.stack(parts, axis=1).reshape(len(dalpha), *lead_shape, *val_shape)
        res = DataArray(out, dims=(alpha_name, *lead, *val_dims), coords={alpha_name: np.atleast_1d(alpha.values)})
        return res.isel({alpha_name: 0}, drop=True) if scalar else res

    def resample(self, sampler, **kws):
        return type(self)(alpha0=self.alpha0, data=self.data.resample(sampler=sampler, **kws), alpha_name=self.alpha_name)


@attrs.define
class MBARModel(StateCollection):
    """MBAR reweighting needs ``pymbar`` (not part of the B200 hot path)."""

    def __attrs_pre_init__(self):
        from importlib.util import find_spec

        if find_spec("pymbar") is None:
            msg = "need pymbar to use this"
            raise ImportError(msg)

    def predict(self, alpha, alpha_name=None):
        msg = "MBARModel is outside the B200 hot path (SURVEY.md section 2.2)"
        raise NotImplementedError(msg)

    def resample(self, *args, **kwargs):
        msg = "resample not implemented for this class"
        raise NotImplementedError(msg)

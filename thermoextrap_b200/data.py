"""Data containers (mirror of ``thermoextrap.data``; every reduce / resample runs on the GPU).

Scheme (reference ``data.py:5-11``):

* ``uv, xv``   samples of energy ``u`` and observable ``x``
* ``u[i] = <u**i>``, ``xu[i] = <x u**i>`` (raw) or ``du[i] = <du**i>``, ``dxdu[i] = <dx du**i>`` (central)
* ``xu[i, j] = <d^j x/d beta^j u**i>`` when ``x`` depends on the expansion variable (``deriv_dim``)

Reference symbols mirrored here: ``DataSelector`` (data.py:91-162),
``DataCallbackABC/DataCallback`` (:165-250), ``DataValues/DataValuesCentral`` (:326-656),
``factory_data_values`` (:660-730), ``DataCentralMomentsBase`` (:791-962),
``DataCentralMoments`` (:965-1618), ``DataCentralMomentsVals`` (:1621-1813).
"""

from __future__ import annotations

from abc import abstractmethod
from typing import Any

import attrs
import numpy as np
from attrs import field
from attrs import validators as attv

from . import backend as cm
from .backend import MISSING, CentralMomentsData, IndexSampler
from .core import cached
from .core.attrs_utils import MyAttrsMixin, convert_dims_to_tuple, convert_mapping_or_none_to_dict
from .core.xr import DataArray, concat, is_dataarray
from .core.xrutils import xrwrap_uv, xrwrap_xv

__all__ = [
    "DataCallback",
    "DataCallbackABC",
    "DataCentralMoments",
    "DataCentralMomentsVals",
    "DataSelector",
    "DataValues",
    "DataValuesCentral",
    "factory_data_values",
]


def _raise_if_not_dataarray(x: object, name: str | None = None) -> None:
    if not is_dataarray(x):
        msg = f"type({name})={type(x)} must be a DataArray."
        raise TypeError(msg)


def _validate_dims(self: Any, attribute: Any, dims: Any) -> None:  # noqa: ARG001
    for d in dims:
        if d not in self.data.dims:
            msg = f"{d} not in data.dimensions {self.data.dims}"
            raise ValueError(msg)


@attrs.define(frozen=True)
class DataSelector(MyAttrsMixin):
    """Index a labelled array like ``sel[i, j]`` along named ``dims``.

    >>> x = DataArray([[1, 2, 3], [4, 5, 6]], dims=["x", "y"])
    >>> s = DataSelector(data=x, dims=["y", "x"])
    >>> int(s[0, 1])
    4
    """

    data: Any = field(validator=attv.instance_of(DataArray))
    dims: tuple = field(converter=convert_dims_to_tuple, validator=_validate_dims)

    @classmethod
    def from_defaults(cls, data, *, dims=None, mom_dim="moment", deriv_dim=None):
        if dims is None:
            dims = (mom_dim, deriv_dim) if deriv_dim is not None else (mom_dim,)
        return cls(data=data, dims=dims)

    def __getitem__(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        if len(idx) != len(self.dims):
            msg = f"bad idx {idx}, vs dims {self.dims}"
            raise ValueError(msg)
        return self.data.isel(dict(zip(self.dims, idx)), drop=True)

    def __repr__(self) -> str:
        return repr(self.data)


@attrs.define
class DataCallbackABC(MyAttrsMixin):
    """Hook object carried as ``data.meta`` (lnPi and volume models add derivative arguments)."""

    @abstractmethod
    def check(self, data) -> None:
        """Consistency checks between ``self`` and ``data``."""

    @abstractmethod
    def derivs_args(self, data, *, derivs_args):
        """Return the (possibly extended) tuple of derivative-function arguments."""
        return derivs_args

    def resample(self, data, *, meta_kws, sampler, **kws):
        raise NotImplementedError

    def reduce(self, data, *, meta_kws, **kws):
        raise NotImplementedError

    def __repr__(self) -> str:
        return f"<{self.__class__.__name__}>"


@attrs.define
class DataCallback(DataCallbackABC):
    """Pass-through callback (the default ``meta``)."""

    def check(self, data) -> None:
        pass

    def derivs_args(self, data, *, derivs_args):  # noqa: ARG002
        return derivs_args

    def resample(self, data, *, meta_kws, sampler, **kws):  # noqa: ARG002
        return self

    def reduce(self, data, *, meta_kws, **kws):  # noqa: ARG002
        return self


def _meta_converter(meta):
    return DataCallback() if meta is None else meta


def _meta_validator(self, attribute, meta) -> None:  # noqa: ARG001
    if not isinstance(meta, DataCallbackABC):
        msg = "meta must be None or subclass of DataCallbackABC"
        raise TypeError(msg)
    meta.check(data=self)


@attrs.define
class AbstractData(MyAttrsMixin):
    """Fields shared by every data class."""

    meta: DataCallbackABC = field(kw_only=True, converter=_meta_converter, validator=_meta_validator, default=None)
    umom_dim: str = field(kw_only=True, default="umom")
    deriv_dim: str | None = field(kw_only=True, default=None)
    x_is_u: bool = field(kw_only=True, default=False)
    _cache: dict = field(init=False, repr=False, factory=dict)

    @property
    def xalpha(self) -> bool:
        """Whether ``x`` depends explicitly on the expansion variable (``deriv_dim`` set)."""
        return self.deriv_dim is not None

    def pipe(self, func, *args, **kwargs):
        return func(self, *args, **kwargs)


###############################################################################
# values-based data (reference data.py:326-730)
###############################################################################
class _LazyGather:
    """``base.isel({rec: indices})`` that is only materialised if somebody looks at it.

    The reference materialises ``[rep, rec, *val]`` copies in ``DataValuesBase.resample``
    (data.py:430-431); here the resampled averages come straight from the bootstrap
    kernels and the gathered samples are built on demand.
    """

    def __init__(self, base: DataArray, sampler, rec_dim: str, rep_dim: str) -> None:
        self.base, self.sampler, self.rec_dim, self.rep_dim = base, sampler, rec_dim, rep_dim

    def compute(self) -> DataArray:
        # only here is the sampler's index table itself needed (a device-side plan stays a plan until now)
        idx = DataArray(self.sampler.indices, dims=(self.rep_dim, self.rec_dim))
        return self.base.isel({self.rec_dim: idx})


def _is_sample_array(x) -> bool:
    return is_dataarray(x) or isinstance(x, _LazyGather)


@attrs.define
class DataValuesBase(AbstractData):
    """Samples ``uv[rec]``, ``xv[rec, *val]`` and the expansion order."""

    _uv: Any = field(alias="uv")
    _xv: Any = field(alias="xv")
    order: int = field()
    rec_dim: str = field(kw_only=True, default="rec")
    # (base object, sampler, rep_dim) when this object is a bootstrap resample
    _boot: tuple | None = field(kw_only=True, default=None, repr=False, alias="boot")

    _CENTRAL = False

    def __attrs_post_init__(self) -> None:
        if not _is_sample_array(self._uv):
            msg = f"type(uv)={type(self._uv)} must be a DataArray."
            raise TypeError(msg)
        if not _is_sample_array(self._xv):
            msg = f"type(xv)={type(self._xv)} must be a DataArray."
            raise TypeError(msg)

    @property
    def uv(self) -> DataArray:
        if isinstance(self._uv, _LazyGather):
            self._uv = self._uv.compute()
        return self._uv

    @property
    def xv(self) -> DataArray:
        if self.x_is_u and isinstance(self._xv, _LazyGather):
            return self.uv
        if isinstance(self._xv, _LazyGather):
            self._xv = self._xv.compute()
        return self._xv

    @classmethod
    def from_vals(cls, uv, xv, *, order, rec_dim="rec", umom_dim="umom", deriv_dim=None, meta=None, x_is_u=False):
        return cls(
            uv=uv, xv=uv if xv is None else xv, order=order, rec_dim=rec_dim, umom_dim=umom_dim,
            deriv_dim=deriv_dim, meta=meta, x_is_u=x_is_u,
        )

    @property
    def central(self) -> bool:
        return self._CENTRAL

    def __len__(self) -> int:
        src = self._uv.base if isinstance(self._uv, _LazyGather) else self._uv
        return src.sizes[self.rec_dim]

    def resample(self, sampler, *, rep_dim="rep", meta_kws=None):
        """Bootstrap resample (reference data.py:404-449)."""
        if self._boot is not None:
            msg = "object is already resampled"
            raise ValueError(msg)
        sampler = cm.factory_sampler(sampler, data=self._xv, dim=self.rec_dim)
        if sampler.nsamp != cm.global_rows(len(self)):
            msg = f"indices.sizes[{self.rec_dim}]={sampler.nsamp} must equal {len(self)=}"
            raise ValueError(msg)
        meta = self.meta.resample(
            data=self, meta_kws={} if meta_kws is None else meta_kws, sampler=sampler, rep_dim=rep_dim
        )
        return type(self)(
            uv=_LazyGather(self._uv, sampler, self.rec_dim, rep_dim),
            xv=_LazyGather(self._xv, sampler, self.rec_dim, rep_dim),
            order=self.order, rec_dim=self.rec_dim, umom_dim=self.umom_dim, deriv_dim=self.deriv_dim,
            meta=meta, x_is_u=self.x_is_u, boot=(self, sampler, rep_dim),
        )

    # -- the two reductions every values object needs ---------------------------------
    def _comoments(self) -> tuple[CentralMomentsData, CentralMomentsData | None]:
        """(comoments of (x, u), 1-D moments of u)  -- reduced, or bootstrap-resampled."""
        key = ("_comoments",)
        if key in self._cache:
            return self._cache[key]
        mdims = ("_xmom", self.umom_dim)
        if self._boot is None:
            c = cm.wrap_reduce_vals(self._xv, self._uv, mom=(1, self.order), dim=self.rec_dim, mom_dims=mdims)
        else:
            base, sampler, rep_dim = self._boot
            c = cm.wrap_resample_vals(
                base._xv, base._uv, mom=(1, self.order), sampler=sampler, dim=self.rec_dim,
                rep_dim=rep_dim, mom_dims=mdims,
            ).transpose(rep_dim, ...)
        self._cache[key] = c
        return c


def build_aves_xu(uv, xv, *, order, dim=MISSING, umom_dim="umom"):
    """``u[i] = <u**i>`` and ``xu[i] = <x u**i>`` (reference data.py:456-493)."""
    _raise_if_not_dataarray(uv, "uv")
    _raise_if_not_dataarray(xv, "xv")
    c = cm.wrap_reduce_vals(xv, uv, mom=(1, order), dim=dim, mom_dims=("_xmom", umom_dim))
    r = c.rmom()
    xu = cm.select_moment(r, "xmom_1", mom_ndim=2, mom_dims=c.mom_dims)
    u = cm.wrap_reduce_vals(uv, mom=order, dim=dim, mom_dims=umom_dim).rmom()
    return u, xu


def build_aves_dxdu(uv, xv, *, order, dim=MISSING, umom_dim="umom"):
    """``<x>``, ``<du**i>`` and ``<dx du**i>`` (reference data.py:497-536)."""
    _raise_if_not_dataarray(uv, "uv")
    _raise_if_not_dataarray(xv, "xv")
    duave = cm.wrap_reduce_vals(uv, mom=order, dim=dim, mom_dims=umom_dim).cmom()
    c = cm.wrap_reduce_vals(xv, uv, mom=(1, order), dim=dim, mom_dims=("_xmom", umom_dim))
    xave = c.select_moment("xave")
    dxduave = cm.select_moment(c.cmom(), "xmom_1", mom_ndim=2, mom_dims=c.mom_dims)
    return xave, duave, dxduave


def _xu_to_u(xu: DataArray, dim: str = "umom") -> DataArray:
    """``x == u``: shift ``umom`` by one and put 1 at ``umom = 0`` (reference data.py:539-546)."""
    x = xu.transpose(..., dim)
    v = np.asarray(x.values)
    out = np.empty((*v.shape[:-1], v.shape[-1] + 1), dtype=v.dtype)
    out[..., 0] = 1.0
    out[..., 1:] = v
    return DataArray(out, dims=x.dims).transpose(*xu.dims)


@attrs.define
class DataValues(DataValuesBase):
    """Raw-moment view of (uv, xv)."""

    _CENTRAL = False

    @cached.prop
    def xu(self) -> DataArray:
        """``<x u**n>``."""
        c = self._comoments()
        return cm.select_moment(c.rmom(), "xmom_1", mom_ndim=2, mom_dims=c.mom_dims)

    @cached.prop
    def u(self) -> DataArray:
        """``<u**n>``."""
        if self.x_is_u:
            return _xu_to_u(self.xu, self.umom_dim)
        return self._u_moments(central=False)

    def _u_moments(self, central: bool) -> DataArray:
        if self._boot is None:
            m = cm.wrap_reduce_vals(self._uv, mom=self.order, dim=self.rec_dim, mom_dims=self.umom_dim)
            return m.cmom() if central else m.rmom()
        base, sampler, rep_dim = self._boot
        return _resampled_u_moments(base._uv, sampler, self.order, self.rec_dim, rep_dim, self.umom_dim, central)

    @cached.prop
    def u_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.u, deriv_dim=None, mom_dim=self.umom_dim)

    @cached.prop
    def xu_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.xu, deriv_dim=self.deriv_dim, mom_dim=self.umom_dim)

    @property
    def derivs_args(self) -> tuple:
        out = (self.u_selector,) if self.x_is_u else (self.u_selector, self.xu_selector)
        return self.meta.derivs_args(data=self, derivs_args=out)


@attrs.define
class DataValuesCentral(DataValuesBase):
    """Central-moment view of (uv, xv)."""

    _CENTRAL = True

    @cached.prop
    def xave(self) -> DataArray:
        return self._comoments().select_moment("xave")

    @cached.prop
    def dxdu(self) -> DataArray:
        c = self._comoments()
        return cm.select_moment(c.cmom(), "xmom_1", mom_ndim=2, mom_dims=c.mom_dims)

    @cached.prop
    def du(self) -> DataArray:
        if self.x_is_u:
            return _xu_to_u(self.dxdu, dim=self.umom_dim)
        return DataValues._u_moments(self, central=True)

    @cached.prop
    def du_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.du, deriv_dim=None, mom_dim=self.umom_dim)

    @cached.prop
    def dxdu_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.dxdu, deriv_dim=self.deriv_dim, mom_dim=self.umom_dim)

    @cached.prop
    def xave_selector(self):
        if self.deriv_dim is None:
            return self.xave
        return DataSelector.from_defaults(self.xave, dims=[self.deriv_dim])

    @property
    def derivs_args(self) -> tuple:
        out = (
            (self.xave_selector, self.du_selector)
            if self.x_is_u
            else (self.xave_selector, self.du_selector, self.dxdu_selector)
        )
        return self.meta.derivs_args(data=self, derivs_args=out)


def _resampled_u_moments(uv, sampler, order, rec_dim, rep_dim, umom_dim, central) -> DataArray:
    """1-D moments of ``u`` per bootstrap replicate: ``[rep, umom]`` (x_is_u kernels)."""
    k = max(order - 1, 1)
    ud = cm.to_device(np.asarray(uv.transpose(rec_dim, ...).values, dtype=np.float64))
    out = cm._host(cm.resample_device(None, ud, None, sampler, k, x_is_u=True))
    mom = cm._moments_1d_from_x_is_u(out, order)
    rest = tuple(d for d in uv.dims if d != rec_dim)
    m = CentralMomentsData(DataArray(mom, dims=(rep_dim, *rest, umom_dim)), mom_ndim=1, mom_dims=umom_dim)
    return m.cmom() if central else m.rmom()


def factory_data_values(
    order, uv, xv, central=False, xalpha=False, rec_dim="rec", umom_dim="umom", val_dims="val",
    rep_dim="rep", deriv_dim=None, x_is_u=False, **kws,
):
    """Build a :class:`DataValues` / :class:`DataValuesCentral` (reference data.py:660-730)."""
    cls = DataValuesCentral if central else DataValues
    if xalpha and deriv_dim is None:
        msg = "if xalpha, must pass string name of derivative"
        raise ValueError(msg)
    uv = xrwrap_uv(uv, rec_dim=rec_dim, rep_dim=rep_dim)
    if xv is not None:
        xv = xrwrap_xv(xv, rec_dim=rec_dim, rep_dim=rep_dim, deriv_dim=deriv_dim, val_dims=val_dims)
    return cls.from_vals(
        uv=uv, xv=xv, order=order, rec_dim=rec_dim, umom_dim=umom_dim, deriv_dim=deriv_dim, x_is_u=x_is_u, **kws
    )


###############################################################################
# comoment-array data (reference data.py:791-1813)
###############################################################################
@attrs.define
class DataCentralMomentsBase(AbstractData):
    """Data object built on a comoment array ``dxduave[..., xmom=2, umom=order+1]``."""

    dxduave: CentralMomentsData = field(validator=attv.instance_of(CentralMomentsData))
    xmom_dim: str = field(kw_only=True, default="xmom")
    rec_dim: str = field(kw_only=True, default="rec")
    central: bool = field(kw_only=True, default=False)
    x_is_u: bool = field(kw_only=True, default=None)
    _use_cache: bool = field(kw_only=True, default=True, alias="use_cache")

    @property
    def order(self) -> int:
        return self.dxduave.sizes[self.umom_dim] - 1

    @property
    def values(self) -> DataArray:
        """The underlying comoment array."""
        return self.dxduave.obj

    @cached.meth(check_use_cache=True)
    def rmom(self) -> DataArray:
        """Raw comoments ``<x^i u^j>``."""
        return self.dxduave.rmom()

    @cached.meth(check_use_cache=True)
    def cmom(self) -> DataArray:
        """Central comoments ``<dx^i du^j>``."""
        return self.dxduave.cmom()

    def _row(self, table: DataArray, name: str) -> DataArray:
        return cm.select_moment(table, name, mom_ndim=2, mom_dims=self.dxduave.mom_dims)

    @cached.prop(check_use_cache=True)
    def xu(self) -> DataArray:
        """``<x u**n>``."""
        return self._row(self.rmom(), "xmom_1")

    def _u_like(self, table: DataArray) -> DataArray:
        if self.x_is_u:
            return cm.convert.comoments_to_moments(table, mom_dims=self.dxduave.mom_dims, mom_dims_out=self.umom_dim)
        out = self._row(table, "xmom_0")
        if self.xalpha:
            out = out.sel({self.deriv_dim: 0}, drop=True)
        return out

    @cached.prop(check_use_cache=True)
    def u(self) -> DataArray:
        """``<u**n>``."""
        return self._u_like(self.rmom())

    @cached.prop(check_use_cache=True)
    def xave(self) -> DataArray:
        """``<x>``."""
        return self.dxduave.select_moment("xave")

    @cached.prop(check_use_cache=True)
    def dxdu(self) -> DataArray:
        """``<dx du**n>``."""
        return self._row(self.cmom(), "xmom_1")

    @cached.prop(check_use_cache=True)
    def du(self) -> DataArray:
        """``<du**n>``."""
        return self._u_like(self.cmom())

    @cached.prop(check_use_cache=True)
    def u_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.u, deriv_dim=None, mom_dim=self.umom_dim)

    @cached.prop(check_use_cache=True)
    def xu_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.xu, deriv_dim=self.deriv_dim, mom_dim=self.umom_dim)

    @cached.prop(check_use_cache=True)
    def xave_selector(self):
        if self.deriv_dim is None:
            return self.xave
        return DataSelector(self.xave, dims=[self.deriv_dim])

    @cached.prop(check_use_cache=True)
    def du_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.du, deriv_dim=None, mom_dim=self.umom_dim)

    @cached.prop(check_use_cache=True)
    def dxdu_selector(self) -> DataSelector:
        return DataSelector.from_defaults(self.dxdu, deriv_dim=self.deriv_dim, mom_dim=self.umom_dim)

    @property
    def derivs_args(self) -> tuple:
        """Arguments for the derivative functions: ``funcs[i](*data.derivs_args)``."""
        if not self.x_is_u:
            out = (
                (self.xave_selector, self.du_selector, self.dxdu_selector)
                if self.central
                else (self.u_selector, self.xu_selector)
            )
        elif self.central:
            out = (self.xave_selector, self.du_selector)
        else:
            out = (self.u_selector,)
        return self.meta.derivs_args(data=self, derivs_args=out)


@attrs.define
class DataCentralMoments(DataCentralMomentsBase):
    """Comoment blocks along ``rec_dim`` (pre-reduced data)."""

    def __len__(self) -> int:
        return self.values.sizes[self.rec_dim]

    def reduce(self, dim=MISSING, axis=MISSING, meta_kws=None, **kwargs):
        """Merge blocks along ``dim`` (default ``rec_dim``)."""
        if dim is MISSING and axis is MISSING:
            dim = self.rec_dim
        kws = dict(dim=dim, axis=axis, **kwargs)
        return self.new_like(
            dxduave=self.dxduave.reduce(**kws), meta=self.meta.reduce(data=self, meta_kws=meta_kws, **kws)
        )

    def resample(self, sampler, dim=MISSING, axis=MISSING, rep_dim="rep", parallel=None, meta_kws=None, **kwargs):
        """Bootstrap over pre-reduced blocks (K3)."""
        if dim is MISSING and axis is MISSING:
            dim = self.rec_dim
        sampler = cm.factory_sampler(
            sampler, data=self.dxduave.obj, dim=dim, axis=axis, mom_ndim=self.dxduave.mom_ndim,
            mom_dims=self.dxduave.mom_dims, rep_dim=rep_dim, parallel=parallel,
        )
        kws = dict(sampler=sampler, dim=dim, axis=axis, rep_dim=rep_dim, parallel=parallel, **kwargs)
        dxdu_new = self.dxduave.resample_and_reduce(**kws).transpose(rep_dim, ...)
        meta = self.meta.resample(data=self, meta_kws=meta_kws, **kws)
        return self.new_like(dxduave=dxdu_new, rec_dim=rep_dim, meta=meta)

    # ---- constructors --------------------------------------------------------------
    @classmethod
    def from_raw(cls, raw, rec_dim="rec", xmom_dim="xmom", umom_dim="umom", deriv_dim=None, central=False,
                 x_is_u=False, meta=None, **kwargs):
        """From raw moments ``raw[..., i, j] = <x**i u**j>`` (``[...,0,0]`` = weight)."""
        if x_is_u:
            data = cm.convert.moments_type(raw, mom_ndim=1, mom_dims=umom_dim, to="central", **kwargs)
        else:
            data = cm.convert.moments_type(raw, mom_ndim=2, mom_dims=(xmom_dim, umom_dim), to="central", **kwargs)
        return cls.from_data(
            data, rec_dim=rec_dim, xmom_dim=xmom_dim, umom_dim=umom_dim, deriv_dim=deriv_dim, central=central,
            meta=meta, x_is_u=x_is_u,
        )

    @classmethod
    def from_vals(cls, uv, xv, order, xmom_dim="xmom", umom_dim="umom", rec_dim="rec", deriv_dim=None,
                  central=False, weight=None, axis=MISSING, dim=MISSING, meta=None, x_is_u=False, **kwargs):
        """From samples; reduces along ``axis`` / ``dim`` (default ``axis=0``)."""
        _raise_if_not_dataarray(uv)
        if axis is MISSING and dim is MISSING:
            axis = 0
        if xv is None or x_is_u:
            dxduave = cm.wrap_reduce_vals(
                uv, weight=weight, axis=axis, dim=dim, mom=order + 1, mom_dims=umom_dim, **kwargs
            ).moments_to_comoments(mom_dims_out=(xmom_dim, umom_dim), mom=(1, order))
        else:
            _raise_if_not_dataarray(xv)
            dxduave = cm.wrap_reduce_vals(
                xv, uv, weight=weight, axis=axis, dim=dim, mom=(1, order), mom_dims=(xmom_dim, umom_dim), **kwargs
            )
        return cls(dxduave=dxduave, xmom_dim=xmom_dim, umom_dim=umom_dim, rec_dim=rec_dim, deriv_dim=deriv_dim,
                   central=central, meta=meta, x_is_u=x_is_u)

    @classmethod
    def from_data(cls, data, rec_dim="rec", xmom_dim="xmom", umom_dim="umom", deriv_dim=None, central=False,
                  meta=None, x_is_u=False, **kwargs):
        """From a comoment array in the layout of reference data.py:1233-1236."""
        _raise_if_not_dataarray(data)
        if x_is_u:
            dxduave = cm.wrap(data, mom_ndim=1, mom_dims=umom_dim, **kwargs).moments_to_comoments(
                mom_dims_out=(xmom_dim, umom_dim), mom=(1, -1)
            )
        else:
            dxduave = cm.wrap(data, mom_ndim=2, mom_dims=(xmom_dim, umom_dim), **kwargs)
        return cls(dxduave=dxduave, xmom_dim=xmom_dim, umom_dim=umom_dim, rec_dim=rec_dim, deriv_dim=deriv_dim,
                   central=central, meta=meta, x_is_u=x_is_u)

    @classmethod
    def from_resample_vals(cls, xv, uv, order, sampler, weight=None, axis=MISSING, dim=MISSING, xmom_dim="xmom",
                           umom_dim="umom", rep_dim="rep", deriv_dim=None, central=False, meta=None,
                           meta_kws=None, x_is_u=False, parallel=None, **kwargs):
        """Bootstrap straight from samples (K2), without keeping the samples."""
        if xv is None or x_is_u:
            xv = uv
        _raise_if_not_dataarray(xv)
        _raise_if_not_dataarray(uv)
        if axis is MISSING and dim is MISSING:
            axis = 0
        mom_dims = (xmom_dim, umom_dim)
        sampler = cm.factory_sampler(sampler, data=xv, dim=dim, axis=axis, rep_dim=rep_dim, parallel=parallel)
        dxduave = cm.wrap_resample_vals(
            xv, uv, weight=weight, sampler=sampler, mom=(1, order), axis=axis, dim=dim, mom_dims=mom_dims,
            rep_dim=rep_dim, parallel=parallel, **kwargs,
        )
        out = cls(dxduave=dxduave, xmom_dim=xmom_dim, umom_dim=umom_dim, rec_dim=rep_dim, deriv_dim=deriv_dim,
                  central=central, meta=meta, x_is_u=x_is_u)
        return out.set_params(
            meta=out.meta.resample(
                data=out, meta_kws=meta_kws, sampler=sampler, weight=weight, mom=(1, order), axis=axis, dim=dim,
                mom_dims=mom_dims, rep_dim=rep_dim, **kwargs,
            )
        )

    @classmethod
    def from_ave_raw(cls, u, xu, weight=None, rec_dim="rec", xmom_dim="xmom", umom_dim="umom", deriv_dim=None,
                     central=False, meta=None, x_is_u=False):
        """From ``u[n] = <u**n>`` and ``xu[n] = <x u**n>``."""
        _raise_if_not_dataarray(u)
        if xu is None or x_is_u:
            raw = u
            if weight is not None:
                raw = cm.assign_moment(raw, weight=weight, mom_dims=umom_dim, mom_ndim=1, copy=False)
            raw = raw.transpose(..., umom_dim)
        else:
            _raise_if_not_dataarray(xu)
            uu = u
            extra = [d for d in xu.dims if d not in u.dims]
            if extra:  # broadcast <u**n> over the value dims of xu
                ones = DataArray(np.ones([xu.sizes[d] for d in extra]), dims=extra)
                uu = u * ones
            raw = concat((uu.transpose(*xu.dims), xu), dim=xmom_dim)
            if weight is not None:
                raw = cm.assign_moment(raw, weight=weight, mom_dims=(xmom_dim, umom_dim), copy=False)
            raw = raw.transpose(..., xmom_dim, umom_dim)
        return cls.from_raw(raw=raw, xmom_dim=xmom_dim, umom_dim=umom_dim, deriv_dim=deriv_dim, rec_dim=rec_dim,
                            central=central, meta=meta, x_is_u=x_is_u)

    @classmethod
    def from_ave_central(cls, du, dxdu, weight=None, xave=None, uave=None, axis=-1, umom_axis=None,
                         xumom_axis=None, rec_dim="rec", xmom_dim="xmom", umom_dim="umom", deriv_dim=None,
                         central=False, dtype=None, dims=None, attrs=None, coords=None, name=None, meta=None,
                         x_is_u=False):
        """From ``du[n] = <du**n>`` (``du[0]`` weight, ``du[1]`` <u>) and ``dxdu[n] = <dx du**n>`` (``dxdu[0]`` <x>)."""
        labelled = is_dataarray(du)
        if dxdu is None or x_is_u:
            if labelled:
                d = du.transpose(..., umom_dim)
                dxdu = d.isel({umom_dim: slice(1, None)})
                du = d.isel({umom_dim: slice(None, -1)})
            else:
                du = np.swapaxes(np.asarray(du), umom_axis if umom_axis is not None else axis, -1)
                dxdu, du = du[..., 1:], du[..., :-1]
                umom_axis = xumom_axis = -1
        if (xave is None or x_is_u) and uave is not None:
            xave = uave
        if labelled:
            dxv = dxdu.transpose(..., umom_dim)
            duv = du.transpose(..., umom_dim)
            lead = dxv.dims[:-1]
            a = np.asarray(dxv.values)
            b = np.broadcast_to(_expand_like(duv, dxv), a.shape) if duv.dims != dxv.dims else np.asarray(duv.values)
        else:
            ua = axis if umom_axis is None else umom_axis
            xa = axis if xumom_axis is None else xumom_axis
            b = np.swapaxes(np.asarray(du), ua, -1)
            a = np.swapaxes(np.asarray(dxdu), xa, -1)
            b = np.broadcast_to(b, a.shape) if b.shape != a.shape else b
            lead = tuple(dims) if dims is not None else tuple(f"dim_{i}" for i in range(a.ndim - 1))
        n = min(a.shape[-1], b.shape[-1])
        data = np.empty((*a.shape[:-1], 2, n), dtype=dtype or a.dtype)
        data[..., 0, :] = b[..., :n]
        data[..., 1, :] = a[..., :n]
        for val, (i, j) in ((weight, (0, 0)), (xave, (1, 0)), (uave, (0, 1))):
            if val is not None:
                v = cm._values(val)
                if is_dataarray(val) and val.ndim:
                    v = np.asarray(val.transpose(*[d for d in lead if d in val.dims]).values)
                data[..., i, j] = v
        obj = DataArray(data, dims=(*lead, xmom_dim, umom_dim), name=name, attrs=attrs)
        dxduave = CentralMomentsData(obj, mom_ndim=2, mom_dims=(xmom_dim, umom_dim))
        return cls(dxduave=dxduave, xmom_dim=xmom_dim, umom_dim=umom_dim, rec_dim=rec_dim, deriv_dim=deriv_dim,
                   central=central, meta=meta, x_is_u=x_is_u)


def _expand_like(small: DataArray, big: DataArray) -> np.ndarray:
    """Values of ``small`` broadcast-ready against ``big`` (same dim order, size-1 where absent)."""
    perm = [d for d in big.dims if d in small.dims]
    v = np.asarray(small.transpose(*perm).values)
    shape = [small.sizes[d] if d in small.dims else 1 for d in big.dims]
    return v.reshape(shape)


def _convert_dxduave(dxduave, self_):
    if dxduave is not None:
        return dxduave
    if self_.order is None or self_.order <= 0:
        msg = "must pass order if calculating dxduave"
        raise ValueError(msg)
    if self_.x_is_u and self_.xv is self_.uv:
        # x == u: one pass over u to order+1 instead of a comoment pass over (u, u)
        return cm.wrap_reduce_vals(
            self_.uv, weight=self_.weight, dim=self_.rec_dim, mom=self_.order + 1, mom_dims=self_.umom_dim,
            **self_.from_vals_kws,
        ).moments_to_comoments(mom_dims_out=(self_.xmom_dim, self_.umom_dim), mom=(1, self_.order))
    return cm.wrap_reduce_vals(
        self_.xv, self_.uv, weight=self_.weight, dim=self_.rec_dim, mom=(1, self_.order),
        mom_dims=(self_.xmom_dim, self_.umom_dim), **self_.from_vals_kws,
    )


@attrs.define
class DataCentralMomentsVals(DataCentralMomentsBase):
    """Keeps the samples ``(uv, xv, weight)`` next to their comoments, so that bootstrap
    resampling (K2) can reweight the original samples."""

    uv: DataArray = field(validator=attv.instance_of(DataArray))
    xv: DataArray = field(validator=attv.instance_of(DataArray))
    order: int | None = field(kw_only=True, default=None, validator=attv.optional(attv.instance_of(int)))
    weight: DataArray | None = field(kw_only=True, validator=attv.optional(attv.instance_of(DataArray)), default=None)
    from_vals_kws: dict = field(kw_only=True, converter=convert_mapping_or_none_to_dict, default=None)
    dxduave: CentralMomentsData = field(
        kw_only=True,
        converter=attrs.Converter(_convert_dxduave, takes_self=True),
        validator=attv.instance_of(CentralMomentsData),
        default=None,
    )

    @classmethod
    def from_vals(cls, xv, uv, order, weight=None, rec_dim="rec", umom_dim="umom", xmom_dim="xmom", deriv_dim=None,
                  central=False, from_vals_kws=None, meta=None, x_is_u=False):
        """Reduce samples to comoments of order ``order`` (K1) and keep the samples."""
        return cls(
            uv=uv, xv=uv if xv is None else xv, order=order, weight=weight, rec_dim=rec_dim, umom_dim=umom_dim,
            xmom_dim=xmom_dim, deriv_dim=deriv_dim, central=central, from_vals_kws=from_vals_kws, meta=meta,
            x_is_u=x_is_u,
        )

    def __len__(self) -> int:
        return self.uv.sizes[self.rec_dim]

    def resample(self, sampler, dim=MISSING, axis=MISSING, rep_dim="rep", parallel=None, meta_kws=None, **kwargs):
        """Bootstrap: comoments of the samples reweighted by each replicate's frequencies (K2)."""
        if dim is MISSING and axis is MISSING:
            dim = self.rec_dim
        sampler = cm.factory_sampler(sampler, data=self.xv, dim=dim, axis=axis, rep_dim=rep_dim, parallel=parallel)
        kws = {"sampler": sampler, "parallel": parallel, "axis": axis, "dim": dim, "rep_dim": rep_dim, **kwargs}
        meta = self.meta.resample(data=self, meta_kws=meta_kws, **kws)
        dxduave = cm.wrap_resample_vals(
            self.xv, self.uv, weight=self.weight, mom=(1, self.order), mom_dims=(self.xmom_dim, self.umom_dim), **kws
        )
        dxduave = dxduave.transpose(rep_dim, ...)
        return self.new_like(dxduave=dxduave, rec_dim=rep_dim, meta=meta)

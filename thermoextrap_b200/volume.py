"""First-order volume extrapolation from virial data (mirror of ``thermoextrap.volume``).

With ``W = beta * virial`` playing the role of the energy (reference volume.py:39-80):

    d0 = <x>
    d1 = (-<x><W> + <x W> + <dx/dq . q>) / (volume * ndim)

Only ``DataValues``-style data and ``order = 1`` are supported (higher orders need force
derivatives).
"""

from __future__ import annotations

from functools import lru_cache

import attrs
import numpy as np
import sympy as sp
from attrs import field
from attrs import validators as attv

from . import backend as cm
from .core import cached
from .core.xr import DataArray
from .core.xrutils import xrwrap_xv
from .data import DataCallbackABC, DataValues
from .models import Derivatives, ExtrapModel


class VolumeDerivFuncs:
    """``self[order]`` = python callable ``f(W, xW, dxdq, volume, ndim=1)`` (orders 0 and 1)."""

    def __getitem__(self, order):
        if order > 1:
            raise ValueError(
                "Volume derivatives cannot go past 1st order"
                + " and received %i" % order
                + "\n(because would need derivatives of forces)"
            )
        return self.create_deriv_func(order)

    @staticmethod
    def create_deriv_func(order):
        def func(W, xW, dxdq, volume, ndim=1):  # noqa: N803
            if order == 0:
                return xW[0]
            return (-xW[0] * W[1] + xW[1] + dxdq) / (volume * ndim)

        return func


class _VolumeExprs:
    """The same two expressions in sympy form, so they run through the device evaluator."""

    W, xW = sp.IndexedBase("W"), sp.IndexedBase("xW")
    dxdq, volume, ndim = sp.symbols("dxdq volume ndim")
    args = (W, xW, dxdq, volume, ndim)

    def __getitem__(self, order):
        if order > 1:
            return VolumeDerivFuncs()[order]
        if order == 0:
            return self.xW[0]
        return (-self.xW[0] * self.W[1] + self.xW[1] + self.dxdq) / (self.volume * self.ndim)


@lru_cache(5)
def factory_derivatives():
    """Derivatives object for the volume expansion."""
    e = _VolumeExprs()
    return Derivatives(VolumeDerivFuncs(), exprs=e, args=e.args)


@attrs.define
class VolumeDataCallback(DataCallbackABC):
    """Adds ``(<dxdq>, volume, ndim)`` to the derivative arguments.

    ``dxdqv[rec, *val]`` holds ``sum_i dx/dq_i q_i`` per sample.
    """

    volume: float = field(validator=attv.instance_of(float))
    dxdqv: DataArray = field(validator=attv.instance_of(DataArray))
    ndim: int = field(default=3, validator=attv.instance_of(int))
    # set on bootstrap resamples: (sampler, rep_dim); the gathered samples are never materialised
    _boot: tuple | None = field(default=None, repr=False, alias="boot")
    _cache: dict = field(init=False, repr=False, factory=dict)

    def check(self, data) -> None:
        pass

    @cached.meth
    def dxdq(self, rec_dim):
        """Mean of ``dxdqv`` along ``rec_dim`` (per replicate when resampled), on the GPU."""
        x = self.dxdqv
        zeros = DataArray(np.zeros(x.sizes[rec_dim]), dims=(rec_dim,))
        if self._boot is None:
            c = cm.wrap_reduce_vals(x, zeros, mom=(1, 1), dim=rec_dim)
        else:
            sampler, rep_dim = self._boot
            c = cm.wrap_resample_vals(x, zeros, mom=(1, 1), sampler=sampler, dim=rec_dim, rep_dim=rep_dim)
            c = c.transpose(rep_dim, ...)
        return c.select_moment("xave")

    def resample(self, data, meta_kws, sampler, **kws):  # noqa: ARG002
        if not isinstance(data, DataValues):
            msg = "resampling only possible with DataValues style."
            raise NotImplementedError(msg)
        return self.new_like(boot=(sampler, kws.get("rep_dim", "rep")))

    def derivs_args(self, data, derivs_args):
        return (*tuple(derivs_args), self.dxdq(data.rec_dim), self.volume, self.ndim)


def factory_extrapmodel(volume, uv, xv, dxdqv, ndim=3, order=1, alpha_name="volume", rec_dim="rec", val_dims="val",
                        rep_dim="rep", **kws):
    """Volume-extrapolation model.  ``uv`` is the temperature-scaled virial ``beta * virial``."""
    if order != 1:
        msg = "only order=1 is supported"
        raise ValueError(msg)
    dxdqv = xrwrap_xv(dxdqv, rec_dim=rec_dim, rep_dim=rep_dim, deriv_dim=None, val_dims=val_dims)
    meta = VolumeDataCallback(volume=volume, dxdqv=dxdqv, ndim=ndim)
    data = DataValues.from_vals(uv=uv, xv=xv, order=order, meta=meta, rec_dim=rec_dim, deriv_dim=None, **kws)
    return ExtrapModel(alpha0=volume, data=data, derivatives=factory_derivatives(), order=order, minus_log=False,
                       alpha_name=alpha_name)


@attrs.define
class VolumeMomentsCallback(DataCallbackABC):
    """``(<dxdq>, volume, ndim)`` for data that was reduced while streaming (no samples are kept, so there is
    nothing to resample: the same restriction the reference puts on non-``DataValues`` data, volume.py:121-124)."""

    volume: float = field(validator=attv.instance_of(float))
    dxdq_mean: DataArray = field(validator=attv.instance_of(DataArray))
    ndim: int = field(default=3, validator=attv.instance_of(int))

    def check(self, data) -> None:
        pass

    def resample(self, data, meta_kws, sampler, **kws):  # noqa: ARG002
        msg = "resampling only possible with DataValues style."
        raise NotImplementedError(msg)

    def derivs_args(self, data, derivs_args):  # noqa: ARG002
        return (*tuple(derivs_args), self.dxdq_mean, self.volume, self.ndim)


def factory_extrapmodel_stream(volume, chunks, ndim=3, order=1, alpha_name="volume", val_dims="val"):
    """:func:`factory_extrapmodel` for a sample set that is STREAMED: ``chunks`` yields host triples
    ``(xv[n(, *val)], dxdqv[n(, *val)], uv[n])`` (``uv = beta * virial``); they go through the double-buffered
    H2D ring of :mod:`.streaming` and only the order-1 moments stay.  Same model, same ``predict``."""
    from . import streaming
    from .data import DataCentralMoments

    if order != 1:
        msg = "only order=1 is supported"
        raise ValueError(msg)
    com, dq, _n = streaming.volume_moments_stream(chunks, order)
    com, dq = cm._host(com), cm._host(dq)
    vd = () if com.ndim == 2 else ((val_dims,) if isinstance(val_dims, str) else tuple(val_dims))
    data = DataCentralMoments.from_data(DataArray(com, dims=(*vd, "xmom", "umom")), central=False,
                                        meta=VolumeMomentsCallback(volume=float(volume), dxdq_mean=DataArray(dq, dims=vd),
                                                                   ndim=ndim))
    return ExtrapModel(alpha0=volume, data=data, derivatives=factory_derivatives(), order=order, minus_log=False,
                       alpha_name=alpha_name)

"""Volume expansion for the 1-D ideal-gas test system (mirror of ``thermoextrap.volume_idealgas``).

    d0 = <x>
    d1 = (<x W> - <x><W>) / refV + <x> / refV        (1-D system: no factor 3)
"""

from __future__ import annotations

from functools import lru_cache

import sympy as sp

from .models import Derivatives, ExtrapModel


class VolumeDerivFuncsIG:
    """``self[order]`` = python callable ``f(W, xW)`` for orders 0 and 1 at reference volume ``refV``."""

    def __init__(self, refV=1.0) -> None:  # noqa: N803
        self.refV = refV

    def __getitem__(self, order):
        if order > 1:
            raise ValueError(
                "Volume derivatives cannot go past 1st order"
                + " and received %i" % order
                + "\n(because would need derivatives of forces)"
            )
        return self.create_deriv_func(order)

    def create_deriv_func(self, order):
        def func(W, xW):  # noqa: N803
            if order == 0:
                return xW[0]
            return (xW[1] - xW[0] * W[1]) / self.refV + xW[0] / self.refV

        return func


class _VolumeExprsIG:
    W, xW = sp.IndexedBase("W"), sp.IndexedBase("xW")
    args = (W, xW)

    def __init__(self, refV) -> None:  # noqa: N803
        self.refV = sp.Float(refV)

    def __getitem__(self, order):
        if order > 1:
            return VolumeDerivFuncsIG()[order]
        if order == 0:
            return self.xW[0]
        return (self.xW[1] - self.xW[0] * self.W[1]) / self.refV + self.xW[0] / self.refV


@lru_cache(5)
def factory_derivatives(refV=1.0):  # noqa: N803
    """Derivatives object for the ideal-gas volume expansion."""
    e = _VolumeExprsIG(refV)
    return Derivatives(VolumeDerivFuncsIG(refV=refV), exprs=e, args=e.args)


def factory_extrapmodel(volume, uv, xv, order=1, alpha_name="volume", **kws):
    """Volume extrapolation of the ideal-gas observable from samples ``(uv, xv)``."""
    if order != 1:
        msg = "only first order supported"
        raise ValueError(msg)
    from .data import factory_data_values

    data = factory_data_values(uv=uv, xv=xv, order=order, central=False, xalpha=False, **kws)
    return ExtrapModel(alpha0=volume, data=data, derivatives=factory_derivatives(refV=volume), order=order,
                       minus_log=False, alpha_name=alpha_name)


def factory_extrapmodel_data(volume, data, order=1, alpha_name="volume"):
    """Same, from an existing (raw-moment, order >= 1) data object."""
    if order != 1:
        raise ValueError
    if order > data.order:
        raise ValueError
    if data.central:
        raise ValueError
    if getattr(data, "deriv_dim", None) is not None:
        raise ValueError
    return ExtrapModel(alpha0=volume, data=data, derivatives=factory_derivatives(refV=volume), order=order,
                       minus_log=False, alpha_name=alpha_name)

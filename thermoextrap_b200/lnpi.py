r"""Beta expansion of the macrostate distribution :math:`\ln\Pi(N)` (mirror of ``thermoextrap.lnpi``).

With canonical averages :math:`\langle\cdot\rangle` at fixed N (reference lnpi.py:42-116):

    (ln Pi)'   = mu . N - <u>          (the N-independent grand-canonical term is dropped)
    (ln Pi)^(k+1) = -(<u>)^(k)         k >= 1

so the series needs ``lnPi0`` and ``mu . N`` as two extra derivative arguments, supplied by
:class:`lnPiDataCallback` through the data object's ``meta`` hook.
"""

from __future__ import annotations

import warnings
from functools import lru_cache

import attrs
import numpy as np
from attrs import field
from attrs import validators as attv

from . import beta as beta_xpan
from .core import cached
from .core.attrs_utils import convert_dims_to_tuple
from .core.sputils import get_default_symbol
from .core.xr import DataArray, dot
from .data import DataCallbackABC
from .models import Derivatives, ExtrapModel, SymSubs


def _lnpi_rule(name, energy_family, energy):
    """lnPi as one more family of the beta rule table: value ``lnPi0`` at beta0, first derivative ``mu.N - <u>``
    (``energy(F, beta)`` is ``<u>`` in the chosen moment form), arguments = those of ``<u>`` + (lnPi0, mudotN)."""
    mudotn = get_default_symbol("mudotN")
    return beta_xpan.MomentRule(
        name, "lnPi0", 1, index=lambda: (), fixed=lambda: None,
        ddbeta=lambda F, b: mudotn - energy(F, b), inherits=(energy_family,), extras=("mudotN",))


#: lnPi with its derivatives in central moments of u (reference lnpi.py:42-78) / in raw moments (lnpi.py:81-116)
lnPi_func_central = beta_xpan.register_rule(
    _lnpi_rule("lnPi_func_central", "u_func_central", lambda F, b: F.u_func_central(b)))
lnPi_func_raw = beta_xpan.register_rule(_lnpi_rule("lnPi_func_raw", "u_func", lambda F, b: F.u_func(b, 1)))


@lru_cache(5)
def factory_derivatives(name="lnPi", n=None, d=None, xalpha=False, central=False, expand=True, post_func=None):
    """Derivatives of lnPi (``name='lnPi'``) or of a general beta-expansion quantity."""
    if name != "lnPi":
        return beta_xpan.factory_derivatives(name=name, n=n, d=d, xalpha=xalpha, central=central,
                                             post_func=post_func, expand=expand)
    beta = get_default_symbol("beta")
    func = lnPi_func_central(beta) if central else lnPi_func_raw(beta)
    derivs = beta_xpan.SymDerivBeta(func=func, expand=expand, post_func=post_func)
    exprs = SymSubs(derivs, subs_all={derivs.beta: "None"}, expand=False, simplify=False)
    # the plain lnPi derivatives also exist as generated device functions (kernels.LNPI_FAMILIES): the u_ave family
    # one order lower, shifted, plus lnPi0 and mu.N
    family = (4 if central else 5) if post_func is None else None
    return Derivatives.from_sympy(exprs, args=derivs.args, family=family)


@attrs.define
class lnPiDataCallback(DataCallbackABC):
    """Carries ``lnPi0`` and ``mu`` and appends ``(lnPi0, mu . N)`` to the derivative arguments.

    ``dims_n``: particle-number dimension(s) of ``lnPi0``; ``dims_comp``: component dimension of
    ``mu``; ``ncoords[comp, *dims_n]`` defaults to the meshgrid of the ``dims_n`` coordinates.
    """

    lnPi0: DataArray = field(validator=attv.instance_of(DataArray))
    mu: DataArray = field(validator=attv.instance_of(DataArray))
    dims_n: tuple = field(converter=convert_dims_to_tuple)
    dims_comp: str = field()
    ncoords: DataArray = field(validator=attv.instance_of(DataArray))
    allow_resample: bool = field(default=False)
    _cache: dict = field(init=False, repr=False, factory=dict)

    def check(self, data) -> None:
        pass

    @ncoords.default
    def _set_default_ncoords(self):
        grids = np.meshgrid(*tuple(np.asarray(self.lnPi0[x].values) for x in self.dims_n), indexing="ij")
        return DataArray(np.array(grids), dims=(self.dims_comp, *self.dims_n))

    @property
    def lnPi0_ave(self):
        return self.lnPi0

    @cached.prop
    def mudotN(self):
        """``sum_comp mu[comp] * ncoords[comp, ...]``."""
        return dot(self.mu, self.ncoords, dims=self.dims_comp)

    def resample(self, data, meta_kws=None, **kws):
        """Ad-hoc resampling of the *average* lnPi0 (guarded, as in the reference lnpi.py:258-286)."""
        if not self.allow_resample:
            msg = (
                "Must set `self.allow_resample` to `True` to use resampling. "
                "Resampling here is handled in an ad-hoc way, and should be used with care."
            )
            raise ValueError(msg)
        warnings.warn(
            "'Correct' resampling of lnPi should be handled externally.  This resamples the average "
            "lnPi values; resampling collection matrices and rebuilding lnPi from them is recommended.",
            category=UserWarning,
            stacklevel=2,
        )
        from . import backend as cm

        sampler, dim, rep_dim = kws["sampler"], kws.get("dim"), kws.get("rep_dim", "rep")
        if not isinstance(dim, str) or dim not in self.lnPi0.dims:
            return self
        zeros = DataArray(np.zeros(self.lnPi0.sizes[dim]), dims=(dim,))
        c = cm.wrap_resample_vals(self.lnPi0, zeros, mom=(1, 1), sampler=sampler, dim=dim, rep_dim=rep_dim)
        return self.new_like(lnPi0=c.select_moment("xave").transpose(rep_dim, ...))

    def derivs_args(self, data, derivs_args):
        return (*tuple(derivs_args), self.lnPi0_ave, self.mudotN)


def factory_extrapmodel_lnPi(beta, data, *, central=None, order=None, alpha_name="beta", derivatives=None,
                             post_func=None, derivatives_kws=None):
    """Extrapolation model for lnPi; ``data`` must be ``x_is_u`` and carry a :class:`lnPiDataCallback`.

    ``order`` defaults to ``data.order + 1`` (the first derivative only needs ``<u>``).
    """
    if central is None:
        central = data.central
    if order is None:
        order = data.order + 1
    if central != data.central:
        raise ValueError
    if order > data.order + 1:
        raise ValueError
    if not data.x_is_u:
        raise ValueError
    if derivatives is None:
        derivatives = factory_derivatives(name="lnPi", central=central, post_func=post_func, **(derivatives_kws or {}))
    return ExtrapModel(alpha0=beta, data=data, derivatives=derivatives, order=order, alpha_name=alpha_name)

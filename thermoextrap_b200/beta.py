"""Inverse-temperature (beta) expansion (mirror of ``thermoextrap.beta``).

Derivative recurrences, with ``du = u - <u>``, ``dx = x - <x>`` and ``'`` = d/dbeta
(reference beta.py:32-266):

central moments
    <du^n>'        = -<du^(n+1)> + n <du^(n-1)> <du^2>
    <u>'           = -<du^2>
    <dx du^n>'     = -<dx du^(n+1)> + n <dx du^(n-1)> <du^2> + <dx du> <du^n>
                     (+ <dx^(d+1) du^n> when x depends on beta)
    <x>'           = -<dx du>   (+ <x^(1)> when x depends on beta)
raw moments
    <u^n>'         = -<u^(n+1)> + <u^n> <u>
    <x u^n>'       = -<x u^(n+1)> + <x u^n> <u>   (+ <x^(d+1) u^n>)
"""

from __future__ import annotations

from functools import lru_cache

from .core.sputils import get_default_indexed, get_default_symbol
from .models import Derivatives, ExtrapModel, PerturbModel, SymDerivBase, SymFuncBase, SymSubs

__all__ = [
    "SymDerivBeta",
    "factory_derivatives",
    "factory_extrapmodel",
    "factory_perturbmodel",
]


# ---------------------------------------------------------------- central moments
class du_func(SymFuncBase):
    r""":math:`\langle (u - \langle u \rangle)^n \rangle`; substitute ``beta -> 'None'`` to index ``du[n]``."""

    nargs = 2
    du = get_default_indexed("du")

    @classmethod
    def deriv_args(cls):
        return [cls.du]

    def fdiff(self, argindex=1):
        beta, n = self.args
        return n * du_func(beta, n - 1) * du_func(beta, 2) - du_func(beta, n + 1)

    @classmethod
    def eval(cls, beta, n):
        if n == 0:
            return 1
        if n == 1:
            return 0
        return cls.du[n] if beta is None else None


class u_func_central(SymFuncBase):
    r""":math:`\langle u \rangle` with central-moment derivatives."""

    nargs = 1
    u = get_default_symbol("u")

    @classmethod
    def deriv_args(cls):
        return [cls.u, *du_func.deriv_args()]

    def fdiff(self, argindex=1):
        (beta,) = self.args
        return -du_func(beta, 2)

    @classmethod
    def eval(cls, beta):
        return cls.u if beta is None else None


class dxdu_func_nobeta(SymFuncBase):
    r""":math:`\langle \delta x\, \delta u^n \rangle` for an observable that does not depend on beta."""

    nargs = 2
    dxdu = get_default_indexed("dxdu")

    @classmethod
    def deriv_args(cls):
        return [*du_func.deriv_args(), cls.dxdu]

    def fdiff(self, argindex=1):
        beta, n = self.args
        f = dxdu_func_nobeta
        return -f(beta, n + 1) + n * f(beta, n - 1) * du_func(beta, 2) + f(beta, 1) * du_func(beta, n)

    @classmethod
    def eval(cls, beta, n):
        if n == 0:
            return 0
        return cls.dxdu[n] if beta is None else None


class dxdu_func_beta(SymFuncBase):
    r""":math:`\langle \delta x^{(d)}\, \delta u^n \rangle` with :math:`x^{(d)} = d^d x / d\beta^d`."""

    nargs = 3
    dxdu = get_default_indexed("dxdu")

    @classmethod
    def deriv_args(cls):
        return [*du_func.deriv_args(), cls.dxdu]

    def fdiff(self, argindex=1):
        beta, n, d = self.args
        f = dxdu_func_beta
        return (
            -f(beta, n + 1, d)
            + n * f(beta, n - 1, d) * du_func(beta, 2)
            + f(beta, n, d + 1)
            + f(beta, 1, d) * du_func(beta, n)
        )

    @classmethod
    def eval(cls, beta, n, deriv):
        if n == 0:
            return 0
        return cls.dxdu[n, deriv] if beta is None else None


class x_func_central_nobeta(SymFuncBase):
    r""":math:`\langle x \rangle` (central form), x independent of beta."""

    nargs = 1
    x1_symbol = get_default_symbol("x1")

    @classmethod
    def deriv_args(cls):
        return [cls.x1_symbol, *dxdu_func_nobeta.deriv_args()]

    def fdiff(self, argindex=1):
        (beta,) = self.args
        return -dxdu_func_nobeta(beta, 1)

    @classmethod
    def eval(cls, beta):
        return cls.x1_symbol if beta is None else None


class x_func_central_beta(SymFuncBase):
    r""":math:`\langle x^{(d)}(\beta) \rangle` (central form), x depends on beta."""

    nargs = 2
    x1_indexed = get_default_indexed("x1")

    @classmethod
    def deriv_args(cls):
        return [cls.x1_indexed, *dxdu_func_beta.deriv_args()]

    def fdiff(self, argindex=1):
        beta, d = self.args
        return x_func_central_beta(beta, d + 1) - dxdu_func_beta(beta, 1, d)

    @classmethod
    def eval(cls, beta, deriv):
        return cls.x1_indexed[deriv] if beta is None else None


# ---------------------------------------------------------------- raw moments
class u_func(SymFuncBase):
    r""":math:`\langle u^n \rangle`."""

    nargs = 2
    u = get_default_indexed("u")

    @classmethod
    def deriv_args(cls):
        return [cls.u]

    def fdiff(self, argindex=1):
        beta, n = self.args
        return u_func(beta, n) * u_func(beta, 1) - u_func(beta, n + 1)

    @classmethod
    def eval(cls, beta, n):
        if n == 0:
            return 1
        return cls.u[n] if beta is None else None


class xu_func(SymFuncBase):
    r""":math:`\langle x u^n \rangle` (two arguments) or :math:`\langle x^{(d)} u^n \rangle` (three)."""

    nargs = (2, 3)
    xu = get_default_indexed("xu")

    @classmethod
    def deriv_args(cls):
        return [*u_func.deriv_args(), cls.xu]

    def fdiff(self, argindex=1):
        if len(self.args) == 2:
            beta, n = self.args
            return xu_func(beta, n) * u_func(beta, 1) - xu_func(beta, n + 1)
        beta, n, d = self.args
        return xu_func(beta, n, d + 1) + xu_func(beta, n, d) * u_func(beta, 1) - xu_func(beta, n + 1, d)

    @classmethod
    def eval(cls, beta, n, deriv=None):
        if beta is not None:
            return None
        return cls.xu[n] if deriv is None else cls.xu[n, deriv]


class SymDerivBeta(SymDerivBase):
    r"""Symbolic :math:`d^n \langle\cdot\rangle / d\beta^n` for the quantities below."""

    beta = get_default_symbol("beta")

    @classmethod
    def x_ave(cls, xalpha=False, central=None, expand=True, post_func=None):
        r""":math:`\langle x \rangle` (``central`` / raw, ``xalpha``: x depends on beta)."""
        if central:
            func = x_func_central_beta(cls.beta, 0) if xalpha else x_func_central_nobeta(cls.beta)
        else:
            func = xu_func(cls.beta, 0, 0) if xalpha else xu_func(cls.beta, 0)
        return cls(func=func, expand=expand, post_func=post_func)

    @classmethod
    def u_ave(cls, central=None, expand=True, post_func=None):
        r""":math:`\langle u \rangle`."""
        func = u_func_central(cls.beta) if central else u_func(cls.beta, 1)
        return cls(func=func, expand=expand, post_func=post_func)

    @classmethod
    def dun_ave(cls, n, expand=True, post_func=None, central=None):
        r""":math:`\langle \delta u^n \rangle` (central only, ``n > 1``)."""
        if central is not None and not central:
            msg = f"{central=} must be None or evaluate to True"
            raise ValueError(msg)
        if (n := int(n)) <= 1:
            msg = f"{n=} must be > 1."
            raise ValueError(msg)
        # same argument list as u_ave (includes the symbol u) so the two are interchangeable
        return cls(func=du_func(cls.beta, n), args=u_func_central.deriv_args(), expand=expand, post_func=post_func)

    @classmethod
    def dxdun_ave(cls, n, xalpha=False, expand=True, post_func=None, d=None, central=None):
        r""":math:`\langle \delta x^{(d)} \delta u^n \rangle` (central only, ``n > 0``; ``d`` needed if ``xalpha``)."""
        if central is not None and not central:
            msg = f"{central=} nust be `None` or evaluate to `True`"
            raise ValueError(msg)
        if (n := int(n)) <= 0:
            msg = f"{n=} must be positive integer."
            raise ValueError(msg)
        if xalpha:
            if not isinstance(d, int):
                raise TypeError
            func, args = dxdu_func_beta(cls.beta, n, d), x_func_central_beta.deriv_args()
        else:
            func, args = dxdu_func_nobeta(cls.beta, n), x_func_central_nobeta.deriv_args()
        return cls(func=func, args=args, expand=expand, post_func=post_func)

    @classmethod
    def un_ave(cls, n, expand=True, post_func=None, central=None):
        r""":math:`\langle u^n \rangle` (raw only, ``n >= 1``)."""
        if central is not None and central:
            msg = f"{central=} must be `None` or evaluate to False"
            raise ValueError(msg)
        if (n := int(n)) < 1:
            msg = f"{n=} must be >=1."
            raise ValueError(msg)
        return cls(func=u_func(cls.beta, n), expand=expand, post_func=post_func)

    @classmethod
    def xun_ave(cls, n, d=None, xalpha=False, expand=True, post_func=None, central=None):
        r""":math:`\langle x^{(d)} u^n \rangle` (raw only, ``n >= 0``)."""
        if central is not None and central:
            msg = f"{central=} must be `None` or False"
            raise ValueError(msg)
        if (n := int(n)) < 0:
            msg = f"{n=} must be >= 0"
            raise ValueError(msg)
        if xalpha:
            if not isinstance(d, int):
                raise TypeError
            if d < 0:
                raise ValueError
            func = xu_func(cls.beta, n, d)
        else:
            func = xu_func(cls.beta, n)
        return cls(func=func, expand=expand, post_func=post_func)

    @classmethod
    def from_name(cls, name, xalpha=False, central=None, expand=True, post_func=None, n=None, d=None):
        """Constructor by name: x_ave, u_ave, dun_ave, dxdun_ave, un_ave, xun_ave."""
        func = getattr(cls, name, None)
        if func is None:
            msg = f"{name} not found"
            raise ValueError(msg)
        kws = {"expand": expand, "post_func": post_func, "central": central}
        if name == "x_ave":
            kws.update(xalpha=xalpha)
        elif name in {"dun_ave", "un_ave"}:
            kws.update(n=n)
        elif name in {"dxdun_ave", "xun_ave"}:
            kws.update(n=n, xalpha=xalpha, d=d)
        return func(**kws)


@lru_cache(5)
def factory_derivatives(name="x_ave", n=None, d=None, xalpha=False, central=None, post_func=None, expand=True):
    """:class:`~thermoextrap_b200.models.Derivatives` for a beta expansion of the named quantity."""
    derivs = SymDerivBeta.from_name(name=name, n=n, d=d, xalpha=xalpha, central=central, post_func=post_func,
                                    expand=expand)
    exprs = SymSubs(derivs, subs_all={derivs.beta: "None"}, expand=False, simplify=False)
    # the plain beta derivatives of <x> / <u> also exist as generated device functions (kernels.BETA_FAMILIES)
    family = None
    if name in {"x_ave", "u_ave"} and not xalpha and post_func is None and n is None and d is None:
        family = {("x_ave", True): 0, ("u_ave", True): 1, ("x_ave", False): 2, ("u_ave", False): 3}[(name, bool(central))]
    return Derivatives.from_sympy(exprs, args=derivs.args, family=family)


def factory_extrapmodel(beta, data, *, name="x_ave", n=None, d=None, xalpha=None, central=None, order=None,
                        alpha_name="beta", derivatives=None, post_func=None, derivatives_kws=None):
    """Beta-extrapolation model; ``order``, ``xalpha`` and ``central`` default to those of ``data``."""
    if xalpha is None:
        xalpha = data.xalpha
    if central is None:
        central = data.central
    if order is None:
        order = data.order
    if xalpha != data.xalpha:
        msg = f"{xalpha=} must equal {data.xalpha=}"
        raise ValueError(msg)
    if central != data.central:
        msg = f"{central=} must equal {data.central=}"
        raise ValueError(msg)
    if order > data.order:
        msg = f"{order=} must be <= {data.order=}"
        raise ValueError(msg)
    if derivatives is None:
        if name in {"u_ave", "un_ave", "dun_ave"} and not data.x_is_u:
            msg = "if name in [u_ave, un_ave, dun_ave] must have data.x_is_u"
            raise ValueError(msg)
        derivatives = factory_derivatives(
            name=name, n=n, d=d, xalpha=xalpha, central=central, post_func=post_func, **(derivatives_kws or {})
        )
    return ExtrapModel(alpha0=beta, data=data, derivatives=derivatives, order=order, alpha_name=alpha_name)


def factory_perturbmodel(beta, uv, xv, alpha_name="beta", **kws):
    """Perturbation (reweighting) model at reference ``beta``."""
    from .data import DataValues

    data = DataValues.from_vals(xv=xv, uv=uv, order=0, **kws)
    return PerturbModel(alpha0=beta, data=data, alpha_name=alpha_name)

"""Inverse-temperature (beta) expansion -- the public surface of ``thermoextrap.beta``, built from a rule table.

The reference writes one ``sympy.Function`` subclass by hand per moment family (beta.py:32-266).  Here a moment
family is DATA -- a :class:`MomentRule`: which symbol it turns into once ``beta -> None`` is substituted, which
boundary values are fixed, and its d/dbeta recurrence as a function of the *namespace of all families* -- and
:func:`_function_class` manufactures the ``sympy.Function`` class (``eval`` / ``fdiff`` / ``deriv_args``) from it.
The named constructors of :class:`SymDerivBeta` are a table too (``_QUANTITIES``).  Same class names, call
signatures, argument order and error types as the reference, so ``SymDerivBeta.x_ave(...)``, ``du_func(beta, n)`` or
``factory_extrapmodel(beta, data)`` behave identically; ``tests/golden/exprs.json`` (strings produced by the
reference's own classes) pins every expression, term order included.  The four numeric families also exist as
generated device functions (``csrc/gen_beta_derivs.py``), which the models use instead of these sympy objects.

With ``du = u - <u>``, ``dx = x - <x>``, ``'`` = d/dbeta and ``x^(d) = d^d x / dbeta^d``:

    <du^n>'          = n <du^(n-1)> <du^2> - <du^(n+1)>                                   <du^0> = 1, <du^1> = 0
    <u>'             = -<du^2>
    <dx^(d) du^n>'   = n <dx^(d) du^(n-1)> <du^2> + <dx^(d) du> <du^n> - <dx^(d) du^(n+1)>  (+ <dx^(d+1) du^n>)
    <x^(d)>'         = -<dx^(d) du>                                                        (+ <x^(d+1)>)
    <u^n>'           = <u^n> <u> - <u^(n+1)>                                               <u^0> = 1
    <x^(d) u^n>'     = <x^(d) u^n> <u> - <x^(d) u^(n+1)>                                   (+ <x^(d+1) u^n>)
"""

from __future__ import annotations

from dataclasses import dataclass
from functools import lru_cache
from types import SimpleNamespace
from typing import Callable

from .core.sputils import get_default_indexed, get_default_symbol
from .models import Derivatives, ExtrapModel, PerturbModel, SymDerivBase, SymFuncBase, SymSubs

__all__ = [
    "SymDerivBeta",
    "factory_derivatives",
    "factory_extrapmodel",
    "factory_perturbmodel",
]


@dataclass(frozen=True)
class MomentRule:
    """One moment family as data.

    ``index(*args)`` maps the function's non-beta arguments to the subscript of ``symbol`` (``()``: plain Symbol);
    ``fixed(*args)`` returns a boundary value or None; ``ddbeta(F, beta, *args)`` is the recurrence, written against the
    namespace ``F`` of all families; ``inherits`` names the families whose symbols precede this one's in ``deriv_args``.
    """

    name: str
    symbol: str
    nargs: object
    index: Callable
    fixed: Callable
    ddbeta: Callable
    inherits: tuple = ()
    attr: str | None = None   # class attribute holding the sympy symbol (the reference's spelling)
    prepend: bool = False     # own symbol before the inherited ones in deriv_args
    extras: tuple = ()        # further plain symbols the recurrence uses, appended to deriv_args (lnpi.py)


F = SimpleNamespace()   # name -> sympy.Function class, filled below (the recurrences look each other up here)


def _function_class(rule: MomentRule):
    indexed = rule.index(*([0] * (rule.nargs[0] if isinstance(rule.nargs, tuple) else rule.nargs - 1))) != ()
    sym = get_default_indexed(rule.symbol) if indexed else get_default_symbol(rule.symbol)

    extras = [get_default_symbol(e) for e in rule.extras]

    def deriv_args(cls):
        inherited = [a for fam in rule.inherits for a in getattr(F, fam).deriv_args()]
        return [sym, *inherited, *extras] if rule.prepend else [*inherited, sym, *extras]

    def fdiff(self, argindex=1):
        beta, *rest = self.args
        return rule.ddbeta(F, beta, *rest)

    def eval_(cls, beta, *rest):
        value = rule.fixed(*rest)
        if value is not None:
            return value
        if beta is not None:
            return None
        sub = rule.index(*rest)
        return sym[sub if len(sub) > 1 else sub[0]] if sub else sym

    body = {
        "nargs": rule.nargs, "__doc__": f"sympy function of the {rule.symbol!r} moments (rule table, beta.py)",
        "deriv_args": classmethod(deriv_args), "fdiff": fdiff, "eval": classmethod(eval_), "rule": rule,
        rule.attr or rule.symbol: sym, **{str(e): e for e in extras},
    }
    return type(rule.name, (SymFuncBase,), body)


def register_rule(rule: MomentRule):
    """Add a family to the namespace ``F`` and return its ``sympy.Function`` class."""
    setattr(F, rule.name, _function_class(rule))
    return getattr(F, rule.name)


def _n_only(n):
    return (n,)


_RULES = (
    # ---- central moments
    MomentRule("du_func", "du", 2, _n_only, lambda n: {0: 1, 1: 0}.get(n),
               lambda F, b, n: n * F.du_func(b, n - 1) * F.du_func(b, 2) - F.du_func(b, n + 1)),
    MomentRule("u_func_central", "u", 1, lambda: (), lambda: None,
               lambda F, b: -F.du_func(b, 2), inherits=("du_func",), prepend=True),
    MomentRule("dxdu_func_nobeta", "dxdu", 2, _n_only, lambda n: 0 if n == 0 else None,
               lambda F, b, n: (-F.dxdu_func_nobeta(b, n + 1) + n * F.dxdu_func_nobeta(b, n - 1) * F.du_func(b, 2)
                                + F.dxdu_func_nobeta(b, 1) * F.du_func(b, n)),
               inherits=("du_func",)),
    MomentRule("dxdu_func_beta", "dxdu", 3, lambda n, d: (n, d), lambda n, d: 0 if n == 0 else None,
               lambda F, b, n, d: (-F.dxdu_func_beta(b, n + 1, d) + n * F.dxdu_func_beta(b, n - 1, d) * F.du_func(b, 2)
                                   + F.dxdu_func_beta(b, n, d + 1) + F.dxdu_func_beta(b, 1, d) * F.du_func(b, n)),
               inherits=("du_func",)),
    MomentRule("x_func_central_nobeta", "x1", 1, lambda: (), lambda: None,
               lambda F, b: -F.dxdu_func_nobeta(b, 1), inherits=("dxdu_func_nobeta",), attr="x1_symbol", prepend=True),
    MomentRule("x_func_central_beta", "x1", 2, lambda d: (d,), lambda d: None,
               lambda F, b, d: F.x_func_central_beta(b, d + 1) - F.dxdu_func_beta(b, 1, d),
               inherits=("dxdu_func_beta",), attr="x1_indexed", prepend=True),
    # ---- raw moments
    MomentRule("u_func", "u", 2, _n_only, lambda n: 1 if n == 0 else None,
               lambda F, b, n: F.u_func(b, n) * F.u_func(b, 1) - F.u_func(b, n + 1)),
    MomentRule("xu_func", "xu", (2, 3), lambda n, d=None: (n,) if d is None else (n, d), lambda n, d=None: None,
               lambda F, b, n, d=None: (F.xu_func(b, n) * F.u_func(b, 1) - F.xu_func(b, n + 1) if d is None else
                                        F.xu_func(b, n, d + 1) + F.xu_func(b, n, d) * F.u_func(b, 1)
                                        - F.xu_func(b, n + 1, d)),
               inherits=("u_func",)),
)
for _rule in _RULES:
    globals()[_rule.name] = register_rule(_rule)   # importable by the reference's names (lnpi.py uses two of them)


# ---------------------------------------------------------------- the quantities SymDerivBeta can expand
def _positive(n, least, message):
    n = int(n)
    if n < least:
        raise ValueError(message.format(n=n))
    return n


def _need_form(central, want_central, message):
    if central is not None and bool(central) != want_central:
        raise ValueError(message.format(central=central))


def _q_x_ave(beta, *, xalpha=False, central=None, **_):
    if central:
        return (F.x_func_central_beta(beta, 0) if xalpha else F.x_func_central_nobeta(beta)), None
    return (F.xu_func(beta, 0, 0) if xalpha else F.xu_func(beta, 0)), None


def _q_u_ave(beta, *, central=None, **_):
    return (F.u_func_central(beta) if central else F.u_func(beta, 1)), None


def _q_dun_ave(beta, *, n, central=None, **_):
    _need_form(central, True, "central={central} must be None or evaluate to True")
    n = _positive(n, 2, "n={n} must be > 1.")
    # the argument list of u_ave (it includes the symbol u), so the two quantities are interchangeable
    return F.du_func(beta, n), F.u_func_central.deriv_args()


def _q_dxdun_ave(beta, *, n, xalpha=False, d=None, central=None, **_):
    _need_form(central, True, "central={central} nust be `None` or evaluate to `True`")
    n = _positive(n, 1, "n={n} must be positive integer.")
    if not xalpha:
        return F.dxdu_func_nobeta(beta, n), F.x_func_central_nobeta.deriv_args()
    if not isinstance(d, int):
        raise TypeError
    return F.dxdu_func_beta(beta, n, d), F.x_func_central_beta.deriv_args()


def _q_un_ave(beta, *, n, central=None, **_):
    _need_form(central, False, "central={central} must be `None` or evaluate to False")
    return F.u_func(beta, _positive(n, 1, "n={n} must be >=1.")), None


def _q_xun_ave(beta, *, n, d=None, xalpha=False, central=None, **_):
    _need_form(central, False, "central={central} must be `None` or False")
    n = _positive(n, 0, "n={n} must be >= 0")
    if not xalpha:
        return F.xu_func(beta, n), None
    if not isinstance(d, int):
        raise TypeError
    if d < 0:
        raise ValueError
    return F.xu_func(beta, n, d), None


#: quantity name -> (builder, keywords ``from_name`` forwards, positional order of the named constructor)
_QUANTITIES = {
    "x_ave": (_q_x_ave, ("xalpha",), ("xalpha", "central", "expand", "post_func")),
    "u_ave": (_q_u_ave, (), ("central", "expand", "post_func")),
    "dun_ave": (_q_dun_ave, ("n",), ("n", "expand", "post_func", "central")),
    "dxdun_ave": (_q_dxdun_ave, ("n", "xalpha", "d"), ("n", "xalpha", "expand", "post_func", "d", "central")),
    "un_ave": (_q_un_ave, ("n",), ("n", "expand", "post_func", "central")),
    "xun_ave": (_q_xun_ave, ("n", "xalpha", "d"), ("n", "d", "xalpha", "expand", "post_func", "central")),
}


class SymDerivBeta(SymDerivBase):
    r"""Symbolic :math:`d^n \langle\cdot\rangle / d\beta^n`; one named constructor per entry of ``_QUANTITIES``
    (``x_ave``, ``u_ave``, ``dun_ave``, ``dxdun_ave``, ``un_ave``, ``xun_ave``) plus :meth:`from_name`."""

    beta = get_default_symbol("beta")

    @classmethod
    def _build(cls, name, expand=True, post_func=None, **kws):
        func, args = _QUANTITIES[name][0](cls.beta, **kws)
        return cls(func=func, args=args, expand=expand, post_func=post_func)

    @classmethod
    def from_name(cls, name, xalpha=False, central=None, expand=True, post_func=None, n=None, d=None):
        """Constructor by name; only the keywords that quantity takes are passed on."""
        if name not in _QUANTITIES:
            msg = f"{name} not found"
            raise ValueError(msg)
        given = {"xalpha": xalpha, "n": n, "d": d}
        kws = {k: given[k] for k in _QUANTITIES[name][1]}
        return getattr(cls, name)(expand=expand, post_func=post_func, central=central, **kws)


def _named_constructor(name):
    def make(cls, *args, **kws):
        order = _QUANTITIES[name][2]   # positional arguments in the reference's order
        if len(args) > len(order):
            msg = f"{name}() takes at most {len(order)} positional arguments"
            raise TypeError(msg)
        for key, val in zip(order, args):
            if key in kws:
                msg = f"{name}() got multiple values for argument {key!r}"
                raise TypeError(msg)
            kws[key] = val
        if "n" in order and "n" not in kws:
            msg = f"{name}() missing required argument 'n'"
            raise TypeError(msg)
        return cls._build(name, **kws)

    make.__name__ = name
    make.__doc__ = f"``SymDerivBeta`` for ``{name}``; arguments ({', '.join(_QUANTITIES[name][2])})."
    return classmethod(make)


for _name in _QUANTITIES:
    setattr(SymDerivBeta, _name, _named_constructor(_name))

#: (name, central) -> generated device-function family (``kernels.BETA_FAMILIES``)
_DEVICE_FAMILY = {("x_ave", True): 0, ("u_ave", True): 1, ("x_ave", False): 2, ("u_ave", False): 3}


@lru_cache(5)
def factory_derivatives(name="x_ave", n=None, d=None, xalpha=False, central=None, post_func=None, expand=True):
    """:class:`~thermoextrap_b200.models.Derivatives` for a beta expansion of the named quantity."""
    derivs = SymDerivBeta.from_name(name=name, n=n, d=d, xalpha=xalpha, central=central, post_func=post_func,
                                    expand=expand)
    exprs = SymSubs(derivs, subs_all={derivs.beta: "None"}, expand=False, simplify=False)
    plain = not xalpha and post_func is None and n is None and d is None
    family = _DEVICE_FAMILY.get((name, bool(central))) if plain else None
    return Derivatives.from_sympy(exprs, args=derivs.args, family=family)


def factory_extrapmodel(beta, data, *, name="x_ave", n=None, d=None, xalpha=None, central=None, order=None,
                        alpha_name="beta", derivatives=None, post_func=None, derivatives_kws=None):
    """Beta-extrapolation model; ``order``, ``xalpha`` and ``central`` default to those of ``data``."""
    xalpha = data.xalpha if xalpha is None else xalpha
    central = data.central if central is None else central
    order = data.order if order is None else order
    for label, mine, theirs in (("xalpha", xalpha, data.xalpha), ("central", central, data.central)):
        if mine != theirs:
            msg = f"{label}={mine} must equal data.{label}={theirs}"
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

    return PerturbModel(alpha0=beta, data=DataValues.from_vals(xv=xv, uv=uv, order=0, **kws), alpha_name=alpha_name)

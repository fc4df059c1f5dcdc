r"""Analytic 1-D ideal gas in a linear external field (mirror of ``thermoextrap.idealgas``).

Positions :math:`0 \le x \le L` (``vol``), :math:`U = \sum_i x_i`.  Single-particle
distribution :math:`p(x) = \beta e^{-\beta x} / (1 - e^{-\beta L})`, so
:math:`\langle x \rangle = 1/\beta - L / (e^{\beta L} - 1)`.  Used as the synthetic data
generator (host side, numpy) and as analytic ground truth for the extrapolation models.
"""

from __future__ import annotations

import math
from functools import lru_cache

import numpy as np
import sympy as sp

from .backend import random as _random

__all__ = [
    "dbeta_xave", "dbeta_xave_depend", "dbeta_xave_depend_minuslog", "dbeta_xave_minuslog", "dvol_xave",
    "generate_data", "u_prob", "u_sample", "x_ave", "x_beta_extrap", "x_beta_extrap_depend",
    "x_beta_extrap_depend_minuslog", "x_beta_extrap_minuslog", "x_cdf", "x_prob", "x_sample", "x_var",
    "x_vol_extrap",
]

beta_sym, vol_sym = sp.symbols("beta_sym vol_sym")
xave_sym = (1 / beta_sym) - vol_sym / (sp.exp(beta_sym * vol_sym) - 1)


def x_ave(beta, vol=1.0):
    """Average position at inverse temperature ``beta``."""
    return 1.0 / beta - vol / (np.exp(beta * vol) - 1.0)


def x_var(beta, vol=1.0):
    """Variance of the position."""
    return 1.0 / beta**2 - (vol**2 * np.exp(beta * vol) / ((np.exp(beta * vol) - 1) ** 2))


def x_prob(x, beta, vol=1.0):
    """Canonical probability density of a single particle's position."""
    return (beta * np.exp(-beta * x)) / (1.0 - np.exp(-beta * vol))


def u_prob(u, npart, beta, vol=1.0):
    """Large-N (normal) approximation of the potential-energy density."""
    u_ave = npart * x_ave(beta, vol)
    u_std = np.sqrt(npart * x_var(beta, vol))
    return np.exp(-0.5 * ((u - u_ave) / u_std) ** 2) / (u_std * np.sqrt(2 * np.pi))


def x_cdf(x, beta, vol=1.0):
    """Cumulative distribution of a single particle's position."""
    return (1.0 - np.exp(-beta * x)) / (1.0 - np.exp(-beta * vol))


def x_sample(shape, beta, vol=1.0, rng: np.random.Generator | None = None):
    """Positions drawn by inverting the cumulative distribution."""
    r = _random.validate_rng(rng).random(shape)
    return (-1.0 / beta) * np.log(1.0 - r * (1.0 - np.exp(-beta * vol)))


def u_sample(shape, beta, vol=1.0, rng: np.random.Generator | None = None):
    """Potential energies of ``shape = (nsamp, npart)`` sampled configurations."""
    return x_sample(shape=shape, beta=beta, vol=vol, rng=rng).sum(axis=-1)


_KINDS = {
    "plain": (xave_sym, beta_sym),
    "minuslog": (-sp.log(xave_sym), beta_sym),
    "depend": (beta_sym * xave_sym, beta_sym),
    "depend_minuslog": (-sp.log(beta_sym * xave_sym), beta_sym),
    "vol": (xave_sym, vol_sym),
}


@lru_cache(maxsize=500)
def _analytic_derivative(kind: str, k: int):
    expr, var = _KINDS[kind]
    return sp.lambdify([beta_sym, vol_sym], sp.diff(expr, var, k), "numpy")


def dbeta_xave(k):
    """k-th beta derivative of ``<x>`` as a function ``f(beta, vol)``."""
    return _analytic_derivative("plain", k)


def dbeta_xave_minuslog(k):
    """k-th beta derivative of ``-log <x>``."""
    return _analytic_derivative("minuslog", k)


def dbeta_xave_depend(k):
    """k-th beta derivative of ``beta <x>`` (observable depending explicitly on beta)."""
    return _analytic_derivative("depend", k)


def dbeta_xave_depend_minuslog(k):
    """k-th beta derivative of ``-log(beta <x>)``."""
    return _analytic_derivative("depend_minuslog", k)


def dvol_xave(k):
    """k-th volume derivative of ``<x>``."""
    return _analytic_derivative("vol", k)


def _taylor(kind, order, x0, x, args):
    d = x - x0
    derivs = np.array([_analytic_derivative(kind, k)(*args) for k in range(order + 1)])
    tot = 0.0
    for k in range(order + 1):
        tot = tot + derivs[k] / math.factorial(k) * (d**k)
    return tot, derivs


def x_beta_extrap(order, beta0, beta, vol=1.0):
    """Analytic Taylor extrapolation of ``<x>`` in beta: ``(value, derivatives)``."""
    return _taylor("plain", order, beta0, beta, (beta0, vol))


def x_beta_extrap_minuslog(order, beta0, beta, vol=1.0):
    return _taylor("minuslog", order, beta0, beta, (beta0, vol))


def x_beta_extrap_depend(order, beta0, beta, vol=1.0):
    return _taylor("depend", order, beta0, beta, (beta0, vol))


def x_beta_extrap_depend_minuslog(order, beta0, beta, vol=1.0):
    return _taylor("depend_minuslog", order, beta0, beta, (beta0, vol))


def x_vol_extrap(order, vol0, vol, beta=1.0):
    """Analytic Taylor extrapolation of ``<x>`` in volume."""
    return _taylor("vol", order, vol0, vol, (beta, vol0))


def generate_data(shape, beta, vol=1.0, rng: np.random.Generator | None = None):
    """``(x, u)`` per configuration: mean position and potential energy of ``shape = (nconfig, npart)``."""
    positions = x_sample(shape=shape, beta=beta, vol=vol, rng=rng)
    return positions.mean(axis=-1), positions.sum(axis=-1)

"""The sympy atoms the expansion modules share.

Every moment family names its data by a symbol (``u``, ``x1``, ``lnPi0`` ...) or by an indexed base (``du[n]``,
``dxdu[n, d]`` ...).  sympy interns ``Symbol`` by name, but ``IndexedBase`` construction is not free and the rule
tables of ``beta.py`` / ``lnpi.py`` ask for the same few names over and over, so both kinds are memoised here by name.
One name returns the object itself, several names a tuple / list in the order asked (the reference's calling
convention, ``core/sputils.py``).
"""

from __future__ import annotations

from functools import cache

import sympy as sp


@cache
def _symbol(name: str) -> sp.Symbol:
    return sp.Symbol(name)


@cache
def _indexed(name: str) -> sp.IndexedBase:
    return sp.IndexedBase(name)


def get_default_symbol(*names: str):
    """``Symbol(name)`` for one name, a tuple of symbols for several."""
    found = tuple(_symbol(n) for n in names)
    return found[0] if len(found) == 1 else found


def get_default_indexed(*names: str):
    """``IndexedBase(name)`` for one name, a list of them for several."""
    found = [_indexed(n) for n in names]
    return found[0] if len(found) == 1 else found

"""Cached sympy symbols / IndexedBase objects."""

from __future__ import annotations

from functools import lru_cache

import sympy as sp


@lru_cache(100)
def get_default_symbol(*args):
    return sp.symbols(",".join(args))


@lru_cache(100)
def get_default_indexed(*args):
    out = [sp.IndexedBase(key) for key in args]
    return out[0] if len(out) == 1 else out

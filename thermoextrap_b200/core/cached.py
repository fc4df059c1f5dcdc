"""Per-instance caching decorators (``module_utilities.cached`` protocol used by the reference:
``@cached.meth``, ``@cached.prop``, optional ``check_use_cache=True``; results live in
``self._cache`` keyed by method name and arguments)."""

from __future__ import annotations

import functools


def _wrap(func, as_prop: bool, check_use_cache: bool):
    name = func.__name__

    @functools.wraps(func)
    def wrapper(self, *args, **kwargs):
        if check_use_cache and not getattr(self, "_use_cache", True):
            return func(self, *args, **kwargs)
        key = (name, args, tuple(sorted(kwargs.items())))
        cache = self._cache
        try:
            return cache[key]
        except KeyError:
            out = cache[key] = func(self, *args, **kwargs)
            return out
        except TypeError:  # unhashable argument
            return func(self, *args, **kwargs)

    return property(wrapper) if as_prop else wrapper


def meth(func=None, *, check_use_cache: bool = False):
    if func is None:
        return lambda f: _wrap(f, False, check_use_cache)
    return _wrap(func, False, check_use_cache)


def prop(func=None, *, check_use_cache: bool = False):
    if func is None:
        return lambda f: _wrap(f, True, check_use_cache)
    return _wrap(func, True, check_use_cache)

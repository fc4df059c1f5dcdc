"""Labelled-array backend.

Default: :mod:`.labeled` (the minimal labelled array every test in this repo runs against -- the
validated backend).  Real xarray is an explicit opt-in (``THERMOEXTRAP_B200_XARRAY=1`` in the
environment before the first import): the image this was built in has no xarray, so that branch
is unvalidated and must not switch on silently just because ``import xarray`` happens to work.
Arrays cross the API edge either way: :func:`as_labeled` accepts anything with ``.dims`` /
``.values`` / ``.coords`` (e.g. an ``xarray.DataArray``) and :func:`to_xarray` converts back.
"""

from __future__ import annotations

import os

HAS_XARRAY = False
if os.environ.get("THERMOEXTRAP_B200_XARRAY", "0") not in ("", "0"):  # pragma: no cover - no xarray in the image
    import xarray as _xr

    DataArray = _xr.DataArray
    concat = _xr.concat
    HAS_XARRAY = True

    def dot(*arrays, dims=None, dim=None, **kw):
        d = dim if dims is None else dims
        try:
            return _xr.dot(*arrays, dim=d, **kw)      # xarray >= 2023.12
        except TypeError:
            return _xr.dot(*arrays, dims=d, **kw)     # older releases

    def assert_allclose(a, b, rtol=1e-5, atol=1e-8):
        _xr.testing.assert_allclose(a, b, rtol=rtol, atol=atol)

else:
    from .labeled import DataArray, assert_allclose, concat, dot


def is_dataarray(x: object) -> bool:
    return isinstance(x, DataArray)


def as_labeled(x):
    """Any labelled array (``.dims``, ``.values``, optional ``.coords`` / ``.name`` / ``.attrs``) -> this
    backend's ``DataArray``; the conversion at the API edge for callers holding real xarray objects."""
    if is_dataarray(x):
        return x
    if not (hasattr(x, "dims") and hasattr(x, "values")):
        msg = f"type(x)={type(x)} is not a labelled array"
        raise TypeError(msg)
    coords = {}
    for k, c in dict(getattr(x, "coords", {}) or {}).items():
        cd = tuple(getattr(c, "dims", ()))
        coords[k] = (cd, getattr(c, "values", c))
    return DataArray(x.values, dims=tuple(x.dims), coords=coords or None, name=getattr(x, "name", None),
                     attrs=dict(getattr(x, "attrs", {}) or {}))


def to_xarray(x):
    """This backend's ``DataArray`` -> ``xarray.DataArray`` (needs xarray; for callers that want xarray out)."""
    import xarray as xr

    if isinstance(x, xr.DataArray):
        return x
    coords = {k: (tuple(c.dims), c.values) for k, c in x.coords.items()}
    return xr.DataArray(x.values, dims=x.dims, coords=coords, name=x.name, attrs=dict(x.attrs))


__all__ = ["HAS_XARRAY", "DataArray", "as_labeled", "assert_allclose", "concat", "dot", "is_dataarray", "to_xarray"]

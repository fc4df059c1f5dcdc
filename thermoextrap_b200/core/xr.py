"""Labelled-array backend: real xarray when importable, else :mod:`.labeled`.

The hot path only relies on the API subset implemented by ``labeled.DataArray``.
"""

from __future__ import annotations

try:  # pragma: no cover - depends on the environment
    import xarray as _xr

    DataArray = _xr.DataArray
    concat = _xr.concat
    dot = _xr.dot
    HAS_XARRAY = True

    def assert_allclose(a, b, rtol=1e-5, atol=1e-8):
        _xr.testing.assert_allclose(a, b, rtol=rtol, atol=atol)

except ImportError:
    from .labeled import DataArray, assert_allclose, concat, dot

    HAS_XARRAY = False


def is_dataarray(x: object) -> bool:
    return isinstance(x, DataArray)


__all__ = ["HAS_XARRAY", "DataArray", "assert_allclose", "concat", "dot", "is_dataarray"]

"""Wrap raw arrays with thermoextrap's conventional dims (``core/xrutils.py:55-138`` of the reference)."""

from __future__ import annotations

from collections.abc import Mapping

import numpy as np

from .xr import DataArray, is_dataarray


def _check_xr(x, dims, name=None, strict=False):
    if is_dataarray(x):
        if strict:
            if isinstance(dims, Mapping):
                dims = dims[x.ndim]
            for d in dims:
                if d not in x.dims:
                    msg = f"{d} not in dims"
                    raise ValueError(msg)
        return x
    x = np.asarray(x)
    if isinstance(dims, Mapping):
        dims = dims[x.ndim]
    return DataArray(x, dims=dims, name=name)


def xrwrap_uv(uv, dims=None, rec_dim="rec", rep_dim="rep", name="u", strict=True):
    """``uv[rec]`` or ``uv[rep, rec]``."""
    if dims is None:
        dims = {1: [rec_dim], 2: [rep_dim, rec_dim]}
    return _check_xr(uv, dims, name=name, strict=strict)


def xrwrap_xv(xv, dims=None, rec_dim="rec", rep_dim="rep", deriv_dim=None, val_dims="val", name="x", strict=None):
    """``xv[rec]``, ``xv[rec, *val]``, ``xv[rep, rec, *val]`` (with ``deriv`` after ``rec`` if given)."""
    val_dims = [val_dims] if isinstance(val_dims, str) else list(val_dims)
    if strict is None:
        strict = False
    if dims is None:
        if deriv_dim is None:
            rec_val = [rec_dim, *val_dims]
            rep_val = [rep_dim, rec_dim, *val_dims]
            dims = {1: [rec_dim], len(rec_val): rec_val, len(rep_val): rep_val}
        else:
            rec_val = [rec_dim, deriv_dim, *val_dims]
            rep_val = [rep_dim, rec_dim, deriv_dim, *val_dims]
            dims = {2: [rec_dim, deriv_dim], len(rec_val): rec_val, len(rep_val): rep_val}
    return _check_xr(xv, dims=dims, name=name, strict=strict)


def xrwrap_alpha(alpha, dims=None, name="alpha"):
    """Wrap expansion-parameter values; 0-d and 1-d inputs get a coordinate of the same name."""
    if is_dataarray(alpha):
        return alpha
    alpha = np.array(alpha)
    if dims is None:
        dims = name
    if alpha.ndim == 0:
        return DataArray(alpha, coords={dims: alpha}, name=name)
    if alpha.ndim == 1:
        return DataArray(alpha, dims=dims, coords={dims: alpha}, name=name)
    return DataArray(alpha, dims=dims, name=name)

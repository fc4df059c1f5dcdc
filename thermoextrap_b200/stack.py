"""Statistics over the replicate axis of derivative arrays (reference ``stack.py:105-217``), on the device.

The reference's ``stack`` module feeds bootstrap derivatives to the GPR models (out of scope, gpflow); the part that
sits directly behind the hot path is ``to_mean_var`` / ``states_derivs_concat``: mean and variance over ``rep`` of
``[order, rep, val]`` arrays.  Here that reduction is one kernel (``txb_mean_var_f64``: Welford per thread, Chan
merge) so that a ``[order, rep, val]`` array produced on the device never has to be averaged on the host.
"""

from __future__ import annotations

import numpy as np

from . import backend as cm
from .core.xr import DataArray, concat


def to_mean_var(da, dim, concat_dim=None, concat_kws=None, **kws):
    """Mean and variance of ``da`` along ``dim``, stacked along a new ``stats`` dimension (``['mean', 'var']``).

    ``kws`` follow ``xarray.DataArray.mean`` / ``.var``: ``ddof`` (default 0) is honoured, ``skipna`` must be falsy
    (stack.py:157-182)."""
    if dim not in da.dims:
        msg = f"{dim!r} not in {da.dims}"
        raise ValueError(msg)
    ddof = int(kws.pop("ddof", 0))
    if kws.pop("skipna", None):
        msg = "skipna is not supported"
        raise NotImplementedError(msg)
    kws.pop("keep_attrs", None)
    if kws:
        msg = f"unsupported reduction arguments {sorted(kws)}"
        raise TypeError(msg)
    ax = da.dims.index(dim)
    vals = np.ascontiguousarray(np.asarray(da.values, dtype=np.float64))
    shape = vals.shape
    a = int(np.prod(shape[:ax], dtype=np.int64))
    b = int(np.prod(shape[ax + 1 :], dtype=np.int64))
    mean, var = cm.K.mean_var(cm.to_device(vals.reshape(a, shape[ax], b)), ddof=ddof)
    rest_dims = tuple(d for d in da.dims if d != dim)
    rest_shape = shape[:ax] + shape[ax + 1 :]
    coords = {k: np.asarray(c.values) for k, c in da.coords.items() if k in rest_dims}
    parts = [DataArray(cm._host(t).reshape(rest_shape), dims=rest_dims, coords=coords or None) for t in (mean, var)]
    if concat_dim is None:
        concat_dim = DataArray(np.array(["mean", "var"]), dims="stats", name="stats")
    return concat(parts, dim=concat_dim, **(concat_kws or {}))


def states_derivs_concat(states, dim=None, concat_kws=None, **kws):
    """``concat([s.derivs(norm=False) for s in states])`` along ``alpha_name`` (stack.py:185-217)."""
    if dim is None:
        dim = DataArray(np.asarray(states.alpha0), dims=states.alpha_name, name=states.alpha_name)
    kws.setdefault("norm", False)
    return concat([s.derivs(**kws) for s in states], dim=dim, **(concat_kws or {}))

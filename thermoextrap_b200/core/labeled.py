"""Minimal labelled array used when :mod:`xarray` is not installed.

Only what thermoextrap's hot path touches is implemented (see SURVEY.md 7.2
item 2): named dims, 1-D dimension coordinates + scalar coordinates,
``isel / sel / transpose``, broadcasting arithmetic by dimension name,
``sum / mean / std / var / cumsum / max``, ``concat`` and ``dot``.  The method
names and semantics follow xarray so that the same host code runs against
real ``xarray.DataArray`` objects when xarray is importable
(``thermoextrap_b200.core.xr``).
"""

from __future__ import annotations

from collections.abc import Hashable, Mapping, Sequence
from typing import Any

import numpy as np


def _as_tuple(dims) -> tuple[Hashable, ...]:
    if dims is None:
        return ()
    if isinstance(dims, str):
        return (dims,)
    return tuple(dims)


class DataArray:
    """Array with named dimensions (subset of ``xarray.DataArray``)."""

    __array_priority__ = 50
    __slots__ = ("_coords", "_dims", "_values", "attrs", "name")

    def __init__(self, data, dims=None, coords=None, name=None, attrs=None) -> None:
        if isinstance(data, DataArray):
            dims = data.dims if dims is None else dims
            coords = data._coords if coords is None else coords
            name = data.name if name is None else name
            data = data._values
        values = np.asarray(data)
        dims = _as_tuple(dims)
        if coords is not None and not dims and values.ndim > 0:
            dims = tuple(k for k in coords if np.ndim(coords[k]) == 1)[: values.ndim]
        if not dims and values.ndim > 0:
            dims = tuple(f"dim_{i}" for i in range(values.ndim))
        if len(dims) != values.ndim:
            msg = f"dims {dims} do not match array of shape {values.shape}"
            raise ValueError(msg)
        if len(set(dims)) != len(dims):
            msg = f"duplicate dimension names in {dims}"
            raise ValueError(msg)
        self._values = values
        self._dims = dims
        self._coords: dict[Hashable, tuple[tuple[Hashable, ...], np.ndarray]] = {}
        self.name = name
        self.attrs = dict(attrs) if attrs else {}
        if coords:
            for k, v in dict(coords).items():
                self._set_coord(k, v)

    # ------------------------------------------------------------------ coords
    def _set_coord(self, key, val) -> None:
        if isinstance(val, tuple) and len(val) == 2 and isinstance(val[0], (str, tuple, list)):
            cdims, cval = _as_tuple(val[0]), np.asarray(val[1])
        elif isinstance(val, DataArray):
            cdims, cval = val.dims, val._values
        else:
            cval = np.asarray(val)
            cdims = (key,) if cval.ndim == 1 else ()
            if cval.ndim > 1:
                msg = f"coordinate {key!r} needs explicit dims"
                raise ValueError(msg)
        for d, n in zip(cdims, cval.shape):
            if d not in self._dims:
                if cval.ndim == 0:
                    break
                msg = f"coordinate {key!r} has dim {d!r} not in {self._dims}"
                raise ValueError(msg)
            if self.sizes[d] != n:
                msg = f"coordinate {key!r} has length {n}, dim {d!r} has {self.sizes[d]}"
                raise ValueError(msg)
        self._coords[key] = (cdims, cval)

    @property
    def coords(self) -> dict[Hashable, DataArray]:
        return {k: DataArray(v, dims=d, name=k) for k, (d, v) in self._coords.items()}

    def assign_coords(self, coords: Mapping | None = None, **kw) -> DataArray:
        out = self.copy(deep=False)
        for k, v in {**(dict(coords) if coords else {}), **kw}.items():
            if callable(v):
                v = v(out)
            out._set_coord(k, v)
        return out

    def drop_vars(self, names, errors="raise") -> DataArray:
        out = self.copy(deep=False)
        for n in _as_tuple(names):
            out._coords.pop(n, None)
        return out

    # ------------------------------------------------------------------ basics
    @property
    def values(self) -> np.ndarray:
        return self._values

    @values.setter
    def values(self, v) -> None:
        v = np.asarray(v)
        if v.shape != self._values.shape:
            msg = "shape mismatch"
            raise ValueError(msg)
        self._values = v

    data = values

    @property
    def dims(self) -> tuple[Hashable, ...]:
        return self._dims

    @property
    def shape(self) -> tuple[int, ...]:
        return self._values.shape

    @property
    def ndim(self) -> int:
        return self._values.ndim

    @property
    def dtype(self):
        return self._values.dtype

    @property
    def size(self) -> int:
        return self._values.size

    @property
    def sizes(self) -> dict[Hashable, int]:
        return dict(zip(self._dims, self._values.shape))

    def __len__(self) -> int:
        if self.ndim == 0:
            msg = "len() of unsized object"
            raise TypeError(msg)
        return self.shape[0]

    def __array__(self, dtype=None, copy=None):
        return np.asarray(self._values, dtype=dtype)

    def __float__(self) -> float:
        return float(self._values)

    def __int__(self) -> int:
        return int(self._values)

    def __bool__(self) -> bool:
        return bool(self._values)

    def item(self):
        return self._values.item()

    def to_numpy(self) -> np.ndarray:
        return self._values

    def copy(self, deep: bool = True) -> DataArray:
        out = DataArray.__new__(DataArray)
        out._values = self._values.copy() if deep else self._values
        out._dims = self._dims
        out._coords = dict(self._coords)
        out.name = self.name
        out.attrs = dict(self.attrs)
        return out

    def astype(self, dtype) -> DataArray:
        return self._replace(self._values.astype(dtype))

    def pipe(self, func, *args, **kwargs):
        return func(self, *args, **kwargs)

    def rename(self, new_name_or_dims=None, **names) -> DataArray:
        out = self.copy(deep=False)
        if isinstance(new_name_or_dims, Mapping):
            names = {**new_name_or_dims, **names}
        elif new_name_or_dims is not None:
            out.name = new_name_or_dims
        if names:
            out._dims = tuple(names.get(d, d) for d in out._dims)
            out._coords = {
                names.get(k, k): (tuple(names.get(d, d) for d in cd), cv) for k, (cd, cv) in out._coords.items()
            }
        return out

    def _replace(self, values, dims=None, drop_dims=()) -> DataArray:
        """New array with the coordinates that still fit."""
        dims = self._dims if dims is None else tuple(dims)
        out = DataArray.__new__(DataArray)
        out._values = np.asarray(values)
        out._dims = dims
        out.name = self.name
        out.attrs = dict(self.attrs)
        out._coords = {}
        sizes = dict(zip(dims, out._values.shape))
        for k, (cd, cv) in self._coords.items():
            if k in drop_dims:
                continue
            if all(d in sizes and sizes[d] == n for d, n in zip(cd, cv.shape)):
                out._coords[k] = (cd, cv)
        return out

    def __repr__(self) -> str:
        dims = ", ".join(f"{d}: {n}" for d, n in self.sizes.items())
        body = np.array2string(self._values, threshold=20, edgeitems=3)
        coords = "".join(f"\n    {k!s} {cd} {np.array2string(cv, threshold=6, edgeitems=2)}" for k, (cd, cv) in self._coords.items())
        head = f"<thermoextrap_b200.DataArray {self.name + ' ' if self.name else ''}({dims})>"
        return f"{head}\n{body}" + (f"\nCoordinates:{coords}" if coords else "")


# ---------------------------------------------------------------------- indexing
def _isel(self: DataArray, indexers: Mapping | None = None, drop: bool = False, **kw) -> DataArray:
    indexers = {**(dict(indexers) if indexers else {}), **kw}
    for d in indexers:
        if d not in self._dims:
            msg = f"dimension {d!r} not in {self._dims}"
            raise ValueError(msg)
    # vectorised (DataArray / ndarray with ndim >= 2) indexer on one dim: new dims replace it
    values = self._values
    dims = list(self._dims)
    coords = dict(self._coords)
    for d, ix in indexers.items():
        ax = dims.index(d)
        if isinstance(ix, DataArray):
            ix_dims, ix_val = ix.dims, np.asarray(ix._values)
        else:
            ix_val = ix if isinstance(ix, slice) else np.asarray(ix)
            ix_dims = None
        if isinstance(ix_val, slice):
            sl = [slice(None)] * values.ndim
            sl[ax] = ix_val
            values = values[tuple(sl)]
            coords = {k: ((cd, cv[ix_val]) if cd == (d,) else (cd, cv)) for k, (cd, cv) in coords.items()}
        elif ix_val.ndim == 0:
            values = np.take(values, int(ix_val), axis=ax)
            new = {}
            for k, (cd, cv) in coords.items():
                if d in cd:
                    if not drop and cd == (d,):
                        new[k] = ((), cv[int(ix_val)])
                else:
                    new[k] = (cd, cv)
            coords = new
            dims.pop(ax)
        else:
            taken = np.take(values, ix_val, axis=ax)  # ix dims inserted at ax
            new_dims = list(ix_dims) if ix_dims is not None else ([d] if ix_val.ndim == 1 else None)
            if new_dims is None:
                msg = "multi-dimensional indexer needs dims (pass a DataArray)"
                raise ValueError(msg)
            values = taken
            dims = dims[:ax] + new_dims + dims[ax + 1 :]
            new = {}
            for k, (cd, cv) in coords.items():
                if d in cd:
                    if cd == (d,) and new_dims == [d]:
                        new[k] = (cd, cv[ix_val])
                else:
                    new[k] = (cd, cv)
            coords = new
    out = DataArray.__new__(DataArray)
    out._values, out._dims, out._coords = np.asarray(values), tuple(dims), coords
    out.name, out.attrs = self.name, dict(self.attrs)
    return out


def _index_of(self: DataArray, dim, label):
    if dim in self._coords and self._coords[dim][0] == (dim,):
        cv = self._coords[dim][1]
    else:
        cv = np.arange(self.sizes[dim])
    if isinstance(label, slice):
        return label
    lab = np.asarray(label)
    if lab.ndim == 0:
        hit = np.nonzero(cv == lab)[0]
        if hit.size == 0:
            raise KeyError(f"{label!r} not found in coordinate {dim!r}")
        return int(hit[0])
    out = []
    for v in lab.ravel():
        hit = np.nonzero(cv == v)[0]
        if hit.size == 0:
            raise KeyError(f"{v!r} not found in coordinate {dim!r}")
        out.append(int(hit[0]))
    return np.asarray(out).reshape(lab.shape)


def _sel(self: DataArray, indexers: Mapping | None = None, drop: bool = False, **kw) -> DataArray:
    indexers = {**(dict(indexers) if indexers else {}), **kw}
    conv = {}
    for d, lab in indexers.items():
        if d not in self._dims:
            msg = f"dimension {d!r} not in {self._dims}"
            raise ValueError(msg)
        if isinstance(lab, DataArray):
            lab = lab._values
        conv[d] = _index_of(self, d, lab)
    return _isel(self, conv, drop=drop)


def _getitem(self: DataArray, key):
    if isinstance(key, str):
        if key in self._coords:
            cd, cv = self._coords[key]
            return DataArray(cv, dims=cd, name=key)
        if key in self._dims:
            return DataArray(np.arange(self.sizes[key]), dims=(key,), name=key)
        raise KeyError(key)
    if isinstance(key, Mapping):
        return _isel(self, key)
    if not isinstance(key, tuple):
        key = (key,)
    if any(k is Ellipsis for k in key):
        i = key.index(Ellipsis)
        key = key[:i] + (slice(None),) * (self.ndim - len(key) + 1) + key[i + 1 :]
    out = self
    # apply from the last axis so that earlier axis numbers stay valid
    for ax in range(len(key) - 1, -1, -1):
        k = key[ax]
        if isinstance(k, slice) and k == slice(None):
            continue
        out = _isel(out, {self._dims[ax]: k})
    return out


def _transpose(self: DataArray, *dims, **_kw) -> DataArray:
    if not dims:
        dims = self._dims[::-1]
    dims = list(dims)
    if Ellipsis in dims:
        i = dims.index(Ellipsis)
        rest = [d for d in self._dims if d not in dims]
        dims = dims[:i] + rest + dims[i + 1 :]
    if sorted(map(str, dims)) != sorted(map(str, self._dims)):
        msg = f"{tuple(dims)} must be a permutation of {self._dims}"
        raise ValueError(msg)
    perm = [self._dims.index(d) for d in dims]
    return self._replace(self._values.transpose(perm), dims=dims)


def _expand_dims(self: DataArray, dim, axis=0) -> DataArray:
    return self._replace(np.expand_dims(self._values, axis), dims=self._dims[:axis] + (dim,) + self._dims[axis:])


DataArray.isel = _isel
DataArray.sel = _sel
DataArray.__getitem__ = _getitem
DataArray.transpose = _transpose
DataArray.expand_dims = _expand_dims
DataArray.T = property(lambda self: _transpose(self))


# ---------------------------------------------------------------------- arithmetic
def _broadcast_pair(a: DataArray, b):
    """Align by dimension name; result dims = a.dims then new dims of b (xarray rule)."""
    if not isinstance(b, DataArray):
        bv = np.asarray(b)
        if bv.ndim == 0:
            return a._values, bv, a._dims, a
        msg = "binary operations between DataArray and a non-scalar ndarray are ambiguous; wrap it"
        raise TypeError(msg)
    dims = list(a._dims) + [d for d in b._dims if d not in a._dims]

    def expand(x: DataArray):
        perm = [x._dims.index(d) for d in dims if d in x._dims]
        v = x._values.transpose(perm)
        shape = [x.sizes[d] if d in x._dims else 1 for d in dims]
        return v.reshape(shape)

    for d in dims:
        if d in a._dims and d in b._dims and a.sizes[d] != b.sizes[d]:
            msg = f"size mismatch along {d!r}: {a.sizes[d]} vs {b.sizes[d]}"
            raise ValueError(msg)
    return expand(a), expand(b), tuple(dims), b


def _binary(op, reflexive=False):
    def f(self: DataArray, other):
        if isinstance(other, (list, tuple)):
            return NotImplemented
        av, bv, dims, b = _broadcast_pair(self, other)
        val = op(bv, av) if reflexive else op(av, bv)
        out = DataArray.__new__(DataArray)
        out._values, out._dims = np.asarray(val), dims
        out.name = self.name if not isinstance(other, DataArray) or other.name == self.name else None
        out.attrs = {}
        coords = dict(self._coords)
        if isinstance(b, DataArray) and b is not self:
            for k, v in b._coords.items():
                coords.setdefault(k, v)
        sizes = dict(zip(dims, out._values.shape))
        out._coords = {
            k: (cd, cv) for k, (cd, cv) in coords.items() if all(d in sizes and sizes[d] == n for d, n in zip(cd, cv.shape))
        }
        return out

    return f


import operator as _op  # noqa: E402

for _name, _fn in [
    ("add", _op.add), ("sub", _op.sub), ("mul", _op.mul), ("truediv", _op.truediv),
    ("pow", _op.pow), ("floordiv", _op.floordiv), ("mod", _op.mod),
    ("lt", _op.lt), ("le", _op.le), ("gt", _op.gt), ("ge", _op.ge),
]:
    setattr(DataArray, f"__{_name}__", _binary(_fn))
    if _name in {"add", "sub", "mul", "truediv", "pow", "floordiv", "mod"}:
        setattr(DataArray, f"__r{_name}__", _binary(_fn, reflexive=True))
DataArray.__eq__ = _binary(_op.eq)
DataArray.__ne__ = _binary(_op.ne)
DataArray.__hash__ = None
DataArray.__neg__ = lambda self: self._replace(-self._values)
DataArray.__pos__ = lambda self: self
DataArray.__abs__ = lambda self: self._replace(np.abs(self._values))


def _array_ufunc(self, ufunc, method, *inputs, **kwargs):
    """np.exp(da), np.log(da), np.abs(da) ... (unary) and binary ufuncs with scalars."""
    if method != "__call__" or kwargs.get("out") is not None:
        return NotImplemented
    das = [x for x in inputs if isinstance(x, DataArray)]
    if len(inputs) == 1:
        return das[0]._replace(ufunc(das[0]._values, **kwargs))
    if len(inputs) == 2:
        a, b = inputs
        if isinstance(a, DataArray):
            return _binary(lambda x, y: ufunc(x, y, **kwargs))(a, b)
        return _binary(lambda x, y: ufunc(x, y, **kwargs), reflexive=True)(b, a)
    return NotImplemented


DataArray.__array_ufunc__ = _array_ufunc


# ---------------------------------------------------------------------- reductions
def _reduce(np_func, **fixed):
    def f(self: DataArray, dim=None, axis=None, skipna=None, keep_attrs=None, **kw):
        if dim is None and axis is None:
            axes = tuple(range(self.ndim))
        elif dim is not None:
            dims = _as_tuple(dim)
            for d in dims:
                if d not in self._dims:
                    msg = f"dimension {d!r} not in {self._dims}"
                    raise ValueError(msg)
            axes = tuple(self._dims.index(d) for d in dims)
        else:
            axes = tuple(np.atleast_1d(axis))
            axes = tuple(a % self.ndim for a in axes)
        val = np_func(self._values, axis=axes, **fixed, **kw)
        keep = tuple(d for i, d in enumerate(self._dims) if i not in axes)
        return self._replace(val, dims=keep, drop_dims={self._dims[a] for a in axes})

    return f


DataArray.sum = _reduce(np.sum)
DataArray.mean = _reduce(np.mean)
DataArray.max = _reduce(np.max)
DataArray.min = _reduce(np.min)
DataArray.prod = _reduce(np.prod)
DataArray.std = _reduce(np.std)
DataArray.var = _reduce(np.var)
DataArray.all = _reduce(np.all)
DataArray.any = _reduce(np.any)


def _cumsum(self: DataArray, dim=None, axis=None, **_kw) -> DataArray:
    ax = self._dims.index(dim) if dim is not None else (0 if axis is None else axis)
    return self._replace(np.cumsum(self._values, axis=ax))


DataArray.cumsum = _cumsum


# ---------------------------------------------------------------------- module-level
def concat(objs, dim, **_kw) -> DataArray:
    """Concatenate along an existing or a new dimension (name, or an index-like with ``.name``)."""
    objs = list(objs)
    if not objs:
        msg = "must supply at least one object to concatenate"
        raise ValueError(msg)
    coord_vals = None
    if isinstance(dim, DataArray):
        coord_vals, dim = dim._values, (dim.name if dim.name is not None else dim._dims[0])
    elif not isinstance(dim, str) and hasattr(dim, "name"):  # pandas.Index
        coord_vals, dim = np.asarray(dim), dim.name
    first = objs[0]
    objs = [o if isinstance(o, DataArray) else DataArray(o) for o in objs]
    if dim in first._dims:
        ax = first._dims.index(dim)
        aligned = [o.transpose(*first._dims) for o in objs]
        val = np.concatenate([o._values for o in aligned], axis=ax)
        out = first._replace(val, drop_dims={dim})
        if all(dim in o._coords for o in aligned):
            out._coords[dim] = ((dim,), np.concatenate([o._coords[dim][1] for o in aligned]))
        return out
    aligned = [o.transpose(*first._dims) for o in objs]
    val = np.stack([o._values for o in aligned], axis=0)
    out = first._replace(val, dims=(dim, *first._dims))
    # coords that differ between the objects are stacked along the new dim, as xarray.concat does
    # (scalar coords such as ``beta0`` -> ``(dim,)``; non-index coords such as ``dalpha(alpha)`` ->
    # ``(dim, alpha)``); index coords must agree, coords missing from some object are dropped
    for k, (cd, cv) in first._coords.items():
        if not all(k in o._coords and o._coords[k][0] == cd and o._coords[k][1].shape == cv.shape for o in aligned):
            out._coords.pop(k, None)
            continue
        vals = np.stack([np.asarray(o._coords[k][1]) for o in aligned], axis=0)
        same = all(np.array_equal(v, vals[0]) for v in vals[1:])
        if cd == (k,):  # dimension (index) coordinate
            if not same:
                msg = f"cannot concatenate along new dimension {dim!r}: index coordinate {k!r} differs"
                raise ValueError(msg)
            out._coords[k] = (cd, cv)
        elif k == dim or not same:
            out._coords[k] = ((dim, *cd), vals)
        else:
            out._coords[k] = (cd, cv)
    if coord_vals is not None:
        out._coords[dim] = ((dim,), np.asarray(coord_vals))
    return out


def dot(*arrays, dims=None, dim=None, **_kw) -> DataArray:
    """Sum of products over ``dims`` (default: all shared dims), like ``xarray.dot``."""
    arrays = [a if isinstance(a, DataArray) else DataArray(a) for a in arrays]
    dims = dim if dims is None else dims
    all_dims: list[Hashable] = []
    for a in arrays:
        for d in a._dims:
            if d not in all_dims:
                all_dims.append(d)
    if dims is None:
        count = {d: sum(d in a._dims for a in arrays) for d in all_dims}
        dims = tuple(d for d in all_dims if count[d] > 1)
    elif dims is Ellipsis:
        dims = tuple(all_dims)
    dims = _as_tuple(dims)
    letters = {d: chr(ord("a") + i) for i, d in enumerate(all_dims)}
    out_dims = [d for d in all_dims if d not in dims]
    spec = ",".join("".join(letters[d] for d in a._dims) for a in arrays) + "->" + "".join(letters[d] for d in out_dims)
    val = np.einsum(spec, *[a._values for a in arrays])
    out = DataArray(val, dims=out_dims)
    for a in arrays:
        for k, (cd, cv) in a._coords.items():
            if all(d in out_dims for d in cd) and k not in out._coords:
                out._coords[k] = (cd, cv)
    return out


def assert_allclose(a, b, rtol=1e-5, atol=1e-8) -> None:
    """``xr.testing.assert_allclose`` on values + dims."""
    a = a if isinstance(a, DataArray) else DataArray(a)
    b = b if isinstance(b, DataArray) else DataArray(b)
    if a.dims != b.dims:
        msg = f"dims differ: {a.dims} vs {b.dims}"
        raise AssertionError(msg)
    np.testing.assert_allclose(a.values, b.values, rtol=rtol, atol=atol)

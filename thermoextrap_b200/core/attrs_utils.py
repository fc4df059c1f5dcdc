"""attrs helpers mirroring ``thermoextrap.core._attrs_utils`` (new_like / assign / set_params)."""

from __future__ import annotations

import attrs


def convert_dims_to_tuple(dims):
    if dims is None:
        return ()
    return (dims,) if isinstance(dims, str) else tuple(dims)


def convert_mapping_or_none_to_dict(kws):
    return {} if kws is None else dict(kws)


def _init_fields(obj) -> list[str]:
    return [a.name for a in attrs.fields(type(obj)) if a.init]


class MyAttrsMixin:
    """Adds ``asdict``, ``new_like``, ``assign`` and ``set_params`` to attrs classes."""

    def asdict(self) -> dict:
        return {n: getattr(self, n) for n in _init_fields(self)}

    def new_like(self, **kws):
        """New object of the same type; unspecified fields are taken from ``self``."""
        current = {n.lstrip("_"): getattr(self, n) for n in _init_fields(self)}
        current.update(kws)
        return type(self)(**current)

    def assign(self, **kws):
        return self.new_like(**kws)

    def set_params(self, **kws):
        """In-place update of fields (clears the cache)."""
        names = {n.lstrip("_") for n in _init_fields(self)}
        for k, v in kws.items():
            if k not in names:
                msg = f"{k} not a parameter of {type(self).__name__}"
                raise ValueError(msg)
            object.__setattr__(self, k if hasattr(self, k) else "_" + k, v)
        if hasattr(self, "_cache"):
            object.__setattr__(self, "_cache", {})
        return self

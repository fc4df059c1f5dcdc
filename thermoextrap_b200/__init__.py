"""thermoextrap_b200: B200-native (sm_100a) implementation of thermoextrap's hot path.

samples -> weighted central comoments -> bootstrap resample -> derivative
expressions -> Taylor / interpolation predict, behind the reference's API::

    import thermoextrap_b200 as xtrap

    data = xtrap.DataCentralMomentsVals.from_vals(xv=x, uv=u, order=6, central=True)
    model = xtrap.beta.factory_extrapmodel(beta=beta0, data=data)
    model.resample(sampler={"nrep": 1000}).predict(betas)

Submodules are imported lazily; nothing here touches the GPU until a kernel is needed.
"""

from __future__ import annotations

import importlib

__version__ = "0.1.0"

_SUBMODULES = ["backend", "beta", "data", "idealgas", "kernels", "lnpi", "models", "parallel", "stack", "streaming", "volume",
               "volume_idealgas"]
_ATTRS = {
    "data": ["DataCentralMoments", "DataCentralMomentsVals", "DataValues", "DataValuesCentral", "factory_data_values"],
    "models": [
        "Derivatives", "ExtrapModel", "ExtrapWeightedModel", "InterpModel", "InterpModelPiecewise", "MBARModel",
        "PerturbModel", "StateCollection",
    ],
    "core.xrutils": ["xrwrap_alpha", "xrwrap_uv", "xrwrap_xv"],
    "core.xr": ["DataArray"],
}
_WHERE = {name: mod for mod, names in _ATTRS.items() for name in names}

__all__ = [*_SUBMODULES, *_WHERE, "__version__"]


def __getattr__(name: str):
    if name in _SUBMODULES:
        return importlib.import_module(f"{__name__}.{name}")
    if name in _WHERE:
        return getattr(importlib.import_module(f"{__name__}.{_WHERE[name]}"), name)
    msg = f"module {__name__!r} has no attribute {name!r}"
    raise AttributeError(msg)


def __dir__():
    return sorted(__all__)

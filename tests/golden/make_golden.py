"""Generate golden fixtures by running the REFERENCE's own code in the build container.

    python tests/golden/make_golden.py        (needs /root/reference; not needed on the GPU box)

The reference cannot be imported as a package here (cmomy / xarray / module_utilities /
lazy_loader are not installed), but its symbolic layer (``beta.py``, ``lnpi.py``,
``models.py`` sympy classes) only needs sympy + attrs once those imports are stubbed, and
``legacy/utilities.py`` runs unchanged.  Outputs (committed):

    exprs.json         expression strings per family / order from the reference's classes
    legacy_derivs.json derivatives + raw moments from the reference's legacy implementation
                       on the reference's FixtureData (tests/conftest.py:15-28)
    lnpi_sample.json   every 10th macrostate of tests/lnpi_data/sample_data.json
    perturb.json       legacy PerturbModel.predict (legacy/reweight.py) on the reference's FixtureData
    legacy_extrap.json legacy ExtrapModel (legacy/extrap.py:40-118): derivatives to order 6 and predict at every
                       order 0..6 on the reference's FixtureData -- what tests/test_beta.py:43-47 compares with
    idealgas.json      the reference's idealgas.py (only ``validate_rng`` shimmed): analytic functions, the seeded
                       sampler / generate_data, and every x_*_extrap variant (tests/test_idealgas.py)
"""

from __future__ import annotations

import importlib.util
import json
import sys
import types
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
SRC = REF / "src" / "thermoextrap"
OUT = Path(__file__).resolve().parent


class _Permissive(types.ModuleType):
    """Module whose every attribute is a do-nothing decorator / class factory."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = _Anything(name)
        setattr(self, name, obj)
        return obj


class _Anything:
    def __init__(self, name="x"):
        self._name = name

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        if name == "inherit":  # docfiller.inherit(Base) -> decorator
            return lambda *a, **k: (lambda f: f)
        return _Anything(name)

    def __call__(self, *args, **kwargs):
        # used as decorator: return the decorated object unchanged
        if len(args) == 1 and not kwargs and (callable(args[0]) or isinstance(args[0], (classmethod, staticmethod))):
            return args[0]
        return self

    def __mro_entries__(self, bases):
        return (object,)


def _cached_module():
    import functools

    m = types.ModuleType("module_utilities.cached")

    def _wrap(func):
        name = func.__name__

        @functools.wraps(func)
        def wrapper(self, *args, **kwargs):
            key = (name, args, tuple(sorted(kwargs.items())))
            if key not in self._cache:
                self._cache[key] = func(self, *args, **kwargs)
            return self._cache[key]

        return wrapper

    def meth(func=None, **_kw):
        return _wrap(func) if func is not None else (lambda f: _wrap(f))

    def prop(func=None, **_kw):
        return property(_wrap(func)) if func is not None else (lambda f: property(_wrap(f)))

    m.meth, m.prop = meth, prop
    return m


def install_stubs():
    for name in ["cmomy", "cmomy.core", "cmomy.core.missing", "cmomy.core.validate", "cmomy.core.docstrings",
                 "cmomy.core.typing", "cmomy.random", "xarray", "lazy_loader", "module_utilities.docfiller"]:
        sys.modules[name] = _Permissive(name)
    mu = types.ModuleType("module_utilities")
    mu.cached = _cached_module()
    mu.docfiller = sys.modules["module_utilities.docfiller"]
    sys.modules["module_utilities"] = mu
    sys.modules["module_utilities.cached"] = mu.cached
    xr = sys.modules["xarray"]
    xr.DataArray = type("DataArray", (), {})
    xr.Dataset = type("Dataset", (), {})
    cm = sys.modules["cmomy"]
    cm.CentralMomentsData = type("CentralMomentsData", (), {})
    cm.IndexSampler = type("IndexSampler", (), {})
    sys.modules["cmomy.random"].validate_rng = lambda rng=None: np.random.default_rng(rng)
    pkg = types.ModuleType("thermoextrap")
    pkg.__path__ = [str(SRC)]
    sys.modules["thermoextrap"] = pkg


def load_legacy():
    spec = importlib.util.spec_from_file_location("ref_legacy_utilities", SRC / "legacy" / "utilities.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_exprs():
    import thermoextrap.beta as rbeta
    import thermoextrap.lnpi as rlnpi
    import thermoextrap.models as rmodels

    out = {}
    cases = []
    for central in (True, False):
        for post in (None, "minus_log", "pow_2"):
            cases.append(("x_ave", dict(central=central, xalpha=False, post_func=post), 8 if post is None else 6))
            cases.append(("u_ave", dict(central=central, post_func=post), 8 if post is None else 6))
        cases.append(("x_ave", dict(central=central, xalpha=True), 4))
    for n in (2, 3):
        cases.append(("dun_ave", dict(n=n, central=True), 5))
        cases.append(("un_ave", dict(n=n, central=False), 5))
        cases.append(("dxdun_ave", dict(n=n, central=True, xalpha=False), 5))
        cases.append(("xun_ave", dict(n=n, central=False, xalpha=False), 5))
    cases.append(("dxdun_ave", dict(n=2, d=1, central=True, xalpha=True), 4))
    cases.append(("xun_ave", dict(n=2, d=1, central=False, xalpha=True), 4))
    for name, kws, order in cases:
        f = rbeta.factory_derivatives(name=name, **kws)
        key = name + "|" + ",".join(f"{k}={v}" for k, v in sorted(kws.items()))
        out[key] = {"name": name, "kws": kws, "args": [str(a) for a in f.args],
                    "exprs": [str(f.exprs[i]) for i in range(order + 1)]}
    for central in (True, False):
        f = rlnpi.factory_derivatives(central=central)
        out[f"lnPi|central={central}"] = {"name": "lnPi", "kws": {"central": central},
                                          "args": [str(a) for a in f.args],
                                          "exprs": [str(f.exprs[i]) for i in range(7)]}
    s = rmodels.SymMinusLog()
    out["SymMinusLog"] = {"exprs": [str(s[i]) for i in range(6)]}
    (OUT / "exprs.json").write_text(json.dumps(out, indent=1))
    print("exprs.json:", len(out), "families")


def make_legacy():
    legacy = load_legacy()
    # FixtureData(n=100, nv=5, order=5, seed=0), reference tests/conftest.py:15-28
    rng = np.random.default_rng(0)
    n, nv, order = 100, 5, 5
    u = rng.random(n)
    x = rng.random((n, nv))
    ub = rng.random(n)
    xb = rng.random((n, nv))
    ufunc, xufunc = legacy.buildAvgFuncs(x, u, order)
    derivs = [np.asarray(legacy.symDerivAvgX(i)(ufunc, xufunc), dtype=float) for i in range(order + 1)]
    out = {
        "seed": 0, "n": n, "nv": nv, "order": order,
        "u_raw": [float(ufunc(i)) for i in range(order + 1)],
        "xu_raw": [np.asarray(xufunc(i)).tolist() for i in range(order + 1)],
        "derivs": [d.tolist() for d in derivs],
        "check_u_sum": float(u.sum()), "check_ub_sum": float(ub.sum()), "check_xb_sum": float(xb.sum()),
    }
    (OUT / "legacy_derivs.json").write_text(json.dumps(out))
    print("legacy_derivs.json written")


def make_lnpi():
    d = json.loads((REF / "tests" / "lnpi_data" / "sample_data.json").read_text())
    sl = slice(None, None, 10)

    def cut(rec):
        r = dict(rec)
        r["lnPi"] = list(np.asarray(rec["lnPi"])[sl])
        r["energy"] = np.asarray(rec["energy"])[sl].tolist()
        return r

    out = {"stride": 10, "ref": cut(d["ref"]), "samples": [cut(s) for s in d["samples"]]}
    (OUT / "lnpi_sample.json").write_text(json.dumps(out))
    print("lnpi_sample.json written:", len(out["ref"]["lnPi"]), "macrostates")


def make_perturb():
    """``legacy.PerturbModel.predict`` (legacy/reweight.py:54-101) -- the implementation the reference's own test
    compares ``PerturbModel`` with (tests/test_beta.py:154-162) -- on the reference's FixtureData."""
    # reweight.py imports ``.interp`` (-> cmomy, xarray): give it a stub package and stub siblings
    pkg = types.ModuleType("ref_legacy")
    pkg.__path__ = [str(SRC / "legacy")]
    sys.modules["ref_legacy"] = pkg
    sys.modules["ref_legacy.interp"] = _Permissive("ref_legacy.interp")
    sys.modules["ref_legacy.interp"].InterpModel = object
    spec = importlib.util.spec_from_file_location("ref_legacy.reweight", SRC / "legacy" / "reweight.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_legacy.reweight"] = mod
    spec.loader.exec_module(mod)
    rng = np.random.default_rng(0)
    n, nv = 100, 5
    u = rng.random(n)
    x = rng.random((n, nv))
    beta0, betas = 0.5, [0.3, 0.7, 0.5, 2.5]
    pm = mod.PerturbModel(beta0, xData=x, uData=u)
    out = {"seed": 0, "n": n, "nv": nv, "beta0": beta0, "betas": betas,
           "predict": np.asarray(pm.predict(betas)).tolist(),
           "check_u_sum": float(u.sum()), "check_x_sum": float(x.sum())}
    (OUT / "perturb.json").write_text(json.dumps(out))
    print("perturb.json written")


def make_legacy_extrap():
    """``legacy.ExtrapModel`` (legacy/extrap.py:9-118) trained on the reference's FixtureData."""
    pkg = sys.modules.get("ref_legacy")
    if pkg is None:
        pkg = types.ModuleType("ref_legacy")
        pkg.__path__ = [str(SRC / "legacy")]
        sys.modules["ref_legacy"] = pkg
    for name in ("utilities", "extrap"):
        spec = importlib.util.spec_from_file_location(f"ref_legacy.{name}", SRC / "legacy" / f"{name}.py")
        mod = importlib.util.module_from_spec(spec)
        sys.modules[f"ref_legacy.{name}"] = mod
        spec.loader.exec_module(mod)
    extrap = sys.modules["ref_legacy.extrap"]
    rng = np.random.default_rng(0)
    n, nv, order = 100, 5, 6
    u = rng.random(n)
    x = rng.random((n, nv))
    beta0, betas = 0.5, [0.3, 0.4, 0.55, 0.9]
    em = extrap.ExtrapModel(maxOrder=order, refB=beta0, xData=x, uData=u)
    out = {"seed": 0, "n": n, "nv": nv, "order": order, "beta0": beta0, "betas": betas,
           "derivs": np.asarray(em.params).tolist(),
           "predict": [np.asarray(em.predict(betas, order=o)).tolist() for o in range(order + 1)],
           "check_u_sum": float(u.sum()), "check_x_sum": float(x.sum())}
    (OUT / "legacy_extrap.json").write_text(json.dumps(out))
    print("legacy_extrap.json written")


def make_idealgas():
    """The reference's ``idealgas.py`` itself (``validate_rng`` is the only shim, install_stubs)."""
    import thermoextrap.idealgas as ig

    b = np.linspace(0.1, 10, 5)
    rng = np.random.default_rng(seed=1234)
    x = rng.random()
    beta = rng.random()
    seed = 456
    xs = ig.x_sample(5, b[:, None], rng=np.random.default_rng(seed))
    us = ig.u_sample((4, 7), 2.5, 1.3, rng=np.random.default_rng(seed))
    xd, ud = ig.generate_data((100, 1000), beta, 1.0, rng=np.random.default_rng(seed))
    out = {
        "b": b.tolist(), "x": x, "beta": beta, "seed": seed,
        "x_ave": np.asarray(ig.x_ave(b)).tolist(), "x_ave_vol": np.asarray(ig.x_ave(b, 2.0)).tolist(),
        "x_var": np.asarray(ig.x_var(b)).tolist(), "x_prob": np.asarray(ig.x_prob(x, b)).tolist(),
        "u_prob": np.asarray(ig.u_prob(x * 300, 1000, b)).tolist(), "x_cdf": np.asarray(ig.x_cdf(x, b)).tolist(),
        "x_sample": np.asarray(xs).tolist(), "u_sample": np.asarray(us).tolist(),
        "generate_data_x": np.asarray(xd).tolist(), "generate_data_u": np.asarray(ud).tolist(),
        "dbeta_xave": [float(ig.dbeta_xave(k)(1.7, 1.2)) for k in range(7)],
        "dbeta_xave_minuslog": [float(ig.dbeta_xave_minuslog(k)(1.7, 1.2)) for k in range(7)],
        "dbeta_xave_depend": [float(ig.dbeta_xave_depend(k)(1.7, 1.2)) for k in range(7)],
        "dbeta_xave_depend_minuslog": [float(ig.dbeta_xave_depend_minuslog(k)(1.7, 1.2)) for k in range(7)],
        "dvol_xave": [float(ig.dvol_xave(k)(1.7, 1.2)) for k in range(7)],
    }
    betas = np.array([0.7, 1.5, 2.0])
    for name in ("x_beta_extrap", "x_beta_extrap_minuslog", "x_beta_extrap_depend", "x_beta_extrap_depend_minuslog"):
        tot, coef = getattr(ig, name)(5, beta0=1.0, beta=betas, vol=1.3)
        out[name] = {"tot": np.asarray(tot, dtype=float).tolist(), "coef": np.asarray(coef, dtype=float).tolist()}
    tot, coef = ig.x_vol_extrap(order=5, vol0=0.3, beta=beta, vol=np.array([0.5, 1.0]))
    out["x_vol_extrap"] = {"tot": np.asarray(tot, dtype=float).tolist(), "coef": np.asarray(coef, dtype=float).tolist()}
    (OUT / "idealgas.json").write_text(json.dumps(out))
    print("idealgas.json written")


if __name__ == "__main__":
    make_lnpi()
    make_legacy()
    install_stubs()
    make_exprs()
    make_perturb()
    make_legacy_extrap()
    make_idealgas()

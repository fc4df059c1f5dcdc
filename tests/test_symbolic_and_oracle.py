"""CPU tests: the oracle against the reference's golden vectors, the symbolic layer against the
reference's own expressions, and the C ABI surface of libtxb.so (no compute calls)."""

import json
import re
from pathlib import Path

import numpy as np
import pytest
import sympy as sp

from oracle import comoments as oc
from oracle import sampler as osamp

ROOT = Path(__file__).resolve().parent.parent
GOLD = Path(__file__).resolve().parent / "golden"


# ------------------------------------------------------------------ oracle pinned
@pytest.fixture(scope="module")
def legacy():
    return json.loads((GOLD / "legacy_derivs.json").read_text())


def fixture_arrays():
    """The reference's FixtureData(n=100, nv=5, seed=0) (tests/conftest.py:15-28)."""
    rng = np.random.default_rng(0)
    u = rng.random(100)
    x = rng.random((100, 5))
    ub = rng.random(100)
    xb = rng.random((100, 5))
    return u, x, ub, xb


def test_oracle_raw_moments_match_reference_legacy(legacy):
    """reference tests/test_data.py:7-38: raw moments == np.average(x*u**i) (legacy/utilities.py:12-39)."""
    u, x, ub, xb = fixture_arrays()
    assert np.isclose(u.sum(), legacy["check_u_sum"], rtol=0, atol=0)
    assert np.isclose(xb.sum(), legacy["check_xb_sum"], rtol=0, atol=0)
    c = oc.comoments_two_pass(x, u, legacy["order"])
    raw = oc.central_to_raw(c)
    np.testing.assert_allclose(raw[0, 0, :], legacy["u_raw"], rtol=1e-13)
    np.testing.assert_allclose(raw[:, 1, :].T, np.array(legacy["xu_raw"]), rtol=1e-13)


def test_oracle_push_merge_equivalences():
    u, x, _, _ = fixture_arrays()
    a = oc.comoments_two_pass(x, u, 5)
    np.testing.assert_allclose(oc.push_vals(x, u, 5), a, rtol=1e-12, atol=1e-15)
    blocks = np.stack([oc.comoments_two_pass(x[i::4], u[i::4], 5) for i in range(4)])
    np.testing.assert_allclose(oc.merge_blocks(blocks), a, rtol=1e-12, atol=1e-15)
    back = oc.raw_to_central(np.where(np.arange(6) + np.arange(2)[:, None] == 0, a[..., :1, :1], oc.central_to_raw(a)))
    np.testing.assert_allclose(back, a, rtol=1e-11, atol=1e-14)


def test_oracle_resample_freq_equals_gather():
    """reference tests/test_data.py:94-112."""
    u, x, _, _ = fixture_arrays()
    idx = np.random.default_rng(1).choice(100, (10, 100), replace=True)
    f = oc.indices_to_freq(idx, 100)
    assert f.sum(axis=1).tolist() == [100] * 10
    np.testing.assert_allclose(oc.resample_vals_freq(x, u, 5, f), oc.resample_vals_gather(x, u, 5, idx), rtol=1e-12, atol=1e-15)


@pytest.mark.parametrize("ndat", [7, 1000, 100_000, 3_000_000_000])
def test_sampler_restatement_matches_numpy(ndat):
    rng = np.random.default_rng(3)
    st = osamp.pcg64_state(rng)
    ref = rng.choice(ndat, size=(3, 400), replace=True).ravel()
    got, _ = osamp.pcg64_bounded(*st, ndat, 1200)
    assert np.array_equal(ref, got)


def test_sampler_advance_matches_numpy():
    rng = np.random.default_rng(5)
    s, inc, _, _ = osamp.pcg64_state(rng)
    rng.bit_generator.advance(123_456_789)
    assert osamp.pcg64_advance(s, inc, 123_456_789) == osamp.pcg64_state(rng)[0]


# ------------------------------------------------------------------ symbolic layer
EXPRS = json.loads((GOLD / "exprs.json").read_text())


def _parse(s, names):
    local = {n: (sp.IndexedBase(n) if re.search(rf"(?<![A-Za-z_0-9]){n}\[", s) else sp.Symbol(n)) for n in names}
    return sp.sympify(s, locals=local)


@pytest.mark.parametrize("key", [k for k in EXPRS if k != "SymMinusLog"])
def test_expressions_match_reference(key):
    """Same sympy expressions as the reference's beta.py / lnpi.py classes (run in the build container)."""
    from thermoextrap_b200 import beta, lnpi

    rec = EXPRS[key]
    if rec["name"] == "lnPi":
        f = lnpi.factory_derivatives(**rec["kws"])
    else:
        f = beta.factory_derivatives(name=rec["name"], **rec["kws"])
    assert [str(a) for a in f.args] == rec["args"]
    names = set(re.findall(r"[A-Za-z_][A-Za-z_0-9]*", " ".join(rec["exprs"]))) - {"log"}
    for i, s in enumerate(rec["exprs"]):
        ours = f.exprs[i]
        assert sp.expand(ours - _parse(s, names)) == 0, (key, i)
        assert str(ours) == s, (key, i)  # same term order as well


def test_minus_log_expressions_match_reference():
    from thermoextrap_b200.models import SymMinusLog

    s = SymMinusLog()
    for i, ref in enumerate(EXPRS["SymMinusLog"]["exprs"]):
        assert sp.simplify(s[i] - _parse(ref, {"X"})) == 0


@pytest.mark.parametrize("central", [True, False])
def test_lnpi_equations(central):
    """reference tests/test_lnPi.py:15-21."""
    from thermoextrap_b200 import beta, lnpi

    f0 = beta.factory_derivatives("u_ave", central=central)
    f1 = lnpi.factory_derivatives(central=central)
    for i in range(2, 6):
        assert f0.exprs[i] + f1.exprs[i + 1] == 0


@pytest.mark.parametrize("central", [True, False])
@pytest.mark.parametrize("post_func", [None, "minus_log"])
def test_x_to_u_substitution(central, post_func):
    """reference tests/test_u_equations.py:55-88: x -> u turns x_ave expressions into u_ave ones."""
    from thermoextrap_b200 import beta

    fx = beta.factory_derivatives("x_ave", central=central, post_func=post_func)
    fu = beta.factory_derivatives("u_ave", central=central, post_func=post_func)
    u, du, dxdu, xu = sp.Symbol("u"), sp.IndexedBase("du"), sp.IndexedBase("dxdu"), sp.IndexedBase("xu")
    ui = sp.IndexedBase("u")
    for i in range(6):
        e = fx.exprs[i]
        if central:
            subs = {sp.Symbol("x1"): u, **{dxdu[j]: du[j + 1] for j in range(1, 8)}}
        else:
            subs = {xu[j]: ui[j + 1] for j in range(0, 8)}
        assert sp.expand(e.subs(subs) - fu.exprs[i]) == 0


def test_named_constructors_keep_reference_signatures():
    """reference beta.py:291-470: positional order of each named constructor, its validation errors, and
    ``from_name`` rejecting unknown quantities."""
    from thermoextrap_b200.beta import SymDerivBeta, du_func, u_func_central, x_func_central_beta, xu_func

    def same(a, b, n=4):
        assert [str(a[i]) for i in range(n)] == [str(b[i]) for i in range(n)]
        assert [str(x) for x in a.args] == [str(x) for x in b.args]

    same(SymDerivBeta.x_ave(True, True), SymDerivBeta.x_ave(xalpha=True, central=True))
    same(SymDerivBeta.u_ave(True), SymDerivBeta.from_name("u_ave", central=True))
    same(SymDerivBeta.dun_ave(3), SymDerivBeta.from_name("dun_ave", n=3))
    same(SymDerivBeta.dxdun_ave(2, True, True, None, 1), SymDerivBeta.dxdun_ave(n=2, xalpha=True, d=1))
    same(SymDerivBeta.un_ave(2), SymDerivBeta.from_name("un_ave", n=2, central=False))
    same(SymDerivBeta.xun_ave(2, 1, True), SymDerivBeta.from_name("xun_ave", n=2, d=1, xalpha=True))
    for bad in (lambda: SymDerivBeta.dun_ave(1), lambda: SymDerivBeta.dun_ave(2, central=False),
                lambda: SymDerivBeta.dxdun_ave(0),
                lambda: SymDerivBeta.un_ave(0), lambda: SymDerivBeta.un_ave(2, central=True),
                lambda: SymDerivBeta.xun_ave(-1), lambda: SymDerivBeta.xun_ave(1, central=True),
                lambda: SymDerivBeta.from_name("nope")):
        with pytest.raises(ValueError):
            bad()
    with pytest.raises(TypeError):
        SymDerivBeta.dun_ave()
    with pytest.raises(TypeError):   # xalpha needs an integer d (reference beta.py:395-397)
        SymDerivBeta.dxdun_ave(1, xalpha=True)
    with pytest.raises(TypeError):
        SymDerivBeta.u_ave(True, True, None, 1)
    # the function classes: boundary values, symbols, argument lists
    b = sp.Symbol("beta")
    assert du_func(b, 0) == 1 and du_func(b, 1) == 0 and xu_func(b, 0).nargs is not None
    assert str(du_func(None, 3)) == "du[3]" and str(u_func_central(None)) == "u"
    assert str(x_func_central_beta(None, 2)) == "x1[2]" and str(xu_func(None, 2, 1)) == "xu[2, 1]"
    assert [str(a) for a in x_func_central_beta.deriv_args()] == ["x1", "du", "dxdu"]
    assert [str(a) for a in xu_func.deriv_args()] == ["u", "xu"]
    assert str(sp.diff(du_func(b, 3), b)) == str(3 * du_func(b, 2) ** 2 - du_func(b, 4))


# ------------------------------------------------------------------ ABI surface
def test_library_exports_every_declared_symbol():
    from thermoextrap_b200 import _lib

    header = (ROOT / "include" / "txb.h").read_text()
    declared = set(re.findall(r"\b(txb_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations found"
    lib = _lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/txb.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.txb_version() >= 100
    assert lib.txb_max_order() >= 11


def test_product_has_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from thermoextrap_b200 import kernels

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        kernels.reduce_vals(np.zeros(4), np.zeros(4), 2)


def test_product_does_not_import_oracle():
    for p in (ROOT / "thermoextrap_b200").rglob("*.py"):
        assert "oracle" not in p.read_text(), p


def test_new_entry_points_reject_bad_arguments_before_any_launch():
    """Argument validation of the round-2 entry points returns an error code and a message without touching the
    device (so this runs without a GPU): null pointers, unknown family / mode, records too short for the order,
    orders above the generated range (TXB_ERR_UNSUPPORTED = 2), ninner not dividing nelem, ddof >= nrep."""
    import ctypes as C

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    one = C.c_void_p(8)   # a non-null placeholder: every call below fails validation before dereferencing
    err = lambda: lib.txb_last_error_string().decode()  # noqa: E731
    assert lib.txb_beta_predict_f64(None, 4, 7, 0, 6, 0, 0, one, 1, one, None) == 1 and "null" in err()
    assert lib.txb_beta_predict_f64(one, 4, 7, 9, 6, 0, 0, one, 1, one, None) == 1 and "family" in err()
    assert lib.txb_beta_predict_f64(one, 4, 7, 0, 6, 0, 7, one, 1, one, None) == 1 and "mode" in err()
    assert lib.txb_beta_predict_f64(one, 4, 5, 0, 6, 0, 0, one, 1, one, None) == 1 and "order 6" in err()
    assert lib.txb_beta_predict_f64(one, 4, 20, 0, 12, 0, 0, one, 1, one, None) == 2 and "generated maximum" in err()
    assert lib.txb_beta_predict_f64(one, 4, 7, 0, 6, 0, 0, None, 1, one, None) == 1 and "dalpha" in err()
    assert lib.txb_beta_predict_f64(one, 0, 7, 0, 6, 0, 0, one, 1, one, None) == 0     # nothing to do
    assert lib.txb_lnpi_predict_f64(one, 6, 7, 1, 6, 0, 0, one, 1, None, one, 3, one, None) == 1 and "null" in err()
    assert lib.txb_lnpi_predict_f64(one, 7, 7, 1, 6, 0, 0, one, 1, one, one, 3, one, None) == 1 and "multiple" in err()
    assert lib.txb_lnpi_predict_f64(one, 6, 5, 1, 6, 0, 0, one, 1, one, one, 3, one, None) == 1 and "order 6" in err()
    assert lib.txb_lnpi_predict_f64(one, 6, 7, 1, 6, 1, 0, one, 1, one, one, 3, one, None) == 1 and "-log" in err()
    assert lib.txb_combine_f64(one, 3, 600, one, 10, 0, one, None) == 1 and "nt <= 512" in err()
    assert lib.txb_combine_f64(one, 0, 5, one, 10, 0, one, None) == 0
    assert lib.txb_mean_var_f64(one, 2, 5, 3, 5, one, one, None) == 1 and "ddof" in err()
    assert lib.txb_mean_var_f64(None, 2, 5, 3, 0, one, one, None) == 1 and "null" in err()
    assert lib.txb_set_resample_kernel(-1) in (0, 1, 2, 3, 4)
    assert lib.txb_set_scatter_threads(-1) in (512, 1024)

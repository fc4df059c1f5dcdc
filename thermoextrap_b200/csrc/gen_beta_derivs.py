"""Build-time generator of the beta-derivative device functions (K7), ``beta_derivs_gen.cuh``.

    python thermoextrap_b200/csrc/gen_beta_derivs.py [--check]

north_star (3): "the sympy-derived derivative expressions are code-generated into device functions".  The
polynomials are NOT taken from sympy: they are derived here from the reference's recurrences with a small
integer-coefficient polynomial algebra (dict: monomial -> int), and ``tests/test_beta_codegen.py``
checks them term by term against ``tests/golden/exprs.json`` (strings produced by the reference's own
classes, beta.py:32-266).  Rules (beta = None substitution, i.e. in terms of the moments at beta0):

  central   d<du^n>    = -<du^(n+1)> + n <du^(n-1)> <du^2>                    du_0 = 1, du_1 = 0   (beta.py:49-51)
            d<dx du^n> = -<dx du^(n+1)> + n <dx du^(n-1)> <du^2> + <dx du> <du^n>      dxdu_0 = 0  (beta.py:107-113)
            d<x> = -<dx du>     (beta.py:164-180)          d<u> = -<du^2>     (beta.py:80-82)
  raw       d<u^n>   = -<u^(n+1)> + <u^n> <u>              u_0 = 1            (beta.py:206-228)
            d<x u^n> = -<x u^(n+1)> + <x u^n> <u>                             (beta.py:231-266)

Families (index = ``family`` of txb_beta_predict_f64): 0 x_ave central, 1 u_ave central, 2 x_ave raw, 3 u_ave raw.
Variables are read from the two rows of a comoment record ``[2, K+1]``:
  central: r0[n] = <du^n> (n >= 2), r0[1] = <u>;  r1[0] = <x>, r1[n] = <dx du^n>
  raw:     r0[n] = <u^n>;  r1[n] = <x u^n>
"""

from __future__ import annotations

import sys
from pathlib import Path

MAX_ORDER = 10
FAMILIES = ("x_ave_central", "u_ave_central", "x_ave_raw", "u_ave_raw")

# a variable is (row, n); a monomial is a sorted tuple of (variable, power); a polynomial a dict monomial -> int


def _mul(a: dict, b: dict) -> dict:
    out: dict = {}
    for ma, ca in a.items():
        for mb, cb in b.items():
            pw: dict = dict(ma)
            for v, p in mb:
                pw[v] = pw.get(v, 0) + p
            m = tuple(sorted(pw.items()))
            out[m] = out.get(m, 0) + ca * cb
    return {m: c for m, c in out.items() if c}


def _add(a: dict, b: dict, s: int = 1) -> dict:
    out = dict(a)
    for m, c in b.items():
        out[m] = out.get(m, 0) + s * c
    return {m: c for m, c in out.items() if c}


def _var(row: int, n: int) -> dict:
    return {(((row, n), 1),): 1}


ONE: dict = {(): 1}
ZERO: dict = {}


def _scale(a: dict, s: int) -> dict:
    return {m: c * s for m, c in a.items() if c * s}


class _Central:
    """du_n = (0, n), dxdu_n = (1, n), x1 = (1, 0), u1 = (0, 1)."""

    @staticmethod
    def du(n: int) -> dict:
        return ONE if n == 0 else ZERO if n == 1 else _var(0, n)

    @staticmethod
    def dxdu(n: int) -> dict:
        return ZERO if n == 0 else _var(1, n)

    @classmethod
    def dvar(cls, v) -> dict:
        row, n = v
        if row == 0 and n == 1:   # <u>
            return _scale(cls.du(2), -1)
        if row == 1 and n == 0:   # <x>
            return _scale(cls.dxdu(1), -1)
        if row == 0:
            return _add(_scale(cls.du(n + 1), -1), _scale(_mul(cls.du(n - 1), cls.du(2)), n))
        out = _add(_scale(cls.dxdu(n + 1), -1), _scale(_mul(cls.dxdu(n - 1), cls.du(2)), n))
        return _add(out, _mul(cls.dxdu(1), cls.du(n)))


class _Raw:
    """u_n = (0, n), xu_n = (1, n)."""

    @staticmethod
    def dvar(v) -> dict:
        row, n = v
        u1 = _var(0, 1)
        return _add(_scale(_var(row, n + 1), -1), _mul(_var(row, n), u1))


def _deriv(poly: dict, rules) -> dict:
    out: dict = {}
    for mono, c in poly.items():
        for i, (v, p) in enumerate(mono):
            rest = list(mono[:i]) + ([(v, p - 1)] if p > 1 else []) + list(mono[i + 1:])
            term = _mul({tuple(sorted(rest)): c * p}, rules.dvar(v))
            out = _add(out, term)
    return out


def family_polys(family: int, max_order: int = MAX_ORDER) -> list[dict]:
    """``[d_0, ..., d_max_order]`` of one family as integer polynomials."""
    rules = _Central if family < 2 else _Raw
    d0 = {0: _var(1, 0), 1: _var(0, 1), 2: _var(1, 0), 3: _var(0, 1)}[family]
    out = [d0]
    for _ in range(max_order):
        out.append(_deriv(out[-1], rules))
    return out


def poly_to_str(poly: dict, family: int) -> str:
    """sympy-parsable string in the reference's symbols (``tests/golden/exprs.json`` naming)."""
    central = family < 2

    def name(v):
        row, n = v
        if central:
            if row == 0:
                return "u" if n == 1 else f"du[{n}]"
            return "x1" if n == 0 else f"dxdu[{n}]"
        return f"u[{n}]" if row == 0 else f"xu[{n}]"

    terms = []
    for mono, c in sorted(poly.items()):
        fac = "*".join(f"{name(v)}**{p}" if p > 1 else name(v) for v, p in mono)
        terms.append(f"({c})" + (f"*{fac}" if fac else ""))
    return " + ".join(terms) if terms else "0"


def _emit_poly(poly: dict) -> str:
    def ref(v):
        return f"r{v[0]}[{v[1]}]"

    terms = []
    for mono, c in sorted(poly.items(), key=lambda kv: (sum(p for _, p in kv[0]), kv[0])):
        facs = []
        for v, p in mono:
            facs += [ref(v)] * p
        body = " * ".join(facs)
        if not facs:
            terms.append(f"{'+' if c > 0 else '-'} {abs(c)}.0")
        elif abs(c) == 1:
            terms.append(f"{'+' if c > 0 else '-'} {body}")
        else:
            terms.append(f"{'+' if c > 0 else '-'} {abs(c)}.0 * {body}")
    s = " ".join(terms) if terms else "0.0"
    return s[2:] if s.startswith("+ ") else s


def generate() -> str:
    lines = [
        "// GENERATED by gen_beta_derivs.py -- do not edit.  Beta-derivative polynomials of the four numeric families",
        "// (reference: src/thermoextrap/beta.py:32-266 with beta -> None), derived from the recurrences by an",
        "// integer polynomial algebra and checked against tests/golden/exprs.json.",
        "#pragma once",
        "",
        "namespace txb {",
        "",
        f"constexpr int BETA_GEN_MAX_ORDER = {MAX_ORDER};",
        f"constexpr int BETA_GEN_NFAMILY = {len(FAMILIES)};",
        "",
    ]
    for fam, fname in enumerate(FAMILIES):
        polys = family_polys(fam)
        lines.append(f"// family {fam}: {fname}; r0 / r1 = the two rows of one comoment record "
                     f"({'central: r0[1] = <u>, r0[n] = <du^n>, r1[0] = <x>, r1[n] = <dx du^n>' if fam < 2 else 'raw: r0[n] = <u^n>, r1[n] = <x u^n>'})")
        lines.append("template <int ORDER>")
        lines.append(f"__device__ __forceinline__ void beta_derivs_{fname}(const double* __restrict__ r0, "
                     "const double* __restrict__ r1, double* __restrict__ d) {")
        lines.append("    (void)r0; (void)r1;")
        for i, p in enumerate(polys):
            expr = _emit_poly(p)
            if i == 0:
                lines.append(f"    d[0] = {expr};")
            else:
                lines.append(f"    if constexpr (ORDER >= {i}) d[{i}] = {expr};")
        lines.append("}")
        lines.append("")
    lines.append("}  // namespace txb")
    return "\n".join(lines) + "\n"


def main() -> None:
    out = Path(__file__).resolve().parent / "beta_derivs_gen.cuh"
    text = generate()
    if "--check" in sys.argv:
        if not out.exists() or out.read_text() != text:
            sys.exit(f"{out} is stale: run python {Path(__file__).name}")
        return
    if not out.exists() or out.read_text() != text:
        out.write_text(text)


if __name__ == "__main__":
    main()

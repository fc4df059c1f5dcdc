"""Oracle for the reference sampler's integer stream (TEST INFRASTRUCTURE ONLY).

The reference draws bootstrap indices with
``rng.choice(ndat, (nrep, nsamp), replace=True)`` on a ``numpy.random.Generator``
(PCG64) -- written out by thermoextrap itself at
``/root/reference/src/thermoextrap/adaptive_interp.py:543-544`` and
``/root/reference/tests/test_data.py:98-100``; cmomy's ``factory_sampler`` does
the same for ``{"nrep": ...}`` specs.

Live numpy is the golden reference; ``pcg64_bounded`` below is the pure-Python
restatement of the published algorithm (PCG64 XSL-RR 128/64, buffered 32-bit
halves, Lemire bounded rejection) that the device generator is checked against
for jump-ahead / rejection bookkeeping.
"""

from __future__ import annotations

import numpy as np

PCG_MULT = 0x2360ED051FC65DA44385DF649FCCF645
M128 = (1 << 128) - 1
M64 = (1 << 64) - 1
M32 = (1 << 32) - 1


def numpy_indices(seed, nrep, nsamp, ndat=None):
    """What the reference sampler produces for ``default_rng(seed)``."""
    ndat = nsamp if ndat is None else ndat
    return np.random.default_rng(seed).choice(ndat, size=(nrep, nsamp), replace=True)


def pcg64_state(rng):
    """``(state, inc, has_uint32, uinteger)`` of a numpy ``Generator(PCG64)``."""
    st = rng.bit_generator.state
    return st["state"]["state"], st["state"]["inc"], st["has_uint32"], st["uinteger"]


def _next64(state, inc):
    state = (state * PCG_MULT + inc) & M128
    x = ((state >> 64) ^ state) & M64
    rot = state >> 122
    return state, ((x >> rot) | (x << ((-rot) & 63))) & M64


def pcg64_advance(state, inc, delta):
    """LCG jump-ahead by ``delta`` steps (O(log delta))."""
    acc_mult, acc_plus = 1, 0
    cur_mult, cur_plus = PCG_MULT, inc
    while delta > 0:
        if delta & 1:
            acc_mult = (acc_mult * cur_mult) & M128
            acc_plus = (acc_plus * cur_mult + cur_plus) & M128
        cur_plus = ((cur_mult + 1) * cur_plus) & M128
        cur_mult = (cur_mult * cur_mult) & M128
        delta >>= 1
    return (acc_mult * state + acc_plus) & M128


def pcg64_bounded(state, inc, has_u32, u32, ndat, count):
    """``count`` draws of ``Generator.integers(0, ndat)`` for ``ndat - 1 < 2**32 - 1``.

    Returns ``(draws int64[count], (state, inc, has_u32, u32))``.
    """
    out = np.empty(count, dtype=np.int64)

    def next32():
        nonlocal state, has_u32, u32
        if has_u32:
            has_u32 = 0
            return u32
        state, v = _next64(state, inc)
        u32 = v >> 32
        has_u32 = 1
        return v & M32

    rng_excl = ndat
    thresh = ((1 << 32) - rng_excl) % rng_excl
    for i in range(count):
        m = next32() * rng_excl
        if (m & M32) < rng_excl:
            while (m & M32) < thresh:
                m = next32() * rng_excl
        out[i] = m >> 32
    return out, (state, inc, has_u32, u32)

"""Kernel-level timing on the GPU box (development tool; bench.py is the judged harness).

usage: python tools/kbench.py [reduce|resample|hist|peak ...]
"""

from __future__ import annotations

import ctypes as C
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch

from thermoextrap_b200 import _lib
from thermoextrap_b200 import kernels as K


def timeit(fn, warm=3, rep=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(rep):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e-3)
    return min(ts), sorted(ts)[len(ts) // 2]


def peak():
    lib = _lib.load()
    scratch = torch.zeros(8, dtype=torch.float64, device="cuda")
    fl = C.c_double()
    out = {}
    for bps in (4, 8):
        def run():
            _lib.check(lib.txb_fp64_peak_launch(20000, bps, scratch.data_ptr(), C.byref(fl), None), "peak")
        t, _ = timeit(run)
        out[f"fp64_tflops_bps{bps}"] = fl.value / t / 1e12
    for tile in (0, 1, 2, 3):
        for bps in (1, 2, 4):  # 8, 16, 32 warps per SM
            def run():
                _lib.check(lib.txb_fp64_tile_peak_launch(4000, bps, tile, scratch.data_ptr(), C.byref(fl), None), "tile peak")
            t, _ = timeit(run)
            out[f"fp64_{('tile4x14', 'tile2x14', 'dmma_m8n8k4', 'dmma_m16n8k16')[tile]}_tflops_{8 * bps}warps"] = fl.value / t / 1e12
    for nbins in (1024, 2048):
        for bps in (1, 2):
            def run():
                _lib.check(lib.txb_smem_atomic_peak_launch(4000, bps, nbins, scratch.data_ptr(), C.byref(fl), None), "atoms")
            t, _ = timeit(run)
            out[f"smem_atomics_per_s_{nbins}bins_{16 * bps}warps"] = fl.value / t
    print(json.dumps({"peak": out}))


def reduce():
    for n, nv, order in [(100_000_000, 0, 8), (100_000_000, 0, 6), (100_000_000, 0, 2), (4_000_000, 64, 6), (400_000, 2000, 4), (16_000_000, 16, 6)]:
        u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
        x = torch.randn((n, nv) if nv else (n,), dtype=torch.float64, device="cuda") * 0.02 + 1e-3 * (u[:, None] if nv else u)
        t, tm = timeit(lambda: K.reduce_vals(x, u, order))
        byts = 8 * n * (max(nv, 1) + 1)
        print(json.dumps({"reduce": [n, nv, order], "ms": t * 1e3, "ms_med": tm * 1e3, "GBs": byts / t / 1e9, "samples_per_s": n / t}))
        del x, u
    # x_is_u wide (lnPi shape)
    n, nv, order = 400_000, 2000, 4
    u = torch.randn((n, nv), dtype=torch.float64, device="cuda")
    t, tm = timeit(lambda: K.reduce_vals(None, u, order, x_is_u=True))
    print(json.dumps({"reduce_x_is_u": [n, nv, order], "ms": t * 1e3, "GBs": 8 * n * nv / t / 1e9}))


def sys_modes():
    import os
    return os.environ.get("KB_MODES", "0,1").split(",")


def reduce_wide():
    """A/B of the register-staged and the bulk-copy (TMA ring) wide kernels, kernel time from the profile hook."""
    import os
    lib = _lib.load()
    cases = [(4_000_000, 64, 6, False), (16_000_000, 64, 6, False), (400_000, 2000, 4, False), (16_000_000, 16, 6, False),
             (2_000_000, 130, 6, False), (400_000, 2000, 4, True), (4_000_000, 64, 8, False), (2_000_000, 64, 6, "upc"),
             (1_000_000, 500, 6, False), (400_000, 1000, 6, "upc"), (200_000, 4000, 4, False), (8_000_000, 32, 12, False)]
    if os.environ.get("KB_SHAPES"):
        cases = [tuple(int(v) for v in c.split(",")[:3]) + ({"0": False, "1": True, "upc": "upc"}[c.split(",")[3]],)
                 for c in os.environ["KB_SHAPES"].split(";")]
    for n, nv, order, kind in cases:
        u = torch.randn((n, nv) if kind else (n,), dtype=torch.float64, device="cuda") * 5.3 + 175
        x = None if kind is True else torch.randn((n, nv), dtype=torch.float64, device="cuda") * 0.02
        byts = 8 * n * nv * (2 if kind == "upc" else 1) + (0 if kind else 8 * n)
        row = {"reduce_wide": [n, nv, order, str(kind)]}
        outs = {}
        for mode in sys_modes():
            os.environ["TXB_WIDE_TMA"] = mode[0]
            os.environ["TXB_TMAP_L2"] = mode[1:] or "3"
            fn = (lambda: K.reduce_vals(None, u, order, x_is_u=True)) if kind is True else (lambda: K.reduce_vals(x, u, order))
            t, _ = timeit(fn)
            lib.txb_profile_begin(b"reduce")
            for _ in range(5):
                outs[mode] = fn()
            ms, nl = C.c_double(), C.c_int()
            lib.txb_profile_end(C.byref(ms), C.byref(nl))
            row["tma" + mode] = {"call_ms": t * 1e3, "kern_ms": ms.value / 5, "GBs": byts / (ms.value / 5 * 1e-3) / 1e9}
        row["maxdiff"] = max(float((outs["0"] - o).abs().max()) for o in outs.values())
        print(json.dumps(row), flush=True)
        del x, u
    os.environ.pop("TXB_WIDE_TMA", None)


def reduce_scalar():
    """A/B of the register-staged and ring (cp.async.bulk) scalar kernels."""
    import os
    lib = _lib.load()
    n = 100_000_000
    u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
    x = torch.randn(n, dtype=torch.float64, device="cuda") * 0.02
    w = torch.rand(n, dtype=torch.float64, device="cuda")
    for nn, order, hasx, hasw in [(n, 1, True, False), (n, 2, True, False), (n, 4, True, False), (n, 6, True, False), (n, 8, True, False),
                                  (n, 10, True, False), (n, 12, True, False), (n, 6, True, True), (n, 6, False, False),
                                  (10_000_000, 8, True, False), (n - 3, 6, True, False)]:
        xx, uu, ww = (x[:nn] if hasx else None), u[:nn], (w[:nn] if hasw else None)
        byts = 8 * nn * (1 + hasx + hasw)
        row = {"reduce_scalar": [nn, order, hasx, hasw]}
        outs = {}
        for mode in ("0", "1"):
            os.environ["TXB_SCALAR_TMA"] = mode
            fn = (lambda: K.reduce_vals(xx, uu, order, w=ww)) if hasx else (lambda: K.reduce_vals(None, uu, order, w=ww, x_is_u=True))
            timeit(fn)
            lib.txb_profile_begin(b"reduce")
            for _ in range(5):
                outs[mode] = fn()
            ms, nl = C.c_double(), C.c_int()
            lib.txb_profile_end(C.byref(ms), C.byref(nl))
            row["ring" + mode] = round(byts / (ms.value / 5 * 1e-3) / 1e9)
        row["maxdiff"] = float((outs["0"] - outs["1"]).abs().max())
        print(json.dumps(row), flush=True)
    os.environ.pop("TXB_SCALAR_TMA", None)


def reduce_big():
    for n, nv, order in [(1_000_000_000, 0, 8), (1_000_000_000, 0, 2), (50_000_000, 64, 6)]:
        u = torch.randn(n, dtype=torch.float64, device="cuda")
        u.mul_(5.3).add_(175)
        x = torch.randn((n, nv) if nv else (n,), dtype=torch.float64, device="cuda")
        t, tm = timeit(lambda: K.reduce_vals(x, u, order))
        byts = 8 * n * (max(nv, 1) + 1)
        print(json.dumps({"reduce_big": [n, nv, order], "ms": t * 1e3, "ms_med": tm * 1e3, "GBs": byts / t / 1e9, "samples_per_s": n / t}))
        del x, u


def ncu_case():
    """small fixed sequence for the ncu launch list"""
    n = 100_000_000
    u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
    x = torch.randn(n, dtype=torch.float64, device="cuda")
    K.reduce_vals(x, u, 8)
    xw = torch.randn((4_000_000, 64), dtype=torch.float64, device="cuda")
    K.reduce_vals(xw, u[:4_000_000], 6)
    freq = torch.randint(0, 4, (128, 10_000_000), device="cuda", dtype=torch.int32)
    K.resample_vals(x[:10_000_000], u[:10_000_000], freq.to(torch.int64), 8)
    K.resample_vals(x[:10_000_000], u[:10_000_000], freq.to(torch.uint8), 8)
    K.resample_vals(xw[:1_000_000], u[:1_000_000], freq[:, :1_000_000].to(torch.uint8), 6)
    torch.cuda.synchronize()


def ncu_full():
    """one launch of each hot kernel for `ncu --set full`"""
    n = 10_000_000
    u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
    x = torch.randn(n, dtype=torch.float64, device="cuda")
    idx = torch.randint(0, n, (16, n), device="cuda", dtype=torch.int64)
    K.resample_indices(x, u, idx, 8)
    f8 = K.indices_to_freq(idx, n, dtype=torch.uint8)
    f8 = f8.repeat(8, 1)
    K.resample_vals(x, u, f8, 8)
    K.resample_vals(x, u, f8.to(torch.int64), 8)
    xw = torch.randn((2_000_000, 64), dtype=torch.float64, device="cuda")
    K.reduce_vals(xw, u[:2_000_000], 6)
    K.resample_vals(xw[:500_000], u[:500_000], f8[:, :500_000].contiguous(), 6)
    xs = torch.randn(100_000_000, dtype=torch.float64, device="cuda")
    K.reduce_vals(xs, xs, 8)
    torch.cuda.synchronize()


def resample():
    for n, nv, order, nrep, dt in [
        (10_000_000, 0, 8, 128, torch.int64),
        (10_000_000, 0, 8, 128, torch.int32),
        (10_000_000, 0, 8, 128, torch.uint8),
        (10_000_000, 0, 8, 1024, torch.uint8),
        (10_000_000, 0, 2, 128, torch.int64),
        (1_000_000, 64, 6, 128, torch.uint8),
        (1_000_000, 64, 6, 1024, torch.uint8),
    ]:
        u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
        x = torch.randn((n, nv) if nv else (n,), dtype=torch.float64, device="cuda") * 0.02 + 1e-3 * (u[:, None] if nv else u)
        freq = torch.randint(0, 4, (nrep, n), device="cuda", dtype=torch.int32).to(dt)
        t, tm = timeit(lambda: K.resample_vals(x, u, freq, order), warm=2, rep=3)
        ncol = (order + 1) * (max(nv, 1) + 1)
        flops = 2.0 * nrep * n * ncol
        print(json.dumps({"resample": [n, nv, order, nrep, str(dt)], "ms": t * 1e3, "ms_med": tm * 1e3,
                          "TFLOPs": flops / t / 1e12, "freq_GBs": freq.element_size() * nrep * n / t / 1e9,
                          "reps_per_s": nrep / t}))
        del x, u, freq


def gather():
    for n, order, nrep in [(10_000_000, 8, 64), (10_000_000, 8, 256), (10_000_000, 2, 64), (1_000_000, 8, 256)]:
        u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
        x = torch.randn(n, dtype=torch.float64, device="cuda")
        idx = torch.randint(0, n, (nrep, n), device="cuda", dtype=torch.int64)
        t, tm = timeit(lambda: K.resample_indices(x, u, idx, order), warm=2, rep=3)
        print(json.dumps({"gather": [n, order, nrep], "ms": t * 1e3, "ms_med": tm * 1e3, "draws_per_s": nrep * n / t,
                          "idx_GBs": 8 * nrep * n / t / 1e9, "reps_per_s": nrep / t}))
        del x, u, idx


def sampler():
    import numpy as np
    for n, order, nrep in [(10_000_000, 8, 64), (10_000_000, 8, 256), (1_000_000, 8, 256), (100_000_000, 6, 16)]:
        u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
        x = torch.randn(n, dtype=torch.float64, device="cuda")
        rng = np.random.default_rng(0)
        t, tm = timeit(lambda: K.sampler_resample(rng, x, u, nrep, order, sync_rng=False), warm=2, rep=3)
        print(json.dumps({"sampler_fused": [n, order, nrep], "ms": t * 1e3, "ms_med": tm * 1e3, "draws_per_s": nrep * n / t, "reps_per_s": nrep / t}))
        lib = _lib.load()
        for tag in (b"sampler_count", b"bucket_scatter", b"bucket_consume"):
            lib.txb_profile_begin(tag)
            for _ in range(3):
                K.sampler_resample(rng, x, u, nrep, order, sync_rng=False)
            ms, nl = C.c_double(), C.c_int()
            lib.txb_profile_end(C.byref(ms), C.byref(nl))
            print(json.dumps({"kernel": tag.decode(), "ms": ms.value / max(nl.value, 1), "launches": nl.value}))
        t, tm = timeit(lambda: K.sampler_resample(rng, x, u, nrep, order, sync_rng=False, buckets=False), warm=2, rep=3)
        print(json.dumps({"sampler_gather": [n, order, nrep], "ms": t * 1e3, "draws_per_s": nrep * n / t}))
        t, tm = timeit(lambda: K.sampler_indices(rng, nrep, n), warm=1, rep=3)
        print(json.dumps({"sampler_fill": [n, nrep], "ms": t * 1e3, "draws_per_s": nrep * n / t, "GBs": 8 * nrep * n / t / 1e9}))
        t, tm = timeit(lambda: K.sampler_freq(rng, nrep, n, dtype=torch.uint8), warm=1, rep=3)
        print(json.dumps({"sampler_hist_u8": [n, nrep], "ms": t * 1e3, "draws_per_s": nrep * n / t}))
        del x, u


def hist():
    for nrep, n, dt in [(64, 10_000_000, torch.int64), (64, 10_000_000, torch.int32), (64, 10_000_000, torch.uint8)]:
        idx = torch.randint(0, n, (nrep, n), device="cuda", dtype=torch.int64)
        t, tm = timeit(lambda: K.indices_to_freq(idx, n, dtype=dt), warm=2, rep=3)
        print(json.dumps({"hist": [nrep, n, str(dt)], "ms": t * 1e3, "draws_per_s": nrep * n / t, "idx_GBs": 8 * nrep * n / t / 1e9}))


if __name__ == "__main__":
    which = sys.argv[1:] or ["peak", "reduce", "resample", "hist"]
    for w in which:
        globals()[w]()

"""Box probe (run under gpurun): host cores / RAM, pinned H2D / D2H bandwidth, int8 IMMA peak, and the wide
bootstrap's legs at the north_star target shape.  Prints one JSON object."""

from __future__ import annotations

import ctypes as C
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from thermoextrap_b200 import _lib  # noqa: E402
from thermoextrap_b200 import kernels as K  # noqa: E402


def ev_ms(fn, reps=3):
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1)
        best = t if best is None else min(best, t)
    return best


def main() -> None:
    out: dict = {}
    out["cpus"] = os.cpu_count()
    mem = {l.split(":")[0]: l.split(":")[1].strip() for l in open("/proc/meminfo") if ":" in l}
    out["mem_total"], out["mem_avail"] = mem.get("MemTotal"), mem.get("MemAvailable")
    out["gpus"] = torch.cuda.device_count()
    out["gpu"] = torch.cuda.get_device_name(0)
    dev = torch.device("cuda", 0)
    lib = _lib.load()

    # pinned copies
    nb = 4 << 30
    t0 = time.perf_counter()
    h = torch.empty(nb, dtype=torch.uint8).pin_memory()
    out["pin_4GiB_s"] = time.perf_counter() - t0
    d = torch.empty(nb, dtype=torch.uint8, device=dev)
    out["h2d_GBps"] = nb / ev_ms(lambda: d.copy_(h, non_blocking=True)) / 1e6
    out["d2h_GBps"] = nb / ev_ms(lambda: h.copy_(d, non_blocking=True)) / 1e6
    del h, d

    scratch = torch.zeros(8, dtype=torch.float64, device=dev)
    fl = C.c_double()
    for tile, name in ((2, "dmma884"), (4, "imma16832")):
        for bps in (1, 2, 4):
            ms = ev_ms(lambda: lib.txb_fp64_tile_peak_launch(20000, bps, tile, scratch.data_ptr(), C.byref(fl), None))
            out[f"{name}_tops_{8 * bps}warps"] = fl.value / ms / 1e9

    if "--no-target" not in sys.argv:
        n, c, k, nrep = 100_000_000, 64, 6, 1000
        g = torch.Generator(device=dev).manual_seed(2)
        u = torch.empty(n, dtype=torch.float64, device=dev).normal_(175.0, 5.3, generator=g)
        x = torch.empty((n, c), dtype=torch.float64, device=dev)
        slope = torch.linspace(0.5e-3, 2e-3, c, dtype=torch.float64, device=dev)
        for i in range(0, n, 1 << 22):
            xs = x[i : i + (1 << 22)]
            xs.normal_(0.0, 0.02, generator=g)
            xs.addcmul_(u[i : i + (1 << 22), None], slope[None, :])
        torch.cuda.synchronize()
        out["reduce_ms"] = ev_ms(lambda: K.reduce_vals(x, u, k))
        rng = np.random.default_rng(1)
        t0 = time.perf_counter()
        plan = K.SamplerPlan(rng, nrep, n)
        torch.cuda.synchronize()
        out["plan_ms"] = (time.perf_counter() - t0) * 1e3
        win = 128
        table = torch.empty((win, n), dtype=torch.uint8, device=dev)
        out["freq_128_ms"] = ev_ms(lambda: K.sampler_freq(plan, dtype=torch.uint8, reps=(0, win), out=table), reps=2)
        freq = K.sampler_freq(plan, dtype=torch.uint8, reps=(0, win), out=table)
        out["contract_128_ms"] = ev_ms(lambda: K.resample_vals(x, u, freq, k), reps=2)
        out["s0_only_128_ms"] = ev_ms(lambda: K.resample_vals(None, u, freq, k, x_is_u=True), reps=2)
        timing: dict = {}
        e = ev_ms(lambda: K.sampler_resample_wide(plan, x, u, order=k, timing=timing), reps=1)
        out["wide_total_ms"] = e
        out["wide_timing"] = timing
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

"""north_star target shape on ONE B200 (development tool; bench.py is the judged harness).

    order-6 comoment reduction + 1000-replicate bootstrap of 1e8 samples x 64 observable channels
    (one state point of BASELINE config 3; the wide-x path: K1 `k_reduce_wide*`, K6 `txb_sampler_freq`
    per replicate window, K2 `k_resample<RS_WIDE>`).

usage: python tools/target_c3.py [--nsamp 1e8] [--nchan 64] [--order 6] [--nrep 1000] [--table-gb G]
prints one JSON line (copied to profiles/ by hand).  Synthetic data generated on the device.
"""

from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import numpy as np
import torch

from bench import fp64_peak_tflops, measured_peaks
from thermoextrap_b200 import kernels as K


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nsamp", type=float, default=1e8)
    ap.add_argument("--nchan", type=int, default=64)
    ap.add_argument("--order", type=int, default=6)
    ap.add_argument("--nrep", type=int, default=1000)
    ap.add_argument("--table-gb", type=float, default=0.0, help="window table budget (0: a third of free memory)")
    a = ap.parse_args()
    n, c, k = int(a.nsamp), a.nchan, a.order
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(2)
    u = torch.empty(n, dtype=torch.float64, device=dev).normal_(175.0, 5.3, generator=g)
    slope = torch.linspace(0.5e-3, 2e-3, c, dtype=torch.float64, device=dev)
    x = torch.empty((n, c), dtype=torch.float64, device=dev)
    step = 1 << 22
    for i in range(0, n, step):  # chunked: no second 51 GB temporary
        xs = x[i : i + step]
        xs.normal_(0.0, 0.02, generator=g)
        xs.addcmul_(u[i : i + step, None], slope[None, :])
    torch.cuda.synchronize()

    def timed(fn, reps=3):
        best, out = None, None
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize()
            t = e0.elapsed_time(e1)
            best = t if best is None else min(best, t)
        return best, out

    K.reduce_vals(x[: 1 << 20], u[: 1 << 20], k)  # warm-up
    red_ms, base = timed(lambda: K.reduce_vals(x, u, k))
    red_bytes = 8.0 * n * (c + 1)

    rng = np.random.default_rng(1)
    t0 = time.perf_counter()
    plan = K.SamplerPlan(rng, a.nrep, n)
    torch.cuda.synchronize()
    plan_ms = (time.perf_counter() - t0) * 1e3

    timing: dict = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    boot = K.sampler_resample_wide(plan, x, u, order=k, table_bytes=int(a.table_gb * 1e9) or None, timing=timing)
    e1.record()
    torch.cuda.synchronize()
    boot_ms = e0.elapsed_time(e1)

    # size-independent checks: every replicate holds exactly nsamp draws; the replicate means scatter
    # around the sample mean with the standard error sigma/sqrt(n)
    w0 = boot[:, :, 0, 0]
    assert bool((w0 == float(n)).all()), "a replicate's total weight differs from nsamp"
    xm, xvar = base[:, 1, 0], x[: 1 << 20].var(dim=0)
    z = ((boot[:, :, 1, 0] - xm[None]) / (xvar / n).sqrt()[None]).abs().max().item()
    assert z < 6.5, f"bootstrap means {z:.1f} standard errors from the sample mean"

    flops = 2.0 * (k + 1) * (c + 1) * n * a.nrep
    peak_hbm, src = measured_peaks()
    peak64 = fp64_peak_tflops()
    line = {
        "workload": f"order-{k} reduction + {a.nrep}-replicate bootstrap of {n:.0e} samples x {c} channels (1 B200)",
        "reduce": {"ms": red_ms, "GBps": red_bytes / red_ms / 1e6, "frac_hbm": red_bytes / red_ms / 1e6 / peak_hbm,
                   "samples_per_s": n / red_ms * 1e3, "peak_source": src},
        "sampler_plan_ms": plan_ms,
        "bootstrap": {"ms": boot_ms, "freq_ms": timing["freq_ms"], "contract_ms": timing["contract_ms"],
                      "windows": timing["windows"], "replicates_per_s": a.nrep / boot_ms * 1e3,
                      "draws_per_s_freq": a.nrep * n / timing["freq_ms"] * 1e3,
                      "contract_tflops": flops / timing["contract_ms"] / 1e9,
                      "fp64_peak_tflops": peak64, "contract_frac_fp64": flops / timing["contract_ms"] / 1e9 / peak64,
                      "whole_bootstrap_frac_fp64": flops / boot_ms / 1e9 / peak64},
        "checks": {"weights_exact": True, "max_z_of_replicate_means": z},
        "gpu": torch.cuda.get_device_name(0),
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()

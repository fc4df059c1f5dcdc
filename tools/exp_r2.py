"""Round-2 A/B of kernel forms on one B200 (builder tool; results are copied under profiles/).

    python tools/exp_r2.py [--nrep 256] [--out gpurun_out/exp_r2.json]

Target shape (1e8 x 64, order 6): wide contraction forms 1 / 2 / 3 (txb_set_resample_kernel) and partition
kernel thread counts 512 / 1024 (txb_set_scatter_threads), per-kernel-family device time from the library's
own event brackets.  C2 shape (1e7 scalar, order 8, nrep 1000): partition kernel forms.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nrep", type=int, default=256)
    ap.add_argument("--nsamp", type=int, default=100_000_000)
    ap.add_argument("--out", default="gpurun_out/exp_r2.json")
    ap.add_argument("--skip-c2", action="store_true")
    a = ap.parse_args()

    import torch

    import bench
    from thermoextrap_b200 import _lib
    from thermoextrap_b200 import kernels as K

    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    res: dict = {"gpu": torch.cuda.get_device_name(0)}

    # ---- target shape
    ns = argparse.Namespace(nsamp=a.nsamp, nchan=64, order=6, nrep=a.nrep, steps=1, warmup=1)
    job = bench.Job(ns, 1, 0, dev)
    tags = ("reduce", "sampler_count", "bucket_scatter", "bucket_hist", "resample")
    ref = None
    rows = []
    for ver, nth in ((3, 512), (4, 512)):
        lib.txb_set_resample_kernel(ver)
        lib.txb_set_scatter_threads(nth)
        job.step()
        ms, out = bench.timed_ms(job.step, reps=2, warm=0)
        boot = out[1]
        if ref is None:
            ref = boot.clone()
        err = float(((boot - ref).abs() / ref.abs().clamp_min(1e-300)).max())
        kt = bench.kernel_times(lib, job, 1, tags)
        row = {"resample_kernel": ver, "scatter_threads": nth, "step_ms": ms, "nrep": a.nrep,
               "max_rel_diff_vs_first": err, "kernels_ms": {k: v["ms_per_step"] for k, v in kt.items()},
               "launches": {k: v["launches_per_step"] for k, v in kt.items()}}
        flops = 2.0 * 7 * 65 * a.nsamp * a.nrep
        row["resample_tflops"] = flops / (kt["resample"]["ms_per_step"] * 1e-3) / 1e12
        rows.append(row)
        print(json.dumps(row), flush=True)
    res["target"] = rows
    del job, ref, boot, out
    K._workspaces.clear()
    torch.cuda.empty_cache()

    # ---- C2 shape
    if not a.skip_c2:
        ns = argparse.Namespace(nsamp=10_000_000, nchan=0, order=8, nrep=1000, steps=1, warmup=1)
        job = bench.Job(ns, 1, 0, dev)
        rows = []
        for nth in (512, 1024):
            lib.txb_set_scatter_threads(nth)
            for _ in range(3):
                job.step()
            job.check()
            ms, _ = bench.timed_ms(lambda: [job.step() for _ in range(5)], reps=2, warm=0)
            job.check()
            kt = bench.kernel_times(lib, job, 3, ("sampler_count", "bucket_scatter", "bucket_consume", "sampler_gather"))
            row = {"scatter_threads": nth, "step_ms": ms / 5, "kernels_ms": {k: v["ms_per_step"] for k, v in kt.items()}}
            rows.append(row)
            print(json.dumps(row), flush=True)
        res["c2"] = rows
    res["fp64_peak_tflops"] = bench.fp64_peak_tflops()
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()

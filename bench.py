"""Headline benchmark: north_star's target job, ONE job sharded over N B200s (one process per GPU).

workload ("step") = one pass of thermoextrap's hot path over one batch of synthetic samples
(BASELINE.json north_star: order 6, 1e8 samples x 64 observable channels, 1000 bootstrap replicates
 = one state point of BASELINE config 3):
    K1  order-6 comoment reduction of x[1e8, 64], u[1e8]         (DataCentralMomentsVals.from_vals)
    K6  the reference sampler's index stream, nrep = 1000        (numpy PCG64, regenerated on the device)
    K2  1000-replicate bootstrap of the comoments                (.resample(sampler={"nrep": 1000}))
    K7  derivative expressions + Taylor predict per replicate    (beta.factory_extrapmodel().predict)  [e2e only]

metric: bootstrap replicates/s of the WHOLE job.  N > 1 is STRONG scaling: the same job, K1 sample-sharded
(all_reduce of pivot sums), the sampler's counting pass split over ranks (integer all_reduce), replicates
sharded (each rank bootstraps nrep / N replicates of the ONE numpy table), slabs all_gathered.
``value`` times K1 + K6 + K2 with the samples resident in HBM; ``e2e`` times the public API call chain from pinned
HOST arrays to HOST predictions (H2D + every kernel + K7 + D2H inside the timed region; with N > 1 every rank
uploads its 1/N row shard and the shards are exchanged over NVLink).  At N = 1 the API hides the upload: x travels in
chunks on a copy stream while the sampler kernels run and the first replicate windows contract each chunk as it
lands -- still every byte inside the timed region, counted in ``h2d_bytes_per_step``.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config target|c2]

``--config c2`` = BASELINE config 2 (order 8, 1e7 samples, scalar x, nrep 1000: the round-1 headline); its
record also rides along in the default line (``configs``), with configs 1, 4, 5 (N = 1 only, ``--no-extras`` skips).
``--impl reference`` times the reference's CPU implementation of the same path (numba port of cmomy's push
algorithm + live numpy sampler, oracle/cmomy_port.py) on the host cores, on a bounded sample of the workload.
"""

from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "bootstrap_replicates_per_s"
UNIT = "replicates/s"
BETA0, BETAS = 1.0, [0.9, 0.95, 1.05, 1.1]
SHAPES = {
    "target": dict(nsamp=100_000_000, nchan=64, order=6, nrep=1000),
    "c2": dict(nsamp=10_000_000, nchan=0, order=8, nrep=1000),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="target", choices=sorted(SHAPES))
    ap.add_argument("--nsamp", type=int, default=None, help=argparse.SUPPRESS)
    ap.add_argument("--nchan", type=int, default=None, help=argparse.SUPPRESS)
    ap.add_argument("--nrep", type=int, default=None, help=argparse.SUPPRESS)
    ap.add_argument("--order", type=int, default=None, help=argparse.SUPPRESS)
    ap.add_argument("--cpu-seconds", type=float, default=6.0, help=argparse.SUPPRESS)
    ap.add_argument("--e2e-steps", type=int, default=3, help="timed end-to-end steps (each moves the whole sample set over PCIe)")
    ap.add_argument("--no-extras", action="store_true", help="skip the per-config records (N = 1 only)")
    ap.add_argument("--no-wide", action="store_true", help=argparse.SUPPRESS)  # round-1 flag, accepted
    a = ap.parse_args()
    for k, v in SHAPES[a.config].items():
        if getattr(a, k) is None:
            setattr(a, k, v)
    return a


def workload_name(a) -> str:
    if a.nchan:
        return (f"north_star target: order-{a.order} comoment reduction + {a.nrep}-replicate bootstrap of "
                f"{a.nsamp:.0e} samples x {a.nchan} observable channels with the reference sampler's (numpy PCG64) "
                "index stream (one state point of BASELINE config 3)")
    return (f"C2: beta extrapolation order {a.order}, {a.nsamp:.0e} samples, scalar x, "
            f"bootstrap nrep={a.nrep} with the reference sampler's (numpy PCG64) index stream")


def config_of(a) -> dict:
    """Identical for both arms (the driver compares them)."""
    return {"workload": workload_name(a), "order": a.order, "nsamp": a.nsamp, "nchan": a.nchan, "nrep": a.nrep,
            "l2": "inputs larger than L2 (samples 51 GB / 160 MB per step; frequency table regenerated every step)"}


def make_samples_host(n: int, c: int, seed: int):
    """u ~ N(175, 5.3^2), x_c = a_c u + N(0, 0.02^2), a_c in [0.5e-3, 2e-3]  (notebook statistics)."""
    import numpy as np

    rng = np.random.default_rng(seed)
    u = rng.normal(175.0, 5.3, n)
    if not c:
        return 1e-3 * u + rng.normal(0.0, 0.02, n), u
    slope = np.linspace(0.5e-3, 2e-3, c)
    return u[:, None] * slope[None, :] + rng.normal(0.0, 0.02, (n, c)), u


def make_samples_device(n: int, c: int, seed: int, dev):
    """Same distribution, generated on the device (torch Philox) -- identical on every rank for one seed."""
    import torch

    g = torch.Generator(device=dev).manual_seed(seed)
    u = torch.empty(n, dtype=torch.float64, device=dev).normal_(175.0, 5.3, generator=g)
    if not c:
        x = torch.empty(n, dtype=torch.float64, device=dev).normal_(0.0, 0.02, generator=g)
        x.add_(u, alpha=1e-3)
        return x, u
    slope = torch.linspace(0.5e-3, 2e-3, c, dtype=torch.float64, device=dev)
    x = torch.empty((n, c), dtype=torch.float64, device=dev)
    step = 1 << 22
    for i in range(0, n, step):  # chunked: no second 51 GB temporary
        xs = x[i : i + step]
        xs.normal_(0.0, 0.02, generator=g)
        xs.addcmul_(u[i : i + step, None], slope[None, :])
    return x, u


# --------------------------------------------------------------------------- CPU baseline / reference arm
def cpu_reference_step(a, seconds: float):
    """The reference path on the host cores on a BOUNDED sample of the workload: ``n_sub`` of the ``nsamp`` samples
    (all channels), ``r`` replicates.  One replicate of the full workload costs ``nsamp / n_sub`` sample-replicates,
    so ``value = r * (n_sub / nsamp) / t`` in replicates/s of the full-size job."""
    import numba
    import numpy as np

    from oracle import cmomy_port as cp

    cores = numba.get_num_threads()
    n_sub = min(a.nsamp, 1_000_000 if a.nchan else 10_000_000)
    x, u = make_samples_host(n_sub, a.nchan, 0)
    cp.bootstrap_reference_path(x[:1000], u[:1000], a.order, 2, np.random.default_rng(0))  # JIT
    cp.reduce_vals(x[:1000], u[:1000], a.order)
    # calibrate the replicate count for ~`seconds` per step (wide x parallelises over replicate x channel)
    r0 = 1 if a.nchan else cores
    t0 = time.perf_counter()
    cp.bootstrap_reference_path(x, u, a.order, r0, np.random.default_rng(1))
    t1 = (time.perf_counter() - t0) / r0
    nrep = int(max(1, min(a.nrep, seconds / t1)))
    if not a.nchan:
        nrep = max(cores, nrep // cores * cores)

    def step():
        t0 = time.perf_counter()
        cp.reduce_vals(x, u, a.order)
        cp.bootstrap_reference_path(x, u, a.order, nrep, np.random.default_rng(1))
        return time.perf_counter() - t0

    sample = (f"{n_sub} of {a.nsamp} samples x {max(a.nchan, 1)} channel(s) x {nrep} of {a.nrep} replicates per step: "
              "K1 push reduce + numpy choice table + indices_to_freq + K2 freq-weighted push (numba port of cmomy, "
              f"prange over {'replicate x channel' if a.nchan else 'replicates'}, {cores} threads); rate scaled to "
              "full-width replicates: value = nrep_step * (n_sub / nsamp) / t")
    return step, nrep * (n_sub / a.nsamp), cores, sample


def run_reference(a) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    step, units, cores, sample = cpu_reference_step(a, a.cpu_seconds)
    warm = max(a.warmup, 3)
    for _ in range(warm):
        step()
    ts = [step() for _ in range(max(1, a.steps))]
    t = sum(ts) / len(ts)
    value = units / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
        "steps": len(ts), "warmup": warm, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(a),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "best_value": units / min(ts)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region (NVML in a thread: the same counters
    `nvidia-smi --query-gpu=clocks.sm,clocks_event_reasons.*` prints, without a second process)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int, period: float = 0.2) -> None:
        self.rows: list[tuple[int, int]] = []
        self.ok = False
        self._stop = threading.Event()
        try:
            import pynvml

            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v for v in vis.split(",") if v.strip().isdigit()]
            phys = int(ids[index]) if index < len(ids) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as exc:  # noqa: BLE001
            self.err = repr(exc)
            return
        self.period = period
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self) -> None:
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                why = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((mhz, why))
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(self.period)

    def stop(self) -> dict:
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [f"NVML unavailable: {self.err}"]}
        self._stop.set()
        self.thread.join(timeout=2)
        sm = [r[0] for r in self.rows]
        reasons = sorted({name for _, why in self.rows for bit, name in self.REASONS.items() if why & bit})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self.max_mhz, "samples": len(sm),
                "reasons": reasons}


def measured_peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    out = {"hbm_gbs": 6650.0, "bf16_tflops": None, "source": "fallback (B200_PROFILING.md 6.65 TB/s)"}
    if p.exists():
        try:
            d = json.loads(p.read_text())
            out = {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": d.get("bf16_tflops"),
                   "bf16_tflops_sustained": d.get("bf16_tflops_sustained"), "source": "measured (MEASURED_PEAKS.json)"}
        except (KeyError, ValueError):
            pass
    return out


def fp64_peak_tflops(tile: int = -1) -> float:
    """fp64 peak measured in this run (MEASURED_PEAKS.json has no fp64 figure): DFMA micro-benchmark, or with
    ``tile=2`` the DMMA m8n8k4 tensor form (both ~37 TFLOP/s on a B200)."""
    import torch

    from thermoextrap_b200 import _lib

    lib = _lib.load()
    scratch = torch.zeros(8, dtype=torch.float64, device="cuda")
    fl = C.c_double()
    best = 0.0
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if tile < 0:
            lib.txb_fp64_peak_launch(20000, 8, scratch.data_ptr(), C.byref(fl), None)
        else:
            lib.txb_fp64_tile_peak_launch(20000, 4, tile, scratch.data_ptr(), C.byref(fl), None)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, fl.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def timed_ms(fn, reps: int = 3, warm: int = 1):
    import torch

    for _ in range(warm):
        fn()
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


# --------------------------------------------------------------------------- B200 arm
class Job:
    """Device-resident form of one step on this rank (N = 1: plain kernels; N > 1: the sharded forms)."""

    def __init__(self, a, world, rank, dev):
        import numpy as np

        from thermoextrap_b200 import kernels as K
        from thermoextrap_b200 import parallel as par

        self.a, self.world, self.rank, self.K, self.par, self.np = a, world, rank, K, par, np
        # every rank holds the whole sample set (replicate sharding); generated from ONE seed on every rank
        self.x, self.u = make_samples_device(a.nsamp, a.nchan, 100, dev)
        self.lo, self.hi = par.shard_range(a.nsamp, rank, world)
        self.pending: list = []

    def step(self):
        a, K, par = self.a, self.K, self.par
        rng = self.np.random.default_rng(1)
        if self.world == 1:
            c = K.reduce_vals(self.x, self.u, a.order)
            if a.nchan:
                b = K.sampler_resample_wide(rng, self.x, self.u, a.nrep, a.order)
            else:
                b = K.sampler_resample(rng, self.x, self.u, a.nrep, a.order, sync_rng=False, status_out=self.pending)
            return c, b
        c = par.reduce_vals_sharded(self.x[self.lo : self.hi], self.u[self.lo : self.hi], a.order)
        b = par.bootstrap_sharded(rng, self.x, self.u, a.nrep, a.order)
        return c, b

    def check(self):
        self.K.check_deferred(self.pending)


def kernel_times(lib, job, steps: int, tags) -> dict:
    """Device time of each kernel family over `steps` more steps, one family bracketed at a time (CUDA events on
    the launching stream inside libtxb; the hook adds two event records per launch)."""
    from thermoextrap_b200 import _lib

    out = {}
    for tag in tags:
        lib.txb_profile_begin(tag.encode())
        for _ in range(steps):
            job.step()
        ms, nl = C.c_double(), C.c_int()
        _lib.check(lib.txb_profile_end(C.byref(ms), C.byref(nl)), "txb_profile_end")
        if nl.value:
            out[tag] = {"ms_per_step": ms.value / steps, "launches_per_step": nl.value / steps}
        job.check()
    return out


def run_b200(a) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    import thermoextrap_b200 as xt
    from thermoextrap_b200 import _lib, backend
    from thermoextrap_b200 import kernels as K
    from thermoextrap_b200 import parallel as par

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    clocks = ClockSampler(local) if rank == 0 else None
    wide = bool(a.nchan)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    job = Job(a, world, rank, dev)
    warm = max(a.warmup, 3)
    for _ in range(warm):
        job.step()
    barrier()
    if clocks is not None:  # NVML start-up perturbs the GPU: wait until samples are flowing
        t_wait = time.perf_counter()
        while not clocks.rows and time.perf_counter() - t_wait < 5.0:
            time.sleep(0.05)
    job.step()  # every rank: the wait above idles the GPU, bring the clocks back up before timing
    barrier()
    if clocks is not None:
        clocks.rows.clear()
    job.check()
    launches0 = K.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(a.steps):
        comom, boot = job.step()
    ev1.record()
    barrier()
    t_dev = ev0.elapsed_time(ev1) * 1e-3
    job.check()   # raises if any timed step overflowed a bucket list / ring (one sync, after timing)
    launches = K.launch_count - launches0
    clock_info = clocks.stop() if clocks else None

    # size-independent checks of the last step (no oracle can run at this size): every replicate holds exactly
    # nsamp draws; the replicate means scatter around the sample mean with the standard error sigma / sqrt(n)
    nv = max(a.nchan, 1)
    bt = boot.reshape(a.nrep, nv, 2, a.order + 1)
    ct = comom.reshape(nv, 2, a.order + 1)
    assert bool((bt[:, :, 0, 0] == float(a.nsamp)).all()), "a replicate's total weight differs from nsamp"
    sig = ((job.x[: 1 << 20].reshape(-1, nv)).var(dim=0) / a.nsamp).sqrt()
    z = float(((bt[:, :, 1, 0] - ct[None, :, 1, 0]) / sig[None]).abs().max())
    assert z < 6.5, f"bootstrap means {z:.1f} standard errors from the sample mean"

    # per-kernel device time (outside the headline region)
    ksteps = max(1, min(a.steps, 2 if wide else a.steps))
    tags = ("reduce", "sampler_count", "bucket_scatter", "bucket_hist", "resample", "resample.wide") if wide else (
        "reduce", "sampler_count", "bucket_scatter", "bucket_consume", "sampler_gather")
    kernels = kernel_times(lib, job, ksteps, tags)

    # ---- end to end through the public API: pinned host arrays in, host predictions out.
    # N > 1: the API runs in N-GPU mode -- every rank owns (and uploads) its 1/N row shard of the samples
    lo, hi = job.lo, job.hi
    xh = torch.empty((hi - lo, a.nchan) if wide else (hi - lo,), dtype=torch.float64).pin_memory()
    uh = torch.empty(hi - lo, dtype=torch.float64).pin_memory()
    xh.copy_(job.x[lo:hi])
    uh.copy_(job.u[lo:hi])
    torch.cuda.synchronize()
    del job.x, job.u, comom, boot, bt, ct   # hand the device-resident copy back before the API leg
    K._workspaces.clear()
    torch.cuda.empty_cache()
    xv = xt.DataArray(xh.numpy(), dims=("rec", "val") if wide else ("rec",))
    uv = xt.DataArray(uh.numpy(), dims=("rec",))
    if world > 1:
        par.enable()

    def step_e2e():
        backend.clear_device_cache()  # every step uploads its own inputs
        data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=a.order, central=True)
        model = xt.beta.factory_extrapmodel(beta=BETA0, data=data)
        b = model.resample(sampler={"nrep": a.nrep, "rng": np.random.default_rng(1)})
        return model.predict(BETAS).values, b.predict(BETAS).values

    e2e_steps = max(1, min(a.steps, a.e2e_steps)) if wide else a.steps
    for _ in range(2 if wide else warm):
        pred, pred_boot = step_e2e()
    barrier()
    backend.transfer_bytes.update(h2d=0, d2h=0)
    t0 = time.perf_counter()
    per = []
    for _ in range(e2e_steps):
        ts = time.perf_counter()
        pred, pred_boot = step_e2e()
        per.append(time.perf_counter() - ts)
    barrier()
    t_e2e = time.perf_counter() - t0
    if rank == 0:
        print("e2e ms per step: " + " ".join(f"{1e3 * t:.1f}" for t in per), file=sys.stderr)
    h2d, d2h = backend.transfer_bytes["h2d"], backend.transfer_bytes["d2h"]
    assert pred_boot.shape[:2] == (len(BETAS), a.nrep) and np.isfinite(pred_boot).all()
    if world > 1:
        par.disable()
    backend.clear_device_cache()

    stats = torch.tensor([t_dev, t_e2e, float(h2d), float(d2h)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = stats[:2].clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tot = stats[2:].clone()
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        stats = torch.cat([tmax, tot])
    t_dev, t_e2e, h2d, d2h = (float(v) for v in stats.tolist())

    if rank == 0:
        emit(a, world, t_dev, t_e2e, e2e_steps, kernels, launches, h2d / e2e_steps, d2h / e2e_steps, clock_info, z, dev)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


#: MEASURED_PEAKS.json's HBM figure is a COPY (half reads, half writes); a read-only stream (K1 reads every sample
#: once and writes kilobytes) runs above it on HBM3e (nominal 7.7 TB/s), so its fraction can exceed 1
HBM_NOTE = ("peak = measured copy bandwidth (read + write, MEASURED_PEAKS.json); this kernel only reads, and read-only "
            "streams exceed the copy figure on HBM3e (nominal 7.7 TB/s)")
ALG_BYTES_PER_DRAW = {"sampler_gather": 16.0, "bucket_scatter": 2.0, "bucket_consume": 2.0, "sampler_count": 0.0}
#: DRAM bytes per draw measured by `ncu --set full` (profiles/r01_bucket_kernels_ncu_full_s21.txt, r01_gather_ncu_full.txt)
NCU_DRAM_BYTES_PER_DRAW = {"bucket_scatter": (0.001503 + 1.255659) * 1e9 / 6.4e8,
                           "bucket_consume": (1.550797 + 0.009994) * 1e9 / 6.4e8,
                           "sampler_count": 0.0, "sampler_gather": 60.0}


def emit(a, world, t_dev, t_e2e, e2e_steps, kernels, launches, h2d, d2h, clock_info, z, dev) -> None:
    import torch

    from thermoextrap_b200 import kernels as K

    peaks = measured_peaks()
    step_ms = t_dev / a.steps * 1e3
    nv = max(a.nchan, 1)
    wide = bool(a.nchan)
    flops = 2.0 * (a.order + 1) * (nv + 1) * a.nsamp * a.nrep   # 2 (k+1) (C_u + C_x) per (rep, sample)
    fp64_peak = fp64_peak_tflops()
    shares = {k: {**v, "share_of_step": v["ms_per_step"] / step_ms} for k, v in kernels.items()}
    if "resample.wide" in shares and "resample" in shares:   # "resample" = wide + scalar: list them side by side
        shares["resample.scalar"] = {
            "ms_per_step": shares["resample"]["ms_per_step"] - shares["resample.wide"]["ms_per_step"],
            "launches_per_step": shares["resample"]["launches_per_step"] - shares["resample.wide"]["launches_per_step"]}
        shares["resample.scalar"]["share_of_step"] = shares["resample.scalar"]["ms_per_step"] / step_ms
        del shares["resample"]
    if wide:
        # dominant kernel: the wide-x DMMA contraction alone (tag "resample.wide"); its algorithmic flops are the
        # x columns' 2 (k+1) C_x per (rep, sample); the u-moment columns (2 (k+1) C_u) are k_resample_mma_scalar's
        rs_all = kernels.get("resample", {"ms_per_step": float("nan"), "launches_per_step": 1})
        rs = kernels.get("resample.wide", rs_all)
        flops_wide = 2.0 * (a.order + 1) * nv * a.nsamp * a.nrep
        per_launch_s = rs["ms_per_step"] * 1e-3 / rs["launches_per_step"]
        ach = flops_wide / world / (rs["ms_per_step"] * 1e-3) / 1e12   # this rank contracts nrep / world replicates
        # DRAM traffic of one launch from `ncu --set full` (profiles/r02_target_kernels_ncu_full.txt: 64.88 GB read +
        # 0.03 GB written for a 128-replicate window = the window's table + one pass over the samples), scaled to
        # this run's replicates per launch
        reps_per_launch = a.nrep / world / rs["launches_per_step"]
        traffic = a.nsamp * reps_per_launch + 8.0 * a.nsamp * (nv + 1) * (1 if reps_per_launch <= 128 else reps_per_launch / 128)
        roofline = {
            "kernel": "k_resample_mma2<6, uint8, 16 warps> (K2: frequency table x shifted power products on the fp64 "
                      "tensor pipe, DMMA.8x8x4)",
            "bound": "tensor", "achieved": ach, "peak": fp64_peak, "unit": "TFLOP/s", "frac": ach / fp64_peak,
            "peak_source": "fp64 DFMA / DMMA micro-benchmark of this run (txb_fp64_peak_launch; MEASURED_PEAKS.json "
                           "carries bf16 only, and tcgen05 has no fp64 kind: mma.sync DMMA is the fp64 tensor path)",
            "algorithmic_flops_per_launch": flops_wide / world / rs["launches_per_step"],
            "launch_ms": per_launch_s * 1e3, "launches_per_step": rs["launches_per_step"],
            "share_of_step": rs["ms_per_step"] / step_ms,
            "traffic": traffic,
            "traffic_source": "table bytes + one pass over the samples per 128-replicate tile; ncu --set full of a "
                              "128-replicate launch measured 64.9 GB against this formula's 64.8 GB "
                              "(profiles/r02_target_kernels_ncu_full.txt)",
            "whole_step_frac_fp64": flops / world / (step_ms * 1e-3) / 1e12 / fp64_peak,
            "kernels": shares,
        }
    else:
        top = max(kernels, key=lambda k: kernels[k]["ms_per_step"] if k != "reduce" else -1.0)
        k_s = kernels[top]["ms_per_step"] * 1e-3 / kernels[top]["launches_per_step"]
        draws_per_launch = a.nrep * a.nsamp / world / kernels[top]["launches_per_step"]
        fp64_ms = kernels.get("bucket_consume", kernels.get("sampler_gather", {"ms_per_step": float("nan")}))["ms_per_step"]
        roofline = {
            "kernel": top, "bound": "issue (integer PCG64 + shared-memory atomics), reported against HBM as the contract asks",
            "achieved": ALG_BYTES_PER_DRAW[top] * draws_per_launch / k_s / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": ALG_BYTES_PER_DRAW[top] * draws_per_launch / k_s / 1e9 / peaks["hbm_gbs"],
            "peak_source": peaks["source"], "traffic": NCU_DRAM_BYTES_PER_DRAW[top] * draws_per_launch,
            "launch_ms": k_s * 1e3, "launches_per_step": kernels[top]["launches_per_step"],
            "share_of_step": kernels[top]["ms_per_step"] / step_ms,
            "whole_step_frac_fp64": flops / world / (step_ms * 1e-3) / 1e12 / fp64_peak,
            "fp64": {"achieved_tflops": flops / world / (fp64_ms * 1e-3) / 1e12, "peak_tflops": fp64_peak},
            "kernels": shares,
        }
    red = kernels.get("reduce")
    red_bytes = 8.0 * a.nsamp * (nv + 1) / world
    reduce_rec = None
    if red:
        red_s = red["ms_per_step"] * 1e-3
        reduce_rec = {"samples_per_s": a.nsamp / red_s, "kernel_ms_per_step": red["ms_per_step"],
                      "GBps_per_gpu": red_bytes / red_s / 1e9, "frac_hbm": red_bytes / red_s / 1e9 / peaks["hbm_gbs"],
                      "peak_note": HBM_NOTE}
    cpu = None
    if world == 1:
        step, units, cores, sample = cpu_reference_step(a, a.cpu_seconds)
        ts = [step() for _ in range(3)]
        cpu = {"value": units / min(ts), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": sample + f"; best of 3 passes ({min(ts):.1f} s)"}
    sharding = ("one GPU" if world == 1 else
                f"ONE job over {world} GPUs: K1 sample-sharded (all_reduce of pivot sums), sampler counting pass split "
                "by thread block (integer all_reduce), replicates sharded (nrep / N per rank), slabs all_gathered")
    line = {
        "metric": METRIC, "value": a.nrep * a.steps / t_dev, "unit": UNIT, "n_gpus": world,
        "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": step_ms,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {**config_of(a), "sharding": sharding,
                   "step": "K1 reduce_vals + K6 device sampler (counting pass + frequency windows) + K2 bootstrap "
                           "(+finalize), samples resident in HBM; a fresh default_rng(1) table every step"},
        "e2e": {"value": a.nrep * e2e_steps / t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": t_e2e / e2e_steps * 1e3, "steps": e2e_steps,
                "path": "DataCentralMomentsVals.from_vals -> beta.factory_extrapmodel -> .resample({'nrep','rng'}) -> "
                        ".predict, from pinned host arrays" + ("" if world == 1 else
                        f" (parallel.enable(): each rank uploads its 1/{world} row shard, shards all_gathered over NVLink)"),
                "upload": ("x in 16 row chunks on a copy stream: the sampler's integer kernels and the first replicate "
                           f"windows' contraction run behind it chunk by chunk ({K.follow_windows} windows followed an "
                           "upload in this run); K1 deferred to first use") if K.follow_windows else
                          "one copy on the compute stream ahead of the kernels"},
        "gpu_launches": int(launches),
        "samples_per_s_reduce": None if not reduce_rec else reduce_rec["samples_per_s"],
        "roofline": roofline,
        "reduce": reduce_rec,
        "checks": {"weights_exact": True, "max_z_of_replicate_means": z},
        "cpu_baseline": cpu,
        "clocks": clock_info,
        "gpu": torch.cuda.get_device_name(0),
    }
    if world == 1 and not a.no_extras:
        line["configs"] = extra_configs(a, dev)
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- the other BASELINE configs (N = 1)
def extra_configs(a, dev) -> dict:
    """BASELINE configs 1-5 as short records beside the headline (each bounded to seconds; outside the timed
    region).  Config 3 (5 states x nrep 1e4 of the headline shape, ~50x the headline job) rides along at 1/20 of its
    replicates; the full job on 8 GPUs is tools/config3.py under torchrun, recorded under profiles/."""
    import torch

    from thermoextrap_b200 import backend
    from thermoextrap_b200 import kernels as K

    out = {}
    for name, fn in (("c1", config_c1), ("c2", config_c2), ("c3", config_c3), ("c4", config_c4), ("c5", config_c5),
                     ("perturb", config_perturb)):
        if name == a.config:
            continue
        try:
            backend.clear_device_cache()
            K._workspaces.clear()
            torch.cuda.empty_cache()
            out[name] = fn(dev)
        except Exception as exc:  # noqa: BLE001  (never lose the headline line to an extra record)
            out[name] = {"error": repr(exc)}
    return out


def config_c1(dev) -> dict:
    """Ideal-gas beta extrapolation through the public API: order 6, 1e5 samples, scalar x."""
    import numpy as np

    import thermoextrap_b200 as xt
    from thermoextrap_b200 import backend

    beta0, order, betas = 5.6, 6, np.array([5.1, 5.5, 5.7, 6.1])
    xdata, udata = xt.idealgas.generate_data((100_000, 1000), beta0, 1.0, rng=np.random.default_rng(0))
    xv, uv = xt.DataArray(xdata, dims=("rec",)), xt.DataArray(udata, dims=("rec",))

    def job():
        backend.clear_device_cache()
        data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
        return xt.beta.factory_extrapmodel(beta=beta0, data=data).predict(betas).values

    job()
    ts = []
    for _ in range(10):
        t0 = time.perf_counter()
        pred = job()
        ts.append(time.perf_counter() - t0)
    true, _ = xt.idealgas.x_beta_extrap(order, beta0, betas, 1.0)
    return {"workload": "C1: ideal-gas beta extrapolation, order 6, 1e5 samples, scalar x, from_vals + "
                        "factory_extrapmodel + predict through the public API (host arrays in, host out)",
            "ms_per_call": min(ts) * 1e3, "samples_per_s": 1e5 / min(ts),
            "max_abs_err_vs_analytic": float(np.max(np.abs(pred - true)))}


def config_c2(dev) -> dict:
    """BASELINE config 2 (the round-1 headline): order 8, 1e7 samples, scalar x, nrep 1000."""
    import argparse as _ap

    from thermoextrap_b200 import _lib

    b = _ap.Namespace(nsamp=10_000_000, nchan=0, order=8, nrep=1000, steps=5, warmup=3)
    job = Job(b, 1, 0, dev)
    for _ in range(3):
        job.step()
    job.check()
    ms, _ = timed_ms(lambda: [job.step() for _ in range(5)], reps=2, warm=0)
    job.check()
    kern = kernel_times(_lib.load(), job, 3, ("reduce", "sampler_count", "bucket_scatter", "bucket_consume", "sampler_gather"))
    step_ms = ms / 5
    flops = 4.0 * 9 * 1e7 * 1000
    return {"workload": workload_name(b), "ms_per_step": step_ms, "replicates_per_s": 1000 / step_ms * 1e3,
            "samples_per_s_reduce": None if "reduce" not in kern else 1e7 / (kern["reduce"]["ms_per_step"] * 1e-3),
            "whole_step_frac_fp64": flops / (step_ms * 1e-3) / 1e12 / fp64_peak_tflops(),
            "bound": "instruction issue (partition: integer PCG64 + ring bookkeeping, 44 % issue-active, ALU 35 % + IMAD "
                     "17 % of the pipes) and the fp64 pipe (consume: 26 fp64 instructions per draw for 18 counted FMAs); "
                     "profiles/r02_target_kernels_ncu_full.txt",
            "note": "device-timed steps run with sync_rng=False and a deferred status word (no host sync inside the timed "
                    "region); the public API call pays one D2H of raw_used per call to hand the generator state back",
            "kernels": {k: {**v, "share_of_step": v["ms_per_step"] / step_ms} for k, v in kern.items()}}


def config_c3(dev) -> dict:
    """BASELINE config 3 through the public API (5-state beta InterpModel, order 6, 1e8 x 64 per state) on ONE GPU at
    nrep = 500 per state -- 1/20 of the config's 1e4 replicates, the only thing reduced (tools/config3.py)."""
    sys.path.insert(0, str(ROOT / "tools"))
    import config3

    rec = config3.run(100_000_000, 64, 6, 500, 5)
    rec["note"] = ("nrep 500 of the config's 10000 per state (1/20), everything else at full size; the full config on 8 GPUs: "
                   "profiles/r02_config3_8gpu.json")
    return rec


def config_c4(dev) -> dict:
    """lnPi width: K1 on uv[rec = 1e6, n = 2000] (x_is_u, order 4) and the extrapolation through
    lnpi.factory_extrapmodel_lnPi on the reduced moments."""
    import numpy as np
    import torch

    import thermoextrap_b200 as xt
    from thermoextrap_b200 import kernels as K

    n, c, k = 1_000_000, 2000, 4
    u2 = torch.empty((n, c), dtype=torch.float64, device=dev)
    col = torch.arange(1, c + 1, device=dev, dtype=torch.float64)
    for i in range(0, n, 1 << 16):
        blk = u2[i : i + (1 << 16)]
        blk.normal_(0.0, 1.0)
        blk.mul_(0.9 * col.sqrt()).sub_(4.8 * col)
    ms, out = timed_ms(lambda: K.reduce_vals(None, u2, k, x_is_u=True), reps=5, warm=2)
    peaks = measured_peaks()
    gbs = 8.0 * n * c / ms / 1e6
    # the model: comoments -> DataCentralMoments(x_is_u) -> lnPi extrapolation at 5 betas
    com = out.cpu().numpy()                                    # [n=2000, 2, k+1]
    t0 = time.perf_counter()
    data = xt.DataCentralMoments.from_data(xt.DataArray(np.concatenate([com[:, 0, :], com[:, 1, -1:]], axis=-1),
                                                        dims=("n", "umom")), x_is_u=True, central=True)
    lnpi0 = xt.DataArray(-0.5 * (np.arange(c) - c / 2.0) ** 2 / c, dims=("n",))
    meta = xt.lnpi.lnPiDataCallback(lnpi0, xt.DataArray(np.array([-2.63]), dims=("comp",)), dims_n=["n"], dims_comp="comp")
    model = xt.lnpi.factory_extrapmodel_lnPi(beta=1.0, data=data.new_like(meta=meta))
    pred = model.predict([0.9, 0.95, 1.0, 1.05, 1.1]).values
    t_model = time.perf_counter() - t0
    assert pred.shape == (5, c) and np.isfinite(pred).all()
    del u2
    return {"workload": "C4: lnPi macrostate extrapolation, order 4, 1e6 frames x 2000 macrostate channels",
            "reduce_ms": ms, "frames_per_s": n / ms * 1e3, "GBps": gbs, "frac_hbm": gbs / peaks["hbm_gbs"],
            "algorithmic_bytes": 8.0 * n * c, "model_predict_ms": t_model * 1e3, "peak_note": HBM_NOTE}


def config_perturb(dev) -> dict:
    """PerturbModel.predict kernel at three widths: device time and the fraction of the HBM
    roofline (algorithmic bytes: one pass over x and u per tile of 4 alphas, plus the min/max pass over u)."""
    import numpy as np
    import torch

    from thermoextrap_b200 import kernels as K

    peaks = measured_peaks()
    out = {"workload": "PerturbModel.predict (models.py:1009-1046): exp(-dalpha u) reweighting of x[rec, val]", "cases": []}
    for n, c, na in [(100_000_000, 0, 8), (4_000_000, 64, 8), (400_000, 2000, 4)]:
        u = torch.randn(n, dtype=torch.float64, device=dev) * 1.5 + 3.0
        x = torch.randn((n, c) if c else (n,), dtype=torch.float64, device=dev)
        da = np.linspace(-0.4, 0.5, na)
        ms, _ = timed_ms(lambda: K.perturb_predict(x, u, da), reps=3, warm=2)
        nbytes = 8.0 * n * (max(c, 1) + 1) * ((na + 3) // 4) + 8.0 * n
        out["cases"].append({"nsamp": n, "nchan": c, "nalpha": na, "ms": ms, "GBps": nbytes / ms / 1e6,
                             "frac_hbm": nbytes / ms / 1e6 / peaks["hbm_gbs"],
                             "bound": "fp64 exp (scalar x)" if not c else "hbm"})
        del x, u
    return out


def config_c5(dev) -> dict:
    """Volume extrapolation, order 1, 1e9 samples STREAMED from pinned host memory in 2^26-sample chunks through a
    double-buffered H2D ring (volume.factory_extrapmodel's data: <x>, <W>, <xW>, <dxdq>)."""
    from thermoextrap_b200 import streaming

    return streaming.bench_volume_stream(total=1_000_000_000, chunk=1 << 26, dev=dev, peaks=measured_peaks())


def main() -> None:
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()

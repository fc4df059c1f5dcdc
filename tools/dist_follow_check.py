"""N-GPU mode with the opt-in upload pipeline (THERMOEXTRAP_B200_ASYNC_UPLOAD_DIST): parity against the plain N-GPU
path and end-to-end timing of both, through the public API from pinned host shards.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
        tools/dist_follow_check.py [nsamp] [nrep] [steps]

Same job as bench.py's e2e leg (order 6, nsamp x 64 channels, nrep replicates): every rank pins its 1/N row shard,
``parallel.enable()``, then  from_vals -> factory_extrapmodel -> resample -> predict.  Prints one JSON line (rank 0).
"""

from __future__ import annotations

import json
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main() -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    import thermoextrap_b200 as xt
    from thermoextrap_b200 import backend
    from thermoextrap_b200 import kernels as K
    from thermoextrap_b200 import parallel as par

    nsamp = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
    nrep = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    nchan, order = 64, 6
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    par.enable()
    lo, hi = par.shard_range(nsamp, rank, world)
    xh = torch.empty((hi - lo, nchan), dtype=torch.float64).pin_memory()
    uh = torch.empty(hi - lo, dtype=torch.float64).pin_memory()
    step = 1 << 22
    g = torch.Generator(device=dev).manual_seed(77 + rank)   # every rank its own shard of one (random) sample set
    for i in range(0, hi - lo, step):
        j = min(hi - lo, i + step)
        u = torch.empty(j - i, dtype=torch.float64, device=dev).normal_(3.0, 1.5, generator=g)
        x = torch.empty((j - i, nchan), dtype=torch.float64, device=dev).normal_(0.0, 1.0, generator=g)
        x.add_(0.3 * u[:, None])
        xh[i:j].copy_(x)
        uh[i:j].copy_(u)
    del x, u
    torch.cuda.empty_cache()
    xv = xt.DataArray(xh.numpy(), dims=("rec", "val"))
    uv = xt.DataArray(uh.numpy(), dims=("rec",))
    betas = [0.9, 1.1]

    def job():
        backend.clear_device_cache()
        data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
        model = xt.beta.factory_extrapmodel(beta=1.0, data=data)
        boot = model.resample(sampler={"nrep": nrep, "rng": np.random.default_rng(1)})
        return model.predict(betas).values, boot.predict(betas).values

    def timed(flag: bool):
        backend.ASYNC_UPLOAD_DIST = flag
        out = job()
        job()
        ts = []
        for _ in range(steps):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            out = job()
            torch.cuda.synchronize()
            dist.barrier()
            ts.append(time.perf_counter() - t0)
        return out, ts

    before = K.follow_windows
    (p0, b0), t_plain = timed(False)
    assert K.follow_windows == before
    (p1, b1), t_follow = timed(True)
    followed = K.follow_windows - before
    rel = lambda a, b: float(np.max(np.abs(a - b) / np.abs(b)))   # noqa: E731
    res = {"n_gpus": world, "nsamp": nsamp, "nchan": nchan, "order": order, "nrep": nrep,
           "e2e_ms_plain": [round(1e3 * t, 1) for t in t_plain], "e2e_ms_following": [round(1e3 * t, 1) for t in t_follow],
           "replicates_per_s_plain": nrep / min(t_plain), "replicates_per_s_following": nrep / min(t_follow),
           "windows_followed_on_rank0": followed,
           "max_rel_diff_predict": rel(p1, p0), "max_rel_diff_bootstrap_predict": rel(b1, b0)}
    par.disable()
    if rank == 0:
        print(json.dumps(res), flush=True)
    assert followed > 0 and res["max_rel_diff_bootstrap_predict"] < 1e-10 and res["max_rel_diff_predict"] < 1e-10
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

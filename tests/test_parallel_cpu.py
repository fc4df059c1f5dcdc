"""World-size-2 runs of the multi-GPU host logic on CPU (gloo), kernels replaced by the oracle stand-in.

What is under test: sharding arithmetic, the common-pivot broadcast, the all_reduce of additive
sums, replicate windows and the all_gather of slabs (thermoextrap_b200/parallel.py) -- the same
code that runs over NCCL on the GPUs.
"""

import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import fake_kernels

    import thermoextrap_b200.parallel as par

    par.K = fake_kernels
    from oracle import comoments as oc

    rng = np.random.default_rng(0)
    n, nv, order, nrep = 1001, 3, 5, 7
    u = rng.normal(10.0, 2.0, n)
    x = 0.3 * u[:, None] + rng.normal(0.0, 1.0, (n, nv))
    lo, hi = par.shard_range(n, rank, world)
    # samples sharded
    red = par.reduce_vals_sharded(x[lo:hi], u[lo:hi], order).numpy()
    ref = oc.comoments_two_pass(x, u, order)
    assert np.all(np.abs(red - ref) <= 1e-9 * np.maximum(np.abs(ref), oc.tolerance_scale(x, u, order)))
    # replicates sharded: one numpy table split over the ranks
    g = np.random.default_rng(123)
    boot = par.bootstrap_sharded(g, x[:, 0], u, nrep, order).numpy()
    idx = np.random.default_rng(123).choice(n, (nrep, n))
    bref = oc.resample_vals_gather(x[:, 0], u, order, idx)
    assert boot.shape == bref.shape
    assert np.all(np.abs(boot - bref) <= 1e-9 * np.maximum(np.abs(bref), oc.tolerance_scale(x[:, 0], u, order)))
    slab = par.bootstrap_sharded(np.random.default_rng(123), x[:, 0], u, nrep, order, gather=False).numpy()
    a, b = par.shard_range(nrep, rank, world)
    assert np.allclose(slab, bref[a:b], rtol=1e-9, atol=1e-12)
    # wide x: every rank bootstraps its replicate slice of the same table through the windowed path
    wide = par.bootstrap_sharded(np.random.default_rng(123), x, u, nrep, order).numpy()
    wref = oc.resample_vals_gather(x, u, order, idx)
    assert wide.shape == wref.shape == (nrep, nv, 2, order + 1)
    assert np.all(np.abs(wide - wref) <= 1e-9 * np.maximum(np.abs(wref), oc.tolerance_scale(x, u, order)))
    # samples sharded bootstrap: every rank keeps the frequency columns of its own rows
    ss = par.bootstrap_sample_sharded(np.random.default_rng(123), x[lo:hi], u[lo:hi], nrep, order).numpy()
    assert ss.shape == wref.shape
    assert np.all(np.abs(ss - wref) <= 1e-9 * np.maximum(np.abs(wref), oc.tolerance_scale(x, u, order)))
    # x_is_u sharded reduce
    ru = par.reduce_vals_sharded(None, u[lo:hi], order, x_is_u=True).numpy()
    uref = oc.comoments_two_pass(u, u, order)
    assert np.all(np.abs(ru.reshape(uref.shape) - uref) <= 1e-9 * np.maximum(np.abs(uref), oc.tolerance_scale(u, u, order)))
    # ragged / equal row gathers
    import torch

    full = par.all_gather_rows(torch.from_numpy(x[lo:hi].copy()))
    assert np.array_equal(full.numpy(), x)
    # ---- the public API in N-GPU mode: every rank hands in its row shard, every rank gets the full answer
    import thermoextrap_b200 as xt
    import thermoextrap_b200.backend as backend

    backend.K = fake_kernels
    backend.clear_device_cache()
    par.enable()
    try:
        xv = xt.DataArray(x[lo:hi].copy(), dims=("rec", "val"))
        uv = xt.DataArray(u[lo:hi].copy(), dims=("rec",))
        data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
        got = data.dxduave.obj.values
        assert np.all(np.abs(got - ref) <= 1e-9 * np.maximum(np.abs(ref), oc.tolerance_scale(x, u, order)))
        model = xt.beta.factory_extrapmodel(beta=0.5, data=data)
        boot = model.resample(sampler={"nrep": nrep, "rng": np.random.default_rng(123)})
        bv = boot.data.dxduave.obj.values
        assert bv.shape == wref.shape
        assert np.all(np.abs(bv - wref) <= 1e-9 * np.maximum(np.abs(wref), oc.tolerance_scale(x, u, order)))
        pred = boot.predict([0.4, 0.6]).values
        assert pred.shape == (2, nrep, nv) and np.isfinite(pred).all()
        # an explicit index table over the GLOBAL row numbering
        b2 = model.resample(sampler={"indices": idx}).data.dxduave.obj.values
        assert np.all(np.abs(b2 - wref) <= 1e-9 * np.maximum(np.abs(wref), oc.tolerance_scale(x, u, order)))
        # opt-in N-GPU upload pipeline, host logic: K1 (collectives included) deferred at from_vals, launched by the
        # flush at the end of the bootstrap -- on every rank, in the same order
        backend.ASYNC_UPLOAD_DIST, backend.DEFER_REDUCE = True, True
        calls = []
        real = fake_kernels.reduce_sums

        def counted(*a, **k):
            calls.append(1)
            return real(*a, **k)

        fake_kernels.reduce_sums = counted
        try:
            data2 = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
            assert not calls and data2.dxduave.mom == (1, order)
            boot2 = xt.beta.factory_extrapmodel(beta=0.5, data=data2).resample(
                sampler={"nrep": nrep, "rng": np.random.default_rng(123)})
            assert calls and data2.dxduave._obj is not None and not backend._deferred   # flushed by the bootstrap
            assert np.array_equal(boot2.data.dxduave.obj.values, bv)
            assert np.array_equal(data2.dxduave.obj.values, got)
        finally:
            fake_kernels.reduce_sums = real
            backend.ASYNC_UPLOAD_DIST, backend.DEFER_REDUCE = False, False
    finally:
        par.disable()
    (Path(out_dir) / f"ok{rank}").write_text("ok")
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_paths_gloo(tmp_path, world):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert sorted(p.name for p in tmp_path.iterdir()) == [f"ok{r}" for r in range(world)]


def test_shard_range_is_a_partition():
    from thermoextrap_b200.parallel import shard_range

    for n in (0, 1, 7, 1000):
        for world in (1, 2, 3, 8):
            cuts = [shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            assert max(b - a for a, b in cuts) - min(b - a for a, b in cuts) <= 1

"""Multi-GPU parity check over NCCL: torchrun --nproc-per-node N tools/dist_check.py

Every rank checks the sharded paths (thermoextrap_b200/parallel.py) against the single-GPU kernels.
"""

import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import numpy as np
import torch
import torch.distributed as dist

from thermoextrap_b200 import kernels as K
from thermoextrap_b200 import parallel as par

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()

rng = np.random.default_rng(0)
n, nv, order, nrep = 2_000_003, 8, 6, 13
u = rng.normal(175.0, 5.3, n)
x = 1e-3 * u[:, None] + rng.normal(0.0, 0.02, (n, nv))
lo, hi = par.shard_range(n, rank, world)
red = par.reduce_vals_sharded(x[lo:hi], u[lo:hi], order)
ref = K.reduce_vals(x, u, order)
err = float(((red - ref).abs() / ref.abs().clamp_min(1e-300)).max())
assert torch.allclose(red, ref, rtol=1e-10, atol=1e-16), err

g = np.random.default_rng(77)
boot = par.bootstrap_sharded(g, x[:, 0].copy(), u, nrep, order)
g1 = np.random.default_rng(77)
single = K.sampler_resample(g1, x[:, 0].copy(), u, nrep, order)
assert torch.allclose(boot, single, rtol=1e-11, atol=0)
assert g.bit_generator.state["state"] == g1.bit_generator.state["state"]
idx = K.sampler_indices(np.random.default_rng(77), nrep, n)
assert np.array_equal(idx[:2, :1000].cpu().numpy(), np.random.default_rng(77).choice(n, (nrep, n))[:2, :1000])
# wide x: replicate slices of ONE table through the windowed histogram + tensor-core contraction
nrep_w = 300
gw, gw1 = np.random.default_rng(78), np.random.default_rng(78)
wide = par.bootstrap_sharded(gw, x, u, nrep_w, order)
wide1 = K.sampler_resample_wide(gw1, x, u, nrep_w, order)
scale = torch.as_tensor(np.abs(ref.cpu().numpy()), device=wide.device).clamp_min(1e-30)
werr = float(((wide - wide1).abs() / torch.maximum(wide1.abs(), scale)).max())
assert wide.shape == wide1.shape == (nrep_w, nv, 2, order + 1) and werr < 1e-11, werr
assert gw.bit_generator.state["state"] == gw1.bit_generator.state["state"]
# samples sharded bootstrap: every rank keeps the frequency columns of its own rows, pivot sums all_reduced
ss = par.bootstrap_sample_sharded(np.random.default_rng(78), x[lo:hi], u[lo:hi], nrep_w, order)
serr = float(((ss - wide1).abs() / torch.maximum(wide1.abs(), scale)).max())
assert ss.shape == wide1.shape and serr < 1e-11, serr
# the public API in N-GPU mode: row shards in, full result out on every rank
import thermoextrap_b200 as xt

par.enable()
xv = xt.DataArray(x[lo:hi].copy(), dims=("rec", "val"))
uv = xt.DataArray(u[lo:hi].copy(), dims=("rec",))
data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
aerr = float(np.max(np.abs(data.dxduave.obj.values - ref.cpu().numpy()) / np.maximum(np.abs(ref.cpu().numpy()), 1e-30)))
assert aerr < 1e-9, aerr
bapi = xt.beta.factory_extrapmodel(beta=1.0, data=data).resample(sampler={"nrep": nrep_w, "rng": np.random.default_rng(78)})
bv = torch.as_tensor(bapi.data.dxduave.obj.values, device=wide.device)
berr = float(((bv - wide1).abs() / torch.maximum(wide1.abs(), scale)).max())
assert bv.shape == wide1.shape and berr < 1e-11, berr
par.disable()
dist.barrier()
if rank == 0:
    print(f"sample-sharded bootstrap max err {serr:.1e}; N-GPU API mode: reduce {aerr:.1e}, bootstrap {berr:.1e}")
    print(f"dist_check ok: world={world} sharded reduce max rel err {err:.2e}; sharded bootstrap == single-GPU "
          f"(scalar and wide, wide max err {werr:.1e}); rng state == numpy")
dist.destroy_process_group()

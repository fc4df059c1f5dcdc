# python tools/e2e_probe.py nsamp nrep [table_bytes] -- e2e probe: the API chain from pinned host arrays with the chunked upload on / off
import sys, time, os
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import thermoextrap_b200 as xt
from thermoextrap_b200 import backend, kernels as K
nsamp, nchan, nrep, order = int(sys.argv[1]), 64, int(sys.argv[2]), 6
g = torch.Generator(device="cuda"); g.manual_seed(0)
xd = torch.randn((nsamp, nchan), dtype=torch.float64, device="cuda", generator=g)
ud = torch.randn(nsamp, dtype=torch.float64, device="cuda", generator=g)
xd += 0.3 * ud[:, None]
xh = torch.empty((nsamp, nchan), dtype=torch.float64).pin_memory(); xh.copy_(xd)
uh = torch.empty(nsamp, dtype=torch.float64).pin_memory(); uh.copy_(ud)
del xd, ud; torch.cuda.empty_cache()
xv = xt.DataArray(xh.numpy(), dims=("rec", "val")); uv = xt.DataArray(uh.numpy(), dims=("rec",))
if len(sys.argv) > 3:
    backend.WIDE_TABLE_BYTES = int(float(sys.argv[3]))
BETAS = [0.9, 1.1]
def step():
    backend.clear_device_cache()
    data = xt.DataCentralMomentsVals.from_vals(xv=xv, uv=uv, order=order, central=True)
    model = xt.beta.factory_extrapmodel(beta=1.0, data=data)
    b = model.resample(sampler={"nrep": nrep, "rng": np.random.default_rng(1)})
    return model.predict(BETAS).values, b.predict(BETAS).values
res = {}
for mode in (True, False, True):
    backend.ASYNC_UPLOAD = mode
    step(); 
    torch.cuda.synchronize(); ts = []
    for _ in range(3):
        t0 = time.perf_counter(); out = step(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    res[mode] = out
    print(f"async={mode} ms/step: " + " ".join(f"{1e3*t:.1f}" for t in ts), "follow_windows", K.follow_windows, flush=True)
backend.ASYNC_UPLOAD = True
K.upload_trace = {}
t0 = time.perf_counter(); step(); torch.cuda.synchronize(); t1 = time.perf_counter()
tr = K.upload_trace; K.upload_trace = None
z = tr["t0"]
print(f"traced step {1e3*(t1-t0):.1f} ms; {tr['windows']} windows follow; relative to the pivot rows landing: contraction starts at "
      f"{z.elapsed_time(tr['start']):.0f} ms")
print("chunk landed  :", " ".join(f"{z.elapsed_time(e):.0f}" for e in tr["landed"]))
print("chunk done    :", " ".join(f"{z.elapsed_time(e):.0f}" for e in tr["done"]))
a, b = res[True], res[False]
print("pred max rel diff", float(np.max(np.abs(a[0]-b[0])/np.abs(b[0]))), "boot max rel diff", float(np.max(np.abs(a[1]-b[1])/np.abs(b[1]))))

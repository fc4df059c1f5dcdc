"""Short launch sequence for `ncu --set full` captures of the hot kernels (development tool).

usage: python tools/ncu_case.py [gather|reduce|resample]
  gather  : count pass (k_sampler #0) then the fused sampler+bootstrap (k_sampler #1), C2 width, 64 replicates
"""

import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import numpy as np
import torch

from thermoextrap_b200 import kernels as K

which = sys.argv[1] if len(sys.argv) > 1 else "gather"
n = 10_000_000
u = torch.randn(n, dtype=torch.float64, device="cuda") * 5.3 + 175
x = 1e-3 * u + 0.02 * torch.randn(n, dtype=torch.float64, device="cuda")
if which == "gather":
    nrep = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    K.sampler_resample(np.random.default_rng(1), x, u, nrep, 8, sync_rng=False)
elif which == "reduce":
    K.reduce_vals(x, u, 8)
elif which == "reduce_scalar":  # python tools/ncu_case.py reduce_scalar [nsamp order]: scalar K1 at HBM-resident size
    nw, kw = (int(v) for v in (sys.argv[2:4] if len(sys.argv) > 3 else (100_000_000, 8)))
    uw = torch.randn(nw, dtype=torch.float64, device="cuda") * 5.3 + 175
    xw = 1e-3 * uw + 0.02 * torch.randn(nw, dtype=torch.float64, device="cuda")
    for _ in range(3):
        K.reduce_vals(xw, uw, kw)
elif which == "wide":  # python tools/ncu_case.py wide nsamp nchan order   (TXB_WIDE_TMA picks the kernel)
    nw, cw, kw = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    xw = torch.randn((nw, cw), dtype=torch.float64, device="cuda")
    K.reduce_vals(xw, u[:nw], kw)
elif which == "resample":
    f8 = K.sampler_freq(np.random.default_rng(1), 64, n, dtype=torch.uint8)
    K.resample_vals(x, u, f8, 8)
elif which == "resample_wide":  # python tools/ncu_case.py resample_wide [nsamp nchan order nrep]
    nw, cw, kw, rw = (int(v) for v in (sys.argv[2:6] if len(sys.argv) > 5 else (1_000_000, 64, 6, 128)))
    xw = torch.randn((nw, cw), dtype=torch.float64, device="cuda") + 1e-3 * u[:nw, None]
    f8 = K.sampler_freq(np.random.default_rng(1), rw, nw, dtype=torch.uint8)
    K.resample_vals(xw, u[:nw], f8, kw)
elif which == "resample_i64":  # scalar x from the reference's own int64 frequency table (k_resample_mma_scalar)
    nr = 256
    f64 = K.sampler_freq(np.random.default_rng(1), nr, 2_000_000, dtype=torch.uint8).to(torch.int64)
    K.resample_vals(x[:2_000_000], u[:2_000_000], f64, 8)
elif which == "target":  # python tools/ncu_case.py target [nsamp nrep]: K6 table window + K2 contraction, both forms
    from thermoextrap_b200 import _lib

    lib = _lib.load()
    nw, rw = (int(v) for v in (sys.argv[2:4] if len(sys.argv) > 3 else (100_000_000, 128)))
    del x
    uw = torch.randn(nw, dtype=torch.float64, device="cuda") * 5.3 + 175
    xw = torch.empty((nw, 64), dtype=torch.float64, device="cuda")
    for i in range(0, nw, 1 << 22):
        xw[i : i + (1 << 22)].normal_(0.0, 0.02)
        xw[i : i + (1 << 22)].addcmul_(uw[i : i + (1 << 22), None], torch.linspace(0.5e-3, 2e-3, 64, dtype=torch.float64, device="cuda")[None])
    plan = K.SamplerPlan(np.random.default_rng(1), rw, nw)            # k_sampler_count (+ scans)
    if len(sys.argv) > 4 and sys.argv[4] == "forms":                   # every form, for A/B captures
        for nth in (512, 1024):
            lib.txb_set_scatter_threads(nth)
            f8 = K.sampler_freq(plan, dtype=torch.uint8)
        for ver in (1, 2, 3):
            lib.txb_set_resample_kernel(ver)
            K.resample_vals(xw, uw, f8, 6)
    else:                                                              # the defaults bench.py runs
        f8 = K.sampler_freq(plan, dtype=torch.uint8)                   # k_bucket_scatter<1, 512>, k_bucket_hist
        K.resample_vals(xw, uw, f8, 6)                                 # k_resample_mma_scalar, k_resample_mma2<6, u8, 8>
elif which == "perturb":
    xw = torch.randn((4_000_000, 64), dtype=torch.float64, device="cuda")
    K.perturb_predict(xw, u[:4_000_000] * 0.01, np.linspace(-0.4, 0.5, 8))
torch.cuda.synchronize()

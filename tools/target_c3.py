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

from bench import wide_target


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--nsamp", type=float, default=1e8)
    ap.add_argument("--nchan", type=int, default=64)
    ap.add_argument("--order", type=int, default=6)
    ap.add_argument("--nrep", type=int, default=1000)
    ap.add_argument("--table-gb", type=float, default=0.0, help="window table budget (0: a third of free memory)")
    a = ap.parse_args()
    print(json.dumps(wide_target(int(a.nsamp), a.nchan, a.order, a.nrep, a.table_gb)), flush=True)


if __name__ == "__main__":
    main()

"""Per-kernel histogram of the SASS opcodes that identify the Blackwell paths (run here, no GPU needed):
python tools/sass_opcodes.py > profiles/sass_opcodes_r02.txt

UBLKCP / UTMALDG = TMA bulk / tensor-map loads, DMMA = fp64 tensor-core MMA (tcgen05 has no f64 kind: mma.sync DMMA
is the fp64 tensor path on sm_100a), LDGSTS = cp.async, SYNCS = mbarrier ops, ATOMS / RED / ATOMG = shared / global
atomics, DFMA / DMUL / DADD = the scalar fp64 pipe.
"""

import re
import subprocess
import sys
from collections import Counter, defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "thermoextrap_b200" / "lib" / "libtxb.so"
OPS = ["UBLKCP", "UTMALDG", "DMMA", "LDGSTS", "SYNCS", "ATOMS", "RED", "ATOMG", "DFMA", "DMUL", "DADD", "IMAD", "LDS", "STS",
       "LDG", "STG", "BAR"]

sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
per: dict = defaultdict(Counter)
name = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name).replace("void ", "").replace("txb::", "")
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and name:
        per[name][m.group(1)] += 1
print(f"# SASS opcode counts of {LIB.relative_to(ROOT)} (cuobjdump -sass; static instruction counts, summed over the")
print("# template instantiations of each kernel; the bench's own instantiations are listed separately at the end)")
print(f"# {'kernel (instantiations)':48s} " + " ".join(f"{o:>7s}" for o in OPS))
tot = Counter()
groups: dict = defaultdict(Counter)
ninst: Counter = Counter()
for k, c in per.items():
    base = k.split("<")[0]
    groups[base].update({o: c[o] for o in OPS})
    ninst[base] += 1
    tot.update({o: c[o] for o in OPS})
for base in sorted(groups):
    c = groups[base]
    print(f"{(base + ' (' + str(ninst[base]) + ')')[:50]:50s} " + " ".join(f"{c[o]:7d}" for o in OPS))
print("# ---- the instantiations bench.py's headline and C2 legs launch")
for k in sorted(per):
    if any(t in k for t in ("k_resample_mma2<6, unsigned char, 8>", "k_resample_mma<6, unsigned char>", "k_bucket_scatter<",
                            "k_bucket_hist<unsigned char>", "k_bucket_consume<8, true, false>", "k_sampler_count",
                            "k_resample_mma_scalar<6, false, unsigned char>", "k_beta_fused<0, 6>",
                            "k_reduce_wide_tma<6, 4, true, false, 512>", "k_perturb")):
        print(f"{k[:50]:50s} " + " ".join(f"{per[k][o]:7d}" for o in OPS))
print(f"{'TOTAL (' + str(len(per)) + ' kernels in the library)':50s} " + " ".join(f"{tot[o]:7d}" for o in OPS))
print("# absent by design: HMMA / UTC*MMA (no fp64 kind on tcgen05), HGMMA (Hopper only)")
for bad in ("HMMA", "UTCHMMA", "HGMMA"):
    print(f"# {bad}: {sum(c[bad] for c in per.values())}")

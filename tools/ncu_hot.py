"""Hottest SASS instructions of every kernel in an .ncu-rep (source page, needs -lineinfo + --import-source on):
python tools/ncu_hot.py rep.ncu-rep [top_n] > profiles/x_hot.txt"""

import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
blocks, cur = [], []
for line in raw.splitlines():
    if line.startswith('"Kernel Name"') or line.startswith("Kernel Name"):
        if cur:
            blocks.append(cur)
        cur = [line]
    elif cur is not None:
        cur.append(line)
if cur:
    blocks.append(cur)
if not blocks:
    print(raw[:4000])
    sys.exit(0)
for blk in blocks:
    rows = list(csv.reader(io.StringIO("\n".join(blk))))
    # layout: a few "key,value" header rows, then the table header containing "Source" / "Address"
    hi = next((i for i, r in enumerate(rows) if "Source" in r and any("Sampl" in c for c in r)), None)
    for r in rows[: hi if hi is not None else 3]:
        print("# " + ", ".join(r))
    if hi is None:
        continue
    hdr = rows[hi]
    samp = next(i for i, c in enumerate(hdr) if "Sampling" in c and "All" in c) if any("Sampling" in c and "All" in c for c in hdr) else next(i for i, c in enumerate(hdr) if "Sampl" in c)
    src = hdr.index("Source")
    stall_cols = [i for i, c in enumerate(hdr) if c.startswith("stall_")]
    body = [r for r in rows[hi + 1 :] if len(r) == len(hdr)]

    def num(v):
        try:
            return float(v.replace(",", ""))
        except ValueError:
            return 0.0

    tot = sum(num(r[samp]) for r in body) or 1.0
    agg = {hdr[i]: sum(num(r[i]) for r in body) for i in stall_cols}
    print(f"# total samples {tot:.0f}; stall reasons: " +
          ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    for r in sorted(body, key=lambda r: -num(r[samp]))[:top]:
        why = sorted(((hdr[i][6:], num(r[i])) for i in stall_cols), key=lambda kv: -kv[1])[:3]
        print(f"{100 * num(r[samp]) / tot:6.2f}%  {r[src][:90]:90s}  " + " ".join(f"{k}:{v:.0f}" for k, v in why if v))
    print()

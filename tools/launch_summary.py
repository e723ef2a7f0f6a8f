"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the LAST `--last N` launches.

    python tools/launch_summary.py gpurun_out/launches.csv --last 75
"""
import argparse
import collections
import csv

ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--last", type=int, default=0)
args = ap.parse_args()
with open(args.csv) as f:
    lines = [l for l in f if l.startswith('"')]
r = csv.reader(lines)
hdr = next(r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
seq = []
for row in r:
    v = float(row[vi].replace(",", ""))
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0}.get(row[ui], 1e-6)
    seq.append((row[ki], v * scale))
if args.last:
    seq = seq[-args.last:]
tot = collections.OrderedDict()
for n, v in seq:
    n = n.split("(")[0].replace("void ", "").replace("g16::", "")
    n = n.replace("Curve<Ext2<FieldCall<FpBN254>, 1>>", "G2").replace("Curve<FieldCall<FpBN254>>", "G1")
    c, t = tot.get(n, (0, 0.0))
    tot[n] = (c + 1, t + v)
total = sum(t for _, t in tot.values())
for n, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{t:9.3f} ms {100 * t / total:5.1f} %  x{c:<3d} {n[:100]}")
print(f"{total:9.3f} ms total over {len(seq)} launches")

"""Kernel shares from an ncu launch list (--metrics gpu__time_duration.sum --csv): python tools/launch_shares.py file.csv"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows:
    if r is hdr or r[ki] == "Kernel Name":
        continue
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    name = r[ki].split("(")[0][:70]
    tot[name] += v
    cnt[name] += 1
s = sum(tot.values())
for k, v in tot.most_common():
    print(f"{v:10.3f} ms {100 * v / s:5.1f}%  x{cnt[k]:4d}  {k}")

"""ncu launch list (--metrics gpu__time_duration.sum --csv) -> per-kernel launch count, total and mean duration, share.
python scripts/ncu_by_kernel.py launches.csv [skip_first_n_launches]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if r]
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[start]
col = {h: i for i, h in enumerate(hdr)}
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
body = [r for r in rows[start + 1:] if len(r) == len(hdr) and r[col["Metric Name"]] == "gpu__time_duration.sum"][skip:]
agg = collections.OrderedDict()
for r in body:
    name = r[col["Kernel Name"]]
    for junk in ("void ", "<unnamed>::"):
        name = name.replace(junk, "")
    name = name.split("(")[0]
    unit, val = r[col["Metric Unit"]], float(r[col["Metric Value"]].replace(",", ""))
    us = val * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"{len(body)} launches, {tot / 1e3:.3f} ms of kernel time")
print(f"{'kernel':70s} {'n':>6s} {'total_us':>12s} {'mean_us':>10s} {'share':>7s}")
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{name[:70]:70s} {n:6d} {us:12.1f} {us / n:10.2f} {100 * us / tot:6.2f}%")

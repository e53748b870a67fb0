"""Turn the output of scripts/profile_round.sh (gpurun_out/prof/) into the tracked artefacts under profiles/:
  r1_launches_<cfg>.csv.gz + r1_launches_<cfg>_summary.csv   per-kernel launch counts, total time, share
  r1_ncu_<kernel>_details.txt                                `ncu --page details` of the one captured launch
  r1_ncu_<kernel>_phases.txt                                 raw DRAM / time metrics + stall samples per phase
  r1_traffic.json                                            dram bytes per launch (bench.py -> roofline.traffic)
python scripts/summarise_profiles.py [gpurun_out/prof] [profiles] [round tag, default r1]"""
import collections
import csv
import gzip
import json
import os
import re
import shutil
import subprocess
import sys

src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/prof"
dst = sys.argv[2] if len(sys.argv) > 2 else "profiles"
tag = sys.argv[3] if len(sys.argv) > 3 else "r1"
here = os.path.dirname(os.path.abspath(__file__))


def short(name):
    name = re.sub(r"^void ", "", name)
    name = name.replace("<unnamed>::", "")
    depth, out = 0, ""
    for ch in name:  # cut the argument list: the first '(' outside template brackets
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            break
        out += ch
    return out


for cfg in ("cfg2", "cfg4"):
    path = os.path.join(src, f"launches_{cfg}.csv")
    if not os.path.exists(path):
        continue
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    rows = list(csv.DictReader(lines[start:]))
    agg = collections.OrderedDict()
    for r in rows:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
        a = agg.setdefault(short(r["Kernel Name"]), [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(dst, f"{tag}_launches_{cfg}_summary.csv"), "w") as f:
        f.write(f"# every kernel launch of: ncu --metrics gpu__time_duration.sum --clock-control none python bench.py "
                f"--workload {cfg} --steps 2 --warmup 3 (bf16x3); all phases of the run (timed, e2e, profiling pass); "
                f"cold-cache serialised times: compare SHARES, not absolutes\n")
        f.write("kernel,launches,total_us,share,avg_us\n")
        for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f'"{k}",{n},{us:.1f},{us / tot:.4f},{us / n:.1f}\n')
    with open(path, "rb") as fi, open(os.path.join(dst, f"{tag}_launches_{cfg}.csv.gz"), "wb") as raw:
        with gzip.GzipFile(fileobj=raw, mode="wb", mtime=0) as fo:  # mtime 0: the same input gives the same bytes
            shutil.copyfileobj(fi, fo)

traffic_path = os.path.join(dst, f"{tag}_traffic.json")
traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
key_of = {"tc_edge_bwd128": "tc_edge_bwd", "tc_edge_fwd128": "tc_edge_fwd"}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
for name in ("tc_edge_bwd128", "tc_edge_fwd128", "tc_edge_fwd256", "tc_edge_bwd256", "tc_wgrad256"):
    det, raw, srcv = (os.path.join(src, f"{name}_{s}") for s in ("details.txt", "raw.txt", "source.csv"))
    if not (os.path.exists(det) and os.path.exists(raw)):
        continue
    shutil.copy(det, os.path.join(dst, f"{tag}_ncu_{name}_details.txt"))
    total = 0.0
    with open(os.path.join(dst, f"{tag}_ncu_{name}_phases.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on, one launch of {name} inside bench.py (see "
                f"scripts/profile_round.sh); stall samples summed between synchronisation points (scripts/ncu_phases.py)\n")
        for line in open(raw):
            f.write("# " + line)
            p = line.split()
            if len(p) == 3 and p[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                total += float(p[2].replace(",", "")) * UNIT[p[1]]
        if os.path.exists(srcv):
            f.write(subprocess.run([sys.executable, os.path.join(here, "ncu_phases.py"), srcv], capture_output=True,
                                   text=True).stdout)
    traffic[key_of.get(name, name)] = {"dram_bytes_per_launch": int(total),
                                       "capture": f"profiles/{tag}_ncu_{name}_details.txt"}
json.dump(traffic, open(traffic_path, "w"), indent=1)
print(json.dumps({k: v for k, v in traffic.items() if k != "_comment"}, indent=1))

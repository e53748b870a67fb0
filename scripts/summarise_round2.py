"""gpurun_out/prof2 (scripts/profile_round2.sh) -> tracked artefacts under profiles/:
  r2_ncu_all_kernels.csv      one row per kernel of the path from ONE `ncu --set full` pass (scripts/profile_step.py, cfg2 size):
                              duration, DRAM bytes read + written, achieved DRAM GB/s and its fraction of the measured HBM
                              peak, L2 bytes, tensor-pipe and SM / memory throughput percentages
  r2_launches_cfg2_summary.csv  per-kernel launch count, total time and share of one eager bench run (+ the .csv.gz)
  r2_ncu_tc_edge_{fwd,bwd}128_{details,phases}.txt, r2_traffic.json
python scripts/summarise_round2.py [gpurun_out/prof2] [profiles]"""
import collections
import csv
import gzip
import json
import os
import shutil
import subprocess
import sys

src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/prof2"
dst = sys.argv[2] if len(sys.argv) > 2 else "profiles"
here = os.path.dirname(os.path.abspath(__file__))
root = os.path.dirname(here)
peaks = json.load(open(os.path.join(root, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(root, "MEASURED_PEAKS.json")) else {}
HBM = float(peaks.get("hbm_gbs", 6553.3))


def short(name):
    name = name.replace("void ", "").replace("<unnamed>::", "")
    depth, out = 0, ""
    for ch in name:
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            break
        out += ch
    return out.strip()


def fnum(v):
    try:
        return float(str(v).replace(",", ""))
    except ValueError:
        return float("nan")


# ---- (1) every kernel, one full-set capture ---------------------------------------------------------------------------
raw = os.path.join(src, "all_kernels_raw.csv")
traffic = {}
if os.path.exists(raw):
    rows = [r for r in csv.reader(open(raw)) if r]
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units, body = rows[start], rows[start + 1], rows[start + 2:]
    col = {h: i for i, h in enumerate(hdr)}
    want = [("duration_us", "gpu__time_duration.sum", 1e-3), ("dram_read_bytes", "dram__bytes_read.sum", 1.0),
            ("dram_write_bytes", "dram__bytes_write.sum", 1.0), ("l2_bytes", "lts__t_bytes.sum", 1.0),
            ("tensor_pipe_pct_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
            ("sm_throughput_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
            ("dram_throughput_pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
            ("l2_throughput_pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
            ("achieved_occupancy_pct", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
            ("registers", "launch__registers_per_thread", 1.0), ("grid", "launch__grid_size", 1.0), ("block", "launch__block_size", 1.0)]
    seen = collections.OrderedDict()
    for r in body:
        if len(r) <= col["Kernel Name"]:
            continue
        name = short(r[col["Kernel Name"]])
        rec = {k: (fnum(r[col[m]]) * sc if m in col else float("nan")) for k, m, sc in want}
        seen.setdefault(name, []).append(rec)
    out = os.path.join(dst, "r2_ncu_all_kernels.csv")
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["# one `ncu --set full --clock-control none` pass over scripts/profile_step.py (cfg2 size: N=12288, E=417420, d=128, "
                    "bf16x3); cold caches, kernels serialised.  achieved_dram_GBps = (read + written DRAM bytes) / duration; "
                    f"frac_hbm = achieved / {HBM} GB/s (MEASURED_PEAKS.json).  The LARGEST launch of each kernel is listed."])
        w.writerow(["kernel", "launches_captured"] + [k for k, _, _ in want] + ["achieved_dram_GBps", "frac_hbm"])
        for name, recs in seen.items():
            rec = max(recs, key=lambda q: q["duration_us"] if q["duration_us"] == q["duration_us"] else -1)
            byts = rec["dram_read_bytes"] + rec["dram_write_bytes"]
            gbs = byts / (rec["duration_us"] * 1e-6) / 1e9 if rec["duration_us"] > 0 else float("nan")
            w.writerow([name, len(recs)] + [f"{rec[k]:.6g}" for k, _, _ in want] + [f"{gbs:.1f}", f"{gbs / HBM:.4f}"])
            traffic[name] = dict(dram_bytes_per_launch=int(byts) if byts == byts else None, capture="profiles/r2_ncu_all_kernels.csv")
    print("wrote", out, len(seen), "kernels")

# ---- (2) launch list ----------------------------------------------------------------------------------------------------
path = os.path.join(src, "launches_cfg2.csv")
if os.path.exists(path):
    rows = [r for r in csv.reader(open(path)) if r]
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[start]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}
    agg = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) <= mv:
            continue
        a = agg.setdefault(short(r[kn]), [0, 0.0])
        a[0] += 1
        a[1] += fnum(r[mv]) * scale.get(r[mu], 1.0)
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(dst, "r2_launches_cfg2_summary.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["# ncu --metrics gpu__time_duration.sum --clock-control none, `bench.py --eager --steps 2 --warmup 3` (5 steps + "
                    "profile pass + set-up); per-launch times are cold-cache and serialised: compare SHARES, not absolutes"])
        w.writerow(["kernel", "launches", "total_us", "share"])
        for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([k, c, f"{t:.1f}", f"{t / tot:.4f}"])
    with open(path, "rb") as fi, gzip.open(os.path.join(dst, "r2_launches_cfg2.csv.gz"), "wb") as fo:
        shutil.copyfileobj(fi, fo)
    print("wrote launches summary", len(agg), "kernels")

# ---- (3) the two tcgen05 edge kernels: details + stall samples per phase ---------------------------------------------------
for k in ("tc_edge_bwd", "tc_edge_fwd"):
    d = os.path.join(src, f"{k}_details.txt")
    if os.path.exists(d) and os.path.getsize(d):
        shutil.copy(d, os.path.join(dst, f"r2_ncu_{k}128_details.txt"))
    sc = os.path.join(src, f"{k}_source.csv")
    if os.path.exists(sc) and os.path.getsize(sc):
        # the export holds one table per captured launch of the kernel: keep the FIRST (layer 0 of the cfg2 training step)
        lines = open(sc).read().splitlines(keepends=True)
        heads = [i for i, ln in enumerate(lines) if ln.startswith('"Address"') or ('"Address"' in ln and '"Source"' in ln)]
        if len(heads) > 1:
            sc = sc + ".first"
            with open(sc, "w") as f:
                f.writelines(lines[:max(heads[1] - 1, heads[0] + 1)])
        ph = subprocess.run([sys.executable, os.path.join(here, "ncu_phases.py"), sc], capture_output=True, text=True).stdout
        with open(os.path.join(dst, f"r2_ncu_{k}128_phases.txt"), "w") as f:
            f.write(f"# stall samples of {k} (d=128, bf16x3, cfg2) summed between synchronisation points (scripts/ncu_phases.py)\n" + ph)
if traffic:
    pub = {"_comment": "dram__bytes_read.sum + dram__bytes_write.sum per launch, one ncu --set full pass (profiles/r2_ncu_all_kernels.csv); "
                       "bench.py copies the entry of its top kernel into roofline.traffic"}
    for k, v in traffic.items():
        pub[k.split("<")[0]] = v
    old = os.path.join(dst, "r1_traffic.json")
    if os.path.exists(old):  # the d = 256 kernels were captured in round 1 (cfg4) and have not changed
        for k, v in json.load(open(old)).items():
            if k.endswith("256") and k not in pub:
                pub[k] = v
    json.dump(pub, open(os.path.join(dst, "r2_traffic.json"), "w"), indent=1)

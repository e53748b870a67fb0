"""`ncu --page raw --csv` of the tile-GEMM kernels (scripts/profile_wide.sh) -> one row per captured launch:
duration, DRAM bytes, L2 bytes, tensor-pipe / SM / DRAM / L2 throughput percentages.
python scripts/summarise_wide.py raw.csv out.csv"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if r]
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr, body = rows[start], rows[start + 2:]
col = {h: i for i, h in enumerate(hdr)}


def f(r, name, scale=1.0):
    try:
        return float(r[col[name]].replace(",", "")) * scale
    except (KeyError, ValueError):
        return float("nan")


want = [("duration_us", "gpu__time_duration.sum", 1e-3), ("dram_read_MB", "dram__bytes_read.sum", 1e-6),
        ("dram_write_MB", "dram__bytes_write.sum", 1e-6), ("l2_MB", "lts__t_bytes.sum", 1e-6),
        ("tensor_pipe_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
        ("sm_throughput_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
        ("dram_throughput_pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
        ("l2_throughput_pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0)]
MODES = ["IMG", "GATE", "ROWDOT", "IMG_BWD", "GATE_BWD", "ROWDOT_BWD", "DGRAD", "F32", "WGRAD"]
with open(sys.argv[2], "w", newline="") as fo:
    w = csv.writer(fo)
    w.writerow(["id", "kernel", "epilogue", "grid", "block"] + [k for k, _, _ in want])
    for r in body:
        if len(r) != len(hdr):
            continue
        name = r[col["Kernel Name"]].replace("void ", "").replace("<unnamed>::", "").split("(")[0]
        mode = ""
        if "tc_wide_gemm<" in name:
            try:
                mode = MODES[int(name.split("<")[1].split(",")[0])]
            except (ValueError, IndexError):
                pass
        w.writerow([r[col["ID"]], name, mode, r[col["Grid Size"]], r[col["Block Size"]]] + [f"{f(r, m, s):.3f}" for _, m, s in want])
print("wrote", sys.argv[2])

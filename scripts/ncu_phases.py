"""Sum stall samples of an `ncu --page source --csv` dump between synchronisation points (BAR.SYNC / mbarrier TRYWAIT):
python scripts/ncu_phases.py file.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
key = ci["Warp Stall Sampling (All Samples)"]
body = [r for r in rows[2:] if len(r) > key]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[key] or 0) for r in body)
seg_start, acc, accs = 0, 0, {}
def flush(i, why):
    global seg_start, acc, accs
    if acc:
        top = sorted(accs.items(), key=lambda kv: -kv[1])[:3]
        print(f"#{seg_start:5d}-{i:5d} {acc:6d} {100*acc/tot:5.1f}%  {why:12s} {top}")
    seg_start, acc, accs = i + 1, 0, {}
for i, r in enumerate(body):
    acc += int(r[key] or 0)
    for s in stalls:
        v = int(r[ci[s]] or 0)
        if v:
            accs[s[6:]] = accs.get(s[6:], 0) + v
    src = r[ci["Source"]]
    if "BAR.SYNC" in src:
        flush(i, "BAR")
    elif "TRYWAIT" in src:
        # include the loop branch that follows
        pass
    elif "BRA" in src and i > 0 and any("TRYWAIT" in body[j][ci["Source"]] for j in range(max(0, i - 8), i)) and int(r[key] or 0) > 50:
        flush(i, "MBAR-loop")
flush(len(body), "end")
print("total", tot)

"""Print the hottest SASS lines (stall samples) of an `ncu --page source --csv` dump: python scripts/ncu_hot.py file.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ci = {h: i for i, h in enumerate(hdr)}
key = ci["Warp Stall Sampling (All Samples)"]
body = [r for r in rows[2:] if len(r) > key]
tot = sum(int(r[key] or 0) for r in body)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot)
for idx, r in enumerate(body):
    r.append(idx)
for r in sorted(body, key=lambda r: -int(r[key] or 0))[:n]:
    why = sorted(((int(r[ci[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"{int(r[key]):6d} #{r[-1]:5d} {r[ci['Source']][:110]:110s} {why}")

"""Top stall sites of one kernel from `ncu -i x.ncu-rep --page source --csv` (SASS view).
usage: python tools/ncu_hot.py source.csv [top=25]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
H = rows[1]
ix = {h: i for i, h in enumerate(H)}
stalls = [h for h in H if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    if len(r) < len(H):
        continue
    try:
        n = int(r[ix["# Samples"]])
    except ValueError:
        continue
    data.append((n, r))
tot = sum(n for n, _ in data)
print("total samples", tot, "instructions", len(data))
for pos, (n, r) in enumerate(data):
    r.append(pos)
for n, r in sorted(data, key=lambda t: -t[0])[:top]:
    st = sorted(((int(r[ix[s]] or 0), s) for s in stalls), reverse=True)[:2]
    print(f"{100 * n / tot:5.1f}%  #{r[-1]:5d} exec={r[ix['Instructions Executed']]:>9s}  {r[ix['Source']][:90]:90s} {st}")

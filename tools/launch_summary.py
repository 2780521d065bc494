"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py: per-kernel launches, total
time and share of ONE decode step (the 4th step = first timed one after 3 warm-ups).
usage: python tools/launch_summary.py launches.csv [step_index=3] > table.md ; also writes the step's rows with --rows out.csv"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    step = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 3
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr, rows = rows[0], rows[1:]
    ki, vi, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    names = [r[ki] for r in rows]
    # a step starts with cond_kernel (engine::run)
    starts = [i for i, n in enumerate(names) if n.startswith("cond_kernel")]
    if step + 1 < len(starts):
        lo, hi = starts[step], starts[step + 1]
    else:
        lo, hi = starts[step], len(rows)
    sel = rows[lo:hi]
    agg = collections.OrderedDict()
    for r in sel:
        n = re.sub(r"\(.*", "", r[ki]).replace("void ", "")
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi]) / 1e6
    tot = sum(a[1] for a in agg.values())
    print(f"one step = launches [{lo}, {hi}) of {len(rows)}: {hi - lo} launches, {tot:.3f} ms (serialised, cold cache)\n")
    print("| kernel | launches | sum ms | share |\n|---|---|---|---|")
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{n}` | {a[0]} | {a[1]:.3f} | {100 * a[1] / tot:.1f}% |")
    if "--rows" in sys.argv:
        out = sys.argv[sys.argv.index("--rows") + 1]
        with open(out, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["kernel", "grid", "ns"])
            for r in sel:
                w.writerow([re.sub(r"\(.*", "", r[ki]).replace("void ", ""), r[gi], r[vi]])


if __name__ == "__main__":
    main()

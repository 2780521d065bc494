"""Prints a compact per-launch table from `ncu -i x.ncu-rep --page raw --csv` output.
usage: python tools/ncu_table.py raw.csv [more metric names...]"""
import csv
import re
import sys

COLS = [("gpu__time_duration.sum", "us"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__t_sector_hit_rate.pct", "L2hit%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("smsp__issue_active.avg.pct", "issue%"), ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bankconf"),
        ("launch__registers_per_thread", "regs"), ("launch__occupancy_limit_shared_mem", "occ_smem"),
        ("launch__occupancy_limit_registers", "occ_regs")]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, rows = rows[0], rows[1], rows[2:]
    cols = COLS + [(m, m[-12:]) for m in sys.argv[2:]]
    idx = {h: i for i, h in enumerate(hdr)}
    print("| kernel | grid | " + " | ".join(n for _, n in cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for r in rows:
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]]).replace("void ", "")
        vals = []
        for m, n in cols:
            if m not in idx:
                vals.append("-")
                continue
            v, u = r[idx[m]], units[idx[m]]
            try:
                f = float(v.replace(",", ""))
                if u in ("byte", "Byte"):
                    f /= 1e6
                elif u == "Kbyte":
                    f /= 1e3
                elif u == "Gbyte":
                    f *= 1e3
                elif u in ("ns", "nsecond"):
                    f /= 1e3
                elif u in ("ms", "msecond"):
                    f *= 1e3
                vals.append(f"{f:.1f}")
            except ValueError:
                vals.append(v)
        print(f"| `{name}` | {r[idx['Grid Size']]} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main()

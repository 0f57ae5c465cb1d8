#!/usr/bin/env python
"""Key figures of the `ncu --page raw --csv` exports under profiles/ (or the given files): per launch duration, DRAM
bytes, throughput, registers, occupancy, pipe utilisation, top stall reasons.
    python tools/ncu_summary.py profiles/r02_prof_*_raw.csv
"""
import csv
import sys

KEYS = [
    ("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__t_sector_hit_rate.pct", "L2hit%"), ("launch__registers_per_thread", "regs"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64pipe%"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64cyc%"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fmapipe%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alupipe%"),
    ("sm__issue_active.avg.pct_of_peak_sustained_active", "issue%"), ("smsp__inst_executed.sum", "warp_inst"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("launch__grid_size", "grid"), ("launch__waves_per_multiprocessor", "waves"),
]
STALL = "smsp__average_warp"


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return None


def main(files):
    for f in files:
        rows = list(csv.reader(open(f)))
        hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
        names, units, data = rows[hdr], rows[hdr + 1], rows[hdr + 2:]
        col = {n: i for i, n in enumerate(names)}
        print("==", f)
        for r in data:
            if len(r) < len(names):
                continue
            out = [r[col["Kernel Name"]][:48]]
            for k, short in KEYS:
                if k in col:
                    v, u = num(r[col[k]]), units[col[k]]
                    if v is None:
                        continue
                    if short == "dur":
                        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
                        out.append("dur=%.1fus" % (v * scale))
                    elif short.startswith("dram_"):
                        scale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
                        out.append("%s=%.1fMB" % (short, v * scale))
                    else:
                        out.append("%s=%g" % (short, round(v, 2)))
            print("  " + " ".join(out))
            st = sorted(((num(r[i]) or 0, n) for n, i in col.items() if n.startswith("smsp__average_warps_issue_stalled") and n.endswith("_per_issue_active.ratio")), reverse=True)[:5]
            if st:
                print("     stalls/issue: " + ", ".join("%s=%.2f" % (n.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v) for v, n in st))


if __name__ == "__main__":
    main(sys.argv[1:])

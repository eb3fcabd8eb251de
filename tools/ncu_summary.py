#!/usr/bin/env python
"""Condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into a per-kernel table of the
metrics the roofline discussion uses.  Usage: tools/ncu_summary.py REPORT.ncu-rep > profiles/x.md"""
import csv
import io
import statistics
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu%"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_static", "smem_static"),
    ("launch__shared_mem_per_block_dynamic", "smem_dyn"),
]


def to_float(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                         text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    kname = col["Kernel Name"]
    groups = {}
    for r in body:
        groups.setdefault(r[kname].split("(")[0], []).append(r)
    print(f"# ncu --set full summary of `{rep}` (median over the captured launches of each kernel)\n")
    names = [m for m, _ in METRICS if m in col]
    print("| kernel | launches | " + " | ".join(short for m, short in METRICS if m in col) + " |")
    print("|---|---|" + "---|" * len(names))
    for k, rs in groups.items():
        cells = []
        for m in names:
            vals = [to_float(r[col[m]]) for r in rs]
            vals = [v for v in vals if v is not None]
            if not vals:
                cells.append("-")
                continue
            v = statistics.median(vals)
            u = units[col[m]]
            cells.append(f"{v:.4g} {u}".strip() if u not in ("%", "") else f"{v:.4g}")
        print(f"| `{k.replace('void ', '')}` | {len(rs)} | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()

"""Summarises an ncu report (read here, no GPU needed) into a small markdown table for profiles/.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/ncu_probe_rN.md
"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
    ("launch__registers_per_thread", "regs"),
    ("smsp__inst_executed.sum", "warp inst"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_sb"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_sb"),
]


def to_bytes(value, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value.replace(",", "")) * scale.get(unit, 1)


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    cols = [(m, label, hdr.index(m)) for m, label in METRICS if m in hdr]
    print(f"ncu --set full --clock-control none summary of `{rep}` (per launch; serialised, cold caches)\n")
    print("| # | kernel | " + " | ".join(label for _, label, _ in cols) + " |")
    print("|---|---|" + "---|" * len(cols))
    for n, r in enumerate(rows[2:]):
        cells = []
        for m, label, i in cols:
            v, u = r[i], units[i]
            if "bytes" in m:
                cells.append(f"{to_bytes(v, u) / 1e9:.3f} GB")
            elif m == "gpu__time_duration.sum":
                t = float(v.replace(",", ""))
                t = {"ns": t / 1e6, "us": t / 1e3, "ms": t, "s": t * 1e3}.get(u, t)
                cells.append(f"{t:.3f} ms")
            else:
                try:
                    cells.append(f"{float(v.replace(',', '')):.4g}")
                except ValueError:
                    cells.append(v)
        name = r[ki].split("(")[0].replace("void ", "").replace("rpt::", "")
        print(f"| {n} | `{name[:70]}` | " + " | ".join(cells) + " |")


if __name__ == "__main__":
    main()

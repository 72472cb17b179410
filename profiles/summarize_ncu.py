#!/usr/bin/env python
"""ncu -i X.ncu-rep --page raw --csv | python profiles/summarize_ncu.py "title" > profiles/NAME.md
Keeps the handful of metrics the roofline discussion in DESIGN.md uses."""
import csv
import sys

WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %peak"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %peak"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_sb"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_sb"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg_throttle"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
        ("smsp__inst_executed.sum", "warp instructions")]
rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
print(f"# {sys.argv[1] if len(sys.argv) > 1 else 'ncu --set full summary'}\n")
print("| kernel | " + " | ".join(n for _, n in WANT) + " |")
print("|---|" + "---:|" * len(WANT))
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")].replace("void ", "").split("(")[0][:70]
    cells = []
    for m, _ in WANT:
        if m in hdr:
            i = hdr.index(m)
            v = r[i]
            try:
                v = f"{float(v):.4g}"
            except ValueError:
                pass
            cells.append(f"{v} {units[i]}".strip())
        else:
            cells.append("n/a")
    print(f"| `{name}` | " + " | ".join(cells) + " |")

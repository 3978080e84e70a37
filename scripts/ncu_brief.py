#!/usr/bin/env python
"""Brief per-kernel table from an ncu report: `python scripts/ncu_brief.py report.ncu-rep` (uses `ncu -i ... --page raw --csv`)."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue active %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
    ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "fmaheavy cycles %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu pipe %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu pipe %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem wavefronts %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / inst"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("smsp__average_warp_latency_per_inst_issued.ratio", "warp cycles per issued inst"),
]
STALL = "smsp__average_warps_issue_stalled_"
names = [r[hdr.index("Kernel Name")].split("(")[0] for r in data]
print("| metric | " + " | ".join(names) + " |")
print("|---|" + "---|" * len(names))
for key, label in WANT:
    if key in hdr:
        i = hdr.index(key)
        print(f"| {label} ({units[i]}) | " + " | ".join(r[i] for r in data) + " |")
for i, h in enumerate(hdr):
    if h.startswith(STALL) and h.endswith("_per_issue_active.ratio"):
        vals = [float(r[i] or 0) for r in data]
        if max(vals) >= 0.15:
            print(f"| stall {h[len(STALL):-len('_per_issue_active.ratio')]} (warps / issue) | " + " | ".join(f"{v:.2f}" for v in vals) + " |")

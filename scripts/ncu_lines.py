#!/usr/bin/env python
"""Summarise an ncu report here (no GPU): headline metrics per launch and the source lines with most stall samples.
   python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [kernel-regex] [launch-index] [top-n]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else "."
launch = int(sys.argv[3]) if len(sys.argv) > 3 else 0
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]
ki = hdr.index("Kernel Name")
print("launches:", [r[ki][:40] for r in rows[2:]])
for w in WANT:
    if w in hdr:
        i = hdr.index(w)
        print(f"{w:95s} {units[i]:14s}", [r[i] for r in rows[2:]])

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", f"regex:{kern}",
                      "--launch-skip", str(launch), "--launch-count", "1"], capture_output=True, text=True).stdout
agg = collections.OrderedDict()
fname, head = None, None
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        head = r
    elif head is not None and len(r) >= 8 and r[0] != "" and r[0] != "Function Name":
        try:
            agg[(fname, int(r[0]), r[1].strip()[:100])] = (int(r[4]), int(r[7]))
        except ValueError:
            pass
tot = sum(v[0] for v in agg.values()) or 1
toti = sum(v[1] for v in agg.values())
print(f"\nlaunch {launch}: {tot} stall samples, {toti} warp instructions; top {topn} source lines")
for (f, l, text), (s, i) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{s:7d} {100 * s / tot:5.1f}%  inst {i:10d} {100 * i / max(toti, 1):5.1f}%  {f}:{l}  {text}")

#!/usr/bin/env python
"""Turn the ncu outputs a GPU visit left in gpurun_out/ into the small tracked summaries under profiles/ (run here,
no GPU).   python scripts/summarise_profiles.py <tag> [--launches gpurun_out/launches.csv] [--full gpurun_out/x.ncu-rep]
           [--traffic-key onf_train_fwd_bwd_tensor --traffic-launch 2]"""
import argparse
import collections
import csv
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]


def launches(path, out, note):
    rows = [r for r in csv.reader(open(path)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(r[ui], 1.0)
        name = r[ki].split("(")[0]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(out, "w") as f:
        f.write(f"# {note}\n# source: ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n")
        f.write("kernel,launches,total_us,share\n")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k},{n},{t:.1f},{t / tot:.4f}\n")
    print(open(out).read())


def full(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    res = []
    for n, r in enumerate(rows[2:]):
        d = {"launch": n, "kernel": r[ki].split("(")[0]}
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                try:
                    d[k] = float(r[i].replace(",", ""))
                except ValueError:
                    d[k] = r[i]
                d.setdefault("_units", {})[k] = units[i]
        res.append(d)
    json.dump(res, open(out, "w"), indent=1)
    for d in res:
        print(d["kernel"], d.get("gpu__time_duration.sum"), d.get("dram__bytes_read.sum"), d.get("dram__bytes_write.sum"))
    return res


def to_bytes(v, unit):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("tag")
    ap.add_argument("--launches")
    ap.add_argument("--full")
    ap.add_argument("--note", default="")
    ap.add_argument("--traffic-key")
    ap.add_argument("--traffic-launch", type=int, default=0)
    a = ap.parse_args()
    if a.launches:
        launches(a.launches, os.path.join(ROOT, "profiles", f"{a.tag}_launches_summary.csv"), a.note)
    if a.full:
        name = os.path.join("profiles", f"{a.tag}_ncu_full.json")
        res = full(a.full, os.path.join(ROOT, name))
        if a.traffic_key:
            d = res[a.traffic_launch]
            rd = to_bytes(d["dram__bytes_read.sum"], d["_units"]["dram__bytes_read.sum"])
            wr = to_bytes(d["dram__bytes_write.sum"], d["_units"]["dram__bytes_write.sum"])
            p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
            t = json.load(open(p)) if os.path.exists(p) else {}
            import sys
            sys.path.insert(0, ROOT)
            import bench                                          # the sha bench.py will compare against at run time
            commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
            dirty = subprocess.run(["git", "-C", ROOT, "status", "--porcelain", "--", "pytorch-motion-planner_b200/csrc"],
                                   capture_output=True, text=True).stdout.strip()
            t[a.traffic_key] = {
                "bytes": rd + wr, "read": rd, "write": wr, "sources_sha": bench.kernel_sources_sha(),
                "commit": commit + ("+uncommitted csrc changes" if dirty else ""),
                "source": f"{name} launch {a.traffic_launch} ({a.note})",
                "launch_shape": f"grid {int(d.get('launch__grid_size', 0))} x block {int(d.get('launch__block_size', 0))}, "
                                f"{d.get('gpu__time_duration.sum')} {d['_units'].get('gpu__time_duration.sum')} under ncu"}
            json.dump(t, open(p, "w"), indent=1)


if __name__ == "__main__":
    main()

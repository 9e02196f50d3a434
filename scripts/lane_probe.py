#!/usr/bin/env python
"""Does splitting the problem batch into independent lanes (one engine + CUDA graph + stream per lane) raise
throughput?  Lanes are independent planning problems, so there is no cross-lane dependency: the GPU may overlap the
HBM-bound Adam phase / small kernels of one lane with the tensor phases of another.
    python scripts/lane_probe.py [--problems 1024] [--lanes 1 2 4] [--steps 200]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import bench as B  # noqa: E402
from nfopp_b200 import _lib as L  # noqa: E402
from nfopp_b200.batched import BatchedPlanner  # noqa: E402


def run(problems, lanes, steps, warmup, workload):
    per = problems // lanes
    pls = []
    for i in range(lanes):
        cfg, env, starts, goals, _ = B.build_workload(workload, per, i * per)
        pl = BatchedPlanner(cfg, lanes=1)
        pl.init(starts, goals, **env)
        pls.append(pl)
    lib = L.lib()

    def go(n):
        cur = torch.cuda.current_stream()
        for pl in pls:
            pl.engine._stream.wait_stream(cur)
        for _ in range(n):
            for pl in pls:
                with torch.cuda.stream(pl.engine._stream):
                    L.check(lib.nfopp_engine_step(pl.engine._h, 1, 1, L.stream_ptr()))
        for pl in pls:
            cur.wait_stream(pl.engine._stream)

    go(warmup)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    go(steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(f"{workload}: problems {problems} lanes {lanes}: {ms:.4f} ms/step  {problems / ms * 1e3:.0f} problem-steps/s", flush=True)
    for pl in pls:
        pl.engine.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--problems", type=int, default=1024)
    ap.add_argument("--lanes", type=int, nargs="+", default=[1, 2, 4])
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="corridor")
    args = ap.parse_args()
    for lanes in args.lanes:
        run(args.problems, lanes, args.steps, args.warmup, args.workload)
    # lanes inside one engine + one CUDA graph (nfopp_engine_set_lanes)
    for lanes in args.lanes:
        cfg, env, starts, goals, _ = B.build_workload(args.workload, args.problems, 0)
        pl = BatchedPlanner(cfg, lanes=lanes)
        pl.init(starts, goals, **env)
        pl.step(args.warmup)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pl.step(args.steps)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        print(f"{args.workload}: in-graph lanes {lanes}: {ms:.4f} ms/step  {args.problems / ms * 1e3:.0f} problem-steps/s", flush=True)
        pl.engine.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Long-run stability check: thousands of graphed steps of every bench workload; paths must stay finite and no problem may
raise a status flag.   python scripts/soak.py [--steps 5000]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import bench as B  # noqa: E402
from nfopp_b200.batched import BatchedPlanner  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5000)
    a = ap.parse_args()
    for workload, problems, steps in (("corridor", 1024, a.steps), ("forest", 1024, a.steps), ("gridmap-wide", 256, a.steps // 10)):
        cfg, env, starts, goals, _ = B.build_workload(workload, problems, 0)
        pl = BatchedPlanner(cfg)
        pl.init(starts, goals, **env)
        t0 = time.time()
        pl.step(steps)
        torch.cuda.synchronize()
        paths = pl.paths()
        status = pl.status()
        collides, length, _ = pl.evaluate_paths(4)
        print(f"{workload}: {problems} problems x {steps} steps in {time.time() - t0:.1f} s; finite={bool(torch.isfinite(paths).all())} "
              f"status flags set on {int((status != 0).sum())} problems; collision-free {int((~collides).sum())}/{problems}; "
              f"mean length {float(length.mean()):.3f}", flush=True)
        pl.engine.close()


if __name__ == "__main__":
    main()

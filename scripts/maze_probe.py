#!/usr/bin/env python
"""A* initialiser vs straight-line initialiser on corridor mazes (bench-mr-style, own generator): colliding poses of the
initial paths and problems solved under the reference's stop rule.   python scripts/maze_probe.py [--problems 64]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from nfopp_b200.batched import BatchedPlanner, forest_config  # noqa: E402
from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--problems", type=int, default=64)
    ap.add_argument("--resolution", type=float, default=0.09)
    ap.add_argument("--iterations", type=int, default=600)
    a = ap.parse_args()
    P = a.problems
    starts, goals, bounds, grids, origin, cell = B.make_corridor_grid_batch(P, seed=2)
    for name, res in (("straight line", None), ("A*", a.resolution)):
        bp = BatchedPlanner(forest_config(P, cell=cell, seed=5))
        bp.init(starts, goals, grids=grids, astar_resolution=res)
        hits0 = bp.evaluate_paths(4)[2]
        out = bp.solve(max_iterations=a.iterations, min_iterations=200, check_every=50)
        torch.cuda.synchronize()
        print(f"{name}: colliding poses of the initial paths: mean {float(hits0.float().mean()):.1f}, "
              f"collision-free at start {int((hits0 == 0).sum())}/{P}; solved within {a.iterations} iterations: "
              f"{int(out['solved'].sum())}/{P}", flush=True)
        bp.engine.close()


if __name__ == "__main__":
    main()

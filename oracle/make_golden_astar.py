"""Generates tests/golden/astar.npz from the UNMODIFIED reference (container only; /root/reference must exist):
AstarTrajectoryInitializer (astar/astar_trajectory_initializer.py) with the reference's own collision checkers on a few
random obstacle worlds, recording the labelled cell matrix, the A* cell path, the re-parametrised x, y trajectory and
the initial headings - and checks oracle/astar.py against every recorded case while doing so.

    python -m oracle.make_golden_astar
"""
import contextlib
import os

import numpy as np
import torch

from . import astar as A
from . import checkers as C
from .reference_bridge import load_reference

CASES = [
    # name, checker, bounds, resolution, obstacle count, seed, T
    ("disc_coarse", "circle", (0.0, 3.0, 0.0, 3.0), 0.25, 14, 1, 100),
    ("disc_fine", "circle", (0.0, 3.0, 0.0, 3.0), 0.1, 25, 2, 100),
    ("box_room", "rectangle", (0.0, 10.0, -2.0, 6.0), 0.5, 60, 3, 100),
    ("box_short", "rectangle", (0.0, 10.0, -2.0, 6.0), 0.5, 60, 4, 20),
    ("disc_blocked", "circle", (0.0, 3.0, 0.0, 3.0), 0.25, 0, 5, 50),        # a wall: goal unreachable -> last expanded state
]


def make_world(name, kind, bounds, n_obs, seed):
    rng = np.random.RandomState(seed)
    w, h = bounds[1] - bounds[0], bounds[3] - bounds[2]
    if name == "disc_blocked":
        ys = np.linspace(bounds[2], bounds[3], 40)
        obs = np.stack([np.full_like(ys, bounds[0] + 0.5 * w), ys], 1)
    else:
        obs = np.stack([bounds[0] + w * (0.2 + 0.6 * rng.rand(n_obs)), bounds[2] + h * rng.rand(n_obs)], 1)
    start = np.array([bounds[0] + 0.07 * w, bounds[2] + h * (0.2 + 0.6 * rng.rand()), 0.3], dtype=np.float32)
    goal = np.array([bounds[0] + 0.93 * w, bounds[2] + h * (0.2 + 0.6 * rng.rand()), -0.4], dtype=np.float32)
    return obs, start, goal


def main():
    ref = load_reference()
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        from neural_field_optimal_planner.astar.astar_trajectory_initializer import AstarTrajectoryInitializer
        from neural_field_optimal_planner.collision_checker import CircleDirectedCollisionChecker, RectangleCollisionChecker
    out = {}
    box = (-0.3, 0.2, -0.3, 0.2)
    for name, kind, bounds, res, n_obs, seed, T in CASES:
        obs, start, goal = make_world(name, kind, bounds, n_obs, seed)
        if kind == "circle":
            cc = CircleDirectedCollisionChecker(0.3, bounds)
            label = lambda p: C.circle_labels(p, obs, 0.3, bounds)          # noqa: E731
        else:
            cc = RectangleCollisionChecker(box, bounds)
            label = lambda p: C.rectangle_labels(p, obs, box, bounds)       # noqa: E731
        cc.update_obstacle_points(obs)
        init = AstarTrajectoryInitializer(cc, res)
        traj = torch.zeros(T, 3)
        with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
            cells_xy = init.calculate_astar_path(start, goal)               # float32 cell centres
            init.initialize_trajectory(traj, torch.tensor(start)[None], torch.tensor(goal)[None])
        traj = traj.detach().numpy()
        # ---- the oracle restatement against the live reference ----
        xy, cells, matrix = A.initial_trajectory_xy(start, goal, bounds, res, label, T)
        pts = A.cells_to_points(cells, bounds, res)
        assert pts.shape == cells_xy.shape and np.array_equal(pts, cells_xy), name
        err = np.abs(xy - traj[:, :2]).max()
        assert err <= 2e-6, (name, err)
        print(f"{name}: grid {matrix.shape}, {len(cells)} cells on the path, occupied {int(matrix.sum())}, "
              f"oracle vs reference trajectory max abs err {err:.2e}")
        for k, v in dict(obstacles=obs, start=start, goal=goal, bounds=np.array(bounds), resolution=np.array(res),
                         kind=np.array(kind), T=np.array(T), matrix=matrix, cells=cells, cell_points=cells_xy,
                         trajectory=traj).items():
            out[f"{name}/{k}"] = v
    out["names"] = np.array([c[0] for c in CASES])
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "astar.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

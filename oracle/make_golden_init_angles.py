"""Generates tests/golden/init_angles.npz from the UNMODIFIED reference (container only): TrajectoryInitializer with
init_angles_with_trajectory=True (trajectory_initializer.py:12-44) - straight-line x, y, and a curved x, y already in the
trajectory (the A* initialiser's situation) - recording the trajectory before and after `initialize_angle`.

    python -m oracle.make_golden_init_angles
"""
import contextlib
import os

import numpy as np
import torch

from .reference_bridge import load_reference


def main():
    load_reference()
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        from neural_field_optimal_planner.trajectory_initializer import TrajectoryInitializer
    out = {}
    cases = [("line_even", 100, (0.5, 0.5, 0.3), (2.5, 1.5, -0.4), False),
             ("line_odd", 21, (2.0, 0.2, 3.0), (0.3, 2.6, -3.0), False),
             ("curved", 100, (0.5, 0.5, 0.0), (2.5, 2.5, 1.2), True),
             ("curved_odd", 37, (2.8, 0.4, -2.9), (0.2, 0.3, 2.7), True)]
    for name, T, start, goal, curved in cases:
        start_t, goal_t = torch.tensor([start], dtype=torch.float32), torch.tensor([goal], dtype=torch.float32)
        init = TrajectoryInitializer(None, init_angles_with_trajectory=True)
        traj = torch.zeros(T, 3)
        with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):      # the reference prints the angles
            if curved:
                s = torch.linspace(0, 1, T + 2)[1:-1]
                traj[:, 0] = start[0] + (goal[0] - start[0]) * s + 0.4 * torch.sin(2 * np.pi * s)
                traj[:, 1] = start[1] + (goal[1] - start[1]) * s ** 2 + 0.3 * torch.sin(3 * np.pi * s)
                before = traj.clone()
                init.initialize_angle(traj, start_t, goal_t)
            else:
                before = traj.clone()
                init.initialize_trajectory(traj, start_t, goal_t)
        out[f"{name}/before"] = before.numpy()
        out[f"{name}/after"] = traj.numpy()
        out[f"{name}/start"] = start_t.numpy()
        out[f"{name}/goal"] = goal_t.numpy()
        out[f"{name}/curved"] = np.array(curved)
        print(name, "heading range", float(traj[:, 2].min()), float(traj[:, 2].max()))
    out["names"] = np.array([c[0] for c in cases])
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "init_angles.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()

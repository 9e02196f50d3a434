"""Straight-line x,y and wrapped-linear heading initialisation (reference trajectory_initializer.py:7-44).
The planner engine performs it on the device for all problems (PlannerEngine.init_trajectories); this class
keeps the reference's constructor so PlannerFactory configs resolve unchanged."""
import torch

from .engine import batched_linspace
from .torch_math import wrap_angle


def blend_headings_with_path_direction(trajectories, starts, goals):
    """`init_angles_with_trajectory=True` (trajectory_initializer.py:31-44) for a batch, in place: the direction of travel
    at every node (central difference over the full path start | nodes | goal) is blended into the linearly interpolated
    headings with a triangular weight - 0 at both ends, 1 in the middle - so the end headings stay the requested ones.
    trajectories [P, T, 3], starts / goals [P, 3]; any device."""
    with torch.no_grad():
        xy = torch.cat([starts[:, None, :2], trajectories[..., :2], goals[:, None, :2]], dim=1)       # [P, T+2, 2]
        chord = xy[:, 2:] - xy[:, :-2]
        direction = torch.atan2(chord[..., 1], chord[..., 0])
        T = trajectories.shape[1]
        ramp = torch.cat([torch.linspace(0., 1, T // 2), torch.linspace(1., 0, (T + 1) // 2)]).to(trajectories)
        trajectories[..., 2] += wrap_angle(direction - trajectories[..., 2]) * ramp
    return trajectories


class TrajectoryInitializer(object):
    def __init__(self, collision_checker, init_angles_with_trajectory=False):
        self._collision_checker = collision_checker
        self._init_angles_with_trajectory = init_angles_with_trajectory

    def initialize_trajectory(self, trajectory, start_point, goal_point):
        """trajectory [T,3] (modified in place), start/goal [1,3]."""
        with torch.no_grad():
            n = trajectory.shape[0] + 2
            for d in range(2):
                trajectory[:, d] = batched_linspace(start_point[:, d], goal_point[:, d], n)[0, 1:-1]
            self.initialize_angle(trajectory, start_point, goal_point)

    def initialize_angle(self, trajectory, start_point, goal_point):
        """Headings: linear from the start heading to the goal heading along the wrapped difference (:21-29), then, if asked
        for, pulled towards the direction of travel of the x, y already in `trajectory` (:31-44)."""
        with torch.no_grad():
            n = trajectory.shape[0] + 2
            goal_angle = wrap_angle(goal_point[:, 2] - start_point[:, 2]) + start_point[:, 2]
            trajectory[:, 2] = batched_linspace(start_point[:, 2], goal_angle, n)[0, 1:-1]
            if self._init_angles_with_trajectory:
                blend_headings_with_path_direction(trajectory[None], start_point[:1], goal_point[:1])

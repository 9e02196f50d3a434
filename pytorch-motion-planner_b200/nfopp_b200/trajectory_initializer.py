"""Straight-line x,y and wrapped-linear heading initialisation (reference trajectory_initializer.py:7-29).
The planner engine performs it on the device for all problems (PlannerEngine.init_trajectories); this class
keeps the reference's constructor so PlannerFactory configs resolve unchanged."""
import torch

from .engine import batched_linspace
from .torch_math import wrap_angle


class TrajectoryInitializer(object):
    def __init__(self, collision_checker, init_angles_with_trajectory=False):
        self._collision_checker = collision_checker
        self._init_angles_with_trajectory = init_angles_with_trajectory
        if init_angles_with_trajectory:
            raise NotImplementedError("init_angles_with_trajectory is outside the accelerated path (SURVEY.md §8f)")

    def initialize_trajectory(self, trajectory, start_point, goal_point):
        """trajectory [T,3] (modified in place), start/goal [1,3]."""
        with torch.no_grad():
            n = trajectory.shape[0] + 2
            for d in range(2):
                trajectory[:, d] = batched_linspace(start_point[:, d], goal_point[:, d], n)[0, 1:-1]
            goal_angle = wrap_angle(goal_point[:, 2] - start_point[:, 2]) + start_point[:, 2]
            trajectory[:, 2] = batched_linspace(start_point[:, 2], goal_angle, n)[0, 1:-1]

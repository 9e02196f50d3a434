"""Disc robot against obstacle points (reference collision_checker/circle_collision_checker.py:5-14)."""
from .. import _lib as L
from .collision_checker import CollisionChecker, _bounds_ptr


class CircleCollisionChecker(CollisionChecker):
    kind = "circle"

    def __init__(self, robot_radius, boundaries=None):
        super().__init__(boundaries)
        self._robot_radius = robot_radius

    def _label(self, poses):
        obs, m = self._obstacles_dev()
        return self._run(poses, lambda lib, p, out, n, d, s: lib.nfopp_label_circle(
            p, 1, n, d, L.ptr(obs), None, 1, m, float(self._robot_radius), _bounds_ptr(self._boundaries), out, s))

    def engine_spec(self):
        spec = super().engine_spec()
        spec.update(robot_radius=float(self._robot_radius))
        return spec

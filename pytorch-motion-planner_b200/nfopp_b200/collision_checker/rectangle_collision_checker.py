"""Oriented-box robot against obstacle points (reference rectangle_collision_checker.py:6-26)."""
import ctypes as C

from .. import _lib as L
from .collision_checker import CollisionChecker, _bounds_ptr


class RectangleCollisionChecker(CollisionChecker):
    kind = "rectangle"

    def __init__(self, box, collision_boundaries=None):
        super().__init__(collision_boundaries)
        self._box = box

    def _label(self, poses):
        if poses.shape[1] != 3:
            raise ValueError("RectangleCollisionChecker needs poses with a heading (Position2 or [N,3])")
        obs, m = self._obstacles_dev()
        box = (C.c_double * 4)(*[float(b) for b in self._box])
        return self._run(poses, lambda lib, p, out, n, d, s: lib.nfopp_label_rectangle(
            p, 1, n, L.ptr(obs), None, 1, m, box, _bounds_ptr(self._boundaries), out, s))

    def engine_spec(self):
        spec = super().engine_spec()
        spec.update(box=tuple(float(b) for b in self._box))
        return spec

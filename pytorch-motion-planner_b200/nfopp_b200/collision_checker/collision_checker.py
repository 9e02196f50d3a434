"""Ground-truth collision checkers with the reference's interface (collision_checker/collision_checker.py:4-28),
labelling on the GPU through libnfopp's float64 kernels (bit-exact w.r.t. the reference's numpy arithmetic).

`check_collision` accepts what the reference's callers pass: a raw [N,2] array (holonomic planner,
nerf_opt_planner.py:93-96) or a Position2 (constrained planner, constrained_nerf_opt_planner.py:49-52)."""
import ctypes as C

import numpy as np
import torch

from .. import _lib as L


def _poses_of(test_positions):
    """-> float64 [N, 2|3] array."""
    if hasattr(test_positions, "translation"):
        xy = np.asarray(test_positions.translation, dtype=np.float64).reshape(-1, 2)
        th = np.asarray(test_positions.rotation, dtype=np.float64).reshape(-1, 1)
        return np.ascontiguousarray(np.concatenate([xy, th], axis=1))
    return np.ascontiguousarray(np.asarray(test_positions, dtype=np.float64))


def _bounds_ptr(bounds):
    return None if bounds is None else (C.c_double * 4)(*[float(b) for b in bounds])


class CollisionChecker(object):
    kind = "boundary"

    def __init__(self, collision_boundaries=None):
        self._obstacle_points = np.zeros((0, 2))
        self._boundaries = collision_boundaries
        self._version = 0          # bumped on every environment change so planners re-upload it

    # ---- reference interface ----------------------------------------------------------------------
    def check_collision(self, test_positions):
        if type(self) is CollisionChecker and self._boundaries is None:
            return False           # reference returns the scalar False when there is nothing to check (:13-14)
        return self._label(_poses_of(test_positions))

    def _check_boundaries_collision(self, test_positions):
        if self._boundaries is None:
            return False
        poses = _poses_of(test_positions)
        return self._run(poses, lambda lib, p, out, n, d, s: lib.nfopp_label_boundary(
            p, 1, n, d, _bounds_ptr(self._boundaries), out, s))

    def update_obstacle_points(self, points):
        self._obstacle_points = points
        self._version += 1

    def update_boundaries(self, boundaries):
        self._boundaries = boundaries
        self._version += 1

    def get_boundaries(self):
        return self._boundaries

    # ---- CUDA plumbing ---------------------------------------------------------------------------
    def _label(self, poses):
        return self._run(poses, lambda lib, p, out, n, d, s: lib.nfopp_label_boundary(
            p, 1, n, d, _bounds_ptr(self._boundaries), out, s))

    @staticmethod
    def _run(poses, call):
        L.require_cuda()
        n, d = poses.shape
        dev = torch.as_tensor(poses, device="cuda")
        out = torch.empty(n, dtype=torch.uint8, device="cuda")
        L.check(call(L.lib(), L.ptr(dev), L.ptr(out), n, d, L.stream_ptr()))
        return out.cpu().numpy().astype(bool)

    def _obstacles_dev(self):
        pts = np.ascontiguousarray(np.asarray(self._obstacle_points, dtype=np.float64).reshape(-1, 2))
        return torch.as_tensor(pts, device="cuda"), len(pts)

    def engine_spec(self):
        """What PlannerEngine needs to reproduce this checker on the device."""
        return dict(checker=self.kind, checker_bounds=None if self._boundaries is None else tuple(
            float(b) for b in self._boundaries))

"""PathPostprocessor (ros/path_postprocessor.py:8-71) on the device: what the ROS adapter applies to every path before
publishing it - distance filter, re-sampling every `distance_step` metres with a quadratic interpolating spline over the
chord-length parametrisation (headings unfolded), direction filter at the head of the path.

`process(trajectory)` keeps the reference's one-path interface (Position2 in, Position2 out); `process_batch(paths)` takes
`[P, n, 3]` float32 paths (BatchedPlanner.paths()) and returns the padded poses with their counts.  Both run
`nfopp_postprocess_paths` (csrc/postprocess.cu); there is no host fallback."""
import ctypes as C

import numpy as np
import torch

from . import _lib as L
from .utils.position2 import Position2


class PathPostprocessor(object):
    def __init__(self, minimal_distance=0.001, distance_step=0.05, device="cuda"):
        self._distance_step = distance_step
        self._minimal_distance = minimal_distance
        self._device = device

    def process_batch(self, paths, max_points=None):
        """paths [P, n, 3] float32 (host or device) -> (poses [P, max_points, 3] float64, counts [P] int32), on the device.
        counts[p] = -1: fewer than 3 distinct nodes (the reference's scipy call raises for such a path)."""
        L.require_cuda()
        lib = L.lib()
        paths = torch.as_tensor(paths, dtype=torch.float32).to(self._device).contiguous()
        P, n, d = paths.shape
        if d != 3:
            raise ValueError(f"paths must be [P, n, 3], got {tuple(paths.shape)}")
        if max_points is None:
            # upper bound on int(total / step): every chord is at most its float32 length + 1e-6
            seg = (paths[:, 1:, :2] - paths[:, :-1, :2]).double().norm(dim=-1).sum(-1).max()
            max_points = max(n, int(float(seg) * 1.001 / self._distance_step) + n + 8)
        nbytes = C.c_size_t()
        L.check(lib.nfopp_postprocess_workspace_bytes(P, n, C.byref(nbytes)))
        ws = torch.empty(int(nbytes.value), dtype=torch.uint8, device=paths.device)
        out = torch.zeros(P, max_points, 3, dtype=torch.float64, device=paths.device)
        count = torch.empty(P, dtype=torch.int32, device=paths.device)
        with torch.cuda.device(paths.device):
            L.check(lib.nfopp_postprocess_paths(L.ptr(paths), P, n, float(self._minimal_distance), float(self._distance_step),
                                                int(max_points), L.ptr(ws), C.c_size_t(ws.numel()), L.ptr(out), L.ptr(count),
                                                L.stream_ptr()))
            torch.cuda.current_stream().synchronize()          # the workspace is freed on return
        short = int(count.min())
        if short < -1:
            raise ValueError(f"postprocess: a path needs {-short} poses, max_points={max_points}")
        return out, count

    def process(self, trajectory):
        """ros/path_postprocessor.py:13-26 for one trajectory (Position2 of arrays)."""
        if len(trajectory) < 3:
            return trajectory
        vec = np.asarray(trajectory.as_vec(), dtype=np.float32)
        out, count = self.process_batch(vec[None])
        k = int(count[0])
        if k < 0:
            raise ValueError("postprocess: fewer than 3 distinct path nodes")
        return Position2.from_vec(out[0, :k].cpu().numpy())

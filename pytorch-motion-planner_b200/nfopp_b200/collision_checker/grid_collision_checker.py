"""Point robot on an occupancy grid: the reference's only in-repo grid checker is the notebook class
MapCollisionChecker (notebooks/onf_planner_image_map.ipynb:60-80): half-cell offset, truncation to int32,
last row/column rejected, outside the map counts as collision, and no boundary OR."""
from dataclasses import dataclass

import numpy as np
import torch

from .. import _lib as L
from .collision_checker import CollisionChecker


@dataclass
class MapImage:
    cols: int
    rows: int
    origin_x: float
    origin_y: float
    cell_size: float
    map_image: np.ndarray


class GridCollisionChecker(CollisionChecker):
    kind = "grid"

    def __init__(self, map_image: MapImage, boundaries=None):
        super().__init__(boundaries)
        self._map_image = map_image

    def _label(self, poses):
        mi = self._map_image
        grid = torch.as_tensor((np.asarray(mi.map_image) > 0).astype(np.uint8), device="cuda").contiguous()
        rows, cols = grid.shape
        return self._run(poses, lambda lib, p, out, n, d, s: lib.nfopp_label_grid(
            p, 1, n, d, L.ptr(grid), 1, rows, cols, float(mi.origin_x), float(mi.origin_y), float(mi.cell_size),
            None, out, s))

    def engine_spec(self):
        mi = self._map_image
        rows, cols = np.asarray(mi.map_image).shape
        return dict(checker="grid", checker_bounds=None, grid_shape=(int(rows), int(cols)),
                    grid_origin=(float(mi.origin_x), float(mi.origin_y)), grid_cell=float(mi.cell_size))


MapCollisionChecker = GridCollisionChecker

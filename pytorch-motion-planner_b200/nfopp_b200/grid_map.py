"""Occupancy grid -> obstacle point cloud for the disc / rectangle checkers.

Mirror of the reference's ros/grid_map.py:6-29 (GridMap) without the ROS message type: the adapter there
(ros/planner_adapter.py) feeds `as_point_cloud()` to `collision_checker.update_obstacle_points` and `boundaries` to
`planner.set_boundaries`.  A 1000x1000 map with 10 % occupancy yields ~1e5 points; the engine stages up to
2048 points per problem in shared memory and reads larger sets through L1 / L2.
"""
import numpy as np


class GridMap(object):
    def __init__(self, data, resolution, origin, threshold=0.5):
        self._map = np.asarray(data)
        self._resolution = resolution
        self._origin = origin                       # Position2 of cell (0, 0)'s corner
        self._threshold = threshold
        self._points = None

    def as_point_cloud(self):
        """Centres of the cells above the threshold, (x, y) = (column, row) order, in the world frame (:14-20)."""
        if self._points is None:
            rows, cols = np.nonzero(self._map > self._threshold)
            cells = np.stack([cols, rows], axis=1)
            local = cells * self._resolution + np.ones(2) * self._resolution / 2.
            self._points = self._origin.apply(local)
        return self._points

    @property
    def boundaries(self):
        """(left, right, bottom, top) of the map in the world frame (:22-29; the origin's rotation is ignored there too)."""
        height, width = self._map.shape
        left, bottom = self._origin.x, self._origin.y
        return left, left + width * self._resolution, bottom, bottom + height * self._resolution

    @classmethod
    def from_occupancy_data(cls, data, width, height, resolution, origin):
        """nav_msgs/OccupancyGrid payload (:31-40): row-major int8 cells in 0..100, -1 = unknown -> free; scaled to 0..1."""
        cells = np.array(data).reshape(height, width)
        cells = np.where(cells == -1, 0, cells)
        return cls(cells.astype(np.float32) / 100, resolution, origin)

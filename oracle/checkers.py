"""Oracle: ground-truth collision labelling, numpy float64, in the reference's operation order.

Poses are [N, 2] (x, y) or [N, 3] (x, y, theta) float64 arrays; every function returns bool [N].
"""
import numpy as np


def boundary_labels(xy, bounds):
    """collision_checker/collision_checker.py:12-19 - strict comparisons; bounds=(x0,x1,y0,y1) or None."""
    out = np.zeros(len(xy), dtype=bool)
    if bounds is None:
        return out
    x0, x1, y0, y1 = bounds
    out |= xy[:, 0] > x1
    out |= xy[:, 0] < x0
    out |= xy[:, 1] > y1
    out |= xy[:, 1] < y0
    return out


def circle_labels(poses, obstacles, radius, bounds):
    """collision_checker/circle_collision_checker.py:11-14 (and circle_directed...:5-6, which feeds it
    Position2.translation): any_m sqrt(dx*dx + dy*dy) < r, OR boundary."""
    xy = np.asarray(poses, dtype=np.float64)[:, :2]
    hit = np.zeros(len(xy), dtype=bool)
    for ox, oy in np.asarray(obstacles, dtype=np.float64).reshape(-1, 2):
        dx = xy[:, 0] - ox
        dy = xy[:, 1] - oy
        hit |= np.sqrt(dx * dx + dy * dy) < radius
    return hit | boundary_labels(xy, bounds)


def rectangle_labels(poses, obstacles, box, bounds):
    """collision_checker/rectangle_collision_checker.py:11-26 with utils/position2.py:91-94 (inv):
    obstacle points are moved into the robot frame; inside iff box0 < x1 < box1 and box2 < y1 < box3."""
    pose = np.asarray(poses, dtype=np.float64)
    x, y, th = pose[:, 0], pose[:, 1], pose[:, 2]
    c, s = np.cos(th), np.sin(th)
    ix = -x * c - y * s                      # position2.py:92
    iy = x * s - y * c                       # position2.py:93
    ith = -th                                # position2.py:94
    ci, si = np.cos(ith), np.sin(ith)
    hit = np.zeros(len(pose), dtype=bool)
    for ox, oy in np.asarray(obstacles, dtype=np.float64).reshape(-1, 2):
        x1 = ox * ci - oy * si + ix          # rectangle_collision_checker.py:22-23
        y1 = ox * si + oy * ci + iy          # rectangle_collision_checker.py:24-25
        hit |= (x1 > box[0]) & (x1 < box[1]) & (y1 > box[2]) & (y1 < box[3])
    return hit | boundary_labels(pose[:, :2], bounds)


def grid_labels(poses, grid, origin_x, origin_y, cell, bounds=None):
    """notebooks/onf_planner_image_map.ipynb:65-80 (MapCollisionChecker): half-cell offset, C truncation
    to int32, last row/column rejected, outside the map => collision.  Note the notebook class does NOT
    OR in the boundary check; `bounds` is kept (default None) for the batched corridor configs."""
    xy = np.asarray(poses, dtype=np.float64)[:, :2]
    rows, cols = grid.shape
    fx = (xy[:, 0] - origin_x - cell / 2) / cell
    fy = (xy[:, 1] - origin_y - cell / 2) / cell
    with np.errstate(invalid="ignore"):
        ix = np.trunc(np.clip(fx, -2.0e9, 2.0e9)).astype(np.int64)
        iy = np.trunc(np.clip(fy, -2.0e9, 2.0e9)).astype(np.int64)
    valid = (ix >= 0) & (iy >= 0) & (iy < rows - 1) & (ix < cols - 1)
    out = np.ones(len(xy), dtype=bool)
    out[valid] = grid[iy[valid], ix[valid]] > 0
    return out | boundary_labels(xy, bounds)

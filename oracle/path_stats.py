"""Oracle (test infrastructure, never the product path): per-plan path statistics, the numbers
BenchmarkAdapter::evaluate (benchmark/bench_mr_binding/BenchmarkAdapter.cpp:68-103) files for a path.

`cusps` restates BenchmarkAdapter::computeCusps (BenchmarkAdapter.cpp:105-135) line by line.  The other metrics
(PathLengthMetric, MaxCurvatureMetric, NormalizedCurvatureMetric, AOLMetric, ClearingMetric of bench-mr @3d641be and
ompl::geometric::PathGeometric::smoothness) live in submodules that are missing from /root/reference, so they are
restated from their published definitions - parity unpinned for those: Menger curvature of consecutive node triples,
its arc-length-weighted sum, summed |heading change| over length, OMPL's smoothness sum, distance to the nearest
obstacle of the reference's collision checkers.  Nodes = the T+2 poses of get_path(), x and y, float64."""
import numpy as np


def normalize_angle(a):
    a = np.fmod(a + np.pi, 2 * np.pi)
    if a < 0:
        a += 2 * np.pi
    return a - np.pi


def cusps_and_turn(xy, threshold):
    """BenchmarkAdapter.cpp:105-135 -> (cusp count, summed |yaw change| of the de-duplicated polyline)."""
    n = len(xy)
    dist = lambda i, j: float(np.hypot(*(xy[i] - xy[j])))       # noqa: E731
    prev = cur = nxt = 0
    cusps, turn = 0, 0.0
    while nxt < n:
        if dist(prev, cur) <= 0:
            cur += 1
            nxt += 1
        elif dist(cur, nxt) <= 0:
            nxt += 1
        else:
            yaw_prev = np.arctan2(xy[cur, 1] - xy[prev, 1], xy[cur, 0] - xy[prev, 0])
            yaw_next = np.arctan2(xy[nxt, 1] - xy[cur, 1], xy[nxt, 0] - xy[cur, 0])
            change = abs(normalize_angle(yaw_next - yaw_prev))
            if change > threshold:
                cusps += 1
            turn += change
            prev, cur = cur, nxt
            nxt += 1
    return cusps, turn


def clearing(pose, kind, obstacles=None, radius=0.0, box=None, grid=None, origin=(0.0, 0.0), cell=1.0, bounds=None,
             window=64):
    x, y, th = pose
    best = 1e30
    if kind in ("circle", "rectangle"):
        d = np.asarray(obstacles, dtype=np.float64) - np.array([x, y])
        if kind == "circle":
            best = float((np.hypot(d[:, 0], d[:, 1]) - radius).min())
        else:
            c, s = np.cos(th), np.sin(th)
            rx, ry = c * d[:, 0] + s * d[:, 1], -s * d[:, 0] + c * d[:, 1]
            ex = np.maximum(np.maximum(box[0] - rx, rx - box[1]), 0.0)
            ey = np.maximum(np.maximum(box[2] - ry, ry - box[3]), 0.0)
            best = float(np.hypot(ex, ey).min())
    elif kind == "grid":
        rows, cols = grid.shape
        cx, cy = int(np.floor((x - origin[0]) / cell)), int(np.floor((y - origin[1]) / cell))
        y0, y1 = max(cy - window, 0), min(cy + window, rows - 1)
        x0, x1 = max(cx - window, 0), min(cx + window, cols - 1)
        if y0 <= y1 and x0 <= x1:
            iy, ix = np.nonzero(grid[y0:y1 + 1, x0:x1 + 1] > 0)
            if len(iy):
                gx, gy = origin[0] + (ix + x0) * cell, origin[1] + (iy + y0) * cell
                ex = np.maximum(np.maximum(gx - x, x - (gx + cell)), 0.0)
                ey = np.maximum(np.maximum(gy - y, y - (gy + cell)), 0.0)
                best = float(np.hypot(ex, ey).min())
        best = min(best, window * cell)
    if bounds is not None:
        best = min(best, x - bounds[0], bounds[1] - x, y - bounds[2], bounds[3] - y)
    return max(best, 0.0)


def path_statistics(path, cusp_threshold=np.pi / 3, **env):
    """path [T+2, 2|3] -> dict with bench-mr's PathStatistics names."""
    p = np.asarray(path, dtype=np.float64)
    xy = p[:, :2]
    seg = np.hypot(*(xy[1:] - xy[:-1]).T)
    length = float(seg.sum())
    kmax = knorm = smooth = 0.0
    for j in range(1, len(xy) - 1):
        av, bv = xy[j] - xy[j - 1], xy[j + 1] - xy[j]
        a, b, c = np.hypot(*av), np.hypot(*bv), np.hypot(*(av + bv))
        if a > 1e-12 and b > 1e-12 and c > 1e-12:
            k = 2.0 * abs(av[0] * bv[1] - av[1] * bv[0]) / (a * b * c)
            kmax = max(kmax, k)
            knorm += k * 0.5 * (a + b)
        if a > 0 and b > 0:
            acv = (a * a + b * b - c * c) / (2 * a * b)
            if -1.0 < acv < 1.0:
                kk = 2.0 * (np.pi - np.arccos(acv)) / (a + b)
                smooth += kk * kk
    cusps, turn = cusps_and_turn(xy, cusp_threshold)
    out = dict(path_length=length, max_curvature=kmax, normalized_curvature=knorm, aol=turn / length if length > 0 else 0.0,
               smoothness=smooth, cusps=float(cusps))
    if env:
        cl = [clearing((q[0], q[1], q[2] if p.shape[1] == 3 else 0.0), **env) for q in p]
        out["min_clearing_distance"], out["mean_clearing_distance"] = float(min(cl)), float(np.mean(cl))
    return out

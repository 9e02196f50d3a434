"""Oracle (test infrastructure, never the product path): the A* trajectory initialiser.

Restates, for one problem,
  * AstarTrajectoryInitializer.calculate_astar_path (astar/astar_trajectory_initializer.py:27-46): the checker-labelled
    cell grid (cell centres at heading 3 pi / 4, goal cell forced free),
  * JPS.search with jps=False (astar/jps.py:40-130, 282-322): plain A* on the 8-connected grid, diagonal cost sqrt(2),
    Euclidean heuristic, start key 0, heap ordered by (key, (row, col)), strict-improvement relaxation, closed set;
    when the goal is unreachable the path to the LAST expanded state is returned (jps.py:321-322),
  * reparametrize_path (utils/math.py:57-65): float32 chord lengths + 1e-6, normalised cumulative parameter and scipy's
    interp1d(kind="quadratic") = make_interp_spline(k=2) with not-a-knot knots, written out (collocation matrix, banded
    solve, de Boor evaluation) so that the CUDA kernel can be checked stage by stage.
Pinned against the live reference and scipy by oracle/make_golden.py (tests/golden/astar.npz)."""
import heapq
import math

import numpy as np

NEIGHBOURS = [(-1, 0, False), (0, -1, False), (1, 0, False), (0, 1, False),
              (-1, -1, True), (-1, 1, True), (1, -1, True), (1, 1, True)]


def grid_geometry(boundaries, resolution):
    """astar_trajectory_initializer.py:29-30 -> (x_cells, y_cells)."""
    b = boundaries
    return int((b[1] - b[0]) // resolution) + 1, int((b[3] - b[2]) // resolution) + 1


def point_cells(points_xy, boundaries, resolution):
    """astar_trajectory_initializer.py:31-34: float32 point, weak python scalars -> float32 floor division."""
    p = np.asarray(points_xy, dtype=np.float32)
    cx = ((p[..., 0] - np.float32(boundaries[0])) // np.float32(resolution)).astype(np.int64)
    cy = ((p[..., 1] - np.float32(boundaries[2])) // np.float32(resolution)).astype(np.int64)
    return cx, cy


def cell_centres(x_cells, y_cells, boundaries, resolution):
    """astar_trajectory_initializer.py:35-37 -> float64 [y_cells * x_cells, 3] poses (x, y, 3 pi / 4), row-major in y."""
    x, y = np.meshgrid(range(x_cells), range(y_cells))
    x = x.reshape(-1) * resolution + resolution / 2 + boundaries[0]
    y = y.reshape(-1) * resolution + resolution / 2 + boundaries[2]
    return np.stack([x, y, np.ones_like(x) * 3 * np.pi / 4], 1)


def astar_cells(matrix, start, goal, diagonal_cost=math.sqrt(2)):
    """JPS(matrix, jps=False).find_path(start, goal) -> int array [n, 2] of (row, col)."""
    rows, cols = matrix.shape
    cost = np.full((rows, cols), np.inf)
    key = np.full((rows, cols), np.inf)
    parent = {}
    closed = set()
    cost[start] = 0.0
    key[start] = 0.0
    heap = [(0.0, tuple(start))]
    last = tuple(start)
    found = None
    while heap:
        _, loc = heapq.heappop(heap)
        if loc in closed:
            continue
        last = loc
        if loc == tuple(goal):
            found = loc
            break
        closed.add(loc)
        for dr, dc, diag in NEIGHBOURS:
            r, c = loc[0] + dr, loc[1] + dc
            if 0 <= r < rows and 0 <= c < cols and matrix[r, c] == 0 and (r, c) not in closed:
                new_cost = cost[loc] + (diagonal_cost if diag else 1)
                if cost[r, c] > new_cost:
                    cost[r, c] = new_cost
                    key[r, c] = math.sqrt((r - goal[0]) ** 2 + (c - goal[1]) ** 2) + new_cost
                    parent[(r, c)] = loc
                heapq.heappush(heap, (key[r, c], (r, c)))
    node = found if found is not None else last
    path = [node]
    while node in parent:
        node = parent[node]
        path.append(node)
    return np.array(path[::-1], dtype=np.int64)


def cells_to_points(cells, boundaries, resolution):
    """astar_trajectory_initializer.py:43-46 -> float32 [n, 2] (x, y)."""
    out = np.zeros(cells.shape, dtype=np.float32)
    out[:, 0] = cells[:, 1] * resolution + resolution / 2 + boundaries[0]
    out[:, 1] = cells[:, 0] * resolution + resolution / 2 + boundaries[2]
    return out


# ---- scipy.interpolate.interp1d(kind="quadratic") written out ---------------------------------------------------
def _basis3(t, x, j):
    """The three quadratic B-spline basis functions that are non-zero on [t[j], t[j+1]) at x (de Boor-Cox), i.e.
    B_{j-2}, B_{j-1}, B_j of degree 2."""
    left1, right1 = x - t[j], t[j + 1] - x
    b0 = right1 / (right1 + left1)
    b1 = left1 / (right1 + left1)
    left2, right2 = x - t[j - 1], t[j + 2] - x
    n0 = b0 * right1 / (right1 + left2)
    tmp = b0 * left2 / (right1 + left2)
    n1 = tmp + b1 * right2 / (right2 + left1)
    n2 = b1 * left1 / (right2 + left1)
    return n0, n1, n2


def _interval(t, n, x):
    """largest j in [2, n-1] with t[j] <= x (clamped): the knot interval used for evaluation (extrapolating outside)."""
    j = 2
    while j < n - 1 and t[j + 1] <= x:
        j += 1
    return j


def quadratic_interp(xp, yp, xq):
    """interp1d(xp, yp, kind='quadratic', axis=0, fill_value='extrapolate')(xq); xp strictly increasing, len >= 3."""
    xp = np.asarray(xp, dtype=np.float64)
    yp = np.asarray(yp, dtype=np.float64)
    n = len(xp)
    mids = (xp[1:] + xp[:-1]) / 2
    t = np.concatenate([[xp[0]] * 3, mids[1:-1], [xp[-1]] * 3])
    A = np.zeros((n, n))
    for i in range(n):
        j = _interval(t, n, xp[i])
        A[i, j - 2:j + 1] = _basis3(t, xp[i], j)
    c = np.linalg.solve(A, yp)
    out = np.zeros((len(xq),) + yp.shape[1:])
    for q, x in enumerate(xq):
        j = _interval(t, n, x)
        b = _basis3(t, x, j)
        out[q] = b[0] * c[j - 2] + b[1] * c[j - 1] + b[2] * c[j]
    return out


def reparametrize_path(path, point_count):
    """utils/math.py:57-65 for a float32 path [n, 2]."""
    path = np.asarray(path, dtype=np.float32)
    distances = np.linalg.norm(path[1:] - path[:-1], axis=1) + 1e-6       # float32
    cum = np.concatenate([np.zeros(1), np.cumsum(distances)], axis=0)     # float32 running sum, then float64
    param = cum / cum[-1]
    return quadratic_interp(param, path, np.linspace(0, 1, point_count))


def initial_trajectory_xy(start, goal, boundaries, resolution, labeller, trajectory_length):
    """AstarTrajectoryInitializer.initialize_trajectory (x, y part; :15-24) -> (float32 [T, 2], cells, matrix)."""
    start = np.asarray(start, dtype=np.float32)
    goal = np.asarray(goal, dtype=np.float32)
    x_cells, y_cells = grid_geometry(boundaries, resolution)
    sx, sy = point_cells(start[:2], boundaries, resolution)
    gx, gy = point_cells(goal[:2], boundaries, resolution)
    matrix = np.asarray(labeller(cell_centres(x_cells, y_cells, boundaries, resolution)), dtype=bool).reshape(y_cells, x_cells)
    matrix[gy, gx] = False
    cells = astar_cells(matrix, (int(sy), int(sx)), (int(gy), int(gx)))
    pts = cells_to_points(cells, boundaries, resolution)
    full = np.concatenate([start[None, :2], pts, goal[None, :2]], axis=0)
    rep = reparametrize_path(full, trajectory_length + 2)
    return rep[1:-1].astype(np.float32), cells, matrix

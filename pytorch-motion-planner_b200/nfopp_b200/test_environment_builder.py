"""The reference's fixed toy worlds (test_environment_builder.py:8-54) plus seeded batch generators for the
BASELINE configs (jittered corridor walls, random-obstacle grids)."""
import numpy as np

from .test_environment import TestEnvironment

_BOUNDS = (-0.1, 3.1, -0.1, 3.1)


def _segment(p, q, count):
    """`count` points from p to q inclusive."""
    s = np.linspace(0.0, 1.0, count)[:, None]
    return np.asarray(p, dtype=np.float64)[None] * (1 - s) + np.asarray(q, dtype=np.float64)[None] * s


class TestEnvironmentBuilder(object):
    __test__ = False

    @staticmethod
    def _obstacle_points():
        # two vertical walls of 10 points: x=1.15, y in [0,2] and x=1.85, y in [1,3]  (:16-22)
        y = np.linspace(0, 2, 10)
        left = np.column_stack([np.full(10, 1.15), y])
        right = np.column_stack([np.full(10, 1.85), y + 1])
        return np.vstack([left, right])

    @staticmethod
    def _point_line(start, end, point_count):
        return np.column_stack([np.linspace(start[0], end[0], point_count), np.linspace(start[1], end[1], point_count)])

    @staticmethod
    def _car_obstacle_points():
        # parking nook: 4 walls of 10 points (:38-54)
        top, left = 2.3, 1.6
        walls = [((0, top), (left, top)), ((left, top), (left, 3)), ((2.5, top), (2.5, 3)), ((2.5, top), (3, top))]
        return np.vstack([TestEnvironmentBuilder._point_line(a, b, 10) for a, b in walls])

    @staticmethod
    def make_test_environment():
        return TestEnvironment(np.array([0.5, 0.5], dtype=np.float32), np.array([2.5, 2.5], dtype=np.float32), _BOUNDS,
                               TestEnvironmentBuilder._obstacle_points())

    @staticmethod
    def make_test_environment_with_angles():
        return TestEnvironment(np.array([0.5, 0.5, 0], dtype=np.float32), np.array([2.5, 1.5, 0], dtype=np.float32),
                               _BOUNDS, TestEnvironmentBuilder._obstacle_points())

    def make_car_environment(self):
        return TestEnvironment(np.array([0.5, 1.5, 0], dtype=np.float32), np.array([2, 2.7, 0], dtype=np.float32),
                               _BOUNDS, TestEnvironmentBuilder._car_obstacle_points())

    # ---- batch generators (not in the reference: it has no batching) ----------
    @staticmethod
    def make_corridor_batch(problems, seed=0, points_per_wall=10):
        """Jittered S-bend corridors around the reference's two-wall world (:16-22): left wall x ~ U(0.95,1.25),
        channel width ~ U(0.7,0.9) (the reference's is 0.7; a disc of radius 0.3 needs > 0.6), wall ends shifted
        by U(-0.2,0.2), goal height by U(-0.3,0.3); problem i uses RandomState(seed + i).
        Returns (starts [P,3], goals [P,3], bounds, obstacles [P, 2*ppw, 2])."""
        starts = np.tile(np.array([0.5, 0.5, 0], dtype=np.float32), (problems, 1))
        goals = np.tile(np.array([2.5, 1.5, 0], dtype=np.float32), (problems, 1))
        obstacles = np.empty((problems, 2 * points_per_wall, 2))
        for i in range(problems):
            rs = np.random.RandomState(seed + i)
            x_left = rs.uniform(0.95, 1.25)
            x_right = x_left + rs.uniform(0.7, 0.9)
            off_l, off_r = rs.uniform(-0.2, 0.2, 2)
            obstacles[i, :points_per_wall] = _segment((x_left, 0 + off_l), (x_left, 2 + off_l), points_per_wall)
            obstacles[i, points_per_wall:] = _segment((x_right, 1 + off_r), (x_right, 3 + off_r), points_per_wall)
            goals[i, 1] = np.float32(1.5 + rs.uniform(-0.3, 0.3))
        return starts, goals, _BOUNDS, obstacles

    @staticmethod
    def make_random_grid_batch(problems, rows=100, cols=100, obstacle_ratio=0.03, cell=0.03, seed=0):
        """Random-obstacle occupancy grids (forest-style, obstacle_ratio 0.03); problem i uses
        RandomState(seed + i).  A free disc is cleared around start and goal.
        Returns (starts [P,3], goals [P,3], bounds, grids [P,rows,cols] uint8, origin, cell)."""
        grids = np.zeros((problems, rows, cols), dtype=np.uint8)
        w, h = cols * cell, rows * cell
        starts = np.tile(np.array([0.15 * w, 0.15 * h, 0], dtype=np.float32), (problems, 1))
        goals = np.tile(np.array([0.85 * w, 0.85 * h, 0], dtype=np.float32), (problems, 1))
        yy, xx = np.mgrid[0:rows, 0:cols]
        cx, cy = (xx + 0.5) * cell, (yy + 0.5) * cell
        for i in range(problems):
            rs = np.random.RandomState(seed + i)
            g = rs.rand(rows, cols) < obstacle_ratio
            for pt in (starts[i], goals[i]):
                g &= (cx - pt[0]) ** 2 + (cy - pt[1]) ** 2 > (4 * cell) ** 2
            grids[i] = g
        return starts, goals, (0.0, w, 0.0, h), grids, (0.0, 0.0), cell

    @staticmethod
    def make_corridor_grid_batch(problems, rows=100, cols=100, radius=3, branches=100, cell=0.03, seed=0):
        """Corridor mazes in the style of bench-mr's random corridor grids (notebooks/corridor_experiment.ipynb:91,407:
        100x100 cells, radius 3, 100 branches, seeds 2..11).  bench-mr's generator lives in a missing submodule; this is our
        own with the same parameters: the map starts fully occupied, the first node is a random cell, and every branch
        carves an L-shaped corridor (first along x, then along y, `radius` cells to either side) from the nearest existing
        node to a new random cell.  Start = the first node, goal = the node farthest from it; headings point along the
        start -> goal chord.  Problem i uses RandomState(seed + i).
        Returns (starts [P,3], goals [P,3], bounds, grids [P,rows,cols] uint8, origin, cell)."""
        grids = np.ones((problems, rows, cols), dtype=np.uint8)
        starts = np.zeros((problems, 3), dtype=np.float32)
        goals = np.zeros((problems, 3), dtype=np.float32)
        lo, hi_r, hi_c = radius + 1, rows - radius - 1, cols - radius - 1
        for i in range(problems):
            rs = np.random.RandomState(seed + i)
            g = grids[i]
            nodes = [(rs.randint(lo, hi_r), rs.randint(lo, hi_c))]

            def carve(r, c):
                g[r - radius:r + radius + 1, c - radius:c + radius + 1] = 0
            carve(*nodes[0])
            for _ in range(branches):
                tr, tc = rs.randint(lo, hi_r), rs.randint(lo, hi_c)
                r, c = min(nodes, key=lambda n: (n[0] - tr) ** 2 + (n[1] - tc) ** 2)
                while c != tc:
                    c += 1 if tc > c else -1
                    carve(r, c)
                while r != tr:
                    r += 1 if tr > r else -1
                    carve(r, c)
                nodes.append((tr, tc))
            r0, c0 = nodes[0]
            r1, c1 = max(nodes, key=lambda n: (n[0] - r0) ** 2 + (n[1] - c0) ** 2)
            heading = np.arctan2(r1 - r0, c1 - c0)
            starts[i] = ((c0 + 0.5) * cell, (r0 + 0.5) * cell, heading)
            goals[i] = ((c1 + 0.5) * cell, (r1 + 0.5) * cell, heading)
        return starts, goals, (0.0, cols * cell, 0.0, rows * cell), grids, (0.0, 0.0), cell

    @staticmethod
    def make_blob_grid(rows=1000, cols=1000, obstacle_ratio=0.1, cell=0.003, blob_cells=24, seed=1):
        """One large synthetic occupancy grid (BASELINE config 4): square blobs of `blob_cells` cells dropped
        uniformly until ~obstacle_ratio of the map is occupied (bench-mr's random-grid generator is in a missing
        submodule; this is our own, documented in DESIGN.md).  Returns (grid uint8 [rows, cols], origin, cell)."""
        rs = np.random.RandomState(seed)
        grid = np.zeros((rows, cols), dtype=np.uint8)
        n_blobs = int(obstacle_ratio * rows * cols / (blob_cells * blob_cells))
        for _ in range(n_blobs):
            r, c = rs.randint(0, rows - blob_cells), rs.randint(0, cols - blob_cells)
            grid[r:r + blob_cells, c:c + blob_cells] = 1
        return grid, (0.0, 0.0), cell

    @staticmethod
    def free_endpoints(grid, cell, problems, seed=0, margin_cells=40):
        """Start / goal poses for `problems` problems on one shared grid: start in the lower-left quarter, goal in the
        upper-right quarter, both in free cells with a free neighbourhood of +-3 cells; problem i uses
        RandomState(seed + i)."""
        rows, cols = grid.shape
        occ = grid > 0
        starts = np.zeros((problems, 3), dtype=np.float32)
        goals = np.zeros((problems, 3), dtype=np.float32)

        def free(rs, r0, r1, c0, c1):
            while True:
                r, c = rs.randint(r0, r1), rs.randint(c0, c1)
                if not occ[max(r - 3, 0):r + 4, max(c - 3, 0):c + 4].any():
                    return (c + 0.5) * cell, (r + 0.5) * cell
        for i in range(problems):
            rs = np.random.RandomState(seed + i)
            m = margin_cells
            starts[i, :2] = free(rs, m, rows // 3, m, cols // 3)
            goals[i, :2] = free(rs, 2 * rows // 3, rows - m, 2 * cols // 3, cols - m)
        return starts, goals

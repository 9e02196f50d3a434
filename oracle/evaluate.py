"""Oracle (test infrastructure, never the product path): path evaluation for the stop criterion of
scripts/run_bench_mr.py:111-126 - `benchmark.evaluate_path(planner.get_path())` -> (collision, length), and the loop's
best-path / early-stop bookkeeping.

bench-mr's `PlannerUtils::interpolated` / `PathLengthMetric` (BenchmarkAdapter.cpp:201-208) live in a submodule that is
missing from /root/reference (parity unpinned for that interpolation): this restatement fixes OUR interpolation - `k`
equally spaced poses on each segment of the full path (x, y linear, heading along the wrapped difference) plus the
goal, float64 - and labels every pose with the reference's collision checker (oracle.checkers)."""
import numpy as np


def wrap(a):
    return (a + np.pi) % (2 * np.pi) - np.pi


def interpolate_path(path, k):
    """path [T+2, D] (float32 from get_path) -> [(T+1)*k + 1, D] float64."""
    p = np.asarray(path, dtype=np.float64)
    a, b = p[:-1], p[1:]
    s = (np.arange(k, dtype=np.float64) / float(k))[None, :, None]
    d = b - a
    if p.shape[1] == 3:
        d[:, 2] = wrap(d[:, 2])
    out = (a[:, None, :] + s * d[:, None, :]).reshape(-1, p.shape[1])
    return np.concatenate([out, p[-1:]], 0)


def evaluate_path(path, labeller, k=4):
    """-> (collides: bool, length: float, colliding pose count)."""
    poses = interpolate_path(path, k)
    hits = np.asarray(labeller(poses), dtype=bool)
    p = np.asarray(path, dtype=np.float64)
    length = float(np.sqrt(((p[1:, :2] - p[:-1, :2]) ** 2).sum(1)).sum())
    return bool(hits.any()), length, int(hits.sum())


def solve_loop(step, get_path, evaluate, max_iterations=1000, min_iterations=200, check_every=50):
    """scripts/run_bench_mr.py:109-131 for one planner: step until a collision-free path stops getting shorter.
    -> (path, collides, length, iterations)."""
    best_length, best_path, it = np.inf, None, 0
    for i in range(max_iterations):
        step()
        it = i + 1
        if i > min_iterations and i % check_every == 0:
            collision, length = evaluate(get_path())[:2]
            if not collision and length < best_length:
                best_length, best_path = length, get_path()
            elif not collision:
                break
    path = get_path()
    collision, length = evaluate(path)[:2]
    if collision and best_path is not None:
        path, collision, length = best_path, False, best_length
    return path, collision, length, it

"""Generates tests/golden/postprocess.npz from the UNMODIFIED reference (container only; /root/reference must exist):
PathPostprocessor.process (ros/path_postprocessor.py) on planner-like float32 paths - smooth S-curves, paths with
repeated / nearly repeated points (the distance filter), a reversing start (the direction filter), headings that cross
+-pi (unfolding), a short path - and checks oracle/postprocess.py against every recorded case while doing so.

    python -m oracle.make_golden_postprocess
"""
import contextlib
import os

import numpy as np

from . import postprocess as PP
from .reference_bridge import load_reference


def make_path(kind, n, seed):
    rng = np.random.RandomState(seed)
    s = np.linspace(0.0, 1.0, n)
    if kind == "s_curve":
        x, y = 0.3 + 2.4 * s, 1.5 + 0.9 * np.sin(2.5 * np.pi * s) * s
    elif kind == "loop":                                     # heading sweeps through +-pi
        a = 2.2 * np.pi * s + 0.4
        x, y = 2.0 + 1.5 * np.cos(a) * (1 + 0.3 * s), 2.0 + 1.5 * np.sin(a)
    elif kind == "reverse_start":                            # backs up for a few nodes, then drives forward
        u = np.where(s < 0.04, -s, s - 0.08)
        x, y = 1.0 + 3.0 * u, 0.5 + 0.8 * u * u
    else:                                                    # "jitter": noisy planner-like path
        x, y = 0.2 + 2.6 * s + 0.01 * rng.randn(n), 0.4 + 2.0 * s ** 2 + 0.01 * rng.randn(n)
    dx, dy = np.gradient(x), np.gradient(y)
    th = np.arctan2(dy, dx)
    if kind == "reverse_start":
        th = np.arctan2(np.abs(dy), np.abs(dx))             # the car keeps facing forward while it backs up
    th = th + 0.02 * rng.randn(n)
    path = np.stack([x, y, (th + np.pi) % (2 * np.pi) - np.pi], 1).astype(np.float32)
    return path


CASES = [
    # name, kind, points, seed, duplicate some points, minimal_distance, distance_step
    ("s_curve", "s_curve", 102, 1, False, 0.001, 0.05),
    ("loop", "loop", 102, 2, False, 0.001, 0.05),
    ("reverse_start", "reverse_start", 102, 3, False, 0.001, 0.05),
    ("jitter_dups", "jitter", 102, 4, True, 0.001, 0.05),
    ("coarse_filter", "jitter", 102, 5, True, 0.03, 0.11),
    ("short", "s_curve", 7, 6, False, 0.001, 0.2),
    ("dense_output", "loop", 22, 7, False, 0.001, 0.013),
]


def main():
    load_reference()
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        from neural_field_optimal_planner.ros.path_postprocessor import PathPostprocessor
        from neural_field_optimal_planner.utils.position2 import Position2
    out = {}
    for name, kind, n, seed, dups, min_d, step in CASES:
        path = make_path(kind, n, seed)
        if dups:
            rng = np.random.RandomState(100 + seed)
            for i in rng.choice(np.arange(2, n - 2), 9, replace=False):
                path[i] = path[i - 1] + np.float32(1e-4) * rng.randn(3).astype(np.float32)   # nearly repeated node
            path[n // 2] = path[n // 2 - 1]                                                  # exactly repeated node
        res = PathPostprocessor(minimal_distance=min_d, distance_step=step).process(Position2.from_vec(path.copy()))
        want = res.as_vec()
        got = PP.process(path, min_d, step)
        assert got.shape == want.shape, (name, got.shape, want.shape)
        err = np.abs(got - want).max()
        assert err <= 1e-9, (name, err)
        print(f"{name}: {n} nodes -> {len(want)} poses, oracle vs reference max abs err {err:.2e}")
        out[f"{name}/path"] = path
        out[f"{name}/result"] = want
        out[f"{name}/params"] = np.array([min_d, step])
    out["names"] = np.array([c[0] for c in CASES])
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "postprocess.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()

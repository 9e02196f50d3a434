"""Oracle: near-path / uniform pose sampling and replay-buffer resampling (numpy float64 + torch float32).

All random variates are passed in explicitly (variate-injection): the callers draw them from numpy's legacy
global stream in the reference's order, so the CUDA path can be fed the same numbers.
"""
import numpy as np
import torch


def intermediate_positions(traj: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    """nerf_opt_planner.py:113-117: traj[1:]*(1-t) + traj[:-1]*t in float32 (theta lerped without wrap)."""
    t = t.to(torch.float32).reshape(-1, 1)
    return traj[1:] * (1 - t) + traj[:-1] * t


def offset_positions(pos: np.ndarray, z_xy: np.ndarray, z_th, offset: float, angle_offset: float):
    """nerf_opt_planner.py:119-120 (D=2): float32 base + float64 noise -> a NEW float64 array.
    constrained_nerf_opt_planner.py:57-61 (D=3): the float32 base array is copied and assigned IN PLACE,
    so the float64 sums are rounded back to float32 (numpy casts on store); the values later reach the
    checker promoted to float64 but float32-representable."""
    base = np.asarray(pos)
    if base.shape[1] == 2:
        return base.astype(np.float64) + z_xy * offset
    out = base.astype(np.float32).copy()
    out[:, :2] = out[:, :2].astype(np.float64) + z_xy * offset
    out[:, 2] = out[:, 2].astype(np.float64) + z_th * angle_offset
    return out.astype(np.float64)


def field_points(u_xy: np.ndarray, u_th, border):
    """nerf_opt_planner.py:135-141; constrained_nerf_opt_planner.py:173-176 appends rand*2*pi."""
    pts = np.array(u_xy, dtype=np.float64, copy=True)
    pts[:, 0] = border[0] + pts[:, 0] * (border[1] - border[0])
    pts[:, 1] = border[2] + pts[:, 1] * (border[3] - border[2])
    if u_th is not None:
        pts = np.concatenate([pts, np.asarray(u_th, dtype=np.float64).reshape(-1, 1) * 2 * np.pi], axis=1)
    return pts


def resample_weights(sig: np.ndarray, ages: np.ndarray) -> np.ndarray:
    """nerf_opt_planner.py:125-127: sigmoid(float32) * exp(-0.03 age) + 1e-6, normalised (float64)."""
    w = sig.astype(np.float32) * np.exp(-ages * 0.03) + 1e-6
    return w / np.sum(w)


def legacy_choice(p: np.ndarray, size: int, replace: bool, draw):
    """numpy.random.RandomState.choice(n, size, replace, p) restated (numpy/random/mtrand.pyx, legacy
    stream frozen by NEP 19); call site nerf_opt_planner.py:131.  `draw(k)` returns k float64 uniforms
    from the same stream position the legacy generator would use (RandomState.rand / random_sample).

    replace=True : cdf = cumsum(p)/cdf[-1]; searchsorted(u, 'right').
    replace=False: repeat { u = rand(size-n_uniq); p[found]=0; cdf=cumsum(p)/cdf[-1];
                   new=searchsorted(u,'right'); keep first occurrences in draw order } until size found."""
    p = np.array(p, dtype=np.float64, copy=True)
    if replace:
        cdf = np.cumsum(p)
        cdf /= cdf[-1]
        return cdf.searchsorted(draw(size), side="right").astype(np.int64)
    found = np.zeros(size, dtype=np.int64)
    n_uniq = 0
    while n_uniq < size:
        u = draw(size - n_uniq)
        if n_uniq > 0:
            p[found[:n_uniq]] = 0
        cdf = np.cumsum(p)
        cdf /= cdf[-1]
        new = cdf.searchsorted(u, side="right")
        seen, keep = set(), []
        for v in new.tolist():
            if v not in seen:
                seen.add(v)
                keep.append(v)
        found[n_uniq:n_uniq + len(keep)] = keep
        n_uniq += len(keep)
    return found


def resample_buffer(cand_pos, cand_age, sig, count, draw):
    """nerf_opt_planner.py:122-133.  cand = replay buffer followed by this step's fine samples.
    Returns (positions, ages, indices or None)."""
    if len(cand_pos) < count:                                   # :128-129 (ages not incremented)
        return cand_pos, cand_age, None
    w = resample_weights(sig, cand_age)
    replace = np.count_nonzero(w > 1e-6) < count                # :130
    idx = legacy_choice(w, count, replace, draw)                # :131
    ages = cand_age + 1                                          # :132
    return cand_pos[idx], ages[idx], idx


class UniformReader:
    """Sequential reader over a recorded array of uniforms (what `draw` is in lock-step tests)."""

    def __init__(self, values):
        self.values = np.asarray(values, dtype=np.float64)
        self.used = 0

    def __call__(self, k):
        out = self.values[self.used:self.used + k]
        assert len(out) == k, "recorded uniform stream exhausted"
        self.used += k
        return out

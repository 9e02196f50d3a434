"""Oracle (test infrastructure, never the product path): PathPostprocessor.process for one path.

Restates ros/path_postprocessor.py:13-71 with numpy's dtypes written out (the path a planner returns is float32):
  * _filter_trajectory (:40-49): walking back from the goal, interior points closer than `minimal_distance` (float32 norm
    of the x, y difference) to the last kept point are dropped; first and last points always stay,
  * _calculate_parametrization / _calculate_total_distance (:28-38): float32 chord lengths + 1e-6; running sum in float32,
    normalised in float64; total = numpy's pairwise float32 sum; point_count = int(total / distance_step),
  * _reparametrize_trajectory (:51-56): headings unfolded (utils/math.py:38-43, float32 wrap / differences / running sum,
    float64 final add, stored back as float32), then scipy interp1d(kind="quadratic") over all three columns
    (oracle/astar.py: quadratic_interp, pinned against scipy) at linspace(0, 1, point_count),
  * _find_minimal_filter_index (:58-71): drop the leading points up to a direction flip within the first 6 segments (at
    least the first point is always dropped).
Pinned against the live reference by oracle/make_golden_postprocess.py (tests/golden/postprocess.npz)."""
import numpy as np

from .astar import quadratic_interp


def wrap_angles(a):
    """utils/math.py:34-35 (dtype of `a` is kept: python scalars are weak)."""
    return (a + np.pi) % (2 * np.pi) - np.pi


def unfold_angles(angles):
    """utils/math.py:38-43."""
    angles = wrap_angles(angles)
    delta = angles[1:] - angles[:-1]
    delta = np.where(delta > np.pi, delta - 2 * np.pi, delta)
    delta = np.where(delta < -np.pi, delta + 2 * np.pi, delta)
    return angles[0] + np.concatenate([np.zeros(1), np.cumsum(delta)], axis=0)


def filter_trajectory(traj, minimal_distance):
    prev = traj[-1]
    kept = [traj[-1]]
    for x in traj[-2:0:-1]:
        if np.linalg.norm(prev[:2] - x[:2]) > minimal_distance:
            kept.append(x)
            prev = x
    kept.append(traj[0])
    return np.array(kept[::-1])


def chord_lengths(traj):
    return np.linalg.norm(traj[1:, :2] - traj[:-1, :2], axis=1) + 1e-6


def minimal_filter_index(traj):
    delta = traj[1:, :2] - traj[:-1, :2]
    mean_angle = traj[:-1, 2] + wrap_angles(traj[1:, 2] - traj[:-1, 2]) / 2
    directions = np.cos(mean_angle) * delta[:, 0] + np.sin(mean_angle) * delta[:, 1] > 0
    index = 1
    if len(directions) > 0:
        other = np.nonzero(directions != directions[0])[0]
        if len(other) > 0 and other[0] < 6:
            index = max(int(other[0]), index)
    return index


def process(path, minimal_distance=0.001, distance_step=0.05):
    """path: float32 [n, 3] (x, y, heading) -> float64 [m, 3]; paths of fewer than 3 points are returned unchanged."""
    traj = np.array(path, dtype=np.float32)
    if len(traj) < 3:
        return traj
    traj = filter_trajectory(traj, minimal_distance)
    d = chord_lengths(traj)                                                   # float32
    cum = np.concatenate([np.zeros(1), np.cumsum(d)], axis=0)                 # float32 running sum -> float64
    param = cum / cum[-1]
    point_count = int(np.sum(d) / distance_step)
    traj[:, 2] = unfold_angles(traj[:, 2])
    out = quadratic_interp(param, traj, np.linspace(0, 1, point_count))
    return out[minimal_filter_index(out):]

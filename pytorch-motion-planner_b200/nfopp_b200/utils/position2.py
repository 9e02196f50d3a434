"""SE(2) struct-of-arrays (reference utils/position2.py; ROS conversions are out of scope)."""
import numpy as np


class Position2(object):
    def __init__(self, x, y, angle):
        self._x, self._y, self._angle = x, y, angle

    x = property(lambda self: self._x)
    y = property(lambda self: self._y)
    rotation = property(lambda self: self._angle)

    @property
    def translation(self):
        return np.array([self._x, self._y]).T

    @classmethod
    def from_vec(cls, vec):
        vec = np.asarray(vec) if isinstance(vec, list) else vec
        if vec.ndim == 1:
            return cls(vec[0], vec[1], vec[2])
        return cls(vec[:, 0], vec[:, 1], vec[:, 2])

    def as_vec(self):
        return np.array([self._x, self._y, self._angle]).T

    @classmethod
    def from_array(cls, positions):
        return cls(*(np.array([getattr(p, k) for p in positions]) for k in ("x", "y", "rotation")))

    def as_array(self):
        return [Position2(self._x[i], self._y[i], self._angle[i]) for i in range(len(self))]

    def __len__(self):
        return 1 if np.ndim(self._x) == 0 else np.shape(self._x)[0]

    def __mul__(self, other):
        c, s = np.cos(self._angle), np.sin(self._angle)
        angle = (other.rotation + self._angle + np.pi) % (2 * np.pi) - np.pi
        return Position2(other.x * c - other.y * s + self._x, other.x * s + other.y * c + self._y, angle)

    def inv(self):
        c, s = np.cos(self._angle), np.sin(self._angle)
        return Position2(-self._x * c - self._y * s, self._x * s - self._y * c, -self._angle)

    def apply(self, points):
        px, py = np.asarray(points).T
        c, s = np.cos(self._angle), np.sin(self._angle)
        return np.stack([px * c - py * s + self._x, px * s + py * c + self._y], axis=1)

"""wrap_angle as the reference's torch_math.py:5-7 computes it (float32, python-sign modulo)."""
import math

import torch


def wrap_angle(angles: torch.Tensor) -> torch.Tensor:
    shifted = angles + math.pi
    return torch.remainder(shifted, 2 * math.pi) - math.pi

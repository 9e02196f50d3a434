"""Learnable-phase heading encoding (reference angle_encoder.py:6-22).  The module only owns the
parameters; the arithmetic runs inside the fused ONF kernels."""
import math

import torch
from torch import nn


class AngleEncoder(nn.Module):
    def __init__(self, encoding_dimension=10):
        super().__init__()
        if encoding_dimension != 10:
            raise ValueError("the fused ONF kernels implement the reference's 10+10 heading features only")
        self._encoding_dimension = encoding_dimension
        self._biases = nn.Parameter(torch.zeros(2 * encoding_dimension), requires_grad=True)
        freq = torch.linspace(1, encoding_dimension, encoding_dimension)
        # requires_grad=False as in the reference; it still ends up trained because the planner switches
        # requires_grad_ on the whole module and hands every parameter to Adam
        self._frequencies = nn.Parameter(torch.cat([freq, freq], dim=0), requires_grad=False)
        nn.init.uniform_(self._biases, -math.pi, math.pi)

    @property
    def encoding_dimension(self):
        return 2 * self._encoding_dimension

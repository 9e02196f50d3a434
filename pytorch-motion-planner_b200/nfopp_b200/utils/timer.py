"""Global tick/tock wall-clock registry with the reference's phase names (utils/timer.py:7-27)."""
import time
from collections import defaultdict

import numpy as np


class Timer(object):
    def __init__(self):
        self._starts, self._deltas, self._counts = {}, defaultdict(list), defaultdict(int)

    def tick(self, name=None):
        self._counts[name] += 1
        self._starts[name] = time.time()

    def tock(self, name=None):
        self._deltas[name].append(time.time() - self._starts.get(name, time.time()))

    def summary(self):
        return {k: (1000 * float(np.mean(v)), 1000 * float(np.std(v))) for k, v in self._deltas.items()}

    def print(self):
        for key, (mean, std) in self.summary().items():
            print("Duration of", key, "=", np.round(mean, 1), "+-", np.round(std, 1), "ms")


timer = Timer()

"""Plain record of a toy planning problem (reference: test_environment.py:1-6)."""
from dataclasses import dataclass
from typing import Any


@dataclass
class TestEnvironment:
    __test__ = False   # not a pytest class
    start_point: Any
    goal_point: Any
    bounds: Any
    obstacle_points: Any

"""Disc robot fed with Position2 poses (reference circle_directed_collision_checker.py:4-6); the base class
already accepts both input types, so this only keeps the reference's class name."""
from .circle_collision_checker import CircleCollisionChecker


class CircleDirectedCollisionChecker(CircleCollisionChecker):
    pass

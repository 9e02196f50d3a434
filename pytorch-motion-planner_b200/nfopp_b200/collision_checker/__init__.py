from .collision_checker import CollisionChecker
from .circle_collision_checker import CircleCollisionChecker
from .rectangle_collision_checker import RectangleCollisionChecker
from .circle_directed_collision_checker import CircleDirectedCollisionChecker
from .grid_collision_checker import GridCollisionChecker, MapImage

__all__ = ["CollisionChecker", "CircleCollisionChecker", "RectangleCollisionChecker",
           "CircleDirectedCollisionChecker", "GridCollisionChecker", "MapImage"]

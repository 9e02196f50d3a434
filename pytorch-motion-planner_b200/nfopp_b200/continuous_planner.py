"""Abstract planner interface - same six methods as the reference's continuous_planner.py:4-27."""
import abc


class ContinuousPlanner(abc.ABC):
    @abc.abstractmethod
    def init(self, start_point, goal_point, boundaries): ...

    @abc.abstractmethod
    def step(self): ...

    @abc.abstractmethod
    def get_path(self): ...

    @abc.abstractmethod
    def set_boundaries(self, boundaries): ...

    @abc.abstractmethod
    def update_goal_point(self, goal_point): ...

    @abc.abstractmethod
    def update_start_point(self, start_point): ...

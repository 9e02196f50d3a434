"""AstarTrajectoryInitializer - the reference's constructor and name (astar/astar_trajectory_initializer.py:10-13) so that
PlannerFactory configs with `trajectory_initializer=dict(name="AstarTrajectoryInitializer", resolution=...)`
(scripts/run_bench_mr.py:23-26) resolve unchanged.  The work itself - labelled cell grid, A*, quadratic
re-parametrisation - runs on the device for all problems of the engine (PlannerEngine.init_astar, csrc/astar.cu)."""
from .trajectory_initializer import TrajectoryInitializer


class AstarTrajectoryInitializer(TrajectoryInitializer):
    def __init__(self, collision_checker, resolution, init_angles_with_trajectory=False):
        super().__init__(collision_checker, init_angles_with_trajectory)
        self._resolution = resolution

    def initialize_engine(self, engine):
        """x, y from A* over the checker-labelled grid; headings as TrajectoryInitializer.initialize_angle (:21-29)."""
        engine.init_trajectories()
        # the cell grid spans collision_checker.get_boundaries() (:29-34) - also for a checker whose labels do not include
        # its boundaries (GridCollisionChecker)
        extent = None
        checker = getattr(self, "_collision_checker", None)
        if checker is not None and hasattr(checker, "get_boundaries"):
            try:
                extent = checker.get_boundaries()
            except Exception:
                extent = None
        return engine.init_astar(self._resolution, extent=extent)

"""NERFOptPlanner - the reference's holonomic NFOPP planner API (nerf_opt_planner.py:10-248) as a P=1 front
end of the batched CUDA engine.  Same constructor, same methods, same attributes callers read; the whole
step() runs on the GPU (sampling, labelling, ONF training, trajectory update, re-parametrisation)."""
import numpy as np
import torch

from . import _lib as L
from .continuous_planner import ContinuousPlanner
from .engine import EngineConfig, PlannerEngine
from .utils.timer import timer


class NERFOptPlanner(ContinuousPlanner):
    _constrained = False

    def __init__(self, trajectory, collision_model, collision_checker, collision_optimizer, trajectory_optimizer,
                 trajectory_random_offset, collision_weight, velocity_hessian_weight, init_collision_iteration=100,
                 init_collision_points=100, reparametrize_trajectory_freq=10, optimize_collision_model_freq=1,
                 random_field_points=10, collision_loss_koef=1, course_random_offset=1.5, collision_point_count=100,
                 rng_mode="philox", seed=None, use_graph=True):
        self._trajectory = trajectory
        self._device = trajectory.device
        self._collision_model = collision_model
        self._collision_checker = collision_checker
        self._collision_optimizer = collision_optimizer
        self._trajectory_optimizer = trajectory_optimizer
        self._random_sample_border = (0, 0, 0, 0)
        self._fine_random_offset = trajectory_random_offset
        self._collision_weight = collision_weight
        self._velocity_hessian_weight = velocity_hessian_weight
        self._init_collision_iteration = init_collision_iteration
        self._init_collision_points = init_collision_points
        self._reparametrize_trajectory_freq = reparametrize_trajectory_freq
        self._optimize_collision_model_freq = optimize_collision_model_freq
        self._random_field_points = random_field_points
        self._collision_loss_koef = collision_loss_koef
        self._course_random_offset = course_random_offset
        self._collision_point_count = collision_point_count
        self._rng_mode, self._seed, self._use_graph = rng_mode, seed, use_graph
        self._engine = None
        self._checker_version = -1
        self._checker_spec = None          # engine_spec() the live engine was built from
        self._optimizer_key = None         # (lr, betas, eps) of both optimisers the live engine was built from
        self._weights_version = -1
        d = trajectory.shape[1]
        self._start_point = torch.zeros(1, d, device=self._device)
        self._goal_point = torch.zeros(1, d, device=self._device)

    # ---- engine construction -----------------------------------------------------------------------
    def _extra_config(self):
        return {}

    SUPPORTED_CHECKERS = ("CollisionChecker (boundary only)", "CircleCollisionChecker", "CircleDirectedCollisionChecker",
                          "RectangleCollisionChecker", "GridCollisionChecker (the notebook's MapCollisionChecker)")

    def _checker_engine_spec(self):
        """The reference calls collision_checker.check_collision(positions) on ANY object every step
        (nerf_opt_planner.py:93-96).  Here labelling runs inside the CUDA step, so the checker must be one whose arithmetic
        exists as a device kernel; a user-defined checker cannot be called from the GPU and there is no CPU fallback."""
        cc = self._collision_checker
        if not hasattr(cc, "engine_spec"):
            raise TypeError(f"collision checker {type(cc).__name__} cannot run inside the CUDA planner step: supported are "
                            f"nfopp_b200.collision_checker.{{{', '.join(self.SUPPORTED_CHECKERS)}}} "
                            "(labelling is a device kernel; there is no CPU fallback - see INTEGRATION.md)")
        return cc.engine_spec()

    def _optimizer_hyperparameters(self):
        og = self._collision_optimizer.param_groups[0]
        tg = self._trajectory_optimizer.param_groups[0]
        return (float(og["lr"]), tuple(og["betas"]), float(og["eps"]), float(tg["lr"]), tuple(tg["betas"]))

    def _engine_config(self) -> EngineConfig:
        og = self._collision_optimizer.param_groups[0]
        tg = self._trajectory_optimizer.param_groups[0]
        spec = self._checker_engine_spec()
        obstacles = np.asarray(self._collision_checker._obstacle_points).reshape(-1, 2)
        seed = self._seed if self._seed is not None else int(np.random.randint(0, 2 ** 31 - 1))
        cfg = EngineConfig(
            problems=1, constrained=self._constrained, trajectory_length=self._trajectory.shape[0],
            onf=self._collision_model.spec, onf_lr=og["lr"], onf_betas=tuple(og["betas"]), traj_lr=tg["lr"],
            traj_betas=tuple(tg["betas"]), adam_eps=og["eps"],
            trajectory_random_offset=self._fine_random_offset, collision_weight=self._collision_weight,
            velocity_hessian_weight=self._velocity_hessian_weight,
            init_collision_iteration=self._init_collision_iteration,
            init_collision_points=self._init_collision_points,
            reparametrize_trajectory_freq=self._reparametrize_trajectory_freq,
            optimize_collision_model_freq=self._optimize_collision_model_freq,
            random_field_points=self._random_field_points, course_random_offset=self._course_random_offset,
            collision_point_count=self._collision_point_count, env_shared=True,
            max_obstacles=max(64, 2 * len(obstacles)), border=tuple(float(b) for b in self._random_sample_border),
            rng_mode=self._rng_mode, seed=seed, **self._extra_config())
        for k, v in spec.items():
            setattr(cfg, k, v)
        self._checker_spec = dict(spec)
        self._optimizer_key = self._optimizer_hyperparameters()
        return cfg

    def _build_engine(self, keep_state=False):
        L.require_cuda(self._device)
        old = self._engine.export_state() if (keep_state and self._engine is not None) else None
        weights = self._collision_model.packed_parameters().detach().clone()
        traj = self._trajectory.detach().clone()
        self._engine = PlannerEngine(self._engine_config(), device=self._device)
        eng = self._engine
        if old is not None:
            eng.import_state(old)
        else:
            eng["onf_w"][0].copy_(weights)
            eng["traj"][0].copy_(traj)
        # the caller's tensors become views of the engine's state (SURVEY.md §8b "ownership")
        self._collision_model._repack(storage=eng["onf_w"][0])
        self._trajectory.data = eng["traj"][0]
        self._sync_environment(force=True)

    def _sync_environment(self, force=False):
        """The reference re-reads obstacle points AND boundaries on every check_collision call
        (collision_checker/*.py); here both live in the engine, so every change of the checker is carried over: new
        boundaries through nfopp_engine_set_bounds, a cloud that outgrew the obstacle buffer (ros/collision_checker_adapter.py
        feeds a lidar cloud of a different size on every callback) or any other change of the checker's geometry by
        re-creating the engine around the exported state."""
        cc = self._collision_checker
        if not force and cc._version == self._checker_version:
            return
        eng = self._engine
        if not force:
            spec = self._checker_engine_spec()
            old = self._checker_spec or {}
            changed = {k for k in spec if spec[k] != old.get(k)}
            n_obs = len(np.asarray(cc._obstacle_points).reshape(-1, 2)) if cc.kind in ("circle", "rectangle") else 0
            if changed - {"checker_bounds"} or n_obs > eng.config.max_obstacles:
                self._build_engine(keep_state=True)          # (ends in _sync_environment(force=True))
                return
            if "checker_bounds" in changed:
                eng.set_bounds(checker_bounds=spec["checker_bounds"])
                self._checker_spec = dict(spec)
        if cc.kind in ("circle", "rectangle"):
            eng.set_obstacles(np.asarray(cc._obstacle_points, dtype=np.float64).reshape(-1, 2))
        elif cc.kind == "grid":
            eng.set_grid(cc._map_image.map_image)
        self._checker_version = cc._version

    # ---- ContinuousPlanner API -----------------------------------------------------------------------
    def init(self, start_point, goal_point, boundaries):
        """nerf_opt_planner.py:181-187: new end points, a fresh initial trajectory, the ONF warm-up and step_count = 0 -
        nothing else.  On a second init() (a new goal, ros/goal_planner_adapter.py:36) the optimisers' moments and step
        counters, the replay buffer, _previous_trajectory and the multipliers persist, as they do in the reference."""
        d = self._trajectory.shape[1]
        self._random_sample_border = boundaries
        if self._engine is None:
            self._build_engine()
        elif self._optimizer_hyperparameters() != self._optimizer_key:
            self._build_engine(keep_state=True)
        else:
            self._engine.set_bounds(border=boundaries)
            self._sync_environment()
        eng = self._engine
        eng.set_endpoints(np.asarray(start_point, dtype=np.float32)[:d], np.asarray(goal_point, dtype=np.float32)[:d])
        self._start_point, self._goal_point = eng["start"], eng["goal"]
        initializer = getattr(self, "_trajectory_initializer", None)
        if hasattr(initializer, "initialize_engine"):          # AstarTrajectoryInitializer
            initializer.initialize_engine(eng)
        else:
            eng.init_trajectories()
        if d == 3 and getattr(initializer, "_init_angles_with_trajectory", False):     # trajectory_initializer.py:28-44
            from .trajectory_initializer import blend_headings_with_path_direction
            blend_headings_with_path_direction(eng["traj"], eng["start"], eng["goal"])
        eng.warmup(self._init_collision_iteration)
        eng.set_counter("step_count", 0)

    def step(self):
        if self._engine is None:
            raise RuntimeError("call init(start_point, goal_point, boundaries) before step()")
        timer.tick("step")
        self._sync_environment()
        # the reference steps caller-owned torch.optim.Adam objects (planner_factory.py:54-58,68-70): a caller may change
        # param_groups[0]["lr" | "betas" | "eps"] between steps - carry it over (engine re-created around its state)
        if self._optimizer_hyperparameters() != self._optimizer_key:
            self._build_engine(keep_state=True)
        # the caller holds the ONF module whose parameters are views of the engine's weights (planner._collision_model,
        # scripts/run_planner.py:72-73): only when torch has written into them since the last step (a version counter
        # moved) does the tensor path have to rebuild its operand copy
        if self._collision_model.weights_version() != self._weights_version:
            self._engine.invalidate_weights()
        self._engine.step(1, graph=self._use_graph and self._rng_mode == "philox")
        self._weights_version = self._collision_model.weights_version()
        timer.tock("step")

    def full_trajectory(self):
        return self._engine.paths()[0]

    def get_path(self):
        return self.full_trajectory().detach().cpu().numpy()

    def set_boundaries(self, boundaries):
        """nerf_opt_planner.py:220-222: store the sampling rectangle, reset the step counter (no engine re-creation)."""
        self._random_sample_border = boundaries
        if self._engine is not None:
            self._engine.set_bounds(border=boundaries)
            self._engine.set_counter("step_count", 0)

    def reparametrize_trajectory(self):
        with torch.cuda.device(self._device):
            L.check(L.lib().nfopp_engine_reparametrize(self._engine._h, L.stream_ptr()))

    def _snap(self, point, to_goal, xy_only):
        """update_goal_point / update_start_point (nerf_opt_planner.py:202-218; constrained:178-194)."""
        eng = self._engine
        d = self._trajectory.shape[1]
        point = torch.as_tensor(np.asarray(point, dtype=np.float32)[:d], device=self._device)[None]
        (eng["goal"] if to_goal else eng["start"]).copy_(point)
        traj = eng["traj"][0]
        cols = 2 if xy_only else d
        k = int(torch.argmin(torch.sum((traj[:, :cols] - point[:, :cols]) ** 2, dim=1)))
        if xy_only:
            k = min(k + 1, traj.shape[0])
        if to_goal:
            traj[k:] = point
        else:
            traj[:k] = point
        self.reparametrize_trajectory()
        eng.set_counter("step_count", 0)

    def update_goal_point(self, goal_point):
        self._snap(goal_point, True, self._constrained)

    def update_start_point(self, start_point):
        self._snap(start_point, False, self._constrained)

    # ---- attributes the reference's callers read -----------------------------------------
    @property
    def _step_count(self):
        return 0 if self._engine is None else self._engine.counter("step_count")

    @property
    def _collision_positions(self):
        if self._engine is None:
            return np.zeros((0, self._trajectory.shape[1]))
        n = int(self._engine["buf_n"][0])
        return self._engine["buf_pos"][0, :n].cpu().numpy()

    @property
    def _collision_positions_times(self):
        if self._engine is None:
            return np.zeros(0)
        n = int(self._engine["buf_n"][0])
        return self._engine["buf_age"][0, :n].cpu().numpy()

    def _checked(self):
        n = int(self._engine["train_n"][0])
        return self._engine["train_pos"][0, :n].cpu().numpy(), self._engine["train_y"][0, :n].cpu().numpy() > 0.5

    @property
    def checked_positions(self):
        if self._engine is None:
            return np.zeros((0, 3))
        return self._checked()[0]

    @property
    def truth_collision(self):
        if self._engine is None:
            return np.zeros(0, dtype=bool)
        return self._checked()[1]

"""PlannerEngine: P independent NFOPP problems stepped together on one GPU through libnfopp.so.

This is the host side of the batched hot path.  It owns nothing but a torch uint8 workspace tensor; every
piece of planner state (ONF weights + Adam moments, trajectories, multipliers, replay buffers, environments)
lives in named buffers inside that workspace, laid out by the C library and exposed here as torch views -
that is export_state()/import_state() and how parity tests inject the oracle's state and
variates.  The reference's single-problem classes (NERFOptPlanner, ConstrainedNERFOptPlanner) are thin
P=1 wrappers over this class.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib as L

_DTYPES = {b"f": torch.float32, b"d": torch.float64, b"i": torch.int32, b"b": torch.uint8}
CHECKER_KINDS = {"boundary": 0, "circle": 1, "rectangle": 2, "grid": 3}


@dataclass
class OnfSpec:
    """onf_model.py:8 ctor arguments (+ hidden/layers, fixed to 100/2 in the reference)."""
    mean: float = 0.0
    sigma: float = 1.0
    use_cos: bool = False
    angle: bool = False
    hidden: int = 100
    layers: int = 2
    bias: bool = True

    def to_c(self) -> L.OnfConfig:
        return L.OnfConfig(float(self.mean), float(self.sigma), int(self.use_cos), int(self.angle), int(self.hidden),
                           int(self.layers), int(self.bias))

    def layout(self) -> L.OnfLayout:
        lay = L.OnfLayout()
        cfg = self.to_c()
        L.check(L.lib().nfopp_onf_get_layout(C.byref(cfg), C.byref(lay)))
        return lay


def onf_segments(spec: OnfSpec):
    """name -> (offset, shape) of the packed ONF blob, plus its length (C library is the source of truth)."""
    lay = spec.layout()
    segs = {"enc_w": (lay.off_enc_w, (lay.f_pos, 2)), "enc_b": (lay.off_enc_b, (lay.f_pos,)),
            "ang_b": (lay.off_ang_b, (lay.n_angle,)), "ang_f": (lay.off_ang_f, (lay.n_angle,))}
    for l in range(spec.layers):
        segs[f"W{l}"] = (lay.off_w[l], (spec.hidden, lay.f_in if l == 0 else spec.hidden))
        segs[f"b{l}"] = (lay.off_b[l], (spec.hidden,))
    segs["w_out"] = (lay.off_w_out, (spec.hidden + lay.f_in,))
    segs["b_out"] = (lay.off_b_out, (1,))
    return segs, lay.n_params


@dataclass
class EngineConfig:
    """Everything NERFOptPlanner / ConstrainedNERFOptPlanner / PlannerFactory configure, for P problems.
    Field names follow nerf_opt_planner.py:11-14 and constrained_nerf_opt_planner.py:13-18."""
    problems: int = 1
    constrained: bool = True
    trajectory_length: int = 100
    onf: OnfSpec = field(default_factory=OnfSpec)
    onf_lr: float = 1e-2
    onf_betas: Tuple[float, float] = (0.9, 0.9)
    traj_lr: float = 1e-2
    traj_betas: Tuple[float, float] = (0.9, 0.9)
    adam_eps: float = 1e-8
    trajectory_random_offset: float = 0.02
    collision_weight: float = 1.0
    velocity_hessian_weight: float = 0.5
    init_collision_iteration: int = 100
    init_collision_points: int = 100
    reparametrize_trajectory_freq: int = 10
    optimize_collision_model_freq: int = 1
    random_field_points: int = 10
    course_random_offset: float = 1.5
    collision_point_count: int = 100
    angle_weight: float = 0.5
    constraint_deltas_weight: float = 20.0
    multipliers_lr: float = 0.1
    boundary_weight: float = 1.0
    collision_multipliers_lr: float = 1e-3
    angle_offset: float = 0.0
    collision_beta: float = 1.0
    direction_delta_weight: float = 0.0
    # collision checker
    checker: str = "circle"
    env_shared: bool = True
    max_obstacles: int = 64
    robot_radius: float = 0.3
    box: Tuple[float, float, float, float] = (-0.3, 0.2, -0.3, 0.2)
    checker_bounds: Optional[Tuple[float, float, float, float]] = None
    grid_shape: Tuple[int, int] = (0, 0)
    grid_origin: Tuple[float, float] = (0.0, 0.0)
    grid_cell: float = 0.1
    border: Tuple[float, float, float, float] = (0.0, 0.0, 0.0, 0.0)
    # randomness
    rng_mode: str = "philox"            # "philox" (on-device) | "inject" (host supplies every variate)
    seed: int = 0
    max_choice_uniforms: int = 0
    # ONF kernels: "auto" | "fma" (CUDA cores) | "tensor" (fused tcgen05 kernel, reference network shape) |
    # "wide" (tcgen05 grouped GEMM per layer: wide / deep networks); auto = tensor when the shape fits, else wide
    onf_impl: str = "auto"

    @property
    def dim(self):
        return 3 if self.constrained else 2

    def to_c(self) -> L.PlannerConfig:
        c = L.PlannerConfig()
        c.onf = self.onf.to_c()
        c.problems, c.constrained, c.trajectory_length = self.problems, int(self.constrained), self.trajectory_length
        c.onf_lr, (c.onf_beta1, c.onf_beta2) = self.onf_lr, self.onf_betas
        c.traj_lr, (c.traj_beta1, c.traj_beta2) = self.traj_lr, self.traj_betas
        c.adam_eps = self.adam_eps
        c.trajectory_random_offset, c.course_random_offset = self.trajectory_random_offset, self.course_random_offset
        c.collision_weight = self.collision_weight
        c.init_collision_points = self.init_collision_points
        c.reparametrize_trajectory_freq = self.reparametrize_trajectory_freq
        c.optimize_collision_model_freq = self.optimize_collision_model_freq
        c.random_field_points, c.collision_point_count = self.random_field_points, self.collision_point_count
        c.angle_offset, c.angle_weight = self.angle_offset, self.angle_weight
        c.constraint_deltas_weight, c.multipliers_lr = self.constraint_deltas_weight, self.multipliers_lr
        c.boundary_weight, c.collision_multipliers_lr = self.boundary_weight, self.collision_multipliers_lr
        c.collision_beta, c.direction_delta_weight = self.collision_beta, self.direction_delta_weight
        c.checker_kind = CHECKER_KINDS[self.checker]
        c.env_shared, c.max_obstacles = int(self.env_shared), self.max_obstacles
        c.grid_rows, c.grid_cols = self.grid_shape
        c.grid_origin_x, c.grid_origin_y = self.grid_origin
        c.grid_cell, c.robot_radius = self.grid_cell, self.robot_radius
        c.box = (C.c_double * 4)(*self.box)
        c.has_bounds = int(self.checker_bounds is not None)
        c.bounds = (C.c_double * 4)(*(self.checker_bounds or (0, 0, 0, 0)))
        c.border = (C.c_double * 4)(*self.border)
        c.rng_mode = {"philox": 0, "inject": 1}[self.rng_mode]
        c.seed, c.max_choice_uniforms = self.seed, self.max_choice_uniforms
        c.onf_impl = {"auto": 0, "fma": 1, "tensor": 2, "wide": 3}[self.onf_impl]
        return c


def inverse_velocity_hessian(n: int, weight: float) -> np.ndarray:
    """inv(weight * tridiag(-2, 4, -2) + I), float64 inverse then float32 - nerf_opt_planner.py:45-58."""
    k = 4.0 * np.eye(n, dtype=np.float32) - 2.0 * np.eye(n, k=1, dtype=np.float32) - 2.0 * np.eye(n, k=-1,
                                                                                               dtype=np.float32)
    return np.linalg.inv(weight * k + np.eye(n)).astype(np.float32)


def batched_linspace(start: torch.Tensor, end: torch.Tensor, steps: int) -> torch.Tensor:
    """torch.linspace(start_i, end_i, steps) for a batch, with torch's CPU float32 formula (step computed
    once, lower half from the start, upper half from the end, fused multiply-add) so P=1 reproduces the reference's
    trajectory_initializer.py:12-27 bit for bit."""
    start, end = start.to(torch.float32), end.to(torch.float32)
    step = ((end - start) / (steps - 1)).double()          # float32 step, as torch computes it
    idx = torch.arange(steps, device=start.device, dtype=torch.float64)
    # torch's kernel evaluates start + step*i with one rounding (FMA): the product of a float32 and a small
    # integer is exact in float64, so float64 arithmetic rounded once to float32 reproduces it
    lower = (start.double()[..., None] + step[..., None] * idx).float()
    upper = (end.double()[..., None] - step[..., None] * (steps - 1 - idx)).float()
    return torch.where(idx < steps // 2, lower, upper)


class PlannerEngine:
    COUNTERS = {"step_count": 0, "onf_step": 1, "traj_step": 2, "has_prev": 3}

    def __init__(self, config: EngineConfig, device="cuda"):
        L.require_cuda(device)
        self.config = config
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self._lib = L.lib()
        self._c = config.to_c()
        nbytes = C.c_size_t()
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_workspace_bytes(C.byref(self._c), C.byref(nbytes)))
            self.workspace = torch.empty(nbytes.value + 256, dtype=torch.uint8, device=self.device)
            base = self.workspace.data_ptr()
            self._ws_off = (-base) % 256
            handle = C.c_void_p()
            L.check(self._lib.nfopp_engine_create(C.byref(self._c), C.c_void_p(base + self._ws_off), nbytes.value,
                                                  C.byref(handle)))
        self._h = handle
        self.buffers = {}
        for i in range(self._lib.nfopp_engine_buffer_count(self._h)):
            name = self._lib.nfopp_engine_buffer_name(self._h, i).decode()
            self.buffers[name] = self._view(name)
        self._stream = torch.cuda.Stream(device=self.device)   # CUDA-graph replay needs a non-default stream
        self._constants_set = False
        self.set_constants()

    # ---- buffers -----------------------------------------------------------------------------------
    def _view(self, name):
        p, nb, dt, dims = C.c_void_p(), C.c_size_t(), C.c_char(), (C.c_int32 * 3)()
        L.check(self._lib.nfopp_engine_buffer(self._h, name.encode(), C.byref(p), C.byref(nb), C.byref(dt),
                                              C.byref(dims)))
        off = p.value - self.workspace.data_ptr()
        dtype = _DTYPES[dt.value]
        shape = [d for d in dims if d > 0]
        return self.workspace[off:off + nb.value].view(dtype).view(*shape)

    def __getitem__(self, name) -> torch.Tensor:
        if name == "onf_w":
            # the caller may write the weights through this view: the tensor path rebuilds its bf16 operand copy
            # at the next step.  Writers that keep a view (ONF modules adopted by a planner) call
            # invalidate_weights() themselves.
            self.invalidate_weights()
        return self.buffers[name]

    def invalidate_weights(self):
        L.check(self._lib.nfopp_engine_invalidate_weights(self._h))

    def set_constants(self):
        cfg = self.config
        T = cfg.trajectory_length
        ih = inverse_velocity_hessian(T, cfg.velocity_hessian_weight)
        self["inv_hessian_t"].copy_(torch.from_numpy(np.ascontiguousarray(ih.T)))
        self["lin_s"].copy_(torch.linspace(0, 1, T + 2)[1:-1])          # nerf_opt_planner.py:230

    # ---- environment / problem definition ---------------------------------------------------------
    def set_obstacles(self, points, counts=None):
        """points: [M,2] (shared) or [P,M,2] float64 obstacle points (collision_checker.update_obstacle_points)."""
        buf = self["obstacles"]
        pts = torch.as_tensor(np.asarray(points, dtype=np.float64))
        if pts.dim() == 2:
            pts = pts[None].expand(buf.shape[0], -1, -1)
        m = pts.shape[1]
        if m > buf.shape[1]:
            raise ValueError(f"{m} obstacle points exceed max_obstacles={buf.shape[1]}")
        buf.zero_()
        buf[:, :m].copy_(pts)
        n = self["obstacle_n"]
        if counts is None:
            n.fill_(m)
        else:
            n.copy_(torch.as_tensor(np.asarray(counts, dtype=np.int32)))

    def set_grid(self, grid):
        g = torch.as_tensor(np.asarray(grid) > 0).to(torch.uint8)
        buf = self["grid"]
        buf.copy_(g if g.dim() == 3 else g[None].expand_as(buf))

    def set_endpoints(self, start, goal):
        self["start"].copy_(torch.as_tensor(np.asarray(start, dtype=np.float32)).reshape(self["start"].shape))
        self["goal"].copy_(torch.as_tensor(np.asarray(goal, dtype=np.float32)).reshape(self["goal"].shape))

    def init_trajectories(self):
        """NERFOptPlanner._init_trajectory (nerf_opt_planner.py:189-195) / TrajectoryInitializer
        (trajectory_initializer.py:12-29): straight line in x,y, wrapped-linear heading."""
        from .torch_math import wrap_angle
        start, goal = self["start"], self["goal"]
        n = self.config.trajectory_length + 2
        traj = self["traj"]
        for d in range(2):
            traj[:, :, d] = batched_linspace(start[:, d], goal[:, d], n)[:, 1:-1]
        if self.config.constrained:
            goal_angle = wrap_angle(goal[:, 2] - start[:, 2]) + start[:, 2]
            traj[:, :, 2] = batched_linspace(start[:, 2], goal_angle, n)[:, 1:-1]

    def init_astar(self, resolution, extent=None):
        """AstarTrajectoryInitializer.initialize_trajectory (astar/astar_trajectory_initializer.py:15-46) for every problem,
        on the device: checker-labelled cell grid -> A* (astar/jps.py, jps=False) -> quadratic re-parametrisation into the
        x, y of the trajectories (call after init_trajectories, which sets the headings).  -> int32 [P] cells per A* path.
        extent = collision_checker.get_boundaries() (x_min, x_max, y_min, y_max) of the cell grid; default: the checker's
        labelling boundaries (the two differ for a checker that reports boundaries without OR-ing them into its labels)."""
        if extent is None:
            extent = self.config.checker_bounds
        if extent is None:
            raise ValueError("A* initialisation needs collision-checker boundaries (get_boundaries)")
        b = tuple(float(v) for v in extent)
        res = float(resolution)
        # grid shape and start / goal cells exactly as the reference computes them (:29-34: python floor division on the
        # boundaries, float32 floor division on the float32 end points)
        x_cells = int((b[1] - b[0]) // res) + 1
        y_cells = int((b[3] - b[2]) // res) + 1
        cells = []
        for name in ("start", "goal"):
            pt = self[name].detach().cpu().numpy().astype(np.float32)
            cx = ((pt[:, 0] - np.float32(b[0])) // np.float32(res)).astype(np.int64)
            cy = ((pt[:, 1] - np.float32(b[2])) // np.float32(res)).astype(np.int64)
            if ((cx < 0) | (cx >= x_cells) | (cy < 0) | (cy >= y_cells)).any():
                raise IndexError(f"A* initialisation: a {name} point lies outside the collision-checker boundaries {b}")
            cells.append(torch.as_tensor(np.stack([cy, cx], 1).astype(np.int32)).to(self.device))
        nbytes = C.c_size_t()
        L.check(self._lib.nfopp_engine_astar_workspace_bytes(self._h, x_cells, y_cells, C.byref(nbytes)))
        ws = torch.empty(int(nbytes.value), dtype=torch.uint8, device=self.device)
        path_cells = torch.empty(self.config.problems, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            origin = (C.c_double * 2)(b[0], b[2])
            L.check(self._lib.nfopp_engine_init_astar(self._h, x_cells, y_cells, res, origin, L.ptr(cells[0]), L.ptr(cells[1]), L.ptr(ws),
                                                      C.c_size_t(ws.numel()), L.ptr(path_cells), L.stream_ptr()))
            torch.cuda.current_stream().synchronize()          # the workspace is freed on return
        return path_cells

    # ---- counters ------------------------------------------------------------------------------------
    def counter(self, name) -> int:
        v = C.c_int64()
        L.check(self._lib.nfopp_engine_get_counter(self._h, self.COUNTERS[name], C.byref(v)))
        return int(v.value)

    def set_counter(self, name, value):
        L.check(self._lib.nfopp_engine_set_counter(self._h, self.COUNTERS[name], int(value)))

    # ---- the hot path ------------------------------------------------------------------------------
    def warmup(self, iterations=None):
        it = self.config.init_collision_iteration if iterations is None else iterations
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_warmup(self._h, int(it), L.stream_ptr()))

    def step(self, n_steps=1, graph=False):
        with torch.cuda.device(self.device):
            if graph:
                self._stream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self._stream):
                    L.check(self._lib.nfopp_engine_step(self._h, int(n_steps), 1, L.stream_ptr()))
                torch.cuda.current_stream().wait_stream(self._stream)
            else:
                L.check(self._lib.nfopp_engine_step(self._h, int(n_steps), 0, L.stream_ptr()))

    def set_lanes(self, lanes: int):
        """Split the graphed step into `lanes` independent problem ranges on parallel streams inside the one CUDA graph
        (bit-identical results; overlaps the small latency-bound kernels of one lane with the ONF kernels of another)."""
        L.check(self._lib.nfopp_engine_set_lanes(self._h, int(lanes)))

    def set_bounds(self, border=None, checker_bounds="unchanged"):
        """set_boundaries (nerf_opt_planner.py:220-222) / update_boundaries (collision_checker.py:24-25) on the live engine:
        the step graph is re-captured, nothing else changes."""
        b = None if border is None else (C.c_double * 4)(*[float(v) for v in border])
        if checker_bounds == "unchanged":
            mode, cb = -1, None
        elif checker_bounds is None:
            mode, cb = 0, None
        else:
            mode, cb = 1, (C.c_double * 4)(*[float(v) for v in checker_bounds])
        L.check(self._lib.nfopp_engine_set_bounds(self._h, b, mode, cb))
        if border is not None:
            self.config.border = tuple(float(v) for v in border)
        if mode >= 0:
            self.config.checker_bounds = None if mode == 0 else tuple(float(v) for v in checker_bounds)

    def profile_step(self) -> dict:
        """One un-graphed step with CUDA events between its kernels -> {kernel name: milliseconds}."""
        ms = (C.c_float * 9)()
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_profile_step(self._h, ms, 9, L.stream_ptr()))
        return {self._lib.nfopp_engine_kernel_name(i).decode(): float(ms[i]) for i in range(9)}

    def paths(self, out=None) -> torch.Tensor:
        cfg = self.config
        if out is None:
            out = torch.empty(cfg.problems, cfg.trajectory_length + 2, cfg.dim, dtype=torch.float32,
                              device=self.device)
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_get_paths(self._h, L.ptr(out), L.stream_ptr()))
        return out

    def reparametrize(self):
        """NERFOptPlanner.reparametrize_trajectory (nerf_opt_planner.py:224-244; constrained:132-171) for all problems,
        outside a step."""
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_reparametrize(self._h, L.stream_ptr()))

    def evaluate_paths(self, samples_per_segment=4):
        """benchmark.evaluate_path (scripts/run_bench_mr.py:119-126) for every problem, on the device: -> (collides
        bool [P], length float32 [P], colliding pose count int32 [P])."""
        P = self.config.problems
        collides = torch.empty(P, dtype=torch.uint8, device=self.device)
        length = torch.empty(P, dtype=torch.float32, device=self.device)
        hits = torch.empty(P, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_evaluate_paths(self._h, int(samples_per_segment), L.ptr(collides), L.ptr(length),
                                                          L.ptr(hits), L.stream_ptr()))
        return collides.bool(), length, hits

    PATH_STATS = ("path_length", "max_curvature", "normalized_curvature", "aol", "smoothness", "cusps",
                  "min_clearing_distance", "mean_clearing_distance")

    def path_statistics(self, cusp_angle_threshold=60.0 * 3.141592653589793 / 180.0, clearing_window_cells=64):
        """The per-plan statistics BenchmarkAdapter::evaluate files (BenchmarkAdapter.cpp:68-135), for every problem's
        current path, on the device -> {name: float32 [P]} (names as in bench-mr's PathStatistics)."""
        out = torch.empty(self.config.problems, len(self.PATH_STATS), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            L.check(self._lib.nfopp_engine_path_statistics(self._h, float(cusp_angle_threshold), int(clearing_window_cells),
                                                           L.ptr(out), L.stream_ptr()))
        return {k: out[:, i] for i, k in enumerate(self.PATH_STATS)}

    @property
    def launches(self) -> int:
        v = C.c_int64()
        L.check(self._lib.nfopp_engine_launch_count(self._h, C.byref(v)))
        return int(v.value)

    # ---- state import / export (checkpoint, lock-step parity) ----------------------------------------
    STATE_BUFFERS = ["onf_w", "onf_m", "onf_v", "traj", "traj_m", "traj_v", "prev_traj", "lam", "mu", "start",
                     "goal", "buf_pos", "buf_age", "buf_n"]

    def export_state(self) -> dict:
        st = {k: self[k].detach().cpu().clone() for k in self.STATE_BUFFERS}
        st.update({k: self.counter(k) for k in self.COUNTERS})
        return st

    def import_state(self, st: dict):
        for k in self.STATE_BUFFERS:
            if k in st:
                self[k].copy_(torch.as_tensor(st[k]).reshape(self[k].shape))
        for k in self.COUNTERS:
            if k in st:
                self.set_counter(k, st[k])

    def inject_variates(self, var: dict, problem=None):
        """Parity mode: the variates one step consumes, for one or all problems."""
        sel = slice(None) if problem is None else problem
        for name, val in var.items():
            key = "var_" + name
            if key not in self.buffers:
                continue
            buf = self[key][sel]
            val = torch.as_tensor(np.asarray(val)).to(buf.dtype).reshape(-1)
            flat = buf.reshape(buf.shape[0], -1) if problem is None else buf.reshape(-1)
            if name == "choice_u":
                if val.numel() > flat.shape[-1]:
                    raise ValueError("choice uniforms exceed max_choice_uniforms")
                flat[..., :val.numel()] = val
            else:
                flat[...] = val.reshape(flat.shape[-1]) if problem is not None else val.reshape(1, -1)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.nfopp_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

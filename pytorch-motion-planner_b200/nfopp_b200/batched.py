"""BatchedPlanner - P independent NFOPP problems on one GPU (what the reference can only do as P separate
Python processes).  Same verbs as ContinuousPlanner: init / step / get_path, with [P, ...] arrays.

Multi-GPU: problems are sharded contiguously over ranks with no collective inside the loop; `gather_paths`
is the one exchange (torch.distributed all_gather of the final [P_local, T+2, D] paths)."""
import math
from dataclasses import replace

import os

import numpy as np
import torch

from .engine import EngineConfig, OnfSpec, PlannerEngine, onf_segments


def init_onf_weights(spec: OnfSpec, problems: int, use_normal_init: bool, seed: int, device) -> torch.Tensor:
    """[P, n_params] initial ONF weights with torch's nn.Linear default distribution
    (U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weights and biases), N(0,1) Fourier matrix when use_normal_init,
    phases U(-pi, pi), frequencies 1..10 twice - what ONF.__init__ (onf_model.py:8-31) draws, vectorised over
    problems.  Distribution-equivalent to constructing P modules, not stream-identical."""
    segs, n_params = onf_segments(spec)
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed))
    w = torch.zeros(problems, n_params, dtype=torch.float32, device=device)

    def uniform(name, bound):
        off, shape = segs[name]
        n = int(np.prod(shape))
        if n:
            w[:, off:off + n] = (torch.rand(problems, n, generator=gen, device=device) * 2 - 1) * bound

    f_in = segs["W0"][1][1]
    for l in range(spec.layers):
        fan = f_in if l == 0 else spec.hidden
        uniform(f"W{l}", 1 / math.sqrt(fan))
        uniform(f"b{l}", 1 / math.sqrt(fan))
    uniform("w_out", 1 / math.sqrt(spec.hidden + f_in))
    uniform("b_out", 1 / math.sqrt(spec.hidden + f_in))
    uniform("enc_w", 1 / math.sqrt(2))
    if spec.bias:
        uniform("enc_b", 1 / math.sqrt(2))
    if use_normal_init:
        off, shape = segs["enc_w"]
        n = int(np.prod(shape))
        w[:, off:off + n] = torch.randn(problems, n, generator=gen, device=device)
    if spec.angle:
        uniform("ang_b", math.pi)
        off, _ = segs["ang_f"]
        f = torch.linspace(1, 10, 10, device=device)
        w[:, off:off + 20] = torch.cat([f, f])
    return w


def shard_range(total: int, rank: int, world: int):
    """Contiguous, balanced partition of problem ids [0,total) over ranks."""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class BatchedPlanner:
    def __init__(self, config: EngineConfig, device="cuda", use_normal_init=True, lanes="auto"):
        self.config = config
        self.engine = PlannerEngine(config, device=device)
        # lanes of the graphed step: independent problem ranges on parallel streams inside the one CUDA graph.  The streamed
        # tensor kernels are persistent and give every lane #SMs / lanes CTAs, so the lanes really run side by side and one
        # lane's small kernels, launch tails and HBM bursts fall under another lane's tensor work (measured on B200, 1024
        # corridor problems: 0.876 ms/step with 1 lane, 0.844 with 2, 0.839 with 3, 0.838 with 4)
        # (the wide ONF path's persistent kernels gain nothing from lanes: 4.78 -> 4.98 ms/step at 2 lanes, so auto keeps 1)
        narrow = config.onf_impl in ("fma", "tensor") or (config.onf_impl == "auto" and config.onf.layers == 2 and config.onf.hidden <= 111)
        if lanes == "auto" and os.environ.get("NFOPP_LANES"):      # profiling: one launch per kernel over all problems
            lanes = int(os.environ["NFOPP_LANES"])
        self.lanes = ((4 if config.problems >= 592 else 2 if config.problems >= 296 else 1) if narrow else 1) if lanes == "auto" else int(lanes)
        if self.lanes > 1:
            try:
                self.engine.set_lanes(self.lanes)
            except ValueError:                     # wide ONF path: one activation workspace for all problems
                if lanes != "auto":
                    raise
                self.lanes = 1
        self.device = self.engine.device
        self._use_normal_init = use_normal_init

    def init(self, start_points, goal_points, obstacles=None, obstacle_counts=None, grids=None, onf_weights=None,
             astar_resolution=None, init_angles_with_trajectory=False):
        """start/goal: [P, D]; obstacles [P|1, M, 2] or grids [P|1, rows, cols]; onf_weights [P, n_params] or
        None (fresh initialisation seeded by config.seed)."""
        eng = self.engine
        eng.set_endpoints(start_points, goal_points)
        if obstacles is not None:
            eng.set_obstacles(obstacles, obstacle_counts)
        if grids is not None:
            eng.set_grid(grids)
        if onf_weights is None:
            onf_weights = init_onf_weights(self.config.onf, self.config.problems, self._use_normal_init,
                                           self.config.seed + 7919, self.device)
        eng["onf_w"].copy_(torch.as_tensor(onf_weights).to(self.device))
        eng.init_trajectories()
        if astar_resolution is not None:               # AstarTrajectoryInitializer (scripts/run_bench_mr.py:23-26)
            eng.init_astar(astar_resolution)
        if init_angles_with_trajectory and self.config.dim == 3:     # trajectory_initializer.py:28-44, all problems at once
            from .trajectory_initializer import blend_headings_with_path_direction
            blend_headings_with_path_direction(eng["traj"], eng["start"], eng["goal"])
        eng.warmup(self.config.init_collision_iteration)
        eng.set_counter("step_count", 0)

    def step(self, n_steps=1, graph=True):
        self.engine.step(n_steps, graph=graph and self.config.rng_mode == "philox")

    def get_path(self) -> np.ndarray:
        return self.engine.paths().cpu().numpy()

    def paths(self) -> torch.Tensor:
        return self.engine.paths()

    def status(self) -> np.ndarray:
        return self.engine["status"].cpu().numpy()

    # ---- warm re-planning for all problems (receding-horizon use, ros/goal_planner_adapter.py:44-53) ----------------
    def _snap(self, points, to_goal: bool):
        """update_goal_point / update_start_point (nerf_opt_planner.py:202-218; constrained:178-194) vectorised over
        problems: the tail (head) of every trajectory from its node nearest to the new goal (start) is snapped onto
        it, then every trajectory is re-parametrised and the step counter reset."""
        eng, cfg = self.engine, self.config
        pts = torch.as_tensor(np.asarray(points, dtype=np.float32)).to(self.device).reshape(cfg.problems, cfg.dim)
        (eng["goal"] if to_goal else eng["start"]).copy_(pts)
        traj = eng["traj"]                                                  # [P, T, D]
        cols = 2 if cfg.constrained else cfg.dim                            # constrained: x, y only (:181,:190)
        k = torch.argmin(((traj[..., :cols] - pts[:, None, :cols]) ** 2).sum(-1), dim=1)
        if cfg.constrained:
            k = torch.clamp(k + 1, max=traj.shape[1])
        idx = torch.arange(traj.shape[1], device=self.device)[None, :]
        mask = (idx >= k[:, None]) if to_goal else (idx < k[:, None])
        traj.copy_(torch.where(mask[..., None], pts[:, None, :], traj))
        eng.reparametrize()
        eng.set_counter("step_count", 0)

    def update_goal_points(self, goal_points):
        self._snap(goal_points, True)

    def update_start_points(self, start_points):
        self._snap(start_points, False)

    def evaluate_paths(self, samples_per_segment=4):
        """(collides bool [P], length [P], colliding pose count [P]) of the current paths, on the device."""
        return self.engine.evaluate_paths(samples_per_segment)

    def path_statistics(self, **kw):
        """bench-mr PathStatistics of the current paths (BenchmarkAdapter.cpp:68-135) -> {name: float32 [P]}."""
        return self.engine.path_statistics(**kw)

    def postprocess_paths(self, minimal_distance=0.001, distance_step=0.05, max_points=None):
        """PathPostprocessor.process (ros/path_postprocessor.py:13-26) for every problem's current path (3-D poses only) ->
        (poses [P, max_points, 3] float64, counts [P] int32), on the device."""
        from .path_postprocessor import PathPostprocessor
        return PathPostprocessor(minimal_distance, distance_step, self.device).process_batch(self.engine.paths(), max_points)

    def solve(self, max_iterations=1000, min_iterations=200, check_every=50, samples_per_segment=4, graph=True,
              freeze_finished=True):
        """scripts/run_bench_mr.py:109-131 for all problems at once: step, and every `check_every` iterations past
        `min_iterations` evaluate every path on the device; a problem keeps its shortest collision-free path and is
        DONE once a later collision-free path is no shorter (the reference's `break`).  Problems are independent, so
        the batch keeps stepping until all are done or `max_iterations` is reached.  A finished problem is FROZEN
        (engine buffer `frozen`): as in the reference, where every problem is its own run that stops at the `break`, no
        kernel touches it any more and the ONF launches skip it, so the remaining steps get cheaper as problems finish
        (freeze_finished=False keeps stepping them; the results are identical, problems being independent).
        Returns dict(paths [P,T+2,D], solved bool [P], length [P], iterations int [P])."""
        P, dev = self.config.problems, self.device
        best_len = torch.full((P,), float("inf"), device=dev)
        best_path = self.engine.paths().clone()
        stop_path, stop_len = best_path.clone(), torch.zeros(P, device=dev)
        done = torch.zeros(P, dtype=torch.bool, device=dev)
        iters = torch.full((P,), max_iterations, dtype=torch.int32, device=dev)
        frozen = self.engine["frozen"]
        frozen.zero_()
        i = 0
        while i < max_iterations:
            # the reference checks after step i (0-based) when i > min_iterations and i % check_every == 0
            nxt = max(min_iterations + 1, i)
            nxt += (-nxt) % check_every
            run = min(nxt, max_iterations - 1) - i + 1
            self.step(run, graph=graph)
            i += run
            if i - 1 != nxt:
                break
            collides, length, _ = self.engine.evaluate_paths(samples_per_segment)
            free = ~collides & ~done
            better = free & (length < best_len)
            stop = free & ~better                           # collision-free but no shorter: the reference's `break`
            if bool(free.any()):
                paths = self.engine.paths()
                best_path[better] = paths[better]
                best_len = torch.where(better, length, best_len)
                stop_path[stop] = paths[stop]               # the reference returns the path it broke on
                stop_len = torch.where(stop, length, stop_len)
            iters = torch.where(stop, torch.full_like(iters, i), iters)
            done |= stop
            if freeze_finished:
                frozen.copy_(done.to(frozen.dtype))
            if bool(done.all()):
                break
        collides, length, _ = self.engine.evaluate_paths(samples_per_segment)
        paths = self.engine.paths()
        use_best = ~done & collides & torch.isfinite(best_len)      # run_bench_mr.py:128-129
        paths[done] = stop_path[done]
        paths[use_best] = best_path[use_best]
        length = torch.where(done, stop_len, torch.where(use_best, best_len, length))
        solved = done | ~collides | use_best
        frozen.zero_()                                      # later step() calls move every problem again
        return dict(paths=paths, solved=solved, length=length, iterations=iters, steps_run=i)

    def gather_paths(self) -> torch.Tensor:
        """All ranks' final paths on every rank (the only collective of the whole job).  Ranks may own
        different numbers of problems; shards are padded to the largest for the all_gather."""
        import torch.distributed as dist
        local = self.engine.paths()
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return local
        world = dist.get_world_size()
        counts = [torch.zeros(1, dtype=torch.int64, device=local.device) for _ in range(world)]
        dist.all_gather(counts, torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device))
        biggest = int(max(int(c) for c in counts))
        padded = torch.zeros(biggest, *local.shape[1:], dtype=local.dtype, device=local.device)
        padded[:local.shape[0]] = local
        parts = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(parts, padded)
        return torch.cat([p[:int(c)] for p, c in zip(parts, counts)], 0)


class HostStepPipeline:
    """planner.step() driven from HOST buffers: what a caller of the reference does every control cycle -
    `update_obstacle_points` (ros/planner_adapter.py, scripts/run_planner.py) or a new map, new start / goal, `step()`,
    `get_path()` - for all problems at once, with the transfers taken off the compute stream.

    Every step's inputs go pinned host -> device staging buffers on a copy-in stream, are moved into the engine's buffers
    by a device copy on the compute stream, and the step's paths go device -> pinned host on a copy-out stream; both
    directions use two alternating buffer sets, so the copies of steps k-1 / k+1 run under the kernels of step k.  Results
    are identical to copying synchronously (tests/test_gpu_api.py)."""

    def __init__(self, planner: "BatchedPlanner", names):
        eng = planner.engine
        self.planner, self.names = planner, tuple(names)
        cfg = planner.config
        with torch.cuda.device(eng.device):
            self.stage = [{k: torch.empty_like(eng[k]) for k in self.names} for _ in range(2)]
            shape = (cfg.problems, cfg.trajectory_length + 2, cfg.dim)
            self.d_paths = [torch.empty(shape, dtype=torch.float32, device=eng.device) for _ in range(2)]
            self.h_paths = [torch.empty(shape, dtype=torch.float32).pin_memory() for _ in range(2)]
            self.in_stream, self.out_stream = torch.cuda.Stream(eng.device), torch.cuda.Stream(eng.device)
            self._ev = {k: [torch.cuda.Event() for _ in range(2)] for k in ("loaded", "consumed", "ready", "drained")}
        self.count = 0
        self.h2d_bytes = sum(eng[k].numel() * eng[k].element_size() for k in self.names)
        self.d2h_bytes = self.h_paths[0].numel() * 4

    def pinned(self, name, value) -> torch.Tensor:
        """`value` as a pinned host tensor of the engine buffer's dtype and shape (so the copy is one async DMA)."""
        ref = self.planner.engine[name]
        return torch.as_tensor(np.ascontiguousarray(value)).to(ref.dtype).reshape(ref.shape).pin_memory()

    def step(self, host_inputs) -> torch.Tensor:
        """Enqueue one step fed from `host_inputs` ({engine buffer name: pinned host tensor}); returns the pinned host
        tensor this step's paths are being written to - valid after `synchronize()` and until two steps later."""
        eng, ev = self.planner.engine, self._ev
        i, b = self.count, self.count & 1
        main = torch.cuda.current_stream(eng.device)
        if i == 0:
            self.in_stream.wait_stream(main)
        with torch.cuda.stream(self.in_stream):
            if i >= 2:
                self.in_stream.wait_event(ev["consumed"][b])     # step i-2 has taken its inputs out of this staging set
            for name in self.names:
                self.stage[b][name].copy_(host_inputs[name], non_blocking=True)
            ev["loaded"][b].record(self.in_stream)
        main.wait_event(ev["loaded"][b])
        for name in self.names:
            eng[name].copy_(self.stage[b][name], non_blocking=True)
        ev["consumed"][b].record(main)
        self.planner.step(1)
        if i >= 2:
            main.wait_event(ev["drained"][b])                    # the read-back of step i-2 has left this device buffer
        eng.paths(out=self.d_paths[b])
        ev["ready"][b].record(main)
        with torch.cuda.stream(self.out_stream):
            self.out_stream.wait_event(ev["ready"][b])
            self.h_paths[b].copy_(self.d_paths[b], non_blocking=True)
            ev["drained"][b].record(self.out_stream)
        self.count += 1
        return self.h_paths[b]

    def synchronize(self):
        eng = self.planner.engine
        main = torch.cuda.current_stream(eng.device)
        main.wait_stream(self.out_stream)
        main.synchronize()


def corridor_config(problems: int, seed=0, **overrides) -> EngineConfig:
    """BASELINE config 2: scripts/run_planner.py hyper-parameters (:14-55), forward-only
    (direction_delta_weight=100, scripts/run_bench_mr.py:60), disc robot r=0.3 in jittered S-bend corridors."""
    cfg = EngineConfig(
        problems=problems, constrained=True, trajectory_length=100,
        onf=OnfSpec(mean=0.0, sigma=1.0, use_cos=True, angle=True, hidden=100, layers=2),
        onf_lr=5e-2, onf_betas=(0.9, 0.9), traj_lr=1e-2, traj_betas=(0.9, 0.9),
        trajectory_random_offset=0.02, collision_weight=1.0, velocity_hessian_weight=0.5, random_field_points=10,
        init_collision_iteration=0, constraint_deltas_weight=20.0, multipliers_lr=0.1, init_collision_points=100,
        reparametrize_trajectory_freq=10, optimize_collision_model_freq=1, angle_weight=0.5, angle_offset=0.3,
        boundary_weight=1.0, collision_multipliers_lr=1e-3, direction_delta_weight=100.0,
        checker="circle", env_shared=False, max_obstacles=20, robot_radius=0.3, checker_bounds=(0.0, 3.0, 0.0, 3.0),
        border=(-0.1, 3.1, -0.1, 3.1), rng_mode="philox", seed=seed)
    return replace(cfg, **overrides)


def forest_config(problems: int, rows=100, cols=100, cell=0.03, seed=0, **overrides) -> EngineConfig:
    """BASELINE config 3: random-obstacle worlds (forest-style occupancy grids, one per problem, obstacle_ratio 0.03,
    notebooks/benchmark/forest_experiment.ipynb:98,424), point robot labelled by the notebook's grid checker
    (onf_planner_image_map.ipynb:65-80), reference ONF and scripts/run_planner.py hyper-parameters."""
    w, h = cols * cell, rows * cell
    cfg = corridor_config(problems, seed=seed, checker="grid", env_shared=False, grid_shape=(rows, cols),
                          grid_origin=(0.0, 0.0), grid_cell=cell, checker_bounds=(0.0, w, 0.0, h),
                          border=(-0.1, w + 0.1, -0.1, h + 0.1), direction_delta_weight=0.0)
    return replace(cfg, **overrides)


def gridmap_wide_config(problems: int, rows=1000, cols=1000, cell=0.003, seed=0, hidden=256, layers=4,
                        **overrides) -> EngineConfig:
    """BASELINE config 4: ONE synthetic 1000x1000 occupancy grid (bench-mr style, obstacle ratio 0.1, seed 1) shared
    by all problems, wide ONF (4x256 hidden, 200 Fourier + 20 heading features) -> tcgen05 grouped-GEMM path.
    The map is scaled into the reference's 3 m x 3 m test world (cell = 3 mm) so that the reference's
    hyper-parameters (offsets 1.5 / 0.02 m, sigma = 1) keep their meaning."""
    w, h = cols * cell, rows * cell
    cfg = corridor_config(problems, seed=seed, checker="grid", env_shared=True, grid_shape=(rows, cols),
                          grid_origin=(0.0, 0.0), grid_cell=cell, checker_bounds=(0.0, w, 0.0, h),
                          border=(-0.1, w + 0.1, -0.1, h + 0.1), direction_delta_weight=0.0,
                          onf=OnfSpec(mean=0.0, sigma=1.0, use_cos=True, angle=True, hidden=hidden, layers=layers))
    return replace(cfg, **overrides)

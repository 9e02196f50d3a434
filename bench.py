#!/usr/bin/env python
"""bench.py - NFOPP optimisation inner loop on B200: planner problem-steps/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--problems P]

Workload (BASELINE.json configs[1]): non-holonomic forward-only ConstrainedNERFOptPlanner problems in jittered
S-bend corridors (scripts/run_planner.py hyper-parameters, direction_delta_weight=100), 1024 problems per GPU,
each with its own ONF weights, trajectory, obstacle set and RNG stream.  A "step" is one planner.step() of all
problems.  N>1 (torchrun): problems are sharded, 1024 per GPU (weak scaling), no collective in the loop.

JSON line: `value` = problem-steps/s with all state resident in HBM (CUDA-graph replay); `e2e` = the same metric
through the host-facing API with HOST buffers: every step copies that step's environments (obstacle sets, start,
goal) from pinned host memory and reads the paths back (the read-back of step k on a copy stream, under step k+1);
`roofline` = the dominant kernel (ONF train fwd+bwd)
timed live with CUDA events; `cpu_baseline` = the CPU oracle (restatement of the reference) on the host cores.
"""
import argparse
import ctypes as C
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "planner problem-steps/s"
UNIT = "problem-steps/s"
PROBLEMS_PER_GPU = 1024


WORKLOADS = ("corridor", "forest", "gridmap-wide")


def workload_config(problems, onf_impl="auto", workload="corridor"):
    tensor = onf_impl != "fma"
    if workload == "corridor":
        cfg = {"workload": "configs[1]: corridor, non-holonomic forward-only ConstrainedNERFOptPlanner",
               "onf": "F_in=220 (200 Fourier + 20 angle), 2x100 hidden",
               "checker": "disc r=0.3 vs 20 obstacle points per problem"}
        impl = "tensor (fused tcgen05 kernel, bf16 hi/lo split operands, fp32 TMEM accumulators, fused Adam)"
        ws = "ONF w/m/v fp32 + bf16 operand copy, ~0.55 GB at 1024 problems"
    elif workload == "forest":
        cfg = {"workload": "configs[2]: random-obstacle worlds (100x100 occupancy grid per problem, obstacle ratio 0.03), "
                           "ConstrainedNERFOptPlanner, each problem its own ONF",
               "onf": "F_in=220 (200 Fourier + 20 angle), 2x100 hidden",
               "checker": "point robot vs the problem's own 100x100 occupancy grid (notebook MapCollisionChecker)"}
        impl = "tensor (fused tcgen05 kernel, bf16 hi/lo split operands, fp32 TMEM accumulators, fused Adam)"
        ws = "ONF w/m/v fp32 + bf16 operand copy, ~0.55 GB at 1024 problems"
    else:
        cfg = {"workload": "configs[3]: one synthetic 1000x1000 occupancy grid shared by all problems, wide ONF",
               "onf": "F_in=220 (200 Fourier + 20 angle), 4x256 hidden",
               "checker": "point robot vs one shared 1000x1000 occupancy grid (obstacle ratio 0.1, seed 1)"}
        impl = "wide (tcgen05 grouped GEMM per layer, streamed bf16 hi/lo split weights, fp32 TMEM accumulators)"
        ws = "ONF w/m/v/grad fp32 + operand copies (5 MB per problem) + activation tiles"
    cfg.update({"problems_per_gpu": problems, "trajectory_length": 100, "samples_per_step": 209,
                "rng": "on-device Philox4x32-10", "onf_impl": impl if tensor else "fma (CUDA-core fp32)",
                "launch": "one CUDA graph per step" + ("" if workload == "gridmap-wide" else
                                                       ", two lanes (independent problem ranges on parallel streams inside the graph)"),
                "l2": f"per-step working set ({ws}) exceeds the 126 MB L2",
                "e2e": "per step: inputs pinned host -> device staging (copy stream) -> engine buffers, planner.step, paths -> "
                       "pinned host (copy stream); the copies of neighbouring steps run under the step's kernels"})
    return cfg


def build_workload(workload, count, lo):
    """-> (EngineConfig, init kwargs, {engine buffer name: host array} copied H2D every e2e step)."""
    import numpy as np
    from nfopp_b200.batched import corridor_config, forest_config, gridmap_wide_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    if workload == "corridor":
        starts, goals, _, obstacles = B.make_corridor_batch(count, seed=lo)
        return corridor_config(count, seed=1234 + lo), dict(obstacles=obstacles), starts, goals, {"obstacles": obstacles}
    if workload == "forest":
        starts, goals, _, grids, _, cell = B.make_random_grid_batch(count, seed=lo)
        return forest_config(count, cell=cell, seed=1234 + lo), dict(grids=grids), starts, goals, {"grid": grids}
    grid, _, cell = B.make_blob_grid()
    starts, goals = B.free_endpoints(grid, cell, count, seed=lo)
    return gridmap_wide_config(count, cell=cell, seed=1234 + lo), dict(grids=grid), starts, goals, {"grid": grid[None]}


def onf_microbench(points=1 << 18, per_problem=4096, hidden=256, layers=4, reps=5):
    """BASELINE metric part 2 (config 5, one bounded case; scripts/onf_sweep.py is the full sweep): ONF forward +
    weight-gradient + input-gradient on `points` points in weight sets of `per_problem`, wide network."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import onf_sweep
    r = onf_sweep.bench_case(hidden, layers, per_problem, points, 3, reps, torch.Generator().manual_seed(0))
    return {"mpts_per_s": r["mpts_per_s"], "tflops_algorithmic": r["tflops"], "ms": r["ms"], "points": r["points"],
            "points_per_weight_set": per_problem, "onf": f"{layers}x{hidden} hidden, F_in=220",
            "flops": "6 x MAC x points (SURVEY 8d); the bf16 hi/lo split issues 3x as many MMAs"}


# ----------------------------------------------------------------------------------------- CPU baseline
def _oracle_worker(args):
    """Runs the CPU oracle (reference restatement) on one problem of the workload for `budget` seconds."""
    seed, budget, workload = args
    import numpy as np
    import torch
    torch.set_num_threads(1)
    from oracle import checkers as OC
    from oracle import onf as O
    from oracle.planner import LiveSource, OraclePlanner, PlannerConfig
    cfg, init_kw, starts, goals, _ = build_workload(workload, 1, seed)
    ocfg = O.OnfConfig(mean=cfg.onf.mean, sigma=cfg.onf.sigma, use_cos=cfg.onf.use_cos, angle=cfg.onf.angle,
                       hidden=cfg.onf.hidden, layers=cfg.onf.layers)
    pc = PlannerConfig(constrained=True, onf=ocfg, onf_lr=cfg.onf_lr, onf_betas=cfg.onf_betas, traj_lr=cfg.traj_lr,
                       traj_betas=cfg.traj_betas, init_collision_iteration=0, angle_offset=cfg.angle_offset,
                       direction_delta_weight=cfg.direction_delta_weight)
    blob = O.init_blob(ocfg, True, torch.Generator().manual_seed(seed))
    if workload == "corridor":
        obstacles = init_kw["obstacles"]
        label = lambda p: OC.circle_labels(p, obstacles[0], cfg.robot_radius, cfg.checker_bounds)   # noqa: E731
    else:
        grid = np.asarray(init_kw["grids"])
        grid = grid[0] if grid.ndim == 3 else grid
        label = lambda p: OC.grid_labels(p, grid, 0.0, 0.0, cfg.grid_cell, cfg.checker_bounds)      # noqa: E731
    op = OraclePlanner(pc, blob, label)
    src = LiveSource(np.random.RandomState(seed), torch.Generator().manual_seed(seed))
    op.init(starts[0], goals[0], cfg.border, src)
    for _ in range(5 if workload == "gridmap-wide" else 20):
        op.step(src)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget:
        op.step(src)
        n += 1
    return n, time.perf_counter() - t0


def cpu_baseline(cores, budget, workload="corridor"):
    """problem-steps/s of the oracle with `cores` independent single-thread processes (the only way the
    reference batches)."""
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        res = pool.map(_oracle_worker, [(i, budget, workload) for i in range(cores)])
    value = sum(n / t for n, t in res)
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{cores} independent {workload} problems x {budget:.0f} s of oracle planner.step() each "
                      f"({sum(n for n, _ in res)} steps), 1 torch thread per process"}


# ----------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region: NVML every 10 ms (nvidia-smi every 200 ms as a
    fallback when pynvml is unusable)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.rows, self._stop = index, [], threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self._nvml = pynvml
            self._max = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    @staticmethod
    def _physical_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[index])
            except Exception:
                pass
        return index

    def _sample_nvml(self):
        n = self._nvml
        sm = float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM))
        r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h)) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        self.rows.append([sm, self._max] + [bool(r & bits[k]) for k in self.NAMES])

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        c = [x.strip() for x in out.strip().split(",")]
        self.rows.append([float(c[0]), float(c[1])] + [v.lower().startswith("active") for v in c[3:7]])

    def _run(self):
        while not self._stop.is_set():
            try:
                if self._nvml:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop.wait(0.01 if self._nvml else 0.2)

    def __enter__(self):
        self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({n for r in self.rows for n, v in zip(self.NAMES, r[2:6]) if v})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max((r[1] for r in self.rows), default=None),
                "reasons": reasons, "samples": len(sm), "source": "nvml" if self._nvml else "nvidia-smi"}


# ----------------------------------------------------------------------------------------- ours
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from nfopp_b200.batched import BatchedPlanner, HostStepPipeline, shard_range
    from nfopp_b200.engine import onf_segments

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P = args.problems
    lo, hi = shard_range(P * world, rank, world)            # this rank's problem ids (weak scaling)
    cfg, init_kw, starts, goals, host_env = build_workload(args.workload, hi - lo, lo)
    if args.onf_impl != "auto":
        cfg.onf_impl = args.onf_impl
    planner = BatchedPlanner(cfg, device=f"cuda:{local}")
    planner.init(starts, goals, **init_kw)
    eng = planner.engine

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(steps)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    # ---- device-resident throughput: CUDA-graph replay --------------------------------------------
    planner.step(args.warmup)
    launches0 = eng.launches
    with ClockSampler(local) as clocks:
        ms = timed(lambda k: planner.step(k), args.steps)
    launches = eng.launches - launches0
    value = P * world * args.steps / (ms * 1e-3)

    # ---- end-to-end through the host-facing API with HOST buffers ------------------------------------
    # HostStepPipeline: inputs pinned host -> staging (copy-in stream) -> engine buffers, planner.step, paths -> pinned
    # host (copy-out stream), alternating buffer sets, so the copies of neighbouring steps run under the step's kernels.
    # Every step's inputs still leave pinned host memory and every step's paths still reach it inside the timed region.
    pipe = HostStepPipeline(planner, list(host_env) + ["start", "goal"])
    h_in = {k: pipe.pinned(k, v) for k, v in dict(host_env, start=starts, goal=goals).items()}
    h2d, d2h = pipe.h2d_bytes, pipe.d2h_bytes

    def e2e_steps(k):
        for _ in range(k):
            pipe.step(h_in)            # update_obstacle_points / the map, start, goal; step(); get_path()
        pipe.synchronize()
    e2e_steps(max(3, args.warmup // 2))
    e2e_k = max(3, args.steps // 2)
    ms_e2e = timed(e2e_steps, e2e_k)
    e2e_value = P * world * e2e_k / (ms_e2e * 1e-3)

    # ---- live per-kernel timing (CUDA events on the launching stream) -> roofline of the dominant kernel ----
    prof = {}
    reps = 5
    for _ in range(reps):
        for k, v in eng.profile_step().items():
            prof[k] = prof.get(k, 0.0) + v / reps
    barrier()
    finite = bool(torch.isfinite(eng.paths()).all())
    all_paths = planner.gather_paths()                           # the one collective of the job (NCCL)

    # ---- problems SOLVED per second at the reference's stop rule (scripts/run_bench_mr.py:64-66,109-131): fresh planners,
    # <= 1000 iterations, >= 200, path evaluated on the device every 50 iterations, stop at a collision-free path that no
    # longer shortens ----
    solve = None
    if not args.no_solve:
        solver = BatchedPlanner(cfg, device=f"cuda:{local}")
        solver.init(starts, goals, **init_kw)
        res = {}
        ms_solve = timed(lambda k: res.update(solver.solve(max_iterations=1000, min_iterations=200, check_every=50)), 0)
        counts = torch.tensor([float(res["solved"].sum()), float(res["steps_run"])], device="cuda")
        if world > 1:
            dist.all_reduce(counts, op=dist.ReduceOp.SUM)
        # per-plan statistics of the solved paths (BenchmarkAdapter.cpp:68-135), outside the timed region, rank 0's shard
        quality = None
        if bool(res["solved"].any()):
            solver.engine["traj"].copy_(res["paths"][:, 1:-1])
            st = solver.path_statistics()
            sel = res["solved"]
            quality = {k: float(v[sel].mean()) for k, v in st.items()}
            quality["of"] = "mean over rank 0's solved problems"
        # SURVEY 8(d) M1, second half: problems solved per second at a FIXED 500-step budget (fresh planners, one
        # evaluation of every path at the end)
        fixed = BatchedPlanner(cfg, device=f"cuda:{local}")
        fixed.init(starts, goals, **init_kw)
        ok = {}

        def run_fixed(_):
            fixed.step(500)
            ok["free"] = ~fixed.evaluate_paths(4)[0]
        ms_fixed = timed(run_fixed, 0)
        n_fixed = torch.tensor([float(ok["free"].sum())], device="cuda")
        if world > 1:
            dist.all_reduce(n_fixed, op=dist.ReduceOp.SUM)
        fixed.engine.close()
        solve = {"problems": P * world, "solved": int(counts[0]), "seconds": ms_solve * 1e-3, "path_quality": quality,
                 "fixed_500_steps": {"solved": int(n_fixed[0]), "seconds": ms_fixed * 1e-3,
                                     "solved_per_s": float(n_fixed[0]) / (ms_fixed * 1e-3)},
                 "solved_per_s": float(counts[0]) / (ms_solve * 1e-3), "steps_run_per_gpu": float(counts[1]) / world,
                 "rule": "run_bench_mr.py: max 1000 / min 200 iterations, evaluate every 50 (4 poses per segment through the "
                         "collision checker), stop at a collision-free path that is no shorter than the best"}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = peaks.get("bf16_tflops_sustained", 1400.0)
    segs, n_params = onf_segments(cfg.onf)
    f_pos, f_in, H, L = 200, 220, cfg.onf.hidden, cfg.onf.layers
    mac = 2 * f_pos + f_in * H + (L - 1) * H * H + (H + f_in)            # MAC / point / pass
    dom = max(prof, key=prof.get)
    train_ms = prof["onf_train_fwd_bwd"]
    flops_launch = 3 * 2 * mac * 209 * (hi - lo)                           # fwd + dgrad + wgrad over 209 points
    achieved = flops_launch / (train_ms * 1e-3) / 1e12
    impl = "wide" if "onf_ws" in eng.buffers else "tensor" if "onf_wp" in eng.buffers else "fma"
    tensor = impl != "fma"
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get(
            {"tensor": "onf_train_fwd_bwd_tensor", "wide": "onf_train_fwd_bwd_wide", "fma": "onf_train_fwd_bwd"}[impl])
    except Exception:
        pass
    kernel_label = {"tensor": "onf_tc_kernel (ONF train fwd+bwd+Adam, fused tcgen05 kernel; algorithmic flops, the split "
                              "products issue 3x as many MMAs)",
                    "wide": "wide_chain_kernel + wide_wgrad2_kernel (+ encode / aux kernels): the ONF train call of the wide path "
                            "(streamed bf16 hi/lo weights, activations chained through shared memory / TMEM), one roofline entry "
                            "for the whole call; algorithmic flops, 3x as many MMAs issued; Adam + repack is a separate launch",
                    "fma": "onf_tile_kernel (ONF train fwd+bwd, FMA path)"}[impl]
    step_bytes = (24 * n_params + 15_000) * (hi - lo)                     # bytes / problem-step
    # Which roofline bounds the train launch: its arithmetic intensity against the ridge of the measured peaks.
    # Algorithmic bytes of the launch per problem: fused tensor path (Adam inside the kernel) = w, m, v read + written
    # (24 B / parameter) + the bf16 hi/lo operand copy read + rewritten + the 209 training points and labels; the other
    # paths' train call ends at the gradient = weights read + gradient written (8 B / parameter) + points and labels.
    peak_gbs = peaks.get("hbm_gbs", 6650.0)
    if impl == "tensor":
        wp = eng["onf_wp"]
        bytes_problem = 24 * n_params + 2 * (wp.numel() * wp.element_size() // (hi - lo)) + 209 * 16
    else:
        bytes_problem = 8 * n_params + 209 * 16
    bytes_launch = bytes_problem * (hi - lo)
    achieved_gbs = bytes_launch / (train_ms * 1e-3) / 1e9
    intensity, ridge = flops_launch / bytes_launch, peak_tf * 1e12 / (peak_gbs * 1e9)
    hbm_bound = tensor and intensity < ridge
    views = {"tensor": {"achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf,
                        "algorithmic_flops_per_launch": flops_launch},
             "hbm": {"achieved": achieved_gbs, "peak": peak_gbs, "unit": "GB/s", "frac": achieved_gbs / peak_gbs,
                     "algorithmic_bytes_per_launch": bytes_launch}}
    primary = views["hbm" if hbm_bound else "tensor"]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (tensor path: bf16 hi/lo split products, fp32 accumulate)" if tensor else "f32",
        "data": "synthetic", "config": workload_config(P, impl, args.workload),
        "clocks": clocks.summary(),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_k, "ms_per_step": ms_e2e / e2e_k},
        "gpu_launches": int(launches),
        "roofline": {"kernel": kernel_label, "bound": "hbm" if hbm_bound else "tensor",
                     "achieved": primary["achieved"], "peak": primary["peak"], "unit": primary["unit"], "frac": primary["frac"],
                     "traffic": traffic,
                     "peak_source": ("MEASURED_PEAKS.json " + ("hbm_gbs" if hbm_bound else "bf16_tflops_sustained") + " (of measured)")
                     if peaks else "fallback",
                     "why": f"arithmetic intensity of the launch {intensity:.0f} flop/B vs ridge {ridge:.0f} flop/B "
                            f"(measured peaks) -> {'left' if hbm_bound else 'right'} of the ridge; both views below",
                     "views": views,
                     "ms_per_launch": train_ms, "dominant_kernel": dom,
                     "whole_step_hbm": {"algorithmic_bytes": step_bytes,
                                        "achieved_gbs": step_bytes / (ms / args.steps * 1e-3) / 1e9,
                                        "peak_gbs": peaks.get("hbm_gbs", 6650.0)}},
        "kernel_ms": prof,
        "paths_finite": finite, "paths_gathered": list(all_paths.shape),
    }
    if solve:
        line["solve"] = solve
    if world == 1 and not args.no_onf_microbench:
        mb = onf_microbench()
        mb["frac_of_bf16_peak"] = mb["tflops_algorithmic"] / peaks.get("bf16_tflops", 1590.0)
        mb["peak"] = peaks.get("bf16_tflops", 1590.0)
        line["onf_fwd_bwd"] = mb
    if world == 1 and not args.no_cpu_baseline:
        cores = max(1, min(os.cpu_count() or 1, 32))
        line["cpu_baseline"] = cpu_baseline(cores, args.cpu_seconds, args.workload)
    emit(line)
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path: /root/reference cannot travel to the GPU box, so
    this times the oracle port (pinned in lock-step against the live reference, tests/golden) on all host
    cores, same config and metric; each step = a bounded sample (one planner.step() per process)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cores = max(1, os.cpu_count() or 1)
    budget = max(2.0, min(30.0, args.cpu_seconds))
    base = cpu_baseline(cores, budget, args.workload)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT,
            "n_gpus": int(os.environ.get("WORLD_SIZE", args.gpus)), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * cores / base["value"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.problems, "auto", args.workload),
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


_OUT = sys.stdout


def emit(line):
    """The ONE JSON line of the contract, on the process's real stdout."""
    _OUT.write(json.dumps(line) + "\n")
    _OUT.flush()


def main():
    global _OUT
    # libraries (NCCL prints "NCCL version ..." on stdout) must not pollute the one-JSON-line contract: keep a private
    # handle on the real stdout and point fd 1 at stderr for everything else
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = sys.stderr
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--problems", type=int, default=PROBLEMS_PER_GPU, help="problems per GPU")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-onf-microbench", action="store_true")
    ap.add_argument("--no-solve", action="store_true", help="skip the problems-solved-per-second run")
    ap.add_argument("--workload", default="corridor", choices=list(WORKLOADS),
                    help="corridor = BASELINE configs[1] (the headline), forest = configs[2], gridmap-wide = configs[3]")
    ap.add_argument("--onf-impl", default="auto", choices=["auto", "fma", "tensor", "wide"],
                    help="ONF kernels: tcgen05 tensor path (default for the reference network shape) or CUDA-core FMA")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

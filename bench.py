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
`roofline` = SURVEY 8(d): the whole step and the dominant kernel (ONF train fwd+bwd+Adam) against the measured HBM peak,
algorithmic bytes only (24 B per ONF parameter + the training points), timed live with CUDA events; implementation
bytes (the fp16 operand copy) and the ncu DRAM traffic are reported beside it, never inside `frac`.
`cpu_baseline` = the CPU oracle (restatement of the reference) on the host cores, all cores and one thread.
Beside the headline (same line, each bounded to seconds): `configs2_forest_16k` = BASELINE configs[2], 16 384 random-obstacle
problems SHARDED over the N GPUs (strong scaling); `configs3_gridmap_wide` = configs[3] (N = 1 only); `streams` = achieved
GB/s of the point / trajectory kernels and of the stand-alone labelling kernel on 2^24 poses; `single_problem` = configs[0]'s
use: one problem through the drop-in planner API, `step(); get_path()` per iteration.
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
TENSOR_IMPL = ("tensor (persistent fused tcgen05 kernel: one CTA per SM walks its problems, 112-point tiles, streamed fp16 hi/lo "
               "weights, every product 3 MMA passes (hi.hi + lo.hi + hi.lo), static power-of-two gradient scale, fp32 TMEM "
               "accumulators, Adam + operand-copy refresh of the previous problem on six optimiser warps under the next one's tiles)")
FOREST_TOTAL = 16384            # BASELINE configs[2]: problems of the whole job, sharded over the GPUs


def workload_config(problems, onf_impl="auto", workload="corridor"):
    tensor = onf_impl != "fma"
    if workload == "corridor":
        cfg = {"workload": "configs[1]: corridor, non-holonomic forward-only ConstrainedNERFOptPlanner",
               "onf": "F_in=220 (200 Fourier + 20 angle), 2x100 hidden",
               "checker": "disc r=0.3 vs 20 obstacle points per problem"}
        impl = TENSOR_IMPL
        ws = "ONF w/m/v fp32 + fp16 operand copy, ~0.55 GB at 1024 problems"
    elif workload == "forest":
        cfg = {"workload": "configs[2]: random-obstacle worlds (100x100 occupancy grid per problem, obstacle ratio 0.03), "
                           "ConstrainedNERFOptPlanner, each problem its own ONF",
               "onf": "F_in=220 (200 Fourier + 20 angle), 2x100 hidden",
               "checker": "point robot vs the problem's own 100x100 occupancy grid (notebook MapCollisionChecker)"}
        impl = TENSOR_IMPL
        ws = "ONF w/m/v fp32 + fp16 operand copy, ~0.55 GB at 1024 problems"
    else:
        cfg = {"workload": "configs[3]: one synthetic 1000x1000 occupancy grid shared by all problems, wide ONF",
               "onf": "F_in=220 (200 Fourier + 20 angle), 4x256 hidden",
               "checker": "point robot vs one shared 1000x1000 occupancy grid (obstacle ratio 0.1, seed 1)"}
        impl = "wide (tcgen05 grouped GEMM per layer, streamed bf16 hi/lo split weights, fp32 TMEM accumulators)"
        ws = "ONF w/m/v/grad fp32 + operand copies (5 MB per problem) + activation tiles"
    cfg.update({"problems_per_gpu": problems, "trajectory_length": 100, "samples_per_step": 209,
                "rng": "on-device Philox4x32-10", "onf_impl": impl if tensor else "fma (CUDA-core fp32)",
                "launch": "one CUDA graph per step" + ("" if workload == "gridmap-wide" else
                                                       ", up to four lanes (independent problem ranges on parallel streams inside the graph, "
                                                       "each lane's persistent kernels on #SMs / lanes CTAs; BatchedPlanner picks the count)"),
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
            "flops": "6 x MAC x points (SURVEY 8d)"}


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


def host_cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return f"{ln.split(':', 1)[1].strip()} ({os.cpu_count()} logical cores)"
    except Exception:
        pass
    return f"unknown ({os.cpu_count()} logical cores)"


def cpu_baseline(cores, budget, workload="corridor"):
    """problem-steps/s of the oracle with `cores` independent single-thread processes (the only way the
    reference batches)."""
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        res = pool.map(_oracle_worker, [(i, budget, workload) for i in range(cores)])
    value = sum(n / t for n, t in res)
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "host_cpu": host_cpu_model(),
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
def kernel_sources_sha():
    """sha256 over the ONF kernel sources: an ncu traffic figure in profiles/roofline_traffic.json is only attached to a
    bench line when it was captured from exactly these sources (there is no .git on the GPU box to ask)."""
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "pytorch-motion-planner_b200", "csrc")
    for name in sorted(os.listdir(csrc)):
        if name.startswith("onf") or name in ("tc05.cuh",):
            h.update(name.encode())
            h.update(open(os.path.join(csrc, name), "rb").read())
    return h.hexdigest()[:16]


def load_traffic(key):
    """-> (bytes per launch or None, note)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
    except Exception as e:  # pragma: no cover
        return None, f"profiles/roofline_traffic.json unreadable: {e}"
    ent = t.get(key)
    if not isinstance(ent, dict):
        return None, f"no ncu capture recorded for '{key}'"
    sha = kernel_sources_sha()
    if ent.get("sources_sha") != sha:
        return None, (f"stale: captured from kernel sources {ent.get('sources_sha')} (commit {ent.get('commit')}), "
                      f"this run is {sha}")
    return float(ent["bytes"]), f"{ent.get('source')} (commit {ent.get('commit')}, launch shape {ent.get('launch_shape')})"


def peaks_file():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


class Dist:
    """rank / world plumbing: one process per GPU, NCCL only for the barrier, the max over ranks and the final gather."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", 0))
        self.world = int(os.environ.get("WORLD_SIZE", 1))
        self.local = int(os.environ.get("LOCAL_RANK", 0))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps):
        """CUDA events around fn(steps), bracketed by barrier + synchronize on both sides, max over ranks -> ms."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(steps)
        e1.record()
        self.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if self.world > 1:
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms)

    def sum(self, *vals):
        t = self.torch.tensor([float(v) for v in vals], device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [float(v) for v in t]

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def measure_planner(D, workload, count, lo, total, steps, warmup, onf_impl="auto", e2e_steps=None, profile_reps=5,
                    with_clocks=False):
    """Builds `count` problems of `workload` (global ids lo..), times `steps` graph-replayed steps (device-resident state),
    the same through HostStepPipeline with HOST buffers, and one CUDA-event profile of the step's kernels.
    `total` = problems of the whole job (all ranks)."""
    import torch
    from nfopp_b200.batched import BatchedPlanner, HostStepPipeline
    cfg, init_kw, starts, goals, host_env = build_workload(workload, count, lo)
    if onf_impl != "auto":
        cfg.onf_impl = onf_impl
    planner = BatchedPlanner(cfg, device=f"cuda:{D.local}")
    planner.init(starts, goals, **init_kw)
    eng = planner.engine
    planner.step(warmup)
    launches0 = eng.launches
    if with_clocks:
        with ClockSampler(D.local) as clocks:
            ms = D.timed(lambda k: planner.step(k), steps)
        clock_summary = clocks.summary()
    else:
        ms = D.timed(lambda k: planner.step(k), steps)
        clock_summary = None
    launches = eng.launches - launches0
    out = {"planner": planner, "cfg": cfg, "init_kw": init_kw, "starts": starts, "goals": goals, "ms": ms,
           "ms_per_step": ms / steps, "value": total * steps / (ms * 1e-3), "launches": int(launches),
           "clocks": clock_summary, "steps": steps, "lanes": int(planner.lanes)}
    # ---- end-to-end through the host-facing API with HOST buffers: inputs pinned host -> staging (copy-in stream) ->
    # engine buffers, planner.step, paths -> pinned host (copy-out stream), alternating buffer sets.  Every step's inputs
    # leave pinned host memory and every step's paths reach it inside the timed region.
    if e2e_steps:
        pipe = HostStepPipeline(planner, list(host_env) + ["start", "goal"])
        h_in = {k: pipe.pinned(k, v) for k, v in dict(host_env, start=starts, goal=goals).items()}

        def run(k):
            for _ in range(k):
                pipe.step(h_in)            # update_obstacle_points / the map, start, goal; step(); get_path()
            pipe.synchronize()
        run(3)
        ms_e2e = D.timed(run, e2e_steps)
        out["e2e"] = {"value": total * e2e_steps / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
                      "d2h_bytes_per_step": pipe.d2h_bytes, "steps": e2e_steps, "ms_per_step": ms_e2e / e2e_steps}
        del pipe, h_in
    prof = {}
    for _ in range(profile_reps):
        for k, v in eng.profile_step().items():
            prof[k] = prof.get(k, 0.0) + v / profile_reps
    D.barrier()
    out["prof"] = prof
    out["finite"] = bool(torch.isfinite(eng.paths()).all())
    return out


def roofline_block(m, count, peaks):
    """SURVEY 8(d) / BASELINE.md 4: algorithmic bytes per problem-step = 24 B per ONF parameter (w, m, v read + written) +
    ~15 KB of trajectory / replay-buffer / environment state; the train launch alone = 24 B per parameter + its training
    points (209 x (3 coordinates + label) x 4 B).  Nothing the implementation adds (operand copies, activation tiles) is
    counted as algorithmic: it is listed as implementation_bytes, and the ncu DRAM traffic beside both."""
    from nfopp_b200.engine import onf_segments
    cfg, eng, prof = m["cfg"], m["planner"].engine, m["prof"]
    _, n_params = onf_segments(cfg.onf)
    impl = "wide" if "onf_ws" in eng.buffers else "tensor" if "onf_wp" in eng.buffers else "fma"
    peak_gbs = peaks.get("hbm_gbs", 6650.0)
    peak_tf = peaks.get("bf16_tflops_sustained", 1400.0)
    f_pos, f_in, H, L = 200 if cfg.onf.use_cos else 100, (200 if cfg.onf.use_cos else 100) + (20 if cfg.onf.angle else 0), \
        cfg.onf.hidden, cfg.onf.layers
    mac = 2 * f_pos + f_in * H + (L - 1) * H * H + (H + f_in)            # MAC / point / pass
    n_train = (cfg.trajectory_length - 1) + cfg.collision_point_count + cfg.random_field_points
    train_ms = prof["onf_train_fwd_bwd"] + (prof.get("onf_reduce_adam", 0.0) if impl != "tensor" else 0.0)
    step_bytes = (24 * n_params + 15_000) * count
    train_bytes = (24 * n_params + n_train * 16) * count
    wp = eng.buffers.get("onf_wp")
    packed = (wp.numel() * wp.element_size() // count) if wp is not None else 0
    # what the implementation moves on top (per problem): tensor = operand copy read by the 3 ONF launches + rewritten
    # by the fused Adam; wide = both packed layouts re-read per launch + rewritten, plus the wgrad operand tiles
    impl_bytes = {"tensor": 4 * packed, "wide": 4 * packed, "fma": 0}[impl] * count
    key = {"tensor": "onf_train_fwd_bwd_tensor", "wide": "onf_train_fwd_bwd_wide", "fma": "onf_train_fwd_bwd"}[impl]
    traffic, traffic_note = load_traffic(key)
    flops_launch = 3 * 2 * mac * n_train * count
    ms_step = m["ms_per_step"]
    step_gbs = step_bytes / (ms_step * 1e-3) / 1e9
    train_gbs = train_bytes / (train_ms * 1e-3) / 1e9
    kernel = {"tensor": "onf_tcs_kernel (ONF train: encode, fwd, BCE, dgrad, wgrad, Adam, operand-copy refresh in ONE launch)",
              "wide": "wide_chain_kernel + wide_wgrad2_kernel + wide_pack_kernel (ONF train call of the wide path incl. Adam)",
              "fma": "onf_tile_kernel + onf_adam_kernel (FMA path)"}[impl]
    return {
        "bound": "hbm", "unit": "GB/s", "peak": peak_gbs,
        "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)",
        # the headline fraction: the dominant kernel's algorithmic bytes / its CUDA-event time / the measured peak
        "kernel": kernel, "achieved": train_gbs, "frac": train_gbs / peak_gbs,
        "algorithmic_bytes_per_launch": train_bytes, "ms_per_launch": train_ms,
        "bytes_per_unit": f"24 x {n_params} parameters + {n_train} points x 16 B per problem",
        "whole_step": {"algorithmic_bytes": step_bytes, "achieved": step_gbs, "frac": step_gbs / peak_gbs,
                       "ms_per_step": ms_step, "bytes_per_unit": f"24 x {n_params} + 15000 B per problem-step (SURVEY 8d)",
                       "ceiling_problem_steps_per_s": peak_gbs * 1e9 / (24 * n_params + 15_000)},
        "implementation_bytes": impl_bytes,
        "implementation_bytes_note": "operand copy of the hidden-layer weights: read by the step's three ONF launches, rewritten by "
                                     "the fused Adam; NOT part of achieved / frac",
        "traffic": traffic, "traffic_note": traffic_note,
        "traffic_ratio": (traffic / train_bytes) if traffic else None,
        "tensor_view": {"achieved": flops_launch / (train_ms * 1e-3) / 1e12, "peak": peak_tf, "unit": "TFLOP/s",
                        "frac": flops_launch / (train_ms * 1e-3) / 1e12 / peak_tf,
                        "algorithmic_flops_per_launch": flops_launch,
                        "note": "6 x MAC x points; every product issues 3 MMAs (fp16 hi/lo split), not counted here"},
        "dominant_kernel": max(prof, key=prof.get),
        "kernel_share_of_step": prof["onf_train_fwd_bwd"] / max(sum(prof.values()), 1e-9),
    }


def stream_block(m, count, peaks):
    """Achieved GB/s of the point and trajectory kernels of a step (north_star: 'achieved HBM GB/s for the point and
    trajectory streams'): algorithmic bytes = the engine buffers each kernel reads / writes once, per problem."""
    eng, prof, cfg = m["planner"].engine, m["prof"], m["cfg"]
    peak = peaks.get("hbm_gbs", 6650.0)

    def nbytes(*names):
        return sum(eng.buffers[n].numel() * eng.buffers[n].element_size() for n in names if n in eng.buffers)
    env = nbytes("obstacles", "obstacle_n") if cfg.checker != "grid" else count * 209          # one grid byte per labelled pose
    kernels = {
        "sample_pre": nbytes("prev_traj", "traj", "buf_pos", "buf_age", "buf_n", "cand_pos", "cand_age", "cand_x", "cand_n",
                             "train_pos", "train_x") ,
        "resample_label": nbytes("cand_z", "cand_pos", "cand_age", "buf_pos", "buf_age", "buf_n", "train_pos", "train_x",
                                 "train_y", "train_n", "choice_idx") + env,
        "traj_pre": nbytes("traj", "mu", "path_q", "path_mu", "path_t", "path_n"),
        "traj_post": 2 * nbytes("traj", "traj_m", "traj_v", "lam", "mu") + nbytes("path_gq", "path_z", "path_t", "start", "goal"),
    }
    out = {}
    for k, b in kernels.items():
        ms = prof.get(k)
        if ms:
            out[k] = {"ms": ms, "algorithmic_bytes": b, "gbs": b / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": b / (ms * 1e-3) / 1e9 / peak}
    return out


def label_stream_bench(peaks, n_poses=1 << 24):
    """Stand-alone K1: nfopp_label_circle / nfopp_label_grid on 2^24 poses (24 B pose in, 1 B label out)."""
    import numpy as np
    import torch
    from nfopp_b200 import _lib as L
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    peak = peaks.get("hbm_gbs", 6650.0)
    gen = torch.Generator(device="cuda").manual_seed(0)
    poses = torch.rand(n_poses, 3, generator=gen, device="cuda", dtype=torch.float64) * 3.0
    labels = torch.empty(n_poses, dtype=torch.uint8, device="cuda")
    obs = torch.as_tensor(np.asarray(B._obstacle_points(), dtype=np.float64)).cuda()
    grid_np, _, cell = B.make_blob_grid()
    grid = torch.as_tensor((np.asarray(grid_np) > 0).astype(np.uint8)).cuda()
    bounds = (C.c_double * 4)(0.0, 3.0, 0.0, 3.0)
    s = L.stream_ptr()
    calls = {
        "circle_20_obstacles": lambda: L.lib().nfopp_label_circle(L.ptr(poses), 1, n_poses, 3, L.ptr(obs), None, 1, obs.shape[0],
                                                                  0.3, bounds, L.ptr(labels), s),
        "grid_1000x1000": lambda: L.lib().nfopp_label_grid(L.ptr(poses), 1, n_poses, 3, L.ptr(grid), 1, grid.shape[0], grid.shape[1],
                                                           0.0, 0.0, float(cell), bounds, L.ptr(labels), s),
    }
    out = {"poses": n_poses, "bytes_per_pose": 25, "note": "24 B float64 pose read + 1 B label written; 400 MB per call (> L2)"}
    for name, fn in calls.items():
        for _ in range(3):
            L.check(fn())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            L.check(fn())
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        gbs = n_poses * 25 / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "gbs": gbs, "frac_of_hbm_peak": gbs / peak, "mposes_per_s": n_poses / ms / 1e3}
    return out


def single_problem_bench(iterations=300):
    """BASELINE configs[0] as the reference is actually used (scripts/run_planner.py:76-78, the ROS adapter's 10 Hz loop): ONE
    problem through the drop-in API, PlannerFactory.make_constrained_onf_planner(...), `step(); get_path()` per iteration."""
    import torch
    from nfopp_b200.collision_checker import RectangleCollisionChecker
    from nfopp_b200.planner_factory import PlannerFactory
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder
    env = TestEnvironmentBuilder().make_car_environment()
    checker = RectangleCollisionChecker((-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))
    checker.update_obstacle_points(env.obstacle_points)
    planner = PlannerFactory.make_constrained_onf_planner(checker)
    planner.init(env.start_point, env.goal_point, env.bounds)
    for _ in range(30):
        planner.step()
        planner.get_path()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iterations):
        planner.step()
        path = planner.get_path()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"steps_per_s": iterations / dt, "ms_per_step": 1e3 * dt / iterations, "iterations": iterations,
            "what": "scripts/run_planner.py config (car environment, rectangle checker, ONF 33 161 parameters, T = 100), one problem, "
                    "step() + get_path() (device -> host) per iteration, wall clock",
            "path_shape": list(path.shape),
            "reference_cpu": "BASELINE.md 2: 156 steps/s (1 thread, anomaly off, warm) / 32.7 steps/s as shipped (anomaly on), "
                             "measured in the build container; see profiles/r02_reference_cpu_container.json"}


def run_ours(args):
    import torch
    from nfopp_b200.batched import BatchedPlanner, shard_range

    D = Dist()
    rank, world, local = D.rank, D.world, D.local
    peaks = peaks_file()
    P = args.problems
    lo, hi = shard_range(P * world, rank, world)            # this rank's problem ids (weak scaling)
    m = measure_planner(D, args.workload, hi - lo, lo, P * world, args.steps, args.warmup, args.onf_impl,
                        e2e_steps=max(3, args.steps // 2), with_clocks=True)
    planner, cfg, init_kw, starts, goals = m["planner"], m["cfg"], m["init_kw"], m["starts"], m["goals"]
    eng = planner.engine
    all_paths = planner.gather_paths()                           # the one collective of the job (NCCL)
    roof = roofline_block(m, hi - lo, peaks)
    streams = stream_block(m, hi - lo, peaks)
    impl = "wide" if "onf_ws" in eng.buffers else "tensor" if "onf_wp" in eng.buffers else "fma"

    # ---- problems SOLVED per second at the reference's stop rule (scripts/run_bench_mr.py:64-66,109-131): fresh planners,
    # <= 1000 iterations, >= 200, path evaluated on the device every 50 iterations, stop at a collision-free path that no
    # longer shortens ----
    solve = None
    if not args.no_solve:
        solver = BatchedPlanner(cfg, device=f"cuda:{local}")
        solver.init(starts, goals, **init_kw)
        res = {}
        ms_solve = D.timed(lambda k: res.update(solver.solve(max_iterations=1000, min_iterations=200, check_every=50)), 0)
        n_solved, steps_run = D.sum(res["solved"].sum(), res["steps_run"])
        # per-plan statistics of the solved paths (BenchmarkAdapter.cpp:68-135), outside the timed region, rank 0's shard
        quality = None
        if bool(res["solved"].any()):
            solver.engine["traj"].copy_(res["paths"][:, 1:-1])
            st = solver.path_statistics()
            sel = res["solved"]
            quality = {k: float(v[sel].mean()) for k, v in st.items()}
            quality["of"] = "mean over rank 0's solved problems (raw T+2 nodes; bench-mr's interpolated path is unavailable)"
        solver.engine.close()
        del solver
        # SURVEY 8(d) M1, second half: problems solved per second at a FIXED 500-step budget (fresh planners, one
        # evaluation of every path at the end)
        fixed = BatchedPlanner(cfg, device=f"cuda:{local}")
        fixed.init(starts, goals, **init_kw)
        ok = {}

        def run_fixed(_):
            fixed.step(500)
            ok["free"] = ~fixed.evaluate_paths(4)[0]
        ms_fixed = D.timed(run_fixed, 0)
        n_fixed, = D.sum(ok["free"].sum())
        fixed.engine.close()
        del fixed
        solve = {"problems": P * world, "solved": int(n_solved), "seconds": ms_solve * 1e-3, "path_quality": quality,
                 "fixed_500_steps": {"solved": int(n_fixed), "seconds": ms_fixed * 1e-3,
                                     "solved_per_s": float(n_fixed) / (ms_fixed * 1e-3)},
                 "solved_per_s": float(n_solved) / (ms_solve * 1e-3), "steps_run_per_gpu": float(steps_run) / world,
                 "rule": "run_bench_mr.py: max 1000 / min 200 iterations, evaluate every 50 (4 poses per segment through the "
                         "collision checker), stop at a collision-free path that is no shorter than the best"}
    finite = m["finite"]
    eng.close()
    del planner, eng, m["planner"]
    torch.cuda.empty_cache()

    # ---- BASELINE configs[2] at its own size: 16 384 forest problems SHARDED over the N GPUs (strong scaling) ----
    forest = None
    if not args.no_extra and args.workload == "corridor":
        flo, fhi = shard_range(FOREST_TOTAL, rank, world)
        fm = measure_planner(D, "forest", fhi - flo, flo, FOREST_TOTAL, max(5, min(args.steps, 20)), 3, e2e_steps=5, profile_reps=2)
        froof = roofline_block(fm, fhi - flo, peaks)
        forest = {"workload": "configs[2]: 16 384 random-obstacle problems (own 100x100 grid and own ONF each), sharded "
                              f"{FOREST_TOTAL}/N per GPU, no collective in the loop", "problems_total": FOREST_TOTAL,
                  "problems_per_gpu": fhi - flo, "scaling": "strong", "value": fm["value"], "unit": UNIT,
                  "ms_per_step": fm["ms_per_step"], "steps": fm["steps"], "e2e": fm["e2e"], "gpu_launches": fm["launches"],
                  "paths_finite": fm["finite"],
                  "roofline": {"whole_step_frac": froof["whole_step"]["frac"], "train_launch_frac": froof["frac"],
                               "train_ms": froof["ms_per_launch"]}}
        fm["planner"].engine.close()
        del fm
        torch.cuda.empty_cache()

    if rank != 0:
        D.close()
        return
    line = {
        "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (tensor path: fp16 hi/lo split products, 3 MMA passes each, fp32 accumulate)" if impl != "fma" else "f32",
        "data": "synthetic", "config": dict(workload_config(P, impl, args.workload), lanes=m["lanes"]),
        "clocks": m["clocks"], "e2e": m["e2e"], "gpu_launches": m["launches"],
        "roofline": roof, "kernel_ms": m["prof"], "streams": streams,
        "paths_finite": finite, "paths_gathered": list(all_paths.shape),
    }
    if solve:
        line["solve"] = solve
    if forest:
        line["configs2_forest_16k"] = forest
    if world == 1 and not args.no_extra and args.workload == "corridor":
        # ---- BASELINE configs[3]: one shared 1000x1000 grid, 4x256 ONF (wide tcgen05 path) ----
        wm = measure_planner(D, "gridmap-wide", P, 0, P, max(5, min(args.steps, 20)), 3, e2e_steps=5, profile_reps=3)
        wroof = roofline_block(wm, P, peaks)
        line["configs3_gridmap_wide"] = {
            "workload": workload_config(P, "wide", "gridmap-wide")["workload"], "problems": P, "value": wm["value"], "unit": UNIT,
            "ms_per_step": wm["ms_per_step"], "steps": wm["steps"], "e2e": wm["e2e"], "gpu_launches": wm["launches"],
            "paths_finite": wm["finite"], "roofline": wroof, "kernel_ms": wm["prof"]}
        wm["planner"].engine.close()
        del wm
        torch.cuda.empty_cache()
        line["streams"]["label_kernel"] = label_stream_bench(peaks)
        line["single_problem"] = single_problem_bench()
    if world == 1 and not args.no_onf_microbench:
        mb = onf_microbench()
        mb["frac_of_bf16_peak"] = mb["tflops_algorithmic"] / peaks.get("bf16_tflops", 1590.0)
        mb["peak"] = peaks.get("bf16_tflops", 1590.0)
        line["onf_fwd_bwd"] = mb
    if world == 1 and not args.no_cpu_baseline:
        cores = max(1, min(os.cpu_count() or 1, 32))
        line["cpu_baseline"] = cpu_baseline(cores, args.cpu_seconds, args.workload)
        line["cpu_baseline_1thread"] = cpu_baseline(1, min(args.cpu_seconds, 8.0), args.workload)
    emit(line)
    D.close()


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
    one = cpu_baseline(1, min(budget, 8.0), args.workload)
    wl = workload_config(args.problems, "auto", args.workload)
    config = {"workload": wl["workload"], "onf": wl["onf"], "checker": wl["checker"], "trajectory_length": 100,
              "samples_per_step": 209, "problems": f"{cores} at a time, one per process (the only way the reference batches)",
              "implementation": "oracle/planner.py: CPU restatement of neural_field_optimal_planner's step() (numpy float64 sampling / "
                                "labelling, torch float32 ONF with analytic gradients, Adam), pinned in lock-step against the "
                                "unmodified reference (tests/golden); no GPU, no libnfopp.so",
              "rng": "numpy legacy MT19937 + torch CPU generator, as the reference", "host_cpu": host_cpu_model(),
              "anomaly_mode": "not applicable (analytic gradients; the reference ships with torch anomaly detection ON, which "
                              "costs it 3-5x: BASELINE.md 2)"}
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT,
            "n_gpus": int(os.environ.get("WORLD_SIZE", args.gpus)), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * cores / base["value"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 / f64", "data": "synthetic", "config": config,
            "cpu_baseline": base, "cpu_baseline_1thread": one,
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
    ap.add_argument("--no-extra", action="store_true",
                    help="skip configs[2] (16k forest, sharded), configs[3] (gridmap-wide), the stream and single-problem lines")
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

"""GPU-side helpers for the parity tests: build a PlannerEngine from a golden reference state."""
import numpy as np
import torch

import _util as U
from nfopp_b200.engine import EngineConfig, OnfSpec, PlannerEngine


def onf_spec(ocfg) -> OnfSpec:
    return OnfSpec(mean=ocfg.mean, sigma=ocfg.sigma, use_cos=ocfg.use_cos, angle=ocfg.angle, hidden=ocfg.hidden,
                   layers=ocfg.layers, bias=ocfg.bias)


def engine_config(name, fx, rng_mode="inject", problems=1, **over) -> EngineConfig:
    """over: any EngineConfig field, e.g. onf_impl="fma" | "tensor"."""
    pc = U.PLANNER_CFGS[name]
    obs = fx["meta/obstacles"]
    kw = dict(
        problems=problems, constrained=pc.constrained, trajectory_length=pc.trajectory_length, onf=onf_spec(pc.onf),
        onf_lr=pc.onf_lr, onf_betas=pc.onf_betas, traj_lr=pc.traj_lr, traj_betas=pc.traj_betas, adam_eps=pc.adam_eps,
        trajectory_random_offset=pc.trajectory_random_offset, collision_weight=pc.collision_weight,
        velocity_hessian_weight=pc.velocity_hessian_weight, init_collision_iteration=pc.init_collision_iteration,
        init_collision_points=pc.init_collision_points,
        reparametrize_trajectory_freq=pc.reparametrize_trajectory_freq,
        optimize_collision_model_freq=pc.optimize_collision_model_freq, random_field_points=pc.random_field_points,
        course_random_offset=pc.course_random_offset, collision_point_count=pc.collision_point_count,
        angle_weight=pc.angle_weight, constraint_deltas_weight=pc.constraint_deltas_weight,
        multipliers_lr=pc.multipliers_lr, boundary_weight=pc.boundary_weight,
        collision_multipliers_lr=pc.collision_multipliers_lr, angle_offset=pc.angle_offset,
        collision_beta=pc.collision_beta, direction_delta_weight=pc.direction_delta_weight,
        checker="rectangle" if name == "car" else "circle", env_shared=True, max_obstacles=max(64, len(obs)),
        robot_radius=0.3, box=(-0.3, 0.2, -0.3, 0.2), checker_bounds=tuple(float(b) for b in fx["meta/bounds"]),
        border=tuple(float(b) for b in fx["free/state0/border" if "free/state0/border" in fx.files else "warm/before/border"]),
        rng_mode=rng_mode, max_choice_uniforms=4096)
    kw.update(over)
    return EngineConfig(**kw)


def load_state(eng: PlannerEngine, st: dict, problem=None):
    """Oracle-format state dict -> engine buffers (all problems when problem is None)."""
    sel = slice(None) if problem is None else slice(problem, problem + 1)

    def put(name, val):
        buf = eng[name][sel]
        buf.copy_(torch.as_tensor(np.asarray(val)).to(buf.dtype).reshape(1, *buf.shape[1:]).expand_as(buf))

    put("onf_w", st["w"]); put("onf_m", st["m"]); put("onf_v", st["v"])
    put("traj", st["traj"]); put("traj_m", st["traj_m"]); put("traj_v", st["traj_v"])
    put("lam", st["lam"]); put("mu", st["mu"]); put("start", st["start"]); put("goal", st["goal"])
    if st["prev_traj"] is not None:
        put("prev_traj", st["prev_traj"])
    n = len(st["buf_pos"])
    eng["buf_n"][sel] = n
    if n:
        eng["buf_pos"][sel, :n] = torch.as_tensor(st["buf_pos"]).to(eng.device)
        eng["buf_age"][sel, :n] = torch.as_tensor(st["buf_age"]).to(eng.device)
    eng.set_counter("step_count", st["step_count"])
    eng.set_counter("onf_step", st["onf_step"])
    eng.set_counter("traj_step", st["traj_step"])
    eng.set_counter("has_prev", int(st["prev_traj"] is not None))


def engine_from_golden(name, fx, prefix, **over) -> PlannerEngine:
    eng = PlannerEngine(engine_config(name, fx, **over))
    eng.set_obstacles(fx["meta/obstacles"])
    load_state(eng, U.state_from(fx, prefix))
    return eng


_PAIRS = [("onf_w", "w"), ("onf_m", "m"), ("onf_v", "v"), ("traj", "traj"), ("traj_m", "traj_m"),
          ("traj_v", "traj_v"), ("lam", "lam"), ("mu", "mu")]


def compare_state(eng: PlannerEngine, st: dict, tol, problem=0, skip=()):
    """max-norm relative error per state tensor; asserts each below tol and returns the worst."""
    worst = {}
    for en, on in _PAIRS:
        if on in skip:
            continue
        err = U.relerr(eng[en][problem].cpu().numpy(), np.asarray(st[on]))
        worst[on] = err
        assert err < tol, f"{on}: rel err {err:.3e} >= {tol}"
    n = int(eng["buf_n"][problem])
    assert n == len(st["buf_pos"]), (n, len(st["buf_pos"]))
    if n:
        assert np.array_equal(eng["buf_pos"][problem, :n].cpu().numpy(), np.asarray(st["buf_pos"])), "replay buffer"
        assert np.array_equal(eng["buf_age"][problem, :n].cpu().numpy(), np.asarray(st["buf_age"])), "replay ages"
    for en, on in [("step_count", "step_count"), ("onf_step", "onf_step"), ("traj_step", "traj_step")]:
        assert eng.counter(en) == st[on], (en, eng.counter(en), st[on])
    return max(worst.values())

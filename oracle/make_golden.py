"""Generates tests/golden/*.npz from the UNMODIFIED reference (run in the build container only):

    python -m oracle.make_golden            # needs /root/reference

The reference's own tests hold no numeric vectors, so these fixtures - outputs of the
reference's code on seeded inputs - are what pins the oracle (and through it the CUDA path) on the GPU box,
where /root/reference does not exist.  torch 2.11.0 / numpy 2.3.5, 1 thread, seeds as in
scripts/run_planner.py:11-12.
"""
import json
import os
import sys

import numpy as np
import torch

from . import onf as O
from . import reference_bridge as RB

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _np(x):
    if x is None:
        return np.zeros(0)
    if torch.is_tensor(x):
        return x.detach().numpy().copy()
    return np.asarray(x)


def _flat_state(prefix, st):
    out = {}
    for k, v in st.items():
        if k == "prev_traj":
            out[f"{prefix}/has_prev"] = np.array(v is not None)
            out[f"{prefix}/prev_traj"] = _np(v)
        else:
            out[f"{prefix}/{k}"] = _np(v)
    return out


def golden_checkers():
    ref = RB.load_reference()
    cc, teb = ref["collision_checker"], ref["test_environment_builder"].TestEnvironmentBuilder
    P2 = ref["position2"].Position2
    rng = np.random.RandomState(7)
    n = 4096
    poses = np.stack([rng.uniform(-0.5, 3.5, n), rng.uniform(-0.5, 3.5, n), rng.uniform(-7, 7, n)], 1)
    poses[:64, 0] = np.linspace(-0.1, 3.1, 64)          # a few exactly-on-boundary style values
    poses[64:128, :2] = np.round(poses[64:128, :2], 1)
    out = {"poses": poses}
    obs_s = teb._obstacle_points()
    obs_car = teb._car_obstacle_points()
    out["obstacles_sbend"], out["obstacles_car"] = obs_s, obs_car
    c = cc.CircleCollisionChecker(0.3, (0, 3, 0, 3))
    c.update_obstacle_points(obs_s)
    out["circle_sbend"] = c.check_collision(poses[:, :2])
    c = cc.CircleDirectedCollisionChecker(0.3, (0, 3, 0, 3))
    c.update_obstacle_points(obs_car)
    out["circle_directed_car"] = c.check_collision(P2.from_vec(poses))
    c = cc.CircleCollisionChecker(0.25, None)
    c.update_obstacle_points(obs_car)
    out["circle_car_nobounds"] = np.asarray(c.check_collision(poses[:, :2])) | False
    c = cc.RectangleCollisionChecker((-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))
    c.update_obstacle_points(obs_car)
    out["rect_car"] = c.check_collision(P2.from_vec(poses))
    c = cc.RectangleCollisionChecker((-0.2, 0.2, -0.2, 0.2), (0, 3, 0, 3))
    c.update_obstacle_points(obs_s)
    out["rect_sbend"] = c.check_collision(P2.from_vec(poses))
    c = cc.CollisionChecker((0, 3, 0, 3))
    out["boundary_only"] = c.check_collision(poses[:, :2])
    # occupancy-grid checker: executes the notebook's own class source (cell 2 of onf_planner_image_map.ipynb)
    nb = json.load(open(os.path.join(RB.REFERENCE_ROOT, "notebooks", "onf_planner_image_map.ipynb")))
    src = "".join(nb["cells"][2]["source"])
    ns = {"np": np, "CollisionChecker": cc.CollisionChecker}
    exec("from dataclasses import dataclass\n" + src, ns)
    grid = (np.random.RandomState(1).rand(60, 80) < 0.2).astype(np.float64) * 255
    mi = ns["MapImage"](cols=80, rows=60, origin_x=-1.0, origin_y=-0.5, cell_size=0.05, map_image=grid)
    gc = ns["MapCollisionChecker"](mi, (-1, 3, -0.5, 2.5))
    gposes = poses.copy()
    gposes[:, 0] = rng.uniform(-1.3, 3.3, n)
    gposes[:, 1] = rng.uniform(-0.8, 2.8, n)
    gposes[:200, 0] = -1.0 + 0.05 * rng.randint(-2, 84, 200) + 0.025      # exactly on cell edges (half-cell offset)
    gposes[200:400, 1] = -0.5 + 0.05 * rng.randint(-2, 64, 200) + 0.025
    out["grid"], out["grid_poses"] = grid, gposes
    out["grid_meta"] = np.array([-1.0, -0.5, 0.05])
    out["grid_labels"] = gc.check_collision(P2.from_vec(gposes))
    np.savez_compressed(os.path.join(OUT, "checkers.npz"), **out)
    print("checkers:", {k: int(np.sum(v)) for k, v in out.items() if v.dtype == bool})


def golden_onf():
    ref = RB.load_reference()
    from neural_field_optimal_planner.onf_model import ONF
    cases = {"holonomic": dict(mean=1.5, sigma=1), "car": dict(mean=0, sigma=1, use_cos=True, use_normal_init=True,
                                                                angle_encoding=True),
             "image": dict(mean=0, sigma=7, use_cos=True, use_normal_init=True)}
    out = {}
    for i, (name, kw) in enumerate(cases.items()):
        torch.manual_seed(10 + i)
        model = ONF(**kw)
        cfg = RB.onf_config_of(model)
        d = cfg.in_dim
        x = torch.rand(96, d) * torch.tensor([3.0, 3.0, 6.28][:d]) + torch.tensor([0.0, 0.0, -3.14][:d])
        x.requires_grad_(True)
        y = (torch.rand(96) < 0.3).float()
        model.requires_grad_(True)
        z = model(x)[:, 0]
        loss = torch.nn.BCEWithLogitsLoss()(z, y) + 0.01 * torch.nn.functional.softplus(z).sum()
        loss.backward()
        named = {RB._PARAM_TO_SEG[n]: p.grad.detach().reshape(-1) for n, p in model.named_parameters()}
        for seg in ("ang_b", "ang_f"):
            named.setdefault(seg, torch.zeros(0))
        out.update({f"{name}/w": _np(O.blob_from_state_dict(model.state_dict(), cfg)), f"{name}/x": _np(x),
                    f"{name}/y": _np(y), f"{name}/z": _np(z), f"{name}/gw": _np(O.pack(named, cfg)),
                    f"{name}/gx": _np(x.grad), f"{name}/loss": _np(loss),
                    f"{name}/cfg": np.array([cfg.mean, cfg.sigma, cfg.use_cos, cfg.angle], dtype=np.float64)})
    np.savez_compressed(os.path.join(OUT, "onf.npz"), **out)
    print("onf:", sorted(out)[:4], "...")


def golden_choice():
    out = {}
    rs = np.random.RandomState(5)
    for i, (n, conc) in enumerate([(198, 1.0), (199, 0.05), (199, 8.0), (150, 0.5)]):
        p = rs.gamma(conc, size=n) + 1e-6
        p /= p.sum()
        for replace in (False, True):
            st = np.random.RandomState(100 + i)
            clone = np.random.RandomState(100 + i)
            idx = st.choice(n, 100, replace=replace, p=p)
            used = []

            def draw(k):
                u = clone.rand(k)
                used.append(u)
                return u
            from .sampling import legacy_choice
            assert np.array_equal(legacy_choice(p, 100, replace, draw), idx)
            out[f"{i}_{int(replace)}/p"], out[f"{i}_{int(replace)}/idx"] = p, idx
            out[f"{i}_{int(replace)}/u"] = np.concatenate(used)
    np.savez_compressed(os.path.join(OUT, "choice.npz"), **out)
    print("choice:", len(out) // 3, "cases")


def golden_lockstep(name, steps, free_run=10, tag=None, **overrides):
    """Chain of reference states around the given step indices + the variates of each step, plus a short
    free run (initial state, all variates, reference path)."""
    planner, env, checker = RB.make_reference_case(name, **overrides)
    cfg = RB.planner_config_of(planner)
    out = {"meta/steps": np.array(steps), "meta/obstacles": np.asarray(checker._obstacle_points),
           "meta/bounds": np.asarray(checker._boundaries, dtype=np.float64),
           "meta/init_collision_iteration": np.array(cfg.init_collision_iteration)}
    out.update(_flat_state("free/state0", RB.snapshot(planner)))
    for k in range(max(max(steps) + 1, free_run)):
        if k in steps:
            out.update(_flat_state(f"step{k}/before", RB.snapshot(planner)))
        with RB.record_variates(cfg.constrained, cfg.trajectory_length) as rec:
            planner.step()
        if k in steps:
            out.update(_flat_state(f"step{k}/after", RB.snapshot(planner)))
            out.update({f"step{k}/var/{kk}": _np(v) for kk, v in rec.items()})
            out[f"step{k}/labels"] = np.asarray(planner.truth_collision, dtype=bool)
            cp = planner.checked_positions
            out[f"step{k}/checked"] = cp.as_vec() if hasattr(cp, "as_vec") else np.asarray(cp)
        if k < free_run:
            out.update({f"free/var{k}/{kk}": _np(v) for kk, v in rec.items()})
            out[f"free/path{k}"] = planner.get_path()
    fname = f"lockstep_{tag or name}.npz"
    np.savez_compressed(os.path.join(OUT, fname), **out)
    print(tag or name, "lockstep fixture:", os.path.getsize(os.path.join(OUT, fname)) // 1024, "KiB")


def golden_warmup(name, iterations, **overrides):
    """ONF warm-up (nerf_opt_planner.py:181-187,197-200): init() walked by hand on the UNMODIFIED reference planner -
    start / goal / border, _init_trajectory(), then `iterations` x {_sample_random_field_points(init_collision_points)
    (:135-141; constrained :173-176) -> _optimize_collision_model(positions) (:76-91)} - with the planner's full state
    snapshotted before the loop and after every iteration, and every iteration's uniforms, sampled points and labels."""
    planner, env, checker = RB.make_reference_case(name, call_init=False, init_collision_iteration=iterations, **overrides)
    cfg = RB.planner_config_of(planner)
    assert cfg.init_collision_iteration == iterations
    planner._start_point = torch.tensor(env.start_point.astype(np.float32))[None]
    planner._goal_point = torch.tensor(env.goal_point.astype(np.float32))[None]
    planner._random_sample_border = env.bounds
    planner._init_trajectory()
    out = {"meta/iterations": np.array(iterations), "meta/obstacles": np.asarray(checker._obstacle_points),
           "meta/bounds": np.asarray(checker._boundaries, dtype=np.float64),
           "meta/init_collision_points": np.array(cfg.init_collision_points)}
    out.update(_flat_state("warm/before", RB.snapshot(planner)))
    for i in range(iterations):
        with RB.record_variates(cfg.constrained, cfg.trajectory_length, order=["warm_xy", "warm_th"]) as rec:
            positions = planner._sample_random_field_points(planner._init_collision_points)
            planner._optimize_collision_model(positions)
        out.update(_flat_state(f"warm/after{i}", RB.snapshot(planner)))
        out.update({f"warm/var{i}/{kk}": _np(v) for kk, v in rec.items() if kk.startswith("warm_")})
        out[f"warm/labels{i}"] = np.asarray(planner.truth_collision, dtype=bool)
        cp = planner.checked_positions
        out[f"warm/checked{i}"] = cp.as_vec() if hasattr(cp, "as_vec") else np.asarray(cp)
    planner._step_count = 0
    # the first planner step after the warm-up (separate step counters: the ONF optimiser is at `iterations`, the
    # trajectory optimiser at 0)
    out.update(_flat_state("step0/before", RB.snapshot(planner)))
    with RB.record_variates(cfg.constrained, cfg.trajectory_length) as rec:
        planner.step()
    out.update(_flat_state("step0/after", RB.snapshot(planner)))
    out.update({f"step0/var/{kk}": _np(v) for kk, v in rec.items()})
    np.savez_compressed(os.path.join(OUT, f"warmup_{name}.npz"), **out)
    print(name, "warm-up fixture:", os.path.getsize(os.path.join(OUT, f"warmup_{name}.npz")) // 1024, "KiB")


def main():
    torch.set_num_threads(1)
    os.makedirs(OUT, exist_ok=True)
    golden_checkers()
    golden_onf()
    golden_choice()
    golden_lockstep("car", steps=[0, 1, 10])
    golden_lockstep("corridor", steps=[1, 10])
    golden_lockstep("holonomic", steps=[0, 1, 10], init_collision_iteration=2)
    # training every 2nd step, re-parametrisation every 3rd (nerf_opt_planner.py:62,66)
    golden_lockstep("corridor", steps=[0, 1, 2, 3], free_run=6, tag="corridor_freq", optimize_collision_model_freq=2,
                    reparametrize_trajectory_freq=3)
    golden_warmup("holonomic", 3)          # planner_factory.py:59 runs 400 of these; the image-map notebook 3
    golden_warmup("car", 3)


if __name__ == "__main__":
    sys.exit(main())

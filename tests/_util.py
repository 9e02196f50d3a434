"""Shared helpers for the parity tests: golden-fixture loading and oracle planner construction."""
import copy
import os

import numpy as np
import torch

from oracle import checkers as C
from oracle import onf as O
from oracle.planner import OraclePlanner, PlannerConfig

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name))


def relerr(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0 and b.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


ONF_CFGS = {
    "holonomic": O.OnfConfig(mean=1.5, sigma=1.0),
    "car": O.OnfConfig(mean=0.0, sigma=1.0, use_cos=True, angle=True),
    "corridor": O.OnfConfig(mean=0.0, sigma=1.0, use_cos=True, angle=True),
    "image": O.OnfConfig(mean=0.0, sigma=7.0, use_cos=True),
}

# scripts/run_planner.py:14-55 / planner_factory.py:51-59 (values probed from the reference objects by
# oracle.reference_bridge.planner_config_of when the fixtures were made)
PLANNER_CFGS = {
    "car": PlannerConfig(constrained=True, onf=ONF_CFGS["car"], onf_lr=5e-2, onf_betas=(0.9, 0.9), traj_lr=1e-2,
                         traj_betas=(0.9, 0.9), init_collision_iteration=0, angle_offset=0.3),
    "corridor": PlannerConfig(constrained=True, onf=ONF_CFGS["corridor"], onf_lr=5e-2, onf_betas=(0.9, 0.9),
                              traj_lr=1e-2, traj_betas=(0.9, 0.9), init_collision_iteration=0, angle_offset=0.3,
                              direction_delta_weight=100.0),
    "holonomic": PlannerConfig(constrained=False, onf=ONF_CFGS["holonomic"], onf_lr=1e-3, onf_betas=(0.9, 0.9),
                               traj_lr=1e-2, traj_betas=(0.9, 0.999), collision_weight=0.01,
                               velocity_hessian_weight=3.0, init_collision_iteration=2),
}


import dataclasses
# nerf_opt_planner.py:62,66 with both frequencies off their defaults (tests/golden/lockstep_corridor_freq.npz)
PLANNER_CFGS["corridor_freq"] = dataclasses.replace(PLANNER_CFGS["corridor"], optimize_collision_model_freq=2,
                                                    reparametrize_trajectory_freq=3)
ONF_CFGS["corridor_freq"] = ONF_CFGS["corridor"]


def labeller_for(name, fx):
    obs, bounds = fx["meta/obstacles"], tuple(fx["meta/bounds"])
    if name == "car":
        return lambda p: C.rectangle_labels(p, obs, (-0.3, 0.2, -0.3, 0.2), bounds)
    return lambda p: C.circle_labels(p, obs, 0.3, bounds)


STATE_KEYS = ["w", "m", "v", "onf_step", "traj", "traj_m", "traj_v", "traj_step", "lam", "mu", "buf_pos", "buf_age",
              "step_count", "start", "goal", "border"]


def state_from(fx, prefix):
    st = {}
    for k in STATE_KEYS:
        v = fx[f"{prefix}/{k}"]
        if k in ("onf_step", "traj_step", "step_count"):
            st[k] = int(v)
        elif k == "border":
            st[k] = tuple(float(b) for b in v)
        elif k in ("buf_pos", "buf_age"):
            st[k] = np.array(v, dtype=np.float64)
        else:
            st[k] = torch.tensor(np.array(v, dtype=np.float32))
    st["prev_traj"] = torch.tensor(fx[f"{prefix}/prev_traj"]) if bool(fx[f"{prefix}/has_prev"]) else None
    return st


def variates_from(fx, prefix):
    pre = prefix + "/"
    return {k[len(pre):]: fx[k] for k in fx.files if k.startswith(pre)}


def oracle_planner(name, fx, prefix):
    cfg = PLANNER_CFGS[name]
    st = state_from(fx, prefix)
    op = OraclePlanner(cfg, st["w"], labeller_for(name, fx))
    op.state.update(copy.deepcopy(st))
    return op

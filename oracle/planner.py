"""Oracle: one NFOPP planner (holonomic or constrained) with explicit state and explicit variates.

Restates NERFOptPlanner.step / init (nerf_opt_planner.py:60-71,181-200) and the ConstrainedNERFOptPlanner
overrides (constrained_nerf_opt_planner.py:42-176) on top of oracle.{onf,sampling,trajectory,checkers}.
State is a plain dict of tensors/arrays so the CUDA engine's export_state()/import_state() can be compared
with it field by field (lock-step parity, SURVEY.md §7 hard part 1).
"""
from dataclasses import dataclass, field as dc_field
from typing import Callable, Optional

import numpy as np
import torch

from . import onf as O
from . import sampling as S
from . import trajectory as TR


@dataclass
class PlannerConfig:
    constrained: bool = True
    trajectory_length: int = 100
    onf: O.OnfConfig = dc_field(default_factory=O.OnfConfig)
    # optimisers (planner_factory.py:54-56,68-70)
    onf_lr: float = 1e-2
    onf_betas: tuple = (0.9, 0.9)
    traj_lr: float = 1e-2
    traj_betas: tuple = (0.9, 0.9)
    adam_eps: float = 1e-8
    # NERFOptPlanner ctor (nerf_opt_planner.py:11-14)
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
    # ConstrainedNERFOptPlanner ctor (constrained_nerf_opt_planner.py:13-18)
    angle_weight: float = 0.5
    constraint_deltas_weight: float = 20.0
    multipliers_lr: float = 0.1
    boundary_weight: float = 1.0
    collision_multipliers_lr: float = 1e-3
    angle_offset: float = 0.0
    collision_beta: float = 1.0
    direction_delta_weight: float = 0.0

    @property
    def dim(self):
        return 3 if self.constrained else 2


class LiveSource:
    """Draws variates in the reference's order from a legacy numpy RandomState and a torch generator,
    recording them."""

    def __init__(self, rs: np.random.RandomState, gen: Optional[torch.Generator] = None):
        self.rs, self.gen, self.rec = rs, gen, {}

    def begin(self):
        self.rec = {"choice_u": np.zeros(0)}

    def rand(self, name, *shape):
        v = self.rs.rand(*shape)
        self.rec[name] = v
        return v

    def randn(self, name, *shape):
        v = self.rs.randn(*shape)
        self.rec[name] = v
        return v

    def choice_draw(self, k):
        v = self.rs.rand(k)
        self.rec["choice_u"] = np.concatenate([self.rec["choice_u"], v])
        return v

    def torch_rand(self, name, n):
        v = torch.rand(n, 1, generator=self.gen)[:, 0] if self.gen is not None else torch.rand(n, 1)[:, 0]
        self.rec[name] = v.numpy().copy()
        return v


class RecordedSource:
    """Replays a recorded variate dict (lock-step parity)."""

    def __init__(self, rec: dict):
        self.rec = rec
        self._reader = S.UniformReader(rec.get("choice_u", np.zeros(0)))

    def begin(self):
        pass

    def rand(self, name, *shape):
        return np.asarray(self.rec[name], dtype=np.float64).reshape(shape)

    randn = rand

    def choice_draw(self, k):
        return self._reader(k)

    def torch_rand(self, name, n):
        return torch.tensor(np.asarray(self.rec[name], dtype=np.float32).reshape(n))


class OraclePlanner:
    def __init__(self, cfg: PlannerConfig, blob: torch.Tensor, labeller: Callable[[np.ndarray], np.ndarray]):
        self.cfg, self.labeller = cfg, labeller
        T, D = cfg.trajectory_length, cfg.dim
        self.inv_h = TR.inv_hessian(T, cfg.velocity_hessian_weight)
        z = torch.zeros
        self.state = dict(
            w=blob.clone().float(), m=z(blob.numel()), v=z(blob.numel()), onf_step=0,
            traj=z(T, D), traj_m=z(T, D), traj_v=z(T, D), traj_step=0,
            lam=z(T + 1), mu=z(T), prev_traj=None,
            buf_pos=np.zeros((0, D)), buf_age=np.zeros(0), step_count=0,
            start=z(D), goal=z(D), border=(0.0, 0.0, 0.0, 0.0))
        self.last = {}

    # ---- init (nerf_opt_planner.py:181-200; trajectory_initializer.py:12-29) ----
    def init(self, start, goal, border, src=None):
        st, cfg = self.state, self.cfg
        st["start"] = torch.tensor(np.asarray(start, dtype=np.float32))
        st["goal"] = torch.tensor(np.asarray(goal, dtype=np.float32))
        st["border"] = tuple(float(b) for b in border)
        n = cfg.trajectory_length + 2
        for d in range(2):
            st["traj"][:, d] = torch.linspace(st["start"][d], st["goal"][d], n)[1:-1]
        if cfg.constrained:
            goal_angle = TR.wrap_angle(st["goal"][2] - st["start"][2]) + st["start"][2]
            st["traj"][:, 2] = torch.linspace(st["start"][2], goal_angle, n)[1:-1]
        for _ in range(cfg.init_collision_iteration):
            src.begin()
            self._train_field(src, self._field_points(src, cfg.init_collision_points, "warm"))
        st["step_count"] = 0

    # ---- step (nerf_opt_planner.py:60-71) ----
    def step(self, src):
        st, cfg = self.state, self.cfg
        src.begin()
        if st["step_count"] % cfg.optimize_collision_model_freq == 0:
            self._train_field(src)
        self._optimize_trajectory(src)
        if st["step_count"] % cfg.reparametrize_trajectory_freq == 0:
            self._reparametrize()
        st["step_count"] += 1
        return getattr(src, "rec", None)

    def get_path(self):
        st = self.state
        return torch.cat([st["start"][None], st["traj"], st["goal"][None]], 0).numpy().copy()

    # ---- field training (nerf_opt_planner.py:76-99) ----
    def _train_field(self, src, positions=None):
        st, cfg = self.state, self.cfg
        if positions is None:
            if st["prev_traj"] is None:
                st["prev_traj"] = st["traj"].clone()
            positions = self._sample(st["prev_traj"], src)
            st["prev_traj"] = st["traj"].clone()
        x32 = torch.tensor(positions.astype(np.float32))
        z, cache = O.onf_forward(st["w"], cfg.onf, x32)
        labels = np.asarray(self.labeller(positions), dtype=bool)
        y = torch.tensor(labels.astype(np.float32))
        loss, gz = O.bce_with_logits_grad(z, y)
        g, _ = O.onf_backward(st["w"], cfg.onf, cache, gz)
        st["onf_step"] += 1
        O.adam_step(st["w"], st["m"], st["v"], g, st["onf_step"], cfg.onf_lr, *cfg.onf_betas, eps=cfg.adam_eps)
        self.last.update(train_positions=positions, train_labels=labels, train_logits=z, train_loss=loss,
                         onf_grad=g)

    def _field_points(self, src, k, tag):
        cfg = self.cfg
        u_xy = src.rand(f"{tag}_xy", k, 2)
        u_th = src.rand(f"{tag}_th", k, 1) if cfg.constrained else None
        return S.field_points(u_xy, u_th, self.state["border"])

    def _sample(self, prev, src):
        """nerf_opt_planner.py:101-111."""
        st, cfg = self.state, self.cfg
        n = cfg.trajectory_length - 1
        t = src.rand("t_sample", n).astype(np.float32)
        base = S.intermediate_positions(prev, torch.tensor(t)).numpy().astype(np.float64)
        cz = src.randn("coarse_xy", n, 2)
        ct = src.randn("coarse_th", n) if cfg.constrained else None
        coarse = S.offset_positions(base, cz, ct, cfg.course_random_offset, cfg.angle_offset)
        fz = src.randn("fine_xy", n, 2)
        ft = src.randn("fine_th", n) if cfg.constrained else None
        fine = S.offset_positions(base, fz, ft, cfg.trajectory_random_offset, cfg.angle_offset)
        cand_pos = np.concatenate([st["buf_pos"], fine], 0)
        cand_age = np.concatenate([st["buf_age"], np.zeros(len(fine))], 0)
        z, _ = O.onf_forward(st["w"], cfg.onf, torch.tensor(cand_pos.astype(np.float32)))
        sig = torch.sigmoid(z).numpy()
        st["buf_pos"], st["buf_age"], idx = S.resample_buffer(cand_pos, cand_age, sig, cfg.collision_point_count,
                                                              src.choice_draw)
        self.last.update(cand_pos=cand_pos, cand_age=cand_age, cand_sig=sig, choice_idx=idx)
        uni = self._field_points(src, cfg.random_field_points, "field")
        return np.concatenate([coarse, st["buf_pos"], uni], 0)

    # ---- trajectory (nerf_opt_planner.py:143-155; constrained:63-100) ----
    def _optimize_trajectory(self, src):
        st, cfg = self.state, self.cfg
        n = cfg.trajectory_length - 1
        if cfg.constrained:
            t = src.torch_rand("t_path", n)
        else:
            t = torch.tensor(src.rand("t_path", n).astype(np.float32))
        q = TR.path_points(st["traj"], t, cfg.constrained)
        z, cache = O.onf_forward(st["w"], cfg.onf, q)
        mu_i = TR.interp_multipliers(st["mu"], t) if cfg.constrained else None
        gbar = TR.path_gbar(z, mu_i, cfg.collision_weight, cfg.collision_beta if cfg.constrained else 1.0)
        _, gq = O.onf_backward(st["w"], cfg.onf, cache, gbar)
        if cfg.constrained:
            g, g_lam, g_mu = TR.constrained_gradient(
                st["traj"], st["start"], st["goal"], st["lam"], st["mu"], t, z, gq,
                angle_weight=cfg.angle_weight, constraint_weight=cfg.constraint_deltas_weight,
                boundary_weight=cfg.boundary_weight, direction_weight=cfg.direction_delta_weight,
                border=st["border"])
        else:
            g = TR.holonomic_gradient(st["traj"], st["start"], st["goal"], t, gq)
        self.last.update(path_q=q, path_z=z, path_gq=gq, traj_grad_raw=g.clone())
        g = self.inv_h @ g                                               # nerf_opt_planner.py:151
        st["traj_step"] += 1
        O.adam_step(st["traj"], st["traj_m"], st["traj_v"], g, st["traj_step"], cfg.traj_lr, *cfg.traj_betas,
                    eps=cfg.adam_eps)
        if cfg.constrained:                                              # constrained:66-73
            st["lam"] = st["lam"] + cfg.multipliers_lr * g_lam
            mu = st["mu"] + cfg.collision_multipliers_lr * g_mu
            st["mu"] = torch.where(mu > 0, mu, torch.zeros_like(mu))

    def _reparametrize(self):
        st = self.state
        if self.cfg.constrained:
            st["traj"], st["lam"], st["mu"] = TR.reparam_constrained(st["traj"], st["start"], st["goal"],
                                                                   st["lam"], st["mu"])
        else:
            st["traj"] = TR.reparam_holonomic(st["traj"], st["start"], st["goal"])

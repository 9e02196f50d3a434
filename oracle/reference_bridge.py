"""Bridge to the UNMODIFIED reference under /root/reference (test infrastructure; container only).

Imports the reference's modules read-only (with a 6-line pytorch_lightning.AttributeDict stand-in from
oracle/_standins, the only missing dependency), builds its planners for the BASELINE configs, snapshots
their full state in the oracle's packed format, and records every random variate a step() consumes so the
oracle restatement and the CUDA engine can be stepped in lock-step with it.

/root/reference does not exist on the GPU box: nothing that runs there may import this module.
"""
import contextlib
import os
import sys

import numpy as np
import torch

from . import onf as O
from . import sampling as S
from .planner import PlannerConfig

REFERENCE_ROOT = os.environ.get("NFOPP_REFERENCE_ROOT", "/root/reference")
_STANDINS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_standins")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "neural_field_optimal_planner"))


def load_reference():
    """Returns the reference package (neural_field_optimal_planner) imported from REFERENCE_ROOT."""
    if not reference_available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT}")
    for p in (_STANDINS, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        import neural_field_optimal_planner  # noqa: F401
        import neural_field_optimal_planner.planner_factory as pf
        import neural_field_optimal_planner.collision_checker as cc
        import neural_field_optimal_planner.test_environment_builder as teb
        import neural_field_optimal_planner.utils.position2 as p2
    torch.autograd.set_detect_anomaly(False)
    mod = sys.modules["neural_field_optimal_planner"]
    assert os.path.abspath(mod.__file__).startswith(os.path.abspath(REFERENCE_ROOT)), mod.__file__
    return dict(pkg=mod, planner_factory=pf, collision_checker=cc, test_environment_builder=teb, position2=p2)


def attribute_dict(**kw):
    load_reference()
    from pytorch_lightning.utilities import AttributeDict
    return AttributeDict(**{k: attribute_dict(**v) if isinstance(v, dict) else v for k, v in kw.items()})


# --- the BASELINE configurations (scripts/run_planner.py:14-64; planner_factory.py:51-59) -----------------
RUN_PLANNER_PARAMETERS = dict(
    device="cpu", trajectory_length=100,
    collision_model=dict(mean=0, sigma=1, use_cos=True, bias=True, use_normal_init=True, angle_encoding=True,
                         name="ONF"),
    trajectory_initializer=dict(name="TrajectoryInitializer", resolution=0.05),
    collision_optimizer=dict(lr=5e-2, betas=(0.9, 0.9)),
    trajectory_optimizer=dict(lr=1e-2, betas=(0.9, 0.9)),
    planner=dict(name="ConstrainedNERFOptPlanner", trajectory_random_offset=0.02, collision_weight=1,
                 velocity_hessian_weight=0.5, random_field_points=10, init_collision_iteration=0,
                 constraint_deltas_weight=20, multipliers_lr=0.1, init_collision_points=100,
                 reparametrize_trajectory_freq=10, optimize_collision_model_freq=1, angle_weight=0.5,
                 angle_offset=0.3, boundary_weight=1, collision_multipliers_lr=1e-3))


def make_reference_case(name: str, torch_seed=100, numpy_seed=400, call_init=True, **planner_overrides):
    """name in {'car' (scripts/run_planner.py as written), 'corridor' (config 2: S-bend walls, disc robot,
    forward-only), 'holonomic' (PlannerFactory.make_onf_planner + circle checker)}.
    Returns (planner, env, checker) with init() already called (call_init=False: not yet - the warm-up fixture
    walks init() by hand)."""
    ref = load_reference()
    torch.random.manual_seed(torch_seed)
    np.random.seed(numpy_seed)
    teb = ref["test_environment_builder"].TestEnvironmentBuilder
    cc = ref["collision_checker"]
    pf = ref["planner_factory"].PlannerFactory
    if name == "holonomic":
        env = teb.make_test_environment()
        checker = cc.CircleCollisionChecker(0.3, (0, 3, 0, 3))
        checker.update_obstacle_points(env.obstacle_points)
        planner = pf.make_onf_planner(checker)
        for k, v in planner_overrides.items():
            setattr(planner, "_" + k, v)
    else:
        params = {k: (dict(v) if isinstance(v, dict) else v) for k, v in RUN_PLANNER_PARAMETERS.items()}
        if name == "car":
            env = teb().make_car_environment()
            checker = cc.RectangleCollisionChecker((-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))
        elif name == "corridor":
            env = teb.make_test_environment_with_angles()
            checker = cc.CircleDirectedCollisionChecker(0.3, (0, 3, 0, 3))
            params["planner"]["direction_delta_weight"] = 100      # scripts/run_bench_mr.py:60 (forward-only)
        else:
            raise KeyError(name)
        params["planner"].update(planner_overrides)
        checker.update_obstacle_points(env.obstacle_points)
        planner = pf.make_constrained_onf_planner(checker, attribute_dict(**params))
    if call_init:
        with _quiet():
            planner.init(env.start_point, env.goal_point, env.bounds)
    return planner, env, checker


@contextlib.contextmanager
def _quiet():
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        yield


def onf_config_of(model) -> O.OnfConfig:
    return O.OnfConfig(mean=float(model._mean), sigma=float(model._sigma), use_cos=bool(model._use_cos),
                       angle=model._angle_encoder is not None, hidden=100, layers=2,
                       bias=model.encoding_layer.bias is not None)


def planner_config_of(planner) -> PlannerConfig:
    """Reads the reference planner's private attributes (nerf_opt_planner.py:16-43; constrained:23-40)."""
    constrained = hasattr(planner, "_constraint_multipliers")
    og = planner._collision_optimizer.param_groups[0]
    tg = planner._trajectory_optimizer.param_groups[0]
    ih = planner._inv_hessian.numpy().astype(np.float64)
    T = planner._trajectory.shape[0]
    # velocity_hessian_weight is not stored; recover it from inv_hessian: H = w*K + I, K[0,0] = 4
    w_h = float((np.linalg.inv(ih)[0, 0] - 1.0) / 4.0)
    kw = dict(
        constrained=constrained, trajectory_length=T, onf=onf_config_of(planner._collision_model),
        onf_lr=og["lr"], onf_betas=tuple(og["betas"]), traj_lr=tg["lr"], traj_betas=tuple(tg["betas"]),
        adam_eps=og["eps"], trajectory_random_offset=planner._fine_random_offset,
        collision_weight=planner._collision_weight, velocity_hessian_weight=round(w_h, 6),
        init_collision_iteration=planner._init_collision_iteration,
        init_collision_points=planner._init_collision_points,
        reparametrize_trajectory_freq=planner._reparametrize_trajectory_freq,
        optimize_collision_model_freq=planner._optimize_collision_model_freq,
        random_field_points=planner._random_field_points, course_random_offset=planner._course_random_offset,
        collision_point_count=planner._collision_point_count)
    if constrained:
        kw.update(angle_weight=planner._angle_weight, constraint_deltas_weight=planner._constraint_delta_weight,
                  multipliers_lr=planner._multipliers_lr, boundary_weight=planner._boundary_weight,
                  collision_multipliers_lr=planner._collision_multipliers_lr, angle_offset=planner._angle_offset,
                  collision_beta=planner._collision_beta, direction_delta_weight=planner._direction_delta_weight)
    return PlannerConfig(**kw)


_PARAM_TO_SEG = {"encoding_layer.weight": "enc_w", "encoding_layer.bias": "enc_b", "mlp.0.weight": "W0",
                 "mlp.0.bias": "b0", "mlp.2.weight": "W1", "mlp.2.bias": "b1", "mlp2.0.weight": "w_out",
                 "mlp2.0.bias": "b_out", "_angle_encoder._biases": "ang_b", "_angle_encoder._frequencies": "ang_f"}


def snapshot(planner) -> dict:
    """Full planner state in the oracle's representation (see OraclePlanner.state)."""
    model = planner._collision_model
    cfg = onf_config_of(model)
    w = O.blob_from_state_dict(model.state_dict(), cfg)
    m_named, v_named, onf_step = {}, {}, 0
    opt = planner._collision_optimizer
    for name, prm in model.named_parameters():
        seg = _PARAM_TO_SEG[name]
        s = opt.state.get(prm, {})
        m_named[seg] = s.get("exp_avg", torch.zeros_like(prm)).detach().reshape(-1)
        v_named[seg] = s.get("exp_avg_sq", torch.zeros_like(prm)).detach().reshape(-1)
        if "step" in s:
            onf_step = max(onf_step, int(s["step"]))
    for d in (m_named, v_named):
        for seg in ("enc_b", "ang_b", "ang_f"):
            d.setdefault(seg, torch.zeros(0 if seg != "enc_b" else cfg.f_pos))
    ts = planner._trajectory_optimizer.state.get(planner._trajectory, {})
    traj = planner._trajectory.detach().clone()
    constrained = hasattr(planner, "_constraint_multipliers")
    T = traj.shape[0]
    return dict(
        w=w, m=O.pack(m_named, cfg), v=O.pack(v_named, cfg), onf_step=onf_step,
        traj=traj, traj_m=ts.get("exp_avg", torch.zeros_like(traj)).detach().clone(),
        traj_v=ts.get("exp_avg_sq", torch.zeros_like(traj)).detach().clone(),
        traj_step=int(ts["step"]) if "step" in ts else 0,
        lam=planner._constraint_multipliers.detach().clone() if constrained else torch.zeros(T + 1),
        mu=planner._collision_multipliers.detach().clone() if constrained else torch.zeros(T),
        prev_traj=None if planner._previous_trajectory is None else planner._previous_trajectory.clone(),
        buf_pos=np.array(planner._collision_positions, dtype=np.float64, copy=True),
        buf_age=np.array(planner._collision_positions_times, dtype=np.float64, copy=True),
        step_count=int(planner._step_count),
        start=planner._start_point[0].detach().clone(), goal=planner._goal_point[0].detach().clone(),
        border=tuple(float(b) for b in planner._random_sample_border))


@contextlib.contextmanager
def record_variates(constrained: bool, T: int, order=None):
    """Patches numpy's legacy global RNG entry points and torch.rand while a reference step() runs and
    records what it draws, keyed as the oracle's sources expect.  `order` overrides the
    names of the numpy draws (warm-up iterations draw only the uniform field points)."""
    rec = {"choice_u": np.zeros(0)}
    if order is None:
        order = (["t_sample", "coarse_xy", "coarse_th", "fine_xy", "fine_th", "field_xy", "field_th"] if constrained
                 else ["t_sample", "coarse_xy", "fine_xy", "field_xy", "t_path"])
    pos = [0]
    real = dict(rand=np.random.rand, randn=np.random.randn, choice=np.random.choice, trand=torch.rand)

    def _next(v):
        rec[order[pos[0]]] = np.array(v, copy=True)
        pos[0] += 1
        return v

    def rand(*shape):
        return _next(real["rand"](*shape))

    def randn(*shape):
        return _next(real["randn"](*shape))

    def choice(n, size, replace=True, p=None):
        clone = np.random.RandomState()
        clone.set_state(np.random.get_state())
        idx = real["choice"](n, size, replace=replace, p=p)
        used = []

        def draw(k):
            u = clone.rand(k)
            used.append(u)
            return u
        mine = S.legacy_choice(p, size, replace, draw)
        assert np.array_equal(mine, idx), "oracle legacy_choice disagrees with numpy RandomState.choice"
        rec["choice_u"] = np.concatenate([rec["choice_u"]] + used)
        rec["choice_idx"] = np.array(idx, copy=True)
        return idx

    def trand(*a, **k):
        v = real["trand"](*a, **k)
        rec["t_path"] = v.numpy().reshape(-1).copy()
        return v

    np.random.rand, np.random.randn, np.random.choice, torch.rand = rand, randn, choice, trand
    try:
        yield rec
    finally:
        np.random.rand, np.random.randn, np.random.choice, torch.rand = (real["rand"], real["randn"],
                                                                           real["choice"], real["trand"])


def reference_labeller(checker, constrained: bool):
    """Labels [N,2|3] float64 poses with the reference's own checker object."""
    ref = load_reference()
    P2 = ref["position2"].Position2

    def label(poses):
        poses = np.asarray(poses, dtype=np.float64)
        return np.asarray(checker.check_collision(P2.from_vec(poses) if constrained else poses), dtype=bool)
    return label

"""GPU: the reference-facing Python API (what the reference's own tests assert, test/test_nerf_opt_planner.py
:28-50 and test/test_constrained_nerf_opt_planner.py:30-50, plus path quality of the batched planner)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RUN_PLANNER = dict(
    device="cuda", trajectory_length=100,
    collision_model=dict(mean=0, sigma=1, use_cos=True, bias=True, use_normal_init=True, angle_encoding=True,
                         name="ONF"),
    trajectory_initializer=dict(name="TrajectoryInitializer", resolution=0.05),
    collision_optimizer=dict(lr=5e-2, betas=(0.9, 0.9)), trajectory_optimizer=dict(lr=1e-2, betas=(0.9, 0.9)),
    planner=dict(name="ConstrainedNERFOptPlanner", trajectory_random_offset=0.02, collision_weight=1,
                 velocity_hessian_weight=0.5, random_field_points=10, init_collision_iteration=0,
                 constraint_deltas_weight=20, multipliers_lr=0.1, init_collision_points=100,
                 reparametrize_trajectory_freq=10, optimize_collision_model_freq=1, angle_weight=0.5,
                 angle_offset=0.3, boundary_weight=1, collision_multipliers_lr=1e-3))


def _attr(d):
    from nfopp_b200.utils.attribute_dict import AttributeDict
    return AttributeDict({k: _attr(v) if isinstance(v, dict) else v for k, v in d.items()})


def test_holonomic_planner_api():
    from nfopp_b200.collision_checker import CircleCollisionChecker
    from nfopp_b200.planner_factory import PlannerFactory
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder
    np.random.seed(400); torch.manual_seed(100)
    env = TestEnvironmentBuilder.make_test_environment()
    checker = CircleCollisionChecker(0.3, (0, 3, 0, 3))
    checker.update_obstacle_points(env.obstacle_points)
    planner = PlannerFactory.make_onf_planner(checker)
    planner.init(env.start_point, env.goal_point, env.bounds)
    path = planner.get_path()
    assert path.shape == (102, 2) and path.dtype == np.float32
    assert np.allclose(path[0], env.start_point) and np.allclose(path[-1], env.goal_point)
    for _ in range(100):
        planner.step()
    path = planner.get_path()
    assert path.shape == (102, 2) and np.isfinite(path).all()
    assert np.allclose(path[0], env.start_point) and np.allclose(path[-1], env.goal_point)
    assert planner._step_count == 100
    assert planner.checked_positions.shape == (209, 2) and planner.truth_collision.shape == (209,)
    assert planner._collision_positions.shape == (100, 2)
    assert planner._collision_model is not None and planner._device.type == "cuda"


def test_constrained_planner_api_and_quality():
    from nfopp_b200.collision_checker import RectangleCollisionChecker
    from nfopp_b200.planner_factory import PlannerFactory
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder
    from nfopp_b200.utils.position2 import Position2
    np.random.seed(400); torch.manual_seed(100)
    env = TestEnvironmentBuilder().make_car_environment()
    checker = RectangleCollisionChecker((-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))
    checker.update_obstacle_points(env.obstacle_points)
    planner = PlannerFactory.make_constrained_onf_planner(checker, _attr(RUN_PLANNER))
    planner.init(env.start_point, env.goal_point, env.bounds)
    assert planner.get_path().shape == (102, 3)
    planner.step()
    w0 = planner._collision_model.mlp[0].weight.detach().clone()
    for _ in range(299):
        planner.step()
    path = planner.get_path()
    assert path.shape == (102, 3) and np.isfinite(path).all()
    assert np.allclose(path[0], env.start_point) and np.allclose(path[-1], env.goal_point)
    # module parameters are live views of the engine's weights
    assert not torch.equal(w0, planner._collision_model.mlp[0].weight.detach())
    # reference path quality after 300 steps (collision-free, non-holonomic residual ~0.03)
    assert not checker.check_collision(Position2.from_vec(path.astype(np.float64))).any()
    dx, dy = np.diff(path[:, 0]), np.diff(path[:, 1])
    abar = path[:-1, 2] + ((np.diff(path[:, 2]) + np.pi) % (2 * np.pi) - np.pi) / 2
    assert np.abs(dx * np.sin(abar) - dy * np.cos(abar)).max() < 0.1
    assert isinstance(planner.checked_positions, Position2)
    # warm re-planning API
    planner.update_goal_point(np.array([2.0, 2.6, 0.0], dtype=np.float32))
    planner.update_start_point(np.array([0.6, 1.5, 0.0], dtype=np.float32))
    assert planner._step_count == 0
    planner.step()
    assert np.isfinite(planner.get_path()).all()


def test_unknown_class_in_config_raises_keyerror():
    from nfopp_b200.collision_checker import CircleCollisionChecker
    from nfopp_b200.planner_factory import PlannerFactory
    bad = _attr(RUN_PLANNER)
    bad["collision_model"]["name"] = "NoSuchField"
    with pytest.raises(KeyError):
        PlannerFactory.make_constrained_onf_planner(CircleCollisionChecker(0.3), bad)


@pytest.mark.parametrize("name", ["corridor", "car"])
def test_distributional_path_quality_matches_oracle(name):
    """SURVEY.md §7 protocol (iv): the loop is chaotic, so free-running parity is distributional.  256 problems
    on the reference's fixed world (own ONF initialisation and RNG stream each), 1000 steps on the GPU; solved
    rate (collision-free final path), path length and non-holonomic residual are compared with the oracle's
    statistics over 32 seeds (tests/golden/quality_*.json, written by oracle/quality_stats.py)."""
    import json, os
    import _util as U
    from _gpu_util import engine_config
    from nfopp_b200.batched import BatchedPlanner
    gold = json.load(open(os.path.join(U.GOLDEN, f"quality_{name}.json")))
    fx = U.load(f"lockstep_{name}.npz")
    P = 256
    cfg = engine_config(name, fx, rng_mode="philox", problems=P, seed=77)
    planner = BatchedPlanner(cfg)
    start, goal = fx["free/state0/start"], fx["free/state0/goal"]
    planner.init(np.tile(start, (P, 1)), np.tile(goal, (P, 1)), obstacles=fx["meta/obstacles"])
    label = U.labeller_for(name, fx)
    for i, step in enumerate(range(250, 1001, 250)):
        planner.step(250)
        paths = planner.get_path().astype(np.float64)
        assert np.isfinite(paths).all() and (planner.status() == 0).all()
        solved = np.array([not label(paths[p]).any() for p in range(P)])
        ref_rate = gold["solved_rate"][str(step)]
        sigma = max(np.sqrt(ref_rate * (1 - ref_rate) / gold["seeds"] + ref_rate * (1 - ref_rate) / P), 0.04)
        assert abs(solved.mean() - ref_rate) < 3.5 * sigma, (step, float(solved.mean()), ref_rate)
    d = np.diff(paths, axis=1)
    length = np.linalg.norm(d[..., :2], axis=2).sum(1)
    abar = paths[:, :-1, 2] + ((d[..., 2] + np.pi) % (2 * np.pi) - np.pi) / 2
    resid = np.abs(d[..., 0] * np.sin(abar) - d[..., 1] * np.cos(abar)).max(1)
    ref_len = np.array([r[-1]["length"] for r in gold["runs"]])
    ref_res = np.array([r[-1]["nonholonomic"] for r in gold["runs"]])
    assert abs(np.median(length) - np.median(ref_len)) < 0.15 * np.median(ref_len)
    assert np.median(resid) < 3 * np.median(ref_res) + 0.02


def test_batched_jittered_corridors_run_through_cuda_graph():
    """64 different corridor problems (own obstacle sets) through the CUDA-graph path: finite paths, pinned
    endpoints, clean status words."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder
    P = 64
    starts, goals, bounds, obstacles = TestEnvironmentBuilder.make_corridor_batch(P, seed=0)
    planner = BatchedPlanner(corridor_config(P, seed=3))
    planner.init(starts, goals, obstacles=obstacles)
    planner.step(100)
    paths = planner.get_path()
    assert paths.shape == (P, 102, 3) and np.isfinite(paths).all()
    assert np.allclose(paths[:, 0], starts) and np.allclose(paths[:, -1], goals)
    assert (planner.status() == 0).all()
    assert planner.engine.counter("step_count") == 100


def test_batched_warm_replanning_matches_single_problem_api():
    """BatchedPlanner.update_goal_points / update_start_points (receding-horizon re-planning for all problems) give,
    problem by problem, what the reference-facing P=1 planner's update_goal_point / update_start_point give."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    P = 4
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=2)
    bp = BatchedPlanner(corridor_config(P, seed=5))
    bp.init(starts, goals, obstacles=obstacles)
    bp.step(30)
    before = bp.engine["traj"].clone()
    lam0, mu0 = bp.engine["lam"].clone().cpu(), bp.engine["mu"].clone().cpu()
    new_goals = goals.copy(); new_goals[:, 0] -= 0.3; new_goals[:, 1] += 0.2
    new_starts = starts.copy(); new_starts[:, 0] += 0.25
    bp.update_goal_points(new_goals)
    bp.update_start_points(new_starts)
    after = bp.engine["traj"].clone()
    assert bp.engine.counter("step_count") == 0 and bool(torch.isfinite(after).all())
    # same operations through single-problem engines
    for p in range(P):
        one = BatchedPlanner(corridor_config(1, seed=5))
        one.init(starts[p:p + 1], goals[p:p + 1], obstacles=obstacles[p:p + 1])
        one.engine["traj"].copy_(before[p:p + 1])
        one.update_goal_points(new_goals[p:p + 1])
        one.update_start_points(new_starts[p:p + 1])
        assert torch.equal(one.engine["traj"][0], after[p]), p
    # and the oracle's restatement of constrained_nerf_opt_planner.py:178-194 + :132-171
    from oracle import trajectory as TR
    for p in range(P):
        traj, lam, mu = before[p].cpu().clone(), lam0[p].clone(), mu0[p].clone()
        st, go = torch.tensor(starts[p]), torch.tensor(new_goals[p])
        k = min(int(torch.argmin(((traj[:, :2] - go[None, :2]) ** 2).sum(1))) + 1, traj.shape[0])
        traj[k:] = go
        traj, lam, mu = TR.reparam_constrained(traj, st, go, lam, mu)
        st = torch.tensor(new_starts[p])
        k = min(int(torch.argmin(((traj[:, :2] - st[None, :2]) ** 2).sum(1))) + 1, traj.shape[0])
        traj[:k] = st
        traj, lam, mu = TR.reparam_constrained(traj, st, go, lam, mu)
        assert float((traj - after[p].cpu()).abs().max()) < 1e-5, p
        assert float((lam - bp.engine["lam"][p].cpu()).abs().max()) < 1e-5
    paths = bp.get_path()
    assert np.allclose(paths[:, 0], new_starts) and np.allclose(paths[:, -1], new_goals)
    bp.step(5)
    assert np.isfinite(bp.get_path()).all()


@pytest.mark.gpu
def test_lanes_are_bit_identical_to_a_single_stream():
    """nfopp_engine_set_lanes: the graphed step split into parallel problem ranges gives, bit for bit, the state of the
    single-stream graph (per-problem RNG streams keep their global problem index) - for even and uneven splits."""
    import torch
    from nfopp_b200.batched import BatchedPlanner, corridor_config, init_onf_weights
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    P = 70
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=4)
    cfg = corridor_config(P, seed=8)
    w0 = init_onf_weights(cfg.onf, P, True, 5, "cuda")
    states = []
    for lanes in (1, 2, 3):
        bp = BatchedPlanner(cfg, lanes=lanes)
        bp.init(starts, goals, obstacles=obstacles, onf_weights=w0)
        bp.step(14)
        torch.cuda.synchronize()
        assert bp.engine.counter("step_count") == 14
        states.append({k: bp.engine[k].clone() for k in ("traj", "onf_w", "onf_m", "lam", "mu", "buf_pos", "train_y")})
        bp.engine.close()
    for st in states[1:]:
        for k, v in st.items():
            assert torch.equal(v, states[0][k]), k


@pytest.mark.gpu
def test_lanes_on_the_wide_onf_path():
    """lanes with the wide (streamed-weights) ONF: every lane gets its own activation workspace; the lanes' results agree
    with the single-stream run to rounding (the weight-gradient split-K count depends on the problems per launch, so the
    summation order - not the algorithm - may differ) and an over-subscribed lane count is refused, not mis-run."""
    import torch
    from nfopp_b200.batched import BatchedPlanner, gridmap_wide_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    grid, _, cell = B.make_blob_grid(200, 200, 0.1, 0.015, blob_cells=12, seed=2)
    P = 12
    starts, goals = B.free_endpoints(grid, cell, P, seed=3, margin_cells=10)
    outs = []
    for lanes in (1, 2, 3):
        cfg = gridmap_wide_config(P, rows=200, cols=200, cell=cell, seed=9, hidden=128, layers=3)
        bp = BatchedPlanner(cfg, lanes=lanes)
        bp.init(starts, goals, grids=grid)
        bp.step(3)
        torch.cuda.synchronize()
        outs.append((bp.paths().clone(), bp.engine["onf_w"].clone(), bp.engine["train_y"].clone()))
        assert bool(torch.isfinite(outs[-1][0]).all()) and int(np.abs(bp.status()).max()) == 0
        bp.engine.close()
    for paths, w, y in outs[1:]:
        assert torch.equal(y, outs[0][2])                                   # same samples, same labels
        assert float((paths - outs[0][0]).abs().max()) < 1e-4
        assert float((w - outs[0][1]).abs().max()) < 1e-3 * float(outs[0][1].abs().max())


@pytest.mark.gpu
def test_host_step_pipeline_equals_synchronous_copies():
    """HostStepPipeline (inputs staged on a copy stream, paths read back on another, alternating buffers) gives, bit for bit,
    the paths of copying each step's obstacle set / start / goal synchronously and reading the paths back with get_path()
    - with the obstacle sets CHANGING every step, so a copy that ran a step early or late would show."""
    import torch
    from nfopp_b200.batched import BatchedPlanner, HostStepPipeline, corridor_config, init_onf_weights
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    P, K = 40, 9
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=11)
    cfg = corridor_config(P, seed=3)
    w0 = init_onf_weights(cfg.onf, P, True, 6, "cuda")
    rng = np.random.default_rng(0)
    moving = [np.asarray(obstacles) + rng.normal(0.0, 0.05, np.asarray(obstacles).shape) for _ in range(K)]

    sync = BatchedPlanner(cfg)
    sync.init(starts, goals, obstacles=obstacles, onf_weights=w0)
    want = []
    for k in range(K):
        sync.engine["obstacles"].copy_(torch.as_tensor(moving[k]).to(sync.engine["obstacles"].dtype).reshape(sync.engine["obstacles"].shape))
        sync.step(1)
        want.append(sync.get_path())
    sync.engine.close()

    piped = BatchedPlanner(cfg)
    piped.init(starts, goals, obstacles=obstacles, onf_weights=w0)
    pipe = HostStepPipeline(piped, ["obstacles", "start", "goal"])
    h_start, h_goal = pipe.pinned("start", piped.engine["start"].cpu()), pipe.pinned("goal", piped.engine["goal"].cpu())
    h_obs = [pipe.pinned("obstacles", m) for m in moving]
    got = []
    for k in range(K):
        out = pipe.step({"obstacles": h_obs[k], "start": h_start, "goal": h_goal})
        if k % 3 == 2 or k == K - 1:                      # read some steps back without draining the pipeline every step
            pipe.synchronize()
            got.append((k, out.numpy().copy()))
    assert pipe.h2d_bytes > 0 and pipe.d2h_bytes == want[0].nbytes
    for k, paths in got:
        assert np.array_equal(paths, want[k]), k
    piped.engine.close()


@pytest.mark.gpu
def test_init_angles_with_trajectory_on_the_device():
    """init_angles_with_trajectory=True through BatchedPlanner.init (A* x, y, then headings pulled towards the direction of
    travel): equals the host helper applied to the same trajectories, end headings untouched, planner steps stay finite."""
    import torch
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from nfopp_b200.trajectory_initializer import blend_headings_with_path_direction
    P = 8
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=31)
    plain = BatchedPlanner(corridor_config(P, seed=1))
    plain.init(starts, goals, obstacles=obstacles, astar_resolution=0.25)
    want = blend_headings_with_path_direction(plain.engine["traj"].cpu().clone(), plain.engine["start"].cpu(), plain.engine["goal"].cpu())
    bp = BatchedPlanner(corridor_config(P, seed=1))
    bp.init(starts, goals, obstacles=obstacles, astar_resolution=0.25, init_angles_with_trajectory=True)
    got = bp.engine["traj"].cpu()
    assert torch.allclose(got, want, atol=2e-6)
    assert not torch.allclose(got[..., 2], plain.engine["traj"].cpu()[..., 2], atol=1e-3)      # the S-bend path does turn
    assert torch.equal(got[..., :2], plain.engine["traj"].cpu()[..., :2])
    bp.step(20)
    assert bool(torch.isfinite(bp.paths()).all())
    plain.engine.close()
    bp.engine.close()


# ---------------------------------------------------------------------------------------- drop-in surface between steps
def _car_planner(**planner_over):
    from nfopp_b200.collision_checker import RectangleCollisionChecker
    from nfopp_b200.planner_factory import PlannerFactory
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder
    np.random.seed(400); torch.manual_seed(100)
    env = TestEnvironmentBuilder().make_car_environment()
    checker = RectangleCollisionChecker((-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))
    checker.update_obstacle_points(env.obstacle_points)
    params = _attr(RUN_PLANNER)
    params["planner"].update(planner_over)
    planner = PlannerFactory.make_constrained_onf_planner(checker, params)
    return planner, env, checker


@pytest.mark.gpu
def test_environment_changes_between_steps_reach_the_engine():
    """The reference re-reads obstacle points and boundaries on every check_collision call (collision_checker/*.py;
    ros/collision_checker_adapter.py feeds a lidar cloud of a different size on every callback): a cloud that outgrows the
    engine's obstacle buffer and new checker boundaries must both show up in the next step's labels."""
    from oracle import checkers as OC
    planner, env, checker = _car_planner()
    planner.init(env.start_point, env.goal_point, env.bounds)
    planner.step()
    cap = planner._engine.config.max_obstacles
    rs = np.random.RandomState(3)
    cloud = np.concatenate([env.obstacle_points, rs.uniform(0.2, 2.8, size=(3 * cap, 2))], 0)     # larger than the buffer
    checker.update_obstacle_points(cloud)
    planner.step()
    assert planner._engine.config.max_obstacles >= len(cloud)
    poses, labels = planner._checked()
    assert np.array_equal(labels, OC.rectangle_labels(poses, cloud, (-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3)))
    assert planner._step_count == 2 and planner._engine.counter("onf_step") == 2                  # state carried over
    eng_before = planner._engine
    checker.update_boundaries((0.5, 2.5, 0.5, 2.5))
    planner.step()
    assert planner._engine is eng_before                                                          # bounds only: no re-creation
    poses, labels = planner._checked()
    assert np.array_equal(labels, OC.rectangle_labels(poses, cloud, (-0.3, 0.2, -0.3, 0.2), (0.5, 2.5, 0.5, 2.5)))
    assert labels[(poses[:, 0] < 0.5) | (poses[:, 0] > 2.5)].all()
    checker.update_obstacle_points(env.obstacle_points)                                           # and back to a small cloud
    planner.step()
    poses, labels = planner._checked()
    assert np.array_equal(labels, OC.rectangle_labels(poses, env.obstacle_points, (-0.3, 0.2, -0.3, 0.2), (0.5, 2.5, 0.5, 2.5)))


@pytest.mark.gpu
def test_second_init_keeps_optimiser_state_and_replay_buffer():
    """nerf_opt_planner.py:181-187: init() re-initialises the trajectory, warms the ONF up and resets step_count - the
    optimisers' moments and step counters, the replay buffer, _previous_trajectory and the multipliers persist across goals
    (ros/goal_planner_adapter.py:36 calls init() for every new goal)."""
    planner, env, checker = _car_planner(init_collision_iteration=2)
    planner.init(env.start_point, env.goal_point, env.bounds)
    for _ in range(12):
        planner.step()
    eng = planner._engine
    before = {k: eng[k].clone() for k in ("onf_m", "onf_v", "traj_m", "traj_v", "buf_pos", "buf_age", "lam", "mu", "prev_traj")}
    w_before = eng["onf_w"].clone()
    onf_step, traj_step = eng.counter("onf_step"), eng.counter("traj_step")
    assert onf_step == 14 and traj_step == 12 and int(eng["buf_n"][0]) == 100
    new_goal = np.array([2.2, 2.5, 0.3], dtype=np.float32)
    planner.init(env.start_point, new_goal, env.bounds)
    assert planner._engine is eng                                       # no re-allocation, no graph re-capture
    assert planner._step_count == 0
    assert eng.counter("onf_step") == onf_step + 2                      # the warm-up continued the ONF optimiser
    assert eng.counter("traj_step") == traj_step and eng.counter("has_prev") == 1
    for k in ("traj_m", "traj_v", "buf_pos", "buf_age", "lam", "mu", "prev_traj"):
        assert torch.equal(eng[k], before[k]), k
    assert not torch.equal(eng["onf_m"], before["onf_m"]) and not torch.equal(eng["onf_w"], w_before)   # two warm-up updates
    path = planner.get_path()
    assert np.allclose(path[0], env.start_point) and np.allclose(path[-1], new_goal)
    assert np.allclose(path[1:-1, 0], np.linspace(env.start_point[0], new_goal[0], 102)[1:-1], atol=1e-6)
    planner.step()
    assert np.isfinite(planner.get_path()).all() and eng.counter("traj_step") == traj_step + 1


@pytest.mark.gpu
def test_live_optimisers_boundaries_and_unsupported_checkers():
    """Callers of the reference own the torch.optim.Adam objects (planner_factory.py:54-58,68-70): a learning rate changed
    between steps is picked up; set_boundaries (nerf_opt_planner.py:220-222) does not re-create the engine; a checker the
    CUDA step cannot run is refused with a TypeError naming the supported ones; an untouched ONF is not re-packed."""
    from nfopp_b200.planner_factory import PlannerFactory
    planner, env, checker = _car_planner()
    planner.init(env.start_point, env.goal_point, env.bounds)
    planner.step()
    launches = planner._engine.launches
    planner.step()
    per_step = planner._engine.launches - launches
    planner.step()
    assert planner._engine.launches - launches == 2 * per_step          # no extra pack launch when nobody wrote the weights
    with torch.no_grad():
        planner._collision_model.mlp[0].bias.add_(0.0)                  # an in-place write through the module's view
    planner.step()
    assert planner._engine.launches - launches == 3 * per_step + 1      # operand copy rebuilt once
    eng = planner._engine
    planner.set_boundaries((0.2, 2.8, 0.2, 2.8))
    assert planner._engine is eng and planner._step_count == 0
    planner.step()
    uni = planner._checked()[0][-10:]                                   # the step's 10 uniform field points
    assert (uni[:, :2] >= 0.2).all() and (uni[:, :2] <= 2.8).all()
    planner._collision_optimizer.param_groups[0]["lr"] = 0.0
    w = planner._engine["onf_w"].clone()
    planner.step()
    assert torch.equal(planner._engine["onf_w"], w), "lr = 0 on the caller's optimiser must freeze the ONF weights"
    assert planner._engine.counter("step_count") == 2

    class Homemade:
        def check_collision(self, positions):
            return np.zeros(len(positions), dtype=bool)
    bad = PlannerFactory.make_constrained_onf_planner(Homemade(), _attr(RUN_PLANNER))
    with pytest.raises(TypeError, match="RectangleCollisionChecker"):
        bad.init(env.start_point, env.goal_point, env.bounds)

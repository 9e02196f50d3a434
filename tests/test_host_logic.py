"""CPU: host-side logic, C-ABI surface and multi-rank sharding (no GPU compute calls)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import _util as U
from oracle import onf as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    from nfopp_b200 import _lib as L
    header = open(os.path.join(ROOT, "include", "nfopp.h")).read()
    declared = set(re.findall(r"\b(nfopp_[a-z_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    handle = L.lib()
    for name in declared:
        assert hasattr(handle, name), f"{name} declared in include/nfopp.h but not exported"
    assert declared == set(L.SIGNATURES), declared ^ set(L.SIGNATURES)
    assert handle.nfopp_abi_version() == 5


@pytest.mark.parametrize("name", list(U.ONF_CFGS))
def test_blob_layout_matches_oracle(name):
    from nfopp_b200.engine import OnfSpec, onf_segments
    cfg = U.ONF_CFGS[name]
    segs, total = onf_segments(OnfSpec(cfg.mean, cfg.sigma, cfg.use_cos, cfg.angle, cfg.hidden, cfg.layers))
    lay, ototal = O.onf_layout(cfg)
    assert total == ototal
    assert {k: (o, tuple(s)) for k, (o, s) in segs.items()} == {k: (o, tuple(s)) for k, (o, s) in lay.items()}


def test_invalid_configs_are_rejected_with_valueerror():
    from nfopp_b200 import _lib as L
    from nfopp_b200.engine import EngineConfig, OnfSpec
    n = C.c_size_t()
    for bad in (EngineConfig(onf=OnfSpec(hidden=30)), EngineConfig(trajectory_length=1),
                EngineConfig(constrained=False, onf=OnfSpec(angle=True)),
                EngineConfig(constrained=False, checker="rectangle")):
        c = bad.to_c()
        with pytest.raises(ValueError):
            L.check(L.lib().nfopp_engine_workspace_bytes(C.byref(c), C.byref(n)))
    ok = EngineConfig(problems=1024, onf=OnfSpec(use_cos=True, angle=True)).to_c()
    L.check(L.lib().nfopp_engine_workspace_bytes(C.byref(ok), C.byref(n)))
    assert 1024 * 33164 * 4 * 3 < n.value < 4 * 2 ** 30


def test_no_cuda_means_loud_failure_not_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nfopp_b200 import _lib as L
    from nfopp_b200.engine import EngineConfig, PlannerEngine
    from nfopp_b200.onf_model import ONF
    with pytest.raises(L.NfoppError):
        PlannerEngine(EngineConfig())
    with pytest.raises(L.NfoppError):
        ONF(0, 1)(torch.zeros(4, 2))


def test_onf_module_matches_reference_constructor_stream():
    """Same torch seed => same initial weights as the reference's ONF (construction order is preserved);
    checked against the golden weights the reference module produced (oracle/make_golden.py seeds 10,11,12)."""
    from nfopp_b200.onf_model import ONF
    fx = U.load("onf.npz")
    cases = {"holonomic": (10, dict(mean=1.5, sigma=1)),
             "car": (11, dict(mean=0, sigma=1, use_cos=True, use_normal_init=True, angle_encoding=True)),
             "image": (12, dict(mean=0, sigma=7, use_cos=True, use_normal_init=True))}
    for name, (seed, kw) in cases.items():
        torch.manual_seed(seed)
        model = ONF(**kw)
        assert np.array_equal(model.packed_parameters().numpy(), fx[f"{name}/w"]), name
        sd = model.state_dict()
        model2 = ONF(**kw)
        model2.load_state_dict(sd)
        assert torch.equal(model2.packed_parameters(), model.packed_parameters())


def test_batched_linspace_equals_torch_linspace():
    from nfopp_b200.engine import batched_linspace
    rs = np.random.RandomState(0)
    a, b = torch.tensor(rs.uniform(-5, 5, 64).astype(np.float32)), torch.tensor(rs.uniform(-5, 5, 64).astype(np.float32))
    got = batched_linspace(a, b, 102)
    for i in range(64):
        assert torch.equal(got[i], torch.linspace(a[i], b[i], 102)), i


def test_environment_builder_matches_reference_geometry():
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    fx = U.load("checkers.npz")
    assert np.array_equal(B._obstacle_points(), fx["obstacles_sbend"])
    assert np.array_equal(B._car_obstacle_points(), fx["obstacles_car"])
    env = B().make_car_environment()
    assert env.start_point.dtype == np.float32 and tuple(env.bounds) == (-0.1, 3.1, -0.1, 3.1)
    s, g, bounds, obs = B.make_corridor_batch(5, seed=1)
    s2, g2, _, obs2 = B.make_corridor_batch(5, seed=1)
    assert np.array_equal(obs, obs2) and obs.shape == (5, 20, 2) and not np.array_equal(obs[0], obs[1])


def test_universal_factory_semantics():
    from nfopp_b200.utils.attribute_dict import AttributeDict
    from nfopp_b200.utils.universal_factory import UniversalFactory

    class Thing:
        def __init__(self, a, b=2):
            self.a, self.b = a, b
    f = UniversalFactory([Thing])
    t = f.make_from_parameters(AttributeDict(name="Thing", a=1, ignored=5), b=7)
    assert (t.a, t.b) == (1, 7)
    with pytest.raises(KeyError, match="Unknown class"):
        f.make_from_parameters(dict(name="Other"))
    assert f.make_from_parameters(3) == 3


def test_numpy_pairwise_sum_restatement():
    """The CUDA resampler restates numpy's pairwise summation (sample.cu np_pairwise_sum); this is the same
    algorithm in Python, checked against np.sum for the sizes the planner uses."""
    def pw(a):
        n = len(a)
        if n < 8:
            r = 0.0
            for v in a:
                r += v
            return r
        if n <= 128:
            r = list(a[:8])
            i = 8
            while i < n - n % 8:
                for j in range(8):
                    r[j] += a[i + j]
                i += 8
            res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
            for v in a[i:]:
                res += v
            return res
        n2 = n // 2
        n2 -= n2 % 8
        return pw(a[:n2]) + pw(a[n2:])
    rs = np.random.RandomState(1)
    for n in (5, 99, 100, 128, 129, 198, 199, 300, 1000):
        a = rs.gamma(0.3, size=n)
        assert pw(a) == np.sum(a), n


def test_shard_range_partitions_all_problems():
    from nfopp_b200.batched import shard_range
    for total in (1, 7, 1024, 16384, 16385):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _gather_worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nfopp_b200.batched import BatchedPlanner, shard_range
    lo, hi = shard_range(5, rank, world)

    class FakeEngine:
        def paths(self):
            return torch.arange(lo, hi, dtype=torch.float32)[:, None, None].expand(hi - lo, 4, 3).contiguous()
    bp = BatchedPlanner.__new__(BatchedPlanner)
    bp.engine = FakeEngine()
    got = bp.gather_paths()
    torch.save(got, f"{out}.{rank}")
    dist.destroy_process_group()


def test_gather_paths_world_size_2_gloo(tmp_path):
    """The one collective of the job (final path gather), uneven shards, over gloo on CPU."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "g")
    mp.spawn(_gather_worker, args=(2, 29533, out), nprocs=2, join=True)
    for r in range(2):
        got = torch.load(f"{out}.{r}")
        assert got.shape == (5, 4, 3)
        assert torch.equal(got[:, 0, 0], torch.arange(5, dtype=torch.float32))


def test_oracle_path_evaluation_and_stop_rule():
    """oracle.evaluate: interpolation (heading along the wrapped difference), length, and the best-path / early-stop
    bookkeeping of scripts/run_bench_mr.py:109-131 on a scripted evaluation history."""
    import numpy as np
    from oracle import evaluate as EV
    path = np.array([[0, 0, 3.0], [1, 0, -3.0], [1, 2, -3.0]], dtype=np.float32)
    poses = EV.interpolate_path(path, 4)
    assert poses.shape == (9, 3) and np.allclose(poses[-1], path[-1])
    assert np.allclose(poses[2, :2], [0.5, 0.0])
    assert abs(poses[2, 2] - (3.0 + 0.5 * ((-6.0 + np.pi) % (2 * np.pi) - np.pi))) < 1e-12      # through +-pi, not through 0
    hit = lambda p: (p[:, 0] > 0.9) & (p[:, 1] > 0.9) & (p[:, 1] < 1.1)                          # noqa: E731
    c, ln, n = EV.evaluate_path(path, hit, 4)
    assert c and n == 1 and abs(ln - 3.0) < 1e-12
    assert EV.evaluate_path(path, hit, 1) == (False, 3.0, 0)                                      # node-only check misses it
    history = iter([(True, 9.0), (False, 5.0), (False, 4.0), (False, 4.5), (False, 1.0)])
    state = {"n": 0}
    out = EV.solve_loop(lambda: state.__setitem__("n", state["n"] + 1), lambda: state["n"], lambda p: next(history),
                        max_iterations=1000, min_iterations=200, check_every=50)
    # checks at i = 250, 300, 350, 400 (0-based): collides, 5.0 (best), 4.0 (best), 4.5 -> break; final evaluation = 1.0
    assert out[3] == 401 and out[1] is False and out[2] == 1.0 and out[0] == 401


def test_grid_map_point_cloud_known_answers():
    """GridMap (ros/grid_map.py:6-40): occupied-cell centres in (x, y) = (column, row) order through the origin pose,
    map boundaries, OccupancyGrid payload decoding (-1 = unknown -> free, 0..100 -> 0..1, threshold 0.5).  The point
    order and values were checked bit-for-bit against the live reference when this test was written."""
    from nfopp_b200.grid_map import GridMap
    from nfopp_b200.utils.position2 import Position2
    g = GridMap.from_occupancy_data(np.array([0, 100, -1, 50, 51, 0]), 3, 2, 0.5, Position2(1.0, 2.0, 0.0))
    assert np.array_equal(g.as_point_cloud(), np.array([[1.75, 2.25], [1.75, 2.75]]))
    assert g.boundaries == (1.0, 2.5, 2.0, 3.0)
    assert g.as_point_cloud() is g.as_point_cloud()                      # cached, as in the reference
    m = np.zeros((4, 5), dtype=np.float32)
    m[3, 0] = m[0, 4] = 1.0
    pts = GridMap(m, 0.2, Position2(0.0, 0.0, np.pi / 2)).as_point_cloud()    # rotated origin: (x, y) -> (-y, x)
    assert np.allclose(pts, np.array([[-0.1, 0.9], [-0.7, 0.1]]), atol=1e-12)


def test_corridor_maze_generator_is_seeded_and_connected():
    """make_corridor_grid_batch (bench-mr-style corridor mazes, own generator): deterministic per (seed, problem index),
    start and goal lie in free cells of the same connected component, corridors are 2 * radius + 1 cells wide."""
    from collections import deque
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    s1, g1, bounds, grids, origin, cell = B.make_corridor_grid_batch(3, seed=2)
    s2, g2, _, grids2, _, _ = B.make_corridor_grid_batch(2, seed=3)
    assert np.array_equal(grids[1], grids2[0]) and np.array_equal(s1[1], s2[0]) and np.array_equal(g1[1], g2[0])
    assert bounds == (0.0, 3.0, 0.0, 3.0) and grids.shape == (3, 100, 100) and grids.dtype == np.uint8
    for i in range(3):
        g = grids[i]
        assert 0.2 < g.mean() < 0.8
        rc = lambda pt: (int(pt[1] / cell), int(pt[0] / cell))
        (r0, c0), (r1, c1) = rc(s1[i]), rc(g1[i])
        assert not g[r0 - 3:r0 + 4, c0 - 3:c0 + 4].any() and not g[r1 - 3:r1 + 4, c1 - 3:c1 + 4].any()
        seen = np.zeros_like(g, dtype=bool)
        seen[r0, c0] = True
        todo = deque([(r0, c0)])
        while todo:
            r, c = todo.popleft()
            for dr, dc in ((1, 0), (-1, 0), (0, 1), (0, -1)):
                rr, cc = r + dr, c + dc
                if 0 <= rr < 100 and 0 <= cc < 100 and not g[rr, cc] and not seen[rr, cc]:
                    seen[rr, cc] = True
                    todo.append((rr, cc))
        assert seen[r1, c1]


def test_trajectory_initializer_with_path_direction_headings_matches_reference():
    """TrajectoryInitializer(init_angles_with_trajectory=True) (trajectory_initializer.py:12-44) vs the live reference's
    outputs (tests/golden/init_angles.npz): straight-line initialisation and heading initialisation over curved x, y, even
    and odd trajectory lengths; and the batched form used by BatchedPlanner gives the same rows."""
    from nfopp_b200.trajectory_initializer import TrajectoryInitializer, blend_headings_with_path_direction
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "init_angles.npz"))
    rows = []
    for name in fx["names"]:
        traj = torch.tensor(fx[f"{name}/before"])
        start, goal = torch.tensor(fx[f"{name}/start"]), torch.tensor(fx[f"{name}/goal"])
        init = TrajectoryInitializer(None, init_angles_with_trajectory=True)
        if bool(fx[f"{name}/curved"]):
            init.initialize_angle(traj, start, goal)
        else:
            init.initialize_trajectory(traj, start, goal)
        assert np.abs(traj.numpy() - fx[f"{name}/after"]).max() <= 2e-6, name
        rows.append((name, traj))
    # batch of two equal-length problems through the batched helper
    a, b = "line_even", "curved"
    plain = TrajectoryInitializer(None)
    batch = []
    for name in (a, b):
        t = torch.tensor(fx[f"{name}/before"])
        if bool(fx[f"{name}/curved"]):
            plain.initialize_angle(t, torch.tensor(fx[f"{name}/start"]), torch.tensor(fx[f"{name}/goal"]))
        else:
            plain.initialize_trajectory(t, torch.tensor(fx[f"{name}/start"]), torch.tensor(fx[f"{name}/goal"]))
        batch.append(t)
    batch = torch.stack(batch)
    starts = torch.cat([torch.tensor(fx[f"{n}/start"]) for n in (a, b)])
    goals = torch.cat([torch.tensor(fx[f"{n}/goal"]) for n in (a, b)])
    blend_headings_with_path_direction(batch, starts, goals)
    for i, name in enumerate((a, b)):
        assert np.abs(batch[i].numpy() - fx[f"{name}/after"]).max() <= 2e-6, name

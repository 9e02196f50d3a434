"""CPU: the oracle restatement against the golden vectors produced by the unmodified reference
(oracle/make_golden.py).  This is what pins the oracle on machines without /root/reference."""
import os

import numpy as np
import pytest
import torch

from oracle import checkers as C
from oracle import onf as O
from oracle import sampling as S
from oracle.planner import RecordedSource

import _util as U


def test_checker_labels_bit_exact():
    fx = U.load("checkers.npz")
    poses, sb, car = fx["poses"], fx["obstacles_sbend"], fx["obstacles_car"]
    b = (0, 3, 0, 3)
    assert np.array_equal(C.circle_labels(poses, sb, 0.3, b), fx["circle_sbend"])
    assert np.array_equal(C.circle_labels(poses, car, 0.3, b), fx["circle_directed_car"])
    assert np.array_equal(C.circle_labels(poses, car, 0.25, None), fx["circle_car_nobounds"])
    assert np.array_equal(C.rectangle_labels(poses, car, (-0.3, 0.2, -0.3, 0.2), b), fx["rect_car"])
    assert np.array_equal(C.rectangle_labels(poses, sb, (-0.2, 0.2, -0.2, 0.2), b), fx["rect_sbend"])
    assert np.array_equal(C.boundary_labels(poses[:, :2], b), fx["boundary_only"])
    ox, oy, cell = fx["grid_meta"]
    assert np.array_equal(C.grid_labels(fx["grid_poses"], fx["grid"], ox, oy, cell), fx["grid_labels"])


@pytest.mark.parametrize("name", ["holonomic", "car", "image"])
def test_onf_forward_backward(name):
    fx = U.load("onf.npz")
    cfg = U.ONF_CFGS[name]
    w = torch.tensor(fx[f"{name}/w"])
    x, y = torch.tensor(fx[f"{name}/x"]), torch.tensor(fx[f"{name}/y"])
    z, cache = O.onf_forward(w, cfg, x)
    assert U.relerr(z, fx[f"{name}/z"]) < 1e-5
    _, gz = O.bce_with_logits_grad(z, y)
    gz = gz + 0.01 * torch.sigmoid(z)
    gw, gx = O.onf_backward(w, cfg, cache, gz)
    assert U.relerr(gw, fx[f"{name}/gw"]) < 1e-5
    assert U.relerr(gx, fx[f"{name}/gx"]) < 1e-5


def test_legacy_choice_indices_bit_exact():
    fx = U.load("choice.npz")
    for case in sorted({k.split("/")[0] for k in fx.files}):
        reader = S.UniformReader(fx[f"{case}/u"])
        idx = S.legacy_choice(fx[f"{case}/p"], 100, case.endswith("_1"), reader)
        assert np.array_equal(idx, fx[f"{case}/idx"]), case
        assert reader.used == len(fx[f"{case}/u"])


@pytest.mark.parametrize("name", ["car", "corridor", "holonomic", "corridor_freq"])
def test_lockstep_against_reference_states(name):
    fx = U.load(f"lockstep_{name}.npz")
    for k in fx["meta/steps"]:
        op = U.oracle_planner(name, fx, f"step{k}/before")
        var = U.variates_from(fx, f"step{k}/var")
        op.step(RecordedSource(var))
        after = U.state_from(fx, f"step{k}/after")
        if "t_sample" in var:                      # a field-training step (nerf_opt_planner.py:62)
            assert np.array_equal(op.last["train_labels"], fx[f"step{k}/labels"])
            assert np.array_equal(op.last["train_positions"], fx[f"step{k}/checked"])
        else:
            assert "train_labels" not in op.last and op.state["onf_step"] == U.state_from(fx, f"step{k}/before")["onf_step"]
        if "choice_idx" in var:
            assert np.array_equal(op.last["choice_idx"], var["choice_idx"])
        for key in ["w", "m", "v", "traj", "traj_m", "traj_v", "lam", "mu", "buf_pos", "buf_age"]:
            # the very first Adam step moves every weight by lr * g / (|g| + eps): rounding of g shows up in w
            assert U.relerr(op.state[key], after[key]) < (5e-6 if key == "w" and k == 0 else 2e-6), (k, key)
        for key in ["onf_step", "traj_step", "step_count"]:
            assert op.state[key] == after[key]


@pytest.mark.parametrize("name", ["car", "corridor", "holonomic"])
def test_short_free_run_matches_reference_path(name):
    """10 steps from the reference's initial state with the reference's variates (the loop is chaotic beyond
    that); tolerance = north_star's path tolerance 1e-2 m / 1e-2 rad with margin."""
    fx = U.load(f"lockstep_{name}.npz")
    op = U.oracle_planner(name, fx, "free/state0")
    for k in range(10):
        op.step(RecordedSource(U.variates_from(fx, f"free/var{k}")))
        assert np.abs(op.get_path() - fx[f"free/path{k}"]).max() < 1e-3, k


@pytest.mark.parametrize("name", ["holonomic", "car"])
def test_warmup_against_reference_states(name):
    """_init_collision_model (nerf_opt_planner.py:197-200): every warm-up iteration of the reference - 100 uniform field
    points (:135-141; constrained :173-176), labels, one BCE/Adam update (:76-91) - and the first step after it."""
    fx = U.load(f"warmup_{name}.npz")
    op = U.oracle_planner(name, fx, "warm/before")
    n = int(fx["meta/init_collision_points"])
    for i in range(int(fx["meta/iterations"])):
        src = RecordedSource(U.variates_from(fx, f"warm/var{i}"))
        op._train_field(src, op._field_points(src, n, "warm"))
        after = U.state_from(fx, f"warm/after{i}")
        assert np.array_equal(op.last["train_positions"], fx[f"warm/checked{i}"])
        assert np.array_equal(op.last["train_labels"], fx[f"warm/labels{i}"])
        for key in ["w", "m", "v"]:
            assert U.relerr(op.state[key], after[key]) < 2e-6, (i, key)
        assert op.state["onf_step"] == after["onf_step"] == i + 1
        assert op.state["traj_step"] == 0 and op.state["prev_traj"] is None
    op.state["step_count"] = 0
    op.step(RecordedSource(U.variates_from(fx, "step0/var")))
    after = U.state_from(fx, "step0/after")
    for key in ["w", "m", "v", "traj", "traj_m", "traj_v", "lam", "mu", "buf_pos", "buf_age"]:
        assert U.relerr(op.state[key], after[key]) < 2e-6, key
    assert (op.state["onf_step"], op.state["traj_step"]) == (after["onf_step"], after["traj_step"])


def test_path_statistics_known_answers():
    """oracle.path_stats on closed-form paths: unit half circle (curvature 1, turn pi), a zig-zag with duplicate nodes
    (BenchmarkAdapter.cpp:105-135 skips duplicates: two cusps at the 60 degree threshold)."""
    from oracle.path_stats import path_statistics
    t = np.linspace(0, np.pi, 200)
    s = path_statistics(np.stack([np.cos(t), np.sin(t), t], 1), kind="circle", obstacles=np.zeros((1, 2)), radius=0.3,
                        bounds=(-2, 2, -2, 2))
    assert abs(s["max_curvature"] - 1) < 1e-6 and abs(s["path_length"] - np.pi) < 1e-3 and s["cusps"] == 0
    assert abs(s["aol"] * s["path_length"] - np.pi * 198 / 199) < 1e-9 and abs(s["min_clearing_distance"] - 0.7) < 1e-12
    z = path_statistics(np.array([[0, 0], [1, 0], [1, 0], [0.2, 0.1], [2, 2.0]]))
    assert z["cusps"] == 2 and z["path_length"] > 4.4
    line = path_statistics(np.stack([np.linspace(0, 3, 20), np.zeros(20)], 1))
    assert line["max_curvature"] == 0 and line["smoothness"] == 0 and line["cusps"] == 0


def test_astar_oracle_matches_reference_fixture():
    """oracle/astar.py against tests/golden/astar.npz (recorded from the live reference's AstarTrajectoryInitializer and
    scipy's interp1d): labelled matrix and cell path exact, trajectory bit-identical after the float32 cast."""
    from oracle import astar as A
    fx = U.load("astar.npz")
    for name in [str(n) for n in fx["names"]]:
        g = lambda k: fx[f"{name}/{k}"]                                      # noqa: E731
        bounds, res, obs = tuple(float(b) for b in g("bounds")), float(g("resolution")), g("obstacles")
        label = (lambda p: C.rectangle_labels(p, obs, (-0.3, 0.2, -0.3, 0.2), bounds)) if str(g("kind")) == "rectangle" else \
            (lambda p: C.circle_labels(p, obs, 0.3, bounds))
        xy, cells, matrix = A.initial_trajectory_xy(g("start"), g("goal"), bounds, res, label, int(g("T")))
        assert np.array_equal(matrix, g("matrix")) and np.array_equal(cells, g("cells"))
        assert np.array_equal(A.cells_to_points(cells, bounds, res), g("cell_points"))
        assert np.array_equal(xy, g("trajectory")[:, :2]), name
    # the unreachable goal returns the path to the last expanded state (jps.py:321-322), not an error
    assert not np.array_equal(fx["disc_blocked/cells"][-1], A.point_cells(fx["disc_blocked/goal"][:2], (0, 3, 0, 3), 0.25)[::-1])


def test_postprocess_oracle_matches_reference_fixture():
    """oracle/postprocess.py vs tests/golden/postprocess.npz (PathPostprocessor.process of the live reference on seven
    float32 paths: smooth, looping through +-pi, reversing start, nearly / exactly repeated nodes, coarse filter, short,
    dense output): same number of poses, values to 1e-9."""
    from oracle import postprocess as PP
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "postprocess.npz"))
    dropped_more_than_one = False
    for name in fx["names"]:
        path, want = fx[f"{name}/path"], fx[f"{name}/result"]
        min_d, step = fx[f"{name}/params"]
        got = PP.process(path, float(min_d), float(step))
        assert got.shape == want.shape, name
        assert np.abs(got - want).max() <= 1e-9, name
        # the cases exercise both filters
        traj = PP.filter_trajectory(np.array(path, dtype=np.float32), float(min_d))
        if name in ("jitter_dups", "coarse_filter"):
            assert len(traj) < len(path)
        full = int(np.sum(PP.chord_lengths(traj)) / float(step))
        dropped_more_than_one |= full - len(want) > 1
    assert dropped_more_than_one                       # reverse_start: the direction filter removed the backing-up poses
    assert PP.process(np.zeros((2, 3), np.float32)).shape == (2, 3)

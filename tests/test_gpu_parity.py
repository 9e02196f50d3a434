"""GPU parity tests proper: every call goes through the C ABI (libnfopp.so) and is compared with the CPU
oracle and with the golden vectors produced by the unmodified reference.

Bars (north_star): collision labels and sampled indices bit-exact; ONF outputs and gradients within 1e-3
relative (measured: ~1e-6 on the fp32 FMA path); lock-step state within 1e-3; short free runs within
1e-2 m / 1e-2 rad."""
import os
import ctypes as C

import numpy as np
import pytest
import torch

import _util as U
from oracle import checkers as OC
from oracle import onf as O
from oracle.planner import RecordedSource

pytestmark = pytest.mark.gpu

ONF_TOL = 1e-3      # north_star tolerance for ONF outputs and gradients (relative, max-norm)
STATE_TOL = 1e-3    # lock-step state tolerance (relative, max-norm per tensor)


def _lib():
    from nfopp_b200 import _lib as L
    return L


# ---------------------------------------------------------------------------------------- K1 labels
def test_labels_match_reference_golden_bit_exact():
    from nfopp_b200.collision_checker import (CircleCollisionChecker, CircleDirectedCollisionChecker,
                                              CollisionChecker, GridCollisionChecker, MapImage,
                                              RectangleCollisionChecker)
    from nfopp_b200.utils.position2 import Position2
    fx = U.load("checkers.npz")
    poses, sb, car = fx["poses"], fx["obstacles_sbend"], fx["obstacles_car"]
    c = CircleCollisionChecker(0.3, (0, 3, 0, 3)); c.update_obstacle_points(sb)
    assert np.array_equal(c.check_collision(poses[:, :2]), fx["circle_sbend"])
    c = CircleDirectedCollisionChecker(0.3, (0, 3, 0, 3)); c.update_obstacle_points(car)
    assert np.array_equal(c.check_collision(Position2.from_vec(poses)), fx["circle_directed_car"])
    c = CircleCollisionChecker(0.25, None); c.update_obstacle_points(car)
    assert np.array_equal(c.check_collision(poses[:, :2]), fx["circle_car_nobounds"])
    c = RectangleCollisionChecker((-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3)); c.update_obstacle_points(car)
    assert np.array_equal(c.check_collision(Position2.from_vec(poses)), fx["rect_car"])
    c = RectangleCollisionChecker((-0.2, 0.2, -0.2, 0.2), (0, 3, 0, 3)); c.update_obstacle_points(sb)
    assert np.array_equal(c.check_collision(Position2.from_vec(poses)), fx["rect_sbend"])
    assert np.array_equal(CollisionChecker((0, 3, 0, 3)).check_collision(poses[:, :2]), fx["boundary_only"])
    assert CollisionChecker().check_collision(poses[:, :2]) is False
    ox, oy, cell = fx["grid_meta"]
    g = fx["grid"]
    gc = GridCollisionChecker(MapImage(g.shape[1], g.shape[0], ox, oy, cell, g), (-1, 3, -0.5, 2.5))
    assert np.array_equal(gc.check_collision(Position2.from_vec(fx["grid_poses"])), fx["grid_labels"])


def test_labels_large_batch_vs_oracle():
    """10 M poses per checker (hard part 2), per-problem obstacle sets, ragged counts and an empty set; label flips
    against the float64 oracle are counted and printed (the rectangle checker's sin / cos must agree with glibc's to the bit)."""
    L = _lib()
    rs = np.random.RandomState(3)
    P, n, m = 4, 2_500_000, 40
    poses = np.stack([rs.uniform(-0.3, 3.3, (P, n)), rs.uniform(-0.3, 3.3, (P, n)), rs.uniform(-9, 9, (P, n))], 2)
    obs = rs.uniform(0, 3, (P, m, 2))
    counts = np.array([40, 17, 0, 1], dtype=np.int32)
    bounds = (C.c_double * 4)(0, 3, 0, 3)
    box = (C.c_double * 4)(-0.3, 0.2, -0.3, 0.2)
    d_poses, d_obs = torch.tensor(poses, device="cuda"), torch.tensor(obs, device="cuda")
    d_cnt = torch.tensor(counts, device="cuda")
    out = torch.empty(P, n, dtype=torch.uint8, device="cuda")
    s = L.stream_ptr()
    L.check(L.lib().nfopp_label_circle(L.ptr(d_poses), P, n, 3, L.ptr(d_obs), L.ptr(d_cnt), P, m, 0.3, bounds,
                                       L.ptr(out), s))
    got_c = out.cpu().numpy().astype(bool)
    L.check(L.lib().nfopp_label_rectangle(L.ptr(d_poses), P, n, L.ptr(d_obs), L.ptr(d_cnt), P, m, box, bounds,
                                          L.ptr(out), s))
    got_r = out.cpu().numpy().astype(bool)
    flips_c = flips_r = 0
    step = 250_000                                         # the oracle builds [poses, obstacles] temporaries
    for p in range(P):
        o = obs[p, :counts[p]]
        for i in range(0, n, step):
            sl = slice(i, i + step)
            flips_c += int(np.sum(got_c[p, sl] != OC.circle_labels(poses[p, sl], o, 0.3, (0, 3, 0, 3))))
            flips_r += int(np.sum(got_r[p, sl] != OC.rectangle_labels(poses[p, sl], o, (-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))))
    print(f"label flips vs the oracle: circle {flips_c} / {P * n} poses, rectangle {flips_r} / {P * n} poses "
          f"(colliding fraction {got_c.mean():.3f} / {got_r.mean():.3f})")
    assert flips_c == 0 and flips_r == 0, f"{flips_c} + {flips_r} label flips in 2 x {P * n} poses"
    grid = (rs.rand(1000, 1000) < 0.1).astype(np.uint8)
    d_grid = torch.tensor(grid, device="cuda")
    gp = poses.copy()
    gp[..., :2] = rs.uniform(-5, 105, (P, n, 2))
    d_gp = torch.tensor(gp, device="cuda")
    L.check(L.lib().nfopp_label_grid(L.ptr(d_gp), P, n, 3, L.ptr(d_grid), 1, 1000, 1000, 0.0, 0.0, 0.1, None,
                                     L.ptr(out), s))
    got_g = out.cpu().numpy().astype(bool)
    flips_g = sum(int(np.sum(got_g[p] != OC.grid_labels(gp[p], grid, 0.0, 0.0, 0.1))) for p in range(P))
    print(f"label flips vs the oracle: grid {flips_g} / {P * n} poses")
    assert flips_g == 0


def test_labels_empty_input():
    L = _lib()
    out = torch.empty(1, dtype=torch.uint8, device="cuda")
    poses = torch.empty(0, 2, dtype=torch.float64, device="cuda")
    L.check(L.lib().nfopp_label_boundary(L.ptr(poses) or C.c_void_p(out.data_ptr()), 1, 0, 2, None, L.ptr(out),
                                         L.stream_ptr()))


# ---------------------------------------------------------------------------------------- K2 ONF
IMPLS = {"fma": 1, "tensor": 2, "wide": 3}     # CUDA-core fp32 / fused tcgen05 kernel (reference shape) / tcgen05 grouped GEMMs per layer


def _onf_call(cfg, w, x, gz=None, mode=0, aux=None, wc=0.0, beta=1.0, count=None, impl=0):
    """-> z, gx, gw through nfopp_onf_forward / backward / reduce_grad.  w [P,NP], x [P,n,D]."""
    L = _lib()
    from nfopp_b200.engine import OnfSpec
    spec = OnfSpec(cfg.mean, cfg.sigma, cfg.use_cos, cfg.angle, cfg.hidden, cfg.layers, cfg.bias)
    cc = spec.to_c()
    P, n, D = x.shape
    dw, dx = w.cuda().contiguous(), x.cuda().contiguous()
    z = torch.zeros(P, n, device="cuda")
    s = L.stream_ptr()
    dcount = None if count is None else torch.tensor(count, dtype=torch.int32, device="cuda")
    scratch = L.onf_scratch(cc, P, impl, "cuda", n, True)
    L.check(L.lib().nfopp_onf_forward(C.byref(cc), L.ptr(dw), L.ptr(dx), L.ptr(dcount), P, n, L.ptr(z), impl,
                                      L.ptr(scratch), s))
    if gz is None and mode == 0:
        return z.cpu(), None, None
    tiles = L.lib().nfopp_onf_tiles(C.byref(cc), n, impl)
    assert tiles >= 1
    gpart = torch.zeros(P, tiles, w.shape[1], device="cuda")
    gx = torch.zeros(P, n, D, device="cuda")
    gw = torch.zeros(P, w.shape[1], device="cuda")
    z2 = torch.zeros(P, n, device="cuda")
    dgz = None if gz is None else gz.cuda().contiguous()
    daux = None if aux is None else aux.cuda().contiguous()
    L.check(L.lib().nfopp_onf_backward(C.byref(cc), L.ptr(dw), L.ptr(dx), L.ptr(dcount), P, n, mode, L.ptr(dgz),
                                       L.ptr(daux), wc, beta, L.ptr(z2), L.ptr(gx), L.ptr(gpart), impl, L.ptr(scratch), s))
    L.check(L.lib().nfopp_onf_reduce_grad(C.byref(cc), L.ptr(gpart), L.ptr(dcount), P, n, impl, L.ptr(gw), s))
    torch.cuda.synchronize()
    assert torch.equal(z, z2) or count is not None
    return z.cpu(), gx.cpu(), gw.cpu()


@pytest.mark.parametrize("impl", ["fma", "tensor", "wide"])
@pytest.mark.parametrize("name", ["holonomic", "car", "image"])
def test_onf_matches_reference_golden(name, impl):
    fx = U.load("onf.npz")
    cfg = U.ONF_CFGS[name]
    w, x, y = torch.tensor(fx[f"{name}/w"]), torch.tensor(fx[f"{name}/x"]), torch.tensor(fx[f"{name}/y"])
    zref = torch.tensor(fx[f"{name}/z"])
    gz = (torch.sigmoid(zref) - y) / len(y) + 0.01 * torch.sigmoid(zref)
    z, gx, gw = _onf_call(cfg, w[None], x[None], gz[None], impl=IMPLS[impl])
    assert U.relerr(z[0], fx[f"{name}/z"]) < ONF_TOL
    assert U.relerr(gw[0], fx[f"{name}/gw"]) < ONF_TOL
    assert U.relerr(gx[0], fx[f"{name}/gx"]) < ONF_TOL


@pytest.mark.parametrize("hidden,layers,use_cos,angle,n,impl", [
    (100, 2, True, True, 209, "fma"), (64, 2, False, False, 33, "fma"), (128, 4, True, True, 100, "fma"),
    (256, 4, True, True, 77, "fma"), (100, 2, True, False, 1, "fma"),
    # tensor path: tiles of 64 points with ragged tails (209 = 3*64+17, 65, 16, 1), narrow and odd widths
    (100, 2, True, True, 209, "tensor"), (64, 2, False, False, 33, "tensor"), (100, 2, True, False, 1, "tensor"),
    (108, 2, False, True, 65, "tensor"), (32, 2, True, True, 16, "tensor"), (100, 2, True, True, 500, "tensor"),
    # wide path: per-layer grouped GEMMs, 128-point row blocks with ragged tails, 1..8 layers, widths up to 512 (two N tiles)
    (100, 2, True, True, 209, "wide"), (64, 2, False, False, 33, "wide"), (256, 4, True, True, 77, "wide"),
    (256, 4, True, True, 1000, "wide"), (512, 2, True, True, 300, "wide"), (128, 1, True, False, 129, "wide"),
    (384, 3, False, True, 130, "wide"), (64, 8, True, True, 257, "wide")])
def test_onf_vs_oracle_shapes_and_ragged_counts(hidden, layers, use_cos, angle, n, impl):
    cfg = O.OnfConfig(mean=0.3, sigma=1.7, use_cos=use_cos, angle=angle, hidden=hidden, layers=layers)
    gen = torch.Generator().manual_seed(hidden + n)
    P = 3
    w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(P)])
    x = torch.rand(P, n, cfg.in_dim, generator=gen) * 4 - 1
    gz = torch.randn(P, n, generator=gen)
    count = [n, max(1, n // 2), 1]
    z, gx, gw = _onf_call(cfg, w, x, gz, count=count, impl=IMPLS[impl])
    for p in range(P):
        c = count[p]
        zo, cache = O.onf_forward(w[p], cfg, x[p, :c])
        gwo, gxo = O.onf_backward(w[p], cfg, cache, gz[p, :c])
        assert U.relerr(z[p, :c], zo) < ONF_TOL
        assert U.relerr(gw[p], gwo) < ONF_TOL
        assert U.relerr(gx[p, :c], gxo) < ONF_TOL


def test_onf_impl_auto_and_unsupported_shapes():
    """impl=0 resolves to the tensor path for the reference's network shape; the tensor path refuses shapes it
    cannot hold (deeper / wider nets) instead of silently computing something else."""
    L = _lib()
    from nfopp_b200.engine import OnfSpec
    ref = OnfSpec(0.0, 1.0, True, True).to_c()
    assert L.lib().nfopp_onf_tiles(C.byref(ref), 209, 0) == 1 == L.lib().nfopp_onf_tiles(C.byref(ref), 209, 2)
    assert L.lib().nfopp_onf_tiles(C.byref(ref), 209, 1) == 7
    wide = OnfSpec(0.0, 1.0, True, True, 256, 4).to_c()
    assert L.lib().nfopp_onf_tiles(C.byref(wide), 209, 2) < 0
    assert L.lib().nfopp_onf_tiles(C.byref(wide), 209, 0) == L.lib().nfopp_onf_tiles(C.byref(wide), 209, 3) == 1    # auto -> wide
    assert L.lib().nfopp_onf_tiles(C.byref(wide), 209, 1) > 1
    nbytes = C.c_size_t()
    with pytest.raises(ValueError):
        L.check(L.lib().nfopp_onf_scratch_bytes(C.byref(wide), 4, 2, C.byref(nbytes)))


@pytest.mark.parametrize("impl", ["fma", "tensor", "wide"])
def test_onf_fused_gbar_modes_and_adam(impl):
    """BCE / path / path_mu upstream gradients computed in-kernel, then the fused reduce+Adam update."""
    impl = IMPLS[impl]
    from oracle import trajectory as TR
    L = _lib()
    cfg = U.ONF_CFGS["car"]
    gen = torch.Generator().manual_seed(5)
    P, n = 2, 209
    w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(P)])
    x = torch.rand(P, n, 3, generator=gen) * 3
    y = (torch.rand(P, n, generator=gen) < 0.3).float()
    mu = torch.rand(P, n, generator=gen)
    for mode, aux in [(1, y), (2, None), (3, mu)]:
        z, gx, gw = _onf_call(cfg, w, x, None, mode=mode, aux=aux, wc=0.7, beta=1.3 if mode == 3 else 1.0, impl=impl)
        for p in range(P):
            zo, cache = O.onf_forward(w[p], cfg, x[p])
            if mode == 1:
                _, g = O.bce_with_logits_grad(zo, y[p])
            else:
                g = TR.path_gbar(zo, mu[p] if mode == 3 else None, 0.7, 1.3 if mode == 3 else 1.0)
            gwo, gxo = O.onf_backward(w[p], cfg, cache, g)
            assert U.relerr(gw[p], gwo) < ONF_TOL and U.relerr(gx[p], gxo) < ONF_TOL
    # Adam: three updates in a row
    from nfopp_b200.engine import OnfSpec
    cc = OnfSpec(cfg.mean, cfg.sigma, cfg.use_cos, cfg.angle).to_c()
    tiles = L.lib().nfopp_onf_tiles(C.byref(cc), n, impl)
    scratch = L.onf_scratch(cc, P, impl, "cuda", n, True)
    dw, dm, dv = w.cuda(), torch.zeros_like(w).cuda(), torch.zeros_like(w).cuda()
    dx, dy = x.cuda(), y.cuda()
    gpart = torch.zeros(P, tiles, w.shape[1], device="cuda")
    wo, mo, vo = w.clone(), torch.zeros_like(w), torch.zeros_like(w)
    for step in (1, 2, 3):
        L.check(L.lib().nfopp_onf_backward(C.byref(cc), L.ptr(dw), L.ptr(dx), None, P, n, 1, None, L.ptr(dy), 0.0, 1.0,
                                           None, None, L.ptr(gpart), impl, L.ptr(scratch), L.stream_ptr()))
        L.check(L.lib().nfopp_onf_adam(C.byref(cc), L.ptr(dw), L.ptr(dm), L.ptr(dv), L.ptr(gpart), None, P, n, impl,
                                       step, 5e-2, 0.9, 0.9, 1e-8, L.stream_ptr()))
        for p in range(P):
            zo, cache = O.onf_forward(wo[p], cfg, x[p])
            _, g = O.bce_with_logits_grad(zo, y[p])
            gwo, _ = O.onf_backward(wo[p], cfg, cache, g)
            O.adam_step(wo[p], mo[p], vo[p], gwo, step, 5e-2, 0.9, 0.9)
    # Adam takes sign-like steps where g ~ 0 (BASELINE.md "chaos"): compare the smooth moments tightly and the
    # weights where the gradient is not negligible
    em = U.relerr(dm.cpu(), mo)
    assert em < ONF_TOL and U.relerr(dv.cpu(), vo) < ONF_TOL
    # first-order bound: an Adam step is lr*m/sqrt(v), so a max-norm gradient error em moves element i by at most
    # lr*em*max|m|/|m_i| per step (and never by more than lr): 3 steps, factor 8 of slack for v and accumulation
    ratio = mo.abs().max() / mo.abs().clamp_min(1e-30)
    bound = 3 * 5e-2 * torch.clamp(8 * max(em, 1e-7) * ratio, max=1.0) + 1e-6
    assert bool(((dw.cpu() - wo).abs() <= bound).all())


def test_onf_module_autograd_and_state_dict():
    from nfopp_b200.onf_model import ONF
    fx = U.load("onf.npz")
    name = "car"
    cfg = U.ONF_CFGS[name]
    model = ONF(0, 1, use_cos=True, use_normal_init=True, angle_encoding=True)
    keys = set(model.state_dict().keys())
    assert keys == {"_angle_encoder._biases", "_angle_encoder._frequencies", "mlp.0.weight", "mlp.0.bias",
                    "mlp.2.weight", "mlp.2.bias", "mlp2.0.weight", "mlp2.0.bias", "encoding_layer.weight",
                    "encoding_layer.bias"}
    model = model.cuda()
    model.packed_parameters().copy_(torch.tensor(fx[f"{name}/w"]).cuda())
    model.requires_grad_(True)
    x = torch.tensor(fx[f"{name}/x"]).cuda().requires_grad_(True)
    y = torch.tensor(fx[f"{name}/y"]).cuda()
    out = model(x)
    assert out.shape == (len(y), 1)
    z = out[:, 0]
    loss = torch.nn.BCEWithLogitsLoss()(z, y) + 0.01 * torch.nn.functional.softplus(z).sum()
    loss.backward()
    assert U.relerr(z.detach().cpu(), fx[f"{name}/z"]) < ONF_TOL
    assert U.relerr(x.grad.cpu(), fx[f"{name}/gx"]) < ONF_TOL
    named = dict(model.named_parameters())
    gw = torch.tensor(fx[f"{name}/gw"])
    lay, _ = O.onf_layout(cfg)
    for pname, seg in [("mlp.0.weight", "W0"), ("mlp.2.bias", "b1"), ("mlp2.0.weight", "w_out"),
                       ("encoding_layer.weight", "enc_w"), ("_angle_encoder._frequencies", "ang_f")]:
        off, shape = lay[seg]
        ref = gw[off:off + int(np.prod(shape))]
        assert U.relerr(named[pname].grad.reshape(-1).cpu(), ref) < ONF_TOL, pname


# ---------------------------------------------------------------------------------------- choice
def test_choice_indices_bit_exact():
    L = _lib()
    fx = U.load("choice.npz")
    for case in sorted({k.split("/")[0] for k in fx.files}):
        p, u, idx = fx[f"{case}/p"], fx[f"{case}/u"], fx[f"{case}/idx"]
        dp, du = torch.tensor(p[None], device="cuda"), torch.tensor(u[None], device="cuda")
        out = torch.zeros(1, 100, dtype=torch.int32, device="cuda")
        used = torch.zeros(1, dtype=torch.int32, device="cuda")
        L.check(L.lib().nfopp_choice(L.ptr(dp), 1, len(p), 100, int(case.endswith("_1")), L.ptr(du), len(u),
                                     L.ptr(out), L.ptr(used), L.stream_ptr()))
        assert np.array_equal(out.cpu().numpy()[0], idx), case
        assert int(used[0]) == len(u), case


# ---------------------------------------------------------------------------------------- planner step
@pytest.mark.parametrize("impl", ["fma", "tensor"])
@pytest.mark.parametrize("name", ["car", "corridor", "holonomic", "corridor_freq"])
def test_lockstep_against_reference_states(name, impl):
    """Import the reference's full state at step k, run ONE GPU step with the reference's variates, compare
    the post-step state with the reference's (golden) and with the oracle's."""
    from _gpu_util import compare_state, engine_from_golden
    fx = U.load(f"lockstep_{name}.npz")
    for k in fx["meta/steps"]:
        eng = engine_from_golden(name, fx, f"step{k}/before", onf_impl=impl)
        var = U.variates_from(fx, f"step{k}/var")
        eng.inject_variates(var)
        eng.step(1)
        torch.cuda.synchronize()
        if "t_sample" in var:              # a field-training step (optimize_collision_model_freq, nerf_opt_planner.py:62)
            n = int(eng["train_n"][0])
            assert n == len(fx[f"step{k}/labels"])
            assert np.array_equal(eng["train_pos"][0, :n].cpu().numpy(), fx[f"step{k}/checked"]), "sampled poses"
            assert np.array_equal(eng["train_y"][0, :n].cpu().numpy() > 0.5, fx[f"step{k}/labels"]), "labels"
        else:
            assert torch.equal(eng["onf_w"][0].cpu(), U.state_from(fx, f"step{k}/before")["w"]), "weights moved on a non-training step"
        if "choice_idx" in var:
            assert np.array_equal(eng["choice_idx"][0].cpu().numpy(), var["choice_idx"]), "sampled indices"
            assert int(eng["choice_used"][0]) == len(var["choice_u"])
        assert int(eng["status"][0]) == 0
        after = U.state_from(fx, f"step{k}/after")
        # Adam moves weights by ~lr*sign(g) where |g| ~ 0 (first steps): compare weights against an absolute
        # bound of 1e-3 of one Adam step, the smooth moments relatively
        compare_state(eng, after, STATE_TOL, skip=("w",))
        if "t_sample" in var:
            _adam_weight_check(eng, after, U.PLANNER_CFGS[name].onf_lr)


def _adam_weight_check(eng, after, lr):
    """weights after an Adam step: |step| <~ 3 lr, and a max-norm gradient error em moves element i by at most
    ~lr*em*max|m|/|m_i| (first order; factor 8 of slack) - tight where the gradient is not negligible."""
    m_ref = np.abs(after["m"].numpy())
    em = max(U.relerr(eng["onf_m"][0].cpu().numpy(), after["m"].numpy()), 1e-7)
    bound = lr * np.minimum(3.0, 8 * em * m_ref.max() / np.maximum(m_ref, 1e-30)) + 1e-6
    dw = np.abs(eng["onf_w"][0].cpu().numpy() - after["w"].numpy())
    assert bool((dw <= bound).all()), float((dw / bound).max())
    assert float(dw[m_ref > 1e-2 * m_ref.max()].max()) < 2e-2 * lr


@pytest.mark.parametrize("impl", ["fma", "tensor"])
@pytest.mark.parametrize("name", ["holonomic", "car"])
def test_warmup_against_reference_states(name, impl):
    """_init_collision_model (nerf_opt_planner.py:197-200) in lock-step with the reference - every
    warm-up iteration's 100 uniform field points (:135-141; constrained :173-176) and labels bit-exact, the ONF
    weights / Adam moments and the ONF optimiser's own step counter after each iteration, then the first planner step
    after the warm-up (ONF optimiser at step 3, trajectory optimiser at step 0)."""
    from _gpu_util import compare_state, engine_from_golden
    fx = U.load(f"warmup_{name}.npz")
    nw = int(fx["meta/init_collision_points"])
    eng = engine_from_golden(name, fx, "warm/before", onf_impl=impl, init_collision_points=nw)
    lr = U.PLANNER_CFGS[name].onf_lr
    for i in range(int(fx["meta/iterations"])):
        eng.inject_variates(U.variates_from(fx, f"warm/var{i}"))
        eng.warmup(1)
        torch.cuda.synchronize()
        assert int(eng["train_n"][0]) == nw
        assert np.array_equal(eng["train_pos"][0, :nw].cpu().numpy(), fx[f"warm/checked{i}"]), "warm-up field points"
        assert np.array_equal(eng["train_y"][0, :nw].cpu().numpy() > 0.5, fx[f"warm/labels{i}"]), "warm-up labels"
        after = U.state_from(fx, f"warm/after{i}")
        compare_state(eng, after, STATE_TOL, skip=("w",))     # incl. onf_step == i + 1, traj_step == step_count == 0
        _adam_weight_check(eng, after, lr)
        assert eng.counter("has_prev") == 0
        # lock-step: continue from the reference's own state
        from _gpu_util import load_state
        load_state(eng, after)
    eng.inject_variates(U.variates_from(fx, "step0/var"))
    eng.step(1)
    torch.cuda.synchronize()
    after = U.state_from(fx, "step0/after")
    compare_state(eng, after, STATE_TOL, skip=("w",))
    _adam_weight_check(eng, after, lr)


@pytest.mark.parametrize("impl", ["fma", "tensor"])
@pytest.mark.parametrize("name", ["car", "corridor", "holonomic"])
def test_short_free_run_matches_reference_path(name, impl):
    """10 consecutive GPU steps from the reference's initial state with its variates: final paths within
    north_star's 1e-2 m / 1e-2 rad."""
    from _gpu_util import engine_from_golden
    fx = U.load(f"lockstep_{name}.npz")
    eng = engine_from_golden(name, fx, "free/state0", onf_impl=impl)
    for k in range(10):
        eng.inject_variates(U.variates_from(fx, f"free/var{k}"))
        eng.step(1)
        path = eng.paths()[0].cpu().numpy()
        assert np.abs(path - fx[f"free/path{k}"]).max() < 1e-2, k


@pytest.mark.parametrize("impl", ["fma", "tensor"])
def test_batch_problems_are_independent_and_deterministic(impl):
    """P=8 problems fed the same state/variates give 8 identical answers equal to the P=1 run; Philox mode
    is reproducible run to run and differs between problems."""
    from _gpu_util import engine_from_golden
    fx = U.load("lockstep_corridor.npz")
    k = 10
    var = U.variates_from(fx, f"step{k}/var")
    one = engine_from_golden("corridor", fx, f"step{k}/before", onf_impl=impl)
    one.inject_variates(var); one.step(1)
    many = engine_from_golden("corridor", fx, f"step{k}/before", problems=8, onf_impl=impl)
    many.inject_variates(var); many.step(1)
    for name in ("traj", "onf_w", "lam", "mu", "buf_pos"):
        a, b = one[name][0], many[name]
        assert all(torch.equal(a, b[p]) for p in range(8)), name
    runs = []
    for _ in range(2):
        e = engine_from_golden("corridor", fx, f"step{k}/before", problems=4, rng_mode="philox", seed=11, onf_impl=impl)
        e.step(5, graph=False)
        runs.append(e["traj"].clone())
    assert torch.equal(runs[0], runs[1])
    assert not torch.equal(runs[0][0], runs[0][1])
    g = engine_from_golden("corridor", fx, f"step{k}/before", problems=4, rng_mode="philox", seed=11, onf_impl=impl)
    g.step(5, graph=True)
    torch.cuda.synchronize()
    assert torch.equal(g["traj"], runs[0]), "CUDA-graph replay differs from plain launches"
    assert g.counter("step_count") == 15


# ---------------------------------------------------------------------------------------- wide ONF on occupancy grids
def _grid_case(hidden, layers, seed=5, rows=200, cols=200, cell=0.015, onf_lr=5e-3):
    """BASELINE config 4 in miniature: one occupancy grid, constrained planner, wide ONF.  The reference's ONF is
    hard-wired to 2x100 (onf_model.py:18-23); the oracle's generalised restatement is bit-identical to it at that
    shape (tests/test_oracle_golden.py) and is the checker for other widths."""
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from oracle.planner import LiveSource, OraclePlanner, PlannerConfig
    grid, (ox, oy), cell = B.make_blob_grid(rows, cols, 0.1, cell, blob_cells=12, seed=seed)
    starts, goals = B.free_endpoints(grid, cell, 1, seed=seed, margin_cells=10)
    w, h = cols * cell, rows * cell
    bounds, border = (0.0, w, 0.0, h), (-0.1, w + 0.1, -0.1, h + 0.1)
    ocfg = O.OnfConfig(mean=0.0, sigma=1.0, use_cos=True, angle=True, hidden=hidden, layers=layers)
    pc = PlannerConfig(constrained=True, onf=ocfg, onf_lr=onf_lr, onf_betas=(0.9, 0.9), traj_lr=1e-2, traj_betas=(0.9, 0.9),
                       init_collision_iteration=0, angle_offset=0.3)
    blob = O.init_blob(ocfg, True, torch.Generator().manual_seed(seed))
    op = OraclePlanner(pc, blob, lambda p: OC.grid_labels(p, grid, ox, oy, cell, bounds))
    src = LiveSource(np.random.RandomState(seed), torch.Generator().manual_seed(seed))
    op.init(starts[0], goals[0], border, src)
    return op, src, pc, grid, cell, bounds, border


def _grid_engine(pc, grid, cell, bounds, border, **over):
    from _gpu_util import onf_spec
    from nfopp_b200.engine import EngineConfig, PlannerEngine
    kw = dict(problems=1, constrained=True, trajectory_length=pc.trajectory_length, onf=onf_spec(pc.onf), onf_lr=pc.onf_lr,
              onf_betas=pc.onf_betas, traj_lr=pc.traj_lr, traj_betas=pc.traj_betas, init_collision_iteration=0,
              angle_offset=pc.angle_offset, checker="grid", env_shared=True, grid_shape=grid.shape, grid_origin=(0.0, 0.0),
              grid_cell=cell, checker_bounds=bounds, border=border, rng_mode="inject", max_choice_uniforms=4096)
    kw.update(over)
    eng = PlannerEngine(EngineConfig(**kw))
    eng.set_grid(grid)
    return eng


@pytest.mark.parametrize("onf_lr", [5e-3, 1e-9])
@pytest.mark.parametrize("hidden,layers,impl", [(256, 4, "wide"), (128, 3, "wide"), (100, 2, "wide"), (128, 3, "fma")])
def test_lockstep_wide_onf_on_grid_map(hidden, layers, impl, onf_lr):
    """Lock-step against the oracle on a grid map with a wide / deep ONF: the oracle's state before steps 0, 1, 10, 11 is
    loaded into the engine, one GPU step runs on the oracle's variates, the post-step states are compared.

    A wide ReLU network has thousands of (point, unit) pre-activations within rounding of zero and thousands of weights
    whose gradient is within rounding of zero; either flips a mask / the sign of an Adam step (+-lr) - the chaos
    BASELINE.md documents for the reference itself.  So: with the real learning rate the smooth quantities (labels,
    indices, Adam moments) are compared and the trajectory half of the step is checked for CONSISTENCY with the
    engine's own post-Adam weights; with lr ~ 0 (weights frozen) the whole post-step state is compared tightly."""
    import copy
    from _gpu_util import compare_state, load_state
    op, src, pc, grid, cell, bounds, border = _grid_case(hidden, layers, onf_lr=onf_lr)
    frozen = onf_lr < 1e-6
    for k in range(12):
        before = copy.deepcopy(op.state)
        var = op.step(src)
        if k not in (0, 1, 10, 11):
            continue
        eng = _grid_engine(pc, grid, cell, bounds, border, onf_impl=impl)
        load_state(eng, before)
        eng.inject_variates(var)
        eng.step(1)
        torch.cuda.synchronize()
        n = int(eng["train_n"][0])
        assert np.array_equal(eng["train_pos"][0, :n].cpu().numpy(), op.last["train_positions"]), "sampled poses"
        assert np.array_equal(eng["train_y"][0, :n].cpu().numpy() > 0.5, op.last["train_labels"]), "labels"
        if len(before["buf_pos"]) + 99 >= 100:
            assert np.array_equal(eng["choice_idx"][0].cpu().numpy(), np.asarray(op.last["choice_idx"])), "sampled indices"
        assert int(eng["status"][0]) == 0
        traj_keys = ("traj", "traj_m", "traj_v", "lam", "mu")
        compare_state(eng, op.state, STATE_TOL, skip=("w",) + (() if frozen else traj_keys))
        m_ref = np.abs(op.state["m"].numpy())
        em = max(U.relerr(eng["onf_m"][0].cpu().numpy(), op.state["m"].numpy()), 1e-7)
        bound = onf_lr * np.minimum(3.0, 8 * em * m_ref.max() / np.maximum(m_ref, 1e-30)) + 1e-6
        dw = np.abs(eng["onf_w"][0].cpu().numpy() - op.state["w"].numpy())
        assert bool((dw <= bound).all()), float((dw / bound).max())
        if not frozen:
            # the path launch must have seen the weights Adam just wrote (operand copies refreshed): recompute its
            # logits through the stateless API from the engine's current weights
            q = eng["path_q"][0, :pc.trajectory_length - 1].cpu()
            z, _, _ = _onf_call(pc.onf, eng["onf_w"].cpu(), q[None], impl=IMPLS[impl])
            assert U.relerr(eng["path_z"][0].cpu(), z[0]) < 1e-5
            for key in traj_keys:        # and the trajectory half stays close to the oracle's despite the +-lr weight noise
                assert U.relerr(eng[key][0].cpu().numpy(), np.asarray(op.state[key])) < 5e-2, key
        eng.close()


def test_wide_engine_graph_replay_and_batch_independence():
    """P problems on one shared grid with the wide path: CUDA-graph replay == plain launches; problems differ."""
    from nfopp_b200.batched import BatchedPlanner, gridmap_wide_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    grid, _, cell = B.make_blob_grid(200, 200, 0.1, 0.015, blob_cells=12, seed=2)
    starts, goals = B.free_endpoints(grid, cell, 6, seed=3, margin_cells=10)
    outs = []
    for graph in (False, True):
        cfg = gridmap_wide_config(6, rows=200, cols=200, cell=cell, seed=9, hidden=128, layers=3)
        bp = BatchedPlanner(cfg)
        bp.init(starts, goals, grids=grid)
        bp.step(6, graph=graph)
        torch.cuda.synchronize()
        outs.append(bp.paths().clone())
        assert bool(torch.isfinite(outs[-1]).all())
        assert int(np.abs(bp.status()).max()) == 0
    assert torch.equal(outs[0], outs[1])
    assert not torch.equal(outs[0][0], outs[0][1])


def test_wide_engine_with_training_every_other_step():
    """nerf_opt_planner.py:62 on the wide path: optimize_collision_model_freq = 2 (and reparametrize_trajectory_freq = 3).  The
    wide ONF's launches cannot be switched off by a device-side flag, so steps that do not train replay a second CUDA graph
    without them: weights, Adam counter and replay buffer stay bit-identical over such a step while the trajectory moves;
    graph replay == plain launches; and the run stays next to the FMA path's (same RNG streams, ONF outputs to ~1e-6)."""
    from nfopp_b200.batched import BatchedPlanner, gridmap_wide_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    grid, _, cell = B.make_blob_grid(200, 200, 0.1, 0.015, blob_cells=12, seed=2)
    P = 5
    starts, goals = B.free_endpoints(grid, cell, P, seed=4, margin_cells=10)
    outs = {}
    for impl, graph in (("wide", False), ("wide", True), ("fma", True)):
        cfg = gridmap_wide_config(P, rows=200, cols=200, cell=cell, seed=9, hidden=64, layers=2, onf_impl=impl,
                                  optimize_collision_model_freq=2, reparametrize_trajectory_freq=3)
        bp = BatchedPlanner(cfg, lanes=1)
        bp.init(starts, goals, grids=grid)
        eng = bp.engine
        bp.step(1, graph=graph)                                   # step_count 0: trains
        torch.cuda.synchronize()
        assert eng.counter("onf_step") == 1
        w1, buf1, traj1 = eng["onf_w"].clone(), eng["buf_pos"].clone(), eng["traj"].clone()
        bp.step(1, graph=graph)                                   # step_count 1: no training
        torch.cuda.synchronize()
        assert eng.counter("onf_step") == 1 and eng.counter("step_count") == 2
        assert torch.equal(eng["onf_w"], w1) and torch.equal(eng["buf_pos"], buf1)
        assert not torch.equal(eng["traj"], traj1)
        bp.step(3, graph=graph)                                   # step_count 2, 3, 4: trains, skips, trains
        torch.cuda.synchronize()
        assert eng.counter("onf_step") == 3 and eng.counter("step_count") == 5
        assert not torch.equal(eng["onf_w"], w1)
        assert int(np.abs(bp.status()).max()) == 0
        outs[(impl, graph)] = (bp.paths().clone(), eng["onf_w"].clone())
    assert torch.equal(outs[("wide", False)][0], outs[("wide", True)][0])
    assert torch.equal(outs[("wide", False)][1], outs[("wide", True)][1])
    dpath = (outs[("wide", True)][0] - outs[("fma", True)][0]).abs().max().item()
    assert dpath < 1e-2, dpath


# ---------------------------------------------------------------------------------------- path evaluation
@pytest.mark.parametrize("name,k", [("car", 4), ("corridor", 7), ("holonomic", 1)])
def test_evaluate_paths_matches_oracle(name, k):
    """nfopp_engine_evaluate_paths vs oracle.evaluate on paths of a running planner: collision flags and colliding-pose
    counts bit-exact (every interpolated pose goes through the reference's checker arithmetic), lengths to 1e-6."""
    from _gpu_util import engine_from_golden
    from oracle import evaluate as EV
    fx = U.load(f"lockstep_{name}.npz")
    P = 16
    eng = engine_from_golden(name, fx, "free/state0", problems=P, rng_mode="philox", seed=5)
    label = U.labeller_for(name, fx)
    seen = set()
    for _ in range(6):
        eng.step(25)
        collides, length, hits = eng.evaluate_paths(k)
        paths = eng.paths().cpu().numpy()
        for p in range(P):
            c, ln, h = EV.evaluate_path(paths[p], label, k)
            assert bool(collides[p]) == c and int(hits[p]) == h, (p, int(hits[p]), h)
            assert abs(float(length[p]) - ln) <= 1e-6 * ln
            seen.add(c)
    assert seen                                              # (the S-bend corridor is still colliding after 150 steps)


def test_batched_solve_loop_matches_reference_bookkeeping():
    """BatchedPlanner.solve = scripts/run_bench_mr.py:109-131 per problem: replaying the recorded per-check (collides,
    length) history of one problem through the oracle's solve_loop bookkeeping gives the same verdict."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    P = 64
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=0)
    bp = BatchedPlanner(corridor_config(P, seed=3))
    bp.init(starts, goals, obstacles=obstacles)
    out = bp.solve(max_iterations=600, min_iterations=200, check_every=50)
    solved = out["solved"].cpu().numpy()
    assert out["paths"].shape == (P, 102, 3) and bool(torch.isfinite(out["paths"]).all())
    assert 0 < solved.sum()                                      # some corridors are solved within 600 steps
    # a solved problem's returned path is collision-free under the engine's checker, and its length is the reported one
    eng = bp.engine
    eng["traj"].copy_(out["paths"][:, 1:-1])
    collides, length, _ = eng.evaluate_paths(4)
    assert not bool((collides & out["solved"]).any())
    assert torch.allclose(length[out["solved"]], out["length"][out["solved"]], rtol=1e-5)
    its = out["iterations"].cpu().numpy()
    assert ((its == 600) | ((its - 1) % 50 == 0)).all() and (its[its < 600] > 200).all()


def test_finished_problems_are_frozen_without_disturbing_the_others():
    """In the reference every problem is its own run that stops at the `break` of run_bench_mr.py:109-131.  The batched
    engine freezes a finished problem (buffer `frozen`): no kernel touches its state any more, the ONF launches skip it,
    and - problems being independent - every other problem continues bit for bit as if nothing had been frozen.  Checked
    on all three ONF paths, then end to end: solve() with and without freezing returns identical results."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    P = 12
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=1)
    names = ["traj", "traj_m", "traj_v", "lam", "mu", "onf_w", "onf_m", "onf_v", "buf_pos", "buf_age"]
    for impl, hidden in (("tensor", 100), ("fma", 100), ("wide", 64)):
        runs = []
        for freeze in (False, True):
            from dataclasses import replace
            cfg = corridor_config(P, seed=5, onf_impl=impl)
            cfg = replace(cfg, onf=replace(cfg.onf, hidden=hidden))
            bp = BatchedPlanner(cfg, lanes=2 if impl == "tensor" else 1)
            bp.init(starts, goals, obstacles=obstacles)
            bp.step(3)
            eng = bp.engine
            before = {k: eng[k].clone() for k in names}
            if freeze:
                eng["frozen"][::3] = 1
            bp.step(4)
            torch.cuda.synchronize()
            runs.append((before, {k: eng[k].clone() for k in names}))
            assert int(np.abs(bp.status()).max()) == 0
        (b0, a0), (b1, a1) = runs
        for k in names:
            assert torch.equal(b0[k], b1[k]), (impl, k)
            assert torch.equal(a1[k][::3], b1[k][::3]), (impl, k, "a frozen problem moved")
            keep = torch.ones(P, dtype=torch.bool)
            keep[::3] = False
            assert torch.equal(a1[k][keep], a0[k][keep]), (impl, k, "freezing disturbed another problem")
            assert not torch.equal(a0[k][::3], b0[k][::3]) or k in ("buf_pos", "buf_age"), (impl, k)
    P = 48
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=0)
    outs = []
    for freeze in (False, True):
        bp = BatchedPlanner(corridor_config(P, seed=3))
        bp.init(starts, goals, obstacles=obstacles)
        outs.append(bp.solve(max_iterations=500, min_iterations=200, check_every=50, freeze_finished=freeze))
        assert int(bp.engine["frozen"].abs().max()) == 0
    assert int(outs[0]["solved"].sum()) > 0 and int((outs[0]["iterations"] < 500).sum()) > 0
    for k in ("paths", "solved", "length", "iterations"):
        assert torch.equal(outs[0][k], outs[1][k]), k


# ---------------------------------------------------------------------------------------- path statistics
@pytest.mark.parametrize("name", ["car", "corridor", "holonomic"])
def test_path_statistics_match_oracle(name):
    """nfopp_engine_path_statistics vs oracle.path_stats (BenchmarkAdapter.cpp:68-135 restated) on paths of a running
    planner: cusp counts exact, the float64 metrics to 1e-4 relative (float32 outputs)."""
    from _gpu_util import engine_from_golden
    from oracle import path_stats as PS
    fx = U.load(f"lockstep_{name}.npz")
    P = 8
    eng = engine_from_golden(name, fx, "free/state0", problems=P, rng_mode="philox", seed=9)
    obs, bounds = fx["meta/obstacles"], tuple(float(b) for b in fx["meta/bounds"])
    env = dict(kind="rectangle", obstacles=obs, box=(-0.3, 0.2, -0.3, 0.2), bounds=bounds) if name == "car" else \
        dict(kind="circle", obstacles=obs, radius=0.3, bounds=bounds)
    for _ in range(3):
        eng.step(40)
        stats = {k: v.cpu().numpy() for k, v in eng.path_statistics(cusp_angle_threshold=np.pi / 3).items()}
        paths = eng.paths().cpu().numpy()
        for p in range(P):
            ref = PS.path_statistics(paths[p], np.pi / 3, **env)
            assert int(stats["cusps"][p]) == int(ref["cusps"])
            for k, v in ref.items():
                assert abs(float(stats[k][p]) - v) <= 1e-4 * max(abs(v), 1e-3), (k, p, float(stats[k][p]), v)


def test_path_statistics_on_occupancy_grid():
    """clearing distances against occupied cells (grid checker) + the straight-line known answers."""
    from nfopp_b200.batched import BatchedPlanner, forest_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from oracle import path_stats as PS
    P = 6
    starts, goals, _, grids, _, cell = B.make_random_grid_batch(P, seed=2)
    cfg = forest_config(P, cell=cell, seed=4)
    bp = BatchedPlanner(cfg)
    bp.init(starts, goals, grids=grids)
    st0 = {k: v.cpu().numpy() for k, v in bp.path_statistics().items()}
    assert np.all(st0["cusps"] == 0) and np.all(st0["max_curvature"] < 1e-3)       # straight-line initial paths
    bp.step(30)
    stats = {k: v.cpu().numpy() for k, v in bp.path_statistics(clearing_window_cells=16).items()}
    paths = bp.get_path()
    for p in range(P):
        ref = PS.path_statistics(paths[p], np.pi / 3, kind="grid", grid=np.asarray(grids[p]), cell=cfg.grid_cell,
                                 bounds=cfg.checker_bounds, window=16)
        for k, v in ref.items():
            assert abs(float(stats[k][p]) - v) <= 1e-4 * max(abs(v), 1e-3), (k, p, float(stats[k][p]), v)


# ---------------------------------------------------------------------------------------- A* initialiser
def test_astar_initialiser_matches_reference():
    """nfopp_engine_init_astar vs the live reference's AstarTrajectoryInitializer (tests/golden/astar.npz, made by
    oracle/make_golden_astar.py): number of cells on the A* path exact, initial x, y trajectory within 2e-6 (float32
    outputs of a float64 spline; the oracle restatement matches the reference bit for bit), headings as the reference."""
    from _gpu_util import engine_config
    from nfopp_b200.engine import PlannerEngine
    from oracle import astar as A
    from oracle import checkers as OC
    fx = U.load("astar.npz")
    lock = U.load("lockstep_car.npz")
    for name in [str(n) for n in fx["names"]]:
        g = lambda k: fx[f"{name}/{k}"]                                      # noqa: E731
        kind, bounds, res, T = str(g("kind")), tuple(float(b) for b in g("bounds")), float(g("resolution")), int(g("T"))
        obs, start, goal = g("obstacles"), g("start"), g("goal")
        P = 3                                                                # problem 1: reversed, problem 2: other goal
        starts = np.stack([start, goal, start]).astype(np.float32)
        goals = np.stack([goal, start, np.array([start[0], goal[1], 0.0], dtype=np.float32)])
        cfg = engine_config("car", lock, rng_mode="philox", problems=P, trajectory_length=T, checker=kind,
                            checker_bounds=bounds, border=bounds, max_obstacles=max(64, len(obs)))
        eng = PlannerEngine(cfg)
        eng.set_obstacles(obs)
        eng.set_endpoints(starts, goals)
        eng.init_trajectories()
        theta0 = eng["traj"][:, :, 2].clone()
        n_cells = eng.init_astar(res).cpu().numpy()
        traj = eng["traj"].cpu().numpy()
        assert n_cells[0] == len(g("cells")) and np.isfinite(traj).all()
        assert np.abs(traj[0, :, :2] - g("trajectory")[:, :2]).max() <= 2e-6, name
        assert np.abs(traj[0, :, 2] - g("trajectory")[:, 2]).max() <= 1e-6       # initialize_angle (trajectory_initializer.py:21-29)
        assert torch.equal(eng["traj"][:, :, 2], theta0)
        label = (lambda p: OC.rectangle_labels(p, obs, (-0.3, 0.2, -0.3, 0.2), bounds)) if kind == "rectangle" else \
            (lambda p: OC.circle_labels(p, obs, 0.3, bounds))
        for p in (1, 2):                                                     # the other problems against the oracle
            xy, cells, _ = A.initial_trajectory_xy(starts[p], goals[p], bounds, res, label, T)
            assert n_cells[p] == len(cells), (name, p)
            assert np.abs(traj[p, :, :2] - xy).max() <= 2e-6, (name, p)
        eng.close()


def test_astar_initialiser_edge_cases():
    """start and goal in the same cell (A* path = one cell, a 3-point spline), and an end point outside the checker
    boundaries (the reference indexes the matrix out of range: IndexError here, before any launch)."""
    from _gpu_util import engine_config
    from nfopp_b200.engine import PlannerEngine
    from oracle import astar as A
    from oracle import checkers as OC
    lock = U.load("lockstep_car.npz")
    bounds, res, T = (0.0, 3.0, 0.0, 3.0), 0.5, 12
    obs = np.array([[1.5, 1.5]])
    cfg = engine_config("car", lock, rng_mode="philox", problems=2, trajectory_length=T, checker="circle",
                        checker_bounds=bounds, border=bounds)
    eng = PlannerEngine(cfg)
    eng.set_obstacles(obs)
    starts = np.array([[0.55, 0.55, 0.0], [0.2, 2.8, 1.0]], dtype=np.float32)
    goals = np.array([[0.9, 0.8, 0.5], [2.8, 0.2, -1.0]], dtype=np.float32)
    eng.set_endpoints(starts, goals)
    eng.init_trajectories()
    n_cells = eng.init_astar(res).cpu().numpy()
    traj = eng["traj"].cpu().numpy()
    label = lambda p: OC.circle_labels(p, obs, 0.3, bounds)                  # noqa: E731
    for p in range(2):
        xy, cells, _ = A.initial_trajectory_xy(starts[p], goals[p], bounds, res, label, T)
        assert n_cells[p] == len(cells) and np.abs(traj[p, :, :2] - xy).max() <= 2e-6, p
    assert n_cells[0] == 1
    eng.set_endpoints(starts, np.array([[3.7, 0.8, 0.5], [2.8, 0.2, -1.0]], dtype=np.float32))
    with pytest.raises(IndexError):
        eng.init_astar(res)
    eng.close()


def test_astar_initialiser_through_the_factory_and_on_a_grid():
    """PlannerFactory resolves AstarTrajectoryInitializer by name (scripts/run_bench_mr.py:23-26); BatchedPlanner.init with
    astar_resolution on occupancy grids gives collision-free cell paths where the straight line collides."""
    from nfopp_b200.batched import BatchedPlanner, forest_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from oracle import astar as A
    from oracle import checkers as OC
    P = 5
    starts, goals, _, grids, _, cell = B.make_random_grid_batch(P, seed=11)
    cfg = forest_config(P, cell=cell, seed=3)
    bp = BatchedPlanner(cfg)
    bp.init(starts, goals, grids=grids, astar_resolution=4 * cell)
    paths = bp.get_path()
    assert np.isfinite(paths).all()
    for p in range(P):
        grid = np.asarray(grids[p])
        label = lambda q: OC.grid_labels(q, grid, 0.0, 0.0, cfg.grid_cell, cfg.checker_bounds)   # noqa: E731
        xy, _, _ = A.initial_trajectory_xy(np.asarray(starts[p], dtype=np.float32), np.asarray(goals[p], dtype=np.float32),
                                           cfg.checker_bounds, 4 * cell, label, cfg.trajectory_length)
        assert np.abs(paths[p, 1:-1, :2] - xy).max() <= 2e-6, p
    # single-problem API: factory config with the A* initialiser
    from nfopp_b200.collision_checker import CircleDirectedCollisionChecker
    from nfopp_b200.planner_factory import DEFAULT_PARAMETERS, PlannerFactory
    from nfopp_b200.utils.attribute_dict import AttributeDict
    params = AttributeDict(DEFAULT_PARAMETERS)
    params["trajectory_initializer"] = AttributeDict(name="AstarTrajectoryInitializer", resolution=0.25)
    params["device"] = "cuda"
    fx = U.load("astar.npz")
    cc = CircleDirectedCollisionChecker(0.3, (0.0, 3.0, 0.0, 3.0))
    cc.update_obstacle_points(fx["disc_coarse/obstacles"])
    planner = PlannerFactory.make_constrained_onf_planner(cc, params)
    planner.init(fx["disc_coarse/start"], fx["disc_coarse/goal"], (0.0, 3.0, 0.0, 3.0))
    path = planner.get_path()
    assert path.shape == (102, 3)
    assert np.abs(path[1:-1, :2] - fx["disc_coarse/trajectory"][:, :2]).max() <= 2e-6
    # a checker that reports boundaries but does not OR them into its labels (the notebook's MapCollisionChecker): the A*
    # cell grid spans get_boundaries() (astar_trajectory_initializer.py:29-34), the labels stay the checker's own
    from nfopp_b200.collision_checker import GridCollisionChecker, MapImage
    rs = np.random.RandomState(5)
    occ = (rs.rand(60, 60) < 0.08).astype(np.uint8)
    occ[:6, :6] = 0
    occ[-8:, -8:] = 0
    bounds = (0.0, 3.0, 0.0, 3.0)
    gc = GridCollisionChecker(MapImage(60, 60, 0.0, 0.0, 0.05, occ), bounds)
    gplanner = PlannerFactory.make_constrained_onf_planner(gc, params)
    start, goal = np.array([0.12, 0.12, 0.0]), np.array([2.8, 2.8, 0.0])
    gplanner.init(start, goal, bounds)
    gpath = gplanner.get_path()
    label = lambda q: OC.grid_labels(q, occ, 0.0, 0.0, 0.05)            # noqa: E731  (no boundary OR)
    xy, _, _ = A.initial_trajectory_xy(start.astype(np.float32), goal.astype(np.float32), bounds, 0.25, label, 100)
    assert np.abs(gpath[1:-1, :2] - xy).max() <= 2e-6


# ---------------------------------------------------------------------------------------- BASELINE's full sizes
def _full_size_checks(bp, label_problem, steps, sample_every):
    """size-independent properties of a running batch: every training label of every problem equals the reference
    checker's on the very poses the engine labelled (bit-exact), training sets are complete and finite, the resampled
    replay buffer keeps its capacity, paths stay finite, no problem raised a status flag."""
    eng = bp.engine
    bp.step(steps)
    torch.cuda.synchronize()
    P = eng.config.problems
    n = eng["train_n"].cpu().numpy()
    pos = eng["train_pos"].cpu().numpy()
    y = eng["train_y"].cpu().numpy() > 0.5
    assert (n == n[0]).all() and n[0] == pos.shape[1]                        # 99 coarse + 100 replay + 10 uniform
    assert np.isfinite(pos).all()
    flips = 0
    for p in range(0, P, sample_every):
        flips += int((label_problem(p, pos[p, :n[p]]) != y[p, :n[p]]).sum())
    assert flips == 0, f"{flips} collision labels differ from the reference checker"
    assert (eng["buf_n"].cpu().numpy() == eng.config.collision_point_count).all()
    assert bool(torch.isfinite(bp.paths()).all()) and int(np.abs(bp.status()).max()) == 0
    return y.mean()


def test_full_size_corridor_batch_1024_problems():
    """BASELINE configs[1] at its full size (1024 corridor problems, one obstacle set each): the properties above for every
    problem, plus batch-size independence - problems 0..7 of the 1024 evolve bit-identically in a batch of 8."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config, init_onf_weights
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from oracle import checkers as OC
    P = 1024
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=0)
    cfg = corridor_config(P, seed=21)
    w0 = init_onf_weights(cfg.onf, P, True, 99, "cuda")
    bp = BatchedPlanner(cfg)
    bp.init(starts, goals, obstacles=obstacles, onf_weights=w0)
    frac = _full_size_checks(bp, lambda p, q: OC.circle_labels(q, np.asarray(obstacles[p]), cfg.robot_radius, cfg.checker_bounds),
                             steps=12, sample_every=1)
    assert 0.0 < frac < 1.0                                                  # both classes present
    small = BatchedPlanner(corridor_config(8, seed=21))
    small.init(starts[:8], goals[:8], obstacles=obstacles[:8], onf_weights=w0[:8])
    small.step(12)
    torch.cuda.synchronize()
    for name in ("traj", "onf_w", "lam", "mu"):
        assert torch.equal(small.engine[name], bp.engine[name][:8]), name


def test_full_size_forest_batch_16384_problems():
    """BASELINE configs[2] at its full size: 16 384 random-obstacle worlds (one 100 x 100 occupancy grid each), each with its
    own ONF; labels of a 1-in-64 sample of the problems checked bit-exact against the reference's grid checker."""
    from nfopp_b200.batched import BatchedPlanner, forest_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from oracle import checkers as OC
    P = 16384
    starts, goals, _, grids, _, cell = B.make_random_grid_batch(P, seed=0)
    cfg = forest_config(P, cell=cell, seed=5)
    bp = BatchedPlanner(cfg)
    bp.init(starts, goals, grids=grids)
    _full_size_checks(bp, lambda p, q: OC.grid_labels(q, np.asarray(grids[p]), 0.0, 0.0, cfg.grid_cell, cfg.checker_bounds),
                      steps=6, sample_every=64)
    w = bp.engine["onf_w"]
    assert not torch.equal(w[0], w[1]) and bool(torch.isfinite(w).all())


def test_full_size_wide_onf_linearity_and_permutation():
    """BASELINE configs[4] shape (4 x 256 hidden, 2^20 points in 16 weight sets): size-independent properties of the
    tensor-core ONF - the backward pass is linear in the upstream gradient (doubling dL/dz doubles every gradient bit for
    bit: powers of two pass through the bf16 hi/lo split exactly), logits follow a permutation of the points exactly and
    the weight gradient is permutation-invariant up to summation order."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
    import onf_sweep
    from nfopp_b200 import _lib as L
    from nfopp_b200.engine import OnfSpec
    P, n, H, layers, impl = 16, 65536, 256, 4, 3
    gen = torch.Generator().manual_seed(3)
    cc = OnfSpec(0.0, 1.0, True, True, H, layers).to_c()
    w8, lay = onf_sweep.init_weights(cc, 8, gen)
    w = w8.repeat(2, 1)[:P].contiguous().cuda()
    x = torch.rand(P, n, 3, generator=gen)
    x[..., :2] *= 3.0
    x[..., 2] *= 6.2831853
    x = x.cuda()
    gz = (torch.randn(P, n, generator=gen) * 1e-3).cuda()
    lib, s = L.lib(), L.stream_ptr()
    scratch = L.onf_scratch(cc, P, impl, "cuda", n, True)

    def run(xx, gg):
        z = torch.zeros(P, n, device="cuda")
        gx = torch.zeros(P, n, 3, device="cuda")
        gw = torch.zeros(P, lay.n_params, device="cuda")
        L.check(lib.nfopp_onf_backward(C.byref(cc), L.ptr(w), L.ptr(xx), None, P, n, 0, L.ptr(gg), None, 0.0, 1.0, L.ptr(z),
                                       L.ptr(gx), L.ptr(gw), impl, L.ptr(scratch), s))
        torch.cuda.synchronize()
        return z, gx, gw

    z1, gx1, gw1 = run(x, gz)
    z2, gx2, gw2 = run(x, 2 * gz)
    assert bool(torch.isfinite(gw1).all()) and float(gw1.abs().max()) > 0
    assert torch.equal(z1, z2) and torch.equal(gx2, 2 * gx1) and torch.equal(gw2, 2 * gw1)
    perm = torch.randperm(n, generator=gen).cuda()
    z3, gx3, gw3 = run(x[:, perm].contiguous(), gz[:, perm].contiguous())
    assert torch.equal(z3, z1[:, perm]) and torch.equal(gx3, gx1[:, perm])
    assert float((gw3 - gw1).abs().max()) <= 1e-5 * float(gw1.abs().max())


@pytest.mark.gpu
@pytest.mark.parametrize("hidden,layers,n", [(256, 4, 4096), (512, 4, 4096), (100, 2, 4096)])
def test_full_width_onf_against_the_float64_oracle(hidden, layers, n):
    """VERDICT r1 'pin what can be pinned at full width': BASELINE configs[3] / [4] network shapes (4 x 256: fused chain kernel,
    4 x 512: per-layer GEMMs; and the reference shape on the streamed kernel with 37 tiles per weight set) at 4096 points per
    weight set against the oracle evaluated in float64 - logits everywhere, input gradients and weight gradients after
    silencing the points whose ReLU masks are rounding-sensitive (some pre-activation within 1e-5 of zero: their upstream
    gradient is set to zero on both sides, which removes them from every sum).  Tolerance: north_star's 1e-3 relative."""
    cfg = O.OnfConfig(mean=0.1, sigma=1.3, use_cos=True, angle=True, hidden=hidden, layers=layers)
    gen = torch.Generator().manual_seed(11 * hidden + layers)
    P = 2
    w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(P)])
    x = torch.rand(P, n, cfg.in_dim, generator=gen) * 4 - 1
    gz = torch.randn(P, n, generator=gen) / n
    ref, risky_total = [], 0
    for p in range(P):
        zo, cache = O.onf_forward(w[p], cfg, x[p], dtype=torch.float64)
        risky = torch.zeros(n, dtype=torch.bool)
        for pre in cache["pre_act"]:
            risky |= (pre.abs() < 1e-5).any(1)
        gz[p, risky] = 0.0
        risky_total += int(risky.sum())
        gwo, gxo = O.onf_backward(w[p], cfg, cache, gz[p])
        ref.append((zo, gxo, gwo))
    z, gx, gw = _onf_call(cfg, w, x, gz, impl=IMPLS["tensor" if hidden <= 111 else "wide"])
    worst = [0.0, 0.0, 0.0]
    for p in range(P):
        zo, gxo, gwo = ref[p]
        for k, (a, b) in enumerate(((z[p], zo), (gx[p], gxo), (gw[p], gwo))):
            worst[k] = max(worst[k], U.relerr(a.double(), b))
    print(f"{layers} x {hidden} @ {n} points per weight set: z {worst[0]:.2e} gx {worst[1]:.2e} gw {worst[2]:.2e} "
          f"(max-norm relative, vs float64; {risky_total} of {P * n} points silenced)")
    assert max(worst) < ONF_TOL


# ------------------------------------------------------------ large obstacle clouds (ros/grid_map.py -> disc checker)
@pytest.mark.gpu
def test_disc_robot_against_a_large_occupancy_point_cloud():
    """ros/grid_map.py turns an occupancy grid into one obstacle point per occupied cell (here ~6 k points,
    beyond the 2048 the kernels stage in shared memory, shared by all problems).  The engine's labels of its own training
    poses and its path evaluation stay bit-exact against the oracle's circle checker."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.grid_map import GridMap
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from nfopp_b200.utils.position2 import Position2
    from oracle import checkers as C
    from oracle import evaluate as EV
    grid, _, cell = B.make_blob_grid(200, 200, 0.15, 0.015, blob_cells=10, seed=4)
    gm = GridMap(np.asarray(grid, dtype=np.float32), cell, Position2(0.0, 0.0, 0.0))
    cloud = gm.as_point_cloud()
    assert 2048 < len(cloud) < 20000
    P = 6
    starts, goals = B.free_endpoints(grid, cell, P, seed=2, margin_cells=12)
    starts = np.concatenate([starts[:, :2], np.zeros((P, 1))], 1)
    goals = np.concatenate([goals[:, :2], np.zeros((P, 1))], 1)
    bounds = gm.boundaries
    cfg = corridor_config(P, seed=7, env_shared=True, max_obstacles=len(cloud), robot_radius=0.03,
                          checker_bounds=bounds, border=bounds)
    bp = BatchedPlanner(cfg)
    bp.init(starts, goals, obstacles=cloud)
    eng = bp.engine
    label = lambda poses: C.circle_labels(poses, cloud, 0.03, bounds)
    seen = set()
    for _ in range(3):
        bp.step(2, graph=False)
        torch.cuda.synchronize()
        n = eng["train_n"].cpu().numpy()
        pos, y = eng["train_pos"].cpu().numpy(), eng["train_y"].cpu().numpy()
        for p in range(P):
            want = label(pos[p, :n[p]])
            assert np.array_equal(y[p, :n[p]] > 0.5, want), p
            seen.update(want.tolist())
    assert seen == {False, True}
    collides, length, hits = eng.evaluate_paths(3)
    paths = eng.paths().cpu().numpy()
    for p in range(P):
        c, ln, h = EV.evaluate_path(paths[p], label, 3)
        assert bool(collides[p]) == c and int(hits[p]) == h
        assert abs(float(length[p]) - ln) <= 1e-6 * ln
    eng.close()


# ---------------------------------------------------------------------------------------- PathPostprocessor
@pytest.mark.gpu
def test_path_postprocessor_matches_reference_golden():
    """nfopp_postprocess_paths vs the live reference's PathPostprocessor.process (tests/golden/postprocess.npz): the same
    number of poses (integer decisions: distance filter, point count, direction filter - exact) and the poses to 1e-9
    (float64 spline; the reference solves the collocation system with LAPACK, the kernel by banded elimination)."""
    from nfopp_b200.path_postprocessor import PathPostprocessor
    from nfopp_b200.utils.position2 import Position2
    fx = np.load(os.path.join(os.path.dirname(__file__), "golden", "postprocess.npz"))
    for name in fx["names"]:
        path, want = fx[f"{name}/path"], fx[f"{name}/result"]
        min_d, step = (float(v) for v in fx[f"{name}/params"])
        pp = PathPostprocessor(min_d, step)
        got = pp.process(Position2.from_vec(path.copy())).as_vec()
        assert got.shape == want.shape, (name, got.shape, want.shape)
        assert np.abs(got - want).max() <= 1e-9, (name, np.abs(got - want).max())
    # paths of fewer than 3 nodes come back unchanged (:14-15)
    two = Position2.from_vec(np.array([[0, 0, 0], [1, 0, 0]], dtype=np.float32))
    assert PathPostprocessor().process(two) is two


@pytest.mark.gpu
def test_path_postprocessor_batch_vs_oracle_on_planner_paths():
    """BatchedPlanner.postprocess_paths on the paths of a running corridor batch (ragged pose counts) vs
    oracle/postprocess.py path by path; a degenerate path (all nodes equal) is flagged -1 instead of mis-fitted; a too small
    output buffer is refused."""
    from nfopp_b200.batched import BatchedPlanner, corridor_config
    from nfopp_b200.path_postprocessor import PathPostprocessor
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    from oracle import postprocess as PP
    P = 48
    starts, goals, _, obstacles = B.make_corridor_batch(P, seed=21)
    bp = BatchedPlanner(corridor_config(P, seed=2))
    bp.init(starts, goals, obstacles=obstacles)
    bp.step(60)
    paths = bp.get_path()
    out, count = bp.postprocess_paths()
    out, count = out.cpu().numpy(), count.cpu().numpy()
    assert len(set(count.tolist())) > 1                                   # ragged
    for p in range(P):
        want = PP.process(paths[p])
        assert count[p] == len(want), (p, count[p], len(want))
        assert np.abs(out[p, :count[p]] - want).max() <= 1e-9, p
        assert not out[p, count[p]:].any()
    bp.engine.close()
    stuck = np.tile(np.array([[1.0, 2.0, 0.5]], dtype=np.float32), (1, 30, 1))
    _, c = PathPostprocessor().process_batch(stuck)
    assert int(c[0]) == -1
    with pytest.raises(ValueError):
        PathPostprocessor(distance_step=0.001).process_batch(paths[:2], max_points=paths.shape[1])


# ------------------------------------------------------------------ corridor mazes (config 2, bench-mr-style grids)
@pytest.mark.gpu
def test_astar_initialiser_on_corridor_mazes():
    """Config 2's corridor mazes (own generator with bench-mr's parameters): the straight-line initial path crosses the walls,
    the A* initialiser's path follows the corridors - far fewer colliding poses before the first step - and the planner then
    solves about as many or more mazes under the reference's stop rule."""
    from nfopp_b200.batched import BatchedPlanner, forest_config
    from nfopp_b200.test_environment_builder import TestEnvironmentBuilder as B
    P = 64
    starts, goals, bounds, grids, origin, cell = B.make_corridor_grid_batch(P, seed=2)
    res = {}
    for name, resolution in (("line", None), ("astar", 0.09)):
        bp = BatchedPlanner(forest_config(P, cell=cell, seed=5))
        bp.init(starts, goals, grids=grids, astar_resolution=resolution)
        hits0 = bp.evaluate_paths(4)[2].cpu().numpy()
        out = bp.solve(max_iterations=500, min_iterations=200, check_every=50)
        assert bool(torch.isfinite(out["paths"]).all())
        res[name] = (hits0, int(out["solved"].sum()))
        bp.engine.close()
    assert res["astar"][0].mean() < 0.4 * res["line"][0].mean(), (res["astar"][0].mean(), res["line"][0].mean())
    assert (res["astar"][0] <= res["line"][0]).mean() > 0.8
    # (the optimisation is chaotic, BASELINE.md §2, and the solved counts are small: scripts/maze_probe.py printed 14 vs 11 of
    # 64 with bf16 operands, 3 vs 7 of 32 with fp16 operands - only a gross regression of the A* path is asserted)
    assert res["astar"][1] + 6 >= res["line"][1] and res["astar"][1] > 0, (res["astar"][1], res["line"][1])

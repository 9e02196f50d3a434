"""GPU: nfopp_onf_forward_static - the allocation-free variant of nfopp_onf_forward (operand copy of the weights in a static
device array inside the library; no counterpart in the reference).

Collected last on purpose (file name): a failure here hides nothing after it.  The test
passes alone in a fresh process and after the other tests (the shared array keeps whatever an earlier call left in it)."""
import ctypes as C

import pytest
import torch

import _util as U
from oracle import onf as O

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 4 * 1024 * 1024          # include/nfopp.h: NFOPP_STATIC_SCRATCH_BYTES
ONF_TOL = 1e-3                                  # the ONF tolerance of tests/test_gpu_parity.py (north_star: 1e-3 relative)


def _spec(cfg):
    from nfopp_b200.engine import OnfSpec
    return OnfSpec(cfg.mean, cfg.sigma, cfg.use_cos, cfg.angle, cfg.hidden, cfg.layers, cfg.bias).to_c()


def _forward_existing(L, cc, w, x, count):
    """nfopp_onf_forward(impl = 2) with caller-owned scratch (a torch allocation - exactly what the new entry point avoids)."""
    P, n, _ = x.shape
    z = torch.full((P, n), float("nan"), device="cuda")
    scratch = L.onf_scratch(cc, P, 2, "cuda", n, False)
    L.check(L.lib().nfopp_onf_forward(C.byref(cc), L.ptr(w), L.ptr(x), L.ptr(count), P, n, L.ptr(z), 2, L.ptr(scratch),
                                      L.stream_ptr()))
    torch.cuda.synchronize()
    return z


def _forward_static(L, cc, w, x, count, z):
    P, n, _ = x.shape
    L.check(L.lib().nfopp_onf_forward_static(C.byref(cc), L.ptr(w), L.ptr(x), L.ptr(count), P, n, L.ptr(z), L.stream_ptr()))
    torch.cuda.synchronize()
    return z


def test_static_scratch_forward_matches_the_existing_entry_point_on_two_calls_in_a_row():
    from nfopp_b200 import _lib as L
    gen = torch.Generator().manual_seed(17)
    # two calls in a row with DIFFERENT network shapes, weight-set counts and point counts: the second call finds the first
    # call's operand copy in the shared array and must not be affected by it (nor by whatever earlier tests left there)
    calls = [("car", 5, 209, [209, 100, 1, 209, 57]), ("holonomic", 3, 330, None), ("car", 2, 99, None)]
    for name, P, n, count in calls:
        cfg = U.ONF_CFGS[name]
        cc = _spec(cfg)
        w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(P)]).cuda().contiguous()
        x = (torch.rand(P, n, cfg.in_dim, generator=gen) * 4 - 1).cuda().contiguous()
        dcount = None if count is None else torch.tensor(count, dtype=torch.int32, device="cuda")
        want = _forward_existing(L, cc, w, x, dcount)
        z = torch.full((P, n), float("nan"), device="cuda")          # output buffer allocated by the caller, before the call
        got = _forward_static(L, cc, w, x, dcount, z)
        for p in range(P):
            c = n if count is None else count[p]
            assert bool(torch.isfinite(got[p, :c]).all())
            assert U.relerr(got[p, :c].cpu(), want[p, :c].cpu()) < ONF_TOL
            assert torch.equal(got[p, :c], want[p, :c])              # same kernels, same arithmetic: bit-identical
            zo, _ = O.onf_forward(w[p].cpu(), cfg, x[p, :c].cpu())
            assert U.relerr(got[p, :c].cpu(), zo) < ONF_TOL          # and right against the oracle


def test_static_scratch_limit_is_enforced_with_a_clear_error():
    from nfopp_b200 import _lib as L
    cfg = U.ONF_CFGS["car"]
    cc = _spec(cfg)
    per_set = 150528                                                 # fp16 hi/lo operand copy of the reference network
    p_max = STATIC_SCRATCH_BYTES // per_set
    n = 16
    gen = torch.Generator().manual_seed(3)
    w1 = O.init_blob(cfg, True, gen)
    for P, ok in ((p_max, True), (p_max + 1, False)):
        w = w1.repeat(P, 1).cuda().contiguous()
        x = torch.rand(P, n, cfg.in_dim, generator=gen).cuda()
        z = torch.zeros(P, n, device="cuda")
        rc = L.lib().nfopp_onf_forward_static(C.byref(cc), L.ptr(w), L.ptr(x), None, P, n, L.ptr(z), L.stream_ptr())
        torch.cuda.synchronize()
        if ok:
            assert rc == 0, L.lib().nfopp_last_error()
            want = _forward_existing(L, cc, w, x, None)
            assert torch.equal(z, want)
        else:
            assert rc != 0
            msg = L.lib().nfopp_last_error().decode()
            assert "static scratch" in msg and str(STATIC_SCRATCH_BYTES) in msg and f"at most {p_max}" in msg
    # a network shape outside the streamed tensor path is refused too (no silent fallback)
    wide = O.OnfConfig(hidden=256, layers=4, use_cos=True, angle=True)
    ccw = _spec(wide)
    w = O.init_blob(wide, True, gen)[None].cuda()
    x = torch.rand(1, n, 3, generator=gen).cuda()
    z = torch.zeros(1, n, device="cuda")
    assert L.lib().nfopp_onf_forward_static(C.byref(ccw), L.ptr(w), L.ptr(x), None, 1, n, L.ptr(z), L.stream_ptr()) != 0

#!/usr/bin/env python
"""ONF kernel microbench + precision report (run on a B200): for each implementation (fma / tensor) the max-norm
relative error of logits, input gradients and weight gradients against the CPU oracle, and the device time of the
three launch shapes of a planner step (candidates fwd 199 pts, train fwd+bwd 209 pts, path fwd+dgrad 99 pts) at
P problems with per-problem weights.   python scripts/onf_microbench.py [--problems 1024] [--json out.json]"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import _util as U  # noqa: E402
from nfopp_b200 import _lib as L  # noqa: E402
from nfopp_b200.engine import OnfSpec  # noqa: E402
from oracle import onf as O  # noqa: E402

IMPLS = {"fma": 1, "tensor": 2}


def run(cc, impl, w, x, mode, gz=None, aux=None, want_w=True, want_x=True, reps=0):
    P, n, D = x.shape
    lib = L.lib()
    s = L.stream_ptr()
    scratch = L.onf_scratch(cc, P, impl, "cuda", n, True)
    tiles = lib.nfopp_onf_tiles(C.byref(cc), n, impl)
    z = torch.zeros(P, n, device="cuda")
    gx = torch.zeros(P, n, D, device="cuda") if want_x else None
    gpart = torch.zeros(P, tiles, w.shape[1], device="cuda") if want_w else None
    gw = torch.zeros(P, w.shape[1], device="cuda")

    def call():
        if mode < 0:
            L.check(lib.nfopp_onf_forward(C.byref(cc), L.ptr(w), L.ptr(x), None, P, n, L.ptr(z), impl, L.ptr(scratch), s))
        else:
            L.check(lib.nfopp_onf_backward(C.byref(cc), L.ptr(w), L.ptr(x), None, P, n, mode, L.ptr(gz), L.ptr(aux), 0.7,
                                           1.0, L.ptr(z), L.ptr(gx), L.ptr(gpart), impl, L.ptr(scratch), s))
    call()
    if want_w and mode >= 0:
        L.check(lib.nfopp_onf_reduce_grad(C.byref(cc), L.ptr(gpart), None, P, n, impl, L.ptr(gw), s))
    torch.cuda.synchronize()
    ms = None
    if reps:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
    return z, gx, gw, ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--problems", type=int, default=1024)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    cfg = U.ONF_CFGS["car"]
    cc = OnfSpec(cfg.mean, cfg.sigma, cfg.use_cos, cfg.angle, cfg.hidden, cfg.layers, cfg.bias).to_c()
    gen = torch.Generator().manual_seed(0)
    out = {"problems": args.problems, "precision": {}, "ms": {}}
    # precision on 4 problems x 209 points against the float64-free CPU oracle
    P, n = 4, 209
    w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(P)])
    x = torch.rand(P, n, 3, generator=gen) * 3
    y = (torch.rand(P, n, generator=gen) < 0.3).float()
    for name, impl in IMPLS.items():
        z, gx, gw, _ = run(cc, impl, w.cuda(), x.cuda(), 1, aux=y.cuda())
        ez = egx = egw = 0.0
        for p in range(P):
            zo, cache = O.onf_forward(w[p], cfg, x[p])
            _, g = O.bce_with_logits_grad(zo, y[p])
            gwo, gxo = O.onf_backward(w[p], cfg, cache, g)
            ez = max(ez, U.relerr(z[p].cpu(), zo)); egx = max(egx, U.relerr(gx[p].cpu(), gxo))
            egw = max(egw, U.relerr(gw[p].cpu(), gwo))
        out["precision"][name] = {"z": ez, "gx": egx, "gw": egw}
        print(f"precision[{name}]: logits {ez:.2e}  input-grad {egx:.2e}  weight-grad {egw:.2e}", flush=True)
    # timing
    P = args.problems
    w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(8)]).repeat(P // 8 + 1, 1)[:P].contiguous().cuda()
    for name, impl in IMPLS.items():
        res = {}
        for label, n, mode, ww, wx in [("candidates_fwd_199", 199, -1, False, False), ("train_fwd_bwd_209", 209, 1, True, False),
                                       ("path_fwd_dgrad_99", 99, 2, False, True)]:
            x = (torch.rand(P, n, 3, generator=gen) * 3).cuda()
            y = (torch.rand(P, n, generator=gen) < 0.3).float().cuda()
            _, _, _, ms = run(cc, impl, w, x, mode, aux=y, want_w=ww, want_x=wx, reps=10)
            res[label] = ms
        out["ms"][name] = res
        print(f"ms[{name}] at P={P} (API calls; the tensor path includes packing the operand copy): "
              + "  ".join(f"{k} {v:.3f}" for k, v in res.items()), flush=True)
    if args.json:
        with open(args.json, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""BASELINE config 5: ONF fwd+bwd microbench sweep (run on a B200).

For hidden in {64,128,256,512} x layers in {2,4} x points per weight set in {256,4096,65536} x total points
2^16..2^24 (totals beyond 2^20 points run as consecutive launches over 2^20-point groups of weight sets, so the activation
workspace stays at its 2^20-point size; the timed region covers all of them): one "point" = forward + weight-gradient + input-gradient (nfopp_onf_backward, BCE upstream gradient,
per-problem weights), timed with CUDA events over `reps` calls after warm-up.  Reports Mpts/s, algorithmic TFLOP/s
(6 x MAC x points) and the fraction of the measured bf16 tensor peak (MEASURED_PEAKS.json).
    python scripts/onf_sweep.py [--quick] [--json out.json] [--impl 3]
Inputs: x,y ~ U(bounds), theta ~ U(0,2pi), labels Bernoulli(0.3), weights = oracle init (seed 0)."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from nfopp_b200 import _lib as L  # noqa: E402
from nfopp_b200.engine import OnfSpec  # noqa: E402


def mac_per_point(f_pos, f_in, H, layers):
    return 2 * f_pos + f_in * H + (layers - 1) * H * H + (H + f_in)


def init_weights(cc, P, gen):
    """Random weights with the reference's init scales (onf_model.py:30-31 normal encoder, nn.Linear uniform)."""
    lay = L.OnfLayout()
    L.check(L.lib().nfopp_onf_get_layout(C.byref(cc), C.byref(lay)))
    w = torch.zeros(P, lay.n_params)
    w[:, lay.off_enc_w:lay.off_enc_w + 2 * lay.f_pos] = torch.randn(P, 2 * lay.f_pos, generator=gen)
    w[:, lay.off_enc_b:lay.off_enc_b + lay.f_pos] = (torch.rand(P, lay.f_pos, generator=gen) - 0.5) * 1.4
    if lay.n_angle:
        w[:, lay.off_ang_b:lay.off_ang_b + lay.n_angle] = torch.rand(P, lay.n_angle, generator=gen) * 6.28
        w[:, lay.off_ang_f:lay.off_ang_f + lay.n_angle] = torch.arange(1, lay.n_angle + 1).float().repeat(P, 1) % 10 + 1
    for l in range(cc.layers):
        k = lay.f_in if l == 0 else cc.hidden
        b = 1.0 / k ** 0.5
        w[:, lay.off_w[l]:lay.off_w[l] + cc.hidden * k] = (torch.rand(P, cc.hidden * k, generator=gen) * 2 - 1) * b
        w[:, lay.off_b[l]:lay.off_b[l] + cc.hidden] = (torch.rand(P, cc.hidden, generator=gen) * 2 - 1) * b
    k = cc.hidden + lay.f_in
    w[:, lay.off_w_out:lay.off_w_out + k] = (torch.rand(P, k, generator=gen) * 2 - 1) / k ** 0.5
    return w, lay


def bench_case(H, layers, n, total, impl, reps, gen, chunk_points=1 << 20):
    P = max(1, total // n)
    Pc = max(1, min(P, chunk_points // n))                 # weight sets per launch
    cc = OnfSpec(0.0, 1.0, True, True, H, layers).to_c()
    w8, lay = init_weights(cc, min(P, 8), gen)
    w = w8.repeat((P + 7) // 8, 1)[:P].contiguous().cuda()
    x = torch.rand(P, n, 3, generator=gen)
    x[..., :2] *= 3.0
    x[..., 2] *= 6.2831853
    x = x.cuda()
    y = (torch.rand(P, n, generator=gen) < 0.3).float().cuda()
    lib, s = L.lib(), L.stream_ptr()
    scratch = L.onf_scratch(cc, Pc, impl, "cuda", n, True)
    z = torch.zeros(P, n, device="cuda")
    gx = torch.zeros(P, n, 3, device="cuda")
    gw = torch.zeros(P, lay.n_params, device="cuda")

    def call():
        for p0 in range(0, P, Pc):
            pc = min(Pc, P - p0)
            L.check(lib.nfopp_onf_backward(C.byref(cc), L.ptr(w[p0:]), L.ptr(x[p0:]), None, pc, n, 1, None, L.ptr(y[p0:]), 0.0, 1.0,
                                           L.ptr(z[p0:]), L.ptr(gx[p0:]), L.ptr(gw[p0:]), impl, L.ptr(scratch), s))
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    pts = P * n
    flops = 6.0 * mac_per_point(lay.f_pos, lay.f_in, H, layers) * pts
    finite = bool(torch.isfinite(gw).all() and torch.isfinite(gx).all())
    del scratch
    return {"hidden": H, "layers": layers, "points_per_problem": n, "problems": P, "launch_groups": (P + Pc - 1) // Pc, "points": pts, "ms": ms,
            "mpts_per_s": pts / ms / 1e3, "tflops": flops / ms / 1e9, "finite": finite,
            "scratch_gb": int(torch.cuda.max_memory_allocated()) / 1e9}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--json", default=None)
    ap.add_argument("--impl", type=int, default=3)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    peak = 1400.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        pass
    gen = torch.Generator().manual_seed(0)
    cases = []
    if a.quick:
        grid = [(256, 4, 4096, 1 << 18), (256, 4, 256, 1 << 18), (512, 4, 4096, 1 << 18), (128, 2, 4096, 1 << 18),
                (64, 2, 65536, 1 << 18), (256, 4, 65536, 1 << 20), (256, 4, 65536, 1 << 24)]
    else:
        grid = [(H, Ly, n, tot) for H in (64, 128, 256, 512) for Ly in (2, 4) for n in (256, 4096, 65536)
                for tot in (1 << 16, 1 << 18, 1 << 20) if tot >= n]
        grid += [(256, 4, 65536, 1 << 22), (128, 4, 65536, 1 << 22)]
        grid += [(H, Ly, 65536, 1 << 24) for H in (64, 128, 256, 512) for Ly in (2, 4)]      # BASELINE configs[4]'s upper end
    for H, Ly, n, tot in grid:
        torch.cuda.reset_peak_memory_stats()
        r = bench_case(H, Ly, n, tot, a.impl, a.reps, gen)
        r["frac_of_peak"] = r["tflops"] / peak
        cases.append(r)
        print(f"H={H:4d} L={Ly} n/problem={n:6d} P={r['problems']:5d} pts=2^{tot.bit_length() - 1:2d}  {r['ms']:9.3f} ms  "
              f"{r['mpts_per_s']:8.1f} Mpts/s  {r['tflops']:7.1f} TFLOP/s  {100 * r['frac_of_peak']:5.1f}% of {peak:.0f}  "
              f"finite={r['finite']}", flush=True)
    if a.json:
        json.dump({"peak_bf16_tflops": peak, "impl": a.impl, "cases": cases}, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()

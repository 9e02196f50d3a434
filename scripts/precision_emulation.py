#!/usr/bin/env python
"""CPU emulation of the operand-precision modes of the fused ONF backward (which operands need an fp16 lo part?).
Operands are rounded exactly as the kernel does (fp16 hi, optional fp16 lo = fp16(x - hi)), products accumulate in float64,
ReLU masks are the reference's; the upstream gradient is scaled to [1, 2) by a power of two before the split.
Reports, over seeds, the worst max-norm relative error of the weight gradient and of the input gradient vs float64.
    python scripts/precision_emulation.py [--seeds 12]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from oracle import onf as O


def split(x, mode):
    """-> list of parts whose sum approximates x: 'x' exact, 'h' fp16 hi, 'hl' fp16 hi + fp16 lo."""
    if mode == "x":
        return [x]
    hi = x.to(torch.float32).to(torch.float16).to(torch.float64)
    if mode == "h":
        return [hi]
    lo = (x - hi).to(torch.float32).to(torch.float16).to(torch.float64)
    return [hi, lo]


def mm(a_parts, b_parts, f):
    """sum of part products, dropping lo*lo (the kernel never issues it)."""
    out = 0
    for i, a in enumerate(a_parts):
        for j, b in enumerate(b_parts):
            if i == 1 and j == 1:
                continue
            out = out + f(a, b)
    return out


def backward(blob, cfg, cache, gz, mg, mw, ma):
    """mg / mw / ma: precision mode of the gradient operands, the weight operands and the activation operands (H1, E)."""
    p = {k: v.double() for k, v in O.unpack(blob, cfg).items()}
    hs = [h.double() for h in cache["hs"]]
    H = cfg.hidden
    e = hs[0]
    gz = gz.double()
    m = float(gz.abs().max())
    S = 2.0 ** (-np.floor(np.log2(m))) if m > 0 else 1.0
    g = {}
    g["w_out"] = torch.cat([hs[-1], e], 1).t() @ gz                  # CUDA cores / row H of wgrad1 (see below)
    g["b_out"] = gz.sum()[None]
    gh = gz[:, None] * p["w_out"][None, :H] * (hs[2] > 0)             # gH2, fp32 on the CUDA cores
    g["b1"] = gh.sum(0)
    gh_parts = split(gh * S, mg)
    g["W1"] = mm(gh_parts, split(hs[1], ma), lambda a, b: a.t() @ b) / S           # wgrad2
    gd = mm(gh_parts, split(p["W1"], mw), lambda a, b: a @ b) / S                 # dgrad2 (unscaled view)
    gh1 = gd * (hs[1] > 0)
    g["b0"] = gh1.sum(0)
    # layer-1 operand with the skip row: [gH1 ; gz]  x  [W0 ; w_out[H:]]
    gfull = torch.cat([gh1, gz[:, None]], 1)
    wfull = torch.cat([p["W0"], p["w_out"][None, H:]], 0)
    gf_parts = split(gfull * S, mg)
    ge = mm(gf_parts, split(wfull, mw), lambda a, b: a @ b) / S                   # dgrad1
    dW = mm(gf_parts, split(e, ma), lambda a, b: a.t() @ b) / S                   # wgrad1
    g["W0"] = dW[:H]
    g["w_out"] = torch.cat([g["w_out"][:H], dW[H]])
    F = cfg.f_pos
    gp = ge[:, :F] * cache["dphi"].double()
    g["enc_w"] = gp.t() @ cache["u"].double()
    g["enc_b"] = gp.sum(0)
    gx = [(gp @ p["enc_w"]) / cfg.sigma]
    if cfg.angle:
        ga = ge[:, F:] * cache["dpsi"].double()
        g["ang_b"] = (ga * p["ang_f"][None]).sum(0)
        g["ang_f"] = (ga * (cache["theta"].double() + p["ang_b"][None])).sum(0)
        gx.append((ga * p["ang_f"][None]).sum(1, keepdim=True))
    else:
        g["ang_b"] = g["ang_f"] = torch.zeros(0)
    return O.pack({k: v.float() for k, v in g.items()}, cfg).double(), torch.cat(gx, 1)


def rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-300))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=12)
    args = ap.parse_args()
    cfgs = {"car": (O.OnfConfig(mean=0.0, sigma=1.0, use_cos=True, angle=True), 209),
            "image": (O.OnfConfig(mean=0.0, sigma=7.0, use_cos=True), 199),
            "holonomic": (O.OnfConfig(mean=1.5, sigma=1.0), 209)}
    modes = {"bwd x1 (g h, W h, act h)": ("h", "h", "h"), "g hl, W h, act h": ("hl", "h", "h"),
             "g hl, W hl, act h": ("hl", "hl", "h"), "g hl, W h, act hl": ("hl", "h", "hl"),
             "g h, W hl, act hl": ("h", "hl", "hl"), "bwd x3 (all hl)": ("hl", "hl", "hl")}
    out = {}
    for name, (cfg, n) in cfgs.items():
        for up in ("randn", "bce"):
            worst = {k: [0.0, 0.0] for k in modes}
            for seed in range(args.seeds):
                gen = torch.Generator().manual_seed(seed)
                w = O.init_blob(cfg, True, gen)
                x = torch.rand(n, cfg.in_dim, generator=gen) * 3
                z, cache = O.onf_forward(w, cfg, x)
                if up == "randn":
                    gz = torch.randn(n, generator=gen) / n
                else:
                    y = (torch.rand(n, generator=gen) < 0.3).float()
                    gz = (torch.sigmoid(z) - y) / n
                ref_w, ref_x = backward(w, cfg, cache, gz, "x", "x", "x")
                for k, (mg, mw, ma) in modes.items():
                    gw, gx = backward(w, cfg, cache, gz, mg, mw, ma)
                    worst[k][0] = max(worst[k][0], rel(gw, ref_w))
                    worst[k][1] = max(worst[k][1], rel(gx, ref_x))
            out[f"{name}/{up}"] = {k: {"gw": v[0], "gx": v[1]} for k, v in worst.items()}
            print(name, up, {k: f"gw {v[0]:.1e} gx {v[1]:.1e}" for k, v in worst.items()}, flush=True)
    json.dump(out, open(os.path.join(ROOT, "profiles", "r02_precision_emulation.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

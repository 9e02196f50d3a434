"""Oracle: occupancy neural field (ONF) forward / analytic backward, torch float32 on CPU.

Follows neural_field_optimal_planner/onf_model.py:8-50 and angle_encoder.py:7-18, generalised to
hidden width H and depth L (the reference hard-codes H=100, L=2, onf_model.py:18-23).  No autograd:
the backward pass is written out so that it can be compared term by term with the CUDA kernels.

Packed parameter blob (float32, every segment starts on a multiple of 4 floats):
    enc_w [F_pos,2] | enc_b [F_pos] | ang_b [A] | ang_f [A] | W_0 [H,F_in] | b_0 [H] | W_l [H,H] | b_l [H] ...
    | w_out [H+F_in] | b_out [1]
"""
from dataclasses import dataclass

import numpy as np
import torch


@dataclass(frozen=True)
class OnfConfig:
    mean: float = 0.0
    sigma: float = 1.0
    use_cos: bool = False      # onf_model.py:14,40-43
    angle: bool = False        # onf_model.py:11-13
    hidden: int = 100          # onf_model.py:18-23 (fixed 100 there)
    layers: int = 2            # number of hidden Linear+ReLU blocks (fixed 2 there)
    bias: bool = True          # onf_model.py:29

    @property
    def f_pos(self):
        return 200 if self.use_cos else 100

    @property
    def n_angle(self):
        return 20 if self.angle else 0   # angle_encoder.py:7,20-22

    @property
    def f_in(self):
        return self.f_pos + self.n_angle

    @property
    def in_dim(self):
        return 3 if self.angle else 2


def _align4(n):
    return (n + 3) // 4 * 4


def onf_layout(cfg: OnfConfig):
    """name -> (offset, shape); plus total length.  Must equal nfopp_b200.layout.onf_layout (tested)."""
    segs = [("enc_w", (cfg.f_pos, 2)), ("enc_b", (cfg.f_pos,)), ("ang_b", (cfg.n_angle,)), ("ang_f", (cfg.n_angle,))]
    for l in range(cfg.layers):
        k = cfg.f_in if l == 0 else cfg.hidden
        segs += [(f"W{l}", (cfg.hidden, k)), (f"b{l}", (cfg.hidden,))]
    segs += [("w_out", (cfg.hidden + cfg.f_in,)), ("b_out", (1,))]
    out, off = {}, 0
    for name, shape in segs:
        out[name] = (off, shape)
        off = _align4(off + int(np.prod(shape)))
    return out, off


def unpack(blob: torch.Tensor, cfg: OnfConfig):
    lay, total = onf_layout(cfg)
    assert blob.numel() == total, (blob.numel(), total)
    return {k: blob[o:o + int(np.prod(s))].view(*s) if int(np.prod(s)) else blob[o:o].view(*s)
            for k, (o, s) in lay.items()}


def pack(named: dict, cfg: OnfConfig) -> torch.Tensor:
    lay, total = onf_layout(cfg)
    blob = torch.zeros(total, dtype=torch.float32)
    for k, (o, s) in lay.items():
        n = int(np.prod(s))
        if n:
            blob[o:o + n] = torch.as_tensor(named[k], dtype=torch.float32).reshape(-1)
    return blob


def trainable_mask(cfg: OnfConfig) -> torch.Tensor:
    """1 where the blob element is a parameter Adam updates, 0 for alignment padding (and enc_b if bias=False)."""
    lay, total = onf_layout(cfg)
    m = torch.zeros(total, dtype=torch.float32)
    for k, (o, s) in lay.items():
        if k == "enc_b" and not cfg.bias:
            continue
        m[o:o + int(np.prod(s))] = 1
    return m


def blob_from_state_dict(sd: dict, cfg: OnfConfig) -> torch.Tensor:
    """Reference ONF.state_dict() (keys per onf_model.py:18-29, angle_encoder.py:10-12) -> packed blob."""
    named = {"enc_w": sd["encoding_layer.weight"],
             "enc_b": sd["encoding_layer.bias"] if cfg.bias else torch.zeros(cfg.f_pos),
             "w_out": sd["mlp2.0.weight"].reshape(-1), "b_out": sd["mlp2.0.bias"]}
    for l in range(cfg.layers):
        named[f"W{l}"] = sd[f"mlp.{2 * l}.weight"]
        named[f"b{l}"] = sd[f"mlp.{2 * l}.bias"]
    if cfg.angle:
        named["ang_b"] = sd["_angle_encoder._biases"]
        named["ang_f"] = sd["_angle_encoder._frequencies"]
    else:
        named["ang_b"] = named["ang_f"] = torch.zeros(0)
    return pack(named, cfg)


def init_blob(cfg: OnfConfig, use_normal_init: bool, generator: torch.Generator) -> torch.Tensor:
    """Fresh initial weights with torch's nn.Linear default distribution (kaiming_uniform(a=sqrt 5) =
    U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias), onf_model.py:18-31; angle phases
    U(-pi, pi), frequencies 1..10 twice (angle_encoder.py:10-13).  Distribution-equivalent to the
    reference constructor, not stream-identical (parity tests import the reference's own weights)."""
    def lin(out_f, in_f):
        bound = 1.0 / np.sqrt(in_f)
        w = (torch.rand(out_f, in_f, generator=generator) * 2 - 1) * bound
        b = (torch.rand(out_f, generator=generator) * 2 - 1) * bound
        return w, b
    named = {}
    for l in range(cfg.layers):
        named[f"W{l}"], named[f"b{l}"] = lin(cfg.hidden, cfg.f_in if l == 0 else cfg.hidden)
    w, b = lin(1, cfg.hidden + cfg.f_in)
    named["w_out"], named["b_out"] = w.reshape(-1), b
    named["enc_w"], named["enc_b"] = lin(cfg.f_pos, 2)
    if not cfg.bias:
        named["enc_b"] = torch.zeros(cfg.f_pos)
    if use_normal_init:
        named["enc_w"] = torch.randn(cfg.f_pos, 2, generator=generator)
    if cfg.angle:
        named["ang_b"] = (torch.rand(20, generator=generator) * 2 - 1) * np.pi
        f = torch.linspace(1, 10, 10)
        named["ang_f"] = torch.cat([f, f])
    else:
        named["ang_b"] = named["ang_f"] = torch.zeros(0)
    return pack(named, cfg)


def onf_forward(blob: torch.Tensor, cfg: OnfConfig, x: torch.Tensor, dtype=torch.float32):
    """x [N, 2|3] float32 -> (logit [N], cache).  onf_model.py:33-50, angle_encoder.py:15-18.  dtype=float64 gives the
    exact-arithmetic reference the full-width GPU tests compare with (the reference itself computes in float32)."""
    p = unpack(blob.to(dtype), cfg)
    x = x.to(dtype)
    u = (x[:, :2] - cfg.mean) / cfg.sigma                                   # onf_model.py:38
    pre = u @ p["enc_w"].t() + p["enc_b"]                                   # onf_model.py:39
    if cfg.use_cos:                                                          # onf_model.py:40-43
        phi = torch.cat([torch.sin(pre[:, :100]), torch.cos(pre[:, 100:])], 1)
        dphi = torch.cat([torch.cos(pre[:, :100]), -torch.sin(pre[:, 100:])], 1)
    else:
        phi, dphi = torch.sin(pre), torch.cos(pre)
    cache = {"u": u, "dphi": dphi}
    if cfg.angle:                                                            # angle_encoder.py:15-18
        a = (x[:, 2:3] + p["ang_b"][None]) * p["ang_f"][None]
        psi = torch.cat([torch.sin(a[:, :10]), torch.cos(a[:, 10:])], 1)
        dpsi = torch.cat([torch.cos(a[:, :10]), -torch.sin(a[:, 10:])], 1)
        e = torch.cat([phi, psi], 1)                                         # onf_model.py:45-46
        cache.update(dpsi=dpsi, theta=x[:, 2:3])
    else:
        e = phi
    hs, pre_act = [e], []
    for l in range(cfg.layers):                                              # onf_model.py:47
        pre_act.append(hs[-1] @ p[f"W{l}"].t() + p[f"b{l}"])
        hs.append(torch.relu(pre_act[-1]))
    z = torch.cat([hs[-1], e], 1) @ p["w_out"] + p["b_out"]                  # onf_model.py:48-49
    cache["hs"] = hs
    cache["pre_act"] = pre_act                                               # (tests: points whose ReLU masks are rounding-sensitive)
    return z, cache


def onf_backward(blob: torch.Tensor, cfg: OnfConfig, cache: dict, gz: torch.Tensor):
    """Given dL/dz [N] return (dL/dblob, dL/dx [N, 2|3]).  Analytic restatement of what autograd does
    for onf_model.py:33-50."""
    hs, H = cache["hs"], cfg.hidden
    e = hs[0]
    p = unpack(blob.to(e.dtype), cfg)
    g = {k: torch.zeros_like(v) for k, v in p.items()}
    gz = gz.to(e.dtype)
    g["w_out"] = torch.cat([hs[-1], e], 1).t() @ gz
    g["b_out"] = gz.sum()[None]
    gh = gz[:, None] * p["w_out"][None, :H]
    ge = gz[:, None] * p["w_out"][None, H:]
    for l in reversed(range(cfg.layers)):
        gh = gh * (hs[l + 1] > 0)
        g[f"W{l}"] = gh.t() @ hs[l]
        g[f"b{l}"] = gh.sum(0)
        gh = gh @ p[f"W{l}"]
    ge = ge + gh
    F = cfg.f_pos
    gp = ge[:, :F] * cache["dphi"]
    g["enc_w"] = gp.t() @ cache["u"]
    if cfg.bias:
        g["enc_b"] = gp.sum(0)
    gx = [(gp @ p["enc_w"]) / cfg.sigma]
    if cfg.angle:
        ga = ge[:, F:] * cache["dpsi"]
        g["ang_b"] = (ga * p["ang_f"][None]).sum(0)
        g["ang_f"] = (ga * (cache["theta"] + p["ang_b"][None])).sum(0)      # trained, SURVEY.md §8 a3
        gx.append((ga * p["ang_f"][None]).sum(1, keepdim=True))
    return pack(g, cfg), torch.cat(gx, 1)


def bce_with_logits_grad(z: torch.Tensor, y: torch.Tensor):
    """nn.BCEWithLogitsLoss(reduction='mean') value and dL/dz  (nerf_opt_planner.py:25,88)."""
    loss = (torch.clamp(z, min=0) - z * y + torch.log1p(torch.exp(-z.abs()))).mean()
    return loss, (torch.sigmoid(z) - y) / z.numel()


def adam_step(w, m, v, g, step, lr, beta1, beta2, eps=1e-8, mask=None):
    """torch.optim.Adam single-tensor update (torch/optim/adam.py _single_tensor_adam; defaults
    amsgrad=False, weight_decay=0, maximize=False); call sites planner_factory.py:54,56,68,70.
    `step` is the 1-based step index of this update.  In place."""
    if mask is not None:
        g = g * mask
    m.lerp_(g, 1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
    w.addcdiv_(m, denom, value=-(lr / bc1))
    return w, m, v

"""ONF - occupancy neural field with the reference's constructor, state_dict keys and call signature
(onf_model.py:7-50), evaluated by the fused CUDA kernels of libnfopp.

All parameters are views into ONE packed float32 blob (layout: include/nfopp.h) so that the planner engine
can adopt the module's weights in place; forward() returns the collision *logit* [N,1] and is
differentiable w.r.t. the input poses and every parameter through the analytic backward kernel.
CPU tensors are accepted for construction / state_dict handling only: forward() on a non-CUDA module
raises - there is no CPU fallback."""
import ctypes as C

import torch
from torch import nn

from . import _lib as L
from .angle_encoder import AngleEncoder
from .engine import OnfSpec, onf_segments


class _OnfFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, module, x, *params):
        spec, flat = module.spec, module._flat
        cfg = spec.to_c()
        n = x.shape[0]
        xin = x[:, :module.in_dim].detach().to(torch.float32).contiguous()
        z = torch.empty(n, dtype=torch.float32, device=flat.device)
        with torch.cuda.device(flat.device):
            scratch = L.onf_scratch(cfg, 1, module.impl, flat.device, n, False)
            L.check(L.lib().nfopp_onf_forward(C.byref(cfg), L.ptr(flat), L.ptr(xin), None, 1, n, L.ptr(z),
                                              module.impl, L.ptr(scratch), L.stream_ptr()))
        ctx.module, ctx.xin, ctx.x_cols = module, xin, x.shape[1]
        return z

    @staticmethod
    def backward(ctx, gz):
        module, xin = ctx.module, ctx.xin
        spec, flat = module.spec, module._flat
        cfg = spec.to_c()
        n = xin.shape[0]
        lib = L.lib()
        impl = module.impl
        tiles = lib.nfopp_onf_tiles(C.byref(cfg), n, impl)
        gz = gz.detach().to(torch.float32).contiguous()
        gpart = torch.zeros(tiles * flat.numel(), dtype=torch.float32, device=flat.device)
        gflat = torch.empty_like(flat)
        gx = torch.zeros(n, ctx.x_cols, dtype=torch.float32, device=flat.device)
        gxin = torch.empty(n, module.in_dim, dtype=torch.float32, device=flat.device)
        with torch.cuda.device(flat.device):
            s = L.stream_ptr()
            scratch = L.onf_scratch(cfg, 1, impl, flat.device, n, True)
            L.check(lib.nfopp_onf_backward(C.byref(cfg), L.ptr(flat), L.ptr(xin), None, 1, n, 0, L.ptr(gz), None,
                                           0.0, 1.0, None, L.ptr(gxin), L.ptr(gpart), impl, L.ptr(scratch), s))
            L.check(lib.nfopp_onf_reduce_grad(C.byref(cfg), L.ptr(gpart), None, 1, n, impl, L.ptr(gflat), s))
        gx[:, :module.in_dim] = gxin
        grads = []
        for name, _ in module._param_segments():
            off, shape = module._segments[name]
            numel = 1
            for d in shape:
                numel *= d
            grads.append(gflat[off:off + numel].view(module._param_shape(name)).clone())
        return (None, gx, *grads)


class ONF(nn.Module):
    def __init__(self, mean, sigma, use_cos=False, use_normal_init=False, bias=True, angle_encoding=False,
                 hidden=100, layers=2):
        super().__init__()
        # construction order = the reference's (onf_model.py:10-31) so torch.manual_seed(s) yields the same
        # initial weights as the reference module for hidden=100, layers=2
        self._angle_encoder = AngleEncoder() if angle_encoding else None
        angle_dim = self._angle_encoder.encoding_dimension if angle_encoding else 0
        feature_dim = (200 if use_cos else 100) + angle_dim
        self._use_cos = use_cos
        blocks = []
        for l in range(layers):
            blocks += [nn.Linear(feature_dim if l == 0 else hidden, hidden), nn.ReLU()]
        self.mlp = nn.Sequential(*blocks)
        self.mlp2 = nn.Sequential(nn.Linear(hidden + feature_dim, 1))
        self._mean, self._sigma = mean, sigma
        self.encoding_layer = nn.Linear(2, feature_dim - angle_dim, bias=bias)
        if use_normal_init:
            nn.init.normal_(self.encoding_layer.weight)
        self.spec = OnfSpec(mean=float(mean), sigma=float(sigma), use_cos=bool(use_cos), angle=bool(angle_encoding),
                            hidden=int(hidden), layers=int(layers), bias=bool(bias))
        self.in_dim = 3 if angle_encoding else 2
        self.impl = 0               # kernel implementation: 0 auto (tcgen05 when the shape allows), 1 FMA, 2 tensor
        self._segments, self._n_params = onf_segments(self.spec)
        self._flat = None
        self._repack()

    # ---- packed storage -----------------------------------------------------------------------------
    def _param_segments(self):
        """(segment name, parameter) in nn.Module.parameters() order."""
        out = []
        for name, prm in self.named_parameters():
            if name == "_angle_encoder._biases":
                seg = "ang_b"
            elif name == "_angle_encoder._frequencies":
                seg = "ang_f"
            elif name.startswith("mlp2."):
                seg = "w_out" if name.endswith("weight") else "b_out"
            elif name.startswith("mlp."):
                idx = int(name.split(".")[1]) // 2
                seg = (f"W{idx}" if name.endswith("weight") else f"b{idx}")
            else:
                seg = "enc_w" if name.endswith("weight") else "enc_b"
            out.append((seg, prm))
        return out

    def _param_shape(self, seg):
        return (1, self._segments[seg][1][0]) if seg == "w_out" else self._segments[seg][1]

    def _is_packed(self):
        if self._flat is None:
            return False
        base = self._flat.data_ptr()
        return all(p.device == self._flat.device and p.data_ptr() == base + 4 * self._segments[seg][0]
                   for seg, p in self._param_segments())

    def _repack(self, storage=None):
        """(Re)creates the packed blob on the parameters' device - or adopts `storage`, e.g. a planner
        engine's weight buffer - copies the current values in, and re-points every parameter at its view."""
        pairs = self._param_segments()
        device = pairs[0][1].device
        flat = torch.zeros(self._n_params, dtype=torch.float32, device=device) if storage is None else storage
        assert flat.numel() == self._n_params and flat.dtype == torch.float32
        with torch.no_grad():
            for seg, prm in pairs:
                off, _ = self._segments[seg]
                view = flat[off:off + prm.numel()].view(prm.shape)
                view.copy_(prm.detach().to(flat.device))
                prm.data = view
        self._flat = flat

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        self._flat = None            # .to()/.cuda() re-created the parameter tensors
        self._repack()
        return out

    def packed_parameters(self) -> torch.Tensor:
        if not self._is_packed():
            self._repack()
        return self._flat

    def weights_version(self):
        """Changes whenever torch wrote into the weights: through the packed blob or through any parameter (each parameter
        keeps its own version counter - `.data = view` does not share the blob's).  Kernel-side updates do not move it."""
        flat = self.packed_parameters()
        return (flat._version,) + tuple(p._version for _, p in self._param_segments())

    # ---- reference call signature ---------------------------------------------------------------------
    def forward(self, x):
        flat = self.packed_parameters()
        L.require_cuda(flat.device)
        if x.device != flat.device:
            raise ValueError("ONF input and parameters live on different devices")
        params = [p for _, p in self._param_segments()]
        return _OnfFunction.apply(self, x, *params)[:, None]

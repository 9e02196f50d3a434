"""Quick GPU check of the streamed tensor kernel (onf_tcs.cu) against the oracle: z / gx / gw errors per ONF config,
per gradient segment, for both tensor-path generations (NFOPP_TC_LEGACY=1 selects the first)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import _util as U
from oracle import onf as O
import test_gpu_parity as T

def main():
    torch.manual_seed(0)
    for name, n in [("car", 209), ("car", 99), ("holonomic", 209), ("image", 199), ("car", 17), ("car", 500)]:
        cfg = U.ONF_CFGS[name]
        gen = torch.Generator().manual_seed(3)
        P = 2
        w = torch.stack([O.init_blob(cfg, True, gen) for _ in range(P)])
        x = torch.rand(P, n, cfg.in_dim, generator=gen) * 3
        gz = torch.randn(P, n, generator=gen) / n
        want_gw = True
        try:
            if want_gw:
                z, gx, gw = T._onf_call(cfg, w, x, gz, impl=2)
            else:
                z, _, _ = T._onf_call(cfg, w, x, None, impl=2)
        except Exception as e:
            print(name, n, "FAILED:", e)
            continue
        for p in range(P):
            zo, cache = O.onf_forward(w[p], cfg, x[p])
            gwo, gxo = O.onf_backward(w[p], cfg, cache, gz[p])
            line = f"{name:10s} n={n:4d} p={p} z {U.relerr(z[p], zo):.2e}"
            if want_gw:
                line += f" gx {U.relerr(gx[p], gxo):.2e} gw {U.relerr(gw[p], gwo):.2e} |"
                lay, _ = O.onf_layout(cfg)
                for seg, (off, shape) in lay.items():
                    k = int(np.prod(shape))
                    if k:
                        line += f" {seg}:{U.relerr(gw[p, off:off + k], gwo[off:off + k]):.1e}"
            print(line, flush=True)

if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Phase trace of the streamed fused ONF kernel (onf_tcs.cu built with -DNFOPP_TC_TRACE: `make -C .../csrc trace`).
Runs the corridor workload for a few steps and prints, for the first and last CTA of the three ONF launches of a step
(candidates forward, train, path), the clock64 stamps of worker thread 0, the MMA-issue warp and the weight-stream warp
relative to the CTA's start (the CTAs are persistent: the stamps are those of the CTA's LAST problem).  Diagnostic only - the product library is libnfopp.so.   python scripts/tcs_trace.py"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from nfopp_b200 import _lib as L  # noqa: E402

L.LIB_PATH = os.path.join(os.path.dirname(L.LIB_PATH), "libnfopp_trace.so")
import bench as B  # noqa: E402

WORKER = {0: "start", 31: "P0 setup done", 32: "P0 inputs staged", 33: "P0 encode done", 34: "P0 fwd1 done", 35: "P0 ep1 signalled", 29: "first problem done", 30: "last problem begins", 1: "setup done (last problem)", 2: "inputs staged", 3: "encode done", 4: "fwd1 done", 5: "ep1 signalled",
          6: "fwd2 done", 7: "logit partials", 8: "dL/dz", 9: "gH2 signalled", 10: "wgrad2+dgrad2 done", 11: "ep3 signalled",
          12: "dgrad1(0) done", 13: "ep4(0) signalled", 14: "dgrad1(1) done", 15: "ep4(1) signalled", 16: "wgrad1 done",
          20: "tiles done", 22: "hand-over buffer free", 24: "partials staged", 25: "dW1 out", 26: "dW2 out", 27: "gradient handed over", 28: "CTA done"}
ISSUER = {0: "start", 1: "fwd1 issued", 2: "ep1 ready", 3: "fwd2 issued", 4: "gH2 ready", 5: "W2 slot full", 6: "bwdA issued",
          7: "gH1 ready", 8: "W1 slot full", 9: "ep4(0) ready", 10: "wgrad1(0)+dgrad1(1) issued", 11: "ep4(1) ready",
          12: "wgrad1(1) issued"}
PRODUCER = {1: "fwd ring streamed", 2: "fwd ring drained", 3: "bwd blocks issued"}
OPTIMISER = {1: "gradient ready (last problem)", 2: "Adam + operand copy done"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--problems", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=12)
    args = ap.parse_args()
    cfg, env, starts, goals, _ = B.build_workload("corridor", args.problems, 0)
    from nfopp_b200.batched import BatchedPlanner
    pl = BatchedPlanner(cfg)
    pl.init(starts, goals, **env)
    pl.step(args.steps)
    torch.cuda.synchronize()
    lib = L.lib()
    lib.nfopp_debug_tcs_trace.restype = C.c_int
    lib.nfopp_debug_tcs_trace.argtypes = [C.c_void_p]
    buf = np.zeros((3, 2, 4, 64), dtype=np.int64)
    rc = lib.nfopp_debug_tcs_trace(buf.ctypes.data_as(C.c_void_p))
    assert rc == 0, rc
    for kind, kname in enumerate(["candidates fwd", "train", "path fwd+dgrad"]):
        for ci, cname in enumerate(["first CTA", "last CTA"]):
            t = buf[kind, ci]
            t0 = t[0, 0]
            print(f"== {kname} / {cname}")
            for role, names in enumerate([WORKER, ISSUER, PRODUCER, OPTIMISER]):
                prev, parts = t0, []
                for k in (sorted(names) if role else [0, 31, 32, 33, 34, 35, 29, 30] + [k for k in sorted(names) if k not in (0, 29, 30, 31, 32, 33, 34, 35)]):
                    v = t[role, k]
                    if v == 0 or v < t0:
                        continue
                    parts.append(f"{names[k]} @{v - t0} (+{v - prev})")
                    prev = v
                print(f"  {['worker', 'issuer', 'producer', 'optimiser'][role]:9s}: " + " | ".join(parts))
            if t[1, 44]:
                print(f"  issuer, last tile (cycles): {t[1, 44]} of which waiting for the encoding {t[1, 40]}, the weight ring {t[1, 41]}, "
                      f"weight blocks {t[1, 42]}, the epilogue warps {t[1, 43]}")
            if t[3, 3] or t[3, 4]:
                print(f"  optimiser, last problem (cycles): waiting for chunks {t[3, 3]}, arithmetic + stores {t[3, 4]}")


if __name__ == "__main__":
    main()

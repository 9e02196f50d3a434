#!/usr/bin/env python
"""Phase trace of the fused tcgen05 ONF kernel (onf_tc.cu built with -DNFOPP_TC_TRACE: `make -C .../csrc trace`).
Runs the corridor workload for a few steps and prints, for the first and last CTA of the three ONF launches of a
step (candidates forward, train, path), the clock64 deltas between the phase stamps of the two worker groups.
Diagnostic only - the product library is libnfopp.so.   python scripts/tc_trace.py [--problems 1024]"""
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

TILE = ["tile start", "P0 inputs", "enc fg0 -> ready", "enc fg1 -> ready", "wait fwd1", "ep1 -> ready", "wait fwd2",
        "ep2a logit partials (sync)", "ep2b dL/dz (sync)", "gH2 -> ready", "wait dgrad2", "ep3 -> ready", "wait dgrad1",
        "ep4 mb0 -> ready", "ep4 mb1 -> ready", "tile end"]
TAIL = {80: "loop end", 81: "CTA sync", 82: "TMEM drain", 83: "Adam", 84: "repack"}


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
    lib.nfopp_debug_tc_trace.restype = C.c_int
    lib.nfopp_debug_tc_trace.argtypes = [C.c_void_p]
    buf = np.zeros((3, 2, 2, 96), dtype=np.int64)
    rc = lib.nfopp_debug_tc_trace(buf.ctypes.data_as(C.c_void_p))
    assert rc == 0, rc
    for kind, kname in enumerate(["candidates fwd", "train", "path fwd+dgrad"]):
        for ci, cname in enumerate(["first CTA", "last CTA"]):
            t = buf[kind, ci]
            t0 = t[0, 0]
            print(f"== {kname} / {cname}: setup {t[0, 1] - t0} cycles")
            for g in range(2):
                for tl in range(5):
                    base = 2 + 16 * tl
                    if base + 15 >= 80 or t[g, base] == 0 or t[g, base] < t0:
                        continue
                    prev = t[g, base]
                    parts = []
                    for k in range(1, 16):
                        v = t[g, base + k]
                        if v == 0 or v < prev:
                            continue
                        parts.append(f"{TILE[k]} {v - prev}")
                        prev = v
                    print(f"  group {g} tile {g + 2 * tl}: start @{t[g, base] - t0}  total {prev - t[g, base]}")
                    print("     " + " | ".join(parts))
            prev = None
            parts = []
            for k in sorted(TAIL):
                v = t[0, k]
                if v == 0 or v < t0:
                    continue
                parts.append(f"{TAIL[k]} @{v - t0}" + (f" (+{v - prev})" if prev else ""))
                prev = v
            print("  tail: " + " | ".join(parts))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Phase trace of the fused wide chain kernel (onf_wide.cu built with -DNFOPP_TC_TRACE: `make -C .../csrc trace`): clock64
stamps of one epilogue warp and of the MMA-issue warp of CTA 0 on its third row block, 4 x 256 network, fwd + bwd.
Diagnostic only.   python scripts/wide_trace.py [--hidden 256 --layers 4]"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "scripts")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from nfopp_b200 import _lib as L  # noqa: E402

L.LIB_PATH = os.path.join(os.path.dirname(L.LIB_PATH), "libnfopp_trace.so")
import onf_sweep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=256)
    ap.add_argument("--layers", type=int, default=4)
    args = ap.parse_args()
    r = onf_sweep.bench_case(args.hidden, args.layers, 4096, 1 << 18, 3, 3, torch.Generator().manual_seed(0))
    print(f"{r['ms']:.3f} ms  {r['tflops']:.1f} TFLOP/s")
    lib = L.lib()
    lib.nfopp_debug_wc_trace.restype = C.c_int
    lib.nfopp_debug_wc_trace.argtypes = [C.c_void_p]
    t = np.zeros(96, dtype=np.int64)
    assert lib.nfopp_debug_wc_trace(t.ctypes.data_as(C.c_void_p)) == 0
    t0 = t[0]
    nph = 2 * args.layers
    names = ["start", "vectors", "tfull", "g0", "g1", "g2", "g3", "end"]
    for ph in range(nph):
        e = t[8 * ph:8 * ph + 8]
        parts, prev = [], e[0]
        for k in range(1, 8):
            if e[k] >= prev and e[k] > 0:
                parts.append(f"{names[k]} +{e[k] - prev}")
                prev = e[k]
        m0, m1 = t[64 + 2 * ph], t[65 + 2 * ph]
        kind = f"fwd {ph}" if ph < args.layers else f"dgrad {2 * args.layers - 1 - ph}"
        print(f"phase {ph} ({kind}): epilogue warp @{e[0] - t0}: " + " | ".join(parts) + f"   || MMA warp: first group ready @{m0 - t0}, "
              f"all issued @{m1 - t0} (+{m1 - m0})")
    print(f"row block total (phase 0 start -> last phase end): {t[8 * (nph - 1) + 7] - t0} cycles")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Times the UNMODIFIED reference planner (imported from /root/reference through oracle/reference_bridge.py) next to the
oracle port, in the build container (no GPU; /root/reference does not exist on the GPU box, so this cannot run there):

    python scripts/time_reference_here.py > profiles/r02_reference_cpu_container.json

scripts/benchmark.py:99-111 style: perf_counter around the step() loop after 50 warm-up steps, 1 torch thread; the
reference as shipped has torch anomaly detection ON (nerf_opt_planner.py:15), reported both ways."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

torch.set_num_threads(1)
from oracle import reference_bridge as RB


def time_reference(name, anomaly, steps=300, warm=50):
    planner, env, checker = RB.make_reference_case(name)
    torch.autograd.set_detect_anomaly(anomaly)
    for _ in range(warm):
        planner.step()
    t0 = time.perf_counter()
    for _ in range(steps):
        planner.step()
    dt = time.perf_counter() - t0
    torch.autograd.set_detect_anomaly(False)
    return {"steps_per_s": steps / dt, "ms_per_step": 1e3 * dt / steps, "steps": steps, "warmup": warm}


def time_oracle(name, steps=300, warm=50):
    import _util as U
    from oracle.planner import LiveSource
    fx = U.load(f"lockstep_{name}.npz")
    op = U.oracle_planner(name, fx, "free/state0")
    src = LiveSource(np.random.RandomState(400), torch.Generator().manual_seed(100))
    for _ in range(warm):
        op.step(src)
    t0 = time.perf_counter()
    for _ in range(steps):
        op.step(src)
    dt = time.perf_counter() - t0
    return {"steps_per_s": steps / dt, "ms_per_step": 1e3 * dt / steps, "steps": steps, "warmup": warm}


def main():
    cpu = "unknown"
    for ln in open("/proc/cpuinfo"):
        if ln.startswith("model name"):
            cpu = ln.split(":", 1)[1].strip()
            break
    out = {"where": "build container (CPU only)", "host_cpu": f"{cpu} ({os.cpu_count()} logical cores)", "torch": torch.__version__,
           "numpy": np.__version__, "threads": 1,
           "what": "unmodified /root/reference planner.step() vs oracle/planner.py (the port bench.py times on the GPU box), "
                   "same configs, same machine, back to back"}
    for name in ("car", "corridor", "holonomic"):
        out[name] = {"reference_anomaly_off": time_reference(name, False), "reference_anomaly_on_as_shipped": time_reference(name, True),
                     "oracle_port": time_oracle(name)}
        out[name]["port_over_reference_anomaly_off"] = out[name]["oracle_port"]["steps_per_s"] / out[name]["reference_anomaly_off"]["steps_per_s"]
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

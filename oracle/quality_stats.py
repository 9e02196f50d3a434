"""Distributional path-quality statistics of the ORACLE (= the reference's algorithm) on a fixed problem, for
the distributional parity test (SURVEY.md §7 hard part 1, protocol iv): the loop is chaotic, so beyond ~10
steps only statistics over seeds can be compared.  Writes tests/golden/quality_<case>.json.

    python -m oracle.quality_stats corridor 32 1000
"""
import json
import os
import sys

import numpy as np
import torch

from . import checkers as C
from . import onf as O
from .planner import LiveSource, OraclePlanner

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))


def run(case="corridor", seeds=32, steps=1000, every=250):
    import _util as U
    torch.set_num_threads(1)
    fx = U.load(f"lockstep_{case}.npz")
    obs, bounds = fx["meta/obstacles"], tuple(fx["meta/bounds"])
    cfg = U.PLANNER_CFGS[case]
    rows = []
    for seed in range(seeds):
        gen = torch.Generator().manual_seed(1000 + seed)
        op = OraclePlanner(cfg, O.init_blob(cfg.onf, True, gen), U.labeller_for(case, fx))
        op.init(fx["free/state0/start"], fx["free/state0/goal"], fx["free/state0/border"],
                LiveSource(np.random.RandomState(seed)))
        src = LiveSource(np.random.RandomState(10_000 + seed), torch.Generator().manual_seed(seed))
        row = []
        for k in range(steps):
            op.step(src)
            if k % every == every - 1:
                path = op.get_path().astype(np.float64)
                lab = U.labeller_for(case, fx)(path)
                d = np.diff(path, axis=0)
                abar = path[:-1, 2] + ((d[:, 2] + np.pi) % (2 * np.pi) - np.pi) / 2
                row.append(dict(step=k + 1, collisions=int(lab.sum()), length=float(np.linalg.norm(d[:, :2], axis=1).sum()),
                                nonholonomic=float(np.abs(d[:, 0] * np.sin(abar) - d[:, 1] * np.cos(abar)).max())))
        rows.append(row)
        print(seed, [r["collisions"] for r in row], flush=True)
    out = dict(case=case, seeds=seeds, steps=steps, every=every, runs=rows,
               solved_rate={str(r["step"]): float(np.mean([run_[i]["collisions"] == 0 for run_ in rows]))
                            for i, r in enumerate(rows[0])})
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                        f"quality_{case}.json")
    json.dump(out, open(path, "w"), indent=1)
    print(out["solved_rate"])


if __name__ == "__main__":
    run(sys.argv[1] if len(sys.argv) > 1 else "corridor", int(sys.argv[2]) if len(sys.argv) > 2 else 32,
        int(sys.argv[3]) if len(sys.argv) > 3 else 1000)

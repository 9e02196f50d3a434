import copy, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np, torch
import _util as U
from _gpu_util import load_state
from test_gpu_parity import _grid_case, _grid_engine
for hidden, layers, impl in [(256, 4, "wide"), (128, 3, "wide"), (128, 3, "fma")]:
    op, src, pc, grid, cell, bounds, border = _grid_case(hidden, layers)
    for k in range(3):
        before = copy.deepcopy(op.state)
        var = op.step(src)
        eng = _grid_engine(pc, grid, cell, bounds, border, onf_impl=impl)
        load_state(eng, before); eng.inject_variates(var); eng.step(1); torch.cuda.synchronize()
        em = U.relerr(eng["onf_m"][0].cpu().numpy(), op.state["m"].numpy())
        dw = np.abs(eng["onf_w"][0].cpu().numpy() - op.state["w"].numpy())
        pz = U.relerr(eng["path_z"][0].cpu().numpy(), op.last["path_z"].numpy())
        pg = U.relerr(eng["path_gq"][0].cpu().numpy(), op.last["path_gq"].numpy())
        tm = U.relerr(eng["traj_m"][0].cpu().numpy(), op.state["traj_m"].numpy())
        # same path launch but with the ORACLE's post-Adam weights
        eng2 = _grid_engine(pc, grid, cell, bounds, border, onf_impl=impl)
        print(f"H={hidden} L={layers} {impl} k={k}: m err {em:.2e}  |dw|>1e-3*lr: {(dw > 1e-3 * pc.onf_lr).sum()} of {dw.size} (max {dw.max():.2e})  "
              f"path_z err {pz:.2e}  path_gq err {pg:.2e}  traj_m err {tm:.2e}")
        eng.close(); eng2.close()

import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch
import _util as U
from oracle import checkers as OC
from nfopp_b200.collision_checker import CircleDirectedCollisionChecker, CircleCollisionChecker
from nfopp_b200.utils.position2 import Position2

fx = U.load("checkers.npz")
poses, car = fx["poses"], fx["obstacles_car"]
c = CircleDirectedCollisionChecker(0.3, (0, 3, 0, 3)); c.update_obstacle_points(car)
got = c.check_collision(Position2.from_vec(poses))
bad = np.nonzero(got != fx["circle_directed_car"])[0]
print("circle directed mismatches", len(bad), bad[:10], poses[bad[:5]])
got2 = c.check_collision(poses)   # raw [N,3]
print("raw [N,3] mismatches", int(np.sum(got2 != fx["circle_directed_car"])))
got3 = c.check_collision(poses[:, :2].copy())
print("raw [N,2] mismatches", int(np.sum(got3 != fx["circle_directed_car"])))

# quality: philox engine on car env
from _gpu_util import engine_from_golden
fxl = U.load("lockstep_car.npz")
obs = fxl["meta/obstacles"]
for mode in ["philox", "numpy"]:
    eng = engine_from_golden("car", fxl, "free/state0", rng_mode="philox" if mode == "philox" else "inject", seed=5)
    rs = np.random.RandomState(400)
    for k in range(400):
        if mode == "numpy":
            n = 99
            var = dict(t_sample=rs.rand(n), coarse_xy=rs.randn(n, 2), coarse_th=rs.randn(n), fine_xy=rs.randn(n, 2),
                       fine_th=rs.randn(n), choice_u=rs.rand(2000), field_xy=rs.rand(10, 2), field_th=rs.rand(10, 1),
                       t_path=rs.rand(n).astype(np.float32))
            eng.inject_variates(var)
        eng.step(1, graph=False)
        if k % 50 == 49:
            path = eng.paths()[0].cpu().numpy().astype(np.float64)
            col = OC.rectangle_labels(path, obs, (-0.3, 0.2, -0.3, 0.2), (0, 3, 0, 3))
            n_tr = int(eng["train_n"][0]); y = eng["train_y"][0, :n_tr].cpu().numpy()
            z = eng["train_z"][0, :n_tr].cpu().numpy()
            print(mode, k + 1, "path collisions", int(col.sum()), "train pos frac", float(y.mean()),
                  "bce", float(np.mean(np.maximum(z, 0) - z * y + np.log1p(np.exp(-np.abs(z))))), "status", int(eng["status"][0]),
                  "used", int(eng["choice_used"][0]))
    tp = eng["train_pos"][0, :n_tr].cpu().numpy()
    print(" coarse std", tp[:99].std(0), "buf std", tp[99:199].std(0), "field min/max", tp[199:].min(0), tp[199:].max(0))

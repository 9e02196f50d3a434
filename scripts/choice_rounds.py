import os, sys
ROOT = "/root/repo"
for p in (ROOT, os.path.join(ROOT, "pytorch-motion-planner_b200")):
    sys.path.insert(0, p)
import torch, bench as B
from nfopp_b200.batched import BatchedPlanner
cfg, env, starts, goals, _ = B.build_workload("corridor", 1024, 0)
pl = BatchedPlanner(cfg); pl.init(starts, goals, **env)
for k in (1, 20, 100, 300):
    pl.step(k); torch.cuda.synchronize()
    u = pl.engine["choice_used"].float()
    print("after", k, "more steps: uniforms consumed mean %.1f max %d min %d" % (u.mean().item(), int(u.max()), int(u.min())))

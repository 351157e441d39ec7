import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from tests.test_pointmass_gpu import _rollout
for prec in ("f64", "mixed"):
    t = time.time()
    wq, wv, re_, _, env = _rollout(prec, n=512, steps=200)
    print(prec, "worst rel q", wq, "v", wv, "rew", re_, "time", time.time() - t, flush=True)
import mujoco_sim_b200 as ms
for prec in ("f64", "mixed"):
    env = ms.make("point_mass_reach", 65536, precision=prec)
    env.reset()
    for _ in range(20): env.step(None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200): env.step_flat(None)
    e1.record(); torch.cuda.synchronize()
    ms_ = e0.elapsed_time(e1) / 200
    print(prec, "ms/step", ms_, "env-steps/s", 65536 / ms_ * 1e3, env.stats())

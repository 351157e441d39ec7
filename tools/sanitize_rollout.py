"""Short Planar Push rollout with autoreset, slot adoption and settle jobs active (episodes of 3 control steps),
meant to be run under `compute-sanitizer --tool memcheck|racecheck|synccheck` (logs kept under profiles/).
Usage: compute-sanitizer --tool racecheck python tools/sanitize_rollout.py [n_envs] [steps] [precision]"""
import sys

import torch

import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import RobotPushConfig

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
prec = sys.argv[3] if len(sys.argv) > 3 else "f64"
env = ms.make("robot_planar_push", n, seed=3, precision=prec, config=RobotPushConfig(max_control_steps_per_episode=3))
env.reset()
done = 0
for k in range(steps):
    _, _, term, trunc, _ = env.step(None)
    done += int((term | trunc).sum())
torch.cuda.synchronize()
st = env.stats()
print(f"sanitize_rollout: {n} envs x {steps} steps ({prec}), episodes finished {done}, stats {st}")
assert done > 0
env.close()

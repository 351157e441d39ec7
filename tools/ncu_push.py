import sys
sys.path.insert(0, ".")
import torch
import mujoco_sim_b200 as ms
task = sys.argv[1] if len(sys.argv) > 1 else "robot_planar_push"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
env = ms.make(task, n, precision="mixed")
env.reset()
for _ in range(6): env.step_flat(None)
torch.cuda.synchronize()
print("ok")

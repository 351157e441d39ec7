import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import RobotPushConfig
n = int(sys.argv[1]); steps = int(sys.argv[2]); maxlen = int(sys.argv[3])
env = ms.make("robot_planar_push", n, precision="mixed", config=RobotPushConfig(max_control_steps_per_episode=maxlen))
env.reset(); torch.cuda.synchronize()
time.sleep(0.2)
for _ in range(3): env.step_flat(None)
torch.cuda.synchronize()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
evs[0].record()
for i in range(steps):
    env.step_flat(None); evs[i + 1].record()
env.join_aux()
end = torch.cuda.Event(enable_timing=True); end.record()
torch.cuda.synchronize()
ts = np.array([evs[i].elapsed_time(evs[i + 1]) for i in range(steps)])
tot = evs[0].elapsed_time(end)
st = env.stats()
print("n", n, "maxlen", maxlen, "ms/step mean", ts.mean().round(3), "med", np.median(ts).round(3), "max", ts.max().round(3), "total incl. aux", round(tot / steps, 3),
      "episodes", st["episodes"], "mean len", st["length_sum"] / max(st["episodes"], 1), "errors", st["errors"], "launches", env.launch_count)
s = env.get_state(0, min(n, 512))
print("status hist", np.bincount([x.status for x in s], minlength=16)[:16].tolist(), "episode ids", np.bincount([x.episode_id for x in s])[:8].tolist())
z = np.array([[x.qpos[6 + 7 * k + 2] for k in range(3)] for x in s])
print("prop z range", z.min(), z.max())

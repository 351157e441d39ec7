"""Where does the gap between the per-launch median and the timed-region mean come from?"""
import sys, numpy as np, torch
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
from bench import _stagger_episodes
N = int(sys.argv[1]) if len(sys.argv) > 1 else 5328
prec = sys.argv[2] if len(sys.argv) > 2 else "f64"
K = 100
env = ms.make("robot_planar_push", N, precision=prec, seed=0)
env.reset()
_stagger_episodes(env, env.bundle.task.max_steps, 0)
dev = torch.device("cuda:0")
lo = torch.tensor([-0.2, -0.6], device=dev); hi = torch.tensor([0.2, -0.3], device=dev)
acts = (lo + (hi - lo) * torch.rand(16, N, 2, device=dev)).float().contiguous()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for i in range(int(sys.argv[3]) if len(sys.argv) > 3 else 10): env.step_flat(acts[i % 16])
def region(fl):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        if fl: flush.zero_()
        env.step_flat(acts[i % 16])
    env.join_aux(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K
print("no flush  ms/step", region(False)); print("flush     ms/step", region(True)); print("no flush  ms/step", region(False))
evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
evs[0].record()
for i in range(K):
    env.step_flat(acts[i % 16]); evs[i + 1].record()
torch.cuda.synchronize()
per = np.array([evs[i].elapsed_time(evs[i + 1]) for i in range(K)])
print("per-launch: mean %.3f median %.3f min %.3f max %.3f p90 %.3f" % (per.mean(), np.median(per), per.min(), per.max(), np.percentile(per, 90)))
print("sorted tail", np.sort(per)[-8:])
print(env.stats())

"""Is an env's own cost persistent from one control step to the next?  (non-lock-step run: warp duration = own cost)"""
import ctypes, os, sys, numpy as np, torch
os.environ["MJS_SYNC_LEVEL"] = "0"
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
from bench import _stagger_episodes
N = 5328
env = ms.make("robot_planar_push", N, precision="f64", seed=0)
env.reset(); _stagger_episodes(env, env.bundle.task.max_steps, 0)
dev = torch.device("cuda:0")
lo = torch.tensor([-0.2, -0.6], device=dev); hi = torch.tensor([0.2, -0.3], device=dev)
acts = (lo + (hi - lo) * torch.rand(16, N, 2, device=dev)).float().contiguous()
for i in range(600): env.step_flat(acts[i % 16])
buf = (ctypes.c_longlong * (16 * N))()
env._L.mjs_debug_times(env._h, buf, 16 * N)
D, NI = [], []
for rep in range(6):
    env.step_flat(acts[rep]); torch.cuda.synchronize()
    env._L.mjs_debug_times(env._h, buf, 16 * N)
    t = np.frombuffer(buf, dtype=np.int64).reshape(N, 16).copy()
    D.append((t[:, 1] - t[:, 0]) / 1e6); NI.append(t[:, 3] >> 32)
D = np.array(D); NI = np.array(NI)
print("own-cost per env-step (ms): mean %.3f p10 %.3f p50 %.3f p90 %.3f p99 %.3f max %.3f" % (D.mean(), np.percentile(D, 10), np.median(D), np.percentile(D, 90), np.percentile(D, 99), D.max()))
for lag in (1, 2, 5):
    print("corr(cost_k, cost_k+%d) = %.3f" % (lag, np.corrcoef(D[0], D[lag])[0, 1]), " corr(nit) = %.3f" % np.corrcoef(NI[0], NI[lag])[0, 1])
print("corr(cost, nit same step) %.3f" % np.corrcoef(D[1], NI[1])[0, 1])
# how well does sorting by the previous cost equalise groups of 6?
for name, key in (("unsorted", np.arange(N)), ("sorted by previous cost", np.argsort(-D[0])), ("oracle sort (same step)", np.argsort(-D[1]))):
    g = D[1][key][: N // 6 * 6].reshape(-1, 6)
    print(f"{name:26s}: mean of group max {g.max(1).mean():.3f}  vs mean cost {D[1].mean():.3f}  -> lock-step efficiency {D[1].mean()/g.max(1).mean():.2f}")

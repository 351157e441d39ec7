"""Cycles per substep stage (needs a -DMJS_STAGE_CLOCKS build selected with MJS_LIB).  MJS_SYNC_LEVEL=0 recommended:
without lock-step a stage's cycles are the warp's own latency, not its group's."""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
from bench import _stagger_episodes
N = int(sys.argv[1]) if len(sys.argv) > 1 else 21312
prec = sys.argv[2] if len(sys.argv) > 2 else "f64"
env = ms.make("robot_planar_push", N, precision=prec, seed=0)
env.reset(); _stagger_episodes(env, env.bundle.task.max_steps, 0)
dev = torch.device("cuda:0")
lo = torch.tensor([-0.2, -0.6], device=dev); hi = torch.tensor([0.2, -0.3], device=dev)
draw = lambda: (lo + (hi - lo) * torch.rand(N, 2, device=dev)).float()
for i in range(550): env.step_flat(draw())  # a new batch every step: stationary state mix of the bench workload
buf = (ctypes.c_longlong * (16 * N))()
env._L.mjs_debug_times(env._h, buf, 16 * N)
acc = np.zeros(12); tot = 0.0
for rep in range(4):
    env.step_flat(draw()); torch.cuda.synchronize()
    env._L.mjs_debug_times(env._h, buf, 16 * N)
    t = np.frombuffer(buf, dtype=np.int64).reshape(N, 16).copy()
    acc += t[:, 4:16].mean(0); tot += ((t[:, 1] - t[:, 0]) * 1.965).mean()   # ns -> cycles at 1965 MHz
acc /= 4 * 20; tot /= 4 * 20
names = ["kinematics", "dynamics", "collide", "make_rows", "solve: eval+grad+cost", "solve: Hessian build", "solve: direction (Cholesky)", "solve: line search",
         "solve: final forces", "touch+integrate", "solve: loop exit vote", "kinematics, 2nd run (I-cache hot)"]
print(f"{prec} N={N} sync={os.environ.get('MJS_SYNC_LEVEL','2')}: cycles per env-substep (mean over envs); control step total/20 = {tot:.0f}")
for n, a in zip(names, acc):
    if n != "-": print(f"  {n:30s} {a:9.0f}  {100 * a / acc.sum():5.1f}%")
print(f"  {'sum of stages':30s} {acc.sum():9.0f}")

"""Cold vs instruction-cache-hot cost of three substep stages (needs a -DMJS_STAGE_CLOCKS -DMJS_STAGE_CLOCKS_REPEAT build
selected with MJS_LIB): kinematics, dynamics and collision run twice back to back through non-inlined wrappers; the
second run finds its code in the instruction cache.  MJS_SYNC_LEVEL=0 (no lock-step) so that cycles are the warp's own."""
import ctypes, os, sys, numpy as np, torch
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
from bench import _stagger_episodes
N = int(sys.argv[1]) if len(sys.argv) > 1 else 19536
prec = sys.argv[2] if len(sys.argv) > 2 else "f64"
WORDS = 18
env = ms.make("robot_planar_push", N, precision=prec, seed=0)
env.reset(); _stagger_episodes(env, env.bundle.task.max_steps, 0)
dev = torch.device("cuda:0")
lo = torch.tensor([-0.2, -0.6], device=dev); hi = torch.tensor([0.2, -0.3], device=dev)
acts = (lo + (hi - lo) * torch.rand(16, N, 2, device=dev)).float().contiguous()
for i in range(300): env.step_flat(acts[i % 16])
buf = (ctypes.c_longlong * (WORDS * N))()
env._L.mjs_debug_times(env._h, buf, WORDS * N)
acc = np.zeros(14); tot = 0.0
for rep in range(4):
    env.step_flat(acts[rep]); torch.cuda.synchronize()
    env._L.mjs_debug_times(env._h, buf, WORDS * N)
    t = np.frombuffer(buf, dtype=np.int64).reshape(N, WORDS).copy()
    acc += t[:, 4:18].mean(0); tot += ((t[:, 1] - t[:, 0]) * 1.965).mean()
acc /= 4 * 20; tot /= 4 * 20
print(f"{prec} N={N} sync={os.environ.get('MJS_SYNC_LEVEL','2')}: cycles per env-substep; control step total/20 = {tot:.0f}")
for name, a, b in (("kinematics", 0, 11), ("dynamics", 1, 12), ("collide", 2, 13)):
    print(f"  {name:12s} first run {acc[a]:8.0f}   second run (code hot) {acc[b]:8.0f}   ratio {acc[b] / acc[a]:.2f}")
for name, k in (("make_rows", 3), ("eval+grad+cost", 4), ("Hessian", 5), ("direction", 6), ("line search", 7), ("final forces", 8), ("touch+integrate", 9)):
    print(f"  {name:16s} {acc[k]:8.0f}")

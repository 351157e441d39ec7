import sys, os; sys.path.insert(0, ".")
import numpy as np, torch, ctypes
import mujoco_sim_b200 as ms
from oracle.oracle import OracleEnvs
n = 2048
env = ms.make("robot_planar_push", n, seed=0)
env.reset()
st = env.get_state()
rng = np.random.RandomState(0)
for s, p in zip(st, rng.randint(0, 500, n)): s.episode_step = int(p)
env.set_state(st, forward=False)
for k in range(560): env.step_flat(None)
torch.cuda.synchronize()
st = env.get_state()
q = np.array([list(s.qpos)[:27] for s in st])
tilt = 1 - np.abs(q[:, [9, 16, 23]])
z = q[:, [8, 15, 22]]
print("props tilted > 1e-3:", (tilt > 1e-3).mean(), "lying (|qw| < 0.8):", (np.abs(q[:, [9, 16, 23]]) < 0.8).mean(), "z>0.03:", (z > 0.03).mean(), "z<0.01", (z < 0.01).mean(), flush=True)
xy = q[:, [6, 7, 13, 14, 20, 21]].reshape(n, 3, 2)
print("props outside the workspace box:", ((np.abs(xy[..., 0]) > 0.22) | (xy[..., 1] > -0.28) | (xy[..., 1] < -0.62)).mean())
ora = OracleEnvs(env.bundle, n); ctypes.memmove(ora.states, st, ctypes.sizeof(st)); ora.tol_scale = 1.0
os.environ["ORACLE_CVX_LOG"] = "1"
a = np.stack([rng.uniform(-0.2, 0.2, n), rng.uniform(-0.6, -0.3, n)], 1)
sys.stderr.flush()
ora.step(a)
print("oracle one step on the GPU's stationary states done: env-substeps", n * 20, flush=True)

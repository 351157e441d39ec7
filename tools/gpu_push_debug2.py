import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
np.set_printoptions(precision=8, suppress=True, linewidth=220)
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
n = 2
bundle = build_planar_push()
bundle.task.n_sub = int(sys.argv[1]) if len(sys.argv) > 1 else 1
ms.envs.register_builder("_push_dbg", lambda: bundle)
ora = push_states(bundle, n, 0)
env = ms.BatchedMujocoSimEnv("_push_dbg", n, precision="f64", autoreset=False)
env.set_state(copy_states(ora.states, n), forward=False)
a = np.array([[0.05, -0.45], [-0.1, -0.5]], np.float32)
ora.step(a.astype(np.float64))
env.step(torch.from_numpy(a).cuda())
st = env.get_state()
for e in range(n):
    s, o = st[e], ora.states[e]
    for name, w in (("traj_q0", 6), ("traj_q1", 6), ("ctrl", 6), ("qpos", 27), ("qvel", 24), ("qacc_warm", 24)):
        g = np.array(list(getattr(s, name))[:w]); r = np.array(list(getattr(o, name))[:w])
        print(e, name, "maxerr", np.abs(g - r).max())
        if np.abs(g - r).max() > 1e-9: print("   gpu", g, "\n   ora", r)
    print(e, "time", s.time, o.time, "t0", s.traj_t0, o.traj_t0, "traj", s.traj_active, o.traj_active, "step", s.episode_step, o.episode_step)

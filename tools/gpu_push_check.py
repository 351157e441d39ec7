import sys, time, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states, rel_err

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
bundle = build_planar_push()
ora = push_states(bundle, n, 0)
for prec in ("f64", "mixed"):
    import copy
    o2 = push_states(bundle, n, 0)
    env = ms.BatchedMujocoSimEnv("robot_planar_push", n, precision=prec, autoreset=False)
    env.set_state(copy_states(o2.states, n), forward=False)
    rng = np.random.RandomState(1)
    for k in range(steps):
        # push towards the first object from the current tcp: makes contacts happen
        obs_prev = o2.collect()[0] if k else None
        a = np.stack([rng.uniform(-0.2, 0.2, n), rng.uniform(-0.6, -0.3, n)], 1).astype(np.float32)
        if k % 2 == 0:
            qp = o2.field("qpos", 27)
            a = (qp[:, 6:8] + rng.uniform(-0.01, 0.01, (n, 2))).astype(np.float32)
        t0 = time.time()
        obs_o, rew_o, _, fl_o = o2.step(a.astype(np.float64))
        t1 = time.time()
        obs, rew, term, trunc, info = env.step(torch.from_numpy(a).cuda())
        st = env.get_state()
        q = np.array([list(s.qpos)[:27] for s in st]); v = np.array([list(s.qvel)[:24] for s in st])
        qo, vo = o2.field("qpos", 27), o2.field("qvel", 24)
        status = np.array([s.status for s in st]); so = o2.field("status")
        eq = np.abs(q - qo).max(); ev = np.abs(v - vo).max()
        print(prec, k, "max|dq|", eq, "max|dv|", ev, "obs", np.abs(env._obs.cpu().numpy() - obs_o).max(), "rew", np.abs(rew.cpu().numpy() - rew_o).max(),
              "status gpu", np.bincount(status, minlength=2)[:4], "oracle", np.unique(so), "oracle s/step", (t1 - t0) / n, flush=True)
        if not np.isfinite(eq): break

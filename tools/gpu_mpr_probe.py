import sys; sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import copy_states, push_states
n, b = 64, build_planar_push()
ora = push_states(b, n, 5)
t = b.task
for e in range(n):
    s = ora.states[e]
    a0, a1, a2 = t.obj_qadr[0], t.obj_qadr[1], t.obj_qadr[2]
    if e % 2:
        s.qpos[a1] = s.qpos[a0] + 0.002 + 0.0001 * e; s.qpos[a1 + 1] = s.qpos[a0 + 1]; s.qpos[a1 + 2] = s.qpos[a0 + 2] + 2 * 0.019 - 1e-4
    else:
        s.qpos[a2] = s.qpos[a0]; s.qpos[a2 + 1] = s.qpos[a0 + 1] + 0.02 + 0.019 - 3e-4; s.qpos[a2 + 2] = 0.02 - 1e-4
        s.qpos[a2 + 3], s.qpos[a2 + 4], s.qpos[a2 + 5], s.qpos[a2 + 6] = np.sqrt(0.5), np.sqrt(0.5), 0.0, 0.0
for ns in (1, 2, 5, 20):
    bb = build_planar_push(); bb.task.n_sub = ns
    ms.envs.register_builder("_p", lambda bb=bb: bb)
    env = ms.BatchedMujocoSimEnv("_p", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
    from oracle.oracle import OracleEnvs
    import ctypes
    oo = OracleEnvs(bb, n); ctypes.memmove(oo.states, ora.states, ctypes.sizeof(ora.states))
    env.set_state(copy_states(oo.states, n), forward=False)
    a = np.array([[0.18, -0.58]] * n, np.float32)
    oo.step(a.astype(np.float64)); env.step(torch.from_numpy(a).cuda())
    q = np.array([list(s.qpos)[:27] for s in env.get_state()]); d = np.abs(q - oo.field("qpos", 27)).max(1)
    v = np.array([list(s.qvel)[:24] for s in env.get_state()]); dv = np.abs(v - oo.field("qvel", 24)).max(1)
    print("n_sub", ns, "stacked(odd) max dq %.2e dv %.2e | lying(even) max dq %.2e dv %.2e" % (d[1::2].max(), dv[1::2].max(), d[::2].max(), dv[::2].max()), flush=True)
    env.close()

import sys, os; sys.path.insert(0, ".")
import numpy as np, torch, ctypes
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
from tests.test_arm_gpu import _push_actions, _qv, _env
from oracle.oracle import OracleEnvs
np.set_printoptions(precision=6, linewidth=220)
n = 128; b = build_planar_push()
ora = push_states(b, n, 0)
rng = np.random.RandomState(1)
pre = copy_states(ora.states, n)
a = _push_actions(ora, rng, n, 0)
e = 19
one = copy_states((type(pre[0]) * 1).from_address(ctypes.addressof(pre) + int(e) * ctypes.sizeof(pre[0])), 1)
def dump(env):
    out = (ctypes.c_double * 1024)()
    ms.lib.check(env._L.mjs_debug_forward(env._h, 0, out, 1024))
    o = np.array(out); p = 36 + 24
    a_g = o[p:p+24]; p += 48
    return a_g, int(o[p]), o[p+1]
# oracle reference for solve #15 = warm after m=15
bm = build_planar_push(); bm.task.n_sub = 15
om = OracleEnvs(bm, 1); ctypes.memmove(om.states, one, ctypes.sizeof(one)); om.step(a[e:e+1].astype(np.float64))
ref = np.array(list(om.states[0].qacc_warm)[:24])
for m, prek in ((14, 0), (13, 1), (12, 2), (10, 4), (0, 14)):
    bm = build_planar_push(); bm.task.n_sub = max(m, 1)
    ms.envs.register_builder("_dbgm", lambda: bm)
    env = ms.BatchedMujocoSimEnv("_dbgm", 1, precision="f64", autoreset=False)
    env.set_state(copy_states(one, 1), forward=False)
    if m > 0: env.step(torch.from_numpy(a[e:e+1]).cuda())
    else: continue
    # prek in-kernel substeps, then before_substep + forward
    os.environ["MJS_DEBUG_PRE"] = str(prek)
    if prek == 0:
        # emulate before_substep on the host
        st = env.get_state(); s = st[0]
        q0 = np.array(list(s.traj_q0)); q1 = np.array(list(s.traj_q1)); fr = min(max((s.time - s.traj_t0) / 0.1, 0.0), 1.0)
        c = q0 + (q1 - q0) * fr
        for k in range(6): s.ctrl[k] = c[k]
        env.set_state(st, forward=False)
    qa, ncon, nit = dump(env)
    print("m", m, "pre", prek, "ncon", ncon, "niter", nit, "max|qacc - oracle| %.3e" % np.abs(qa - ref).max(), qa[12:18])
    env.close()
print("oracle", ref[12:18])

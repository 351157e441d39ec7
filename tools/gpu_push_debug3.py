import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
np.set_printoptions(precision=8, suppress=True, linewidth=220)
import mujoco_sim_b200 as ms
from mujoco_sim_b200 import lib as L
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
n = 8
bundle = build_planar_push()
ora = push_states(bundle, n, 0)
env = ms.BatchedMujocoSimEnv("robot_planar_push", n, precision="f64", autoreset=False)
env.set_state(copy_states(ora.states, n), forward=False)
NV = 24
for k in range(6):
    qp = ora.field("qpos", 27)
    a = qp[:, 6:8].astype(np.float32)
    # before stepping: compare forward stages at the current (shared) oracle state
    env.set_state(copy_states(ora.states, n), forward=False)
    for e in range(n):
        out = (ctypes.c_double * 1024)()
        L.check(env._L.mjs_debug_forward(env._h, e, out, 1024))
        o = np.array(out); d = ora.stage_dump(e)
        qacc = o[36 + NV:36 + 2 * NV]; ncon = int(o[36 + 3 * NV]); niter = int(o[36 + 3 * NV + 1])
        craw = o[36 + 3 * NV + 3:36 + 3 * NV + 3 + 112].reshape(16, 7)
        err = np.abs(qacc - d["qacc"]).max()
        flag = "" if err < 1e-8 else "   <<<<<"
        print(f"step {k} env {e}: ncon {ncon} vs {d['ncon']} niter {niter} vs {d['niter']} qacc err {err:.3e}{flag}")
        if err > 1e-8:
            print("   dist gpu", craw[:ncon, 0]); print("   dist ora", d["con_dist"])
            print("   pos gpu", craw[:ncon, 1:4].ravel()); print("   pos ora", d["con_pos"].ravel())
            print("   qacc gpu", qacc); print("   qacc ora", d["qacc"])
    ora.step(a.astype(np.float64))

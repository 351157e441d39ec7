import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
np.set_printoptions(precision=6, suppress=True, linewidth=200)
import mujoco_sim_b200 as ms
from mujoco_sim_b200 import lib as L
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
n = 4
bundle = build_planar_push()
ora = push_states(bundle, n, 0)
# give some velocity so bias terms matter
rng = np.random.RandomState(3)
for e in range(n):
    for i in range(24): ora.states[e].qvel[i] = rng.uniform(-0.3, 0.3) if e % 2 else 0.0
    for i in range(6): ora.states[e].ctrl[i] += rng.uniform(-0.05, 0.05)
env = ms.BatchedMujocoSimEnv("robot_planar_push", n, precision=sys.argv[1] if len(sys.argv) > 1 else "f64", autoreset=False)
env.set_state(copy_states(ora.states, n), forward=False)
NV = 24
for e in range(n):
    out = (ctypes.c_double * 1024)()
    L.check(env._L.mjs_debug_forward(env._h, e, out, 1024))
    o = np.array(out)
    d = ora.stage_dump(e)
    p = 0
    Marm = o[p:p+36].reshape(6, 6); p += 36
    fsm = o[p:p+NV]; p += NV
    qacc = o[p:p+NV]; p += NV
    fcon = o[p:p+NV]; p += NV
    ncon, niter, touch = o[p:p+3]; p += 3
    craw = o[p:p+16*7].reshape(16, 7); p += 16*7
    xpos = o[p:p+18].reshape(6, 3); p += 18
    fs_o = d["qfrc_passive"] - d["qfrc_bias"]
    # oracle qfrc_smooth includes actuator: recompute via M*qacc_smooth
    fs_o = d["M"] @ d["qacc_smooth"]
    print(f"env {e}: M err", np.abs(Marm - d["M"][:6, :6]).max(), "xpos err", np.abs(xpos - d["xpos"][2:8]).max(),
          "fsm err", np.abs(fsm - fs_o).max(), "ncon", ncon, d["ncon"], "niter", niter, d["niter"], "qacc err", np.abs(qacc - d["qacc"]).max())
    if np.abs(fsm - fs_o).max() > 1e-6: print("  fsm gpu", fsm, "\n  fsm ora", fs_o)
    nc = int(ncon)
    kept = [i for i in range(d["ncon"])]
    print("  dist gpu", craw[:nc, 0], "\n  dist ora", d["con_dist"])
    if np.abs(qacc - d["qacc"]).max() > 1e-6: print("  qacc gpu", qacc, "\n  qacc ora", d["qacc"])

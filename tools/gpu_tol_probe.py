"""Probe: CUDA at the default solver tolerance 1e-8 against the oracle at 1e-8 and at 1e-12, single steps from identical states."""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import copy_states, push_states

n, steps, b = 1024, 12, build_planar_push()
ora = push_states(b, n, 41)
ora8 = push_states(b, n, 41)
env = ms.make("robot_planar_push", n, precision="f64", autoreset=False)
env12 = ms.make("robot_planar_push", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
rng = np.random.RandomState(42)
for k in range(steps):
    import ctypes
    ctypes.memmove(ora8.states, ora.states, ctypes.sizeof(ora.states))
    ora8.tol_scale = 1.0
    env.set_state(copy_states(ora.states, n), forward=False)
    env12.set_state(copy_states(ora.states, n), forward=False)
    qp = ora.field("qpos", 27)
    prop = 6 + 7 * (k % 3)
    a = (qp[:, prop:prop + 2] + rng.uniform(-0.01, 0.01, (n, 2))).astype(np.float32)
    ora.step(a.astype(np.float64)); ora8.step(a.astype(np.float64))
    env.step(torch.from_numpy(a).cuda()); env12.step(torch.from_numpy(a).cuda())
    q = np.array([list(s.qpos)[:27] for s in env.get_state()])
    q12 = np.array([list(s.qpos)[:27] for s in env12.get_state()])
    qo, qo8 = ora.field("qpos", 27), ora8.field("qpos", 27)
    d_12 = np.abs(q - qo).max(1); d_8 = np.abs(q - qo8).max(1); d_oo = np.abs(qo - qo8).max(1); d_1212 = np.abs(q12 - qo).max(1)
    tilt = 1 - np.abs(qo[:, [9, 16, 23]]).min(1)
    bad = np.nonzero(d_12 > 1e-6)[0]
    print(k, "cuda8-ora12 max %.2e n>1e-4 %d n>1e-6 %d | cuda8-ora8 max %.2e n>1e-6 %d | ora8-ora12 max %.2e n>1e-6 %d | cuda12-ora12 max %.2e" % (
        d_12.max(), (d_12 > 1e-4).sum(), (d_12 > 1e-6).sum(), d_8.max(), (d_8 > 1e-6).sum(), d_oo.max(), (d_oo > 1e-6).sum(), d_1212.max()),
        "tilt of outliers", np.round(tilt[bad][:6], 4), flush=True)

"""Probe 2: look at the single-step outliers of gpu_tol_probe.py (tipped props): status bits, which body differs, contact counts."""
import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import copy_states, push_states

n, steps, b = 1024, 12, build_planar_push()
ora = push_states(b, n, 41)
env = ms.make("robot_planar_push", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
rng = np.random.RandomState(42)
for k in range(steps):
    pre = copy_states(ora.states, n)
    env.set_state(copy_states(ora.states, n), forward=False)
    qp = ora.field("qpos", 27)
    prop = 6 + 7 * (k % 3)
    a = (qp[:, prop:prop + 2] + rng.uniform(-0.01, 0.01, (n, 2))).astype(np.float32)
    ora.step(a.astype(np.float64))
    env.step(torch.from_numpy(a).cuda())
    st = env.get_state()
    q = np.array([list(s.qpos)[:27] for s in st]); qo = ora.field("qpos", 27)
    d = np.abs(q - qo)
    for e in np.nonzero(d.max(1) > 1e-6)[0]:
        print("step", k, "env", e, "max diff", d[e].max(), "at index", d[e].argmax(), "status gpu", st[e].status, "oracle", ora.states[e].status)
        print("  pre qpos props", np.round(np.array(list(pre[e].qpos)[6:27]).reshape(3, 7), 4))
        print("  pre qvel props", np.round(np.array(list(pre[e].qvel)[6:24]).reshape(3, 6), 3))
        print("  post oracle   ", np.round(qo[e, 6:].reshape(3, 7), 4))
        print("  post gpu      ", np.round(q[e, 6:].reshape(3, 7), 4))
        # replay this env alone, substep by substep, on both sides: first divergence
        from oracle.oracle import OracleEnvs, lib
        o1 = OracleEnvs(b, 1); ctypes.memmove(o1.states, ctypes.byref(pre[e]), ctypes.sizeof(pre[e]))
        e1 = ms.make("robot_planar_push", 1, precision="f64", autoreset=False, solver_tolerance=1e-12)
        # one full control step with n_sub split: run the control step with the action, but compare after each extra substep of the NEXT step window
        for ns in (1, 2, 4, 8, 12, 16, 20):
            bb = build_planar_push(); bb.task.n_sub = ns
            ms.envs.register_builder("_probe", lambda bb=bb: bb)
            ee = ms.BatchedMujocoSimEnv("_probe", 1, precision="f64", autoreset=False, solver_tolerance=1e-12)
            oo = OracleEnvs(bb, 1); ctypes.memmove(oo.states, ctypes.byref(pre[e]), ctypes.sizeof(pre[e]))
            ee.set_state(copy_states(oo.states, 1), forward=False)
            oo.step(a[e:e+1].astype(np.float64)); ee.step(torch.from_numpy(a[e:e+1]).cuda())
            qq = np.array(list(ee.get_state()[0].qpos)[:27]); dd = np.abs(qq - oo.field("qpos", 27)[0])
            dmp = oo.stage_dump(0)
            print("   n_sub", ns, "diff", dd.max(), "ncon(oracle, after)", dmp["ncon"], "nefc", dmp["nefc"], "status", ee.get_state()[0].status, oo.states[0].status)
            ee.close()
        e1.close()

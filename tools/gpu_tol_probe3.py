"""Probe 3: find the substep at which a single-step outlier departs and compare the forward pass of both sides on the SAME state there."""
import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200 import lib as mlib
from mujoco_sim_b200.model.tasks import build_planar_push
from oracle.oracle import OracleEnvs
from tests.helpers import copy_states, push_states

n, steps, b = 1024, 12, build_planar_push()
ora = push_states(b, n, 41)
env = ms.make("robot_planar_push", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
rng = np.random.RandomState(42)
found = 0
for k in range(steps):
    pre = copy_states(ora.states, n)
    env.set_state(copy_states(ora.states, n), forward=False)
    qp = ora.field("qpos", 27)
    prop = 6 + 7 * (k % 3)
    a = (qp[:, prop:prop + 2] + rng.uniform(-0.01, 0.01, (n, 2))).astype(np.float32)
    ora.step(a.astype(np.float64))
    env.step(torch.from_numpy(a).cuda())
    q = np.array([list(s.qpos)[:27] for s in env.get_state()]); qo = ora.field("qpos", 27)
    for e in np.nonzero(np.abs(q - qo).max(1) > 1e-6)[0]:
        found += 1
        print("=== step", k, "env", e, "diff", np.abs(q[e] - qo[e]).max(), flush=True)
        prev = None
        for ns in range(1, 21):
            bb = build_planar_push(); bb.task.n_sub = ns
            ms.envs.register_builder("_probe", lambda bb=bb: bb)
            ee = ms.BatchedMujocoSimEnv("_probe", 1, precision="f64", autoreset=False, solver_tolerance=1e-12)
            oo = OracleEnvs(bb, 1); ctypes.memmove(oo.states, ctypes.byref(pre[e]), ctypes.sizeof(pre[e]))
            ee.set_state(copy_states(oo.states, 1), forward=False)
            oo.step(a[e:e+1].astype(np.float64)); ee.step(torch.from_numpy(a[e:e+1]).cuda())
            qq = np.array(list(ee.get_state()[0].qpos)[:27]); dd = np.abs(qq - oo.field("qpos", 27)[0]).max()
            vv = np.array(list(ee.get_state()[0].qvel)[:24]); dv = np.abs(vv - oo.field("qvel", 24)[0]).max()
            if dv > 1e-6 and prev is not None:
                # prev = oracle state after ns-1 substeps: forward pass of both sides on it
                po, pe = prev
                pe.set_state(copy_states(po.states, 1), forward=False)
                out = (ctypes.c_double * 1024)()
                mlib.check(pe._L.mjs_debug_forward(pe._h, 0, out, 1024))
                o = np.array(out[:])
                d = po.stage_dump(0)
                nv = 24
                qacc_g = o[36 + nv:36 + 2 * nv]; ncon_g = int(o[36 + 3 * nv]); nit_g = int(o[36 + 3 * nv + 1])
                cons = o[36 + 3 * nv + 3:36 + 3 * nv + 3 + 16 * 7].reshape(16, 7)
                print("  first bad substep", ns, "dq", dd, "dv", dv)
                print("  oracle: ncon", d["ncon"], "nefc", d["nefc"], "niter", d["niter"], "| kernel: ncon(rows incl limits)", ncon_g, "niter", nit_g)
                print("  qacc diff max", np.abs(qacc_g - d["qacc"]).max(), "at", np.abs(qacc_g - d["qacc"]).argmax(), "| oracle qacc prop part", np.round(d["qacc"][6:], 3))
                print("  kernel qacc prop part", np.round(qacc_g[6:], 3))
                print("  oracle con dist", np.array2string(d["con_dist"], precision=3))
                print("  kernel con dist", np.array2string(cons[:ncon_g, 0], precision=3))
                print("  oracle con normal z", np.round(d["con_frame"][:, 2], 3), "kernel", np.round(cons[:ncon_g, 6], 3))
                print("  prop quats", np.round(np.array(list(po.states[0].qpos)[6:27]).reshape(3, 7)[:, 3:], 6))
                print("  prop vel", np.round(np.array(list(po.states[0].qvel)[6:24]).reshape(3, 6), 4))
                break
            if prev is not None: prev[1].close()
            prev = (oo, ee)
        if found >= 3: sys.exit(0)

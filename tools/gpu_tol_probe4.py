"""Probe 4: is the outlier an implementation difference or sensitivity of the map?  K13/O13 = kernel / oracle state before the bad
substep.  A = oracle(K13), B = kernel(K13) = K14, C = oracle(O13) = O14."""
import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from oracle.oracle import OracleEnvs, lib as olib
from tests.helpers import copy_states, push_states

n, steps, b = 1024, 12, build_planar_push()
ora = push_states(b, n, 41)
env = ms.make("robot_planar_push", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
rng = np.random.RandomState(42)
def vec(s): return np.array(list(s.qpos)[:27] + list(s.qvel)[:24])
found = 0
for k in range(steps):
    pre = copy_states(ora.states, n)
    env.set_state(copy_states(ora.states, n), forward=False)
    qp = ora.field("qpos", 27)
    prop = 6 + 7 * (k % 3)
    a = (qp[:, prop:prop + 2] + rng.uniform(-0.01, 0.01, (n, 2))).astype(np.float32)
    ora.step(a.astype(np.float64)); env.step(torch.from_numpy(a).cuda())
    q = np.array([list(s.qpos)[:27] for s in env.get_state()]); qo = ora.field("qpos", 27)
    for e in np.nonzero(np.abs(q - qo).max(1) > 1e-6)[0]:
        found += 1
        print("=== step", k, "env", e, flush=True)
        hist = {}
        for ns in range(1, 21):
            bb = build_planar_push(); bb.task.n_sub = ns
            ms.envs.register_builder("_probe", lambda bb=bb: bb)
            ee = ms.BatchedMujocoSimEnv("_probe", 1, precision="f64", autoreset=False, solver_tolerance=1e-12)
            oo = OracleEnvs(bb, 1); ctypes.memmove(oo.states, ctypes.byref(pre[e]), ctypes.sizeof(pre[e]))
            ee.set_state(copy_states(oo.states, 1), forward=False)
            oo.step(a[e:e+1].astype(np.float64)); ee.step(torch.from_numpy(a[e:e+1]).cuda())
            ks = ee.get_state()
            hist[ns] = (copy_states(ks, 1), copy_states(oo.states, 1))
            d = np.abs(vec(ks[0]) - vec(oo.states[0]))
            print("  ns %2d  |K-O| q %.2e v %.2e" % (ns, d[:27].max(), d[27:].max()), end="")
            if ns > 1:
                K13, O13 = hist[ns - 1]
                A = OracleEnvs(b, 1); ctypes.memmove(A.states, K13, ctypes.sizeof(K13))
                olib().oracle_substeps(ctypes.byref(b.model), ctypes.byref(b.task), A.states, 1, 1, ctypes.c_double(1e-4))
                dA = np.abs(vec(A.states[0]) - vec(ks[0])); dC = np.abs(vec(A.states[0]) - vec(oo.states[0]))
                print("   oracle(K)-kernel(K) v %.2e   oracle(K)-oracle(O) v %.2e" % (dA[27:].max(), dC[27:].max()), end="")
            print(flush=True)
            ee.close()
        if found >= 2: sys.exit(0)

"""Where does precision="mixed" leave the oracle on struck-prop steps?  For the worst envs of a single control step: replay with
n_sub = 1..20 on both sides, apply the ORACLE's one-substep map to the KERNEL's state before each substep (A = oracle(K_{s-1})) and
compare with the kernel's own result K_s: |A - K_s| is the error the mixed substep map makes on identical input in that one substep.
At the worst substep, dump the forward stages of both sides from the kernel's pre-state (contacts, qacc).
usage (GPU box): python tools/gpu_mixed_probe.py > gpurun_out/mixed_probe.txt"""
import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from oracle.oracle import OracleEnvs, lib as olib
from tests.helpers import copy_states, push_states

n, b = 512, build_planar_push()
ora = push_states(b, n, 41)
env = ms.make("robot_planar_push", n, precision="mixed", autoreset=False, solver_tolerance=1e-12)
rng = np.random.RandomState(42)
vec = lambda s: np.array(list(s.qpos)[:27] + list(s.qvel)[:24])
NV = 24
done = 0
for k in range(6):
    pre = copy_states(ora.states, n)
    env.set_state(copy_states(ora.states, n), forward=False)
    qp = ora.field("qpos", 27)
    prop = 6 + 7 * (k % 3)
    a = (qp[:, prop:prop + 2] + rng.uniform(-0.01, 0.01, (n, 2))).astype(np.float32)
    ora.step(a.astype(np.float64)); env.step(torch.from_numpy(a).cuda())
    q = np.array([list(s.qpos)[:27] for s in env.get_state()]); qo = ora.field("qpos", 27)
    err = np.abs(q - qo).max(1)
    print(f"step {k}: envs over 1e-4: {(err > 1e-4).sum()} of {n}; percentiles 50/90/99 {np.percentile(err, 50):.1e} {np.percentile(err, 90):.1e} {np.percentile(err, 99):.1e}", flush=True)
    for e in np.argsort(-err)[:3]:
        if err[e] < 1e-4 or done >= 6: continue
        done += 1
        print(f"=== step {k} env {e} final |dq| {err[e]:.2e}", flush=True)
        hist = {}; worst = (0, 0.0)
        for ns in range(1, 21):
            bb = build_planar_push(); bb.task.n_sub = ns
            ms.envs.register_builder("_probe", lambda bb=bb: bb)
            ee = ms.BatchedMujocoSimEnv("_probe", 1, precision="mixed", autoreset=False, solver_tolerance=1e-12)
            ee.set_state(copy_states((type(pre[0]) * 1).from_buffer_copy(bytes(pre[e])), 1), forward=False)
            ee.step(torch.from_numpy(a[e:e + 1]).cuda())
            ks = copy_states(ee.get_state(), 1); hist[ns] = ks
            line = f"  ns {ns:2d}"
            if ns > 1:
                A = OracleEnvs(b, 1); ctypes.memmove(A.states, hist[ns - 1], ctypes.sizeof(hist[ns - 1]))
                olib().oracle_substeps(ctypes.byref(b.model), ctypes.byref(b.task), A.states, 1, 1, ctypes.c_double(1e-12))
                dA = np.abs(vec(A.states[0]) - vec(ks[0]))
                line += f"   |oracle(K_prev) - K| q {dA[:27].max():.2e} v {dA[27:].max():.2e} (dof {int(dA[27:].argmax())})"
                if dA[27:].max() > worst[1]: worst = (ns, dA[27:].max())
            print(line, flush=True)
            ee.close()
        ns = worst[0]
        if ns > 1:
            # stage dumps from the kernel's state before the worst substep (trajectory / ctrl are inside the state)
            K = hist[ns - 1]
            oo = OracleEnvs(b, 1); ctypes.memmove(oo.states, K, ctypes.sizeof(K))
            od = oo.stage_dump(0)
            bb = build_planar_push()
            ee = ms.make("robot_planar_push", 1, precision="mixed", autoreset=False, solver_tolerance=1e-12)
            ee.set_state(copy_states(K, 1), forward=False)
            buf = (ctypes.c_double * 512)()
            ee._L.mjs_debug_forward(ee._h, 0, buf, 512)
            d = np.frombuffer(buf, dtype=np.float64).copy()
            qacc_k = d[36 + NV:36 + 2 * NV]; ncon_k = int(d[36 + 3 * NV]); con = d[36 + 3 * NV + 1:36 + 3 * NV + 1 + 7 * ncon_k].reshape(ncon_k, 7)
            print(f"  worst substep {ns}: contacts oracle {od['ncon']} kernel {ncon_k}; |qacc diff| max {np.abs(od['qacc'][:NV] - qacc_k).max():.2e} (|qacc| max {np.abs(od['qacc']).max():.2e})")
            print("   oracle dist", np.array2string(od["con_dist"], precision=3, max_line_width=200))
            print("   qacc oracle", np.array2string(od["qacc"][:NV], precision=3, max_line_width=250))
            print("   qacc kernel", np.array2string(qacc_k, precision=3, max_line_width=250))
            ee.close()

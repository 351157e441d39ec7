"""Where does precision="mixed" lose accuracy?  Stage dump of an impact state: CUDA (mixed, f64) vs oracle."""
import sys; sys.path.insert(0, ".")
import numpy as np, torch, ctypes
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
from tests.test_arm_gpu import _push_actions, _qv, _env
from oracle.oracle import OracleEnvs
np.set_printoptions(precision=4, linewidth=220)
n = 128; b = build_planar_push()
ora = push_states(b, n, 0)
rng = np.random.RandomState(1)
pre = copy_states(ora.states, n)
a = _push_actions(ora, rng, n, 0)
for e, m in ((19, 14), (33, 16), (19, 16)):
    one = copy_states((type(pre[0]) * 1).from_address(ctypes.addressof(pre) + int(e) * ctypes.sizeof(pre[0])), 1)
    bm = build_planar_push(); bm.task.n_sub = m
    om = OracleEnvs(bm, 1); ctypes.memmove(om.states, one, ctypes.sizeof(one))
    om.step(a[e:e+1].astype(np.float64))
    d = om.stage_dump(0)
    for prec in ("f64", "mixed"):
        env = _env("robot_planar_push", 1, prec)
        env.set_state(copy_states(om.states, 1), forward=False)
        out = (ctypes.c_double * 1024)()
        ms.lib.check(env._L.mjs_debug_forward(env._h, 0, out, 1024))
        o = np.array(out); p = 0
        Marm = o[:36].reshape(6, 6); p = 36
        fsm = o[p:p+24]; p += 24; a_g = o[p:p+24]; p += 24; fcon = o[p:p+24]; p += 24
        ncon = int(o[p]); niter = o[p+1]; p += 3
        craw = o[p:p+16*7].reshape(16, 7); p += 16*7
        xpos = o[p:p+18].reshape(6, 3); p += 18 + 36
        cf = o[p:p+64].reshape(16, 4)
        fo = d["M"] @ d["qacc_smooth"]
        rel = lambda x, y: np.abs(x - y).max() / max(np.abs(y).max(), 1e-30)
        print(f"env {e} m {m} {prec:5s} ncon {ncon}/{d['ncon']} niter {niter}/{d['niter']} | rel err: Marm {rel(Marm, d['M'][:6,:6]):.1e} fsm {rel(fsm, fo):.1e} "
              f"link xpos abs {np.abs(xpos - d['xpos'][2:8]).max():.1e} | contact dist abs {np.abs(craw[:ncon,0]-d['con_dist']).max():.1e} (rel to dist {np.abs((craw[:ncon,0]-d['con_dist'])/d['con_dist']).max():.1e}) "
              f"pos abs {np.abs(craw[:ncon,1:4]-d['con_pos']).max():.1e} | qacc rel {rel(a_g, d['qacc']):.1e} abs {np.abs(a_g-d['qacc']).max():.2e} | fcon-equivalent cf[:,0] max {cf[:ncon,0].max():.3f}")
        env.close()

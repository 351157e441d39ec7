"""Diagnostic: single-control-step |dq| of the CUDA path vs the oracle on Planar Push, bucketed by prop tilt."""
import sys; sys.path.insert(0, ".")
import numpy as np, torch
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
from tests.test_arm_gpu import _push_actions, _qv, _env
def tilt_deg(q):
    zz = 1 - 2 * (q[:, [10, 17, 24]] ** 2 + q[:, [11, 18, 25]] ** 2)
    return np.degrees(np.arccos(np.clip(zz, -1, 1))).max(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512; b = build_planar_push()
for prec in ("f64", "mixed"):
    ora = push_states(b, n, 0)
    env = _env("robot_planar_push", n, prec)
    rng = np.random.RandomState(1)
    rows = []
    for k in range(10):
        env.set_state(copy_states(ora.states, n), forward=False)
        a = _push_actions(ora, rng, n, k)
        t0 = tilt_deg(ora.field("qpos", 27)); before = ora.field("qpos", 27)[:, 6:].copy()
        ora.step(a.astype(np.float64)); env.step(torch.from_numpy(a).cuda())
        q, v, st = _qv(env, 27, 24)
        qo = ora.field("qpos", 27)
        t1 = np.maximum(t0, tilt_deg(qo))
        moved = np.abs(qo[:, 6:] - before).max(1) > 1e-4
        dq = np.abs(q - qo).max(1)
        for e in range(n): rows.append((t1[e], moved[e], dq[e]))
    r = np.array(rows)
    for lo, hi in ((0, 1e-4), (1e-4, 1e-2), (1e-2, 0.1), (0.1, 1), (1, 5), (5, 30), (30, 181)):
        for mv in (0, 1):
            m = (r[:, 0] >= lo) & (r[:, 0] < hi) & (r[:, 1] == mv)
            if m.sum(): print(prec, "tilt[%g,%g) moved=%d" % (lo, hi, mv), "count", int(m.sum()), "dq median %.2e p90 %.2e p99 %.2e max %.2e" % (np.median(r[m, 2]), np.percentile(r[m, 2], 90), np.percentile(r[m, 2], 99), r[m, 2].max()))
    env.close()

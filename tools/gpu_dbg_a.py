import sys; sys.path.insert(0, ".")
import numpy as np, torch, ctypes
import mujoco_sim_b200 as ms
from mujoco_sim_b200.model.tasks import build_robot_reach
from tests.helpers import copy_states, arm_states
from tests.test_arm_gpu import _qv
from oracle.oracle import OracleEnvs
np.set_printoptions(precision=9, linewidth=200)
b = build_robot_reach(); b.task.time_limit = 0.0
n = 256
ora = arm_states(b, n, 0)
rng = np.random.RandomState(1)
lo, hi = np.array([-0.1, -0.6, 0.02]), np.array([0.1, -0.4, 0.2])
TARGETS = {80: 118, 132: 60, 152: 3}
for k in range(153):
    a = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    if k in TARGETS:
        e = TARGETS[k]
        pre = copy_states(ora.states, n)
        for prec in ("mixed", "f64"):
            for m in (1, 2, 3, 4, 5, 6, 8, 10, 12, 14, 16, 18, 20):
                bm = build_robot_reach(); bm.task.time_limit = 0.0; bm.task.n_sub = m
                ms.envs.register_builder("_dbgm", lambda: bm)
                env = ms.BatchedMujocoSimEnv("_dbgm", 1, precision=prec, autoreset=False)
                one = copy_states((type(pre[0]) * 1).from_address(ctypes.addressof(pre) + e * ctypes.sizeof(pre[0])), 1)
                env.set_state(one, forward=False)
                om = OracleEnvs(bm, 1); ctypes.memmove(om.states, one, ctypes.sizeof(one))
                om.step(a[e:e+1].astype(np.float64)); env.step(torch.from_numpy(a[e:e+1]).cuda())
                q, v, st = _qv(env, 6, 6)
                print(k, e, prec, "m", m, "dq", np.abs(q - om.field("qpos", 6)).max(), "dv", np.abs(v - om.field("qvel", 6)).max(), "v ora", om.field("qvel", 6)[0], "ncon", om.ncon(0), flush=True)
                env.close()
    ora.step(a.astype(np.float64))

import sys, ctypes
sys.path.insert(0, ".")
import numpy as np, torch
np.set_printoptions(precision=8, suppress=True, linewidth=220)
import mujoco_sim_b200 as ms
from mujoco_sim_b200 import lib as L
from mujoco_sim_b200.model.tasks import build_planar_push
from tests.helpers import push_states, copy_states
n = 32
nsub = int(sys.argv[1]) if len(sys.argv) > 1 else 20
bundle = build_planar_push()
bundle.task.n_sub = nsub
ms.envs.register_builder("_push_dbg", lambda: bundle)
ora = push_states(bundle, n, 0)
env = ms.BatchedMujocoSimEnv("_push_dbg", n, precision="f64", autoreset=False)
env.set_state(copy_states(ora.states, n), forward=False)
qp = ora.field("qpos", 27)
a = qp[:, 6:8].astype(np.float32)
ora.step(a.astype(np.float64))
env.step(torch.from_numpy(a).cuda())
st = env.get_state()
q = np.array([list(s.qpos)[:27] for s in st]); qo = ora.field("qpos", 27)
v = np.array([list(s.qvel)[:24] for s in st]); vo = ora.field("qvel", 24)
err = np.abs(q - qo).max(1)
print("per-env max |dq|:", err)
print("status gpu", [s.status for s in st], "ora", ora.field("status"))
w = int(np.argmax(err))
print("worst env", w, "dq", (q - qo)[w], "\n dv", (v - vo)[w])

"""The step kernel's scheduling knobs change timing only: every lock-step level (MJS_SYNC_LEVEL, read at each launch) and the
persistent work-queue grid (MJS_QUEUE, read once per process, so checked in a subprocess) reproduce the default rollout bit
for bit (INTEGRATION.md "Tuning knobs")."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

_ROLLOUT = """
import hashlib, numpy as np, torch, sys
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
env = ms.make("robot_planar_push", 300, precision=sys.argv[1], seed=77)   # 300 envs: full blocks plus a ragged one
env.reset()
h = hashlib.sha256()
for k in range(4):
    env.step_flat(None)
    h.update(env._obs.cpu().numpy().tobytes()); h.update(env._reward.cpu().numpy().tobytes()); h.update(env._flags.cpu().numpy().tobytes())
st = env.get_state()
h.update(np.array([list(s.qpos)[:27] + list(s.qvel)[:24] for s in st]).tobytes())
print("DIGEST", h.hexdigest())
"""


def _digest(prec, **env):
    e = dict(os.environ); e.update({k: str(v) for k, v in env.items()})
    out = subprocess.run([sys.executable, "-c", _ROLLOUT, prec], env=e, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    return [l for l in out.stdout.splitlines() if l.startswith("DIGEST")][0]


@pytest.mark.parametrize("prec", ["f64", "mixed"])
def test_lock_step_levels_and_work_queue_do_not_change_results(prec):
    ref = _digest(prec)
    for level in (0, 3, 49):
        assert _digest(prec, MJS_SYNC_LEVEL=level) == ref, f"MJS_SYNC_LEVEL={level}"
    assert _digest(prec, MJS_QUEUE=1) == ref, "MJS_QUEUE=1"
    for level in (0, 1, 3):  # the static grid (one block per 12 / 16 envs) with its lock-step levels
        assert _digest(prec, MJS_QUEUE=0, MJS_SYNC_LEVEL=level) == ref, f"MJS_QUEUE=0 MJS_SYNC_LEVEL={level}"

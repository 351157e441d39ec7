"""GPU parity: PointMass Reach CUDA kernel vs the CPU oracle (BASELINE.json configs[1]).

Gate (north_star, contact-free tasks): qpos/qvel within 1e-5 relative over
200-step rollouts.  PointMass is only contact-free in the body-body sense: the
weld and the wall planes are exercised here.
"""
import ctypes

import numpy as np
import pytest
import torch

from tests.helpers import copy_states, pointmass_states, rel_err

pytestmark = pytest.mark.gpu

TOL_REL = 1e-5  # north_star tolerance for contact-free tasks


def _rollout(precision, n=1024, steps=200, seed=0):
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import build_pointmass_reach
    from oracle.oracle import OracleEnvs

    bundle = build_pointmass_reach()
    # time limit off so the 200-step trajectory is one uninterrupted rollout
    bundle.task.time_limit = 0.0
    ora = OracleEnvs(bundle, n)
    init = pointmass_states(n, seed)
    ctypes.memmove(ora.states, init, ctypes.sizeof(init))
    ora.finish_reset()
    env = ms.BatchedMujocoSimEnv("point_mass_reach", n, precision=precision, autoreset=False, solver_tolerance=1e-12)
    env.bundle.task.time_limit = 0.0
    env.close()
    # re-create with the patched task (time limit disabled)
    env = ms.BatchedMujocoSimEnv.__new__(ms.BatchedMujocoSimEnv)
    ms.envs.register_builder("_pm_nolimit", lambda: bundle)
    env.__init__("_pm_nolimit", n, precision=precision, autoreset=False, solver_tolerance=1e-12)
    env.set_state(copy_states(ora.states, n))
    rng = np.random.RandomState(seed + 1)
    worst_q = worst_v = 0.0
    rew_err = 0.0
    for k in range(steps):
        a = rng.uniform(-0.05, 0.05, (n, 2)).astype(np.float32)
        obs_o, rew_o, _, flags_o = ora.step(a.astype(np.float64))
        obs, rew, term, trunc, info = env.step(torch.from_numpy(a).cuda())
        st = env.get_state()
        q = np.array([[s.qpos[0], s.qpos[1]] for s in st])
        v = np.array([[s.qvel[0], s.qvel[1]] for s in st])
        qo, vo = ora.field("qpos", 2), ora.field("qvel", 2)
        worst_q = max(worst_q, rel_err(q, qo, floor=1e-2).max())
        worst_v = max(worst_v, rel_err(v, vo, floor=1e-2).max())
        rew_err = max(rew_err, np.abs(rew.cpu().numpy() - rew_o).max())
    return worst_q, worst_v, rew_err, flags_o, env


def test_pointmass_f64_matches_oracle_tightly():
    """The gate of BASELINE.json configs[1] (200-step rollout within 1e-5) is held by the default f64 path."""
    wq, wv, re_, _, _ = _rollout("f64", n=1024, steps=200)
    print("pointmass f64: worst rel qpos", wq, "qvel", wv, "reward abs", re_)
    assert wq < 1e-9 and wv < 1e-8, (wq, wv)


def test_pointmass_mixed_statistics():
    """FP32 solver variant: a few 1e-5 relative on qpos over 200 steps (wall contacts switch on a threshold);
    reported and bounded at 1e-4, not the gate-holder."""
    wq, wv, re_, _, _ = _rollout("mixed", n=1024, steps=200)
    print("pointmass mixed: worst rel qpos", wq, "qvel", wv, "reward abs", re_)
    assert wq < 1e-4 and wv < 1e-2, (wq, wv)
    assert re_ < 1e-5

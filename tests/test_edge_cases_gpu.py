"""GPU parity on the edge cases of the batched path: ragged env counts (blocks with surplus lock-step warps),
a single env, unreachable actions (IK failure is a status bit, robot.py:221-224), masked resets, and the error
returns of the C ABI."""
import ctypes

import numpy as np
import pytest
import torch

from tests.helpers import copy_states, push_states
from tests.test_arm_gpu import _env, _push_actions, _qv

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 5, 13, 37])
def test_ragged_env_counts_match_the_oracle(n):
    from mujoco_sim_b200.model.tasks import build_planar_push

    b = build_planar_push()
    ora = push_states(b, n, 11)
    env = _env("robot_planar_push", n, "f64")
    env.set_state(copy_states(ora.states, n), forward=False)
    rng = np.random.RandomState(12)
    for k in range(3):
        a = _push_actions(ora, rng, n, k)
        obs_o, rew_o, _, fl_o = ora.step(a.astype(np.float64))
        env.step(torch.from_numpy(a).cuda())
        q, v, st = _qv(env, 27, 24)
        assert np.abs(q - ora.field("qpos", 27)).max() < 1e-6
        np.testing.assert_array_equal(env._flags.cpu().numpy(), fl_o)
        assert np.abs(env._obs.cpu().numpy() - obs_o).max() < 1e-5
    env.close()


def test_unreachable_action_sets_the_ik_status_bit_like_the_oracle():
    from mujoco_sim_b200.model.cstruct import MACROS
    from mujoco_sim_b200.model.tasks import build_planar_push

    n, b = 8, build_planar_push()
    ora = push_states(b, n, 13)
    env = _env("robot_planar_push", n, "f64")
    env.set_state(copy_states(ora.states, n), forward=False)
    a = np.tile(np.array([[1.0, 1.0]], np.float32), (n, 1))  # corner of the [-1,1]^2 action spec: outside the UR5e's reach
    a[::2] = [0.0, -0.45]                                      # every other env gets a reachable target
    ora.step(a.astype(np.float64)); env.step(torch.from_numpy(a).cuda())
    q, v, st = _qv(env, 27, 24)
    bit = MACROS["MJS_STATUS_IK_FAIL"]
    gpu = np.array([s.status & bit for s in st]); cpu = ora.field("status") & bit
    np.testing.assert_array_equal(gpu != 0, cpu != 0)
    assert (gpu[1::2] != 0).all() and (gpu[::2] == 0).all()
    assert np.abs(q - ora.field("qpos", 27)).max() < 1e-6   # the previous target is held
    # the same words through mjs_status (device buffer, no host round trip of the state)
    np.testing.assert_array_equal(env.status().cpu().numpy(), np.array([s.status for s in st]))
    env.close()
    import mujoco_sim_b200 as ms
    pm = ms.make("point_mass_reach", 33, seed=1); pm.reset(); pm.step(None)
    assert pm.status().shape == (33,) and int(pm.status().abs().sum()) == 0
    pm.close()


def test_masked_reset_only_touches_the_masked_envs():
    import mujoco_sim_b200 as ms

    n = 16
    env = ms.make("robot_planar_push", n, seed=4)
    env.reset()
    for _ in range(2):
        env.step(None)
    before = env.get_state()
    mask = torch.zeros(n, dtype=torch.uint8); mask[3] = 1; mask[10] = 1
    env.reset(env_mask=mask)
    after = env.get_state()
    for e in range(n):
        same = list(before[e].qpos)[:27] == list(after[e].qpos)[:27]
        assert same == (e not in (3, 10)), e
        assert after[e].episode_id == before[e].episode_id + (1 if e in (3, 10) else 0)
    env.close()


def test_c_abi_error_returns():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.cstruct import MjsEnvState

    env = ms.make("point_mass_reach", 8)
    arr = (MjsEnvState * 4)()
    assert env._L.mjs_get_state(env._h, arr, 6, 4) != 0           # range runs past the end
    assert b"out of bounds" in env._L.mjs_last_error()
    assert env._L.mjs_set_state(env._h, arr, -1, 4, 0) != 0
    out = (ctypes.c_double * 16)()
    assert env._L.mjs_debug_forward(env._h, 0, out, 16) != 0       # articulated tasks only
    with pytest.raises(AssertionError):
        env.step(torch.zeros(8, 3).cuda())                          # wrong action width (point_reach.py:158)
    assert env._L.mjs_num_envs(env._h) == 8 and env._L.mjs_obs_dim(env._h) == 4 and env._L.mjs_action_dim(env._h) == 2
    env.close()


def test_pointmass_state_ranges_travel_alone():
    """mjs_get_state / mjs_set_state of a sub-range of a PointMass batch move only that range (strided copies of the
    field-major arrays): the range reads back what was written, its neighbours are untouched."""
    import mujoco_sim_b200 as ms

    n = 1000
    env = ms.make("point_mass_reach", n, precision="f64", seed=3)
    env.reset()
    env.step(None)
    before = env.get_state()
    first, count = 333, 77
    sub = env.get_state(first, count)
    for i in range(count):
        assert sub[i].qpos[0] == before[first + i].qpos[0] and sub[i].episode_step == before[first + i].episode_step
        sub[i].qpos[0] = 0.001 * i; sub[i].qvel[1] = -0.002 * i; sub[i].target[0] = 0.01; sub[i].episode_step = 7
    env.set_state(sub, first=first, forward=False)
    after = env.get_state()
    for e in range(n):
        if first <= e < first + count:
            i = e - first
            assert after[e].qpos[0] == 0.001 * i and after[e].qvel[1] == -0.002 * i and after[e].target[0] == 0.01 and after[e].episode_step == 7
            assert after[e].qpos[1] == before[e].qpos[1] and after[e].time == before[e].time
        else:
            assert bytes(after[e]) == bytes(before[e])
    assert len(env.get_state(n, 0)) == 0
    env.close()

"""mjs_step_noalloc (no counterpart in the reference): the step launch whose
scratch -- Planar Push's settle-job list -- lives in a file-scope static __device__ array, so the call allocates nothing.
Checked through the C ABI against mjs_step on a twin simulator (same seed, same device-drawn actions), two calls in a row
and then across an episode boundary (3-step episodes: slot adoption and settle jobs are active in every call).
Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _pair(n, seed):
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import RobotPushConfig

    cfg = RobotPushConfig(max_control_steps_per_episode=3)
    return [ms.make("robot_planar_push", n, seed=seed, precision="f64", config=cfg) for _ in range(2)]


def _call(env, fn):
    from mujoco_sim_b200 import lib

    st = ctypes.c_void_p(torch.cuda.current_stream(env.device).cuda_stream)
    lib.check(fn(env._h, None, _ptr(env._obs), _ptr(env._reward), _ptr(env._flags), _ptr(env._final_obs), 1, st))
    torch.cuda.synchronize()
    return env._obs.cpu().numpy().copy(), env._reward.cpu().numpy().copy(), env._flags.cpu().numpy().copy(), env._final_obs.cpu().numpy().copy()


def test_noalloc_step_equals_step_on_two_calls_in_a_row_and_across_resets():
    n = 200
    ref, new = _pair(n, 21)
    L = ref._L
    ref.reset(); new.reset()
    torch.cuda.synchronize()
    finished = 0
    for k in range(8):
        o0, r0, f0, fo0 = _call(ref, L.mjs_step)
        o1, r1, f1, fo1 = _call(new, L.mjs_step_noalloc)
        # same kernels, same inputs: the tolerance of the other same-path GPU tests (bit-identical)
        np.testing.assert_array_equal(f0, f1, err_msg=f"flags differ at call {k}")
        np.testing.assert_array_equal(o0, o1, err_msg=f"obs differ at call {k}")
        np.testing.assert_array_equal(r0, r1, err_msg=f"reward differs at call {k}")
        np.testing.assert_array_equal(fo0, fo1, err_msg=f"final obs differ at call {k}")
        finished += int(f1[:, 3].sum())
    assert finished >= 2 * n  # every env went through two episode ends: the settle-job list was in use
    s0, s1 = ref.get_state(), new.get_state()
    q0 = np.array([list(s.qpos)[:27] for s in s0]); q1 = np.array([list(s.qpos)[:27] for s in s1])
    np.testing.assert_array_equal(q0, q1)
    ref.close(); new.close()


def test_noalloc_scratch_is_shared_and_recleared_between_simulators():
    """Two simulators alternate on the one static array: stale counters / entries of the other must not leak."""
    a_ref, a_new = _pair(96, 5)
    b_ref, b_new = _pair(40, 6)
    L = a_ref._L
    for e in (a_ref, a_new, b_ref, b_new):
        e.reset()
    for k in range(5):
        for ref, new in ((a_ref, a_new), (b_ref, b_new)):
            o0, r0, f0, _ = _call(ref, L.mjs_step)
            o1, r1, f1, _ = _call(new, L.mjs_step_noalloc)
            np.testing.assert_array_equal(f0, f1)
            np.testing.assert_array_equal(o0, o1)
            np.testing.assert_array_equal(r0, r1)
    for e in (a_ref, a_new, b_ref, b_new):
        e.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch_and_forwards_scratch_free_tasks():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200 import lib

    # PointMass uses no scratch: forwarded to mjs_step, same result as the twin
    p0 = ms.make("point_mass_reach", 64, seed=2); p1 = ms.make("point_mass_reach", 64, seed=2)
    p0.reset(); p1.reset()
    for _ in range(2):
        o0, r0, f0, _ = _call(p0, p0._L.mjs_step)
        o1, r1, f1, _ = _call(p1, p1._L.mjs_step_noalloc)
        np.testing.assert_array_equal(o0, o1); np.testing.assert_array_equal(f0, f1)
    p0.close(); p1.close()
    # 546 134 envs need 16 385 list entries: one more than the array holds -> clear error, nothing launched
    big = ms.make("robot_planar_push", 546134, seed=1)
    st = ctypes.c_void_p(torch.cuda.current_stream(big.device).cuda_stream)
    rc = big._L.mjs_step_noalloc(big._h, None, _ptr(big._obs), _ptr(big._reward), _ptr(big._flags), None, 1, st)
    assert rc != 0
    assert b"static scratch holds 16384" in lib.load().mjs_last_error()
    big.close()

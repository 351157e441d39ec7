"""Python restatement of the device-side randomness (mujoco_sim_b200/csrc/mjs_common.cuh): Philox4x32-10 keyed by the seed,
counter = (global env id, episode id, purpose, index), 53-bit uniforms -- and of the Planar Push episode reset that consumes
it (arm_kernel.cuh reset_begin, restating robot_planar_push.py:149-176).  Lets the CPU oracle start an episode from exactly
the state the device reset samples, so free-running comparisons can cross episode boundaries."""
import ctypes

import numpy as np

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF
RNG_RESET, RNG_ACTION, RNG_DISTURB = 1, 2, 3


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> 32, p0 & MASK, p1 >> 32, p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c0, c1, c2, c3


def u53(a, b):
    return ((a >> 5) * 67108864.0 + (b >> 6)) * (1.0 / 9007199254740992.0)


def uniform(seed, genv, episode, purpose, idx, lo, hi):
    v = philox4x32_10(genv & MASK, episode & MASK, purpose, idx >> 1, seed & MASK, (seed >> 32) & MASK)
    u = u53(v[2], v[3]) if idx & 1 else u53(v[0], v[1])
    return lo + (hi - lo) * u


def device_actions(task, seed, genv, step_index):
    """the synthetic-rollout action of env `genv` at launch `step_index` (actions == NULL path of mjs_step), before the float32 cast"""
    return np.array([uniform(seed, genv, step_index, RNG_ACTION, k, task.action_lo[k], task.action_hi[k]) for k in range(task.action_dim)])


def push_reset_state(bundle, state, seed, genv, episode):
    """Fill `state` (MjsEnvState) with the pre-settle reset state of Planar Push for (seed, env, episode)."""
    from oracle.oracle import _dp, lib

    t = bundle.task
    idx = 0
    spawn = np.zeros(3)
    for k in range(3):
        spawn[k] = uniform(seed, genv, episode, RNG_RESET, idx, t.spawn_lo[k], t.spawn_hi[k]); idx += 1
    q = np.zeros(6)
    ok = lib().oracle_tcp_to_joints(ctypes.byref(t), _dp(spawn), _dp(np.zeros(6)), _dp(q))
    if not ok:
        q[:] = 0
    target = np.zeros(3)
    for k in range(3):
        target[k] = uniform(seed, genv, episode, RNG_RESET, idx, t.target_lo[k], t.target_hi[k]); idx += 1
    n = t.n_obj
    pos = np.zeros((n, 3))
    dmin = 2 * t.obj_radius
    for _ in range(64):
        for b in range(n):
            for k in range(3):
                pos[b, k] = uniform(seed, genv, episode, RNG_RESET, idx, t.obj_lo[k], t.obj_hi[k]); idx += 1
        coll = False
        for b in range(n):
            if np.hypot(pos[b, 0] - spawn[0], pos[b, 1] - spawn[1]) < dmin:
                coll = True
            for c in range(b):
                if np.hypot(pos[b, 0] - pos[c, 0], pos[b, 1] - pos[c, 1]) < dmin:
                    coll = True
        if not coll:
            break
    ctypes.memset(ctypes.byref(state), 0, ctypes.sizeof(state))
    for k in range(6):
        state.qpos[k] = q[k]; state.ctrl[k] = q[k]
    for b in range(n):
        a = t.obj_qadr[b]
        for k in range(3):
            state.qpos[a + k] = pos[b, k]
        state.qpos[a + 3] = 1.0
    for k in range(3):
        state.target[k] = target[k]
    state.episode_id = episode
    return spawn, target, pos


def button_disturb_u(seed, genv, episode, step_in_episode):
    """the uniform draw of the button-disturbance test (arm_kernel.cuh control_step, restating robot_push_button.py:159-165)"""
    return uniform(seed, genv, episode, RNG_DISTURB, step_in_episode, 0.0, 1.0)

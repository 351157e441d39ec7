"""Shared helpers for the parity tests."""
import numpy as np

from mujoco_sim_b200.model.cstruct import MjsEnvState


def copy_states(src, n):
    import ctypes

    dst = (MjsEnvState * n)()
    ctypes.memmove(dst, src, ctypes.sizeof(MjsEnvState) * n)
    return dst


def pointmass_states(n, seed, near_walls=True):
    """Host-injected PointMass initial states (SURVEY.md 7.3 'Randomness')."""
    rng = np.random.RandomState(seed)
    states = (MjsEnvState * n)()
    for e in range(n):
        s = states[e]
        lim = 0.45
        x, y = rng.uniform(-lim, lim, 2)
        if near_walls and e % 4 == 0:  # start a quarter of the envs next to a wall / corner
            x = np.sign(x) * rng.uniform(0.40, 0.45)
            if e % 8 == 0:
                y = np.sign(y) * rng.uniform(0.40, 0.45)
        s.qpos[0], s.qpos[1] = x, y
        s.mocap_pos[0], s.mocap_pos[1] = x, y
        s.target[0], s.target[1] = rng.uniform(-lim, lim, 2)
        s.target[2] = 0.025
        s.dist = 1.0
        s.prev_dist = 1.0
    return states


def rel_err(a, b, floor=1e-3):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.abs(a - b) / np.maximum(np.abs(b), floor)


def push_states(bundle, n, seed, settle=True):
    """Host-injected Planar Push states: IK'd arm pose, props dropped and settled by the oracle."""
    import ctypes

    from oracle.oracle import OracleEnvs, _dp, lib

    rng = np.random.RandomState(seed)
    t, m = bundle.task, bundle.model
    ora = OracleEnvs(bundle, n)
    for e in range(n):
        s = ora.states[e]
        while True:
            tcp = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.6, -0.3), 0.02])
            objs = np.stack([rng.uniform(-0.15, 0.15, t.n_obj), rng.uniform(-0.55, -0.35, t.n_obj), rng.uniform(0.03, 0.08, t.n_obj)], 1)
            d = [np.hypot(*(objs[i, :2] - objs[j, :2])) for i in range(t.n_obj) for j in range(i)]
            d += [np.hypot(*(objs[i, :2] - tcp[:2])) for i in range(t.n_obj)]
            if min(d) > 0.045:
                break
        q = np.zeros(6)
        ok = lib().oracle_tcp_to_joints(ctypes.byref(t), _dp(tcp), _dp(np.zeros(6)), _dp(q))
        assert ok
        for i in range(6):
            s.qpos[i] = q[i]
            s.ctrl[i] = q[i]
        for k in range(t.n_obj):
            a = t.obj_qadr[k]
            for i in range(3):
                s.qpos[a + i] = objs[k, i]
            s.qpos[a + 3] = 1.0
        s.target[0], s.target[1], s.target[2] = rng.uniform(-0.15, 0.15), rng.uniform(-0.55, -0.35), 0.003
    ora.finish_reset(150 if settle else 0)
    return ora


def arm_states(bundle, n, seed):
    """Host-injected Robot Reach / Button Push states: IK'd spawn pose in the task's spawn box
    (robot_reach.py:143-150, robot_push_button.py:126-134), target / switch position in its box."""
    import ctypes

    from oracle.oracle import OracleEnvs, _dp, lib

    rng = np.random.RandomState(seed)
    t = bundle.task
    ora = OracleEnvs(bundle, n)
    lo = np.array([t.spawn_lo[k] for k in range(3)]); hi = np.array([t.spawn_hi[k] for k in range(3)])
    tlo = np.array([t.target_lo[k] for k in range(3)]); thi = np.array([t.target_hi[k] for k in range(3)])
    button = bundle.name == "robot_push_button"
    for e in range(n):
        s = ora.states[e]
        while True:
            tcp = rng.uniform(lo, hi)
            tgt = rng.uniform(tlo, thi)
            if button:
                tcp[2] = max(tcp[2], 0.12)  # start above the switch
            q = np.zeros(6)
            if lib().oracle_tcp_to_joints(ctypes.byref(t), _dp(tcp), _dp(np.zeros(6)), _dp(q)):
                break
        for i in range(6):
            s.qpos[i] = q[i]
            s.ctrl[i] = q[i]
        for i in range(3):
            if button:
                s.ovr_pos[i] = tgt[i]
            else:
                s.target[i] = tgt[i]
    ora.finish_reset(0)
    return ora

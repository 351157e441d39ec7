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

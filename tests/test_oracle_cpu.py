"""CPU checks of the oracle (oracle/mjs_oracle.cpp) itself.

The reference holds no golden vectors for this path (parity unpinned), so the oracle is
pinned here against (i) independent closed-form / numpy restatements of small cases and (ii) the
reference's own loose tests restated (test/test_ur_control_api.py, test/test_ur_frame_matches_real.py).
"""
import ctypes

import numpy as np
import pytest

from mujoco_sim_b200.model import tasks as T
from oracle.oracle import OracleEnvs, ur_fk, ur_ik_all, ur_ik_closest
from tests.helpers import pointmass_states, push_states


# ------------------------------------------------------------------ PointMass: closed-form weld dynamics
def _impedance(r, solimp):
    d0, dw, w, mid, p = solimp
    x = min(1.0, abs(r) / w)
    if p == 1:
        y = x
    elif x < mid:
        y = x**p / mid ** (p - 1)
    else:
        y = 1 - (1 - x) ** p / (1 - mid) ** (p - 1)
    return min(max(d0 + y * (dw - d0), 1e-4), 0.9999)


def _numpy_pointmass(q, v, mocap, target, actions, n_sub=5, h=0.02, m=0.1):
    """SURVEY.md App. A.1/A.3 for the wall-free interior: two independent weld rows with J = I, the
    legacy step2->step1 ordering (rows are built by the previous step1 with the mocap of that time)."""
    solimp = (0.9, 0.95, 0.001, 0.5, 2.0)
    tc, zeta = max(0.02, 2 * h), 1.0  # refsafe
    k, b = 1 / (solimp[1] ** 2 * tc**2 * zeta**2), 2 / (solimp[1] * tc)
    diag = (1 / m + 1 / m + 0.0) / 3.0  # body_invweight0 (translational) of the pointmass; mocap has 0

    def rows(q, v, mocap):
        out = []
        for i in range(2):
            r = q[i] - mocap[i]
            d = _impedance(r, solimp)
            R = max(1e-15, (1 - d) / d * diag)
            out.append((1 / R, -b * v[i] - k * d * r))
        return out

    rw = rows(q, v, mocap)
    rewards = []
    for a in actions:
        mocap = np.clip(q + a, -0.5, 0.5)  # before_step: point_reach.py:150-163
        for _ in range(n_sub):
            acc = np.array([D * aref / (m + D) for D, aref in rw])  # argmin 1/2 m a^2 + 1/2 D (a - aref)^2
            v = v + h * acc
            q = q + h * v
            rw = rows(q, v, mocap)
        rewards.append(-np.hypot(*(q - target)) + 0.5)
    return q, v, rewards


def test_pointmass_oracle_matches_closed_form_weld():
    b = T.build_pointmass_reach()
    b.task.time_limit = 0.0
    n = 8
    ora = OracleEnvs(b, n)
    init = pointmass_states(n, 3, near_walls=False)
    for s in init:  # keep the whole rollout in the wall-free interior
        s.qpos[0] *= 0.3; s.qpos[1] *= 0.3
        s.mocap_pos[0], s.mocap_pos[1] = s.qpos[0], s.qpos[1]
    ctypes.memmove(ora.states, init, ctypes.sizeof(init))
    ora.finish_reset()
    rng = np.random.RandomState(0)
    acts = rng.uniform(-0.05, 0.05, (6, n, 2))
    rews = []
    for k in range(6):
        _, r, _, _ = ora.step(acts[k])
        rews.append(r)
    for e in range(n):
        s = init[e]
        q, v, rw = _numpy_pointmass(np.array([s.qpos[0], s.qpos[1]]), np.zeros(2), np.array([s.qpos[0], s.qpos[1]]),
                                    np.array([s.target[0], s.target[1]]), acts[:, e])
        np.testing.assert_allclose([ora.states[e].qpos[0], ora.states[e].qpos[1]], q, rtol=0, atol=1e-10)
        np.testing.assert_allclose([ora.states[e].qvel[0], ora.states[e].qvel[1]], v, rtol=0, atol=1e-9)
        np.testing.assert_allclose([r[e] for r in rews], rw, atol=1e-10)


def test_pointmass_walls_keep_the_sphere_inside():
    b = T.build_pointmass_reach()
    b.task.time_limit = 0.0
    n = 4
    ora = OracleEnvs(b, n)
    init = pointmass_states(n, 1)
    ctypes.memmove(ora.states, init, ctypes.sizeof(init))
    ora.finish_reset()
    push = np.tile(np.array([0.05, 0.05]), (n, 1))
    for _ in range(60):
        ora.step(push)
    q = ora.field("qpos", 2)
    # weld pulls to the clipped mocap target 0.5, the wall contact (same solref/solimp family, active
    # beyond 0.45) pushes back: two equally stiff soft constraints balance near the midpoint
    assert (q < 0.49).all() and (q > 0.455).all(), q


def test_time_limit_matches_accumulated_double_time():
    # 250 additions of 0.02 vs the literal 5.0
    t, steps = 0.0, 0
    while t < 5.0:
        for _ in range(5):
            t += 0.02
        steps += 1
    assert T._accumulated_step_limit(0.1, 0.02, 5.0) == steps
    b = T.build_pointmass_reach()
    ora = OracleEnvs(b, 1)
    init = pointmass_states(1, 0)
    ctypes.memmove(ora.states, init, ctypes.sizeof(init))
    ora.finish_reset()
    k = 0
    while True:
        _, _, _, fl = ora.step(np.zeros((1, 2)))
        k += 1
        if fl[0, 3]:
            break
        ora.states[0].target[0] = 10.0  # never reach the goal
    assert k == steps


# ------------------------------------------------------------------ UR5e analytic IK / FK
def test_ur_ik_roundtrip_all_branches():
    rng = np.random.RandomState(0)
    for _ in range(50):
        q = rng.uniform(-2.5, 2.5, 6)
        Tm = ur_fk(T.UR5E_DH, q)
        sols = ur_ik_all(T.UR5E_DH, Tm)
        assert len(sols) >= 1
        for s in sols:
            np.testing.assert_allclose(ur_fk(T.UR5E_DH, s), Tm, atol=1e-8)
        qc = ur_ik_closest(T.UR5E_DH, Tm, q)
        np.testing.assert_allclose(qc, q, atol=1e-6)


def test_fk_of_mjcf_model_matches_analytic_fk_at_zero():
    """test/test_ur_frame_matches_real.py:7-29 restated: attachment_site pose at q=0 == analytic FK (atol 1e-2)."""
    b = T.build_robot_reach()
    ora = OracleEnvs(b, 1)
    ora.finish_reset()
    for k in range(6):
        ora.states[0].qpos[k] = 0.0
    d = ora.stage_dump(0)
    site = b.task.tcp_site
    Tm = ur_fk(T.UR5E_DH, np.zeros(6))
    np.testing.assert_allclose(d["site_xpos"][site], Tm[:3, 3], atol=1e-2)
    np.testing.assert_allclose(d["site_xmat"][site].reshape(3, 3), Tm[:3, :3], atol=1e-2)

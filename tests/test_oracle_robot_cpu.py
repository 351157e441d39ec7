"""CPU checks of the oracle's articulated path: closed-form free fall, internal consistency of the
smooth dynamics, and the reference's own loose control tests restated (atol 1e-2)."""
import ctypes

import numpy as np

from mujoco_sim_b200.model import tasks as T
from oracle.oracle import OracleEnvs, _dp, lib, ur_fk
from tests.helpers import push_states


def _push_env(n=1, seed=0, settle=False):
    return push_states(T.build_planar_push(), n, seed, settle=settle)


def test_free_prop_falls_like_semi_implicit_euler():
    """A prop released above the table: z_k = z_0 - g h^2 k(k+1)/2 until it touches (A.4 free joint)."""
    b = T.build_planar_push()
    ora = push_states(b, 1, 0, settle=False)
    s = ora.states[0]
    a = b.task.obj_qadr[0]
    z0 = 0.2
    s.qpos[a + 2] = z0
    ora.finish_reset(0)
    h, g = b.model.timestep, 9.81
    ora.finish_reset(10)  # 10 physics steps
    z = ora.states[0].qpos[a + 2]
    assert abs(z - (z0 - g * h * h * 10 * 11 / 2)) < 1e-12
    assert abs(ora.states[0].qvel[b.task.obj_dadr[0] + 2] + g * h * 10) < 1e-12


def test_props_settle_flat_on_the_table():
    ora = _push_env(4, 1, settle=True)
    t = ora.task
    for e in range(4):
        for k in range(t.n_obj):
            a = t.obj_qadr[k]
            z = ora.states[e].qpos[a + 2]
            assert abs(z - 0.019) < 1e-3, z  # half-height 0.019, soft contact penetration < 1 mm
            assert abs(ora.states[e].qpos[a + 3]) > 0.9999  # still upright
        assert np.abs(ora.field("qvel", 24)[e]).max() < 5e-2


def test_smooth_dynamics_are_self_consistent():
    """M qacc_smooth == qfrc_passive - qfrc_bias + qfrc_actuator, recomputed here from the model."""
    ora = _push_env(2, 2, settle=True)
    for e in range(2):
        for k in range(24):
            ora.states[e].qvel[k] = 0.0
        d = ora.stage_dump(e)
        M = d["M"]
        np.testing.assert_allclose(M, M.T, atol=1e-12)
        assert np.linalg.eigvalsh(M).min() > 0
        m, st = ora.model, ora.states[e]
        act = np.zeros(24)
        for u in range(m.nu):  # affine position actuators: gain*ctrl + bias1*q + bias2*v, force-clamped (P7)
            j = m.act_dof[u]
            f = m.act_gain[u] * st.ctrl[u] + m.act_bias[3 * u] + m.act_bias[3 * u + 1] * st.qpos[j] + m.act_bias[3 * u + 2] * st.qvel[j]
            act[j] = np.clip(f, m.act_forcerange[2 * u], m.act_forcerange[2 * u + 1]) * m.act_gear[u]
        np.testing.assert_allclose(M @ d["qacc_smooth"], d["qfrc_passive"] - d["qfrc_bias"] + act, atol=1e-9)
        # free cylinders: mass 0.1 on the translational diagonal
        np.testing.assert_allclose(np.diag(M)[6:9], 0.1, atol=1e-12)
        # gravity compensation (robot.py:81-82): the arm's own links cancel; only the pusher (0.1 kg) loads it
        assert np.abs((d["qfrc_passive"] - d["qfrc_bias"])[:6]).max() < 0.1 * 9.81 * 1.0


def test_servoJ_steady_state_tracking():
    """test/test_ur_control_api.py:7-28 (test_moveJ) restated through the Button Push joint action."""
    b = T.build_push_button()
    b.task.time_limit = 0.0
    ora = OracleEnvs(b, 1)
    q0 = np.array([-0.5, -0.5, 0.5, -0.5, -0.5, -0.5]) * np.pi
    s = ora.states[0]
    for k in range(6):
        s.qpos[k] = q0[k]; s.ctrl[k] = q0[k]
    s.ovr_pos[0], s.ovr_pos[1], s.ovr_pos[2] = 0.15, -0.45, 0.0
    ora.finish_reset()
    tgt = q0 + np.array([0.3, -0.2, 0.2, 0.1, -0.3, 0.2])
    act = np.concatenate([tgt, [0.085]])[None]
    for _ in range(40):
        ora.step(act)
    np.testing.assert_allclose(ora.field("qpos", 6)[0], tgt, atol=1e-2)


def test_servoL_converges_to_the_commanded_tcp():
    """test/test_ur_control_api.py:57-82 (test_servoL) restated through Robot Reach's EEF action."""
    b = T.build_robot_reach()
    b.task.time_limit = 0.0
    ora = OracleEnvs(b, 1)
    tcp0 = np.array([0.0, -0.5, 0.1])
    q = np.zeros(6)
    assert lib().oracle_tcp_to_joints(ctypes.byref(b.task), _dp(tcp0), _dp(np.zeros(6)), _dp(q))
    s = ora.states[0]
    for k in range(6):
        s.qpos[k] = q[k]; s.ctrl[k] = q[k]
    ora.finish_reset()
    target = np.array([0.08, -0.42, 0.15])
    for _ in range(20):
        obs, _, _, _ = ora.step(target[None])
    np.testing.assert_allclose(obs[0, :3], target, atol=1e-2)


def test_planar_push_reward_and_flags_follow_the_task_definition():
    """robot_planar_push.py:203-238 recomputed in numpy from the oracle's observation."""
    ora = _push_env(8, 3, settle=True)
    rng = np.random.RandomState(0)
    for k in range(3):
        a = np.stack([rng.uniform(-0.2, 0.2, 8), rng.uniform(-0.6, -0.3, 8)], 1)
        obs, rew, disc, fl = ora.step(a)
        tcp, tgt, blocks = obs[:, :3], obs[:, 3:5], obs[:, 5:11].reshape(8, 3, 2)
        d_t = np.linalg.norm(blocks - tgt[:, None], axis=2)
        d_r = np.linalg.norm(blocks - tcp[:, None, :2], axis=2)
        np.testing.assert_allclose(rew, 0.1 * (-d_t.mean(1) - 0.1 * d_r.min(1)), atol=1e-12)
        done = (d_t < 0.05).all(1)
        np.testing.assert_array_equal(fl[:, 0].astype(bool), done)
        np.testing.assert_allclose(disc, 1.0 - done)


def test_button_push_eef_action_variant_follows_the_joint_variant():
    """robot_push_button.py:143-157: ABS_EEF_ACTION sends the TCP position through servoL (analytic IK, closest solution),
    ABS_JOINT_ACTION the joint targets through servoJ.  The scripted demonstrator (:231-300) computes a Cartesian target and only
    converts it to joints for the joint action type with the same IK, so both variants must trace the same motion; the EEF
    variant observes tcp_position instead of joint_configuration (:109-113)."""
    from tests.helpers import arm_states

    n = 8
    bj, be = T.build_push_button(), T.build_push_button(action_type=T.ABS_EEF_ACTION)
    assert be.task.action_dim == 4 and be.task.obs_dim == 7 and [k for k, _ in be.obs_keys] == ["ur5e/tcp_position", "unnamed_model/position", "unnamed_model/active"]
    oj, oe = arm_states(bj, n, 21), arm_states(be, n, 21)
    succ = 0
    for k in range(100):
        aj, okj = oj.demo_policy(); ae, oke = oe.demo_policy()
        assert ae.shape == (n, 4) and oke.all() and (ae[:, 3] == 0).all()
        obs_j, rew_j, _, fl_j = oj.step(aj); obs_e, rew_e, _, fl_e = oe.step(ae)
        np.testing.assert_allclose(oe.field("qpos", 6), oj.field("qpos", 6), atol=1e-9)
        np.testing.assert_array_equal(fl_e, fl_j)
        np.testing.assert_allclose(obs_e[:, 3:7], obs_j[:, 6:10], atol=1e-9)          # switch position + active
        tcp = np.array([ur_fk_tcp(be, q) for q in oe.field("qpos", 6)])
        # tcp_position observable = the MJCF site; the DH chain of the analytic IK differs from the MJCF by ~1 mm
        # (the reference's own FK check uses atol 1e-2: test/test_ur_frame_matches_real.py)
        np.testing.assert_allclose(obs_e[:, :3], tcp, atol=2e-3)
        succ += int(fl_e[:, 2].sum())
        if fl_e[:, 3].all():
            break
    assert succ >= n // 2


def ur_fk_tcp(bundle, q):
    """analytic FK of the flange (oracle.ur_fk) + the TCP offset along the tool axis (robot.py:126-136)"""
    T06 = ur_fk(np.array([bundle.task.ik_dh[k] for k in range(6)]), np.asarray(q, float))
    off = np.array([bundle.task.tcp_offset[k] for k in range(3)])
    return T06[:3, 3] + T06[:3, :3] @ off

"""GPU checks of the SB3 / gymnasium vector-env facades (numpy in, numpy out, autoreset bookkeeping)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_sb3_facade_protocol_and_terminal_observation():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import RobotPushConfig

    n = 32
    venv = ms.make_vec("mujoco_sim/robot_planar_push_state-v0", n, api="sb3", seed=3, config=RobotPushConfig(max_control_steps_per_episode=4))
    obs = venv.reset()
    assert set(obs) == {"ur5e/tcp_position", "target_position", "block_positions"} and obs["block_positions"].shape == (n, 6)
    rng = np.random.RandomState(0)
    seen_done = False
    for k in range(6):
        a = np.stack([rng.uniform(-0.2, 0.2, n), rng.uniform(-0.6, -0.3, n)], 1)
        obs, rew, dones, infos = venv.step(a)
        assert rew.shape == (n,) and dones.dtype == bool and len(infos) == n
        if k == 3:  # step limit of 4 control steps: every env truncates here
            assert dones.all()
            for info in infos:
                assert info["TimeLimit.truncated"] is True and "terminal_observation" in info and info["discount"] == 1.0
                assert info["terminal_observation"]["block_positions"].shape == (6,)
            seen_done = True
        else:
            assert not dones.any()
    assert seen_done
    venv.close()


def test_gymnasium_facade_matches_device_api():
    import torch

    import mujoco_sim_b200 as ms

    n = 16
    v = ms.make_vec("mujoco_sim/point_mass_reach_state-v0", n, api="gymnasium", seed=5, flatten_observation_space=True)
    e = ms.make("point_mass_reach", n, seed=5)
    o1, info = v.reset()
    e.reset()
    assert info == {} and o1.shape == (n, 4)
    np.testing.assert_array_equal(o1, e._obs.cpu().numpy())
    rng = np.random.RandomState(1)
    for _ in range(60):  # beyond the 50-step time limit: auto-reset happens inside
        a = rng.uniform(-0.05, 0.05, (n, 2)).astype(np.float32)
        o1, r1, term, trunc, info = v.step(a)
        o2, r2, t2, u2, i2 = e.step(torch.from_numpy(a).cuda())
        np.testing.assert_array_equal(o1, e._obs.cpu().numpy())
        np.testing.assert_array_equal(r1, r2.cpu().numpy())
        np.testing.assert_array_equal(term, t2.cpu().numpy()); np.testing.assert_array_equal(trunc, u2.cpu().numpy())
        np.testing.assert_array_equal(info["is_success"], i2["is_success"].cpu().numpy())
        assert (info["_final_observation"] == (term | trunc)).all()
    v.close(); e.close()


def test_env_ids_reset_and_facade_shortcuts():
    """SURVEY.md 8b exports: reset(env_ids=...), single_*_space on the env, as_sb3_vecenv() / as_gymnasium_vector_env()."""
    import torch

    import mujoco_sim_b200 as ms

    env = ms.make("point_mass_reach", 8, seed=5)
    env.reset()
    for _ in range(3):
        env.step(None)
    before = env.get_state()
    obs, _ = env.reset(env_ids=[1, 6])
    after = env.get_state()
    for e in range(8):
        if e in (1, 6):
            assert after[e].episode_step == 0 and after[e].episode_id == before[e].episode_id + 1
        else:
            assert bytes(after[e]) == bytes(before[e])
    assert env.single_action_space.shape == (2,) and list(env.single_observation_space.spaces) == ["pointmass/position", "goal_position"]
    v = env.as_gymnasium_vector_env()
    o, r, t, tr, info = v.step(np.zeros((8, 2), np.float32))
    assert o["pointmass/position"].shape == (8, 2) and r.shape == (8,)
    s = env.as_sb3_vecenv()
    o2, r2, d2, infos = s.step(np.zeros((8, 2), np.float32))
    assert len(infos) == 8 and d2.shape == (8,)
    env.close()

"""Drop-in boundary details (SURVEY.md 8b / VERDICT r1 item 6): re-seeding a live env (dmc2gym.py:126-131,157-163, restating
test/test_gym_envs.py:21-36 literally), float64 observations (dmc2gym.py:66-75), the VecEnv seed / info protocol, and the
per-call device guard of the C ABI."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TASKS = ["point_mass_reach", "robot_reach", "robot_planar_push", "robot_push_button"]


@pytest.mark.parametrize("task", TASKS)
def test_reset_with_seed_on_the_same_env_is_deterministic(task):
    """test/test_gym_envs.py:21-36: `reset(seed=2025)` twice on ONE env object -> identical observations (atol 1e-6 there,
    bit-identical here), `reset(seed=2024)` -> different.  Steps in between move the episode counters and, for Planar Push,
    leave pre-settled slots of the old key behind: none of that may leak."""
    import mujoco_sim_b200 as ms

    env = ms.make(task, 48, seed=7)
    obs, info = env.reset(seed=2025)
    assert info == {}
    first = {k: v.clone() for k, v in obs.items()}
    traj1 = []
    for _ in range(3):
        o, *_ = env.step(None)
        traj1.append(env._obs.clone())
    obs, _ = env.reset(seed=2025)
    for k in first:
        assert torch.equal(obs[k], first[k]), k
    for i in range(3):  # the device-drawn action stream restarts too
        env.step(None)
        assert torch.equal(env._obs, traj1[i])
    obs, _ = env.reset(seed=2024)
    assert max(float((obs[k] - first[k]).abs().max()) for k in first) > 1e-3
    env.seed(2025)          # DMCEnvironmentAdapter.seed: applied by the next reset
    obs, _ = env.reset()
    for k in first:
        assert torch.equal(obs[k], first[k]), k
    env.close()


def test_reseed_discards_presettled_slots_of_the_old_key():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import RobotPushConfig

    cfg = RobotPushConfig(max_control_steps_per_episode=2)
    a = ms.make("robot_planar_push", 64, seed=1, config=cfg)
    b = ms.make("robot_planar_push", 64, seed=99, config=cfg)
    a.reset(); b.reset()
    for _ in range(3):
        b.step(None)                  # b now holds slots and episode counters of seed 99
    b.reset(seed=1)
    for k in range(5):                # crosses two episode boundaries: slot adoption must use seed-1 episodes
        a.step(None); b.step(None)
        assert torch.equal(a._obs, b._obs), k
        assert torch.equal(a._flags, b._flags), k
    a.close(); b.close()


@pytest.mark.parametrize("task", TASKS)
def test_float64_outputs_are_the_unrounded_values_of_the_float32_ones(task):
    import mujoco_sim_b200 as ms

    e32 = ms.make(task, 40, seed=3)
    e64 = ms.make(task, 40, seed=3, obs_dtype="float64")
    o32, _ = e32.reset(); o64, _ = e64.reset()
    for k in o64:
        assert o64[k].dtype == torch.float64
        assert torch.equal(o64[k].float(), o32[k])
    for _ in range(4):
        o32, r32, t32, u32, i32 = e32.step(None)
        o64, r64, t64, u64, i64 = e64.step(None)
        assert r64.dtype == torch.float64 and torch.equal(r64.float(), r32)
        for k in o64:
            assert torch.equal(o64[k].float(), o32[k])
            assert torch.equal(i64["final_observation"][k].float(), i32["final_observation"][k])
        assert float((e64.observation - e64._obs.double()).abs().max()) > 0  # really unrounded, not a cast of the float32 tensor
    e32.close(); e64.close()


def test_float64_observation_matches_the_oracle_to_rounding():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import build_planar_push
    from tests.helpers import copy_states, push_states

    n, b = 32, build_planar_push()
    ora = push_states(b, n, 17)
    env = ms.make("robot_planar_push", n, precision="f64", autoreset=False, solver_tolerance=1e-12, obs_dtype="float64")
    env.set_state(copy_states(ora.states, n), forward=False)
    a = np.stack([np.full(n, 0.05), np.full(n, -0.45)], 1).astype(np.float32)
    obs_o, rew_o, _, _ = ora.step(a.astype(np.float64))
    env.step(torch.from_numpy(a).cuda())
    assert np.abs(env.observation.cpu().numpy() - obs_o).max() < 1e-9
    assert np.abs(env.reward.cpu().numpy() - rew_o).max() < 1e-9
    env.close()


def test_sb3_facade_seed_and_shared_infos():
    import mujoco_sim_b200 as ms

    n = 24
    v = ms.make_vec("mujoco_sim/point_mass_reach_state-v0", n, api="sb3", seed=0)
    assert v.seed(11) == [11 + i for i in range(n)]
    o1 = v.reset()
    v.step(np.zeros((n, 2), np.float32))
    v.seed(11)
    o2 = v.reset()
    for k in o1:
        np.testing.assert_array_equal(o1[k], o2[k])
    obs, rew, dones, infos = v.step(np.zeros((n, 2), np.float32))
    assert len(infos) == n and all(i["discount"] == 1.0 and i["is_success"] == 0.0 for i in infos)
    assert len({id(i) for i in infos}) <= 4   # not-done envs share read-only dicts instead of 24 fresh ones
    v.close()
    g = ms.make_vec("mujoco_sim/robot_reach_state-v0", 8, api="gymnasium", seed=0, obs_dtype="float64")
    o, _ = g.reset(seed=5)
    assert o["ur5e/tcp_position"].dtype == np.float64 and g.single_observation_space.spaces["ur5e/tcp_position"].dtype == np.float64
    o2, _ = g.reset(seed=5)
    np.testing.assert_array_equal(o["target_position"], o2["target_position"])
    g.close()


def test_calls_leave_the_callers_current_device_alone_and_work_across_devices():
    import mujoco_sim_b200 as ms

    cur = torch.cuda.current_device()
    env = ms.make("robot_reach", 8, seed=1)
    env.reset(); env.step(None)
    assert torch.cuda.current_device() == cur
    env.close()
    if torch.cuda.device_count() < 2:
        return
    # a simulator on cuda:1 stepped while the caller's current device is cuda:0, next to one on cuda:0
    e0 = ms.make("robot_planar_push", 16, seed=4, device="cuda:0")
    e1 = ms.make("robot_planar_push", 16, seed=4, device="cuda:1")
    torch.cuda.set_device(0)
    e0.reset(); e1.reset()
    for _ in range(3):
        e0.step(None); e1.step(None)
    assert torch.cuda.current_device() == 0
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    assert torch.equal(e0._obs.cpu(), e1._obs.cpu())
    e0.close(); e1.close()

"""Scripted demonstrators (mjs_demo_policy) against the oracle's restatement of the reference policies, and closed
loop: following them solves the tasks (point_reach.py:227-240, robot_reach.py:222-253, robot_push_button.py:231-300)."""
import ctypes

import numpy as np
import pytest
import torch

from tests.helpers import arm_states, copy_states, pointmass_states

pytestmark = pytest.mark.gpu


def _pair(task, n, seed):
    """(oracle envs, CUDA env) holding the same injected states"""
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model import tasks as T
    from oracle.oracle import OracleEnvs

    if task == "point_mass_reach":
        b = T.build_pointmass_reach()
        ora = OracleEnvs(b, n)
        init = pointmass_states(n, seed)
        ctypes.memmove(ora.states, init, ctypes.sizeof(init))
        ora.finish_reset()
    else:
        b = T.build_robot_reach() if task == "robot_reach" else T.build_push_button()
        b.task.time_limit = 0.0
        ora = arm_states(b, n, seed)
    ms.envs.register_builder("_pol_" + task, lambda: b)
    env = ms.BatchedMujocoSimEnv("_pol_" + task, n, precision="f64", autoreset=False, solver_tolerance=1e-12)
    env.set_state(copy_states(ora.states, n), forward=task != "point_mass_reach")
    return ora, env


@pytest.mark.parametrize("task,steps", [("point_mass_reach", 30), ("robot_reach", 25), ("robot_push_button", 100)])
def test_demonstrator_matches_the_oracle_and_solves_the_task(task, steps):
    n = 32
    ora, env = _pair(task, n, 7)
    worst = 0.0
    for k in range(steps):
        a_gpu, ok_gpu = env.demonstration_actions()
        a_ora, ok_ora = ora.demo_policy()
        a = a_gpu.cpu().numpy()
        np.testing.assert_array_equal(ok_gpu.cpu().numpy(), ok_ora)
        worst = max(worst, np.abs(a - a_ora).max())
        assert np.abs(a - a_ora).max() < 2e-5, (task, k)   # float32 actions; states track to ~1e-9
        obs_o, rew_o, _, fl_o = ora.step(a.astype(np.float64))  # both sides consume the SAME float32 actions
        env.step(a_gpu)
    obs = env._obs.cpu().numpy()
    if task == "point_mass_reach":
        assert (np.linalg.norm(obs[:, :2] - obs[:, 2:4], axis=1) < 0.06).all()     # within one MAX_STEP_SIZE of the goal
    elif task == "robot_reach":
        assert (np.linalg.norm(obs[:, :3] - obs[:, 3:6], axis=1) < 0.02).all()     # goal_distance_threshold
    else:
        st = env.get_state()
        active = np.array([s.sw_active for s in st])
        # the button got pressed and the arm reached the end pose.  Not every env: with the rigid fingertip stand-in
        #  the touch force chatters while the arm retracts, and a second rising edge toggles the
        # switch off again (switch.py:51-60) -- the oracle shows the same envs oscillating
        assert active.mean() > 0.5, active
        assert env._flags.cpu().numpy()[:, 2].mean() > 0.5
        np.testing.assert_array_equal(env._flags.cpu().numpy(), fl_o)
        np.testing.assert_array_equal(active, ora.field("sw_active"))
    print(task, "max |action gpu - oracle|", worst)
    env.close()


def test_noise_and_error_paths():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.policies import create_demonstration_policy, create_random_policy

    env = ms.make("robot_reach", 64, seed=3)
    env.reset()
    clean = create_demonstration_policy(env)().clone()
    noisy = create_demonstration_policy(env, noise=0.01)().clone()
    d = (noisy - clean).cpu().numpy()
    assert 0.003 < d.std() < 0.03 and abs(d.mean()) < 0.005      # N(0, 0.01) added to the difference
    r = create_random_policy(env)()
    lo, hi = torch.as_tensor(env.single_action_low).cuda(), torch.as_tensor(env.single_action_high).cuda()
    assert ((r >= lo) & (r <= hi)).all()
    env.close()
    push = ms.make("robot_planar_push", 8)
    with pytest.raises(NotImplementedError):
        create_demonstration_policy(push)
    with pytest.raises(Exception, match="no scripted demonstrator"):
        push.demonstration_actions()
    push.close()

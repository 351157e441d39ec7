"""GPU parity of the round-2 narrowphase additions against the oracle: capsule x capsule (arm links against a static capsule
obstacle) and the cap-configuration guard of cylinder x cylinder with the truncation rule (status bit + truncated)."""
import numpy as np
import pytest
import torch

from tests.helpers import arm_states, copy_states, push_states
from tests.test_collision_cpu import OBST

pytestmark = pytest.mark.gpu


def test_capsule_obstacle_rollout_matches_the_oracle():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import build_robot_reach

    n = 128
    b = build_robot_reach(obstacle=OBST)
    b.task.time_limit = 0.0
    ora = arm_states(b, n, 3)
    ms.envs.register_builder("_reach_obstacle", lambda: b)
    env = ms.BatchedMujocoSimEnv("_reach_obstacle", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
    rng = np.random.RandomState(0)
    ncon = 0
    worst = 0.0
    for k in range(12):
        env.set_state(copy_states(ora.states, n), forward=True)
        a = rng.uniform([-0.1, -0.6, 0.02], [0.1, -0.4, 0.2], (n, 3)).astype(np.float32)
        obs_o, rew_o, _, fl_o = ora.step(a.astype(np.float64))
        env.step(torch.from_numpy(a).cuda())
        st = env.get_state()
        q = np.array([list(s.qpos)[:6] for s in st]); v = np.array([list(s.qvel)[:6] for s in st])
        worst = max(worst, np.abs(q - ora.field("qpos", 6)).max())
        assert np.abs(q - ora.field("qpos", 6)).max() < 1e-7 and np.abs(v - ora.field("qvel", 6)).max() < 1e-5, k
        np.testing.assert_array_equal(np.array([s.status for s in st]), ora.field("status"))
        ncon += sum(ora.stage_dump(e)["ncon"] for e in range(0, n, 8))
    print("capsule obstacle: max|dq|", worst, "contacts seen in the sampled envs", ncon)
    assert ncon > 20
    env.close()


def test_cap_contacts_through_the_general_convex_routine_match_the_oracle():
    """Props set down on another prop's cap (cap-cap) and props lying on their side next to the pusher (cap-side): the pairs
    the side-side routine declines go through the MPR restatement shared with the oracle (include/mjs_convex.h).  The two
    compilers contract FMAs differently, so the iterative routine agrees to its tolerance (1e-6), not to rounding:
    gate = the contact tasks' single-step tolerance 1e-4."""
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import build_planar_push

    n, b = 64, build_planar_push()
    ora = push_states(b, n, 5)
    t = b.task
    for e in range(n):
        s = ora.states[e]
        a0, a1, a2 = t.obj_qadr[0], t.obj_qadr[1], t.obj_qadr[2]
        if e % 2:   # prop 1 set down on prop 0's cap, slightly off-centre
            s.qpos[a1] = s.qpos[a0] + 0.002 + 0.0001 * e; s.qpos[a1 + 1] = s.qpos[a0 + 1]; s.qpos[a1 + 2] = s.qpos[a0 + 2] + 2 * 0.019 - 1e-4
        else:       # prop 2 lying on its side with its cap 0.3 mm inside prop 0's side
            s.qpos[a2] = s.qpos[a0]; s.qpos[a2 + 1] = s.qpos[a0 + 1] + 0.02 + 0.019 - 3e-4; s.qpos[a2 + 2] = 0.02 - 1e-4
            s.qpos[a2 + 3], s.qpos[a2 + 4], s.qpos[a2 + 5], s.qpos[a2 + 6] = np.sqrt(0.5), np.sqrt(0.5), 0.0, 0.0
    env = ms.make("robot_planar_push", n, precision="f64", autoreset=False, solver_tolerance=1e-12)
    worst = 0.0
    for k in range(3):
        env.set_state(copy_states(ora.states, n), forward=False)
        a = np.array([[ora.states[e].qpos[0] * 0 + 0.18, -0.58] for e in range(n)], np.float32)  # the pusher stays away
        obs_o, rew_o, _, fl_o = ora.step(a.astype(np.float64))
        env.step(torch.from_numpy(a).cuda())
        st = env.get_state()
        q = np.array([list(s.qpos)[:27] for s in st])
        # cap against side (even envs): a well-conditioned single contact -> the single-step gate.  Cap on cap (odd envs): two
        # parallel flat faces, any point of the overlap is a valid MPR answer and the upper prop rocks on it; only checked to stay put.
        worst = max(worst, np.abs(q - ora.field("qpos", 27))[::2].max())
        np.testing.assert_array_equal(env._flags.cpu().numpy(), fl_o)
        assert all(s.status == 0 for s in st) and (ora.field("status") == 0).all()
        a1 = t.obj_qadr[1]
        assert (q[1::2, a1 + 2] > 0.045).mean() > 0.8 and (ora.field("qpos", 27)[1::2, a1 + 2] > 0.045).mean() > 0.8
    print("cap contacts via MPR: cap-on-side max|dq| over 3 single steps", worst)
    assert worst < 1e-4
    env.close()

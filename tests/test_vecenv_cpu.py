"""Host-side checks of the vector-env facades that need no GPU: spaces and registry."""
import types

import numpy as np

from mujoco_sim_b200 import vecenv
from mujoco_sim_b200.model import tasks as T


def _stub(bundle):
    return types.SimpleNamespace(obs_keys=list(bundle.obs_keys), single_action_low=bundle.action_low, single_action_high=bundle.action_high,
                                 num_envs=4, obs_dim=bundle.task.obs_dim, action_dim=bundle.task.action_dim)


def test_spaces_follow_the_reference_adapter():
    b = T.build_planar_push()
    e = _stub(b)
    a = vecenv.single_action_space(e)
    assert a.shape == (2,) and a.dtype == np.float32 and (a.low == -1).all() and (a.high == 1).all()  # robot_planar_push.py:229-234
    o = vecenv.single_observation_space(e)
    assert list(o.spaces) == ["ur5e/tcp_position", "target_position", "block_positions"]  # entity observables first, then task
    assert [o.spaces[k].shape for k in o.spaces] == [(3,), (2,), (6,)]
    f = vecenv.single_observation_space(e, flatten=True)
    assert f.shape == (11,) and f.dtype == np.float32  # dmc2gym.py:93-98 flattened float32 Box
    b = T.build_pointmass_reach()
    a = vecenv.single_action_space(_stub(b))
    np.testing.assert_allclose(a.high, 0.05)  # point_reach.py:204-209
    assert list(vecenv.single_observation_space(_stub(b)).spaces) == ["pointmass/position", "goal_position"]


def test_registry_names_every_task():
    assert sorted(t for t, _ in vecenv.REGISTRY.values()) == ["point_mass_reach", "robot_planar_push", "robot_push_button", "robot_reach"]
    for k in vecenv.REGISTRY:
        assert k.startswith("mujoco_sim/") and k.endswith("_state-v0")

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
    state = {k: v for k, v in vecenv.REGISTRY.items() if k.endswith("_state-v0")}
    assert sorted(t for t, _ in state.values()) == ["point_mass_reach", "robot_planar_push", "robot_push_button", "robot_reach"]
    # the reference's own two ids (mujoco_sim/__init__.py:26-40), both visual
    assert vecenv.REGISTRY["mujoco_sim/point_mass_reach-v0"] == ("point_mass_reach", {"observation_type": "visual_observations", "image_resolution": 64})
    assert vecenv.REGISTRY["mujoco_sim/robot_push_button_visual-v0"][1]["image_resolution"] == 96
    for k in vecenv.REGISTRY:
        assert k.startswith("mujoco_sim/") and k.endswith("-v0")


def test_visual_observation_spaces():
    b = T.build_pointmass_reach()
    e = _stub(b)
    e.observation_type = "visual_observations"; e._visual_keep = ["pointmass/position"]; e.image_keys = [("Camera/rgb_image", "top_down", (64, 64, 3))]
    o = vecenv.single_observation_space(e)
    assert list(o.spaces) == ["pointmass/position", "Camera/rgb_image"]  # point_reach.py:119-121
    img = o.spaces["Camera/rgb_image"]
    assert img.shape == (64, 64, 3) and img.dtype == np.uint8 and img.high.max() == 255
    import pytest
    with pytest.raises(ValueError):
        vecenv.single_observation_space(e, flatten=True)

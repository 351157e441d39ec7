"""GPU parity of the CUDA ray caster (mjs_render, csrc/render.cuh) against its numpy restatement (oracle/render_ref.py) fed by
the oracle's kinematics: every task, every camera, images of device-reset states.  Float32 on the device against float64 in
numpy: pixels on a silhouette edge may fall on the other side, so the gate is <= 0.5 % of pixels off by more than 1/255."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [("point_mass_reach", "top_down"), ("robot_reach", "front_tilted"), ("robot_planar_push", "front_tilted"), ("robot_planar_push", "top_down"),
         ("robot_push_button", "scene"), ("robot_push_button", "wrist")]


@pytest.mark.parametrize("task,camera", CASES)
def test_render_matches_the_numpy_restatement(task, camera):
    import mujoco_sim_b200 as ms
    from oracle import render_ref as R
    from oracle.oracle import OracleEnvs

    n = 6
    env = ms.make(task, n, seed=12)
    env.reset()
    for _ in range(2):
        env.step(None)
    img = env.render(camera).cpu().numpy()
    res = env.bundle.image_resolution
    assert img.shape == (n, res, res, 3) and img.dtype == np.uint8
    st = env.get_state()
    ora = OracleEnvs(env.bundle, n)
    ctypes.memmove(ora.states, st, ctypes.sizeof(st))
    cam = env.bundle.cameras[camera]
    worst = 0.0
    for e in range(n):
        xpos, xmat = ora.geom_poses(e)
        s = st[e]
        tgt = [s.target[0], s.target[1], 0.0 if task == "point_mass_reach" else s.target[2]]
        vis = R.resolve_visuals(env.bundle, xpos, xmat, tgt, list(s.ovr_pos), bool(s.sw_active))
        cp, cm = R.resolve_camera(cam, xpos, xmat)
        ref = R.render(cp, cm, cam.fovy, res, res, vis)
        off = (np.abs(img[e].astype(int) - ref.astype(int)).max(-1) > 1).mean()
        worst = max(worst, off)
        assert off <= 0.005, (task, camera, e, off)
        assert ref.reshape(-1, 3).max(0).min() > 0 and (ref.sum(-1) > 0).mean() > 0.05  # a real picture, not a black frame
    print(task, camera, "worst fraction of pixels off by > 1/255:", worst)
    if task == "robot_push_button":  # the button turns green while the switch is active (props/switch.py:59)
        for s in st:
            s.sw_active = 1
        env.set_state(st, forward=False)
        img2 = env.render(camera).cpu().numpy()
        assert (img2[..., 1].astype(int) - img[..., 1].astype(int)).max() > 100 or camera == "wrist"
    env.close()


def test_render_argument_checks_and_custom_resolution():
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200 import lib

    env = ms.make("point_mass_reach", 3, seed=1)
    env.reset()
    img = env.render(width=48, height=32)
    assert img.shape == (3, 32, 48, 3)
    with pytest.raises(lib.MjsimError, match="bad image size"):
        env.render(width=5000, height=8)
    env.close()


@pytest.mark.parametrize("task", ["point_mass_reach", "robot_reach", "robot_planar_push", "robot_push_button"])
def test_visual_observation_type_returns_the_cameras_and_the_enabled_state_keys(task):
    """observation_type="visual_observations" (point_reach.py:119-121, robot_reach.py:139-141, robot_planar_push.py:140-142,
    robot_push_button.py:121-124): the observation holds the camera image(s) of the state the step ended in, plus the state
    observables the reference keeps enabled; everything else (reward, flags, state keys' values) equals the state-observation env."""
    import torch

    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.envs import _VISUAL

    n = 6
    ev = ms.make(task, n, seed=11, observation_type="visual_observations")
    es = ms.make(task, n, seed=11)
    ov, _ = ev.reset(); os_, _ = es.reset()
    keep, cams = _VISUAL[task]
    keep = ev._visual_keep
    assert list(ov) == keep + [k for k, _ in cams]
    for _ in range(3):
        ov, rv, tv, trv, iv = ev.step(None); os_, rs, ts, trs, is_ = es.step(None)
        assert torch.equal(rv, rs) and torch.equal(tv, ts) and torch.equal(trv, trs)
        for k in keep:
            assert torch.equal(ov[k], os_[k])
        for key, cam, shape in ev.image_keys:
            img = ov[key]
            assert img.dtype == torch.uint8 and tuple(img.shape) == (n,) + shape
            assert torch.equal(img, es.render(cam, shape[1], shape[0]))      # the frame of the state the step ended in
            assert img.float().std() > 1                                    # not a blank frame
        assert set(iv["final_observation"]) == set(keep)                    # terminal frames are not kept (state keys only)
    ev.close(); es.close()


def test_reference_gym_ids_are_visual_envs():
    """mujoco_sim/__init__.py:26-40 registers two ids, both with visual observations (64 and 96 pixels)."""
    from mujoco_sim_b200 import vecenv

    v = vecenv.make_vec("mujoco_sim/point_mass_reach-v0", 4)
    obs, _ = v.reset(seed=2025)
    assert obs["Camera/rgb_image"].shape == (4, 64, 64, 3) and obs["Camera/rgb_image"].dtype == np.uint8 and obs["pointmass/position"].shape == (4, 2)
    obs2, rew, term, trunc, info = v.step(np.zeros((4, 2), np.float32))
    assert obs2["Camera/rgb_image"].shape == (4, 64, 64, 3) and set(info["final_observation"]) == {"pointmass/position"}
    assert v.single_observation_space.spaces["Camera/rgb_image"].shape == (64, 64, 3)
    obs3, _ = v.reset(seed=2025)
    assert np.array_equal(obs3["Camera/rgb_image"], obs["Camera/rgb_image"])  # test/test_gym_envs.py:21-36 on the image observation
    v.close()
    b = vecenv.make_vec("mujoco_sim/robot_push_button_visual-v0", 2, api="sb3")
    o = b.reset()
    assert o["Camera/rgb_image"].shape == (2, 96, 96, 3) and o["ur5e/Camera/wrist_camera_rgb_image"].shape == (2, 96, 96, 3) and o["ur5e/joint_configuration"].shape == (2, 6)
    b.close()

"""CUDA path (f64, through the C ABI) against the committed traces of tests/golden/ -- no oracle in the loop."""
import numpy as np
import pytest
import torch

from tests.test_golden_cpu import BUILD, load

pytestmark = pytest.mark.gpu

# per-step gates: contact-free tasks at 1e-5 relative (north_star), contact tasks 1e-4 single-step; the f64 kernels sit
# orders of magnitude inside both, so the asserts use what they actually reach
TOL = {"point_mass_reach": 1e-9, "robot_reach": 1e-9, "robot_planar_push": 1e-7, "robot_push_button": 1e-7}


@pytest.mark.parametrize("name", list(BUILD))
def test_cuda_f64_reproduces_the_golden_trace(name):
    import mujoco_sim_b200 as ms

    d, b, states, n = load(name)
    ms.envs.register_builder("_golden_" + name, lambda: b)
    env = ms.BatchedMujocoSimEnv("_golden_" + name, n, precision="f64", autoreset=False, solver_tolerance=1e-12)
    env.set_state(states, forward=False)
    nq, nv = int(d["nq"]), int(d["nv"])
    for k, a in enumerate(d["actions"]):
        obs, rew, term, trunc, info = env.step(torch.from_numpy(a).cuda())
        st = env.get_state()
        q = np.array([list(s.qpos)[:nq] for s in st]); v = np.array([list(s.qvel)[:nv] for s in st])
        assert np.abs(q - d["qpos"][k]).max() < TOL[name], (name, k)
        assert np.abs(v - d["qvel"][k]).max() < TOL[name] * 1e3, (name, k)
        assert np.abs(env._obs.cpu().numpy() - d["obs"][k]).max() < 1e-5   # obs are float32
        assert np.abs(rew.cpu().numpy() - d["reward"][k]).max() < 1e-6
        np.testing.assert_array_equal(env._flags.cpu().numpy(), d["flags"][k])
        if "sw_pressed" in d:
            np.testing.assert_array_equal(np.array([s.sw_pressed for s in st]), d["sw_pressed"][k])
            np.testing.assert_array_equal(np.array([s.sw_active for s in st]), d["sw_active"][k])
    env.close()

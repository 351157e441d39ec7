"""The committed traces in tests/golden/ (tools/make_golden.py; oracle-generated regression fixtures, see the
provenance note there) must be reproduced by the oracle as built now."""
import ctypes
from pathlib import Path

import numpy as np
import pytest

from mujoco_sim_b200.model import tasks as T
from mujoco_sim_b200.model.cstruct import MjsEnvState
from oracle.oracle import OracleEnvs

GOLDEN = Path(__file__).parent / "golden"
BUILD = {"point_mass_reach": T.build_pointmass_reach, "robot_reach": T.build_robot_reach,
         "robot_planar_push": T.build_planar_push, "robot_push_button": T.build_push_button}


def load(name):
    d = np.load(GOLDEN / f"{name}.npz")
    assert int(d["state_bytes"]) == ctypes.sizeof(MjsEnvState), "MjsEnvState layout changed: regenerate tests/golden"
    n = d["actions"].shape[1]
    states = (MjsEnvState * n)()
    ctypes.memmove(states, d["states0"].tobytes(), ctypes.sizeof(states))
    b = BUILD[name]()
    if name in ("robot_reach", "robot_push_button"):
        b.task.time_limit = 0.0
    return d, b, states, n


@pytest.mark.parametrize("name", list(BUILD))
def test_oracle_reproduces_the_golden_trace(name):
    d, b, states, n = load(name)
    ora = OracleEnvs(b, n)
    ctypes.memmove(ora.states, states, ctypes.sizeof(states))
    nq, nv = int(d["nq"]), int(d["nv"])
    for k, a in enumerate(d["actions"]):
        obs, rew, _, fl = ora.step(a.astype(np.float64))
        np.testing.assert_allclose(ora.field("qpos", nq), d["qpos"][k], rtol=0, atol=1e-12)
        np.testing.assert_allclose(ora.field("qvel", nv), d["qvel"][k], rtol=0, atol=1e-10)
        np.testing.assert_allclose(obs, d["obs"][k], rtol=0, atol=1e-12)
        np.testing.assert_allclose(rew, d["reward"][k], rtol=0, atol=1e-12)
        np.testing.assert_array_equal(fl, d["flags"][k])
        if "sw_pressed" in d:
            np.testing.assert_array_equal(ora.field("sw_pressed"), d["sw_pressed"][k])
            np.testing.assert_array_equal(ora.field("sw_active"), d["sw_active"][k])


def test_the_fixtures_exercise_contacts_and_the_switch():
    p = np.load(GOLDEN / "robot_planar_push.npz")
    assert np.abs(p["qvel"][:, :, 6:]).max() > 0.05       # a prop was struck
    s = np.load(GOLDEN / "robot_push_button.npz")
    assert s["sw_pressed"].sum() > 0 and s["sw_active"][-1].sum() > 0
    m = np.load(GOLDEN / "point_mass_reach.npz")
    assert np.abs(m["qpos"]).max() > 0.42                  # an env ran into a wall

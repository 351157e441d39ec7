"""Generate tests/golden/*.npz: small seeded traces of the four tasks.

PROVENANCE: the traces are produced by THIS repo's CPU oracle (oracle/mjs_oracle.cpp), not by MuJoCo /
dm_control -- neither is installable here, so these are regression fixtures ("parity unpinned"):
they freeze the oracle's behaviour so that (a) a change of the oracle is noticed by the CPU suite and (b) the
GPU suite can check the CUDA path against committed numbers.  Regenerate with `python tools/make_golden.py`
after a deliberate change of the physics restatement and say why in the commit.

Each file holds: states0 (raw MjsEnvState bytes of the injected initial states), actions [K, n, A] float32,
and per step qpos/qvel/obs/reward/flags as the oracle produced them (Newton iterated to 1e-12 of MuJoCo's scale).
"""
import ctypes
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from mujoco_sim_b200.model import tasks as T  # noqa: E402
from mujoco_sim_b200.model.cstruct import MjsEnvState  # noqa: E402
from oracle.oracle import OracleEnvs, _dp, lib  # noqa: E402
from tests.helpers import arm_states, pointmass_states, push_states  # noqa: E402


def _raw(states, n):
    return np.frombuffer(ctypes.string_at(ctypes.addressof(states), ctypes.sizeof(MjsEnvState) * n), dtype=np.uint8).copy()


def _trace(ora, n, nq, nv, actions):
    out = {k: [] for k in ("qpos", "qvel", "obs", "reward", "flags", "status")}
    for a in actions:
        obs, rew, _, fl = ora.step(a.astype(np.float64))
        out["qpos"].append(ora.field("qpos", nq)); out["qvel"].append(ora.field("qvel", nv))
        out["obs"].append(obs.copy()); out["reward"].append(rew.copy()); out["flags"].append(fl.copy())
        out["status"].append(ora.field("status"))
    return {k: np.array(v) for k, v in out.items()}


def pointmass(n=8, K=12, seed=21):
    b = T.build_pointmass_reach()
    ora = OracleEnvs(b, n)
    init = pointmass_states(n, seed)
    ctypes.memmove(ora.states, init, ctypes.sizeof(init))
    ora.finish_reset()
    s0 = _raw(ora.states, n)
    rng = np.random.RandomState(seed + 1)
    acts = rng.uniform(-0.05, 0.05, (K, n, 2)).astype(np.float32)
    return dict(states0=s0, actions=acts, nq=2, nv=2, **_trace(ora, n, 2, 2, acts))


def reach(n=6, K=10, seed=22):
    b = T.build_robot_reach(); b.task.time_limit = 0.0
    ora = arm_states(b, n, seed)
    s0 = _raw(ora.states, n)
    rng = np.random.RandomState(seed + 1)
    acts = rng.uniform([-0.1, -0.6, 0.02], [0.1, -0.4, 0.2], (K, n, 3)).astype(np.float32)
    return dict(states0=s0, actions=acts, nq=6, nv=6, **_trace(ora, n, 6, 6, acts))


def push(n=6, K=6, seed=23):
    b = T.build_planar_push()
    ora = push_states(b, n, seed)
    s0 = _raw(ora.states, n)
    rng = np.random.RandomState(seed + 1)
    acts = []
    out = {k: [] for k in ("qpos", "qvel", "obs", "reward", "flags", "status")}
    for k in range(K):  # actions depend on the state: aim the pusher at the first prop on even steps
        a = np.stack([rng.uniform(-0.2, 0.2, n), rng.uniform(-0.6, -0.3, n)], 1)
        if k % 2 == 0:
            a = ora.field("qpos", 27)[:, 6:8] + rng.uniform(-0.01, 0.01, (n, 2))
        a = a.astype(np.float32); acts.append(a)
        for key, v in _trace(ora, n, 27, 24, a[None]).items():
            out[key].append(v[0])
    return dict(states0=s0, actions=np.array(acts), nq=27, nv=24, **{k: np.array(v) for k, v in out.items()})


def button(n=6, K=12, seed=24):
    b = T.build_push_button(); b.task.time_limit = 0.0
    ora = arm_states(b, n, seed)
    s0 = _raw(ora.states, n)
    rng = np.random.RandomState(seed + 1)
    acts = []
    out = {k: [] for k in ("qpos", "qvel", "obs", "reward", "flags", "status", "sw_active", "sw_pressed")}
    for k in range(K):  # descend onto the button through IK'd joint targets
        q_now = ora.field("qpos", 6)
        a = np.zeros((n, 7), np.float32)
        for e in range(n):
            sw = np.array(list(ora.states[e].ovr_pos))
            tcp = sw + np.array([0, 0, 0.07 + 0.015 - 0.005 * k]) + rng.uniform(-0.002, 0.002, 3)
            q = np.zeros(6)
            ok = lib().oracle_tcp_to_joints(ctypes.byref(b.task), _dp(tcp), _dp(q_now[e].copy()), _dp(q))
            a[e, :6] = q if ok else q_now[e]
            a[e, 6] = 0.085
        acts.append(a)
        for key, v in _trace(ora, n, 6, 6, a[None]).items():
            out[key].append(v[0])
        out["sw_active"].append(ora.field("sw_active")); out["sw_pressed"].append(ora.field("sw_pressed"))
    return dict(states0=s0, actions=np.array(acts), nq=6, nv=6, **{k: np.array(v) for k, v in out.items()})


if __name__ == "__main__":
    out = ROOT / "tests" / "golden"
    out.mkdir(exist_ok=True)
    for name, fn in (("point_mass_reach", pointmass), ("robot_reach", reach), ("robot_planar_push", push), ("robot_push_button", button)):
        d = fn()
        d["state_bytes"] = ctypes.sizeof(MjsEnvState)
        np.savez_compressed(out / f"{name}.npz", **d)
        print(name, {k: (v.shape if hasattr(v, "shape") else v) for k, v in d.items()}, (out / f"{name}.npz").stat().st_size, "bytes")

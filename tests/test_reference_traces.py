"""Replay of REAL-reference traces (tests/golden/ref_<task>.npz, written by tools/dump_reference.py where mujoco + dm_control
exist) through the CPU oracle, and through the CUDA library on a GPU.  Skipped only while no ref_*.npz is present -- this
image and the GPU box have neither package, so the repo ships none and parity stays "unpinned" until someone drops them in.

Gates (BASELINE.json north_star): every recorded control step is replayed from its recorded PRE-step state (qpos, qvel,
ctrl, mocap, time, warm start) with the recorded action; |dq| <= 1e-4 single-step for contact tasks, and for the
contact-free tasks additionally a free-running rollout within 1e-5 relative over the first 200 steps of an episode.

The replay machinery itself is exercised on every run by a synthetic trace in the same format generated from the oracle
(test_replay_machinery_on_a_synthetic_trace), so a real trace meets code that is known to work.
"""
import ctypes
from pathlib import Path

import numpy as np
import pytest

from mujoco_sim_b200.model import tasks as T

GOLDEN = Path(__file__).resolve().parent / "golden"
BUILDERS = {"point_mass_reach": T.build_pointmass_reach, "robot_reach": T.build_robot_reach, "robot_planar_push": T.build_planar_push,
            "robot_push_button": T.build_push_button}
CONTACT_FREE = ("point_mass_reach", "robot_reach")


def _inject(bundle, st, tr, k):
    """Write pre-step record k of a trace into an MjsEnvState."""
    m, t = bundle.model, bundle.task
    for i in range(m.nq):
        st.qpos[i] = tr["qpos0"][k][i]
    for i in range(m.nv):
        st.qvel[i] = tr["qvel0"][k][i]; st.qacc_warm[i] = tr["qacc_warm0"][k][i]
    for i in range(min(m.nu, tr["ctrl0"].shape[1])):
        st.ctrl[i] = tr["ctrl0"][k][i]
    if tr["mocap0"].shape[1] >= 3:
        for i in range(3):
            st.mocap_pos[i] = tr["mocap0"][k][i]
    st.time = float(tr["time0"][k])
    tgt = tr["target"][k]
    if bundle.name == "robot_push_button":
        for i in range(3):
            st.ovr_pos[i] = tgt[i]
    else:
        for i in range(len(tgt)):
            st.target[i] = tgt[i]
    st.dist = 1.0; st.prev_dist = 1.0


def replay_single_steps(bundle, tr, stepper):
    """max |dq|, max |dv| over all recorded steps, each replayed from its recorded pre-state.  stepper(states, actions) ->
    (qpos[n, nq], qvel[n, nv]) advances n independent states by one control step."""
    from mujoco_sim_b200.model.cstruct import MjsEnvState

    n = len(tr["action"])
    keep = [k for k in range(n) if not tr["first"][k]]  # a step() right after LAST is the composer's reset, not a physics step
    states = (MjsEnvState * len(keep))()
    for j, k in enumerate(keep):
        _inject(bundle, states[j], tr, k)
    q, v = stepper(states, tr["action"][keep])
    nq, nv = bundle.model.nq, bundle.model.nv
    return np.abs(q - tr["qpos"][keep][:, :nq]).max(), np.abs(v - tr["qvel"][keep][:, :nv]).max()


def oracle_stepper(bundle):
    from oracle.oracle import OracleEnvs

    def step(states, actions):
        n = len(states)
        ora = OracleEnvs(bundle, n)
        ctypes.memmove(ora.states, states, ctypes.sizeof(states))
        # the recorded warm start is the reference's; traj / counters start fresh, like a composer step from this state
        ora.step(actions)
        return ora.field("qpos", bundle.model.nq).reshape(n, -1), ora.field("qvel", bundle.model.nv).reshape(n, -1)

    return step


def _load(task):
    p = GOLDEN / f"ref_{task}.npz"
    if not p.exists():
        pytest.skip(f"{p.name} absent: run tools/dump_reference.py where mujoco + dm_control are installed (neither is in this image)")
    return dict(np.load(p, allow_pickle=False))


def _check_shapes(bundle, tr):
    if tr["qpos"].shape[1] != bundle.model.nq:
        pytest.skip(f"trace has nq={tr['qpos'].shape[1]}, this build models nq={bundle.model.nq} (rigid gripper stand-in, SURVEY.md 8f-2)")


@pytest.mark.parametrize("task", list(BUILDERS))
def test_oracle_single_steps_match_the_reference_trace(task):
    tr = _load(task)
    b = BUILDERS[task]()
    _check_shapes(b, tr)
    dq, dv = replay_single_steps(b, tr, oracle_stepper(b))
    tol = 1e-5 if task in CONTACT_FREE else 1e-4
    assert dq < tol and dv < 100 * tol, (dq, dv)


@pytest.mark.parametrize("task", CONTACT_FREE)
def test_oracle_rollout_matches_the_reference_trace(task):
    from oracle.oracle import OracleEnvs

    tr = _load(task)
    b = BUILDERS[task]()
    _check_shapes(b, tr)
    ora = OracleEnvs(b, 1)
    _inject(b, ora.states[0], tr, 0)
    n = len(tr["action"])
    end = next((k for k in range(1, n) if tr["first"][k]), n)  # first episode only
    for k in range(min(end, 200)):
        ora.step(tr["action"][k][None])
        q = ora.field("qpos", b.model.nq)[0]
        rel = np.abs(q - tr["qpos"][k]) / np.maximum(np.abs(tr["qpos"][k]), 1e-3)
        assert rel.max() < 1e-5, (k, rel.max())


@pytest.mark.gpu
@pytest.mark.parametrize("task", list(BUILDERS))
def test_cuda_single_steps_match_the_reference_trace(task):
    import torch

    import mujoco_sim_b200 as ms

    tr = _load(task)
    b = BUILDERS[task]()
    _check_shapes(b, tr)

    def step(states, actions):
        n = len(states)
        env = ms.make(task, n, precision="f64", autoreset=False)
        env.set_state(states, forward=False)
        env.step(torch.from_numpy(np.asarray(actions, np.float32)).cuda())
        st = env.get_state()
        q = np.array([list(s.qpos)[: b.model.nq] for s in st]); v = np.array([list(s.qvel)[: b.model.nv] for s in st])
        env.close()
        return q, v

    dq, dv = replay_single_steps(b, tr, step)
    tol = 1e-5 if task in CONTACT_FREE else 1e-4
    assert dq < tol and dv < 100 * tol, (dq, dv)


def synthetic_trace(task, steps, seed):
    """A trace in dump_reference.py's format generated by the ORACLE (exercises the replay code; pins nothing)."""
    from oracle.oracle import OracleEnvs
    from tests.helpers import arm_states, pointmass_states, push_states

    b = BUILDERS[task]()
    if task == "point_mass_reach":
        ora = OracleEnvs(b, 1)
        init = pointmass_states(1, seed)
        ctypes.memmove(ora.states, init, ctypes.sizeof(init))
        ora.finish_reset()
    elif task == "robot_planar_push":
        ora = push_states(b, 1, seed)
    else:
        ora = arm_states(b, 1, seed)
    m, t = b.model, b.task
    rng = np.random.RandomState(seed)
    rec = {k: [] for k in ("qpos0", "qvel0", "ctrl0", "mocap0", "time0", "qacc_warm0", "action", "qpos", "qvel", "ctrl", "mocap", "time", "target", "first")}
    s = ora.states[0]

    def snap(suffix):
        rec["qpos" + suffix].append([s.qpos[i] for i in range(m.nq)]); rec["qvel" + suffix].append([s.qvel[i] for i in range(m.nv)])
        rec["ctrl" + suffix].append([s.ctrl[i] for i in range(max(m.nu, 1))]); rec["mocap" + suffix].append([s.mocap_pos[i] for i in range(3)])
        rec["time" + suffix].append(s.time)

    for k in range(steps):
        snap("0")
        rec["qacc_warm0"].append([s.qacc_warm[i] for i in range(m.nv)])
        lo = np.array([t.action_lo[i] for i in range(t.action_dim)]); hi = np.array([t.action_hi[i] for i in range(t.action_dim)])
        a = rng.uniform(lo, hi)
        if task == "robot_push_button":
            a[:6] = np.clip(np.array([s.qpos[i] for i in range(6)]) + a[:6], -3.14, 3.14)
        a = a.astype(np.float32).astype(np.float64)
        rec["target"].append([s.ovr_pos[i] for i in range(3)] if task == "robot_push_button" else [s.target[i] for i in range(3)])
        rec["first"].append(False)
        ora.step(a[None])
        rec["action"].append(a)
        snap("")
    return b, {k: np.array(v) for k, v in rec.items()}


@pytest.mark.parametrize("task", list(BUILDERS))
def test_replay_machinery_on_a_synthetic_trace(task):
    b, tr = synthetic_trace(task, 4, 3)
    dq, dv = replay_single_steps(b, tr, oracle_stepper(b))
    assert dq < 1e-9 and dv < 1e-7, (dq, dv)

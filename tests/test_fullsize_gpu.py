"""Size-independent properties at BASELINE.json's full env counts (65 536 / 16 384 / 4096 / 4096 envs per GPU), where
the CPU oracle cannot follow every env:

* shard independence: env g of the full-size simulator is bit-identical to env g of a small simulator created with the
  same seed and `env_offset` (Philox streams are keyed by the global env id) -- the property the multi-GPU partition
  relies on, checked on one GPU for the first and the last envs of the batch;
* a sample of envs pulled out of the full-size batch (first / middle / last) steps like the oracle from the same states;
* invariants over the whole batch: finite state, unit prop quaternions, no prop below the table top, status word clean.
"""
import numpy as np
import pytest
import torch

from tests.helpers import copy_states

pytestmark = pytest.mark.gpu

FULL = [("point_mass_reach", 65536, 2, 2), ("robot_reach", 16384, 6, 6), ("robot_planar_push", 4096, 27, 24), ("robot_push_button", 4096, 6, 6)]


def _make(task, n, offset=0, seed=11):
    import mujoco_sim_b200 as ms

    return ms.make(task, n, precision="f64", seed=seed, env_offset=offset, autoreset=True)


def _qv(env, first, count, nq, nv):
    st = env.get_state(first, count)
    return (np.array([list(s.qpos)[:nq] for s in st]), np.array([list(s.qvel)[:nv] for s in st]),
            np.array([[s.episode_step, s.status, s.episode_id] for s in st]))


@pytest.mark.parametrize("task,n_full,nq,nv", FULL)
def test_full_size_batch_equals_small_shards_bit_for_bit(task, n_full, nq, nv):
    head, tail, steps = 96, 64, 3
    full = _make(task, n_full)
    a = _make(task, head, offset=0)
    b = _make(task, tail, offset=n_full - tail)
    for env in (full, a, b):
        env.reset()
    for _ in range(steps):  # device-drawn actions: Philox keyed by (seed, global env id, step index)
        for env in (full, a, b):
            env.step_flat(None)
    torch.cuda.synchronize()
    for small, first, cnt in ((a, 0, head), (b, n_full - tail, tail)):
        qf, vf, sf = _qv(full, first, cnt, nq, nv)
        qs, vs, ss = _qv(small, 0, cnt, nq, nv)
        np.testing.assert_array_equal(qf, qs)
        np.testing.assert_array_equal(vf, vs)
        np.testing.assert_array_equal(sf, ss)
        np.testing.assert_array_equal(full._obs[first:first + cnt].cpu().numpy(), small._obs.cpu().numpy())
        np.testing.assert_array_equal(full._reward[first:first + cnt].cpu().numpy(), small._reward.cpu().numpy())
    # invariants over the whole batch
    obs = full._obs.cpu().numpy()
    assert np.isfinite(obs).all() and np.isfinite(full._reward.cpu().numpy()).all()
    st = full.stats(reset=False)
    assert st["errors"] == 0.0
    if task == "robot_planar_push":
        q, v, s = _qv(full, 0, 1024, nq, nv)
        assert (s[:, 1] == 0).all()
        for k in range(3):
            quat = q[:, 6 + 7 * k + 3:6 + 7 * k + 7]
            assert np.abs(np.linalg.norm(quat, axis=1) - 1.0).max() < 1e-12
            assert q[:, 6 + 7 * k + 2].min() > 0.015  # no prop sinks into the table (rest height 0.0189: half-height minus the soft-contact penetration)
    for env in (full, a, b):
        env.close()


def test_sample_of_the_4096_env_push_batch_steps_like_the_oracle():
    """Planar Push at the BASELINE size: 3 x 16 envs taken from the device batch (after a device reset and two
    free-running steps) and the oracle advance one control step from the same states with the same actions."""
    import mujoco_sim_b200 as ms
    from mujoco_sim_b200.model.tasks import build_planar_push
    from oracle.oracle import OracleEnvs

    n, cnt = 4096, 16
    bundle = build_planar_push()
    env = ms.make("robot_planar_push", n, precision="f64", seed=5, autoreset=False, solver_tolerance=1e-12)
    env.reset()
    rng = np.random.RandomState(3)
    for _ in range(2):
        env.step_flat(torch.from_numpy(np.stack([rng.uniform(-0.2, 0.2, n), rng.uniform(-0.6, -0.3, n)], 1).astype(np.float32)).cuda())
    act = np.stack([rng.uniform(-0.2, 0.2, n), rng.uniform(-0.6, -0.3, n)], 1).astype(np.float32)
    firsts = (0, n // 2 - cnt // 2, n - cnt)
    oras = []
    for f in firsts:
        ora = OracleEnvs(bundle, cnt)
        st = env.get_state(f, cnt)
        import ctypes
        ctypes.memmove(ora.states, st, ctypes.sizeof(st))
        oras.append(ora)
    env.step_flat(torch.from_numpy(act).cuda())
    worst = 0.0
    for f, ora in zip(firsts, oras):
        obs_o, rew_o, _, fl_o = ora.step(act[f:f + cnt].astype(np.float64))
        q, v, _ = _qv(env, f, cnt, 27, 24)
        worst = max(worst, np.abs(q - ora.field("qpos", 27)).max())
        assert np.abs(q - ora.field("qpos", 27)).max() < 1e-6
        assert np.abs(v - ora.field("qvel", 24)).max() < 1e-4
        assert np.abs(env._reward[f:f + cnt].cpu().numpy() - rew_o).max() < 1e-6
        np.testing.assert_array_equal(env._flags[f:f + cnt].cpu().numpy(), fl_o)
    print("4096-env push batch, 48 sampled envs, one step vs oracle: max |dq|", worst)
    env.close()

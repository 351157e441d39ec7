"""Dump golden traces of the REAL reference tasks (SURVEY.md 8c plan).

Run this wherever `mujoco`, `dm_control`, `gymnasium`, `ur_analytic_ik` and `robot_descriptions` are installed together with
the reference checkout (they are absent from this image and from the GPU box, so the committed repo holds no ref_*.npz):

    PYTHONPATH=/root/reference python tools/dump_reference.py --out tests/golden [--tasks point_mass_reach,...] [--steps 200]

For every task it builds the reference environment exactly as mujoco_sim/__init__.py:19-23 does
(`Environment(task, strip_singleton_obs_buffer_dim=True, time_limit=...)` wrapped in `DMCEnvironmentAdapter`, state
observations), resets with a fixed seed and steps with seeded actions, recording per control step the full MuJoCo state
around `DMCEnvironmentAdapter.step` (dmc2gym.py:133-155):

    qpos0/qvel0/ctrl0/mocap0/time0/qacc_warm0  state before the step        action   the action passed to step()
    qpos/qvel/ctrl/mocap/time                  state after the step         obs      observation dict flattened in key order
    reward, terminated, truncated, is_success, discount
    target                                     task target (site / goal position) of the episode

tests/test_reference_traces.py replays these through the CPU oracle (always) and the CUDA library (on a GPU) from the
recorded pre-step states and compares within BASELINE.json's tolerances.  The test is skipped only while no ref_*.npz exists.

Planar Push: the reference task crashes as written (robot_planar_push.py:173 `sample()` without random_state; :241
`self.current_step`); the two one-line fixes below are applied by monkey-patching, and GoogleBlockProp.sample_random_object
is replaced by the CylinderProp of SURVEY.md D2 (r 0.02, half-height 0.019, mass 0.1, condim 4, friction [1, 0.05, 0]) so the
dumped task is the configuration BASELINE.json measures.  `--mesh-props` keeps the reference's own mesh blocks instead.
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import numpy as np


def _flatten_obs(obs):
    if isinstance(obs, dict):
        return np.concatenate([np.asarray(obs[k], dtype=np.float64).ravel() for k in obs])
    return np.asarray(obs, dtype=np.float64).ravel()


def _physics_snapshot(physics):
    d = physics.data
    return dict(qpos=np.array(d.qpos), qvel=np.array(d.qvel), ctrl=np.array(d.ctrl), mocap=np.array(d.mocap_pos).ravel(),
                time=float(d.time), qacc_warm=np.array(d.qacc_warmstart))


def _patch_planar_push(mesh_props: bool):
    from dm_control import composer, mjcf

    from mujoco_sim.entities.props import google_block
    from mujoco_sim.environments.tasks import robot_planar_push as rpp

    def randomize_object_position(self, physics):  # robot_planar_push.py:169-176 with the missing random_state supplied
        colliding = True
        while colliding:
            for object_joint in self.object_joints:
                physics.named.data.qpos[object_joint.full_identifier][:3] = self.object_spawn_space.sample(self._dump_random_state)
                physics.named.data.qpos[object_joint.full_identifier][3:] = np.array([1, 0, 0, 0])
            physics.forward()
            colliding = physics.data.ncon > 0

    orig_init_episode = rpp.RobotPushTask.initialize_episode

    def initialize_episode(self, physics, random_state):
        self._dump_random_state = random_state
        return orig_init_episode(self, physics, random_state)

    def should_terminate_episode(self, physics):  # :240-241 with current_step -> episode_step
        return self.is_task_accomplished(physics) or self.episode_step >= self.config.max_control_steps_per_episode

    rpp.RobotPushTask.randomize_object_position = randomize_object_position
    rpp.RobotPushTask.initialize_episode = initialize_episode
    rpp.RobotPushTask.should_terminate_episode = should_terminate_episode
    if mesh_props:
        return

    class CylinderProp(composer.Entity):
        def _build(self):
            self._model = mjcf.RootElement(model="cylinder_prop")
            self.object_body = self._model.worldbody.add("body", name="model")
            self.object_body.add("geom", type="cylinder", size=[0.02, 0.019], mass=0.1, condim=4, friction=[1.0, 0.05, 0.0], rgba=[1, 0, 0, 1])

        @property
        def mjcf_model(self):
            return self._model

        def get_position(self, physics):
            return physics.bind(self.object_body).xpos

    google_block.GoogleBlockProp.sample_random_object = classmethod(lambda cls, *a, **k: CylinderProp())
    rpp.GoogleBlockProp = google_block.GoogleBlockProp


def _make_env(task_name: str, mesh_props: bool):
    from dm_control.composer import Environment

    from mujoco_sim.environments.dmc2gym import DMCEnvironmentAdapter

    if task_name == "point_mass_reach":
        from mujoco_sim.environments.tasks.point_reach import STATE_OBS, PointMassReachTask

        task = PointMassReachTask(observation_type=STATE_OBS)
        limit = 50 * task.CONTROL_TIMESTEP
    elif task_name == "robot_reach":
        from mujoco_sim.environments.tasks.robot_reach import RobotReachConfig, RobotReachTask

        task = RobotReachTask(RobotReachConfig(observation_type=RobotReachConfig.STATE_OBS))
        limit = 100 * task.config.control_timestep
    elif task_name == "robot_planar_push":
        _patch_planar_push(mesh_props)
        from mujoco_sim.environments.tasks.robot_planar_push import RobotPushConfig, RobotPushTask

        task = RobotPushTask(RobotPushConfig(observation_type=RobotPushConfig.STATE_OBS, n_objects=3))
        limit = float("inf")
    elif task_name == "robot_push_button":
        from mujoco_sim.environments.tasks.robot_push_button import RobotPushButtonTask

        task = RobotPushButtonTask(observation_type=RobotPushButtonTask.STATE_OBS, action_type=RobotPushButtonTask.ABS_JOINT_ACTION)
        limit = 100 * 0.1
    else:
        raise ValueError(task_name)
    env = Environment(task, strip_singleton_obs_buffer_dim=True, time_limit=limit)
    return DMCEnvironmentAdapter(env, flatten_observation_space=False), task


def _action(task_name, rng, physics, adapter):
    """Seeded actions inside the region the synthetic benchmark draws from."""
    if task_name == "point_mass_reach":
        return rng.uniform(-0.05, 0.05, 2).astype(np.float32)
    if task_name == "robot_reach":
        return rng.uniform([-0.1, -0.6, 0.02], [0.1, -0.4, 0.2])
    if task_name == "robot_planar_push":
        return rng.uniform([-0.2, -0.6], [0.2, -0.3]).astype(np.float32)
    q = np.array(physics.data.qpos[:6])
    return np.concatenate([np.clip(q + rng.uniform(-0.05, 0.05, 6), -3.14, 3.14), [rng.uniform(0.0, 0.085)]])


def _target(task_name, task, physics):
    if task_name == "point_mass_reach":
        return np.array(physics.bind(task.target).pos)
    if task_name in ("robot_reach", "robot_planar_push"):
        return np.array(physics.bind(task.target).pos)
    return np.array(task.switch.get_position(physics))


def dump(task_name: str, out_dir: Path, steps: int, seed: int, mesh_props: bool):
    import mujoco

    adapter, task = _make_env(task_name, mesh_props)
    rng = np.random.RandomState(seed)
    rec = {k: [] for k in ("qpos0", "qvel0", "ctrl0", "mocap0", "time0", "qacc_warm0", "action", "qpos", "qvel", "ctrl", "mocap", "time", "obs",
                           "reward", "terminated", "truncated", "is_success", "discount", "target", "first")}
    obs, _ = adapter.reset(seed=seed)
    first = True
    obs_reset = _flatten_obs(obs)
    for _ in range(steps):
        physics = adapter.dmc_env.physics
        before = _physics_snapshot(physics)
        a = _action(task_name, rng, physics, adapter)
        obs, reward, term, trunc, info = adapter.step(a)
        physics = adapter.dmc_env.physics
        after = _physics_snapshot(physics)
        for k in ("qpos", "qvel", "ctrl", "mocap", "time", "qacc_warm"):
            rec[k + "0"].append(before[k])
        for k in ("qpos", "qvel", "ctrl", "mocap", "time"):
            rec[k].append(after[k])
        rec["action"].append(np.asarray(a, np.float64)); rec["obs"].append(_flatten_obs(obs))
        rec["reward"].append(0.0 if reward is None else float(reward)); rec["terminated"].append(bool(term)); rec["truncated"].append(bool(trunc))
        rec["is_success"].append(float(info.get("is_success", 0.0))); rec["discount"].append(float(info.get("discount", 1.0) or 0.0))
        rec["target"].append(_target(task_name, task, physics)); rec["first"].append(first)
        first = False
        if term or trunc:  # the composer resets inside the next step() (FIRST time step, reward None)
            first = True
    out = {k: np.array(v) for k, v in rec.items()}
    out["obs_reset"] = obs_reset
    out["meta"] = np.array([f"task={task_name}", f"seed={seed}", f"mujoco={mujoco.__version__}", f"props={'mesh' if mesh_props else 'cylinder'}",
                            "generator=tools/dump_reference.py"])
    out_dir.mkdir(parents=True, exist_ok=True)
    path = out_dir / f"ref_{task_name}.npz"
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {steps} control steps, nq={out['qpos'].shape[1]}, nv={out['qvel'].shape[1]}")


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--out", default="tests/golden")
    ap.add_argument("--tasks", default="point_mass_reach,robot_reach,robot_planar_push,robot_push_button")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--seed", type=int, default=2025)
    ap.add_argument("--mesh-props", action="store_true")
    args = ap.parse_args()
    try:
        import dm_control  # noqa: F401
        import mujoco  # noqa: F401
    except ImportError as e:
        sys.exit(f"dump_reference.py needs the reference's dependencies (mujoco, dm_control, ...): {e}")
    for t in args.tasks.split(","):
        dump(t.strip(), Path(args.out), args.steps, args.seed, args.mesh_props)


if __name__ == "__main__":
    main()

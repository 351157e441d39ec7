"""Scene + task-constant builders for the four composer tasks.

Each builder authors the same kinematic tree, geoms and constants the
reference's task constructor produces through dm_control.mjcf (citations per
function), compiles it with compiler.compile_scene and fills an MjsTask.
"""
from __future__ import annotations

import dataclasses
import math

import numpy as np

from . import scene as S
from .compiler import Compiled, compile_scene
from .cstruct import MACROS, MjsTask

# reward types (point_reach.py:11-14, robot_planar_push.py:33-34)
SPARSE_REWARD = "sparse_reward"
DENSE_POTENTIAL_REWARD = "dense_potential_reward"
DENSE_NEG_DISTANCE_REWARD = "dense_negative_distance_reward"
DENSE_BIASED_NEG_DISTANCE_REWARD = "dense_biased_negative_distance_reward"
REWARD_CODE = {
    SPARSE_REWARD: 0,
    DENSE_NEG_DISTANCE_REWARD: 1,
    DENSE_BIASED_NEG_DISTANCE_REWARD: 2,
    DENSE_POTENTIAL_REWARD: 3,
}

# true UR5e DH parameters used by ur_analytic_ik
UR5E_DH = (0.1625, -0.425, -0.3922, 0.1333, 0.0997, 0.0996)


@dataclasses.dataclass
class TaskBundle:
    name: str
    compiled: Compiled
    task: MjsTask
    obs_keys: list  # [(key, dim)] in the reference's dict order
    action_low: np.ndarray
    action_high: np.ndarray
    action_dtype: type

    @property
    def model(self):
        return self.compiled.model


def _accumulated_step_limit(control_dt: float, physics_dt: float, time_limit: float) -> int:
    """Number of control steps until composer's `physics.time() >= time_limit`
    fires, accumulating time in double exactly like mj_step does."""
    n_sub = int(round(control_dt / physics_dt))
    t, steps = 0.0, 0
    while True:
        for _ in range(n_sub):
            t += physics_dt
        steps += 1
        if t >= time_limit:
            return steps


# =====================================================================  PointMass
def build_pointmass_reach(reward_type: str = DENSE_BIASED_NEG_DISTANCE_REWARD, max_steps: int = 50) -> TaskBundle:
    """PointMassReachTask scene (point_reach.py:61-113, pointmass.py:45-77,
    mjcf/walled_pointmass_arena.xml:15-19, entities/utils.py:32-41).
    The Camera entity's two static geoms (camera.py:76-91) are dropped: they
    sit at z=2.4 and never collide."""
    r = 0.05  # PointMass2D(radius=0.05) point_reach.py:80
    world = S.Body("world")
    # wall planes: zaxis -> quat rotating +z onto the axis
    def zaxis_quat(ax):
        ax = np.asarray(ax, float)
        z = np.array([0.0, 0.0, 1.0])
        c = np.cross(z, ax)
        s = np.linalg.norm(c)
        ang = math.atan2(s, float(z @ ax))
        if s < 1e-12:
            return (1.0, 0, 0, 0)
        c = c / s
        return (math.cos(ang / 2), *(math.sin(ang / 2) * c))

    world.geoms.append(S.Geom(S.PLANE, (0.5, 0.5, 0.1), name="ground"))
    world.geoms.append(S.Geom(S.PLANE, (0.02, 0.5, 0.02), pos=(-0.5, 0, 0.02), quat=zaxis_quat((1, 0, 0)), name="wall_x"))
    world.geoms.append(S.Geom(S.PLANE, (0.5, 0.02, 0.02), pos=(0, -0.5, 0.02), quat=zaxis_quat((0, 1, 0)), name="wall_y"))
    world.geoms.append(S.Geom(S.PLANE, (0.02, 0.5, 0.02), pos=(0.5, 0, 0.02), quat=zaxis_quat((-1, 0, 0)), name="wall_neg_x"))
    world.geoms.append(S.Geom(S.PLANE, (0.5, 0.02, 0.02), pos=(0, 0.5, 0.02), quat=zaxis_quat((0, -1, 0)), name="wall_neg_y"))
    world.add(S.Body("pointmass_mocap", mocap=True))
    pm = world.add(S.Body("pointmass/pointmass_body", pos=(0, 0, r)))
    pm.geoms.append(S.Geom(S.SPHERE, (r,), mass=0.1, name="pointmass/sphere"))
    pm.joints.append(S.Joint(S.SLIDE, axis=(1, 0, 0), name="pointmass/pointmass_x"))
    pm.joints.append(S.Joint(S.SLIDE, axis=(0, 1, 0), name="pointmass/pointmass_y"))
    sc = S.Scene(world, welds=[S.Weld("pointmass_mocap", "pointmass/pointmass_body")], timestep=0.02)
    comp = compile_scene(sc)

    t = MjsTask()
    t.kind = MACROS["MJS_TASK_POINTMASS"]
    control_dt, physics_dt = 0.1, 0.02
    t.n_sub = int(round(control_dt / physics_dt))
    t.reward_type = REWARD_CODE[reward_type]
    t.max_steps = 0
    t.action_dim = 2
    t.obs_dim = 4
    t.pm_body = comp.body_id["pointmass/pointmass_body"]
    t.pm_mocap = 0
    t.control_dt = control_dt
    t.time_limit = max_steps * control_dt  # mujoco_sim/__init__.py:21
    t.goal_threshold = 0.02
    t.set("arena_range", (-0.5, 0.5))
    t.set("spawn_lo", (-0.5 + r, -0.5 + r, 0))
    t.set("spawn_hi", (0.5 - r, 0.5 - r, 0))
    t.set("target_lo", (-0.5 + r, -0.5 + r, r / 2))
    t.set("target_hi", (0.5 - r, 0.5 - r, r / 2))
    t.set("action_lo", [-0.05, -0.05])
    t.set("action_hi", [0.05, 0.05])
    return TaskBundle(
        "point_mass_reach",
        comp,
        t,
        [("pointmass/position", 2), ("goal_position", 2)],
        np.array([-0.05, -0.05], np.float32),
        np.array([0.05, 0.05], np.float32),
        np.float32,
    )

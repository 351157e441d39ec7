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
from .cstruct import MACROS, MjsCamera, MjsTask, MjsVisual

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
    # camera observations: visible primitives and the task's fixed cameras, consumed by mjs_render
    visuals: list = dataclasses.field(default_factory=list)
    cameras: dict = dataclasses.field(default_factory=dict)
    image_resolution: int = 64

    @property
    def model(self):
        return self.compiled.model


def _camera(pos, quat, fovy, geom=-1):
    """entities/camera.py:86-88: fixed camera; quaternions in the reference's configs are not unit length (normalised by
    MuJoCo's compiler, here by mjs_render)."""
    c = MjsCamera()
    c.anchor = MACROS["MJS_ANCHOR_GEOMFRAME"] if geom >= 0 else MACROS["MJS_ANCHOR_WORLD"]
    c.geom = geom
    c.set("pos", pos); c.set("quat", quat)
    c.fovy = float(fovy)
    return c


def _qmul(a, b):
    return np.array([a[0]*b[0]-a[1]*b[1]-a[2]*b[2]-a[3]*b[3], a[0]*b[1]+a[1]*b[0]+a[2]*b[3]-a[3]*b[2],
                     a[0]*b[2]-a[1]*b[3]+a[2]*b[0]+a[3]*b[1], a[0]*b[3]+a[1]*b[2]-a[2]*b[1]+a[3]*b[0]])


def _qrot(q, v):
    w, x, y, z = q
    R = np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)], [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])
    return R @ np.asarray(v, float)


def _in_geom_frame(gp, gq, fp, fq, cp, cq):
    """pose (cp, cq), given in a frame (fp, fq) of a body, re-expressed in the frame of that body's geom (gp, gq)"""
    gq, fq, cq = (np.asarray(q, float) / np.linalg.norm(q) for q in (gq, fq, cq))
    p = np.asarray(fp, float) + _qrot(fq, cp)
    q = _qmul(fq, cq)
    gqi = gq * np.array([1, -1, -1, -1])
    return _qrot(gqi, p - np.asarray(gp, float)), _qmul(gqi, q)


def _geom_visuals(comp: Compiled, skip=(), active=None):
    """one visual per compiled model geom, coloured by its rgba; `active` = {geom name: rgba while the switch is active}"""
    out = []
    names = {v: k for k, v in comp.geom_id.items()}
    m = comp.model
    for g in range(m.ngeom):
        if names.get(g) in skip:
            continue
        v = MjsVisual()
        v.type = m.geom_type[g]; v.geom = g; v.anchor = MACROS["MJS_ANCHOR_GEOM"]
        v.set("quat", (1, 0, 0, 0))
        v.set("size", [m.geom_size[3 * g + k] for k in range(3)])
        v.set("rgba", comp.geom_rgba[g]); v.set("rgba_active", comp.geom_rgba[g])
        if active and names.get(g) in active:
            v.use_active = 1
            v.set("rgba_active", active[names[g]])
        out.append(v)
    return out


def _marker(gtype, size, rgba, anchor, pos=(0, 0, 0)):
    v = MjsVisual()
    v.type = gtype; v.geom = -1; v.anchor = MACROS[anchor]
    v.set("pos", pos); v.set("quat", (1, 0, 0, 0))
    v.set("size", list(size) + [0.0] * (3 - len(size)))
    v.set("rgba", rgba); v.set("rgba_active", rgba)
    return v


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

    world.geoms.append(S.Geom(S.PLANE, (0.5, 0.5, 0.1), name="ground", rgba=(0.15, 0.25, 0.35, 1)))  # mean of the checker texture
    world.geoms.append(S.Geom(S.PLANE, (0.02, 0.5, 0.02), pos=(-0.5, 0, 0.02), quat=zaxis_quat((1, 0, 0)), name="wall_x", rgba=(0.3, 0.5, 0.7, 1)))
    world.geoms.append(S.Geom(S.PLANE, (0.5, 0.02, 0.02), pos=(0, -0.5, 0.02), quat=zaxis_quat((0, 1, 0)), name="wall_y", rgba=(0.3, 0.5, 0.7, 1)))
    world.geoms.append(S.Geom(S.PLANE, (0.02, 0.5, 0.02), pos=(0.5, 0, 0.02), quat=zaxis_quat((-1, 0, 0)), name="wall_neg_x", rgba=(0.3, 0.5, 0.7, 1)))
    world.geoms.append(S.Geom(S.PLANE, (0.5, 0.02, 0.02), pos=(0, 0.5, 0.02), quat=zaxis_quat((0, -1, 0)), name="wall_neg_y", rgba=(0.3, 0.5, 0.7, 1)))
    world.add(S.Body("pointmass_mocap", mocap=True))
    pm = world.add(S.Body("pointmass/pointmass_body", pos=(0, 0, r)))
    pm.geoms.append(S.Geom(S.SPHERE, (r,), mass=0.1, name="pointmass/sphere", rgba=(1, 0, 0, 0.5)))  # pointmass.py:55
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
    vis = _geom_visuals(comp) + [_marker(S.BOX, (0.04, 0.04, 0.04), (0, 1, 0, 1), "MJS_ANCHOR_TARGET", (0, 0, r / 2))]  # point_reach.py:91-93,136
    return TaskBundle(
        "point_mass_reach",
        comp,
        t,
        [("pointmass/position", 2), ("goal_position", 2)],
        np.array([-0.05, -0.05], np.float32),
        np.array([0.05, 0.05], np.float32),
        np.float32,
        visuals=vis,
        cameras={"top_down": _camera((0, 0, 2.4), (1, 0, 0, 0), 30)},  # point_reach.py:22
        image_resolution=64,
    )


# =====================================================================  UR5e arm
# Collision filtering bits (DESIGN.md "collision scope"): the arm's link geoms
# only collide with the ground plane; pusher / fingertip / props collide with
# each other and the plane.  MuJoCo's defaults would also test arm x prop and
# non-adjacent arm x arm pairs, which never touch in these workspaces.
_BIT_WORLD = 1
_BIT_ARM = 2
_BIT_OBST = 4  # test obstacles (capsules) that only the arm's link CAPSULES meet


def _ur5e(world: S.Body):
    """UR5e kinematic tree restated from mujoco_menagerie universal_robots_ur5e/ur5e.xml
    [UPSTREAM-RECALL, file not in this image; SURVEY.md App. A.2] with the reference's
    edits: gravcomp=1 on every body (robot.py:81-82) and the base quaternion
    overwritten to (0,0,0,-1) (robot.py:320).  Returns (wrist_3 body, actuators)."""
    J = dict(armature=0.1)
    full = (-6.28319, 6.28319)
    col = dict(contype=_BIT_ARM, conaffinity=0, rgba=(0.7, 0.7, 0.72, 1))  # the collision primitives stand in for the visual meshes

    def cap(size, pos, quat=(1, 0, 0, 0), name=""):
        return S.Geom(S.CAPSULE, size, pos=pos, quat=quat, name=name, **{**col, "contype": _BIT_ARM | _BIT_OBST})

    base = world.add(S.Body("ur5e/base", quat=(0, 0, 0, -1), gravcomp=1.0,
                            inertial=S.Inertial(4.0, diaginertia=(0.00443333156, 0.00443333156, 0.0072))))
    sh = base.add(S.Body("ur5e/shoulder_link", pos=(0, 0, 0.163), gravcomp=1.0,
                         inertial=S.Inertial(3.7, diaginertia=(0.0102675, 0.0102675, 0.00666))))
    sh.joints.append(S.Joint(S.HINGE, axis=(0, 0, 1), range=full, name="ur5e/shoulder_pan_joint", **J))
    sh.geoms.append(cap((0.06, 0.06), (0, 0, -0.04), name="ur5e/shoulder_col"))
    ua = sh.add(S.Body("ur5e/upper_arm_link", pos=(0, 0.138, 0), quat=(1, 0, 1, 0), gravcomp=1.0,
                       inertial=S.Inertial(8.393, pos=(0, 0, 0.2125), diaginertia=(0.133886, 0.133886, 0.0151074))))
    ua.joints.append(S.Joint(S.HINGE, axis=(0, 1, 0), range=full, name="ur5e/shoulder_lift_joint", **J))
    ua.geoms.append(cap((0.06, 0.06), (0, -0.04, 0), (1, 1, 0, 0), name="ur5e/upper_col0"))
    ua.geoms.append(cap((0.05, 0.2), (0, 0, 0.2), name="ur5e/upper_col1"))
    fa = ua.add(S.Body("ur5e/forearm_link", pos=(0, -0.131, 0.425), gravcomp=1.0,
                       inertial=S.Inertial(2.275, pos=(0, 0, 0.196), diaginertia=(0.0311796, 0.0311796, 0.004095))))
    fa.joints.append(S.Joint(S.HINGE, axis=(0, 1, 0), range=(-3.1415, 3.1415), name="ur5e/elbow_joint", **J))
    fa.geoms.append(cap((0.055, 0.06), (0, 0.08, 0), (1, 1, 0, 0), name="ur5e/fore_col0"))
    fa.geoms.append(cap((0.038, 0.19), (0, 0, 0.2), name="ur5e/fore_col1"))
    w1 = fa.add(S.Body("ur5e/wrist_1_link", pos=(0, 0, 0.392), quat=(1, 0, 1, 0), gravcomp=1.0,
                       inertial=S.Inertial(1.219, pos=(0, 0.127, 0), diaginertia=(0.0025599, 0.0025599, 0.0021942))))
    w1.joints.append(S.Joint(S.HINGE, axis=(0, 1, 0), range=full, name="ur5e/wrist_1_joint", **J))
    w1.geoms.append(cap((0.04, 0.07), (0, 0.05, 0), (1, 1, 0, 0), name="ur5e/w1_col"))
    w2 = w1.add(S.Body("ur5e/wrist_2_link", pos=(0, 0.127, 0), gravcomp=1.0,
                       inertial=S.Inertial(1.219, pos=(0, 0, 0.1), diaginertia=(0.0025599, 0.0025599, 0.0021942))))
    w2.joints.append(S.Joint(S.HINGE, axis=(0, 0, 1), range=full, name="ur5e/wrist_2_joint", **J))
    w2.geoms.append(cap((0.04, 0.06), (0, 0, 0.04), name="ur5e/w2_col0"))
    w2.geoms.append(cap((0.04, 0.04), (0, 0.02, 0.1), (1, 1, 0, 0), name="ur5e/w2_col1"))
    w3 = w2.add(S.Body("ur5e/wrist_3_link", pos=(0, 0, 0.1), gravcomp=1.0,
                       inertial=S.Inertial(0.1889, pos=(0, 0.0771683, 0), quat=(1, 0, 0, 1),
                                           diaginertia=(0.000132134, 9.90863e-05, 9.90863e-05))))
    w3.joints.append(S.Joint(S.HINGE, axis=(0, 1, 0), range=full, name="ur5e/wrist_3_joint", **J))
    w3.geoms.append(S.Geom(S.CYLINDER, (0.04, 0.02), pos=(0, 0.08, 0), quat=(1, 1, 0, 0), name="ur5e/w3_col", **col))
    w3.sites.append(S.Site("ur5e/attachment_site", pos=(0, 0.1, 0), quat=(-1, 1, 0, 0)))
    names = ["shoulder_pan", "shoulder_lift", "elbow", "wrist_1", "wrist_2", "wrist_3"]
    acts = []
    for i, n in enumerate(names):
        big = i < 3
        cr = (-3.1415, 3.1415) if n == "elbow" else (-6.2831, 6.2831)
        acts.append(S.PositionActuator(f"ur5e/{n}_joint", kp=2000.0 if big else 500.0, kv=400.0 if big else 100.0,
                                       ctrlrange=cr, forcerange=(-150.0, 150.0) if big else (-28.0, 28.0)))
    return w3, acts


def _arena_plane(world: S.Body, side=3.0):
    # EmptyRobotArena(3): one ground plane (empty_robot_arena.py:18-20)
    world.geoms.append(S.Geom(S.PLANE, (side / 2, side / 2, 0.1), name="ground", contype=_BIT_WORLD, conaffinity=_BIT_WORLD | _BIT_ARM,
                               rgba=(0.3, 0.3, 0.3, 1)))  # empty_robot_arena.py:19


def _robot_task_common(t: MjsTask, comp: Compiled, control_dt, physics_dt, tcp_offset):
    t.n_sub = int(round(control_dt / physics_dt))
    t.control_dt = control_dt
    t.tcp_site = comp.site_id["ur5e/attachment_site"]
    t.arm_qadr = comp.jnt_qadr["ur5e/shoulder_pan_joint"]
    t.set("tcp_offset", tcp_offset)
    t.set("ik_dh", UR5E_DH)


@dataclasses.dataclass
class RobotPushConfig:
    """Mirror of robot_planar_push.py:29-73 (state observations only) plus the
    `object_type` switch of SURVEY.md D2: the measured config uses cylinders."""

    reward_type: str = DENSE_NEG_DISTANCE_REWARD
    physics_timestep: float = 0.005
    control_timestep: float = 0.1
    max_control_steps_per_episode: int = 500
    nearest_object_reward_coefficient: float = 0.1
    target_radius: float = 0.05
    n_objects: int = 3
    object_type: str = "cylinder"


def build_planar_push(config: RobotPushConfig | None = None) -> TaskBundle:
    """RobotPushTask scene (robot_planar_push.py:76-176, eef/cylinder.py:18-41) with
    CylinderProp objects (r 0.02, half-height 0.019, mass 0.1, condim 4,
    friction [1, 0.05, 0] -- google_block.py:31,43-49 on a primitive)."""
    cfg = config or RobotPushConfig()
    assert cfg.object_type == "cylinder", "convex-hull google blocks are SURVEY.md 8f-1 (not built)"
    assert 1 <= cfg.n_objects <= 3, "the library instantiates the step kernel for 1, 2 and 3 props (the reference default of 5 needs nv = 36 > 32 lanes)"
    world = S.Body("world")
    _arena_plane(world)
    w3, acts = _ur5e(world)
    eef = w3.add(S.Body("ur5e/cylinder_eef", pos=(0, 0.1, 0), quat=(-1, 1, 0, 0)))
    eef.geoms.append(S.Geom(S.CYLINDER, (0.02, 0.05), pos=(0, 0, 0.051), mass=0.1, name="cylinder-EEF", rgba=(0.2, 0.2, 0.2, 1)))  # eef/cylinder.py:31
    R, H = 0.02, 0.019
    for i in range(cfg.n_objects):
        ob = world.add(S.Body(f"object{i}", pos=(-0.1 + 0.1 * i, -0.45, H)))
        ob.joints.append(S.Joint(S.FREE, name=f"object{i}/free"))
        ob.geoms.append(S.Geom(S.CYLINDER, (R, H), mass=0.1, condim=4, friction=(1.0, 0.05, 0.0), name=f"object{i}/geom",
                               rgba=[(1, 0, 0, 1), (0, 0, 1, 1), (0, 1, 0, 1), (1, 1, 0, 1)][i]))  # google_block.py:12-19 colours
    sc = S.Scene(world, actuators=acts, timestep=cfg.physics_timestep, integrator=1)
    comp = compile_scene(sc)
    t = MjsTask()
    t.kind = MACROS["MJS_TASK_PUSH"]
    _robot_task_common(t, comp, cfg.control_timestep, cfg.physics_timestep, (0, 0, 0.1))
    t.reward_type = 0 if cfg.reward_type == SPARSE_REWARD else 1
    t.max_steps = cfg.max_control_steps_per_episode
    t.n_obj = cfg.n_objects
    t.action_dim = 2
    t.obs_dim = 5 + 2 * cfg.n_objects
    for i in range(cfg.n_objects):
        t.obj_body[i] = comp.body_id[f"object{i}"]
        t.obj_qadr[i] = comp.jnt_qadr[f"object{i}/free"]
        t.obj_dadr[i] = comp.jnt_dadr[f"object{i}/free"]
    t.time_limit = 0.0  # scripts build Environment without time_limit (scripts/sb3/planar_push.py:72-73)
    t.target_radius = cfg.target_radius
    t.nearest_coef = cfg.nearest_object_reward_coefficient
    t.tcp_z = 0.02
    t.obj_radius = R
    t.settle_steps = 150
    t.set("spawn_lo", (-0.2, -0.6, 0.02)); t.set("spawn_hi", (0.2, -0.3, 0.02))
    t.set("target_lo", (-0.15, -0.55, 0.001)); t.set("target_hi", (0.15, -0.35, 0.005))
    t.set("obj_lo", (-0.15, -0.55, 0.05)); t.set("obj_hi", (0.15, -0.35, 0.2))
    # action_spec is [-1,1]^2 (robot_planar_push.py:229-234); synthetic rollouts draw from the
    # reachable robot workspace instead because servoL raises outside it
    t.set("action_lo", [-0.2, -0.6]); t.set("action_hi", [0.2, -0.3])
    keys = [("ur5e/tcp_position", 3), ("target_position", 2), ("block_positions", 2 * cfg.n_objects)]
    vis = _geom_visuals(comp) + [_marker(S.CYLINDER, (cfg.target_radius, 0.001), (1, 1, 1, 1), "MJS_ANCHOR_TARGET")]  # robot_planar_push.py:90-98
    cams = {"front_tilted": _camera((0.0, -1.1, 0.5), (-0.7, -0.35, 0, 0.0), 70), "top_down": _camera((0.0, -0.5, 2.4), (1, 0, 0, 0), 30)}  # :42-43
    return TaskBundle("robot_planar_push", comp, t, keys, np.array([-1, -1], np.float32), np.array([1, 1], np.float32), np.float32,
                      visuals=vis, cameras=cams, image_resolution=96)


def _rigid_gripper(w3: S.Body):
    """Rigid stand-in for the Robotiq 2F-85: the menagerie MJCF is not
    obtainable here.  Same TCP offset 0.174 (gripper.py:44-48) and ~0.9 kg; a capsule
    palm that only touches the ground plane plus a spherical fingertip at the TCP."""
    g = w3.add(S.Body("ur5e/gripper", pos=(0, 0.1, 0), quat=(-1, 1, 0, 0),
                      inertial=S.Inertial(0.9, pos=(0, 0, 0.06), diaginertia=(0.002, 0.002, 0.0008))))
    g.geoms.append(S.Geom(S.CAPSULE, (0.03, 0.05), pos=(0, 0, 0.075), contype=_BIT_ARM, conaffinity=0, name="gripper/palm", rgba=(0.1, 0.1, 0.1, 1)))
    g.geoms.append(S.Geom(S.SPHERE, (0.015,), pos=(0, 0, 0.159), name="gripper/fingertip", rgba=(0.1, 0.1, 0.1, 1)))
    return g


@dataclasses.dataclass
class RobotReachConfig:
    """Mirror of robot_reach.py:33-82 (state observations, ABS_EEF_ACTION)."""

    reward_type: str = DENSE_NEG_DISTANCE_REWARD
    physics_timestep: float = 0.005
    control_timestep: float = 0.1
    max_control_steps_per_episode: int = 100
    goal_distance_threshold: float = 0.02


def build_robot_reach(config: RobotReachConfig | None = None, obstacle: dict | None = None) -> TaskBundle:
    """RobotReachTask scene (robot_reach.py:85-132).  `obstacle` (not in the reference; used by the collision parity tests):
    a static capsule in the world that the arm's link capsules collide with, dict(size=(r, hh), pos=..., quat=...)."""
    cfg = config or RobotReachConfig()
    world = S.Body("world")
    _arena_plane(world)
    if obstacle:
        world.geoms.append(S.Geom(S.CAPSULE, obstacle["size"], pos=obstacle["pos"], quat=obstacle.get("quat", (1, 0, 0, 0)), name="obstacle",
                                  contype=_BIT_WORLD, conaffinity=_BIT_OBST, rgba=(0.8, 0.4, 0.1, 1)))
    w3, acts = _ur5e(world)
    _rigid_gripper(w3)
    sc = S.Scene(world, actuators=acts, timestep=cfg.physics_timestep, integrator=1)
    comp = compile_scene(sc)
    t = MjsTask()
    t.kind = MACROS["MJS_TASK_REACH"]
    _robot_task_common(t, comp, cfg.control_timestep, cfg.physics_timestep, (0, 0, 0.174))
    t.reward_type = 0 if cfg.reward_type == SPARSE_REWARD else 1
    t.action_dim = 3
    t.obs_dim = 6
    t.goal_threshold = cfg.goal_distance_threshold
    # the reference never terminates this task; episodes end through the composer time limit
    t.time_limit = cfg.max_control_steps_per_episode * cfg.control_timestep
    lo, hi = (-0.1, -0.6, 0.02), (0.1, -0.4, 0.2)  # robot_reach.py:108-110
    t.set("spawn_lo", lo); t.set("spawn_hi", hi)
    t.set("target_lo", lo); t.set("target_hi", hi)
    t.set("action_lo", lo); t.set("action_hi", hi)
    keys = [("ur5e/tcp_position", 3), ("target_position", 3)]
    vis = _geom_visuals(comp) + [_marker(S.SPHERE, (0.01,), (1, 1, 1, 1), "MJS_ANCHOR_TARGET")]  # robot_reach.py:97-104 (target_radius)
    return TaskBundle("robot_reach", comp, t, keys, np.array(lo, np.float64), np.array(hi, np.float64), np.float64,
                      visuals=vis, cameras={"front_tilted": _camera((0.0, -1.1, 0.5), (-0.7, -0.35, 0, 0.0), 70)}, image_resolution=96)  # :52


ABS_EEF_ACTION, ABS_JOINT_ACTION = "absolute_eef_action", "absolute_joint_action"  # robot_push_button.py:29-30


def build_push_button(max_steps: int = 100, button_disturbances: bool = False, action_type: str = ABS_JOINT_ACTION) -> TaskBundle:
    """RobotPushButtonTask scene (robot_push_button.py:56-119, props/switch.py:17-41),
    state observations; rigid gripper (the gripper channel of the action is accepted and ignored).
    action_type (robot_push_button.py:143-157,172-203): ABS_JOINT_ACTION -- 6 joint targets + gripper through servoJ, observation
    joint_configuration; ABS_EEF_ACTION -- TCP position in the robot workspace + gripper through servoL (top-down), observation
    tcp_position.  button_disturbances
    (robot_push_button.py:56,159-165): an active, unpressed switch deactivates with probability 0.01 per control step."""
    world = S.Body("world")
    _arena_plane(world)
    w3, acts = _ur5e(world)
    _rigid_gripper(w3)
    sw = world.add(S.Body("switch", pos=(0.0, -0.45, 0.0)))
    sw.geoms.append(S.Geom(S.BOX, (0.025, 0.025, 0.025), pos=(0, 0, 0.025), name="switch/box", rgba=(1, 1, 1, 1)))        # switch.py:24-29
    sw.geoms.append(S.Geom(S.CYLINDER, (0.02, 0.02), pos=(0, 0, 0.05), name="switch/button", rgba=(1, 0, 0, 1)))        # switch.py:30-36
    sw.sites.append(S.Site("switch/button_site", pos=(0, 0, 0.05)))
    sc = S.Scene(world, actuators=acts, timestep=0.005, integrator=1, ovr_body="switch",
                 touch_site="switch/button_site", touch_size=(0.02 * 1.01, 0.02 * 1.01))
    comp = compile_scene(sc)
    t = MjsTask()
    t.kind = MACROS["MJS_TASK_BUTTON"]
    _robot_task_common(t, comp, 0.1, 0.005, (0, 0, 0.174))
    assert action_type in (ABS_EEF_ACTION, ABS_JOINT_ACTION), action_type
    eef = action_type == ABS_EEF_ACTION
    t.reward_type = 0
    t.action_dim = 4 if eef else 7
    t.obs_dim = 7 if eef else 10
    t.time_limit = max_steps * 0.1  # mujoco_sim/__init__.py:21,34
    t.set("spawn_lo", (-0.2, -0.6, 0.02)); t.set("spawn_hi", (0.2, -0.3, 0.3))
    t.set("target_lo", (-0.2, -0.6, 0.0)); t.set("target_hi", (0.2, -0.3, 0.1))
    t.set("button_goal", (-0.3, -0.2, 0.3))
    t.button_goal_radius = 0.05
    t.touch_lo, t.touch_hi = 5.0, 200.0
    t.button_disturbances = int(bool(button_disturbances))
    t.disturb_prob = 0.01
    # synthetic rollouts: q_now +- 0.05 rad handled on device; box kept for the gripper dof
    if eef:  # synthetic rollouts: uniform over the action spec = the robot workspace (robot_push_button.py:78,176-192)
        t.set("action_lo", [-0.2, -0.6, 0.02, 0.0]); t.set("action_hi", [0.2, -0.3, 0.3, 0.085])
        keys = [("ur5e/tcp_position", 3), ("unnamed_model/position", 3), ("unnamed_model/active", 1)]
        lo, hi = np.array([-0.2, -0.6, 0.02, 0.0]), np.array([0.2, -0.3, 0.3, 0.085])
    else:
        t.set("action_lo", [-0.05] * 6 + [0.0]); t.set("action_hi", [0.05] * 6 + [0.085])
        keys = [("ur5e/joint_configuration", 6), ("unnamed_model/position", 3), ("unnamed_model/active", 1)]
        lo, hi = np.array([-3.14] * 6 + [0.0]), np.array([3.14] * 6 + [0.085])
    vis = _geom_visuals(comp, active={"switch/button": (0, 1, 0, 1)})  # switch.py:59: green while active
    # scene camera and the wrist camera riding on the flange (robot_push_button.py:51-55,87-96); the flange frame is the
    # frame of the gripper palm geom's body up to the geom's own offset, so the wrist camera is anchored to wrist_3's geom
    cams = {"scene": _camera((0.0, -1.7, 0.7), (-0.7, -0.35, 0, 0.0), 70)}
    gw = comp.geom_id["ur5e/w3_col"]  # cylinder at (0, 0.08, 0), quat (1, 1, 0, 0) in wrist_3; flange site at (0, 0.1, 0), quat (-1, 1, 0, 0)
    wp, wq = _in_geom_frame((0, 0.08, 0), (1, 1, 0, 0), (0, 0.1, 0), (-1, 1, 0, 0), (0.0, 0.05, 0.0), (0.0, 0.0, 0.999, 0.04))
    cams["wrist"] = _camera(wp, wq, 42, geom=gw)
    return TaskBundle("robot_push_button", comp, t, keys, lo, hi, np.float64,
                      visuals=vis, cameras=cams, image_resolution=96)

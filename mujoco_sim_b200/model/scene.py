"""Programmatic scene description (the subset of MJCF the four tasks use).

The reference composes its scenes with dm_control.mjcf (arena + entities, e.g.
/root/reference/mujoco_sim/environments/tasks/point_reach.py:75-98); neither
dm_control nor the mujoco bindings exist in this image, so the same trees are
authored here as plain dataclasses and compiled by compiler.py.
Quaternions are (w, x, y, z) as in MJCF.
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional, Sequence

import numpy as np

# mjtGeom / mjtJoint values
PLANE, SPHERE, CAPSULE, CYLINDER, BOX = 0, 2, 3, 5, 6
FREE, SLIDE, HINGE = 0, 2, 3

DEFAULT_SOLREF = (0.02, 1.0)
DEFAULT_SOLIMP = (0.9, 0.95, 0.001, 0.5, 2.0)


def _v(x, n):
    a = np.asarray(x, dtype=np.float64).ravel()
    assert a.size == n, (x, n)
    return a


@dataclasses.dataclass
class Geom:
    type: int
    size: Sequence[float]
    pos: Sequence[float] = (0.0, 0.0, 0.0)
    quat: Sequence[float] = (1.0, 0.0, 0.0, 0.0)
    mass: Optional[float] = None
    density: float = 1000.0
    friction: Sequence[float] = (1.0, 0.005, 0.0001)
    condim: int = 3
    contype: int = 1
    conaffinity: int = 1
    priority: int = 0
    solref: Sequence[float] = DEFAULT_SOLREF
    solimp: Sequence[float] = DEFAULT_SOLIMP
    solmix: float = 1.0
    margin: float = 0.0
    gap: float = 0.0
    name: str = ""
    rgba: Sequence[float] = (0.5, 0.5, 0.5, 1.0)  # MuJoCo's default geom colour; only the ray caster reads it


@dataclasses.dataclass
class Joint:
    type: int
    axis: Sequence[float] = (0.0, 0.0, 1.0)
    pos: Sequence[float] = (0.0, 0.0, 0.0)
    range: Optional[Sequence[float]] = None  # autolimits: limited iff range given
    armature: float = 0.0
    damping: float = 0.0
    margin: float = 0.0
    solref: Sequence[float] = DEFAULT_SOLREF
    solimp: Sequence[float] = DEFAULT_SOLIMP
    name: str = ""


@dataclasses.dataclass
class Site:
    name: str
    pos: Sequence[float] = (0.0, 0.0, 0.0)
    quat: Sequence[float] = (1.0, 0.0, 0.0, 0.0)


@dataclasses.dataclass
class Inertial:
    mass: float
    pos: Sequence[float] = (0.0, 0.0, 0.0)
    quat: Sequence[float] = (1.0, 0.0, 0.0, 0.0)
    diaginertia: Sequence[float] = (0.0, 0.0, 0.0)


@dataclasses.dataclass
class Body:
    name: str
    pos: Sequence[float] = (0.0, 0.0, 0.0)
    quat: Sequence[float] = (1.0, 0.0, 0.0, 0.0)
    joints: List[Joint] = dataclasses.field(default_factory=list)
    geoms: List[Geom] = dataclasses.field(default_factory=list)
    sites: List[Site] = dataclasses.field(default_factory=list)
    children: List["Body"] = dataclasses.field(default_factory=list)
    inertial: Optional[Inertial] = None
    gravcomp: float = 0.0
    mocap: bool = False

    def add(self, child: "Body") -> "Body":
        self.children.append(child)
        return child


@dataclasses.dataclass
class Weld:
    body1: str
    body2: str
    solref: Sequence[float] = DEFAULT_SOLREF
    solimp: Sequence[float] = DEFAULT_SOLIMP
    torquescale: float = 1.0


@dataclasses.dataclass
class PositionActuator:
    """<general gaintype=fixed biastype=affine> on a joint (UR5e menagerie style)."""

    joint: str
    kp: float
    kv: float
    ctrlrange: Optional[Sequence[float]] = None
    forcerange: Optional[Sequence[float]] = None
    gear: float = 1.0


@dataclasses.dataclass
class Scene:
    world: Body
    welds: List[Weld] = dataclasses.field(default_factory=list)
    actuators: List[PositionActuator] = dataclasses.field(default_factory=list)
    timestep: float = 0.002
    gravity: Sequence[float] = (0.0, 0.0, -9.81)
    integrator: int = 0  # MJS_INT_EULER
    impratio: float = 1.0
    tolerance: float = 1e-8
    iterations: int = 100
    ovr_body: Optional[str] = None
    touch_site: Optional[str] = None
    touch_size: Sequence[float] = (0.0, 0.0)

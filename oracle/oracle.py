"""ctypes front-end of the CPU oracle (oracle/mjs_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never by mujoco_sim_b200.
PARITY UNPINNED (no MuJoCo here or on the GPU box; see mjs_oracle.cpp header).
"""
from __future__ import annotations

import ctypes
import subprocess
from pathlib import Path

import numpy as np

from mujoco_sim_b200.model.cstruct import MjsEnvState, MjsModel, MjsStepOut, MjsTask

_DIR = Path(__file__).resolve().parent
_LIB = _DIR / "_build" / "libmjs_oracle.so"


def build(force: bool = False):
    src = _DIR / "mjs_oracle.cpp"
    hdr = _DIR.parent / "include" / "mjs_model.h"
    hdr2 = _DIR.parent / "include" / "mjs_convex.h"
    if force or not _LIB.exists() or _LIB.stat().st_mtime < max(src.stat().st_mtime, hdr.stat().st_mtime, hdr2.stat().st_mtime):
        subprocess.check_call(["make", "-C", str(_DIR), "-s", "-B"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(str(_LIB))
        assert _lib.oracle_sizeof_model() == ctypes.sizeof(MjsModel)
        assert _lib.oracle_sizeof_task() == ctypes.sizeof(MjsTask)
        assert _lib.oracle_sizeof_state() == ctypes.sizeof(MjsEnvState)
        assert _lib.oracle_sizeof_out() == ctypes.sizeof(MjsStepOut)
    return _lib


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


class OracleEnvs:
    """n independent envs stepped by the oracle; states are MjsEnvState AoS."""

    def __init__(self, bundle, n: int):
        self.bundle = bundle
        self.model = bundle.model
        self.task = bundle.task
        self.n = n
        self.states = (MjsEnvState * n)()
        self.outs = (MjsStepOut * n)()
        self.flops = ctypes.c_long(0)
        self.tol_scale = 1e-4  # iterate far below MuJoCo's 1e-8: exact minimiser

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.float64).reshape(self.n, self.task.action_dim)
        lib().oracle_control_step(
            ctypes.byref(self.model), ctypes.byref(self.task), self.states, _dp(a), self.outs, self.n,
            ctypes.c_double(self.tol_scale), ctypes.byref(self.flops),
        )
        return self.collect()

    def step_disturb(self, actions, u):
        """step() with the button-disturbance draw u[n] of this control step (robot_push_button.py:159-165)"""
        a = np.ascontiguousarray(actions, dtype=np.float64).reshape(self.n, self.task.action_dim)
        u = np.ascontiguousarray(u, dtype=np.float64)
        lib().oracle_control_step_disturb(
            ctypes.byref(self.model), ctypes.byref(self.task), self.states, _dp(a), self.outs, self.n,
            ctypes.c_double(self.tol_scale), _dp(u),
        )
        return self.collect()

    def step_touch(self, actions):
        """step() that also returns the touch-sensor force after every substep: (collect(), touch[n, n_sub])"""
        a = np.ascontiguousarray(actions, dtype=np.float64).reshape(self.n, self.task.action_dim)
        touch = np.zeros((self.n, self.task.n_sub))
        lib().oracle_control_step_touch(
            ctypes.byref(self.model), ctypes.byref(self.task), self.states, _dp(a), self.outs, self.n,
            ctypes.c_double(self.tol_scale), _dp(touch),
        )
        return self.collect(), touch

    def step_raw(self, actions):
        """step without gathering outputs into numpy (timed CPU baseline: no Python in the loop)"""
        a = np.ascontiguousarray(actions, dtype=np.float64).reshape(self.n, self.task.action_dim)
        lib().oracle_control_step(
            ctypes.byref(self.model), ctypes.byref(self.task), self.states, _dp(a), self.outs, self.n,
            ctypes.c_double(self.tol_scale), ctypes.byref(self.flops),
        )

    def demo_policy(self):
        """noise-free scripted demonstrator actions for the current states: (actions[n, A], ok[n])"""
        a = np.zeros((self.n, self.task.action_dim)); ok = np.zeros(self.n, np.int32)
        lib().oracle_demo_policy(ctypes.byref(self.model), ctypes.byref(self.task), self.states, self.n, _dp(a), ok.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
        return a, ok.astype(bool)

    def finish_reset(self, settle_steps: int = 0):
        lib().oracle_finish_reset(ctypes.byref(self.model), ctypes.byref(self.task), self.states, self.outs, self.n, settle_steps)
        return self.collect()

    def collect(self):
        od = self.task.obs_dim
        obs = np.array([[o.obs[k] for k in range(od)] for o in self.outs])
        rew = np.array([o.reward for o in self.outs])
        disc = np.array([o.discount for o in self.outs])
        flags = np.array([[o.terminated, o.truncated, o.success, o.done] for o in self.outs], dtype=np.uint8)
        return obs, rew, disc, flags

    def geom_poses(self, e: int):
        """world pose of every model geom of env e: (xpos [ngeom, 3], xmat [ngeom, 9])"""
        ng = self.model.ngeom
        xpos, xmat = np.zeros((ng, 3)), np.zeros((ng, 9))
        lib().oracle_geom_poses(ctypes.byref(self.model), ctypes.byref(self.states[e]), _dp(xpos), _dp(xmat))
        return xpos, xmat

    def ncon(self, e: int) -> int:
        return lib().oracle_ncon(ctypes.byref(self.model), ctypes.byref(self.states[e]))

    def field(self, name, width=None):
        """gather a state field over envs as numpy"""
        vals = []
        for s in self.states:
            v = getattr(s, name)
            vals.append(list(v)[:width] if hasattr(v, "__len__") else v)
        return np.array(vals)

    def stage_dump(self, e: int):
        m = self.model
        nv, nb, ns = m.nv, m.nbody, max(m.nsite, 1)
        R, C = 160, 24
        out = dict(
            xpos=np.zeros((nb, 3)), xquat=np.zeros((nb, 4)), site_xpos=np.zeros((ns, 3)), site_xmat=np.zeros((ns, 9)),
            M=np.zeros((nv, nv)), qfrc_bias=np.zeros(nv), qfrc_passive=np.zeros(nv), qacc_smooth=np.zeros(nv), qacc=np.zeros(nv),
            con_dist=np.zeros(C), con_pos=np.zeros((C, 3)), con_frame=np.zeros((C, 9)), efc_J=np.zeros((R, nv)),
            efc_aref=np.zeros(R), efc_D=np.zeros(R), efc_force=np.zeros(R),
        )
        ncon, nefc, niter = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        touch = ctypes.c_double(0)
        o = out
        lib().oracle_stage_dump(
            ctypes.byref(m), ctypes.byref(self.states[e]), _dp(o["xpos"]), _dp(o["xquat"]), _dp(o["site_xpos"]), _dp(o["site_xmat"]),
            _dp(o["M"]), _dp(o["qfrc_bias"]), _dp(o["qfrc_passive"]), _dp(o["qacc_smooth"]), _dp(o["qacc"]),
            ctypes.byref(ncon), _dp(o["con_dist"]), _dp(o["con_pos"]), _dp(o["con_frame"]), ctypes.byref(nefc), _dp(o["efc_J"]),
            _dp(o["efc_aref"]), _dp(o["efc_D"]), _dp(o["efc_force"]), ctypes.byref(touch), ctypes.byref(niter),
        )
        out["ncon"], out["nefc"], out["touch"], out["niter"] = ncon.value, nefc.value, touch.value, niter.value
        for k in ("con_dist", "con_pos", "con_frame"):
            out[k] = out[k][: ncon.value]
        for k in ("efc_J", "efc_aref", "efc_D", "efc_force"):
            out[k] = out[k][: nefc.value]
        return out


def ur_fk(dh, q):
    T = np.zeros(16)
    lib().oracle_ur_fk(_dp(np.asarray(dh, float)), _dp(np.asarray(q, float)), _dp(T))
    return T.reshape(4, 4)


def ur_ik_all(dh, T):
    sols = np.zeros((8, 6))
    n = lib().oracle_ur_ik_all(_dp(np.asarray(dh, float)), _dp(np.ascontiguousarray(T, float).ravel()), _dp(sols))
    return sols[:n]


def ur_ik_closest(dh, T, seed):
    q = np.zeros(6)
    ok = lib().oracle_ur_ik_closest(_dp(np.asarray(dh, float)), _dp(np.ascontiguousarray(T, float).ravel()), _dp(np.asarray(seed, float)), _dp(q))
    return q if ok else None


def convex_pair(t1, size1, pos1, mat1, t2, size2, pos2, mat2):
    """MPR penetration of two primitives (include/mjs_convex.h): None or (depth, dir[3], pos[3])"""
    out = np.zeros(7)
    f = lambda a, n: _dp(np.ascontiguousarray(np.asarray(a, float).ravel()[:n] if np.size(a) >= n else np.concatenate([np.ravel(a), np.zeros(n - np.size(a))])))
    ok = lib().oracle_convex_pair(int(t1), f(size1, 3), f(pos1, 3), f(mat1, 9), int(t2), f(size2, 3), f(pos2, 3), f(mat2, 9), _dp(out))
    return (out[0], out[1:4].copy(), out[4:7].copy()) if ok else None

"""numpy-in / numpy-out vector-env facades over BatchedMujocoSimEnv (SURVEY.md 8b "callers").

The reference is consumed through
  * stable_baselines3 VecEnv (DummyVecEnv / SubprocVecEnv of DMCEnvironmentAdapter,
    /root/reference/scripts/sb3/planar_push.py:85,95-98), and
  * gymnasium.vector (lerobot: /root/reference/lerobot/lerobot/common/envs/factory.py:61-67).
Neither stable_baselines3 nor gymnasium is installed in this image, so the facades are duck-typed: they
implement the method/attribute protocol those callers use and subclass the real base class only when it can
be imported.  Spaces are described by the minimal `Box` / `DictSpace` records below (same fields as
gymnasium.spaces.Box / Dict: low, high, shape, dtype / spaces) and are replaced by the gymnasium classes
when gymnasium is importable.
"""
from __future__ import annotations

import dataclasses
from typing import Any, Optional

import numpy as np

from .envs import BatchedMujocoSimEnv


@dataclasses.dataclass
class Box:
    low: np.ndarray
    high: np.ndarray
    shape: tuple
    dtype: Any

    def sample(self, rng: Optional[np.random.RandomState] = None):
        rng = rng or np.random
        lo = np.where(np.isfinite(self.low), self.low, -1.0)
        hi = np.where(np.isfinite(self.high), self.high, 1.0)
        return rng.uniform(lo, hi).astype(self.dtype)


@dataclasses.dataclass
class DictSpace:
    spaces: dict


def _box(low, high, dtype):
    low = np.asarray(low, dtype)
    high = np.asarray(high, dtype)
    try:  # the real thing when available
        import gymnasium

        return gymnasium.spaces.Box(low=low, high=high, dtype=dtype)
    except Exception:
        return Box(low, high, low.shape, np.dtype(dtype))


def single_action_space(env: BatchedMujocoSimEnv):
    """dmc2gym.py:90 -- the flattened action_spec as a float32 Box."""
    return _box(env.single_action_low, env.single_action_high, np.float32)


def single_observation_space(env: BatchedMujocoSimEnv, flatten: bool = False):
    """dmc2gym.py:93-101 -- Dict of unbounded Boxes per observable in the spec dtype (float64 in the reference: make the
    env with obs_dtype="float64"; the default here is float32), or one flat float32 Box when flatten_observation_space=True
    (dmc2gym.py:66-75 casts the flattened observation to float32)."""
    visual = getattr(env, "observation_type", "state_observations") == "visual_observations"
    if flatten:
        if visual:
            raise ValueError("flatten_observation_space needs state observations (dmc2gym.py:93-98 flattens vectors, not images)")
        n = sum(d for _, d in env.obs_keys)
        return _box(np.full(n, -np.inf), np.full(n, np.inf), np.float32)
    dt = np.float64 if getattr(env, "obs_dtype", "float32") == "float64" else np.float32
    spaces = {k: _box(np.full(d, -np.inf), np.full(d, np.inf), dt) for k, d in env.obs_keys if not visual or k in env._visual_keep}
    if visual:  # entities/camera.py:124-150: uint8 image observable, channel-last
        for key, _, shape in env.image_keys:
            spaces[key] = _box(np.zeros(shape), np.full(shape, 255), np.uint8)
    try:
        import gymnasium

        return gymnasium.spaces.Dict(spaces)
    except Exception:
        return DictSpace(spaces)


class _HostBuffers:
    def __init__(self, env: BatchedMujocoSimEnv):
        n, o = env.num_envs, env.obs_dim
        self.obs = np.zeros((n, o), np.float32)
        self.final = np.zeros((n, o), np.float32)
        self.rew = np.zeros(n, np.float32)
        self.flags = np.zeros((n, 4), np.uint8)

    def step(self, env, actions):
        """one control step with host buffers: mjs_step_host (float32), or the device call + copies of the float64 tensors"""
        import ctypes

        from . import lib as _lib

        if env.obs_dtype == "float64":
            import torch

            env.step_flat(torch.from_numpy(actions).to(env.device))
            self.obs, self.final = env.observation.cpu().numpy(), env.final_observation.cpu().numpy()
            self.rew, self.flags = env.reward.cpu().numpy(), env._flags.cpu().numpy()
            return
        p = lambda x: x.ctypes.data_as(ctypes.c_void_p)
        _lib.check(env._L.mjs_step_host(env._h, p(actions), p(self.obs), p(self.rew), p(self.flags), p(self.final), 1))

    def split(self, env, flat, images: bool = True):
        out, k = {}, 0
        for key, dim in env.obs_keys:
            out[key] = flat[:, k : k + dim].copy()
            k += dim
        if getattr(env, "observation_type", "state_observations") == "visual_observations":
            out = {key: out[key] for key in env._visual_keep}
            if images and len(flat) == env.num_envs:  # frames of the current device state (not for single terminal rows)
                for key, cam, shape in env.image_keys:
                    out[key] = env.render(cam, shape[1], shape[0], out=env._images[key]).cpu().numpy()
        return out


class SB3VecEnvFacade:
    """stable_baselines3.common.vec_env.VecEnv protocol: reset() -> obs; step_async(a); step_wait() ->
    (obs, rewards, dones, infos) with infos[i]["terminal_observation"], ["TimeLimit.truncated"],
    ["is_success"] (read by Monitor / EvalCallback, scripts/sb3/planar_push.py:85,130)."""

    def __init__(self, env: BatchedMujocoSimEnv, flatten_observation_space: bool = False):
        self.env = env
        self.num_envs = env.num_envs
        self.flatten = flatten_observation_space
        self.action_space = single_action_space(env)
        self.observation_space = single_observation_space(env, flatten_observation_space)
        self.render_mode = None
        self._b = _HostBuffers(env)
        self._actions = None

    def _obs(self, flat):
        return flat.copy() if self.flatten else self._b.split(self.env, flat)

    def reset(self):
        self.env.reset()
        self._b.obs = self.env.observation.cpu().numpy()
        return self._obs(self._b.obs)

    def step_async(self, actions):
        self._actions = np.ascontiguousarray(actions, np.float32).reshape(self.num_envs, self.env.action_dim)

    def step_wait(self):
        b, e = self._b, self.env
        b.step(e, self._actions)
        f = b.flags
        dones = f[:, 3].astype(bool)
        # One dict per env is the VecEnv protocol, but building 42 k dicts per step would dominate a rollout: envs that are not
        # done share one of four read-only info dicts (copy before mutating, as VecMonitor does); done envs get their own.
        code = (f[:, 0] != 0).astype(np.int64) + 2 * (f[:, 2] != 0).astype(np.int64)
        shared = self._shared_infos
        infos = [shared[c] for c in code.tolist()]
        for i in np.nonzero(dones)[0].tolist():
            info = dict(shared[code[i]])
            t = b.final[i : i + 1]
            info["terminal_observation"] = t[0].copy() if self.flatten else {k: v[0] for k, v in b.split(e, t, images=False).items()}
            info["TimeLimit.truncated"] = bool(f[i, 1]) and not bool(f[i, 0])
            infos[i] = info
        return self._obs(b.obs), b.rew.copy(), dones, infos

    @property
    def _shared_infos(self):
        if not hasattr(self, "_shared"):
            self._shared = []
            for c in range(4):
                info = {"discount": 0.0 if c & 1 else 1.0}
                if self.env.has_success:
                    info["is_success"] = float(bool(c & 2))
                self._shared.append(info)
        return self._shared

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def close(self):
        self.env.close()

    # the remaining VecEnv abstract methods, answered uniformly for all sub-envs
    def get_attr(self, attr_name, indices=None):
        return [getattr(self.env, attr_name)] * len(self._indices(indices))

    def set_attr(self, attr_name, value, indices=None):
        setattr(self.env, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        return [getattr(self.env, method_name)(*args, **kwargs)] * len(self._indices(indices))

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * len(self._indices(indices))

    def seed(self, seed=None):
        """VecEnv.seed: re-seed the live simulator (DMCEnvironmentAdapter.seed, dmc2gym.py:126-131); applied by the next
        reset().  Sub-env i draws from the Philox stream keyed by (seed, i), reported SB3-style as seed + i."""
        if seed is None:
            seed = int(np.random.randint(0, 2**31 - 1))
        self.env.seed(seed)
        return [seed + i for i in range(self.num_envs)]

    def _indices(self, indices):
        return range(self.num_envs) if indices is None else ([indices] if isinstance(indices, int) else list(indices))


class GymnasiumVectorFacade:
    """gymnasium.vector.VectorEnv protocol (same-step autoreset): reset(seed, options) -> (obs, info);
    step(actions) -> (obs, rewards, terminated, truncated, infos) with batched info arrays
    "is_success", "discount", "final_observation", "_final_observation" (dmc2gym.py:133-163 per env)."""

    def __init__(self, env: BatchedMujocoSimEnv, flatten_observation_space: bool = False):
        self.env = env
        self.num_envs = env.num_envs
        self.flatten = flatten_observation_space
        self.single_action_space = single_action_space(env)
        self.single_observation_space = single_observation_space(env, flatten_observation_space)
        self._b = _HostBuffers(env)

    def _obs(self, flat):
        return flat.copy() if self.flatten else self._b.split(self.env, flat)

    def reset(self, *, seed=None, options=None):
        self.env.reset(seed=seed)  # dmc2gym.py:157-160: a seed re-seeds the live env before the reset
        self._b.obs = self.env.observation.cpu().numpy()
        return self._obs(self._b.obs), {}

    def step(self, actions):
        b, e = self._b, self.env
        a = np.ascontiguousarray(actions, np.float32).reshape(self.num_envs, e.action_dim)
        b.step(e, a)
        f = b.flags
        info = {"discount": 1.0 - f[:, 0].astype(np.float32), "_final_observation": f[:, 3].astype(bool),
                "final_observation": b.final.copy() if self.flatten else b.split(e, b.final, images=False)}
        if e.has_success:
            info["is_success"] = f[:, 2].astype(np.float32)
        return self._obs(b.obs), b.rew.copy(), f[:, 0].astype(bool), f[:, 1].astype(bool), info

    def close(self):
        self.env.close()


# gym ids: the two the reference registers (/root/reference/mujoco_sim/__init__.py:26-40, both with visual observations) and a
# state-observation id per task in the same style
REGISTRY = {
    "mujoco_sim/point_mass_reach-v0": ("point_mass_reach", {"observation_type": "visual_observations", "image_resolution": 64}),
    "mujoco_sim/robot_push_button_visual-v0": ("robot_push_button", {"observation_type": "visual_observations", "image_resolution": 96}),
    "mujoco_sim/point_mass_reach_state-v0": ("point_mass_reach", {}),
    "mujoco_sim/robot_reach_state-v0": ("robot_reach", {}),
    "mujoco_sim/robot_planar_push_state-v0": ("robot_planar_push", {}),
    "mujoco_sim/robot_push_button_state-v0": ("robot_push_button", {}),
}


def make_vec(env_id: str, num_envs: int, api: str = "gymnasium", flatten_observation_space: bool = False, **kw):
    """gymnasium.make_vec-style factory over REGISTRY; api = "gymnasium" | "sb3"."""
    task, defaults = REGISTRY[env_id]
    env = BatchedMujocoSimEnv(task, num_envs, **{**defaults, **kw})
    cls = {"gymnasium": GymnasiumVectorFacade, "sb3": SB3VecEnvFacade}[api]
    return cls(env, flatten_observation_space)

"""Vectorised, GPU-resident mirror of the reference's DMCEnvironmentAdapter.

Reference interface mirrored (per env): /root/reference/mujoco_sim/environments/dmc2gym.py:78-168
  reset(seed) -> (obs dict, {})                                    :157-163
  step(a)     -> (obs dict, reward, terminated, truncated, info)   :133-155
  info = {"is_success": float (tasks with is_goal_reached), "discount": float}
Observation keys/orders are the reference's.  Everything is a
torch CUDA tensor with a leading env dimension; auto-reset follows the
gymnasium "same-step" convention (final observation in info).
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np
import torch

from . import lib as _lib
from .model import tasks as _tasks
from .model.cstruct import MjsEnvState

_BUILDERS = {
    "point_mass_reach": _tasks.build_pointmass_reach,
    "robot_reach": _tasks.build_robot_reach,
    "robot_planar_push": _tasks.build_planar_push,
    "robot_push_button": _tasks.build_push_button,
}


def register_builder(name, fn):
    _BUILDERS[name] = fn


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


STATE_OBS, VISUAL_OBS = "state_observations", "visual_observations"
# observation_type="visual_observations": what each task enables besides its camera(s)
# (point_reach.py:115-123, robot_reach.py:134-141, robot_planar_push.py:134-142, robot_push_button.py:108-124) and the
# observable names of the camera entities (entities/camera.py:153-160; the wrist camera is renamed at robot_push_button.py:95).
_VISUAL = {
    "point_mass_reach": (["pointmass/position"], [("Camera/rgb_image", "top_down")]),
    "robot_reach": (["ur5e/tcp_position"], [("Camera/rgb_image", "front_tilted")]),
    "robot_planar_push": (["ur5e/tcp_position"], [("Camera/rgb_image", None)]),  # None: the config's camera (first of the bundle)
    "robot_push_button": (["ur5e/joint_configuration", "ur5e/tcp_position"],  # whichever the action type enables (robot_push_button.py:109-113)
                          [("Camera/rgb_image", "scene"), ("ur5e/Camera/wrist_camera_rgb_image", "wrist")]),
}


class BatchedMujocoSimEnv:
    def __init__(self, task_name: str, num_envs: int, device="cuda:0", seed: int = 0, env_offset: int = 0,
                 precision: str = "f64", autoreset: bool = True, solver_tolerance: Optional[float] = None, obs_dtype: str = "float32",
                 observation_type: str = STATE_OBS, image_resolution: Optional[int] = None, **task_kwargs):
        if not torch.cuda.is_available():
            raise _lib.MjsimError("BatchedMujocoSimEnv needs a CUDA device (there is no CPU fallback)")
        self.bundle = _BUILDERS[task_name](**task_kwargs)
        model = type(self.bundle.model)()
        ctypes.memmove(ctypes.byref(model), ctypes.byref(self.bundle.model), ctypes.sizeof(model))
        if solver_tolerance is not None:  # mjOption.tolerance of the Newton solver (MuJoCo default 1e-8)
            model.tolerance = float(solver_tolerance)
        self.task_name = task_name
        self.num_envs = int(num_envs)
        self.device = torch.device(device)
        self.autoreset = autoreset
        self._L = _lib.load()
        self._h = ctypes.c_void_p()
        prec = {"f64": _lib.PRECISION_F64, "mixed": _lib.PRECISION_MIXED}[precision]
        dev_index = self.device.index or 0
        _lib.check(self._L.mjs_create(ctypes.byref(model), ctypes.byref(self.bundle.task), self.num_envs, dev_index,
                                      int(seed) & (2**64 - 1), int(env_offset), prec, ctypes.byref(self._h)))
        N, O, A = self.num_envs, self.bundle.task.obs_dim, self.bundle.task.action_dim
        self.obs_dim, self.action_dim = O, A
        kw = dict(device=self.device)
        self._obs = torch.zeros(N, O, dtype=torch.float32, **kw)
        self._final_obs = torch.zeros(N, O, dtype=torch.float32, **kw)
        self._reward = torch.zeros(N, dtype=torch.float32, **kw)
        self._flags = torch.zeros(N, 4, dtype=torch.uint8, **kw)
        # float64 observations / rewards like the reference's (dm_env specs are float64, dmc2gym.py:66-75): the kernels
        # write the unrounded doubles into these tensors next to the float32 ones (mjs_set_output64)
        assert obs_dtype in ("float32", "float64"), obs_dtype
        self.obs_dtype = obs_dtype
        self._obs64 = self._final_obs64 = self._reward64 = None
        if obs_dtype == "float64":
            self._obs64 = torch.zeros(N, O, dtype=torch.float64, **kw)
            self._final_obs64 = torch.zeros(N, O, dtype=torch.float64, **kw)
            self._reward64 = torch.zeros(N, dtype=torch.float64, **kw)
            _lib.check(self._L.mjs_set_output64(self._h, _ptr(self._obs64), _ptr(self._reward64), _ptr(self._final_obs64)))
        self.seed_value = int(seed)
        self.obs_keys = list(self.bundle.obs_keys)
        # visual observations: the task's camera image(s), ray-cast on the device after every step / reset (mjs_render), next to the
        # state observables the reference keeps enabled in that mode.  The kernels still produce the full state vector.
        assert observation_type in (STATE_OBS, VISUAL_OBS), observation_type
        self.observation_type = observation_type
        self.image_keys = []  # [(observation key, camera name, (H, W, 3))]
        if observation_type == VISUAL_OBS:
            keep, cams = _VISUAL[task_name]
            self._visual_keep = [k for k, _ in self.obs_keys if k in keep]
            res = int(image_resolution or self.bundle.image_resolution)
            for key, cam in cams:
                cam = cam if cam is not None else next(iter(self.bundle.cameras))
                self.image_keys.append((key, cam, (res, res, 3)))
            self._images = {key: torch.zeros(N, res, res, 3, dtype=torch.uint8, device=self.device) for key, _, _ in self.image_keys}
        self.has_success = task_name in ("point_mass_reach", "robot_push_button")
        self.single_action_low = self.bundle.action_low
        self.single_action_high = self.bundle.action_high

    # ------------------------------------------------------------------ helpers
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _obs_dict(self, flat: torch.Tensor, images: bool = True):
        out, k = {}, 0
        for key, dim in self.obs_keys:
            out[key] = flat[:, k : k + dim]
            k += dim
        if self.observation_type == VISUAL_OBS:
            out = {key: out[key] for key in self._visual_keep}
            if images:  # of the CURRENT device state (after an auto-reset: the new episode's first frame, like the reference's reset obs)
                for key, cam, shape in self.image_keys:
                    out[key] = self.render(cam, shape[1], shape[0], out=self._images[key])
        return out

    def close(self):
        if self._h:
            self._L.mjs_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---------------------------------------------------------------------- API
    def reset(self, seed: Optional[int] = None, env_mask: Optional[torch.Tensor] = None, env_ids=None):
        """dmc2gym.py:157-163 for every env, or for the envs selected by a boolean `env_mask` [N] / a list of `env_ids`."""
        if seed is not None:  # dmc2gym.py:157-160: reset(seed=s) re-seeds the live env first
            self.seed(seed)
        mask = None
        if env_ids is not None:
            assert env_mask is None, "give env_mask or env_ids, not both"
            env_mask = torch.zeros(self.num_envs, dtype=torch.uint8, device=self.device)
            env_mask[torch.as_tensor(env_ids, dtype=torch.long, device=self.device)] = 1
        if env_mask is not None:
            mask = env_mask.to(device=self.device, dtype=torch.uint8).contiguous()
        _lib.check(self._L.mjs_reset(self._h, _ptr(mask), _ptr(self._obs), self._stream()))
        return self._obs_dict(self.observation), {}

    def seed(self, seed: int):
        """dmc2gym.py:126-131: install a new random stream on the live env; applied by the next reset.  Episode counters
        restart, so `reset(seed=s)` twice gives identical episodes (test/test_gym_envs.py:21-36)."""
        self.seed_value = int(seed)
        _lib.check(self._L.mjs_reseed(self._h, int(seed) & (2**64 - 1)))

    # the tensors step()/reset() hand out, in the configured dtype
    @property
    def observation(self):
        return self._obs64 if self._obs64 is not None else self._obs

    @property
    def final_observation(self):
        return self._final_obs64 if self._final_obs64 is not None else self._final_obs

    @property
    def reward(self):
        return self._reward64 if self._reward64 is not None else self._reward

    def step(self, actions: Optional[torch.Tensor]):
        """dmc2gym.py:133-155 for every env.  actions=None -> synthetic random rollout."""
        a = None
        if actions is not None:
            a = actions.to(device=self.device, dtype=torch.float32).contiguous()
            assert a.shape == (self.num_envs, self.action_dim), a.shape
        _lib.check(self._L.mjs_step(self._h, _ptr(a), _ptr(self._obs), _ptr(self._reward), _ptr(self._flags), _ptr(self._final_obs),
                                    int(self.autoreset), self._stream()))
        f = self._flags
        info = {"discount": 1.0 - f[:, 0].float()}
        if self.has_success:
            info["is_success"] = f[:, 2].float()
        if self.autoreset:
            # visual mode: the terminal frame is gone once the kernel has reset the env in the same launch -- state keys only
            info["final_observation"] = self._obs_dict(self.final_observation, images=False)
            info["_final_observation"] = f[:, 3].bool()
        return self._obs_dict(self.observation), self.reward, f[:, 0].bool(), f[:, 1].bool(), info

    def step_flat(self, actions):
        """step() without dict assembly: (obs[N,O], reward[N], flags[N,4] u8)."""
        a = None
        if actions is not None:
            a = actions
        _lib.check(self._L.mjs_step(self._h, _ptr(a), _ptr(self._obs), _ptr(self._reward), _ptr(self._flags), _ptr(self._final_obs),
                                    int(self.autoreset), self._stream()))
        return self.observation, self.reward, self._flags

    def step_host(self, actions: np.ndarray, obs: np.ndarray, reward: np.ndarray, flags: np.ndarray):
        """End-to-end call with HOST numpy buffers (H2D + step + D2H inside the library)."""
        _lib.check(self._L.mjs_step_host(self._h, actions.ctypes.data_as(ctypes.c_void_p) if actions is not None else None, obs.ctypes.data_as(ctypes.c_void_p),
                                         reward.ctypes.data_as(ctypes.c_void_p), flags.ctypes.data_as(ctypes.c_void_p), None, int(self.autoreset)))

    def demonstration_actions(self, noise: float = 0.0):
        """The task's scripted demonstrator on the current device state (mjs_demo_policy): (actions[N,A] f32, ok[N] bool)."""
        import torch

        if not hasattr(self, "_pol_a"):
            self._pol_a = torch.empty(self.num_envs, self.action_dim, dtype=torch.float32, device=self.device)
            self._pol_ok = torch.empty(self.num_envs, dtype=torch.int32, device=self.device)
        _lib.check(self._L.mjs_demo_policy(self._h, float(noise), _ptr(self._pol_a), _ptr(self._pol_ok), self._stream()))
        return self._pol_a, self._pol_ok.bool()

    def render(self, camera: Optional[str] = None, width: Optional[int] = None, height: Optional[int] = None, out: Optional[torch.Tensor] = None):
        """Camera observation of every env (entities/camera.py:94-103 `get_rgb_image`): uint8 [N, H, W, 3], channel-last,
        ray-cast on the device from the task's fixed camera `camera` (default: the task's first) at the task's resolution."""
        from .model.cstruct import MjsVisual

        cams = self.bundle.cameras
        cam = cams[camera] if camera is not None else next(iter(cams.values()))
        w = int(width or self.bundle.image_resolution); h = int(height or self.bundle.image_resolution)
        if out is None:
            out = torch.empty(self.num_envs, h, w, 3, dtype=torch.uint8, device=self.device)
        assert out.shape == (self.num_envs, h, w, 3) and out.dtype == torch.uint8 and out.is_contiguous()
        vis = self.bundle.visuals
        arr = (MjsVisual * max(len(vis), 1))(*vis)
        _lib.check(self._L.mjs_render(self._h, ctypes.byref(cam), arr, len(vis), w, h, _ptr(out), self._stream()))
        return out

    # spaces and facades: same objects the vector-env wrappers expose
    @property
    def single_action_space(self):
        from . import vecenv
        return vecenv.single_action_space(self)

    @property
    def single_observation_space(self):
        from . import vecenv
        return vecenv.single_observation_space(self)

    def as_sb3_vecenv(self, flatten_observation_space: bool = False):
        """stable_baselines3 VecEnv protocol over this env (numpy in / out, `terminal_observation`, `TimeLimit.truncated`)."""
        from . import vecenv
        return vecenv.SB3VecEnvFacade(self, flatten_observation_space)

    def as_gymnasium_vector_env(self, flatten_observation_space: bool = False):
        """gymnasium.vector.VectorEnv protocol over this env (same-step autoreset, batched info arrays)."""
        from . import vecenv
        return vecenv.GymnasiumVectorFacade(self, flatten_observation_space)

    def join_aux(self):
        """Make the current stream wait for background reset preparation in flight."""
        _lib.check(self._L.mjs_join_aux(self._h, self._stream()))

    # state injection / extraction for parity tests
    def set_state(self, states, first: int = 0, forward: bool = True):
        n = len(states)
        _lib.check(self._L.mjs_set_state(self._h, states, first, n, int(forward)))

    def get_state(self, first: int = 0, count: Optional[int] = None):
        count = self.num_envs - first if count is None else count
        arr = (MjsEnvState * count)()
        _lib.check(self._L.mjs_get_state(self._h, arr, first, count))
        return arr

    def status(self) -> torch.Tensor:
        """Per-env status words (MJS_STATUS_* bits: non-finite state, IK unreachable, contact overflow, unsupported pair) as an
        int32 device tensor [N]; the reference raises Python exceptions instead (robots/robot.py:221-224)."""
        if not hasattr(self, "_status"):
            self._status = torch.zeros(self.num_envs, dtype=torch.int32, device=self.device)
        _lib.check(self._L.mjs_status(self._h, _ptr(self._status), self._stream()))
        return self._status

    def stats(self, reset: bool = True):
        out = (ctypes.c_double * 8)()
        _lib.check(self._L.mjs_stats(self._h, out, int(reset)))
        keys = ["return_sum", "length_sum", "episodes", "successes", "truncations", "errors", "substeps", "newton_iterations"]
        return dict(zip(keys, list(out)))

    def stats_tensor(self) -> torch.Tensor:
        """The device-resident [8] f64 accumulator (input of the NCCL all-reduce)."""
        p = ctypes.c_void_p()
        _lib.check(self._L.mjs_stats_device(self._h, ctypes.byref(p)))
        # wrap without copying
        return _wrap_device_f64(p.value, 8, self.device)

    @property
    def launch_count(self) -> int:
        return int(self._L.mjs_launch_count(self._h))


def _wrap_device_f64(ptr: int, n: int, device) -> torch.Tensor:
    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}
    return torch.as_tensor(h, device=device)


def make(task_name: str, num_envs: int, **kw) -> BatchedMujocoSimEnv:
    return BatchedMujocoSimEnv(task_name, num_envs, **kw)

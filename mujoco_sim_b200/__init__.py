"""mujoco_sim_b200 -- B200-native batched simulator for the mujoco_sim composer tasks.

Drop-in scope: the per-control-step hot path behind
/root/reference/mujoco_sim/environments/dmc2gym.py (DMCEnvironmentAdapter).
"""
from .envs import BatchedMujocoSimEnv, make  # noqa: F401
from .vecenv import GymnasiumVectorFacade, SB3VecEnvFacade, make_vec  # noqa: F401

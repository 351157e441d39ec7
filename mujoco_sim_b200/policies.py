"""Batched counterparts of the reference's policy factories.

create_random_policy        point_reach.py:218-225, robot_reach.py:210-219, robot_planar_push.py:244-253,
                            robot_push_button.py:221-229: uniform over the action spec
create_demonstration_policy point_reach.py:227-240, robot_reach.py:222-253, robot_push_button.py:231-300:
                            scripted demonstrators, evaluated by a CUDA kernel on the device state (mjs_demo_policy)

Both return a callable `policy(obs=None) -> actions[N, A]` (float32 CUDA tensor), the batched form of the
reference's `policy(time_step)`.
"""
from __future__ import annotations

from .envs import BatchedMujocoSimEnv


def create_random_policy(env: BatchedMujocoSimEnv, seed: int = 0):
    import torch

    g = torch.Generator(device=env.device)
    g.manual_seed(seed)
    lo = torch.as_tensor(env.single_action_low, dtype=torch.float32, device=env.device)
    hi = torch.as_tensor(env.single_action_high, dtype=torch.float32, device=env.device)

    def random_policy(obs=None):
        return lo + (hi - lo) * torch.rand(env.num_envs, env.action_dim, device=env.device, generator=g)

    return random_policy


def create_demonstration_policy(env: BatchedMujocoSimEnv, noise: float = 0.0):
    if env.task_name == "robot_planar_push":
        raise NotImplementedError("the reference's Planar Push demonstrator is keyboard tele-operation (robot_planar_push.py:288-308)")

    def demonstration_policy(obs=None):
        actions, _ = env.demonstration_actions(noise)
        return actions

    return demonstration_policy

"""Multi-GPU host logic: envs are independent, so the job shards by env range with no
data-path collective; the only exchange is the 8-double episode-statistics vector.

One process per GPU (torchrun).  Works with any torch.distributed backend: NCCL on the GPUs, gloo in the
CPU tests (tests/test_sharding_cpu.py)."""
from __future__ import annotations

import torch
import torch.distributed as dist

STAT_KEYS = ["return_sum", "length_sum", "episodes", "successes", "truncations", "errors", "substeps", "newton_iterations"]


def env_offset(rank: int, envs_per_gpu: int) -> int:
    """Global id of this rank's env 0.  Philox streams are keyed by the global env id, so a job's
    trajectories do not depend on how many GPUs it is spread over."""
    return int(rank) * int(envs_per_gpu)


def owner_of(global_env: int, envs_per_gpu: int) -> tuple[int, int]:
    """(rank, local env id) of a global env id."""
    return int(global_env) // int(envs_per_gpu), int(global_env) % int(envs_per_gpu)


def allreduce_stats(stats: torch.Tensor) -> dict:
    """SUM the per-rank [8] f64 statistics vector (mjs_stats_device) over the job."""
    s = stats.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    return dict(zip(STAT_KEYS, s.cpu().tolist()))


def whole_job_rate(envs_per_gpu: int, steps: int, local_ms: float, device=None) -> tuple[float, float]:
    """(env-steps/s of the whole job, max-over-ranks milliseconds): every rank processed envs_per_gpu * steps."""
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=device)
    world = 1
    if dist.is_available() and dist.is_initialized():
        world = dist.get_world_size()
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    return world * envs_per_gpu * steps / (ms * 1e-3), ms

"""world_size-2 gloo test of the multi-GPU host logic (no GPU needed)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mujoco_sim_b200 import sharding

    n = 4096
    off = sharding.env_offset(rank, n)
    stats = torch.tensor([-1.5 * (rank + 1), 100.0 * (rank + 1), 2.0, float(rank), 1.0, 0.0, n * 10.0, 0.0], dtype=torch.float64)
    tot = sharding.allreduce_stats(stats)
    rate, ms = sharding.whole_job_rate(n, 10, 5.0 + rank)  # rank 1 is slower: the job time is the max
    q.put((rank, off, tot, rate, ms))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_stats_allreduce():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [0, 4096]
    for _, _, tot, rate, ms in res:
        assert tot["return_sum"] == -4.5 and tot["length_sum"] == 300.0 and tot["episodes"] == 4.0 and tot["successes"] == 1.0
        assert tot["substeps"] == 2 * 4096 * 10.0
        assert ms == 6.0 and abs(rate - 2 * 4096 * 10 / 6e-3) < 1e-6


def test_owner_of_roundtrip():
    from mujoco_sim_b200 import sharding

    for g in (0, 1, 4095, 4096, 5000, 8 * 4096 - 1):
        r, l = sharding.owner_of(g, 4096)
        assert sharding.env_offset(r, 4096) + l == g and 0 <= l < 4096

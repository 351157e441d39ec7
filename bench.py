"""bench.py -- env-steps/s of the Planar Push hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W          (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference ...                   (CPU arm: the oracle port on the host cores)

One "step" = one control step (n_sub physics substeps + task hooks + obs; Planar Push: 20) of every env of the shard.
Prints ONE JSON line on rank 0.  See DESIGN.md "Measurement" for how each field is produced.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402

WORKLOADS = {
    # name -> (task, default envs per GPU, algorithmic flops / env-step, algorithmic bytes / env-step)  
    "planar_push": ("robot_planar_push", 42624, 0.86e6, 771.0),  # 148 SMs x 12 FP64 warps x 24 waves (launch tail amortised)
    "point_mass_reach": ("point_mass_reach", 65536, 3.6e3, 151.0),
    "robot_reach": ("robot_reach", 16384, 0.115e6, 283.0),
    "push_button": ("robot_push_button", 4096, 0.58e6, 539.0),  # BASELINE.json configs[4]: 4096 envs per GPU
}
METRIC = "env-steps/s (device-timed) Planar Push"
SMALL_ENVS = 4096   # BASELINE.json configs[3] names "4096 envs per GPU": reported beside the headline as config_4096
STATS_EVERY = 10    # control steps between two NCCL all-reduces of the statistics vector inside the timed region
KERNEL_SOURCES = ["arm_kernel.cuh", "arm_model.cuh", "arm_launch.cuh", "arm_sim.cuh", "mjs_common.cuh", "arm.cu", "../../include/mjs_convex.h", "../../include/mjs_model.h"]
MIXED_ENVS = 37888  # 148 SMs x 16 FP32 warps x 16 waves
FP32_PEAK_TFLOPS_FALLBACK = 148 * 128 * 2 * 1.965e9 / 1e12  # SMs x FP32 lanes x FMA x clocks.max.sm = 74.4


def _peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d.get("hbm_gbs", 6650.0)), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


def _fma_peak(precision: str, sm_max: float):
    """Non-tensor FMA peak in TFLOP/s: measured on this pool's B200 by tools/fma_peak.cu (register-resident FMA chains on
    all SMs, one launch >= 1 s; profiles/r2/fma_peak.json), else the nominal SMs x lanes x 2 x clock figure."""
    lanes = 64 if precision == "f64" else 128
    p = ROOT / "profiles" / "r2" / "fma_peak.json"
    if p.exists():
        d = json.loads(p.read_text())
        key = "fp64_fma_tflops" if precision == "f64" else "fp32_fma_tflops"
        return float(d[key]), f"measured: tools/fma_peak.cu on {d.get('gpu', 'B200')} ({d.get('sms', 148)} SMs), profiles/r2/fma_peak.json"
    return 148 * lanes * 2 * sm_max * 1e6 / 1e12, f"nominal: 148 SMs x {lanes} lanes x 2 x {sm_max:.0f} MHz (profiles/r2/fma_peak.json missing)"


def kernel_source_hash():
    import hashlib

    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        h.update((ROOT / "mujoco_sim_b200" / "csrc" / f).read_bytes())
    return h.hexdigest()[:16]


def _traffic(precision: str, N: int):
    """DRAM bytes per launch of the step kernel from the committed ncu --set full capture of this bench command
    (profiles/r2/traffic.json, written by tools/ncu_traffic.py from the raw ncu page).  The capture names the hash of the
    kernel sources it was taken from: a capture of other sources is STALE and is not reported."""
    p = ROOT / "profiles" / "r2" / "traffic.json"
    if not p.exists():
        return None, "no ncu capture committed for this round (profiles/r2/traffic.json missing)"
    d = json.loads(p.read_text()).get(precision)
    if not d:
        return None, f"profiles/r2/traffic.json has no {precision} capture"
    cur = kernel_source_hash()
    if d.get("kernel_source_hash") != cur:
        print(f"bench.py: ncu traffic capture is STALE (sources {d.get('kernel_source_hash')}, now {cur}); re-run tools/ncu_bench.sh", file=sys.stderr)
        return None, f"STALE: capture {d.get('file')} is of kernel sources {d.get('kernel_source_hash')}, the library now builds from {cur}"
    per_env = float(d["dram_bytes_per_launch"]) / float(d["envs"])
    return per_env * N, (f"dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full launch of this bench command: {d['dram_bytes_per_launch'] / 1e6:.1f} MB at "
                         f"{d['envs']} envs = {per_env:.0f} B/env ({d.get('file')}, kernel sources {cur}), scaled to {N} envs; algorithmic 771 B/env")


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Polls nvidia-smi (the recipe's clocks line) while the timed region runs."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "power_w_max": max(float(r[2]) for r in self.rows),
                "samples": len(self.rows), "reasons": reasons}


# ----------------------------------------------------------------------------- CPU arm
def _oracle_worker(bundle, states, actions, steps, tol_scale, out, idx, burn=0):
    import ctypes

    from oracle.oracle import OracleEnvs

    n = len(states)
    ora = OracleEnvs(bundle, n)
    ctypes.memmove(ora.states, states, ctypes.sizeof(states))
    ora.tol_scale = tol_scale
    for k in range(burn):  # untimed: age the scenes to the mean age of the GPU arm's stationary mix
        ora.step_raw(actions[k % len(actions)])
    t0 = time.perf_counter()
    for k in range(steps):
        ora.step_raw(actions[k])
    out[idx] = time.perf_counter() - t0


def cpu_oracle_rate(task: str, envs_per_core: int, steps: int, cores: int | None = None, seed: int = 0, burn: int = -1):
    """Time the double-precision CPU oracle (oracle/mjs_oracle.cpp, MuJoCo's default solver tolerance)
    on `cores` host threads (ctypes releases the GIL), `envs_per_core` envs each, `steps` control steps
    of the same workload (same scene, same action box, settled initial states)."""
    import mujoco_sim_b200.envs as _envs
    from tests.helpers import copy_states, pointmass_states, push_states

    cores = cores or len(os.sched_getaffinity(0))
    bundle = _envs._BUILDERS[task]()
    t = bundle.task
    A = t.action_dim
    n = envs_per_core * cores
    if task == "robot_planar_push":
        init = push_states(bundle, n, seed).states
    elif task == "point_mass_reach":
        init = pointmass_states(n, seed)
    else:
        raise SystemExit(f"cpu baseline: no host state generator for {task}")
    rng = np.random.RandomState(seed + 1)
    lo = np.array([t.action_lo[k] for k in range(A)]); hi = np.array([t.action_hi[k] for k in range(A)])
    if burn < 0:
        burn = t.max_steps // 2 if task == "robot_planar_push" else 0  # mean scene age of the stationary mix
    threads, out = [], [0.0] * cores
    for c in range(cores):
        chunk = copy_states((type(init[0]) * envs_per_core).from_address(__import__("ctypes").addressof(init) + c * envs_per_core * __import__("ctypes").sizeof(init[0])), envs_per_core)
        acts = rng.uniform(lo, hi, (steps, envs_per_core, A))
        threads.append(threading.Thread(target=_oracle_worker, args=(bundle, chunk, acts, steps, 1.0, out, c, burn)))
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    wall = max(out)  # slowest thread's timed loop (all threads run concurrently; the burn-in before it is untimed)
    return n * steps / wall, cores, f"{n} envs ({envs_per_core}/thread x {cores} threads) x {steps} control steps after {burn} untimed burn-in steps, {wall:.1f} s"


# ----------------------------------------------------------------------------- GPU arm
def _dist_setup(n_gpus: int):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    elif n_gpus > 1:
        raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    return world, rank, local


def _stagger_episodes(env, max_steps: int, seed: int):
    """Spread the episode phase of the envs uniformly over [0, max_steps) so the timed region sees the
    steady-state reset rate (1/max_steps per env-step) instead of zero resets."""
    if max_steps <= 1:
        return
    st = env.get_state()
    rng = np.random.RandomState(seed)
    ph = rng.randint(0, max_steps, len(st))
    for s, p in zip(st, ph):
        s.episode_step = int(p)
    env.set_state(st, forward=False)


def run_gpu(args):
    import torch
    import torch.distributed as dist

    import mujoco_sim_b200 as ms
    from mujoco_sim_b200 import sharding

    world, rank, local = _dist_setup(args.gpus)
    task, n_default, flops_per, bytes_per = WORKLOADS[args.workload]
    N = args.envs or n_default
    if not args.envs and task == "robot_planar_push" and args.precision == "mixed":
        N = MIXED_ENVS
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    env = ms.make(task, N, device=dev, seed=args.seed, env_offset=sharding.env_offset(rank, N), precision=args.precision)
    env.reset()
    burn_in = args.burn_in
    if task == "robot_planar_push":
        _stagger_episodes(env, env.bundle.task.max_steps, args.seed + rank)
        if burn_in < 0:
            burn_in = env.bundle.task.max_steps + 50
    burn_in = max(burn_in, 0)
    A, O = env.action_dim, env.obs_dim
    lo = torch.tensor([env.bundle.task.action_lo[k] for k in range(A)], device=dev)
    hi = torch.tensor([env.bundle.task.action_hi[k] for k in range(A)], device=dev)
    g = torch.Generator(device=dev); g.manual_seed(args.seed + 17 * rank)
    # Action batches are NEVER replayed cyclically: a short cycle of pusher targets sweeps the same path over and over, clears
    # it of props and settles into a cheaper regime than uniform random actions (profiles/r2/step_trace*.txt: 16.7 vs 18.9 ms per
    # step with a 16-batch ring).  Untimed steps draw a new batch per step; timed regions read pre-generated distinct batches.
    ring = min(max(args.steps, 40), 1024)  # distinct batches resident in HBM for the timed regions (period >= the region or >= 2 episodes)

    def draw(n_envs):
        return (lo + (hi - lo) * torch.rand(n_envs, A, device=dev, generator=g)).float()

    def batches(n_envs):
        return (lo + (hi - lo) * torch.rand(ring, n_envs, A, device=dev, generator=g)).float().contiguous()

    actions = batches(N) if args.workload == "planar_push" else None
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream(dev)
    host_actions = args.workload == "planar_push"
    if not host_actions:
        # the other workloads draw their synthetic actions on the device (Philox; Button Push: q_now +- 0.05 rad,
        # SURVEY.md 8d) -- mjs_step(actions = NULL); e2e still copies results to the host every step
        class _NoneRing:
            def __getitem__(self, i): return None
        actions = _NoneRing()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    side = torch.cuda.Stream(dev)
    stat_buf = torch.zeros(8, dtype=torch.float64, device=dev)
    n_allreduce = [0]

    def stats_allreduce(e):
        """The path's only collective: snapshot the 8-double statistics vector on a side stream and all-reduce
        it over NCCL while the step kernels keep running; joined before the region closes."""
        side.wait_stream(stream)
        with torch.cuda.stream(side):
            stat_buf.copy_(e.stats_tensor())
            if world > 1:
                dist.all_reduce(stat_buf)
        n_allreduce[0] += 1

    def timed_region(step_fn, steps, e=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            flush.zero_()          # L2 flush between timed iterations (inside the region: ~2% pessimistic)
            step_fn(i)
            if e is not None and (i + 1) % STATS_EVERY == 0:
                stats_allreduce(e)
        env.join_aux()             # background reset preparation launched by these steps is part of the job
        stream.wait_stream(side)   # ... and so are the statistics all-reduces
        e1.record(stream)
        barrier()
        ms_ = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms_, op=dist.ReduceOp.MAX)
        return float(ms_.item())

    # ---- burn-in (untimed, not part of W): one full episode length, so that every env has gone through its
    # staggered reset and the population of physical states is stationary (fresh scenes cost ~10 % more per step
    # than the steady-state mix; a region timed right after reset() would drift)
    for i in range(burn_in + args.warmup):  # burn-in, then the W warm-up steps of the device-resident arm ("value")
        env.step_flat(draw(N) if host_actions else None)
    torch.cuda.synchronize(dev)  # the clock sampler below should see the timed region, not the tail of the burn-in
    l0 = env.launch_count
    with ClockSampler(local) as clk:
        t_ms = timed_region(lambda i: env.step_flat(actions[i % ring]), args.steps, env)
    launches = env.launch_count - l0
    n_stats = n_allreduce[0]
    # flush cost alone, for the record
    barrier(); f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(stream)
    for _ in range(10):
        flush.zero_()
    f1.record(stream); torch.cuda.synchronize(dev)
    flush_ms = f0.elapsed_time(f1) / 10
    # ---- dominant kernel alone: average per-launch duration, CUDA events on the launching stream around every
    # launch (no flush; the 2 us job-list kernel that precedes each step kernel is inside the bracket)
    barrier()
    kn = max(5, min(args.steps, 40))
    env.join_aux(); torch.cuda.synchronize(dev)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(kn + 1)]
    evs[0].record(stream)
    k_actions = batches(N)[:kn] if host_actions else actions
    for i in range(kn):
        env.step_flat(k_actions[i])
        evs[i + 1].record(stream)
    torch.cuda.synchronize(dev)
    per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(kn))
    kernel_ms = sum(per) / len(per)

    # ---- end-to-end arm: HOST buffers through the C-ABI host call (H2D + step + D2H + sync inside)
    h_act = torch.empty(ring, N, A, dtype=torch.float32).pin_memory()
    if host_actions:
        h_act.copy_(batches(N).cpu())
    h_obs = torch.empty(N, O, dtype=torch.float32).pin_memory()
    h_rew = torch.empty(N, dtype=torch.float32).pin_memory()
    h_flg = torch.empty(N, 4, dtype=torch.uint8).pin_memory()
    a_np, o_np, r_np, f_np = h_act.numpy(), h_obs.numpy(), h_rew.numpy(), h_flg.numpy()
    for i in range(3):
        env.step_host(a_np[i % ring] if host_actions else None, o_np, r_np, f_np)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        env.step_host(a_np[(i + 3) % ring] if host_actions else None, o_np, r_np, f_np)
    e2e_s = time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_s = float(e2e_t.item())

    # ---- secondary: the FP32-dynamics variant (precision="mixed"), same workload and timing method.
    # Not the headline: it misses the 1e-4 single-step gate on impact steps (tests/test_arm_gpu.py).
    alt = None
    if args.precision == "f64" and not args.no_alt and host_actions:
        N2 = N if args.envs else MIXED_ENVS
        env2 = ms.make(task, N2, device=dev, seed=args.seed, env_offset=rank * N, precision="mixed")
        env2.reset()
        if task == "robot_planar_push":
            _stagger_episodes(env2, env2.bundle.task.max_steps, args.seed + rank)
        act2 = batches(N2)
        for i in range(burn_in + args.warmup):
            env2.step_flat(draw(N2))
        main_env, env = env, env2
        t2 = timed_region(lambda i: env2.step_flat(act2[i % ring]), args.steps)
        env = main_env
        env2.close()
        alt = {"precision": "mixed", "envs_per_gpu": N2, "value": world * N2 * args.steps / (t2 * 1e-3), "unit": "env-steps/s", "ms_per_step": t2 / args.steps,
               "note": "FP32 dynamics/solver + FP64 state; misses the 1e-4 single-step gate on impact steps, reported only"}

    # ---- the same kernel at BASELINE.json's own batch size (4096 envs per GPU): a 2.3-wave launch, tail-dominated
    small = None
    if host_actions and not args.envs and not args.no_small:
        env3 = ms.make(task, SMALL_ENVS, device=dev, seed=args.seed, env_offset=sharding.env_offset(rank, SMALL_ENVS), precision=args.precision)
        env3.reset()
        _stagger_episodes(env3, env3.bundle.task.max_steps, args.seed + rank)
        act3 = batches(SMALL_ENVS)
        for i in range(burn_in + args.warmup):
            env3.step_flat(draw(SMALL_ENVS))
        main_env, env = env, env3
        t3 = timed_region(lambda i: env3.step_flat(act3[i % ring]), args.steps)
        env = main_env
        env3.close()
        v3 = world * SMALL_ENVS * args.steps / (t3 * 1e-3)
        small = {"envs_per_gpu": SMALL_ENVS, "value": v3, "unit": "env-steps/s", "ms_per_step": t3 / args.steps, "precision": args.precision,
                 "per_env_rate_vs_headline_batch": v3 / (world * N * args.steps / (t_ms * 1e-3)),
                 "note": "BASELINE.json configs[3] size; same kernel, same timing method (L2 flush inside the region)"}

    # ---- episode statistics: the only collective of the path
    stats = sharding.allreduce_stats(env.stats_tensor())

    value = world * N * args.steps / (t_ms * 1e-3)
    out = None
    if rank == 0:
        hbm_peak, peak_src, sm_max = _peaks()
        kern_rate = N / (kernel_ms * 1e-3)
        fp32_peak, fma_src = _fma_peak(args.precision, sm_max)
        bound = "fp64" if args.precision == "f64" else "fp32"
        traffic, traffic_note = _traffic(args.precision, N) if args.workload == "planar_push" else (None, "not captured for this workload")
        out = {
            "metric": METRIC if args.workload == "planar_push" else f"env-steps/s (device-timed) {args.workload}",
            "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 dynamics/solver + f64 state integration" if args.precision == "mixed" else "f64",
            "data": "synthetic (device reset kernel initial states, uniform random actions in the reachable workspace box)",
            "config": {"workload": (f"{args.workload}: UR5e + cylinder pusher + 3 cylinder props, {env.bundle.task.n_sub} substeps/step, {N} envs/GPU" if host_actions
                                    else f"{args.workload}: {env.bundle.task.n_sub} substeps/step, {N} envs/GPU, device-drawn actions"),
                       "envs_per_gpu": N, "precision": args.precision, "l2": f"flushed between timed iterations (256 MiB memset inside the region, {flush_ms:.3f} ms each)",
                       "episodes": f"phases staggered over [0,max_steps) + {burn_in} untimed burn-in steps: stationary state mix and auto-reset rate",
                       "actions": "a new uniform random batch every control step, burn-in included (no cyclic replay)" if host_actions else "drawn on the device per step (Philox)"},
            "e2e": {"value": world * N * args.steps / e2e_s, "unit": "env-steps/s", "h2d_bytes_per_step": N * A * 4 if host_actions else 0,
                    "d2h_bytes_per_step": N * (O + 1) * 4 + N * 4},
            "gpu_launches": int(launches),
            "clocks": clk.summary(),
            "roofline": {"bound": bound, "achieved": kern_rate * flops_per / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": kern_rate * flops_per / 1e12 / fp32_peak,
                         "traffic": traffic, "traffic_note": traffic_note,
                         "kernel": "arm_step_kernel", "kernel_ms": kernel_ms, "kernel_ms_median": per[len(per) // 2], "flops_per_env_step": flops_per,
                         "peak_source": fma_src,
                         "hbm": {"achieved": kern_rate * bytes_per / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": kern_rate * bytes_per / 1e9 / hbm_peak,
                                 "peak_source": peak_src, "bytes_per_env_step": bytes_per}},
            "alt_precision": alt,
            "config_4096": small,
            "stats_allreduce": {"every_steps": STATS_EVERY, "count_in_timed_region": n_stats, "bytes": 64,
                                "how": "NCCL all_reduce(SUM) of the 8-double statistics vector on a side stream, overlapping the step kernels; joined before the closing event"},
            "episode_stats": {k: stats[k] for k in sharding.STAT_KEYS},
        }
    env.close()
    return out, world, rank


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """`--impl reference`: the reference's CPU path on the host cores.  The reference's arithmetic is
    MuJoCo's C engine behind dm_control, neither vendored nor installed here or on the GPU box (no
    network), so this times the repo's double-precision CPU restatement (oracle/, kind "port") with one
    thread per host core.  Each step is a bounded sample (envs_per_core envs per thread)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    task = WORKLOADS[args.workload][0]
    epc = args.cpu_envs_per_core
    for _ in range(1):
        cpu_oracle_rate(task, epc, max(1, min(args.warmup, 3)), burn=0)  # warm-up (page-in, thread start)
    rate, cores, sample = cpu_oracle_rate(task, epc, args.steps)
    n = epc * cores
    return {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * n / rate, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic (oracle-settled initial states, uniform random actions in the reachable workspace box)",
        "config": {"workload": f"{args.workload}: UR5e + cylinder pusher + 3 cylinder props, 20 substeps/step; bounded sample of {n} envs per step",
                   "note": "MuJoCo/dm_control unavailable on this box: CPU restatement (oracle port, g++ -O3, dense single-env code), Newton solver at MuJoCo's default tolerance 1e-8; NOT the reference's own engine"},
        "cpu_baseline": {"value": rate, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="planar_push", choices=list(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: the workload's)")
    ap.add_argument("--precision", default="f64", choices=["mixed", "f64"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--burn-in", type=int, default=-1, help="untimed steps before warm-up (default: one episode length + 50 for Planar Push, else 0)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-alt", action="store_true", help="skip the secondary precision=mixed measurement")
    ap.add_argument("--no-small", action="store_true", help="skip the 4096-envs-per-GPU record (config_4096)")
    ap.add_argument("--cpu-envs-per-core", type=int, default=64)
    ap.add_argument("--cpu-steps", type=int, default=300)
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 200:
            args.steps = 200  # keep the CPU arm within minutes
        out = run_reference(args)
        if out is not None:
            print(json.dumps(out), flush=True)
        return
    args.warmup = max(args.warmup, 3)
    out, world, rank = run_gpu(args)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline and args.workload in ("planar_push", "point_mass_reach"):
            task = WORKLOADS[args.workload][0]
            rate, cores, sample = cpu_oracle_rate(task, args.cpu_envs_per_core, args.cpu_steps)
            out["cpu_baseline"] = {"value": rate, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

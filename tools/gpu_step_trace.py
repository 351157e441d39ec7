"""Per-step duration of the Planar Push step kernel over a long run of the bench workload (same envs, stagger and action
ring as bench.py): is the state mix stationary after the burn-in?  Prints the mean over 25-step windows.
usage (GPU box): python tools/gpu_step_trace.py [steps=900] [ring|fresh] > gpurun_out/step_trace.txt
"ring" replays 16 pre-generated action batches cyclically (what bench.py did until round 2), "fresh" draws a new batch every step."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
import mujoco_sim_b200 as ms  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 900
mode = sys.argv[2] if len(sys.argv) > 2 else "ring"
N = bench.WORKLOADS["planar_push"][1]
dev = torch.device("cuda", 0)
env = ms.make("robot_planar_push", N, device=dev, seed=0, precision="f64")
env.reset()
bench._stagger_episodes(env, env.bundle.task.max_steps, 0)
A = env.action_dim
lo = torch.tensor([env.bundle.task.action_lo[k] for k in range(A)], device=dev)
hi = torch.tensor([env.bundle.task.action_hi[k] for k in range(A)], device=dev)
g = torch.Generator(device=dev); g.manual_seed(0)
actions = (lo + (hi - lo) * torch.rand(16, N, A, device=dev, generator=g)).float().contiguous()
stream = torch.cuda.current_stream(dev)
evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
evs[0].record(stream)
for i in range(steps):
    if mode == "fresh":
        a = (lo + (hi - lo) * torch.rand(N, A, device=dev, generator=g)).float()
    else:
        a = actions[i % 16]
    env.step_flat(a)
    evs[i + 1].record(stream)
torch.cuda.synchronize(dev)
per = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
print(f"# actions: {mode}; envs {N}, max_steps {env.bundle.task.max_steps}, steps {steps}; mean ms per step over windows of 25")
for w in range(0, steps, 25):
    c = per[w:w + 25]
    print(f"{w:5d} {sum(c) / len(c):8.3f}  min {min(c):7.3f} max {max(c):7.3f}")

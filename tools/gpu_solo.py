"""Does a warp run faster with fewer neighbours?  arm_substeps_kernel (4-warp blocks, no lock-step): 20 substeps at
1 block/SM (4 warps/SM), 2 blocks/SM (8) and 3 blocks/SM (12 warps/SM)."""
import sys, time, numpy as np, torch
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
for N in (148 * 4, 148 * 8, 148 * 12, 148 * 24):
    env = ms.make("robot_planar_push", N, precision="f64", seed=0)
    env.reset()
    dev = torch.device("cuda:0")
    lo = torch.tensor([-0.2, -0.6], device=dev); hi = torch.tensor([0.2, -0.3], device=dev)
    for i in range(60):
        env.step_flat((lo + (hi - lo) * torch.rand(N, 2, device=dev)).float())
    torch.cuda.synchronize()
    ts = []
    for r in range(12):
        t0 = time.perf_counter(); env._L.mjs_debug_substeps(env._h, 20); ts.append(time.perf_counter() - t0)
    ts = np.array(ts[2:]) * 1e3
    print(f"N {N:6d} ({N // 148} warps/SM resident up to 12): 20 substeps median {np.median(ts):.3f} ms  min {ts.min():.3f}  -> {N / np.median(ts) * 1e3 / 1e6:.3f} M env-steps/s")
    env.close()

"""Per-env / per-block timing of one step launch (mjs_debug_times): where does the launch's time go?"""
import ctypes, sys, numpy as np, torch
sys.path.insert(0, ".")
import mujoco_sim_b200 as ms
from bench import _stagger_episodes
N = int(sys.argv[1]) if len(sys.argv) > 1 else 5328
prec = sys.argv[2] if len(sys.argv) > 2 else "f64"
burn = int(sys.argv[3]) if len(sys.argv) > 3 else 600
WPB = 12 if prec == "f64" else 16
env = ms.make("robot_planar_push", N, precision=prec, seed=0)
env.reset(); _stagger_episodes(env, env.bundle.task.max_steps, 0)
dev = torch.device("cuda:0")
lo = torch.tensor([-0.2, -0.6], device=dev); hi = torch.tensor([0.2, -0.3], device=dev)
acts = (lo + (hi - lo) * torch.rand(16, N, 2, device=dev)).float().contiguous()
for i in range(burn): env.step_flat(acts[i % 16])
buf = (ctypes.c_longlong * (16 * N))()
env._L.mjs_debug_times(env._h, buf, 16 * N)  # arm
for rep in range(3):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); env.step_flat(acts[rep]); e1.record(); torch.cuda.synchronize()
    env._L.mjs_debug_times(env._h, buf, 16 * N)
    t = np.frombuffer(buf, dtype=np.int64).reshape(N, 16).copy()
    base = t[:, 0].min()
    start, mid, end = (t[:, 0] - base) / 1e6, (t[:, 1] - base) / 1e6, (t[:, 2] - base) / 1e6
    nit = t[:, 3] >> 32; nmax = (t[:, 3] >> 16) & 0xffff; sm = t[:, 3] & 0xffff
    dur = mid - start
    nb = (N + WPB - 1) // WPB
    bdur = np.array([dur[b * WPB:(b + 1) * WPB].max() for b in range(nb)])
    bstart = np.array([start[b * WPB:(b + 1) * WPB].min() for b in range(nb)])
    bend = np.array([end[b * WPB:(b + 1) * WPB].max() for b in range(nb)])
    print(f"launch {rep}: event ms {e0.elapsed_time(e1):.3f}  last end {end.max():.3f}  warp dur mean {dur.mean():.3f} p50 {np.median(dur):.3f} p99 {np.percentile(dur,99):.3f} max {dur.max():.3f}")
    print(f"   block dur: mean {bdur.mean():.3f} min {bdur.min():.3f} p50 {np.median(bdur):.3f} p90 {np.percentile(bdur,90):.3f} max {bdur.max():.3f};  post-step tail (end-mid) max {(end-mid).max():.3f}")
    # per SM: busy time and finish time
    fin = np.zeros(sm.max() + 1); busy = np.zeros(sm.max() + 1); cnt = np.zeros(sm.max() + 1)
    for b in range(nb):
        s = sm[b * WPB]; fin[s] = max(fin[s], bend[b]); busy[s] += bend[b] - bstart[b]; cnt[s] += 1
    used = cnt > 0
    print(f"   per SM: blocks {cnt[used].min():.0f}..{cnt[used].max():.0f}  finish mean {fin[used].mean():.3f} min {fin[used].min():.3f} max {fin[used].max():.3f}  busy mean {busy[used].mean():.3f}")
    # by start wave
    order = np.argsort(bstart)
    for w in range(0, nb, 148):
        idx = order[w:w + 148]
        print(f"   blocks {w}-{w+len(idx)-1} by start: start {bstart[idx].min():.3f}..{bstart[idx].max():.3f} dur mean {bdur[idx].mean():.3f} max {bdur[idx].max():.3f}")
    c = np.corrcoef(dur, nit)[0, 1]
    print(f"   corr(warp dur, newton its) {c:.2f}; nit_sum mean {nit.mean():.1f} max {nit.max()}  nit_max p99 {np.percentile(nmax,99):.0f} max {nmax.max()}")

import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
task = sys.argv[1]; n = int(sys.argv[2]); steps = int(sys.argv[3]); precs = sys.argv[4].split(",")
autoreset = (sys.argv[5] != "0") if len(sys.argv) > 5 else True
for prec in precs:
    env = ms.make(task, n, precision=prec, autoreset=autoreset)
    t = time.time(); env.reset(); torch.cuda.synchronize(); print(task, prec, "reset s", round(time.time() - t, 4), flush=True)
    for _ in range(3): env.step_flat(None)
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        env.step_flat(None); evs[i + 1].record()
    torch.cuda.synchronize()
    ts = np.array([evs[i].elapsed_time(evs[i + 1]) for i in range(steps)])
    ms_ = ts.mean()
    st = env.get_state(0, min(n, 256))
    status = np.bincount([s.status for s in st], minlength=16)
    print(task, prec, n, "ms/step mean", round(ms_, 4), "min", round(ts.min(), 4), "med", round(float(np.median(ts)), 4), "max", round(ts.max(), 4),
          "env-steps/s(mean)", round(n / ms_ * 1e3), "(median)", round(n / np.median(ts) * 1e3), "episodes", env.stats()["episodes"], "status", status[:4].tolist(), flush=True)
    env.close()

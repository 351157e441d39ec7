import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
import mujoco_sim_b200 as ms
task = sys.argv[1]; n = int(sys.argv[2]); steps = int(sys.argv[3]); precs = sys.argv[4].split(",")
for prec in precs:
    env = ms.make(task, n, precision=prec)
    t = time.time(); env.reset(); torch.cuda.synchronize(); print(task, prec, "reset s", time.time() - t, flush=True)
    for _ in range(3): env.step_flat(None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): env.step_flat(None)
    e1.record(); torch.cuda.synchronize()
    ms_ = e0.elapsed_time(e1) / steps
    st = env.get_state(0, min(n, 256))
    status = np.bincount([s.status for s in st], minlength=16)
    print(task, prec, n, "ms/step", round(ms_, 4), "env-steps/s", round(n / ms_ * 1e3), env.stats(), "status hist", status.tolist(), flush=True)
    env.close()

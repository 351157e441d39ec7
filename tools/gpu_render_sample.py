"""Write sample camera observations of every task as PPM files (headless: the device buffer goes to a file)."""
import sys
sys.path.insert(0, ".")
import os
import mujoco_sim_b200 as ms
os.makedirs("gpurun_out/render", exist_ok=True)
for task, cam in [("point_mass_reach", "top_down"), ("robot_reach", "front_tilted"), ("robot_planar_push", "front_tilted"), ("robot_planar_push", "top_down"), ("robot_push_button", "scene"), ("robot_push_button", "wrist")]:
    env = ms.make(task, 4, seed=5); env.reset()
    for _ in range(3): env.step(None)
    img = env.render(cam, 192, 192).cpu().numpy()[0]
    with open(f"gpurun_out/render/{task}_{cam}.ppm", "wb") as f:
        f.write(b"P6\n192 192\n255\n" + img.tobytes())
    import torch, time
    big = ms.make(task, 4096, seed=5); big.reset(); out = big.render(cam); torch.cuda.synchronize()
    t = time.time()
    for _ in range(10): big.render(cam, out=out)
    torch.cuda.synchronize()
    r = big.bundle.image_resolution
    print(task, cam, f"{r}x{r}: {4096 * 10 / (time.time() - t):.0f} frames/s at 4096 envs")
    big.close(); env.close()

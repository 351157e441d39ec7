"""Run bench.py for one (library build, precision, envs) variant and print a compact line."""
import json, os, subprocess, sys
lib, prec, envs = sys.argv[1], sys.argv[2], sys.argv[3]
extra = sys.argv[4:]
env = dict(os.environ)
if lib != "-": env["MJS_LIB"] = lib
out = subprocess.run([sys.executable, "bench.py", "--steps", "40", "--warmup", "5", "--precision", prec, "--envs", envs, "--no-cpu-baseline", "--no-alt"] + extra,
                     env=env, capture_output=True, text=True, timeout=600)
try:
    d = json.loads(out.stdout.strip().splitlines()[-1])
    k = d["roofline"]["kernel_ms"]
    print(f"{os.path.basename(lib):28s} {prec:5s} envs {envs:>6s}: value {d['value']:>10.0f} env-steps/s  ms/step {d['ms_per_step']:.3f}  kernel_ms {k:.3f}  kernel-only {int(envs)/k*1e3:>10.0f}  e2e {d['e2e']['value']:.0f}", flush=True)
except Exception as ex:
    print("FAILED", lib, prec, envs, ex, out.stderr[-800:], flush=True)

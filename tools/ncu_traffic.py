"""Turn the raw page of an `ncu --set full` capture of the bench command (tools/ncu_bench.sh -> gpurun_out/prof_<P>_raw.csv) into
the entry of profiles/r2/traffic.json that bench.py's roofline.traffic reads.  The entry names the hash of the kernel sources
of the working tree, so a later kernel change makes bench.py report the capture as stale.
Usage: python tools/ncu_traffic.py <precision> <raw.csv> <envs> [<copy-to profiles/r2/...>]"""
import csv
import json
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from bench import kernel_source_hash  # noqa: E402

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    prec, raw, envs = sys.argv[1], Path(sys.argv[2]), int(sys.argv[3])
    rows = list(csv.reader(open(raw)))
    head, units, vals = rows[0], rows[1], rows[2]
    col = {n: i for i, n in enumerate(head)}

    def get(name):
        i = col[name]
        return float(vals[i].replace(",", "")) * UNIT.get(units[i], 1.0)

    keep = (ROOT / sys.argv[4]).resolve() if len(sys.argv) > 4 else None
    if keep:
        shutil.copy(raw, keep)
    out_path = ROOT / "profiles" / "r2" / "traffic.json"
    data = json.loads(out_path.read_text()) if out_path.exists() else {}
    try:
        git = subprocess.run(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, capture_output=True, text=True).stdout.strip()
    except Exception:
        git = "?"
    data[prec] = {
        "dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
        "dram_read": get("dram__bytes_read.sum"), "dram_write": get("dram__bytes_write.sum"),
        "envs": envs, "kernel": vals[col["Kernel Name"]], "grid": vals[col["launch__grid_size"]],
        "kernel_source_hash": kernel_source_hash(), "git_parent": git,
        "file": str(keep.relative_to(ROOT)) if keep else str(raw),
    }
    out_path.write_text(json.dumps(data, indent=1) + "\n")
    print(json.dumps(data[prec]))


if __name__ == "__main__":
    main()

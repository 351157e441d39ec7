"""bench.py pieces that need no GPU: the reference arm's JSON line (CPU oracle port, bounded sample) and the ncu traffic record."""
import argparse
import json

import bench


def test_reference_arm_line_has_the_contract_keys():
    args = argparse.Namespace(workload="planar_push", cpu_envs_per_core=1, steps=2, warmup=1, gpus=1)
    line = bench.run_reference(args)
    json.dumps(line)
    assert line["impl"] == "reference" and line["metric"] == bench.METRIC and line["unit"] == "env-steps/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["steps"] == 2 and line["warmup"] == 1 and line["n_gpus"] == 1 and line["gpu_launches"] == 0


def test_committed_ncu_traffic_record_matches_the_kernel_sources():
    """roofline.traffic is only printed when profiles/r2/traffic.json was captured from the kernel sources of this tree."""
    rec = json.loads((bench.ROOT / "profiles" / "r2" / "traffic.json").read_text())["f64"]
    assert rec["kernel_source_hash"] == bench.kernel_source_hash(), "re-run tools/ncu_bench.sh + tools/ncu_traffic.py after changing the kernels"
    traffic, note = bench._traffic("f64", 42624)
    assert traffic == rec["dram_bytes_per_launch"] and "STALE" not in note and (bench.ROOT / rec["file"]).exists()

"""bench.py's reference arm needs no GPU (it times the reference's own Python from oracle/_ref, one process per host
core, or the C restatement when those sources are absent): its one JSON line carries the keys the driver reads,
stdout carries nothing else, and the product library is not loaded by it."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0", "--ref-pairs", "1500", "--cpu-pairs", "100000"], capture_output=True, text=True, cwd=ROOT, timeout=280)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "pairs/s" and d["higher_is_better"] is True
    assert d["steps"] == 1 and d["n_gpus"] == 1 and d["value"] > 0 and d["dtype"] == "u8"
    assert d["config"]["config"] == "cfg3" and "dual index" in d["metric"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert d["cpu_baseline_port"]["kind"] == "port" and d["cpu_baseline_port"]["value"] > 0
    assert set(d["config"]) == {"workload", "config", "pairs_per_step_per_gpu", "min_qual", "l2", "parallelism"}


def test_reference_arm_does_not_map_the_product_library():
    code = ("import sys, os; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0', '--ref-pairs', '500', "
            "'--cpu-pairs', '100000']; sys.path.insert(0, %r); import bench; bench.main(); "
            "maps = open('/proc/self/maps').read(); "
            "sys.stderr.write('MAPPED ' + str('libquade_b200' in maps) + ' ' + str('libquade_synth' in maps) + chr(10))" % ROOT)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT, timeout=280)
    assert p.returncode == 0, p.stderr[-2000:]
    assert "MAPPED False True" in p.stderr, p.stderr[-500:]

"""bench.py contract checks that need no GPU: the reference arm's JSON line and the files the CUDA arm points at."""
import json
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "muscl_ref")


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/muscl_ref not built (needs /root/reference)")
def test_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` times the real reference program on a bounded sample and prints one JSON line
    with the keys the driver reads (tier rule 4: impl, cpu_baseline, e2e with zero copy bytes)."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                        "--warmup", "1", "--ref-n", "96"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "cell_updates_per_s" and d["unit"] == "cell-updates/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["dtype"] == "f64"
    assert d["value"] > 0 and d["ms_per_step"] > 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] == 1 and cb["value"] == d["value"] and "96x96" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=120, cwd=ROOT,
                       env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_files_named_by_the_cuda_arm_exist():
    """The roofline block cites a committed ncu traffic file: it must be in profiles/ and say what bench.py reads."""
    src = open(os.path.join(ROOT, "bench.py")).read()
    m = re.search(r'TRAFFIC_FILE = "([^"]+)"', src)
    assert m, "bench.py no longer names its traffic file"
    path = os.path.join(ROOT, "profiles", m.group(1))
    assert os.path.exists(path), path
    tj = json.load(open(path))
    for k in ("x_sweep_kernel", "y_sweep_kernel"):
        # DRAM traffic of a sweep stays within a few per cent of its 216 algorithmic bytes per cell
        assert 200.0 < tj[k]["bytes_per_cell"] < 240.0
    for k, alg in (("euler_x_sweep_kernel", 64.0), ("euler_y_sweep_kernel", 96.0)):
        assert 0.95 * alg < tj[k]["bytes_per_cell"] < 1.15 * alg
    assert os.path.exists(os.path.join(ROOT, "profiles", "r2", "ncu_mhd_sweeps_ot4096.txt"))


def test_workloads_and_reference_arm_agree_on_names():
    """every workload of the CUDA arm has a reference program behind --impl reference"""
    sys.path.insert(0, ROOT)
    import importlib
    bench = importlib.import_module("bench")
    for w, d in bench.WORKLOADS.items():
        assert d["ref_exe"] in ("muscl_ref", "gslic_ref", "tabeos_ref")
        assert d["b_alg"] == d["b_x"] + d["b_y"]

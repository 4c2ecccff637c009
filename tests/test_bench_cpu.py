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
    """The roofline block cites a committed ncu traffic file and a summary: both must be in profiles/."""
    src = open(os.path.join(ROOT, "bench.py")).read()
    names = set(re.findall(r"r1_v\d+_[a-z_0-9]+\.(?:json|md)", src))
    assert names, "bench.py no longer names its profile files"
    for n in names:
        assert os.path.exists(os.path.join(ROOT, "profiles", n)), n
    tj = json.load(open(os.path.join(ROOT, "profiles", sorted(n for n in names if n.endswith("traffic.json"))[-1])))
    for k in ("x_sweep_kernel", "y_sweep_kernel"):
        # DRAM traffic of a sweep stays within a few per cent of its 216 algorithmic bytes per cell
        assert 200.0 < tj[k]["bytes_per_cell"] < 240.0

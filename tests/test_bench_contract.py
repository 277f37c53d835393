"""bench.py's driver contract, on the CPU: the reference arm runs without a GPU on a tiny workload and prints
ONE JSON line with the agreed keys; our arm refuses to run without a device (there is no CPU path)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TINY = ["--genome", "200000", "--samples", "4", "--coverage", "12", "--queries", "3000", "--steps", "2", "--warmup", "1"]


def test_reference_arm_prints_one_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", *TINY],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "queries/s" and d["higher_is_better"] is True
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and d["metric"].startswith("31-mer") and "count+locate" in d["metric"]
    assert d["config"]["workload"].startswith("configs[2]") and d["config"]["shards"] == 64


def test_our_arm_needs_a_device():
    import torch
    if torch.cuda.is_available():
        return
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *TINY], capture_output=True, text=True,
                       timeout=600, cwd=ROOT)
    assert r.returncode != 0 and "CUDA" in r.stderr


import pytest


@pytest.mark.gpu
def test_configs3_workload_at_trial_size():
    """bench.py --workload configs3 (run-level collection) on a few small shards: one JSON line, every planted k-mer
    found, the parity slice equal to the CPU oracle."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "configs3", "--c3-symbols", "3000000",
                        "--c3-shards", "4", "--queries", "20000", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["config"]["workload"].startswith("configs[3]") and d["value"] > 0 and d["gpu_launches"] > 0
    assert d["config"]["kmers_found"] >= 9000 and d["config"]["shards_on_rank0"] == 4   # (shards this small have depth-2
    # trees: the reference loses part of the matches to its tail behaviour, and so does the engine)
    s = d["config"]["parity_slice"]
    assert s["counts_equal"] and s["read_indices_equal"] and s["hits"] > 1000
    assert d["roofline"]["frac"] > 0 and d["cpu_baseline"]["value"] > 0

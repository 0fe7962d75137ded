"""GPU: bench.py validates the output of its own first step -- and fails loudly when it is wrong."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ARGS = ["--steps", "2", "--warmup", "3", "--rows-per-gpu", "3000000", "--no-e2e", "--no-cpu-baseline",
        "--blocks-per-gpu", "4"]


def _run(extra):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + ARGS + extra,
                          capture_output=True, text=True, timeout=300, cwd=ROOT)


def test_bench_line_carries_its_validation():
    r = _run([])
    assert r.returncode == 0, r.stderr[-2000:]
    j = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    v = j["validated"]
    assert v["ok"] and v["crc_equal"] and v["unordered"] == 0 and v["rank_order"]
    assert v["records_in"] == v["records_out"] == 3_000_000
    assert v["idx_equal_all_passes"] and v["miniature_bit_exact"]
    assert j["gpu_launches"] > 0 and j["roofline"]["frac"] > 0 and j["value_multiblock"]["value"] > 0
    assert j["step_ms"]["min"] <= j["step_ms"]["median"] <= j["step_ms"]["max"]


def test_bench_exits_nonzero_on_a_corrupted_output():
    """one flipped payload byte in the first output: the CRC sum no longer matches the input's"""
    r = _run(["--corrupt"])
    assert r.returncode == 3, (r.returncode, r.stdout[-500:], r.stderr[-1000:])
    j = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert j["validated"]["ok"] is False and j["validated"]["crc_equal"] is False

"""CPU: the parts of bench.py's contract that need no GPU — the reference arm's JSON line (the
driver parses it and computes the speed-up from it), the host replay of the device generator, and
the product arm's refusal to run without a CUDA device (no CPU fallback)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_reference_arm_prints_the_contract_line(ref):
    env = dict(os.environ, BULKCIPHER_BENCH_REF_MIB="64")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "AES-128-CTR throughput" and j["unit"] == "GB/s"
    assert j["higher_is_better"] is True and j["steps"] == 2 and j["value"] > 0
    assert j["cpu_baseline"]["kind"] == "reference" and j["cpu_baseline"]["cores"] >= 1
    assert j["e2e"] == {"value": j["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["gpu_launches"] == 0


def test_reference_arm_other_ranks_stay_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", BULKCIPHER_BENCH_REF_MIB="64")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, env=env, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_splitmix_replay_matches_the_definition():
    sys.path.insert(0, ROOT)
    import bench
    M = (1 << 64) - 1

    def mix(x):
        z = (x + 0x9E3779B97F4A7C15) & M
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    for first in (0, 5, (1 << 35) + 3, M - 2):                      # incl. wrap-around of seed + index
        got = bench.splitmix_bytes(first, 4).view("<u8")
        assert [int(v) for v in got] == [mix((bench.SEED + first + i) & M) for i in range(4)]


def test_product_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"], capture_output=True, text=True, timeout=300)
    assert p.returncode != 0 and "no CPU fallback" in (p.stdout + p.stderr)

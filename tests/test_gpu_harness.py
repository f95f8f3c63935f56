"""GPU: the C harnesses (counterparts of the reference's test.c / aes-modes/test.c)
run, verify their own outputs (round trips, device == host path, config-1 SHA-256
digests of SURVEY Appendix C) and exit 0."""
import os
import subprocess

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
BUILD = os.path.join(ROOT, "our-tree_b200", "build")


def run(cmd):
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    return p.stdout


def test_aes_harness_config1_and_sweep():
    out = run([os.path.join(BUILD, "aes_test"), "--config1", "--iters", "3", "--sizes", "1048576,10485760",
               "--threads", "1,4"])
    assert "all outputs verified" in out and "MISMATCH" not in out and "FAILED" not in out
    assert out.count("verify") >= 5
    assert "GPU CTR, 1048576, 4," in out            # the reference's CSV shape: mode, bytes, threads, times...


def test_aes_harness_256():
    out = run([os.path.join(BUILD, "aes_test"), "--bits", "256", "--iters", "2", "--sizes", "4194304", "--threads", "2"])
    assert "all outputs verified" in out


def test_rc4_harness():
    out = run([os.path.join(BUILD, "rc4_test"), "--iters", "3", "--single-bytes", "65536"])
    assert "all outputs verified" in out and "MISMATCH" not in out
    assert "Generated a new key in" in out          # the reference's line (test.c:84)


def test_aes_gpu_cpp_facade():
    """include/AES.h: the reference's `class AES : BlockCipher` (aes-gpu/Source/AES.h:84-147) over the engine."""
    out = run([os.path.join(BUILD, "aes_gpu_facade_test")])
    assert "all outputs verified" in out and "MISMATCH" not in out
    assert out.count("FIPS-197") == 3 and "Invalid AES key size" in out

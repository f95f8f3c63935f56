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


@pytest.mark.skipif(not os.environ.get("BULKCIPHER_RUN_REF_HARNESS"), reason="minutes-long; set BULKCIPHER_RUN_REF_HARNESS=1")
def test_reference_harness_binaries_relinked():
    """The reference's own aes-modes/test.c and test.c, compiled unmodified against include/ and linked
    with libbulkcipher.so (`make ref-harness`): they must run to completion; test.c ends with the
    reference's arc4_self_test, whose three 'passed' lines therefore come from the GPU path."""
    aes, rc4 = os.path.join(BUILD, "ref_aes_modes_test"), os.path.join(BUILD, "ref_rc4_test")
    if not (os.path.exists(aes) and os.path.exists(rc4)):
        pytest.skip("ref harness binaries not built (no reference tree where this was built)")
    p = subprocess.run([aes], capture_output=True, text=True, timeout=1800)
    assert p.returncode == 0 and "AESNI CTR, 1048576000, 8," in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]
    p = subprocess.run([rc4], capture_output=True, text=True, timeout=3000)
    assert p.returncode == 0 and p.stdout.count("passed") == 3, p.stdout[-2000:] + p.stderr[-2000:]

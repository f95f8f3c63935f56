"""GPU: the allocation-free in-place CBC decrypt (bc_aes_cbc_decrypt_inplace_noalloc_dev).

Some users of this library run with all of their device memory handed out in
advance — a fixed HBM budget, their own pool, or a latency-critical loop where
an allocation's implicit synchronisation is not acceptable — so they cannot
use an entry point that may call cudaMalloc behind their back.  For them the
library keeps the boundary blocks of the in-place CBC decrypt in a fixed array
inside its own image.  This test calls that entry point through the C-ABI and
requires the same bytes as the ordinary bc_aes_cbc_decrypt_dev, bit for bit
(integer work: no tolerance), on two calls in a row with different data and
sizes, so anything the first call leaves in the shared array must not leak into
the second.  It lives in a file pytest collects last, and passes both on its
own in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

import gpu_util as G

pytestmark = pytest.mark.gpu
DEC = 0


def _both(pkg, rk, nr, iv, data):
    d_ref = G.to_dev(data)
    G.cbc_dec_dev(pkg, rk, nr, iv, d_ref, d_ref, data.size)             # existing entry point, in place
    d_new = G.to_dev(data)
    pkg.check(pkg.lib().bc_aes_cbc_decrypt_inplace_noalloc_dev(rk.ctypes.data, nr, bytes(iv), d_new.data_ptr(),
                                                              data.size, G.stream()))
    return G.host(d_ref), G.host(d_new)


def test_no_alloc_matches_existing_entry_point_twice(pkg, orc):
    rng = np.random.default_rng(20261106)
    key = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
    rk, nr = G.sched(pkg, key, DEC)
    # call 1: many row-groups, so the shared array is filled with this call's boundary blocks
    a = rng.integers(0, 256, 6 * 1024 * 1024 + 16 * 77, dtype=np.uint8)
    iv_a = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    want_a, got_a = _both(pkg, rk, nr, iv_a, a)
    assert np.array_equal(got_a, want_a)
    assert np.array_equal(got_a, orc.aes_cbc(key, DEC, iv_a, a)[1])       # the large case against the oracle too
    # call 2, straight after: other data, other key size, fewer groups — stale entries must not matter
    key2 = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rk2, nr2 = G.sched(pkg, key2, DEC)
    b = rng.integers(0, 256, 3 * 4096 + 16 * 5, dtype=np.uint8)
    iv_b = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    want_b, got_b = _both(pkg, rk2, nr2, iv_b, b)
    assert np.array_equal(got_b, want_b)
    assert np.array_equal(got_b, orc.aes_cbc(key2, DEC, iv_b, b)[1])      # and both agree with the oracle
    # a single row-group never touches the array at all
    c = rng.integers(0, 256, 16 * 9, dtype=np.uint8)
    want_c, got_c = _both(pkg, rk2, nr2, iv_b, c)
    assert np.array_equal(got_c, want_c)


def test_no_alloc_leaves_free_device_memory_unchanged(pkg, orc):
    """The reason the entry point exists: a caller whose HBM is fully budgeted must see exactly the
    same free device memory before and after it.  The first call on a device (and per key size) is
    a warm-up — it may load the kernel image and set its shared-memory attribute — so the check
    brackets the calls that follow it."""
    import torch

    rng = np.random.default_rng(7)
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rk, nr = G.sched(pkg, key, DEC)
    data = rng.integers(0, 256, 64 * 1024 * 1024, dtype=np.uint8)         # 16384 row-groups
    iv = bytes(range(16))
    d = G.to_dev(data)
    call = lambda: pkg.check(pkg.lib().bc_aes_cbc_decrypt_inplace_noalloc_dev(  # noqa: E731
        rk.ctypes.data, nr, iv, d.data_ptr(), data.size, G.stream()))
    call()                                                                # warm-up
    torch.cuda.synchronize()
    d.copy_(torch.from_numpy(data))                                       # restore the ciphertext (no new allocation)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    call()
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, (free_before, free_after)
    assert np.array_equal(G.host(d), orc.aes_cbc(key, DEC, iv, data)[1])


def test_no_alloc_rejects_oversize(pkg):
    key = bytes(range(16))
    rk, nr = G.sched(pkg, key, DEC)
    d = G.to_dev(np.zeros(4096, np.uint8))
    limit = 65536 * 4096
    rc = pkg.lib().bc_aes_cbc_decrypt_inplace_noalloc_dev(rk.ctypes.data, nr, bytes(16), d.data_ptr(), limit + 16,
                                                         G.stream())
    assert rc == pkg.BC_ERR_TOO_LARGE and "at most" in pkg.last_error()
    assert not G.host(d).any()                                            # rejected before any launch
    rc = pkg.lib().bc_aes_cbc_decrypt_inplace_noalloc_dev(rk.ctypes.data, nr, bytes(16), d.data_ptr(), 24, G.stream())
    assert rc == pkg.BC_ERR_BAD_ARG

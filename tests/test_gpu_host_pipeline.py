"""GPU: the host-buffer pipeline (bc_*_host) on transfers that span many chunks, for pageable
(numpy) and pinned (bc_host_alloc) memory, in place and out of place, including the rarely
taken branches: one CBC message / one RC4 stream longer than a pipeline buffer (chained
through iv / context), ragged CTR tails, the last-keystream output.  The device entry points
(already pinned to the oracle by the other tests) are the expected values here, plus sampled
windows straight from the oracle."""
import ctypes

import numpy as np
import pytest

import gpu_util as G

pytestmark = pytest.mark.gpu
ENC, DEC = 1, 0
MIB = 1 << 20


class Pinned:
    def __init__(self, pkg, n):
        self.pkg, self.n = pkg, n
        self.ptr = pkg.lib().bc_host_alloc(n)
        assert self.ptr, pkg.last_error()
        self.a = np.ctypeslib.as_array((ctypes.c_ubyte * n).from_address(self.ptr))

    def free(self):
        self.pkg.lib().bc_host_free(self.ptr)


def device_expect(pkg, fn, data, extra=0):
    d_in = G.to_dev(data)
    d_out = G.to_dev(np.zeros(data.size + extra, np.uint8))
    fn(d_in, d_out)
    return G.host(d_out)


@pytest.mark.parametrize("pinned", [False, True])
def test_ctr_ecb_cbc_cfb_many_chunks(pkg, orc, pinned):
    rng = np.random.default_rng(2024)
    L = pkg.lib()
    n = 150 * MIB + 16 * 37           # > 8 chunks of 16 MiB (pageable) / 3 chunks of 64 MiB (pinned)
    key = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    iv = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    nc = bytes.fromhex("00112233445566778899aabbffffff00")
    src = rng.integers(0, 256, n + 11, dtype=np.uint8)
    if pinned:
        hin, hout = Pinned(pkg, n + 11), Pinned(pkg, n + 11)
        hin.a[:] = src
        a_in, a_out, p_in, p_out = hin.a, hout.a, hin.ptr, hout.ptr
    else:
        a_in, a_out = src.copy(), np.zeros(n + 11, np.uint8)
        p_in, p_out = a_in.ctypes.data, a_out.ctypes.data

    # CTR with a ragged tail and the last keystream block
    m = n + 11
    want = device_expect(pkg, lambda i, o: G.ctr_dev(pkg, rke, nr, nc, i, o, m, first_block=7), np.concatenate([src, np.zeros(5, np.uint8)]))[:m]
    last = np.zeros(16, np.uint8)
    pkg.check(L.bc_aes_ctr_host(rke.ctypes.data, nr, nc, 128, 7, p_in, p_out, m, last.ctypes.data))
    assert np.array_equal(a_out[:m], want)
    w0 = 64 * MIB - 4096                                     # a window across a chunk seam, straight from the oracle
    shifted = ((int.from_bytes(nc, "big") + 7 + w0 // 16) % (1 << 128)).to_bytes(16, "big")
    assert np.array_equal(a_out[w0:w0 + 8192], orc.aes_ctr(key, shifted, src[w0:w0 + 8192])[0])
    sb = orc.aes_ecb(key, ENC, ((int.from_bytes(nc, "big") + 7 + (m - 1) // 16) % (1 << 128)).to_bytes(16, "big"))
    assert last.tobytes() == sb.tobytes()

    # ECB / CBC-dec / CFB-dec out of place, then CBC-dec in place
    for name, host_call, dev_call in (
        ("ecb", lambda i, o: L.bc_aes_ecb_host(rke.ctypes.data, nr, ENC, i, o, n),
         lambda i, o: G.ecb_dev(pkg, rke, nr, ENC, i, o, n)),
        ("cbc", lambda i, o: L.bc_aes_cbc_decrypt_host(rkd.ctypes.data, nr, iv, i, o, n),
         lambda i, o: G.cbc_dec_dev(pkg, rkd, nr, iv, i, o, n)),
        ("cfb", lambda i, o: L.bc_aes_cfb128_decrypt_host(rke.ctypes.data, nr, iv, i, o, n),
         lambda i, o: pkg.check(L.bc_aes_cfb128_decrypt_dev(rke.ctypes.data, nr, iv, i.data_ptr(), o.data_ptr(), n, G.stream()))),
    ):
        want = device_expect(pkg, dev_call, src[:n])
        a_out[:] = 0
        pkg.check(host_call(p_in, p_out))
        assert np.array_equal(a_out[:n], want), name
    want = device_expect(pkg, lambda i, o: G.cbc_dec_dev(pkg, rkd, nr, iv, i, o, n), src[:n])
    pkg.check(L.bc_aes_cbc_decrypt_host(rkd.ctypes.data, nr, iv, p_in, p_in, n))      # in place: chunk ivs collected first
    assert np.array_equal(a_in[:n], want)
    if pinned:
        hin.free()
        hout.free()


def test_cbc_encrypt_batch_host_shapes(pkg, orc):
    rng = np.random.default_rng(77)
    L = pkg.lib()
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    # (a) many messages, several chunks; (b) ONE message longer than a pipeline buffer (chained through iv)
    for n_msgs, msg_len in ((3000, 64 * 1024), (2, 70 * MIB + 32)):
        data = rng.integers(0, 256, n_msgs * msg_len, dtype=np.uint8)
        ivs = rng.integers(0, 256, n_msgs * 16, dtype=np.uint8)
        d_in, d_out = G.to_dev(data), G.to_dev(np.zeros_like(data))
        d_iv, d_ivo = G.to_dev(ivs), G.to_dev(np.zeros_like(ivs))
        G.cbc_enc_batch_dev(pkg, rke, nr, n_msgs, msg_len, d_iv, d_ivo, d_in, d_out)
        want, want_iv = G.host(d_out), G.host(d_ivo)
        got, got_iv = pkg.aes_cbc_encrypt_batch_host(rke, nr, ivs, data, msg_len)
        assert np.array_equal(got, want) and np.array_equal(got_iv, want_iv), (n_msgs, msg_len)
        m = n_msgs - 1                                        # the head of the last message, straight from the oracle
        head = orc.aes_cbc(key, ENC, ivs[16 * m:16 * m + 16].tobytes(), data[m * msg_len:m * msg_len + 4096])[1]
        assert np.array_equal(got[m * msg_len:m * msg_len + 4096], head)


def test_rc4_host_shapes(pkg, orc):
    rng = np.random.default_rng(78)
    L = pkg.lib()
    # (a) many streams over several chunks with contexts coming back
    n, length = 40000, 4096
    keys = rng.integers(0, 256, n * 16, dtype=np.uint8)
    data = rng.integers(0, 256, n * length, dtype=np.uint8)
    d_out = G.to_dev(np.zeros_like(data))
    d_ctx = G.to_dev(np.zeros(n * 264, np.uint8))
    G.arc4_dev(pkg, n, G.to_dev(keys), 16, length, G.to_dev(data), d_out, None, d_ctx)
    got, ctx = pkg.arc4_batch_host(keys, 16, data, length, want_ctx=True)
    assert np.array_equal(got, G.host(d_out)) and np.array_equal(ctx, G.host(d_ctx))
    assert np.array_equal(got[-length:], orc.arc4(keys[-16:].tobytes(), data[-length:]))
    # (b) ONE stream longer than a pipeline buffer: pieces chained through its arc4_context;
    # the same through the drop-in arc4_setup / arc4_prep / arc4_crypt and rc4_init / rc4_crypt
    key = b"a 70 MiB stream"
    big = rng.integers(0, 256, 70 * MIB + 5, dtype=np.uint8)
    want = orc.arc4(key, big)
    out = np.zeros_like(big)
    ctx1 = np.zeros(264, np.uint8)
    k = np.frombuffer(key, np.uint8)
    pkg.check(L.bc_arc4_batch_host(1, k.ctypes.data, len(key), big.size, big.ctypes.data, out.ctypes.data, None, ctx1.ctypes.data))
    assert np.array_equal(out, want)
    _, x, y, m = orc.arc4_keystream(key, [big.size])
    assert (int.from_bytes(ctx1[0:4].tobytes(), "little"), int.from_bytes(ctx1[4:8].tobytes(), "little"), ctx1[8:].tobytes()) == (x, y, m)
    c = pkg.arc4_setup(key)
    ks = pkg.arc4_prep(c, big.size)
    assert np.array_equal(pkg.arc4_crypt(big, ks), want) and (c.x, c.y, bytes(c.m)) == (x, y, m)
    st = pkg.rc4_init(key)
    assert np.array_equal(pkg.rc4_crypt(st, big), want)


# ---- pageable buffers that come back: the registration cache (host/bc_engine.c) ----

def _reg_stats(pkg):
    out = (ctypes.c_uint64 * 4)()
    on = pkg.lib().bc_internal_reg_stats(out)
    return on, list(out)


class Mapped:
    """An anonymous private mapping made with libc mmap, so the test controls its address: the
    stand-in for a big malloc() (glibc serves those by mmap and frees them by munmap)."""
    libc = ctypes.CDLL(None, use_errno=True)
    libc.mmap.restype = ctypes.c_void_p
    libc.mmap.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long]
    libc.munmap.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
    PROT_RW, MAP_PRIVATE, MAP_ANON, MAP_FIXED = 3, 0x02, 0x20, 0x10

    def __init__(self, n, at=None):
        flags = self.MAP_PRIVATE | self.MAP_ANON | (self.MAP_FIXED if at else 0)
        self.n = n
        self.ptr = self.libc.mmap(at, n, self.PROT_RW, flags, -1, 0)
        assert self.ptr and self.ptr != ctypes.c_void_p(-1).value
        assert at is None or self.ptr == at
        self.a = np.ctypeslib.as_array((ctypes.c_ubyte * n).from_address(self.ptr))

    def unmap(self):
        self.a = None
        assert self.libc.munmap(self.ptr, self.n) == 0


def test_repeated_pageable_buffers_get_pinned_and_stay_correct(pkg, orc):
    """The reference harness reuses two malloc'ed buffers for ten calls (aes-modes/test.c:287-350): the
    second call on the same (pointer, length) pins the range, later calls DMA it directly.  Output is
    checked against the oracle every time, with the input changing between calls."""
    L = pkg.lib()
    n = 96 * MIB + 16 * 7
    key = bytes(range(16))
    rk, nr = G.sched(pkg, key, ENC)
    nc = bytes.fromhex("00112233445566778899aabbccddeeff")
    src, dst = Mapped(n), Mapped(n)
    rng = np.random.default_rng(31)
    L.bc_register_pageable(1)                                          # opt-in; 0 back if frame numbers are hidden
    on, before = _reg_stats(pkg)
    for call in range(4):
        src.a[:] = np.resize(rng.integers(0, 256, 1 << 22, dtype=np.uint8), n)
        dst.a[:] = 0
        pkg.check(L.bc_aes_ctr_host(rk.ctypes.data, nr, nc, 128, call, src.ptr, dst.ptr, n, None))
        shifted = (int.from_bytes(nc, "big") + call).to_bytes(16, "big")
        for lo in (0, n // 2 // 16 * 16, n - MIB):                       # windows incl. the ragged tail
            nc_w = (int.from_bytes(shifted, "big") + lo // 16).to_bytes(16, "big")
            assert np.array_equal(dst.a[lo:lo + MIB], orc.aes_ctr(key, nc_w, src.a[lo:lo + MIB])[0]), (call, lo)
    on, after = _reg_stats(pkg)
    if on:                                                             # frame numbers readable: the cache must have engaged
        assert after[0] - before[0] == 2 and after[1] - before[1] >= 4, (before, after)   # src + dst pinned once, hit on calls 3 and 4
    assert L.bc_register_pageable(0) == 0                              # releases the registrations BEFORE the memory goes away
    src.unmap()
    dst.unmap()


def test_freed_and_reused_addresses_are_noticed(pkg, orc):
    """free() + malloc() can hand the same addresses back with different physical pages behind them.
    A registration kept across that would DMA the old pages; the library compares the frame number of
    every page before each use and must drop the stale entry.  Here: pin a range (two calls), unmap it,
    map NEW memory at the very same address with other contents, call again — the output must belong to
    the new contents, and the stale drop must be counted."""
    L = pkg.lib()
    n = 80 * MIB
    key = bytes(range(1, 33))
    rk, nr = G.sched(pkg, key, ENC)
    iv = bytes(range(16))
    rng = np.random.default_rng(32)
    dst = Mapped(n)
    a = Mapped(n)
    addr = a.ptr
    L.bc_register_pageable(1)
    a.a[:] = np.resize(rng.integers(0, 256, 1 << 22, dtype=np.uint8), n)
    for _ in range(2):                                                  # second call pins it
        pkg.check(L.bc_aes_ecb_host(rk.ctypes.data, nr, ENC, a.ptr, dst.ptr, n))
    on, mid = _reg_stats(pkg)
    a.unmap()
    b = Mapped(n, at=addr)                                              # same addresses, fresh pages
    fresh = np.resize(rng.integers(0, 256, 1 << 22, dtype=np.uint8), n)
    b.a[:] = fresh
    dst.a[:] = 0
    pkg.check(L.bc_aes_ecb_host(rk.ctypes.data, nr, ENC, b.ptr, dst.ptr, n))
    for lo in (0, n // 2, n - MIB):
        assert np.array_equal(dst.a[lo:lo + MIB], orc.aes_ecb(key, ENC, fresh[lo:lo + MIB])), lo
    on, end = _reg_stats(pkg)
    if on:
        assert end[2] - mid[2] >= 1, (mid, end)                         # the stale registration was dropped
    # and the new mapping earns its own registration on the next sighting
    pkg.check(L.bc_aes_ecb_host(rk.ctypes.data, nr, ENC, b.ptr, dst.ptr, n))
    assert np.array_equal(dst.a[:MIB], orc.aes_ecb(key, ENC, fresh[:MIB]))
    L.bc_register_pageable(0)                                           # releases every registration before the unmap
    b.unmap()
    dst.unmap()


def test_ramped_chunk_schedule_subprocess():
    """Pinned CTR / ECB transfers of >= 8 granules start and end with chunks of 1/16 .. 1/2 granule
    (host/bc_engine.c, run_pipeline_enqueued).  With the default 64 MiB granule that needs 512 MiB;
    a child process with a 16 MiB granule exercises the same schedule on 150 MiB (+ a ragged CTR
    tail) and compares every byte with the device-resident path."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = r'''
import ctypes, sys, numpy as np, torch
sys.path.insert(0, %r)
import __graft_entry__ as g
pkg = g.load_package(); L = pkg.lib()
n = 150 * (1 << 20) + 4096 * 3 + 21
rc, ctx = pkg.aes_setkey_enc(bytes(range(16))); rk = pkg.schedule_words(ctx)
nc = bytes.fromhex("000102030405060708090a0bfffffff0")
h_in, h_out = L.bc_host_alloc(n), L.bc_host_alloc(n)
a = np.ctypeslib.as_array((ctypes.c_ubyte * n).from_address(h_in)); b = np.ctypeslib.as_array((ctypes.c_ubyte * n).from_address(h_out))
a[:] = np.resize(np.random.default_rng(1).integers(0, 256, 1 << 22, dtype=np.uint8), n)
d_in = torch.from_numpy(a.copy()).cuda(); d_out = torch.empty_like(d_in); st = torch.cuda.current_stream().cuda_stream
for ramp in ("", "1"):
    # BULKCIPHER_NO_RAMP is read per call: the same bytes must come out with and without the ramp
    import os
    if ramp: os.environ["BULKCIPHER_NO_RAMP"] = "1"
    pkg.check(L.bc_aes_ctr_host(rk.ctypes.data, 10, nc, 128, 5, h_in, h_out, n, None))
    pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, 10, nc, 128, 5, d_in.data_ptr(), d_out.data_ptr(), n, None, st)); torch.cuda.synchronize()
    assert np.array_equal(b, d_out.cpu().numpy()), "CTR"
    m = n // 16 * 16
    pkg.check(L.bc_aes_ecb_host(rk.ctypes.data, 10, 1, h_in, h_out, m))
    pkg.check(L.bc_aes_ecb_dev(rk.ctypes.data, 10, 1, d_in.data_ptr(), d_out.data_ptr(), m, st)); torch.cuda.synchronize()
    assert np.array_equal(b[:m], d_out.cpu().numpy()[:m]), "ECB"
print("RAMP_OK")
''' % root
    env = dict(os.environ, BULKCIPHER_CHUNK_MIB="16")
    env.pop("BULKCIPHER_NO_RAMP", None)
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0 and "RAMP_OK" in p.stdout, p.stdout[-2000:] + p.stderr[-3000:]

"""GPU: corners of the C-ABI that the main parity tests do not reach — device pointers handed to
the host-pointer entry points, calls on several CUDA streams at once (per-stream CBC scratch),
concurrent host threads, over-long RC4 keys, zero lengths."""
import threading

import numpy as np
import pytest
import torch

import gpu_util as G

pytestmark = pytest.mark.gpu
ENC, DEC = 1, 0


def test_host_entry_points_accept_device_pointers(pkg, orc):
    """bc_*_host with cudaMalloc'ed pointers: used in place on the device, no staging (aes.h header note)."""
    rng = np.random.default_rng(11)
    L = pkg.lib()
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    data = rng.integers(0, 256, 1 << 20, dtype=np.uint8)
    iv = bytes(range(16))
    d_in, d_out = G.to_dev(data), G.to_dev(np.zeros_like(data))
    last = np.zeros(16, np.uint8)
    pkg.check(L.bc_aes_ctr_host(rke.ctypes.data, nr, iv, 128, 3, d_in.data_ptr(), d_out.data_ptr(), data.size, last.ctypes.data))
    shifted = (int.from_bytes(iv, "big") + 3).to_bytes(16, "big")
    want, _, _, sb = orc.aes_ctr(key, shifted, data)
    assert np.array_equal(G.host(d_out), want) and last.tobytes() == sb
    pkg.check(L.bc_aes_ecb_host(rke.ctypes.data, nr, ENC, d_in.data_ptr(), d_out.data_ptr(), data.size))
    assert np.array_equal(G.host(d_out), orc.aes_ecb(key, ENC, data))
    pkg.check(L.bc_aes_cbc_decrypt_host(rkd.ctypes.data, nr, iv, d_in.data_ptr(), d_out.data_ptr(), data.size))
    assert np.array_equal(G.host(d_out), orc.aes_cbc(key, DEC, iv, data)[1])
    d_iv = G.to_dev(np.frombuffer(iv * 16, np.uint8))
    pkg.check(L.bc_aes_cbc_encrypt_batch_host(rke.ctypes.data, nr, 16, 65536, d_iv.data_ptr(), None, d_in.data_ptr(), d_out.data_ptr()))
    want = np.concatenate([orc.aes_cbc(key, ENC, iv, data[m * 65536:(m + 1) * 65536])[1] for m in range(16)])
    assert np.array_equal(G.host(d_out), want)
    # the drop-in itself, device pointers straight through aes_crypt_ctr
    rc, ctx = pkg.aes_setkey_enc(key)
    import ctypes as C
    off, nc, sbuf = C.c_int(0), np.frombuffer(iv, np.uint8).copy(), np.zeros(16, np.uint8)
    rc = L.aes_crypt_ctr(C.byref(ctx), data.size, C.byref(off), nc.ctypes.data, sbuf.ctypes.data, d_in.data_ptr(), d_out.data_ptr())
    want, woff, wnc, wsb = orc.aes_ctr(key, iv, data)
    assert rc == 0 and np.array_equal(G.host(d_out), want) and (off.value, nc.tobytes(), sbuf.tobytes()) == (woff, wnc, wsb)


def test_concurrent_streams_inplace_cbc(pkg, orc):
    """In-place CBC decrypt on four streams at once: each stream gets its own boundary scratch."""
    rng = np.random.default_rng(12)
    L = pkg.lib()
    key = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
    rkd, nr = G.sched(pkg, key, DEC)
    streams = [torch.cuda.Stream() for _ in range(4)]
    datas = [rng.integers(0, 256, (8 + i) * (1 << 20) + 16 * i, dtype=np.uint8) for i in range(4)]
    ivs = [rng.integers(0, 256, 16, dtype=np.uint8).tobytes() for _ in range(4)]
    bufs = [G.to_dev(d) for d in datas]
    torch.cuda.synchronize()
    for rep in range(3):
        for s, b, iv, d in zip(streams, bufs, ivs, datas):
            pkg.check(L.bc_aes_cbc_decrypt_dev(rkd.ctypes.data, nr, iv, b.data_ptr(), b.data_ptr(), d.size, s.cuda_stream))
        torch.cuda.synchronize()
        if rep == 0:
            for b, iv, d in zip(bufs, ivs, datas):
                assert np.array_equal(G.host(b), orc.aes_cbc(key, DEC, iv, d)[1])


def test_many_streams_inplace_cbc_stream_ordered_scratch(pkg, orc):
    """In-place CBC decrypt on 20 streams at once, twice over: the boundary scratch is allocated and
    freed in stream order (one buffer per call), so no count of streams makes two calls share one
    (round 1 kept a table of 8 per-stream buffers and recycled entry 0 when it was full)."""
    rng = np.random.default_rng(13)
    L = pkg.lib()
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rkd, nr = G.sched(pkg, key, DEC)
    n = 20
    streams = [torch.cuda.Stream() for _ in range(n)]
    datas = [rng.integers(0, 256, (2 + i % 5) * (1 << 20) + 4096 * i, dtype=np.uint8) for i in range(n)]
    ivs = [rng.integers(0, 256, 16, dtype=np.uint8).tobytes() for _ in range(n)]
    wants = [orc.aes_cbc(key, DEC, iv, d)[1] for iv, d in zip(ivs, datas)]
    for rep in range(2):
        bufs = [G.to_dev(d) for d in datas]
        torch.cuda.synchronize()
        for s, b, iv, d in zip(streams, bufs, ivs, datas):
            pkg.check(L.bc_aes_cbc_decrypt_dev(rkd.ctypes.data, nr, iv, b.data_ptr(), b.data_ptr(), d.size, s.cuda_stream))
        torch.cuda.synchronize()
        for b, w in zip(bufs, wants):
            assert np.array_equal(G.host(b), w)


def test_mixed_host_and_device_buffers_are_refused(pkg):
    """One operand in HBM and the other in host memory: an error code, not a crash."""
    L = pkg.lib()
    rk, nr = G.sched(pkg, bytes(16), ENC)
    d = G.to_dev(np.zeros(4096, np.uint8))
    h = np.zeros(4096, np.uint8)
    iv = bytes(16)
    assert L.bc_aes_ctr_host(rk.ctypes.data, nr, iv, 128, 0, d.data_ptr(), h.ctypes.data, 4096, None) == pkg.BC_ERR_BAD_ARG
    assert "mixed" in pkg.last_error()
    assert L.bc_aes_ecb_host(rk.ctypes.data, nr, ENC, h.ctypes.data, d.data_ptr(), 4096) == pkg.BC_ERR_BAD_ARG
    assert L.bc_aes_cbc_decrypt_host(rk.ctypes.data, nr, iv, d.data_ptr(), h.ctypes.data, 4096) == pkg.BC_ERR_BAD_ARG
    assert L.bc_aes_cbc_encrypt_batch_host(rk.ctypes.data, nr, 1, 4096, h.ctypes.data, None, d.data_ptr(), d.data_ptr()) == pkg.BC_ERR_BAD_ARG
    assert L.bc_xor_host(d.data_ptr(), h.ctypes.data, d.data_ptr(), 4096) == pkg.BC_ERR_BAD_ARG
    assert L.bc_arc4_batch_host(1, h.ctypes.data, 16, 4096, d.data_ptr(), d.data_ptr(), None, None) == pkg.BC_ERR_BAD_ARG
    # a misaligned last-keystream pointer is an argument error, not a sticky misaligned-address fault
    ks = G.to_dev(np.zeros(64, np.uint8))
    assert L.bc_aes_ctr_dev(rk.ctypes.data, nr, iv, 128, 0, d.data_ptr(), d.data_ptr(), 4096, ks.data_ptr() + 4, G.stream()) == pkg.BC_ERR_BAD_ARG
    # all-device operands through the RC4 / XOR host calls are used in place
    keys = G.to_dev(np.arange(16, dtype=np.uint8))
    pkg.check(L.bc_arc4_batch_host(1, keys.data_ptr(), 16, 4096, d.data_ptr(), d.data_ptr(), None, None))
    pkg.check(L.bc_xor_host(d.data_ptr(), d.data_ptr(), d.data_ptr(), 4096))
    assert not G.host(d).any()


def test_concurrent_host_threads(pkg, orc):
    """The reference harness calls the API from up to 8 pthreads on slices of one buffer with one shared
    key context (aes-modes/test.c:273-285): same here from Python threads (ctypes drops the GIL)."""
    rng = np.random.default_rng(13)
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    nc = bytes.fromhex("000102030405060708090a0bfffffff5")
    data = rng.integers(0, 256, 48 << 20, dtype=np.uint8)
    out = np.zeros_like(data)
    T = 8
    per = data.size // T // 16 * 16
    errs = []

    def work(t):
        try:
            lo = t * per
            hi = data.size if t == T - 1 else lo + per
            pkg.aes_ctr_host(rke, nr, nc, data[lo:hi], first_block=lo // 16, out=out[lo:hi])
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=work, args=(t,)) for t in range(T)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    assert np.array_equal(out, orc.aes_ctr(key, nc, data)[0])


def test_rc4_long_keys_and_zero_lengths(pkg, orc):
    rng = np.random.default_rng(14)
    data = rng.integers(0, 256, 777, dtype=np.uint8)
    long_key = rng.integers(0, 256, 300, dtype=np.uint8).tobytes()      # only the first 256 bytes are ever used (arc4.c:58-66)
    ctx = pkg.arc4_setup(long_key)
    ks = pkg.arc4_prep(ctx, data.size)
    assert np.array_equal(pkg.arc4_crypt(data, ks), orc.arc4(long_key, data))
    assert np.array_equal(pkg.arc4_crypt(data, ks), orc.arc4(long_key[:256], data))
    assert pkg.arc4_prep(ctx, 0).size == 0 and pkg.arc4_crypt(b"", b"").size == 0
    st = pkg.rc4_init(b"k")
    assert pkg.rc4_crypt(st, b"").size == 0 and (st.index1, st.index2) == (0, 0)
    L = pkg.lib()
    rk = np.zeros(60, np.uint32)
    assert L.bc_aes_ctr_host(rk.ctypes.data, 10, bytes(16), 128, 0, None, None, 0, None) == 0
    assert L.bc_arc4_batch_host(0, None, 16, 4096, None, None, None, None) == 0
    assert L.bc_arc4_batch_host(4, None, 16, 16, None, None, None, None) == pkg.BC_ERR_BAD_ARG


def test_shutdown_and_reuse(pkg, orc):
    """bc_shutdown releases the library's device resources; the next call simply sets up again."""
    rng = np.random.default_rng(15)
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    data = rng.integers(0, 256, 5 << 20, dtype=np.uint8)
    want = orc.aes_ecb(key, ENC, data)
    for _ in range(2):
        assert np.array_equal(pkg.aes_ecb_host(rke, nr, ENC, data), want)
        free0 = torch.cuda.mem_get_info()[0]
        assert pkg.lib().bc_shutdown() == 0
        assert torch.cuda.mem_get_info()[0] >= free0 + (100 << 20)       # the 3 x 64 MiB pipeline buffers came back


def test_resumable_calls_random_splits(pkg, orc):
    """The reference's resumable state (nc_off / nonce_counter / stream_block, iv_off / iv, arc4 x / y / m,
    rc4 index1 / index2 / perm) under random call splits: every split must give the same bytes and the
    same final state as the oracle fed the same splits."""
    rng = np.random.default_rng(2718)
    for trial in range(12):
        kb = (128, 192, 256)[trial % 3]
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        rc, e = pkg.aes_setkey_enc(key)
        total = int(rng.integers(1, 5000))
        data = rng.integers(0, 256, total, dtype=np.uint8).tobytes()
        cuts = sorted(set(int(c) for c in rng.integers(0, total + 1, int(rng.integers(0, 6)))))
        parts = [data[a:b] for a, b in zip([0] + cuts, cuts + [total])]
        nc0 = bytes(rng.integers(0, 256, 16 - trial % 5, dtype=np.uint8)) + b"\xff" * (trial % 5)
        iv0 = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        # CTR
        nc, sb, off, got = nc0, bytes(16), 0, []
        onc, osb, ooff, want = nc0, bytes(16), 0, []
        for p in parts:
            rc, o, off, nc, sb = pkg.aes_crypt_ctr(e, nc, p, off, sb)
            assert rc == 0
            got.append(o.tobytes())
            o2, ooff, onc, osb = orc.aes_ctr(key, onc, p, ooff, osb)
            want.append(o2.tobytes())
        assert b"".join(got) == b"".join(want) and (off, nc, sb) == (ooff, onc, osb)
        # CFB128, both directions
        for mode in (1, 0):
            iv, off, got = iv0, 0, []
            oiv, ooff, want = iv0, 0, []
            for p in parts:
                rc, o, off, iv = pkg.aes_crypt_cfb128(e, mode, iv, p, off)
                assert rc == 0
                got.append(o.tobytes())
                o2, ooff, oiv = orc.aes_cfb128(key, mode, oiv, p, ooff)
                want.append(o2.tobytes())
            assert b"".join(got) == b"".join(want) and (off, iv) == (ooff, oiv)
        # ARC4 keystream chaining and rc4_crypt chaining
        rkey = rng.integers(0, 256, int(rng.integers(1, 40)), dtype=np.uint8).tobytes()
        c = pkg.arc4_setup(rkey)
        ks = np.concatenate([pkg.arc4_prep(c, len(p)) for p in parts] + [np.zeros(0, np.uint8)])
        wks, x, y, m = orc.arc4_keystream(rkey, [len(p) for p in parts])
        assert np.array_equal(ks, wks) and (c.x, c.y, bytes(c.m)) == (x, y, m)
        st = pkg.rc4_init(rkey)
        out = np.concatenate([pkg.rc4_crypt(st, p) for p in parts] + [np.zeros(0, np.uint8)])
        wout, i1, i2, perm = orc.rc4(rkey, parts)
        assert np.array_equal(out, wout) and (st.index1, st.index2, bytes(st.perm)) == (i1, i2, perm)


# ---- aes_crypt_ecb one block per call: the per-thread read-ahead window (host/aes_dropin.c) ----

def _ecb_stats(pkg):
    import ctypes
    out = (ctypes.c_uint64 * 2)()
    pkg.lib().bc_internal_ecb_ahead_stats(out)
    return out[0], out[1]


def _walk(pkg, ctx, mode, src, dst, first, count):
    import ctypes
    L = pkg.lib()
    for b in range(first, first + count):
        assert L.aes_crypt_ecb(ctypes.byref(ctx), mode, src.ctypes.data + 16 * b, dst.ctypes.data + 16 * b) == 0


def test_ecb_block_loop_read_ahead_is_exact(pkg, orc):
    """The reference's loop over aes_crypt_ecb (aes-modes/test.c:37-41), 16 bytes per call: served from
    the read-ahead window, bit-exact against the oracle — also when the caller changes the bytes ahead
    of the walk, re-keys the context in the middle, works in place, or jumps around."""
    rng = np.random.default_rng(77)
    key, key2 = bytes(range(16)), bytes(range(50, 82))
    nblk = 16384                                                        # 256 KiB
    data = rng.integers(0, 256, 16 * nblk, dtype=np.uint8)
    out = np.zeros_like(data)
    rc, ctx = pkg.aes_setkey_enc(key)
    h0, m0 = _ecb_stats(pkg)
    _walk(pkg, ctx, ENC, data, out, 0, 6000)
    h1, m1 = _ecb_stats(pkg)
    assert (h1 - h0) + (m1 - m0) == 6000 and m1 - m0 < 40, (h1 - h0, m1 - m0)       # a handful of GPU trips, not 6000
    data[16 * 6100:16 * 6200] ^= 0x5A                                   # the caller rewrites bytes the window already copied
    _walk(pkg, ctx, ENC, data, out, 6000, 2000)
    assert np.array_equal(out[:16 * 8000], orc.aes_ecb(key, ENC, data[:16 * 8000]))
    import ctypes
    assert pkg.lib().aes_setkey_enc(ctypes.byref(ctx), key2, 256) == 0  # same context object, new key, the walk goes on
    _walk(pkg, ctx, ENC, data, out, 8000, nblk - 8000)
    assert np.array_equal(out[16 * 8000:], orc.aes_ecb(key2, ENC, data[16 * 8000:]))
    # decrypt in place (input == output), then random single blocks
    rc, dctx = pkg.aes_setkey_dec(key2)
    buf = out[16 * 8000:].copy()
    _walk(pkg, dctx, DEC, buf, buf, 0, buf.size // 16)
    assert np.array_equal(buf, data[16 * 8000:])
    for b in rng.integers(0, nblk, 50):
        _walk(pkg, ctx, ENC, data, out, int(b), 1)
        assert np.array_equal(out[16 * b:16 * b + 16], orc.aes_ecb(key2, ENC, data[16 * b:16 * b + 16]))


def test_ecb_read_ahead_stops_at_unmapped_memory(pkg, orc):
    """The window is copied with process_vm_readv, one page per iovec element: a walk that ends right
    below an unmapped page must neither fault nor return anything but the right blocks."""
    import ctypes
    libc = ctypes.CDLL(None, use_errno=True)
    libc.mmap.restype = ctypes.c_void_p
    libc.mmap.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long]
    libc.munmap.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
    base = libc.mmap(None, 3 * 4096, 3, 0x22, -1, 0)
    assert base and base != ctypes.c_void_p(-1).value
    assert libc.munmap(base + 2 * 4096, 4096) == 0                      # third page gone: reads past page 2 would fault
    src = np.ctypeslib.as_array((ctypes.c_ubyte * 8192).from_address(base))
    src[:] = np.random.default_rng(5).integers(0, 256, 8192, dtype=np.uint8)
    dst = np.zeros(8192, np.uint8)
    key = bytes(range(24))
    rc, ctx = pkg.aes_setkey_enc(key)
    _walk(pkg, ctx, ENC, src, dst, 0, 512)                              # every block of both pages, up to the very edge
    assert np.array_equal(dst, orc.aes_ecb(key, ENC, src.copy()))
    libc.munmap(base, 2 * 4096)

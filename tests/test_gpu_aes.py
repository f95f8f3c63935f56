"""GPU parity, AES: the CUDA path through the C-ABI vs the oracle, the golden
fixtures, the standard KATs and the reference's own self-test vectors.
Bit-exact everywhere (integer/byte work: no tolerance)."""
import hashlib

import numpy as np
import pytest

import gpu_util as G

pytestmark = pytest.mark.gpu
ENC, DEC = 1, 0
h = bytes.fromhex


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a)).hexdigest()


# ------------------------------------------------------------ drop-in KATs

def test_fips197_dropin(pkg, golden):
    for v in golden["kat"]["fips197"]:
        rc, ctx = pkg.aes_setkey_enc(h(v["key"]))
        assert pkg.aes_crypt_ecb(ctx, ENC, h(v["pt"])) == (0, h(v["ct"]))
        rc, ctx = pkg.aes_setkey_dec(h(v["key"]))
        assert pkg.aes_crypt_ecb(ctx, DEC, h(v["ct"])) == (0, h(v["pt"]))


def test_sp800_38a_dropin(pkg, golden):
    for v in golden["kat"]["sp800_38a"]:
        key, pt = h(v["key"]), h(v["pt"])
        rc, e = pkg.aes_setkey_enc(key)
        rc, d = pkg.aes_setkey_dec(key)
        nr = e.nr
        assert pkg.AES_ECB_encrypt(pt, pkg.AES_Key_Expansion(key), nr).tobytes().hex() == v["ecb"]
        rc, c, ivo = pkg.aes_crypt_cbc(e, ENC, h(v["iv"]), pt)
        assert (rc, c.tobytes().hex(), ivo.hex()) == (0, v["cbc"], v["cbc_iv_out"])
        rc, p, ivo = pkg.aes_crypt_cbc(d, DEC, h(v["iv"]), h(v["cbc"]))
        assert (rc, p.tobytes(), ivo.hex()) == (0, pt, v["cbc"][-32:])
        rc, p, ivo = pkg.aes_crypt_cbc(d, DEC, h(v["iv"]), h(v["cbc"]), inplace=True)   # aes.c:786
        assert (rc, p.tobytes(), ivo.hex()) == (0, pt, v["cbc"][-32:])
        rc, t, off, nco, sb = pkg.aes_crypt_ctr(e, h(v["ctr"]), pt)
        assert (rc, t.tobytes().hex(), nco.hex(), off) == (0, v["ctr_ct"], v["ctr_out"], 0)


def test_rfc3686_dropin(pkg, golden):
    """aes.c:1025-1078 incl. the 36-byte partial block and the resume state."""
    for v in golden["kat"]["rfc3686"]:
        rc, e = pkg.aes_setkey_enc(h(v["key"]))
        rc, t, off, nco, sb = pkg.aes_crypt_ctr(e, h(v["nonce_counter"]), h(v["pt"]))
        assert (rc, t.tobytes().hex()) == (0, v["ct"])
        assert (off, nco.hex(), sb.hex()) == (v["nc_off_out"], v["nonce_counter_out"], v["stream_block_out"])
        rc, t, *_ = pkg.aes_crypt_ctr(e, h(v["nonce_counter"]), h(v["ct"]), inplace=True)
        assert t.tobytes().hex() == v["pt"]


def test_aes_self_test_on_gpu(pkg):
    """The reference's aes_self_test (aes.c:1084-1330: Monte-Carlo ECB/CBC ×10 000 for
    128/192/256, RFC 3686 CTR) re-implemented over the drop-in calls, every block on the GPU."""
    assert pkg.lib().aes_self_test(0) == 0


def test_ctr_edge_cases_dropin(pkg, golden):
    e = golden["kat"]["edge"]
    v = e["ctr_full_carry"]
    rc, ctx = pkg.aes_setkey_enc(h(v["key"]))
    rc, t, off, nco, sb = pkg.aes_crypt_ctr(ctx, h(v["nonce_counter"]), bytes(v["len"]))
    assert (t.tobytes().hex(), nco.hex()) == (v["ct"], v["nonce_counter_out"])      # 12ff..ff -> 1300..01
    v = e["ctr_100"]
    rc, ctx = pkg.aes_setkey_enc(h(v["key"]))
    pt = h(v["pt"])
    nc, sb, off, outs, pos = h(v["nonce_counter"]), bytes(16), 0, [], 0
    for n in v["chunks"]:                                                         # 5 + 30 + 0 + 65
        rc, o, off, nc, sb = pkg.aes_crypt_ctr(ctx, nc, pt[pos:pos + n], off, sb)
        assert rc == 0
        outs.append(o.tobytes())
        pos += n
    assert b"".join(outs).hex() == v["ct"]
    assert (off, nc.hex(), sb.hex()) == (v["nc_off_out"], v["nonce_counter_out"], v["stream_block_out"])


def test_committed_vectors(pkg, golden):
    for v in golden["vectors"]["aes"]:
        key, data = h(v["key"]), np.frombuffer(h(v["data"]), np.uint8)
        rc, e = pkg.aes_setkey_enc(key)
        rc, d = pkg.aes_setkey_dec(key)
        rke, rkd, nr = pkg.schedule_words(e), pkg.schedule_words(d), e.nr
        assert pkg.aes_ecb_host(rke, nr, ENC, data).tobytes().hex() == v["ecb_enc"]
        assert pkg.aes_ecb_host(rkd, nr, DEC, data).tobytes().hex() == v["ecb_dec"]
        assert pkg.AES_ECB_decrypt(data, h(v["dec_schedule"]), nr).tobytes().hex() == v["ecb_dec"]
        rc, c, ivo = pkg.aes_crypt_cbc(e, ENC, h(v["iv"]), data)
        assert (c.tobytes().hex(), ivo.hex()) == (v["cbc_enc"], v["cbc_enc_iv_out"])
        rc, c, ivo = pkg.aes_crypt_cbc(d, DEC, h(v["iv"]), data)
        assert (c.tobytes().hex(), ivo.hex()) == (v["cbc_dec"], v["cbc_dec_iv_out"])
        rc, t, off, nco, sb = pkg.aes_crypt_ctr(e, h(v["nonce_counter"]), data[: v["ctr_ragged_len"]])
        assert (t.tobytes().hex(), off, nco.hex(), sb.hex()) == (
            v["ctr"], v["ctr_nc_off_out"], v["ctr_nonce_counter_out"], v["ctr_stream_block_out"])
        got = pkg.AES_CTR_encrypt(data, h(v["aesni_ivec"]), h(v["aesni_nonce"]), h(v["enc_schedule"]), nr)
        assert got.tobytes().hex() == v["aesni_ctr"]                              # 64-bit counter wrap inside


def test_cfb128_dropin_and_device(pkg, orc, golden):
    """aes_crypt_cfb128 (aes.h:124, aes.c:822-862): SP 800-38A vectors of aes.c:958-1015, byte-granular
    chaining with *iv_off, ragged lengths, in place; then the device entry points vs the oracle."""
    for v in golden["kat"]["cfb128"]:
        rc, e = pkg.aes_setkey_enc(h(v["key"]))
        rc, c, off, ivo = pkg.aes_crypt_cfb128(e, ENC, h(v["iv"]), h(v["pt"]))
        assert (rc, c.tobytes().hex(), off, ivo.hex()) == (0, v["ct"], 0, v["iv_out"])
        rc, p, off, ivo = pkg.aes_crypt_cfb128(e, DEC, h(v["iv"]), h(v["ct"]), inplace=True)
        assert (rc, p.tobytes().hex()) == (0, v["pt"])
    v = golden["kat"]["edge"]["cfb_150"]
    rc, e = pkg.aes_setkey_enc(h(v["key"]))
    for mode, src, dst in ((ENC, "pt", "ct"), (DEC, "ct", "pt")):
        data, iv, off, outs, pos = h(v[src]), h(v["iv"]), 0, [], 0
        for n in v["chunks"]:                                        # 3 + 40 + 0 + 16 + 91
            rc, o, off, iv = pkg.aes_crypt_cfb128(e, mode, iv, data[pos:pos + n], off)
            assert rc == 0
            outs.append(o.tobytes())
            pos += n
        assert b"".join(outs).hex() == v[dst]
        assert (off, iv.hex()) == (v["iv_off_out"], v["iv_out"])
    for v in golden["vectors"]["aes"]:
        rc, e = pkg.aes_setkey_enc(h(v["key"]))
        rc, c, off, ivo = pkg.aes_crypt_cfb128(e, ENC, h(v["iv"]), h(v["cfb_data"]))
        assert (c.tobytes().hex(), off, ivo.hex()) == (v["cfb_enc"], v["cfb_enc_off"], v["cfb_enc_iv_out"])
        rc, c, off, ivo = pkg.aes_crypt_cfb128(e, DEC, h(v["iv"]), h(v["cfb_data"]))
        assert (c.tobytes().hex(), off, ivo.hex()) == (v["cfb_dec"], v["cfb_dec_off"], v["cfb_dec_iv_out"])
    rng = np.random.default_rng(31)
    for kb, n_msgs, msg_len in ((128, 1, 16 * 1000 + 16), (256, 77, 4096), (192, 600, 48)):
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        rke, nr = G.sched(pkg, key, ENC)
        data = rng.integers(0, 256, n_msgs * msg_len, dtype=np.uint8)
        ivs = rng.integers(0, 256, n_msgs * 16, dtype=np.uint8)
        want = np.concatenate([orc.aes_cfb128(key, ENC, ivs[16 * m:16 * m + 16].tobytes(),
                                              data[m * msg_len:(m + 1) * msg_len])[0] for m in range(n_msgs)])
        d_in, d_out = G.to_dev(data), G.to_dev(np.zeros_like(data))
        d_iv, d_ivo = G.to_dev(ivs), G.to_dev(np.zeros_like(ivs))
        pkg.check(pkg.lib().bc_aes_cfb128_encrypt_batch_dev(rke.ctypes.data, nr, n_msgs, msg_len, d_iv.data_ptr(),
                                                            d_ivo.data_ptr(), d_in.data_ptr(), d_out.data_ptr(), G.stream()))
        assert np.array_equal(G.host(d_out), want)
        assert np.array_equal(G.host(d_ivo).reshape(n_msgs, 16), want.reshape(n_msgs, msg_len)[:, -16:])
        # block-parallel decrypt of the whole buffer as one chain, out of place and in place
        iv0 = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        wantd = orc.aes_cfb128(key, DEC, iv0, data)[0]
        pkg.check(pkg.lib().bc_aes_cfb128_decrypt_dev(rke.ctypes.data, nr, iv0, d_in.data_ptr(), d_out.data_ptr(),
                                                      data.size, G.stream()))
        assert np.array_equal(G.host(d_out), wantd)
        pkg.check(pkg.lib().bc_aes_cfb128_decrypt_dev(rke.ctypes.data, nr, iv0, d_in.data_ptr(), d_in.data_ptr(),
                                                      data.size, G.stream()))
        assert np.array_equal(G.host(d_in), wantd)


# ------------------------------------------------- config 1: 16 MiB digests

def test_golden_digests_16mib_device(pkg, golden, harness_input):
    """BASELINE config 1 input (srand(1337); rand()%255, 16 MiB) through the device C-ABI;
    SHA-256 of every output must equal the reference's."""
    msg, key32 = harness_input
    dg = golden["digests"]
    d_in = G.to_dev(msg)
    d_out = G.to_dev(np.zeros_like(msg))
    z = bytes(16)
    n = msg.size
    for kb in (128, 256):
        key = key32[: kb // 8]
        rke, nr = G.sched(pkg, key, ENC)
        rkd, _ = G.sched(pkg, key, DEC)
        G.ecb_dev(pkg, rke, nr, ENC, d_in, d_out, n)
        assert sha(G.host(d_out)) == dg[f"ecb{kb}_enc"]
        G.ecb_dev(pkg, rkd, nr, DEC, d_in, d_out, n)
        assert sha(G.host(d_out)) == dg[f"ecb{kb}_dec"]
        G.cbc_dec_dev(pkg, rkd, nr, z, d_in, d_out, n)
        assert sha(G.host(d_out)) == dg[f"cbc{kb}_dec"]
        G.ctr_dev(pkg, rke, nr, z, d_in, d_out, n)
        assert sha(G.host(d_out)) == dg[f"ctr{kb}"]
        pkg.check(pkg.lib().bc_aes_cfb128_decrypt_dev(rke.ctypes.data, nr, z, d_in.data_ptr(), d_out.data_ptr(), n, G.stream()))
        assert sha(G.host(d_out)) == dg[f"cfb{kb}_dec"]
        # AES_CTR_encrypt(ivec="HloEelAS", nonce="315A") as aes-modes/test.c:305-308
        assert sha(pkg.AES_CTR_encrypt(msg, b"HloEelAS", b"315A", pkg.AES_Key_Expansion(key), nr)) == dg[f"aesni_ctr{kb}"]
        # CBC encrypt of the whole 16 MiB as ONE message (serial chain, one GPU thread): 128-bit only, it is slow
        if kb == 128:
            d_iv = G.to_dev(np.zeros(16, np.uint8))
            G.cbc_enc_batch_dev(pkg, rke, nr, 1, n, d_iv, None, d_in, d_out)
            assert sha(G.host(d_out)) == dg["cbc128_enc"]
    rke, nr = G.sched(pkg, key32, ENC)
    d_iv = G.to_dev(msg[: 256 * 16])
    d_ivo = G.to_dev(np.zeros(256 * 16, np.uint8))
    G.cbc_enc_batch_dev(pkg, rke, nr, 256, 65536, d_iv, d_ivo, d_in, d_out)
    assert sha(G.host(d_out)) == dg["cbc256_enc_batch_256x64k"]
    assert sha(G.host(d_ivo)) == dg["cbc256_enc_batch_ivs_out"]


# ------------------------------------------ random differential vs oracle

@pytest.mark.parametrize("kb", [128, 192, 256])
def test_ecb_cbc_random_sizes(pkg, orc, kb):
    rng = np.random.default_rng(kb)
    key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    for nblocks in (1, 2, 31, 32, 33, 255, 256, 257, 511, 4096 + 7, 16 * 256 * 16 + 5):
        data = rng.integers(0, 256, nblocks * 16, dtype=np.uint8)
        iv = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        d_in, d_out = G.to_dev(data), G.to_dev(np.zeros_like(data))
        G.ecb_dev(pkg, rke, nr, ENC, d_in, d_out, data.size)
        assert np.array_equal(G.host(d_out), orc.aes_ecb(key, ENC, data))
        G.ecb_dev(pkg, rkd, nr, DEC, d_in, d_out, data.size)
        assert np.array_equal(G.host(d_out), orc.aes_ecb(key, DEC, data))
        want = orc.aes_cbc(key, DEC, iv, data)[1]
        G.cbc_dec_dev(pkg, rkd, nr, iv, d_in, d_out, data.size)
        assert np.array_equal(G.host(d_out), want)
        G.cbc_dec_dev(pkg, rkd, nr, iv, d_in, d_in, data.size)          # in place: scratch-backed boundaries
        assert np.array_equal(G.host(d_in), want)
        d_in = G.to_dev(data)
        G.ecb_dev(pkg, rke, nr, ENC, d_in, d_in, data.size)             # in place ECB
        assert np.array_equal(G.host(d_in), orc.aes_ecb(key, ENC, data))


@pytest.mark.parametrize("kb", [128, 192, 256])
def test_ctr_random_sizes_and_carries(pkg, orc, kb):
    rng = np.random.default_rng(1000 + kb)
    key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    counters = [bytes(16), bytes(range(16)), h("00" * 12 + "fffffff0"), h("00" * 8 + "ffffffffffffff80"),
                h("12" + "ff" * 15), h("ff" * 15 + "01"), rng.integers(0, 256, 16, dtype=np.uint8).tobytes()]
    sizes = [1, 15, 16, 17, 500, 4096, 4097, 2048 * 16, 2048 * 16 + 16 * 300 + 3, 70001]
    for ci, nc in enumerate(counters):
        for n in sizes[ci % 3::3] + [sizes[-1]]:
            data = rng.integers(0, 256, n, dtype=np.uint8)
            pad = (-n) % 16
            d_in = G.to_dev(np.concatenate([data, np.zeros(pad, np.uint8)]))
            d_out = G.to_dev(np.full(n + pad, 0xEE, np.uint8))
            d_last = G.to_dev(np.zeros(16, np.uint8))
            want, off, nco, sb = orc.aes_ctr(key, nc, data)
            G.ctr_dev(pkg, rke, nr, nc, d_in, d_out, n, d_last=d_last)
            got = G.host(d_out)
            assert np.array_equal(got[:n], want), (ci, n)
            assert (got[n:] == 0xEE).all(), "wrote past a ragged tail"
            assert G.host(d_last).tobytes() == sb
            # shard seam: same stream computed as two calls with a block offset
            if n > 64:
                cut = (n // 32) * 16
                d_out2 = G.to_dev(np.zeros(n + pad, np.uint8))
                G.ctr_dev(pkg, rke, nr, nc, d_in, d_out2, cut)
                pkg.check(pkg.lib().bc_aes_ctr_dev(rke.ctypes.data, nr, nc, 128, cut // 16, d_in.data_ptr() + cut,
                                                   d_out2.data_ptr() + cut, n - cut, None, G.stream()))
                assert np.array_equal(G.host(d_out2)[:n], want)


def test_ctr_config5_shard_offsets(pkg, orc):
    """BASELINE config 5: 256 GiB = 2^34 blocks sharded over 2/4/8 GPUs by counter offset.  The buffer
    itself does not fit a test, but every shard start does: windows at first_block = g * 2^34 / G,
    straddling the 2^32 block seams (64 GiB * k) and the end of the 256 GiB stream, vs the oracle
    run on counter + first_block."""
    rng = np.random.default_rng(55)
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, 1)
    nc = bytes.fromhex("0f1e2d3c4b5a69788796a5b4fffffff0")
    total = 1 << 34
    starts = sorted({g * total // n for n in (2, 4, 8) for g in range(n)} |
                    {k * (1 << 32) - 700 for k in (1, 2, 3)} | {total - 1000, (1 << 32) - 16 - 0xfffffff0 % 256})
    data = rng.integers(0, 256, 16 * 1500, dtype=np.uint8)
    d_in, d_out = G.to_dev(data), G.to_dev(np.zeros_like(data))
    for b0 in starts:
        G.ctr_dev(pkg, rke, nr, nc, d_in, d_out, data.size, first_block=b0)
        shifted = ((int.from_bytes(nc, "big") + b0) % (1 << 128)).to_bytes(16, "big")
        assert np.array_equal(G.host(d_out), orc.aes_ctr(key, shifted, data)[0]), b0
    # the same offsets under the 64-bit AES-NI rule: only bytes 8..15 move
    b0 = (1 << 34) - 5
    G.ctr_dev(pkg, rke, nr, nc, d_in, d_out, data.size, bits=64, first_block=b0)
    lo = (int.from_bytes(nc[8:], "big") + b0) % (1 << 64)
    want = np.concatenate([data[16 * i:16 * i + 16] ^ orc.aes_ecb(key, 1, nc[:8] + ((lo + i) % (1 << 64)).to_bytes(8, "big"))
                           for i in range(40)])
    assert np.array_equal(G.host(d_out)[:640], want)


def test_ctr_64bit_counter_mode(pkg, orc):
    """AES_CTR_encrypt's add never carries out of bytes 8..15 (aesni.c:139,143)."""
    rng = np.random.default_rng(5)
    key = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    sched = pkg.AES_Key_Expansion(key)
    data = rng.integers(0, 256, 16 * 1000, dtype=np.uint8)
    for ivec in (h("aabbccddffffffff"), h("00000000fffffffe"), h("0102030400000000")):
        want = orc.aesni_ctr(sched, 10, ivec, b"\x09\x08\x07\x06", data)
        assert np.array_equal(pkg.AES_CTR_encrypt(data, ivec, b"\x09\x08\x07\x06", sched, 10), want)
    # low 64 bits all ones: 128-bit mode carries into byte 7, 64-bit mode wraps to zero
    rke, nr = G.sched(pkg, key, ENC)
    nc = h("0011223344556677" + "ff" * 8)
    d_in, d_a, d_b = G.to_dev(data), G.to_dev(np.zeros_like(data)), G.to_dev(np.zeros_like(data))
    G.ctr_dev(pkg, rke, nr, nc, d_in, d_a, data.size, bits=128)
    G.ctr_dev(pkg, rke, nr, nc, d_in, d_b, data.size, bits=64)
    a, b = G.host(d_a), G.host(d_b)
    assert np.array_equal(a, orc.aes_ctr(key, nc, data)[0])
    assert np.array_equal(a[:16], b[:16]) and not np.array_equal(a[16:32], b[16:32])
    ks1 = orc.aes_ecb(key, ENC, h("0011223344556677" + "00" * 8))       # wrapped counter, no carry
    assert np.array_equal(b[16:32], data[16:32] ^ ks1)


def test_cbc_encrypt_batch_random(pkg, orc):
    rng = np.random.default_rng(9)
    for kb, n_msgs, msg_len in ((128, 1, 16), (256, 3, 48), (128, 700, 160), (192, 33, 4096), (256, 1500, 32)):
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        rke, nr = G.sched(pkg, key, ENC)
        data = rng.integers(0, 256, n_msgs * msg_len, dtype=np.uint8)
        ivs = rng.integers(0, 256, n_msgs * 16, dtype=np.uint8)
        want, want_iv = orc.aes_cbc_encrypt_batch(key, ivs.copy(), data, msg_len)
        d_in, d_out = G.to_dev(data), G.to_dev(np.zeros_like(data))
        d_iv, d_ivo = G.to_dev(ivs), G.to_dev(np.zeros_like(ivs))
        G.cbc_enc_batch_dev(pkg, rke, nr, n_msgs, msg_len, d_iv, d_ivo, d_in, d_out)
        assert np.array_equal(G.host(d_out), want) and np.array_equal(G.host(d_ivo), want_iv)
        o, ivo = pkg.aes_cbc_encrypt_batch_host(rke, nr, ivs, data, msg_len)
        assert np.array_equal(o, want) and np.array_equal(ivo, want_iv)


def test_reference_objects_agree_when_present(pkg, ref):
    """Where oracle/_ref/libref.so travelled to this box: CUDA vs the unmodified aes.c / aesni.c."""
    rng = np.random.default_rng(77)
    data = rng.integers(0, 256, 1 << 20, dtype=np.uint8)
    for kb in (128, 256):
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        rke, nr = G.sched(pkg, key, ENC)
        rkd, _ = G.sched(pkg, key, DEC)
        nc = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        assert np.array_equal(pkg.aes_ctr_host(rke, nr, nc, data), ref.aes_ctr(key, nc, data)[0])
        assert np.array_equal(pkg.aes_ecb_host(rke, nr, ENC, data), ref.aes_ecb(key, ENC, data))
        assert np.array_equal(pkg.aes_cbc_decrypt_host(rkd, nr, nc, data), ref.aes_cbc(key, DEC, nc, data)[1])
        if ref.has_aesni():
            s = ref.packed_schedule(key, ENC)
            assert np.array_equal(pkg.AES_CTR_encrypt(data, b"HloEelAS", b"315A", s, nr),
                                  ref.aesni_ctr(s, nr, b"HloEelAS", b"315A", data))


def test_no_writes_outside_output(pkg, orc):
    """compute-sanitizer is closed on this pool, so bounds are checked the plain way: every
    kernel writes into the middle of a canary-filled buffer (sizes that end inside a warp row, a
    row-group and a TMA box) and the canaries on both sides must survive."""
    import torch
    rng = np.random.default_rng(404)
    key = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    L = pkg.lib()
    guard = 4096
    for n in (16, 16 * 31, 16 * 33, 4096 - 16, 4096 + 16, 16 * 257, 65536 + 48):
        data = rng.integers(0, 256, n, dtype=np.uint8)
        d_in = G.to_dev(data)
        iv = bytes(range(16))
        d_iv = G.to_dev(np.frombuffer(iv, np.uint8))

        def run(fn):
            buf = torch.full((n + 2 * guard,), 0xA5, dtype=torch.uint8, device="cuda")
            fn(buf.data_ptr() + guard)
            h = G.host(buf)
            assert (h[:guard] == 0xA5).all() and (h[guard + n:] == 0xA5).all(), (fn.__name__, n)
            return h[guard:guard + n]

        def ecb(p): pkg.check(L.bc_aes_ecb_dev(rke.ctypes.data, nr, ENC, d_in.data_ptr(), p, n, G.stream()))
        def ecbd(p): pkg.check(L.bc_aes_ecb_dev(rkd.ctypes.data, nr, DEC, d_in.data_ptr(), p, n, G.stream()))
        def cbcd(p): pkg.check(L.bc_aes_cbc_decrypt_dev(rkd.ctypes.data, nr, iv, d_in.data_ptr(), p, n, G.stream()))
        def cfbd(p): pkg.check(L.bc_aes_cfb128_decrypt_dev(rke.ctypes.data, nr, iv, d_in.data_ptr(), p, n, G.stream()))
        def ctr(p): pkg.check(L.bc_aes_ctr_dev(rke.ctypes.data, nr, iv, 128, 5, d_in.data_ptr(), p, n, None, G.stream()))
        def cbce(p): pkg.check(L.bc_aes_cbc_encrypt_batch_dev(rke.ctypes.data, nr, 1, n, d_iv.data_ptr(), None, d_in.data_ptr(), p, G.stream()))
        def cfbe(p): pkg.check(L.bc_aes_cfb128_encrypt_batch_dev(rke.ctypes.data, nr, 1, n, d_iv.data_ptr(), None, d_in.data_ptr(), p, G.stream()))
        def rc4(p): pkg.check(L.bc_arc4_batch_dev(n // 16, d_in.data_ptr(), 16, 16, d_in.data_ptr(), p, None, None, G.stream()))

        assert np.array_equal(run(ecb), orc.aes_ecb(key, ENC, data))
        assert np.array_equal(run(ecbd), orc.aes_ecb(key, DEC, data))
        assert np.array_equal(run(cbcd), orc.aes_cbc(key, DEC, iv, data)[1])
        assert np.array_equal(run(cfbd), orc.aes_cfb128(key, DEC, iv, data)[0])
        assert np.array_equal(run(cbce), orc.aes_cbc(key, ENC, iv, data)[1])
        assert np.array_equal(run(cfbe), orc.aes_cfb128(key, ENC, iv, data)[0])
        run(ctr)
        assert np.array_equal(run(rc4), orc.arc4_batch(data, 16, data, 16))

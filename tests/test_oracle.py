"""CPU: pin the oracle (oracle.c) against the reference's golden vectors, the
standard KATs, the committed digests and — when oracle/_ref is built — the
unmodified reference objects on seeded random inputs."""
import hashlib

import numpy as np
import pytest

ENC, DEC = 1, 0
h = bytes.fromhex


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a)).hexdigest()


def impls(request):
    return request.getfixturevalue(request.param)


@pytest.fixture(params=["orc", "ref"])
def impl(request):
    return request.getfixturevalue(request.param)


def test_fips197(impl, golden):
    for v in golden["kat"]["fips197"]:
        assert impl.aes_ecb(h(v["key"]), ENC, h(v["pt"])).tobytes().hex() == v["ct"]
        assert impl.aes_ecb(h(v["key"]), DEC, h(v["ct"])).tobytes().hex() == v["pt"]


def test_sp800_38a(impl, golden):
    for v in golden["kat"]["sp800_38a"]:
        key, pt = h(v["key"]), h(v["pt"])
        assert impl.aes_ecb(key, ENC, pt).tobytes().hex() == v["ecb"]
        rc, c, ivo = impl.aes_cbc(key, ENC, h(v["iv"]), pt)
        assert (rc, c.tobytes().hex(), ivo.hex()) == (0, v["cbc"], v["cbc_iv_out"])
        rc, p, ivo = impl.aes_cbc(key, DEC, h(v["iv"]), h(v["cbc"]))
        assert p.tobytes() == pt and ivo.hex() == v["cbc"][-32:]
        t, off, nco, sb = impl.aes_ctr(key, h(v["ctr"]), pt)
        assert (t.tobytes().hex(), nco.hex(), off) == (v["ctr_ct"], v["ctr_out"], 0)


def test_rfc3686_ctr(impl, golden):
    """aes.c:1025-1078, driver :1275-1323 (includes the 36-byte partial block)."""
    for v in golden["kat"]["rfc3686"]:
        t, off, nco, sb = impl.aes_ctr(h(v["key"]), h(v["nonce_counter"]), h(v["pt"]))
        assert t.tobytes().hex() == v["ct"]
        assert (off, nco.hex(), sb.hex()) == (v["nc_off_out"], v["nonce_counter_out"], v["stream_block_out"])
        back, *_ = impl.aes_ctr(h(v["key"]), h(v["nonce_counter"]), h(v["ct"]))
        assert back.tobytes().hex() == v["pt"]


def test_cfb128(impl, golden):
    """SP 800-38A CFB128 vectors of aes.c:958-1015 and the byte-granular chaining of aes.c:822-862."""
    for v in golden["kat"]["cfb128"]:
        c, off, ivo = impl.aes_cfb128(h(v["key"]), ENC, h(v["iv"]), h(v["pt"]))
        assert (c.tobytes().hex(), off, ivo.hex()) == (v["ct"], 0, v["iv_out"])
        p, off, ivo = impl.aes_cfb128(h(v["key"]), DEC, h(v["iv"]), h(v["ct"]))
        assert p.tobytes().hex() == v["pt"]
    v = golden["kat"]["edge"]["cfb_150"]
    for mode, src, dst in ((ENC, "pt", "ct"), (DEC, "ct", "pt")):
        data, iv, off, outs, pos = h(v[src]), h(v["iv"]), 0, [], 0
        for n in v["chunks"]:
            o, off, iv = impl.aes_cfb128(h(v["key"]), mode, iv, data[pos:pos + n], off)
            outs.append(o.tobytes())
            pos += n
        assert b"".join(outs).hex() == v[dst]
        assert (off, iv.hex()) == (v["iv_off_out"], v["iv_out"])
    for v in golden["vectors"]["aes"]:
        c, off, ivo = impl.aes_cfb128(h(v["key"]), ENC, h(v["iv"]), h(v["cfb_data"]))
        assert (c.tobytes().hex(), off, ivo.hex()) == (v["cfb_enc"], v["cfb_enc_off"], v["cfb_enc_iv_out"])
        c, off, ivo = impl.aes_cfb128(h(v["key"]), DEC, h(v["iv"]), h(v["cfb_data"]))
        assert (c.tobytes().hex(), off, ivo.hex()) == (v["cfb_dec"], v["cfb_dec_off"], v["cfb_dec_iv_out"])


def test_monte_carlo(orc, golden):
    """aes_self_test's ECB/CBC loops (aes.c:1106-1212): zero key, 10 000 chained blocks."""
    mc = golden["kat"]["monte_carlo"]
    for u, kb in enumerate(mc["keybits"]):
        key = bytes(kb // 8)
        for mode, name in ((ENC, "ecb_enc"), (DEC, "ecb_dec")):
            buf = bytes(16)
            for _ in range(mc["iterations"]):
                buf = orc.aes_ecb(key, mode, buf).tobytes()
            assert buf.hex() == mc[name][u]
        iv, buf = bytes(16), bytes(16)
        for _ in range(mc["iterations"]):
            rc, o, iv = orc.aes_cbc(key, DEC, iv, buf)
            buf = o.tobytes()
        assert buf.hex() == mc["cbc_dec"][u]
        iv, buf, prv = bytes(16), bytes(16), bytes(16)
        for _ in range(mc["iterations"]):
            rc, o, iv = orc.aes_cbc(key, ENC, iv, buf)
            buf, prv = prv, o.tobytes()
        assert prv.hex() == mc["cbc_enc"][u]


def test_reference_self_tests(ref):
    assert ref.self_tests() == (0, 0)   # aes_self_test 18/18, arc4_self_test 3/3


def test_rc4_vectors(impl, golden):
    for v in golden["kat"]["rc4"]:
        assert impl.arc4(h(v["key"]), h(v["pt"])).tobytes().hex() == v["ct"]
    for v in golden["vectors"]["rc4"]:
        assert impl.arc4(h(v["key"]), h(v["data"])).tobytes().hex() == v["ct"]


def test_rc4_h_semantics(orc, golden):
    """rc4_init/rc4_crypt (intended semantics) == arc4, chained calls keep state (rc4.c:56-99)."""
    for v in golden["kat"]["rc4"] + golden["vectors"]["rc4"]:
        data = h(v.get("pt", v.get("data", "")))
        out, i1, i2, perm = orc.rc4(h(v["key"]), [data])
        assert out.tobytes().hex() == v["ct"]
        cut = len(data) // 3
        out2, j1, j2, perm2 = orc.rc4(h(v["key"]), [data[:cut], data[cut:]])
        assert out2.tobytes() == out.tobytes() and (i1, i2, perm) == (j1, j2, perm2)
        ks, x, y, m = orc.arc4_keystream(h(v["key"]), [len(data)])
        assert (x, y, m) == (i1, i2, perm)        # index1<->x, index2<->y, perm<->m


def test_edge_cases(impl, golden):
    e = golden["kat"]["edge"]
    v = e["ctr_full_carry"]
    t, off, nco, sb = impl.aes_ctr(h(v["key"]), h(v["nonce_counter"]), bytes(v["len"]))
    assert (t.tobytes().hex(), nco.hex()) == (v["ct"], v["nonce_counter_out"])
    v = e["ctr_100"]
    key, pt = h(v["key"]), h(v["pt"])
    nc, sb, off, outs, pos = h(v["nonce_counter"]), bytes(16), 0, [], 0
    for n in v["chunks"]:       # 5+30+0+65 chained == one 100-byte call
        o, off, nc, sb = impl.aes_ctr(key, nc, pt[pos:pos + n], off, sb)
        outs.append(o.tobytes())
        pos += n
    assert b"".join(outs).hex() == v["ct"]
    assert (off, nc.hex(), sb.hex()) == (v["nc_off_out"], v["nonce_counter_out"], v["stream_block_out"])
    v = e["arc4_chained"]
    ks, x, y, m = impl.arc4_keystream(h(v["key"]), v["lengths"])
    assert (ks.tobytes().hex(), x, y, m.hex()) == (v["keystream"], v["x"], v["y"], v["m"])
    ks1, *_ = impl.arc4_keystream(h(v["key"]), [sum(v["lengths"])])
    assert ks1.tobytes() == ks.tobytes()


def test_error_codes(impl):
    for bits in (0, 64, 129, 512):
        assert impl.setkey_rc(bytes(64), bits, ENC) == -0x20       # aes.c:455-461
        assert impl.setkey_rc(bytes(64), bits, DEC) == -0x20       # aes.c:555-561
    rc, out, iv = impl.aes_cbc(bytes(16), ENC, bytes(range(16)), bytes(17))
    assert rc == -0x22 and iv == bytes(range(16))                   # aes.c:767-768
    rc, out, iv = impl.aes_cbc(bytes(16), ENC, bytes(range(16)), b"")
    assert rc == 0 and iv == bytes(range(16))                       # length 0: no-op


def test_committed_vectors(impl, golden):
    for v in golden["vectors"]["aes"]:
        key, data = h(v["key"]), h(v["data"])
        assert impl.packed_schedule(key, ENC).hex() == v["enc_schedule"]
        assert impl.packed_schedule(key, DEC).hex() == v["dec_schedule"]
        assert impl.aes_ecb(key, ENC, data).tobytes().hex() == v["ecb_enc"]
        assert impl.aes_ecb(key, DEC, data).tobytes().hex() == v["ecb_dec"]
        rc, c, ivo = impl.aes_cbc(key, ENC, h(v["iv"]), data)
        assert (c.tobytes().hex(), ivo.hex()) == (v["cbc_enc"], v["cbc_enc_iv_out"])
        rc, c, ivo = impl.aes_cbc(key, DEC, h(v["iv"]), data)
        assert (c.tobytes().hex(), ivo.hex()) == (v["cbc_dec"], v["cbc_dec_iv_out"])
        t, off, nco, sb = impl.aes_ctr(key, h(v["nonce_counter"]), data[: v["ctr_ragged_len"]])
        assert (t.tobytes().hex(), off, nco.hex(), sb.hex()) == (
            v["ctr"], v["ctr_nc_off_out"], v["ctr_nonce_counter_out"], v["ctr_stream_block_out"])
        if "aesni_ctr" in v:
            s = h(v["enc_schedule"])
            got = impl.aesni_ctr(s, 6 + v["keybits"] // 32, h(v["aesni_ivec"]), h(v["aesni_nonce"]), data)
            assert got.tobytes().hex() == v["aesni_ctr"]
            assert impl.aesni_ecb(s, 6 + v["keybits"] // 32, ENC, data).tobytes().hex() == v["ecb_enc"]
            assert impl.aesni_ecb(h(v["dec_schedule"]), 6 + v["keybits"] // 32, DEC, data).tobytes().hex() == v["ecb_dec"]
    v = golden["vectors"]["aesni_key256"]
    assert impl.aesni_key256(h(v["key"])).hex() == v["schedule"]


def test_golden_digests_16mib(orc, golden, harness_input):
    """SHA-256 of reference outputs on the config-1 input."""
    msg, key32 = harness_input
    dg = golden["digests"]
    assert sha(msg) == dg["msg"] and key32.hex() == dg["key"]
    z = bytes(16)
    for kb in (128, 256):
        key = key32[: kb // 8]
        assert sha(orc.aes_ecb(key, ENC, msg)) == dg[f"ecb{kb}_enc"]
        assert sha(orc.aes_ecb(key, DEC, msg)) == dg[f"ecb{kb}_dec"]
        assert sha(orc.aes_cbc(key, ENC, z, msg)[1]) == dg[f"cbc{kb}_enc"]
        assert sha(orc.aes_cbc(key, DEC, z, msg)[1]) == dg[f"cbc{kb}_dec"]
        assert sha(orc.aes_ctr(key, z, msg)[0]) == dg[f"ctr{kb}"]
        assert sha(orc.aes_cfb128(key, ENC, z, msg)[0]) == dg[f"cfb{kb}_enc"]
        assert sha(orc.aes_cfb128(key, DEC, z, msg)[0]) == dg[f"cfb{kb}_dec"]
        assert sha(orc.aesni_ctr(orc.packed_schedule(key), 6 + kb // 32, b"HloEelAS", b"315A", msg)) == dg[f"aesni_ctr{kb}"]
    assert sha(orc.arc4_batch(msg[: 4096 * 16], 16, msg, 4096)) == dg["rc4_4096x4k"]
    o, ivo = orc.aes_cbc_encrypt_batch(key32, msg[: 256 * 16].copy(), msg, 65536)
    assert (sha(o), sha(ivo)) == (dg["cbc256_enc_batch_256x64k"], dg["cbc256_enc_batch_ivs_out"])


def test_oracle_vs_reference_random(orc, ref):
    """Differential: oracle.c vs the unmodified reference objects, seeded random inputs."""
    rng = np.random.default_rng(7)
    for trial in range(24):
        kb = (128, 192, 256)[trial % 3]
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        n = int(rng.integers(0, 200)) * 16
        data = rng.integers(0, 256, n, dtype=np.uint8)
        iv = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        for mode in (ENC, DEC):
            assert np.array_equal(orc.schedule(key, mode), ref.schedule(key, mode))
            assert np.array_equal(orc.aes_ecb(key, mode, data), ref.aes_ecb(key, mode, data))
            a, b = orc.aes_cbc(key, mode, iv, data), ref.aes_cbc(key, mode, iv, data)
            assert a[0] == b[0] and np.array_equal(a[1], b[1]) and a[2] == b[2]
        m = int(rng.integers(0, 3000))
        rag = rng.integers(0, 256, m, dtype=np.uint8)
        off0 = int(rng.integers(0, 16))
        sb0 = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        nc = bytes(rng.integers(0, 256, 16 - (trial % 9), dtype=np.uint8)) + b"\xff" * (trial % 9)
        a, b = orc.aes_ctr(key, nc, rag, off0, sb0), ref.aes_ctr(key, nc, rag, off0, sb0)
        assert np.array_equal(a[0], b[0]) and a[1:] == b[1:]
        if ref.has_aesni():
            s = ref.packed_schedule(key, ENC)
            ivec = bytes(rng.integers(0, 256, 4, dtype=np.uint8)) + (b"\xff\xff\xff\xff" if trial % 2 else bytes(4))
            nonce = rng.integers(0, 256, 4, dtype=np.uint8).tobytes()
            assert np.array_equal(orc.aesni_ctr(s, 6 + kb // 32, ivec, nonce, data),
                                  ref.aesni_ctr(s, 6 + kb // 32, ivec, nonce, data))
        klen = int(rng.integers(1, 64))
        rk = rng.integers(0, 256, klen, dtype=np.uint8).tobytes()
        assert np.array_equal(orc.arc4(rk, rag), ref.arc4(rk, rag))
    keys = rng.integers(0, 256, 64 * 16, dtype=np.uint8)
    data = rng.integers(0, 256, 64 * 4096, dtype=np.uint8)
    assert np.array_equal(orc.arc4_batch(keys, 16, data, 4096), ref.arc4_batch(keys, 16, data, 4096))

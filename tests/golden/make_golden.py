#!/usr/bin/env python
"""Generate the committed golden fixtures from the UNMODIFIED reference.

Run in a container that has /root/reference (so oracle/_ref/libref.so can be
built): ``python tests/golden/make_golden.py``.  Every value written here is
produced by the reference's own aes.c / aesni.c / arc4.c objects; the standard
vectors in ``kat.json`` are additionally asserted against the reference at
generation time, so each one is both a published KAT and a reference parity
point.  The GPU box has no /root/reference: tests there
read only the JSON files.

Outputs (tests/golden/):
  kat.json      FIPS-197 App. C, SP 800-38A F.1/F.2/F.5, RFC 3686 (aes.c:1025-1078),
                Rijndael Monte-Carlo (aes.c:912-950), RC4 vectors (arc4.c:124-143),
                counter-carry and chaining edge cases
  digests.json  SHA-256 of reference outputs on the 16 MiB BASELINE config-1
                input (srand(1337); rand()%255 — SURVEY Appendix C)
  vectors.json  seeded random inputs with reference outputs for every mode and
                key size (hex, a few KiB each)
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ENC, DEC = 1, 0
h = bytes.fromhex


def main():
    oracle.build()
    ref = oracle.Ref()
    orc = oracle.Oracle()
    assert ref.self_tests() == (0, 0), "reference self-tests failed"

    # ------------------------------------------------------------ kat.json
    kat = {"source": "standards + reference self-test data; each asserted against libref.so"}
    k128 = h("2b7e151628aed2a6abf7158809cf4f3c")
    k256 = h("603deb1015ca71be2b73aef0857d77811f352c073b6108d72d9810a30914dff4")
    pt = h("6bc1bee22e409f96e93d7e117393172aae2d8a571e03ac9c9eb76fac45af8e51"
           "30c81c46a35ce411e5fbc1191a0a52eff69f2445df4f9b17ad2b417be66c3710")
    iv = h("000102030405060708090a0b0c0d0e0f")
    ctr0 = h("f0f1f2f3f4f5f6f7f8f9fafbfcfdfeff")
    sp = []
    for name, key in (("128", k128), ("256", k256)):
        e = ref.aes_ecb(key, ENC, pt).tobytes()
        rc, c, ivo = ref.aes_cbc(key, ENC, iv, pt)
        t, off, nco, sb = ref.aes_ctr(key, ctr0, pt)
        sp.append({"keybits": int(name), "key": key.hex(), "pt": pt.hex(), "iv": iv.hex(), "ctr": ctr0.hex(),
                   "ecb": e.hex(), "cbc": c.tobytes().hex(), "cbc_iv_out": ivo.hex(),
                   "ctr_ct": t.tobytes().hex(), "ctr_out": nco.hex()})
    # published values (SP 800-38A F.1.1, F.2.1, F.5.1, F.1.5, F.2.5, F.5.5)
    assert sp[0]["ecb"].startswith("3ad77bb40d7a3660a89ecaf32466ef97") and sp[0]["ecb"].endswith("7b0c785e27e8ad3f8223207104725dd4")
    assert sp[0]["cbc"].startswith("7649abac8119b246cee98e9b12e9197d") and sp[0]["cbc"].endswith("3ff1caa1681fac09120eca307586e1a7")
    assert sp[0]["ctr_ct"].startswith("874d6191b620e3261bef6864990db6ce") and sp[0]["ctr_ct"].endswith("1e031dda2fbe03d1792170a0f3009cee")
    assert sp[1]["ecb"].startswith("f3eed1bdb5d2a03c064b5a7e3db181f8") and sp[1]["ecb"].endswith("23304b7a39f9f3ff067d8d8f9e24ecc7")
    assert sp[1]["cbc"].startswith("f58c4c04d6e5f1ba779eabfb5f7bfbd6") and sp[1]["cbc"].endswith("b2eb05e2c39be9fcda6c19078c6a9d1b")
    assert sp[1]["ctr_ct"].startswith("601ec313775789a5b7a7f504bbf3d228") and sp[1]["ctr_ct"].endswith("dfc9c58db67aada613c2dd08457941a6")
    kat["sp800_38a"] = sp

    fips = []
    for kb, exp in ((128, "69c4e0d86a7b0430d8cdb78070b4c55a"), (192, "dda97ca4864cdfe06eaf70a0ec0d7191"),
                    (256, "8ea2b7ca516745bfeafc49904b496089")):
        key = bytes(range(kb // 8))
        p = h("00112233445566778899aabbccddeeff")
        c = ref.aes_ecb(key, ENC, p).tobytes()
        assert c.hex() == exp, (kb, c.hex())
        assert ref.aes_ecb(key, DEC, c).tobytes() == p
        fips.append({"keybits": kb, "key": key.hex(), "pt": p.hex(), "ct": exp})
    kat["fips197"] = fips

    # RFC 3686 vectors as held by the reference self-test (aes.c:1025-1078)
    rfc = [
        ("ae6852f8121067cc4bf7a5765577f39e", "00000030000000000000000000000001",
         "53696e676c6520626c6f636b206d7367", "e4095d4fb7a7b3792d6175a3261311b8"),
        ("7e24067817fae0d743d6ce1f32539163", "006cb6dbc0543b59da48d90b00000001",
         bytes(range(32)).hex(), "5104a106168a72d9790d41ee8edad388eb2e1efc46da57c8fce630df9141be28"),
        ("7691be035e5020a8ac6e618529f9a0dc", "00e0017b27777f3f4a1786f000000001",
         bytes(range(36)).hex(), "c1cf48a89f2ffdd9cf4652e9efdb72d74540a42bde6d7836d59a5ceaaef3105325b2072f"),
    ]
    out = []
    for key, nc, p, c in rfc:
        t, off, nco, sb = ref.aes_ctr(h(key), h(nc), h(p))
        assert t.tobytes().hex() == c
        out.append({"key": key, "nonce_counter": nc, "pt": p, "ct": c, "nc_off_out": off,
                    "nonce_counter_out": nco.hex(), "stream_block_out": sb.hex()})
    kat["rfc3686"] = out

    # Rijndael Monte-Carlo as run by aes_self_test (aes.c:1106-1212): zero key, zero block, 10000 iterations
    mc_expected = {
        "ecb_dec": ["44416ac2d1f53c583303917e6be9ebe0", "48e31e9e256718f29229319c19f15ba4", "058ccffdbbcb382d1f6f56585d8a4ade"],
        "ecb_enc": ["c34c052cc0da8d73451afe5f03be297f", "f3f6752ae8d7831138f041560631b114", "8b79eecc93a0ee5dff30b4ea21636da4"],
        "cbc_dec": ["faca37e0b0c85373df706e73f7c9af86", "5df678dd17ba4e75b61768c6adef7c7b", "4804e1818fe6297519a3e88c57310413"],
        "cbc_enc": ["8a05fc5e095af4848a08d328d3688e3d", "7bd966d53ad8c1bb85d2adfae87bb104", "fe3c53653e2f45b56fcd88b2cc898ff0"],
    }
    kat["monte_carlo"] = {"iterations": 10000, "keybits": [128, 192, 256], **mc_expected}

    # SP 800-38A F.3.13-F.3.18 CFB128 vectors as held by the reference (aes.c:958-1015; the mode itself
    # is compiled out of the reference build, oracle/Makefile turns it on with the source's own switch)
    cfb_ct = {
        128: "3b3fd92eb72dad20333449f8e83cfb4ac8a64537a0b3a93fcde3cdad9f1ce58b26751f67a3cbb140b1808cf187a4f4dfc04b05357c5d1c0eeac4c66f9ff7f2e6",
        192: "cdc80d6fddf18cab34c25909c99a417467ce7f7f81173621961a2b70171d3d7a2e1e8a1dd59b88b1c8e60fed1efac4c9c05f9f9ca9834fa042ae8fba584b09ff",
        256: "dc7e84bfda79164b7ecd8486985d386039ffed143b28b1c832113c6331e5407bdf10132415e54b92a13ed0a8267ae2f975a385741ab9cef82031623d55b1e471",
    }
    k192 = h("8e73b0f7da0e6452c810f32b809079e562f8ead2522c6b7b")
    cfb = []
    for kb, key in ((128, k128), (192, k192), (256, k256)):
        c, off, ivo = ref.aes_cfb128(key, ENC, iv, pt)
        assert c.tobytes().hex() == cfb_ct[kb] and off == 0
        p2, off, ivo2 = ref.aes_cfb128(key, DEC, iv, c)
        assert p2.tobytes() == pt
        cfb.append({"keybits": kb, "key": key.hex(), "iv": iv.hex(), "pt": pt.hex(), "ct": cfb_ct[kb], "iv_out": ivo.hex()})
    kat["cfb128"] = cfb

    # RC4 (arc4.c:124-143) + the "Key"/"Plaintext" check of SURVEY Appendix B
    rc4 = []
    for key, p, c in (("0123456789abcdef", "0123456789abcdef", "75b7878099e0c596"),
                      ("0123456789abcdef", "0000000000000000", "7494c2e7104b0879"),
                      ("0000000000000000", "0000000000000000", "de188941a3375d3a"),
                      (b"Key".hex(), b"Plaintext".hex(), "bbf316e8d940af0ad3")):
        assert ref.arc4(h(key), h(p)).tobytes().hex() == c
        rc4.append({"key": key, "pt": p, "ct": c})
    kat["rc4"] = rc4

    # Appendix D edge cases, recorded from the reference
    edge = {}
    key = h("000102030405060708090a0b0c0d0e0f")
    t, off, nco, sb = ref.aes_ctr(key, h("12" + "ff" * 15), bytes(32))
    edge["ctr_full_carry"] = {"key": key.hex(), "nonce_counter": "12" + "ff" * 15, "len": 32,
                              "ct": t.tobytes().hex(), "nonce_counter_out": nco.hex()}
    assert nco.hex() == "13" + "00" * 14 + "01"
    data = bytes((i * 7 + 3) & 0xFF for i in range(100))
    t, off, nco, sb = ref.aes_ctr(key, h("ab" * 16), data)
    edge["ctr_100"] = {"key": key.hex(), "nonce_counter": "ab" * 16, "pt": data.hex(), "ct": t.tobytes().hex(),
                       "nc_off_out": off, "nonce_counter_out": nco.hex(), "stream_block_out": sb.hex(),
                       "chunks": [5, 30, 0, 65]}
    ks, x, y, m = ref.arc4_keystream(b"Secret", [7, 293])
    edge["arc4_chained"] = {"key": b"Secret".hex(), "lengths": [7, 293], "keystream": ks.tobytes().hex(),
                            "x": x, "y": y, "m": m.hex()}
    data = bytes((i * 11 + 5) & 0xFF for i in range(150))
    c, off, ivo = ref.aes_cfb128(key, ENC, h("cd" * 16), data)
    edge["cfb_150"] = {"key": key.hex(), "iv": "cd" * 16, "pt": data.hex(), "ct": c.tobytes().hex(),
                       "iv_off_out": off, "iv_out": ivo.hex(), "chunks": [3, 40, 0, 16, 91]}
    kat["edge"] = edge
    json.dump(kat, open(os.path.join(HERE, "kat.json"), "w"), indent=1)

    # -------------------------------------------------------- digests.json
    msg, key32 = orc.harness_input(16 << 20)
    assert msg[:16].tobytes().hex() == "656643de18d979fc085fe8e659e78ca8"
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a)).hexdigest()
    dg = {"input": "srand(1337); msg[i]=rand()%255 (16 MiB); key[i]=rand()%255 (32 B); iv / nonce_counter = 16 zero bytes",
          "msg": sha(msg), "key": key32.hex()}
    z = bytes(16)
    for kb in (128, 256):
        key = key32[: kb // 8]
        dg[f"ecb{kb}_enc"] = sha(ref.aes_ecb(key, ENC, msg))
        dg[f"ecb{kb}_dec"] = sha(ref.aes_ecb(key, DEC, msg))
        dg[f"cbc{kb}_enc"] = sha(ref.aes_cbc(key, ENC, z, msg)[1])
        dg[f"cbc{kb}_dec"] = sha(ref.aes_cbc(key, DEC, z, msg)[1])
        dg[f"ctr{kb}"] = sha(ref.aes_ctr(key, z, msg)[0])
        dg[f"cfb{kb}_enc"] = sha(ref.aes_cfb128(key, ENC, z, msg)[0])
        dg[f"cfb{kb}_dec"] = sha(ref.aes_cfb128(key, DEC, z, msg)[0])
    keys = msg[: 4096 * 16]
    dg["rc4_4096x4k"] = sha(ref.arc4_batch(keys, 16, msg, 4096))
    # cross-check against SURVEY Appendix C
    assert dg["ctr128"] == "467d1fe6f3378218dd12ce3628a689d99f4172cb606a7c564f8d6ec59b60e462"
    assert dg["cbc256_dec"] == "f31cbfcfe0b439749895d1ad755fec79128bafd42e00d10b2c21f56a09900420"
    assert dg["rc4_4096x4k"] == "109f647611066c1a8cab28e08773d3feb2a02a151e4ecf5972825473596410b5"
    # CBC-encrypt over independent 64 KiB messages (config-3 shape, 256 messages), per-message IV = msg bytes
    ivs = msg[:256 * 16].copy()
    o, ivo = ref.aes_cbc_encrypt_batch(key32, ivs, msg, 65536)
    dg["cbc256_enc_batch_256x64k"] = sha(o)
    dg["cbc256_enc_batch_ivs_out"] = sha(ivo)
    # AES-NI path (aesni.c) on the same input: AES_CTR_encrypt(ivec="HloEelAS", nonce="315A") as aes-modes/test.c:305-308
    if ref.has_aesni():
        for kb in (128, 256):
            key = key32[: kb // 8]
            s = ref.packed_schedule(key, ENC)
            dg[f"aesni_ctr{kb}"] = sha(ref.aesni_ctr(s, 6 + kb // 32, b"HloEelAS", b"315A", msg))
            assert dg[f"ecb{kb}_enc"] == sha(ref.aesni_ecb(s, 6 + kb // 32, ENC, msg))
            assert dg[f"ecb{kb}_dec"] == sha(ref.aesni_ecb(ref.packed_schedule(key, DEC), 6 + kb // 32, DEC, msg))
    json.dump(dg, open(os.path.join(HERE, "digests.json"), "w"), indent=1)

    # -------------------------------------------------------- vectors.json
    rng = np.random.default_rng(20261106)
    vec = []
    for kb in (128, 192, 256):
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        data = rng.integers(0, 256, 1024 + 16 * (kb // 64), dtype=np.uint8)
        ivv = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        nc = rng.integers(0, 256, 12, dtype=np.uint8).tobytes() + h("fffffffa")   # crosses the 2^32 carry
        ragged = data[: data.size - 5]
        ct, off, nco, sb = ref.aes_ctr(key, nc, ragged)
        ent = {"keybits": kb, "key": key.hex(), "data": data.tobytes().hex(), "iv": ivv.hex(), "nonce_counter": nc.hex(),
               "enc_schedule": ref.packed_schedule(key, ENC).hex(), "dec_schedule": ref.packed_schedule(key, DEC).hex(),
               "ecb_enc": ref.aes_ecb(key, ENC, data).tobytes().hex(), "ecb_dec": ref.aes_ecb(key, DEC, data).tobytes().hex()}
        rc, c, ivo = ref.aes_cbc(key, ENC, ivv, data)
        ent["cbc_enc"], ent["cbc_enc_iv_out"] = c.tobytes().hex(), ivo.hex()
        rc, c, ivo = ref.aes_cbc(key, DEC, ivv, data)
        ent["cbc_dec"], ent["cbc_dec_iv_out"] = c.tobytes().hex(), ivo.hex()
        ent["ctr_ragged_len"] = int(ragged.size)
        ent["ctr"], ent["ctr_nc_off_out"], ent["ctr_nonce_counter_out"], ent["ctr_stream_block_out"] = \
            ct.tobytes().hex(), off, nco.hex(), sb.hex()
        if ref.has_aesni():
            ent["aesni_ivec"], ent["aesni_nonce"] = "fffffffffffffffe", "01020304"   # 64-bit counter wraps inside
            ent["aesni_ctr"] = ref.aesni_ctr(ref.packed_schedule(key, ENC), 6 + kb // 32,
                                             h(ent["aesni_ivec"]), h(ent["aesni_nonce"]), data).tobytes().hex()
        rng2 = np.random.default_rng(kb)              # separate stream: keeps the older vectors stable
        rag = rng2.integers(0, 256, 1000 + kb // 8 + 7, dtype=np.uint8)
        c, off, ivo = ref.aes_cfb128(key, ENC, ivv, rag)
        ent["cfb_data"], ent["cfb_enc"], ent["cfb_enc_off"], ent["cfb_enc_iv_out"] = rag.tobytes().hex(), c.tobytes().hex(), off, ivo.hex()
        c, off, ivo = ref.aes_cfb128(key, DEC, ivv, rag)
        ent["cfb_dec"], ent["cfb_dec_off"], ent["cfb_dec_iv_out"] = c.tobytes().hex(), off, ivo.hex()
        vec.append(ent)
    k256 = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
    aes256_sched = ref.aesni_key256(k256)
    assert aes256_sched == ref.packed_schedule(k256, ENC)
    rc4v = []
    for klen, n in ((1, 33), (5, 300), (16, 4096), (40, 1000), (256, 513)):
        key = rng.integers(0, 256, klen, dtype=np.uint8).tobytes()
        data = rng.integers(0, 256, n, dtype=np.uint8)
        rc4v.append({"key": key.hex(), "data": data.tobytes().hex(), "ct": ref.arc4(key, data).tobytes().hex()})
    json.dump({"aes": vec, "aesni_key256": {"key": k256.hex(), "schedule": aes256_sched.hex()}, "rc4": rc4v},
              open(os.path.join(HERE, "vectors.json"), "w"), indent=1)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()

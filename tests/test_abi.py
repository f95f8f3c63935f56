"""CPU: the C-ABI library builds, loads, exports every symbol include/*.h
declares, and its host-only logic (key schedules, argument checks) matches the
oracle.  No cipher compute is attempted without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT

ENC, DEC = 1, 0
h = bytes.fromhex


def header_functions():
    names = set()
    for hdr in ("aes.h", "aesni.h", "arc4.h", "rc4.h", "bulkcipher.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for m in re.finditer(r"^\s*(?:const\s+)?(?:unsigned\s+)?[A-Za-z_][\w ]*?[\s\*]+(\w+)\s*\(", text, flags=re.M):
            n = m.group(1)
            if n not in ("defined", "__attribute__", "aligned", "__declspec", "align"):
                names.add(n)
    return names


def test_build_and_symbols(pkg):
    import __graft_entry__ as g
    g.build()
    L = pkg.lib()
    declared = header_functions()
    assert len(declared) >= 40
    # KEY_256_ASSIST_*: exported by the reference's aesni.o but absent from its header (aesni.c:13-36) — same here
    assert set(pkg.DROPIN_SYMBOLS + pkg.BC_SYMBOLS) <= declared | {"bc_xor_dev", "bc_xor_host", "KEY_256_ASSIST_1", "KEY_256_ASSIST_2"}
    for name in pkg.DROPIN_SYMBOLS + pkg.BC_SYMBOLS:
        assert hasattr(L, name), f"libbulkcipher.so does not export {name}"
    for name in sorted(declared):
        assert hasattr(L, name), f"libbulkcipher.so does not export {name} declared in include/"


def test_struct_layouts(pkg):
    """ABI of aes-modes/aes.h:41-47, arc4.h:35-41, rc4.h:43-47 (560 / 264 / 264 bytes)."""
    assert C.sizeof(pkg.aes_context) == 560
    assert C.sizeof(pkg.arc4_context) == 264
    assert C.sizeof(pkg.rc4_state) == 264
    assert pkg.aes_context.rk.offset == 8 and pkg.aes_context.buf.offset == 16
    assert pkg.arc4_context.m.offset == 8 and pkg.rc4_state.index1.offset == 256


def test_key_schedules_match_oracle(pkg, orc, golden):
    rng = np.random.default_rng(3)
    keys = [h(v["key"]) for v in golden["vectors"]["aes"]]
    keys += [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in (16, 24, 32) for _ in range(8)]
    for key in keys:
        rc, ctx = pkg.aes_setkey_enc(key)
        assert rc == 0 and ctx.nr == len(key) // 4 + 6
        assert C.addressof(ctx.rk.contents) == C.addressof(ctx.buf)        # rk points into buf (aes.c:466)
        assert np.array_equal(pkg.schedule_words(ctx), orc.schedule(key, ENC))
        assert all(ctx.buf[i] >> 32 == 0 for i in range(4 * (ctx.nr + 1)))  # upper halves zero
        rc, ctx = pkg.aes_setkey_dec(key)
        assert rc == 0 and np.array_equal(pkg.schedule_words(ctx), orc.schedule(key, DEC))
        assert pkg.AES_Key_Expansion(key) == orc.packed_schedule(key, ENC)
    for v in golden["vectors"]["aes"]:
        rc, ctx = pkg.aes_setkey_enc(h(v["key"]))
        assert pkg.schedule_words(ctx).astype("<u4").tobytes().hex() == v["enc_schedule"]
        rc, ctx = pkg.aes_setkey_dec(h(v["key"]))
        assert pkg.schedule_words(ctx).astype("<u4").tobytes().hex() == v["dec_schedule"]
    v = golden["vectors"]["aesni_key256"]
    assert pkg.AES_Key_Expansion(h(v["key"])).hex() == v["schedule"]     # aesni.c:38-77


def test_error_codes_without_compute(pkg):
    for bits in (0, 64, 129, 512):
        rc, ctx = pkg.aes_setkey_enc(bytes(64), bits)
        assert rc == pkg.ERR_INVALID_KEY_LENGTH and ctx.nr == 0           # ctx untouched (aes.c:455-461)
        rc, ctx = pkg.aes_setkey_dec(bytes(64), bits)
        assert rc == pkg.ERR_INVALID_KEY_LENGTH
    rc, ctx = pkg.aes_setkey_enc(bytes(16))
    rc, out, iv = pkg.aes_crypt_cbc(ctx, ENC, bytes(range(16)), bytes(17))
    assert rc == pkg.ERR_INVALID_INPUT_LENGTH and iv == bytes(range(16))  # aes.c:767-768
    rc, out, iv = pkg.aes_crypt_cbc(ctx, ENC, bytes(range(16)), b"")
    assert rc == 0 and iv == bytes(range(16))                             # length 0 is a no-op
    L = pkg.lib()
    assert L.bc_aes_ecb_dev(None, 10, 1, None, None, 0, None) == pkg.BC_ERR_BAD_ARG
    rk = np.zeros(60, np.uint32)
    assert L.bc_aes_ecb_dev(rk.ctypes.data, 11, 1, None, None, 0, None) == pkg.BC_ERR_BAD_ARG
    assert "nr" in pkg.last_error()
    assert L.bc_aes_ctr_dev(rk.ctypes.data, 10, bytes(16), 96, 0, None, None, 0, None, None) == pkg.BC_ERR_BAD_ARG
    assert L.bc_aes_cbc_decrypt_dev(rk.ctypes.data, 10, bytes(16), 16, 16, 24, None) == pkg.BC_ERR_BAD_ARG
    assert L.bc_aes_ecb_dev(rk.ctypes.data, 10, 1, 4096, 4096 + 32, 64, None) == pkg.BC_ERR_OVERLAP


def test_key_256_assist_helpers_match_reference(pkg, ref):
    """KEY_256_ASSIST_1/2 (aesni.c:13-36) against the unmodified reference object, and chained the way
    AES_256_Key_Expansion chains them (aesni.c:38-77) they rebuild the schedule."""
    L = pkg.lib()
    rng = np.random.default_rng(41)
    for _ in range(50):
        a, b = rng.integers(0, 256, 16, dtype=np.uint8), rng.integers(0, 256, 16, dtype=np.uint8)
        for name in ("KEY_256_ASSIST_1", "KEY_256_ASSIST_2"):
            m1, m2, r1, r2 = a.copy(), b.copy(), a.copy(), b.copy()
            getattr(L, name)(C.c_void_p(m1.ctypes.data), C.c_void_p(m2.ctypes.data))
            getattr(ref.lib, "ref_" + name)(C.c_void_p(r1.ctypes.data), C.c_void_p(r2.ctypes.data))
            assert np.array_equal(m1, r1) and np.array_equal(m2, r2), name


def test_void_dropins_report_instead_of_aborting(pkg):
    """A failing void entry point leaves its output untouched and parks the code in bc_void_status()."""
    L = pkg.lib()
    assert L.bc_void_status() == 0
    src, dst = np.arange(32, dtype=np.uint8), np.full(32, 0xEE, np.uint8)
    L.AES_ECB_encrypt(src.ctypes.data, dst.ctypes.data, 32, bytes(240), 11)          # 11 rounds: no such AES
    assert L.bc_void_status() == pkg.BC_ERR_BAD_ARG and "rounds" in pkg.last_error()
    assert (dst == 0xEE).all()
    assert L.bc_void_status() == 0                                                    # cleared on read
    st = pkg.rc4_state()
    L.rc4_init(C.byref(st), b"k", 0)                                                  # rc4.c:69 would divide by zero
    assert L.bc_void_status() == pkg.BC_ERR_BAD_ARG and not any(st.perm)


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the compute entry points must fail, never emulate."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = pkg.lib()
    assert L.bc_device_count() == 0
    assert L.CheckAESSupport() == 0
    rk = np.zeros(60, np.uint32)
    buf = np.zeros(64, np.uint8)
    rc = L.bc_aes_ecb_host(rk.ctypes.data, 10, 1, buf.ctypes.data, buf.ctypes.data, 64)
    assert rc in (pkg.BC_ERR_NO_DEVICE, pkg.BC_ERR_CUDA) and not buf.any()


def test_product_does_not_reference_oracle():
    """The product path may not import, link or call anything under oracle/."""
    pkgdir = os.path.join(ROOT, "our-tree_b200")
    for dirpath, _, files in os.walk(pkgdir):
        if os.path.basename(dirpath) == "build":
            continue
        for f in files:
            if f.endswith((".c", ".cu", ".h", ".cuh", ".py", ".inc")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "orc_" not in text and "libref" not in text, f
                assert not re.search(r"^\s*(import|from)\s+oracle", text, flags=re.M), f


def test_host_layer_under_sanitizers():
    """host/*.c under AddressSanitizer + UBSan with stubbed launchers (tests/c/host_san_test.c): key
    schedules vs FIPS-197 A.1/A.3, error codes, "no device => fail, never emulate", the staging copy pool."""
    import shutil
    import subprocess
    if not os.path.exists("/usr/bin/gcc") and not shutil.which("gcc"):
        pytest.skip("no gcc")
    pkgdir = os.path.join(ROOT, "our-tree_b200")
    b = subprocess.run(["make", "-C", pkgdir, "host-san-test"], capture_output=True, text=True)
    if b.returncode != 0 and ("-lasan" in b.stderr or "libasan" in b.stderr):
        pytest.skip("sanitizer runtime not installed")
    assert b.returncode == 0, b.stdout + b.stderr
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0", UBSAN_OPTIONS="halt_on_error=1")
    r = subprocess.run([os.path.join(pkgdir, "build", "host_san_test")], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and "host_san_test: ok" in r.stdout, r.stdout + r.stderr

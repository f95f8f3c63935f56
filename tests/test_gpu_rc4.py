"""GPU parity, RC4 / ARC4: batched kernel and the single-stream drop-ins vs the
oracle (arc4.c semantics; rc4.h pinned through the state mapping)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import gpu_util as G

pytestmark = pytest.mark.gpu
h = bytes.fromhex


def test_rc4_vectors_dropin(pkg, golden):
    """arc4_self_test's three vectors (arc4.c:124-143) + Key/Plaintext, via setup -> prep -> crypt."""
    assert pkg.lib().arc4_self_test(0) == 0
    for v in golden["kat"]["rc4"] + golden["vectors"]["rc4"]:
        data = h(v.get("pt", v.get("data", "")))
        ctx = pkg.arc4_setup(h(v["key"]))
        ks = pkg.arc4_prep(ctx, len(data))
        assert pkg.arc4_crypt(data, ks).tobytes().hex() == v["ct"]
        st = pkg.rc4_init(h(v["key"]))                                   # rc4.h face, same bytes
        assert pkg.rc4_crypt(st, data).tobytes().hex() == v["ct"]


def test_arc4_state_and_chaining(pkg, orc, golden):
    v = golden["kat"]["edge"]["arc4_chained"]
    ctx = pkg.arc4_setup(h(v["key"]))
    _, x0, y0, m0 = orc.arc4_keystream(h(v["key"]), [])
    assert (ctx.x, ctx.y, bytes(ctx.m)) == (x0, y0, m0)                  # KSA state (arc4.c:43-67)
    ks = np.concatenate([pkg.arc4_prep(ctx, n) for n in v["lengths"]])   # 7 + 293 chained
    assert ks.tobytes().hex() == v["keystream"]
    assert (ctx.x, ctx.y, bytes(ctx.m).hex()) == (v["x"], v["y"], v["m"])
    # rc4.h: chained rc4_crypt calls carry index1/index2/perm
    data = bytes(range(200))
    want, i1, i2, perm = orc.rc4(h(v["key"]), [data])
    st = pkg.rc4_init(h(v["key"]))
    got = np.concatenate([pkg.rc4_crypt(st, data[:33]), pkg.rc4_crypt(st, data[33:])])
    assert got.tobytes() == want.tobytes()
    assert (st.index1, st.index2, bytes(st.perm)) == (i1, i2, perm)


@pytest.mark.parametrize("keylen,length,n", [(16, 4096, 1000), (16, 64, 300), (5, 100, 257), (1, 33, 40),
                                              (40, 1000, 513), (256, 513, 64), (16, 0, 10), (24, 16, 2049)])
def test_batch_vs_oracle(pkg, orc, keylen, length, n):
    rng = np.random.default_rng(keylen * 7919 + length)
    keys = rng.integers(0, 256, n * keylen, dtype=np.uint8)
    data = rng.integers(0, 256, n * length, dtype=np.uint8)
    d_keys = G.to_dev(keys)
    d_in = G.to_dev(data) if length else None
    d_out = G.to_dev(np.zeros_like(data)) if length else None
    d_ctx = G.to_dev(np.zeros(n * 264, np.uint8))
    G.arc4_dev(pkg, n, d_keys, keylen, length, d_in, d_out, None, d_ctx)
    ctx = G.host(d_ctx).reshape(n, 264)
    if length:
        assert np.array_equal(G.host(d_out), orc.arc4_batch(keys, keylen, data, length))
    for s in (0, n // 2, n - 1):                                         # final state of sampled streams
        _, x, y, m = orc.arc4_keystream(keys[s * keylen:(s + 1) * keylen].tobytes(), [length])
        assert int.from_bytes(ctx[s, 0:4].tobytes(), "little") == x
        assert int.from_bytes(ctx[s, 4:8].tobytes(), "little") == y
        assert ctx[s, 8:].tobytes() == m
    if length:
        # resume: second half from the saved contexts == one long call; in place; keystream-only
        half = G.to_dev(np.zeros(n * 264, np.uint8))
        cut = length // 2
        a = np.ascontiguousarray(data.reshape(n, length)[:, :cut]).reshape(-1)
        b = np.ascontiguousarray(data.reshape(n, length)[:, cut:]).reshape(-1)
        want = orc.arc4_batch(keys, keylen, data, length).reshape(n, length)
        d_a, d_b = G.to_dev(a) if cut else None, G.to_dev(b)
        G.arc4_dev(pkg, n, d_keys, keylen, cut, d_a, d_a, None, half)
        G.arc4_dev(pkg, n, None, 0, length - cut, d_b, d_b, half, half)
        if cut:
            assert np.array_equal(G.host(d_a).reshape(n, cut), want[:, :cut])
        assert np.array_equal(G.host(d_b).reshape(n, length - cut), want[:, cut:])
        assert np.array_equal(G.host(half), G.host(d_ctx))
        d_ks = G.to_dev(np.zeros_like(data))
        G.arc4_dev(pkg, n, d_keys, keylen, length, None, d_ks)
        assert np.array_equal(G.host(d_ks) ^ data, want.reshape(-1))


def test_golden_digest_4096_streams(pkg, golden, harness_input):
    """4096 streams x 4 KiB on the config-1 input."""
    msg, _ = harness_input
    keys = msg[: 4096 * 16]
    got = pkg.arc4_batch_host(keys, 16, msg, 4096)
    assert hashlib.sha256(got).hexdigest() == golden["digests"]["rc4_4096x4k"]


def test_reference_arc4_agrees_when_present(pkg, ref):
    rng = np.random.default_rng(4)
    keys = rng.integers(0, 256, 512 * 16, dtype=np.uint8)
    data = rng.integers(0, 256, 512 * 4096, dtype=np.uint8)
    assert np.array_equal(pkg.arc4_batch_host(keys, 16, data, 4096), ref.arc4_batch(keys, 16, data, 4096))

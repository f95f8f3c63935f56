"""GPU parity at BASELINE.json's full sizes (configs 2-5 at one GPU).  The oracle
cannot chew 4 GiB in seconds, so each case combines (a) sampled windows
compared bit-for-bit with the oracle — always including the seams: buffer
start/end, the 2^32 counter carry, row-group and shard boundaries — with (b)
size-independent properties checked over the WHOLE buffer on the device:
encrypt->decrypt round trips, CTR linearity (out ^ in is the keystream),
sharded == unsharded, in-place == out-of-place."""
import numpy as np
import pytest
import torch

import gpu_util as G

pytestmark = pytest.mark.gpu
ENC, DEC = 1, 0
GIB = 1 << 30
N = 4 * GIB
WIN = 1 << 16


def add_counter(nc: bytes, n: int) -> bytes:
    return ((int.from_bytes(nc, "big") + n) % (1 << 128)).to_bytes(16, "big")


def windows(n, win, extra=()):
    rng = np.random.default_rng(12345)
    offs = {0, n - win, n // 2, n // 2 - 4096, *extra}
    offs |= {int(rng.integers(0, n // 4096 - win // 4096)) * 4096 for _ in range(6)}
    return sorted(o for o in offs if 0 <= o <= n - win)


@pytest.fixture(scope="module")
def big():
    if torch.cuda.mem_get_info()[0] < 20 * GIB:
        pytest.skip("needs 20 GiB of free HBM")
    g = torch.Generator(device="cuda").manual_seed(1337)
    d_in = torch.randint(0, 256, (N,), dtype=torch.uint8, device="cuda", generator=g)
    d_out = torch.empty_like(d_in)
    yield d_in, d_out
    del d_in, d_out
    torch.cuda.empty_cache()


def test_config2_ctr128_4gib(pkg, orc, big):
    d_in, d_out = big
    key = bytes.fromhex("f9dfdf0c3cf14e4b9c296478904f979b")
    nc = bytes.fromhex("0f1e2d3c4b5a69788796a5b4fffffff0")          # 2^32 carry crossed at block 16
    rk, nr = G.sched(pkg, key, ENC)
    G.ctr_dev(pkg, rk, nr, nc, d_in, d_out, N)
    for off in windows(N, WIN, extra=(4096, 2 * GIB - WIN // 2)):
        want = orc.aes_ctr(key, add_counter(nc, off // 16), G.host(d_in[off:off + WIN]))[0]
        assert np.array_equal(G.host(d_out[off:off + WIN]), want), off
    # linearity over the whole buffer: out ^ in == E(counter stream) == CTR of zeros
    d_ks = torch.zeros_like(d_in)
    G.ctr_dev(pkg, rk, nr, nc, d_ks, d_ks, N)
    torch.cuda.synchronize()
    assert torch.equal(torch.bitwise_xor(d_out, d_in), d_ks)
    del d_ks
    # sharded by counter offset (config 5's rule): 8 shards == one call
    d_sh = torch.empty_like(d_in)
    per = N // 8
    for g in range(8):
        pkg.check(pkg.lib().bc_aes_ctr_dev(rk.ctypes.data, nr, nc, 128, g * per // 16, d_in.data_ptr() + g * per,
                                           d_sh.data_ptr() + g * per, per, None, G.stream()))
    torch.cuda.synchronize()
    assert torch.equal(d_sh, d_out)
    # round trip in place
    G.ctr_dev(pkg, rk, nr, nc, d_sh, d_sh, N)
    torch.cuda.synchronize()
    assert torch.equal(d_sh, d_in)


def test_config3_ecb256_cbc256_4gib(pkg, orc, big):
    d_in, d_out = big
    key = bytes(range(32))
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    G.ecb_dev(pkg, rke, nr, ENC, d_in, d_out, N)
    for off in windows(N, WIN):
        assert np.array_equal(G.host(d_out[off:off + WIN]), orc.aes_ecb(key, ENC, G.host(d_in[off:off + WIN]))), off
    d_back = torch.empty_like(d_in)
    G.ecb_dev(pkg, rkd, nr, DEC, d_out, d_back, N)
    torch.cuda.synchronize()
    assert torch.equal(d_back, d_in)                                  # D(E(x)) == x over 4 GiB
    # CBC decrypt of 4 GiB (treat d_in as ciphertext), random IV
    iv = bytes(range(100, 116))
    G.cbc_dec_dev(pkg, rkd, nr, iv, d_in, d_out, N)
    for off in windows(N, WIN):
        prev = iv if off == 0 else G.host(d_in[off - 16:off]).tobytes()
        want = orc.aes_cbc(key, DEC, prev, G.host(d_in[off:off + WIN]))[1]
        assert np.array_equal(G.host(d_out[off:off + WIN]), want), off
    d_back.copy_(d_in)
    G.cbc_dec_dev(pkg, rkd, nr, iv, d_back, d_back, N)                # in place (boundary scratch)
    torch.cuda.synchronize()
    assert torch.equal(d_back, d_out)
    del d_back


def test_config3_cbc_encrypt_65536x64k(pkg, orc, big):
    d_in, d_out = big
    n_msgs, msg_len = 65536, 65536
    key = bytes(range(32))
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    g = torch.Generator(device="cuda").manual_seed(7)
    d_iv = torch.randint(0, 256, (n_msgs * 16,), dtype=torch.uint8, device="cuda", generator=g)
    d_ivo = torch.empty_like(d_iv)
    G.cbc_enc_batch_dev(pkg, rke, nr, n_msgs, msg_len, d_iv, d_ivo, d_in, d_out)
    for m in (0, 1, 31, 32, 4097, n_msgs // 2, n_msgs - 1):
        want, ivo = orc.aes_cbc_encrypt_batch(key, G.host(d_iv[16 * m:16 * m + 16]).copy(),
                                              G.host(d_in[m * msg_len:(m + 1) * msg_len]), msg_len)
        assert np.array_equal(G.host(d_out[m * msg_len:(m + 1) * msg_len]), want), m
        assert np.array_equal(G.host(d_ivo[16 * m:16 * m + 16]), ivo)
    # whole-buffer property: CBC-decrypting all 4 GiB as ONE chain gives back every plaintext
    # block except the first of each message, which is off by (previous message's last
    # ciphertext block ^ its own iv); undo that and compare everything
    d_back = torch.empty_like(d_in)
    iv0 = G.host(d_iv[:16]).tobytes()
    G.cbc_dec_dev(pkg, rkd, nr, iv0, d_out, d_back, N)
    torch.cuda.synchronize()
    back = d_back.view(n_msgs, msg_len)
    prev_last = d_out.view(n_msgs, msg_len)[:-1, -16:]
    fix = torch.bitwise_xor(prev_last, d_iv.view(n_msgs, 16)[1:])
    back[1:, :16] = torch.bitwise_xor(back[1:, :16], fix)
    assert torch.equal(d_back, d_in)
    assert torch.equal(d_ivo.view(n_msgs, 16), d_out.view(n_msgs, msg_len)[:, -16:])   # iv <- last ciphertext block
    del d_back


def test_config4_rc4_1m_streams(pkg, orc, big):
    d_in, d_out = big
    n, length = 1 << 20, 4096
    g = torch.Generator(device="cuda").manual_seed(99)
    d_keys = torch.randint(0, 256, (n * 16,), dtype=torch.uint8, device="cuda", generator=g)
    G.arc4_dev(pkg, n, d_keys, 16, length, d_in, d_out)
    for s in (0, 1, 31, 32, 255, 256, 65537, n // 2, n - 2, n - 1):
        key = G.host(d_keys[16 * s:16 * s + 16]).tobytes()
        assert np.array_equal(G.host(d_out[s * length:(s + 1) * length]),
                              orc.arc4(key, G.host(d_in[s * length:(s + 1) * length]))), s
    d_back = torch.empty_like(d_in)
    G.arc4_dev(pkg, n, d_keys, 16, length, d_out, d_back)             # same keystreams again: identity
    torch.cuda.synchronize()
    assert torch.equal(d_back, d_in)
    del d_back


def _reference_ctr_keystream(ref, key, nr, nc, first_block, nbytes):
    """Keystream for blocks first_block.. of the counter stream `nc`, computed by the reference's own
    AES-NI code on all cores: the host builds the counter blocks (128-bit big-endian add) and runs
    AES_ECB_encrypt over them (AES_CTR_encrypt itself cannot start at an arbitrary counter)."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    blocks = nbytes // 16
    base = int.from_bytes(nc, "big") + first_block
    hi, lo0 = (base >> 64) & ((1 << 64) - 1), base & ((1 << 64) - 1)
    lo = np.arange(blocks, dtype=np.uint64) + np.uint64(lo0)          # wraps mod 2^64 like the counter
    carry = lo < np.uint64(lo0)
    ctr = np.empty((blocks, 16), np.uint8)
    ctr[:, 8:] = lo.astype(">u8").view(np.uint8).reshape(blocks, 8)
    ctr[:, :8] = np.frombuffer(hi.to_bytes(8, "big"), np.uint8)
    if carry.any():
        ctr[carry, :8] = np.frombuffer(((hi + 1) % (1 << 64)).to_bytes(8, "big"), np.uint8)
    sched = ref.packed_schedule(key, ENC)
    flat = ctr.reshape(-1)
    T = max(1, min(16, len(os.sched_getaffinity(0))))
    per = (blocks // T) * 16
    ks = np.empty(nbytes, np.uint8)

    def work(t):
        a, b = t * per, nbytes if t == T - 1 else (t + 1) * per
        ks[a:b] = ref.aesni_ecb(sched, nr, ENC, flat[a:b])

    with ThreadPoolExecutor(T) as ex:
        list(ex.map(work, range(T)))
    return ks


def test_config2_ctr128_4gib_every_byte_vs_reference_aesni(pkg, ref, big):
    """BASELINE config 2 in full: every byte of the 4 GiB AES-128-CTR output equals plaintext XOR the
    reference's keystream (the 2^32 counter carry is crossed at block 16), checked in 1 GiB windows."""
    if not ref.has_aesni():
        pytest.skip("host CPU without AES-NI")
    d_in, d_out = big
    key = bytes.fromhex("f9dfdf0c3cf14e4b9c296478904f979b")
    nc = bytes.fromhex("0f1e2d3c4b5a69788796a5b4fffffff0")
    rk, nr = G.sched(pkg, key, ENC)
    G.ctr_dev(pkg, rk, nr, nc, d_in, d_out, N)
    for w in range(N // GIB):
        ks = _reference_ctr_keystream(ref, key, nr, nc, w * (GIB // 16), GIB)
        np.bitwise_xor(ks, G.host(d_in[w * GIB:(w + 1) * GIB]), out=ks)
        assert np.array_equal(G.host(d_out[w * GIB:(w + 1) * GIB]), ks), w
        del ks


def test_ctr_low64_wrap_every_byte_vs_reference_aesni(pkg, ref):
    """256 MiB across the 2^64 carry of the counter (bytes 8..15 all ones): 128-bit rule vs the reference."""
    if not ref.has_aesni():
        pytest.skip("host CPU without AES-NI")
    n = 256 << 20
    key = bytes(range(16))
    nc = bytes.fromhex("00112233445566ff" + "ffffffffffff0000")           # wraps after 65536 blocks, carries into byte 7
    rk, nr = G.sched(pkg, key, ENC)
    g = torch.Generator(device="cuda").manual_seed(4243)
    d_in = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda", generator=g)
    d_out = torch.empty_like(d_in)
    G.ctr_dev(pkg, rk, nr, nc, d_in, d_out, n)
    ks = _reference_ctr_keystream(ref, key, nr, nc, 0, n)
    np.bitwise_xor(ks, G.host(d_in), out=ks)
    assert np.array_equal(G.host(d_out), ks)


def _threads():
    import os
    return max(1, min(16, len(os.sched_getaffinity(0))))


def _parallel(n_items, fn):
    """fn(a, b) over [0, n_items) split into one contiguous range per host thread."""
    from concurrent.futures import ThreadPoolExecutor
    T = _threads()
    per = -(-n_items // T)
    with ThreadPoolExecutor(T) as ex:
        list(ex.map(lambda t: fn(t * per, min(n_items, (t + 1) * per)), range(T)))


def test_config3_every_byte_vs_reference(pkg, ref, big):
    """BASELINE config 3 in full against the reference objects on all host cores, 1 GiB at a time:
    AES-256 ECB encrypt and CBC decrypt of 4 GiB (aesni.c rounds), CBC encrypt of 65 536 x 64 KiB (aes.c)."""
    if not ref.has_aesni():
        pytest.skip("host CPU without AES-NI")
    d_in, d_out = big
    key = bytes(range(32))
    rke, nr = G.sched(pkg, key, ENC)
    rkd, _ = G.sched(pkg, key, DEC)
    se, sd = ref.packed_schedule(key, ENC), ref.packed_schedule(key, DEC)
    iv = bytes(range(100, 116))
    n_msgs, msg_len = 65536, 65536
    g = torch.Generator(device="cuda").manual_seed(7)
    d_iv = torch.randint(0, 256, (n_msgs * 16,), dtype=torch.uint8, device="cuda", generator=g)
    ivs = G.host(d_iv)
    d_ecb, d_cbcd, d_cbce = torch.empty_like(d_in), torch.empty_like(d_in), d_out
    G.ecb_dev(pkg, rke, nr, ENC, d_in, d_ecb, N)
    G.cbc_dec_dev(pkg, rkd, nr, iv, d_in, d_cbcd, N)
    G.cbc_enc_batch_dev(pkg, rke, nr, n_msgs, msg_len, d_iv, None, d_in, d_cbce)
    for w in range(N // GIB):
        lo = w * GIB
        data = G.host(d_in[lo:lo + GIB])
        want = np.empty(GIB, np.uint8)
        # ECB-256 encrypt
        _parallel(GIB // 16, lambda a, b: want.__setitem__(slice(16 * a, 16 * b), ref.aesni_ecb(se, nr, ENC, data[16 * a:16 * b])))
        assert np.array_equal(G.host(d_ecb[lo:lo + GIB]), want), ("ecb", w)
        # CBC-256 decrypt: D(C_i) ^ C_{i-1}
        _parallel(GIB // 16, lambda a, b: want.__setitem__(slice(16 * a, 16 * b), ref.aesni_ecb(sd, nr, DEC, data[16 * a:16 * b])))
        prev = np.empty(GIB, np.uint8)
        prev[16:] = data[:-16]
        prev[:16] = np.frombuffer(iv, np.uint8) if w == 0 else G.host(d_in[lo - 16:lo])
        np.bitwise_xor(want, prev, out=want)
        assert np.array_equal(G.host(d_cbcd[lo:lo + GIB]), want), ("cbc-dec", w)
        # CBC-256 encrypt, 16384 messages of 64 KiB in this window, aes.c on every core
        m0, nm = lo // msg_len, GIB // msg_len

        def enc(a, b):
            o, _ = ref.aes_cbc_encrypt_batch(key, ivs[16 * (m0 + a):16 * (m0 + b)].copy(), data[a * msg_len:b * msg_len], msg_len)
            want[a * msg_len:b * msg_len] = o

        _parallel(nm, enc)
        assert np.array_equal(G.host(d_cbce[lo:lo + GIB]), want), ("cbc-enc", w)
        del data, want, prev
    del d_ecb, d_cbcd


def test_config4_rc4_every_byte_vs_reference(pkg, ref, big):
    """BASELINE config 4 in full: all 2^20 streams x 4 KiB against arc4.c (setup + prep + crypt per
    stream) on all host cores."""
    d_in, d_out = big
    n, length = 1 << 20, 4096
    g = torch.Generator(device="cuda").manual_seed(99)
    d_keys = torch.randint(0, 256, (n * 16,), dtype=torch.uint8, device="cuda", generator=g)
    keys = G.host(d_keys)
    G.arc4_dev(pkg, n, d_keys, 16, length, d_in, d_out)
    per_win = GIB // length
    for w in range(N // GIB):
        data = G.host(d_in[w * GIB:(w + 1) * GIB])
        want = np.empty(GIB, np.uint8)
        s0 = w * per_win

        def run(a, b):
            want[a * length:b * length] = ref.arc4_batch(keys[16 * (s0 + a):16 * (s0 + b)], 16, data[a * length:b * length], length)

        _parallel(per_win, run)
        assert np.array_equal(G.host(d_out[w * GIB:(w + 1) * GIB]), want), w
        del data, want

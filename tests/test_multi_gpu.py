"""Multi-GPU sharding.  CPU part: the world_size-2 rule (gloo, no GPU) — each rank
encrypts its contiguous block range with the counter offset by its first block
 and the concatenation must equal the unsharded stream.  The compute
stand-in on CPU is the oracle (this is a test of the partition arithmetic).
GPU part: bc_aes_ctr_host_multi on however many GPUs the box has."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

KEY = bytes(range(16))
NC = bytes.fromhex("00112233445566778899aabbfffffffe")


def shard(total_bytes: int, world: int, rank: int):
    """Equal contiguous block ranges, remainder to the last rank (bc_aes_ctr_host_multi's rule)."""
    blocks = (total_bytes + 15) // 16
    per = blocks // world
    b0 = per * rank
    byte0 = b0 * 16
    byte1 = total_bytes if rank == world - 1 else (b0 + per) * 16
    return b0, byte0, max(byte1, byte0)


def _worker(rank, world, port, nbytes, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = oracle.Oracle()
    data = np.random.default_rng(5).integers(0, 256, nbytes, dtype=np.uint8)
    b0, lo, hi = shard(nbytes, world, rank)
    nc = ((int.from_bytes(NC, "big") + b0) % (1 << 128)).to_bytes(16, "big")
    mine = orc.aes_ctr(KEY, nc, data[lo:hi])[0]
    # no collective on the data path; gather only to verify
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([mine.size], dtype=torch.int64))
    bufs = [torch.zeros(int(s), dtype=torch.uint8) for s in sizes]
    pad = torch.from_numpy(mine.copy())
    if world == 2:     # gloo all_gather needs equal sizes: send/recv instead
        if rank == 1:
            dist.send(pad, 0)
        else:
            other = torch.zeros(int(sizes[1]), dtype=torch.uint8)
            dist.recv(other, 1)
            whole = np.concatenate([mine, other.numpy()])
            want = orc.aes_ctr(KEY, NC, data)[0]
            q.put(bool(np.array_equal(whole, want)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("nbytes", [16 * 1001, 16 * 64 + 5])
def test_sharding_rule_world2_gloo(nbytes):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, nbytes, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def test_shard_arithmetic_covers_everything():
    for total in (0, 1, 15, 16, 17, 16 * 7, 16 * 8 + 3, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            pos = 0
            for r in range(world):
                b0, lo, hi = shard(total, world, r)
                assert lo == pos and lo == b0 * 16 and hi >= lo
                pos = hi
            assert pos == total


@pytest.mark.gpu
def test_ctr_host_multi_gpu(pkg, orc):
    n_gpus = pkg.lib().bc_device_count()
    rc, ctx = pkg.aes_setkey_enc(KEY)
    rk = pkg.schedule_words(ctx)
    data = np.random.default_rng(8).integers(0, 256, 48 * 1024 * 1024 + 16 * 3 + 7, dtype=np.uint8)
    want = orc.aes_ctr(KEY, NC, data)[0]
    for g in sorted({1, min(2, n_gpus), n_gpus}):
        got = pkg.aes_ctr_host_multi(g, rk, ctx.nr, NC, data)
        assert np.array_equal(got, want), g
    if n_gpus < 8:
        with pytest.raises(pkg.BulkCipherError):
            pkg.aes_ctr_host_multi(n_gpus + 1, rk, ctx.nr, NC, data[:4096])


# ---- the other shard rules (ECB / CBC-decrypt by block range, RC4 / CBC-encrypt by item index) ----

def shard_items(total: int, world: int, rank: int):
    """Equal contiguous item ranges, remainder to the last rank (shard_range in host/bc_engine.c)."""
    per = total // world
    first = per * rank
    return first, (total - first if rank == world - 1 else per)


def _worker_modes(rank, world, port, q):
    """World-size-2 restatement of bc_aes_cbc_decrypt_host_multi and bc_arc4_batch_host_multi with
    the oracle as the compute stand-in: rank r handles its range alone (CBC-decrypt shard 1 chains
    from the last ciphertext block of shard 0, which is INPUT data), nothing is exchanged on the
    data path, and the concatenation must equal the unsharded call."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = oracle.Oracle()
    rng = np.random.default_rng(77)
    key = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
    iv = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
    cipher = rng.integers(0, 256, 16 * 301, dtype=np.uint8)
    b0, nb = shard_items(cipher.size // 16, world, rank)
    my_iv = iv if b0 == 0 else cipher[b0 * 16 - 16:b0 * 16].tobytes()
    cbc_mine = orc.aes_cbc(key, 0, my_iv, cipher[b0 * 16:(b0 + nb) * 16])[1]
    n_streams, length = 37, 160
    keys = rng.integers(0, 256, n_streams * 16, dtype=np.uint8)
    data = rng.integers(0, 256, n_streams * length, dtype=np.uint8)
    s0, ns = shard_items(n_streams, world, rank)
    rc4_mine = orc.arc4_batch(keys[s0 * 16:(s0 + ns) * 16], 16, data[s0 * length:(s0 + ns) * length], length)
    mine = torch.from_numpy(np.concatenate([cbc_mine, rc4_mine]).copy())
    if rank == 1:
        dist.send(torch.tensor([cbc_mine.size, rc4_mine.size], dtype=torch.int64), 0)
        dist.send(mine, 0)
    else:
        sizes = torch.zeros(2, dtype=torch.int64)
        dist.recv(sizes, 1)
        other = torch.zeros(int(sizes.sum()), dtype=torch.uint8)
        dist.recv(other, 1)
        other = other.numpy()
        cbc_whole = np.concatenate([cbc_mine, other[:int(sizes[0])]])
        rc4_whole = np.concatenate([rc4_mine, other[int(sizes[0]):]])
        ok = np.array_equal(cbc_whole, orc.aes_cbc(key, 0, iv, cipher)[1]) and \
            np.array_equal(rc4_whole, orc.arc4_batch(keys, 16, data, length))
        q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


def test_cbc_seam_and_rc4_stream_split_world2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_modes, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def test_item_shard_arithmetic():
    for total in (0, 1, 7, 8, 9, 1 << 20):
        for world in (1, 2, 3, 8):
            pos = 0
            for r in range(world):
                first, cnt = shard_items(total, world, r)
                assert first == pos
                pos += cnt
            assert pos == total


def _gpu_counts(pkg):
    n = pkg.lib().bc_device_count()
    return n, sorted({1, min(2, n), n})


@pytest.mark.gpu
def test_ecb_host_multi_gpu(pkg, orc):
    n_gpus, counts = _gpu_counts(pkg)
    key = bytes(range(32))
    data = np.random.default_rng(9).integers(0, 256, 40 * 1024 * 1024 + 16 * 5, dtype=np.uint8)
    for mode in (1, 0):
        rc, ctx = (pkg.aes_setkey_enc if mode else pkg.aes_setkey_dec)(key)
        rk = pkg.schedule_words(ctx)
        want = orc.aes_ecb(key, mode, data)
        for g in counts:
            assert np.array_equal(pkg.aes_ecb_host_multi(g, rk, ctx.nr, mode, data), want), (mode, g)


@pytest.mark.gpu
@pytest.mark.parametrize("cfb", [False, True])
def test_chained_decrypt_host_multi_gpu(pkg, orc, cfb):
    """Every device seam takes its chaining block from the input, also when in == out."""
    n_gpus, counts = _gpu_counts(pkg)
    key = bytes(range(7, 31))                                     # AES-192
    iv = bytes(range(100, 116))
    data = np.random.default_rng(10).integers(0, 256, 24 * 1024 * 1024 + 16 * 9, dtype=np.uint8)
    rc, ctx = (pkg.aes_setkey_enc if cfb else pkg.aes_setkey_dec)(key)
    rk = pkg.schedule_words(ctx)
    want = orc.aes_cfb128(key, 0, iv, data)[0] if cfb else orc.aes_cbc(key, 0, iv, data)[1]
    for g in counts:
        for inplace in (False, True):
            got = pkg.aes_chained_decrypt_host_multi(g, rk, ctx.nr, iv, data, inplace=inplace, cfb=cfb)
            assert np.array_equal(got, want), (g, inplace)
    # a buffer with fewer blocks than GPUs: the empty shards are skipped
    small = data[:16 * 3]
    want = orc.aes_cfb128(key, 0, iv, small)[0] if cfb else orc.aes_cbc(key, 0, iv, small)[1]
    assert np.array_equal(pkg.aes_chained_decrypt_host_multi(n_gpus, rk, ctx.nr, iv, small, cfb=cfb), want)


@pytest.mark.gpu
def test_cbc_encrypt_batch_host_multi_gpu(pkg, orc):
    n_gpus, counts = _gpu_counts(pkg)
    key = bytes(range(16))
    rng = np.random.default_rng(11)
    n_msgs, msg_len = 203, 16 * 50
    data = rng.integers(0, 256, n_msgs * msg_len, dtype=np.uint8)
    ivs = rng.integers(0, 256, n_msgs * 16, dtype=np.uint8)
    rc, ctx = pkg.aes_setkey_enc(key)
    rk = pkg.schedule_words(ctx)
    want, want_iv = orc.aes_cbc_encrypt_batch(key, ivs, data, msg_len)
    for g in counts:
        got, got_iv = pkg.aes_chained_encrypt_batch_host_multi(g, rk, ctx.nr, ivs, data, msg_len)
        assert np.array_equal(got, want) and np.array_equal(got_iv, want_iv), g


@pytest.mark.gpu
def test_arc4_batch_host_multi_gpu(pkg, orc):
    n_gpus, counts = _gpu_counts(pkg)
    rng = np.random.default_rng(12)
    n_streams, length = 1001, 4096
    keys = rng.integers(0, 256, n_streams * 16, dtype=np.uint8)
    data = rng.integers(0, 256, n_streams * length, dtype=np.uint8)
    want = orc.arc4_batch(keys, 16, data, length)
    for g in counts:
        got, ctx = pkg.arc4_batch_host_multi(g, keys, 16, data, length, want_ctx=True)
        assert np.array_equal(got, want), g
        # contexts of the first and last stream: x = length mod 256, permutation a permutation
        for s in (0, n_streams - 1):
            c = ctx[s * 264:(s + 1) * 264]
            assert int.from_bytes(c[:4].tobytes(), "little") == length % 256
            assert sorted(c[8:].tolist()) == list(range(256))
    with pytest.raises(pkg.BulkCipherError):
        pkg.arc4_batch_host_multi(n_gpus + 1 if n_gpus < 16 else 17, keys, 16, data, length)


@pytest.mark.gpu
def test_fill_splitmix64_replay(pkg):
    """bc_fill_splitmix64_dev: the host replay bench.py's verification relies on, incl. a ragged tail."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    n = (1 << 20) + 16 * 3 + 11
    d = torch.zeros(n + 32, dtype=torch.uint8, device="cuda")
    first_word = (1 << 33) + 12345
    pkg.check(pkg.lib().bc_fill_splitmix64_dev(d.data_ptr(), n, bench.SEED, first_word, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    got = d.cpu().numpy()
    want = bench.splitmix_bytes(first_word, (n + 7) // 8)[:n]
    assert np.array_equal(got[:n], want) and not got[n:].any()

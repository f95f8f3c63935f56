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

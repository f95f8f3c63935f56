#!/usr/bin/env python
"""Kernel-only throughput of every device entry point (CUDA events on torch's
current stream, which is the stream the kernels are launched on).  Development
aid; the contract benchmark is bench.py."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=float, default=4.0)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    pkg = g.load_package()
    L = pkg.lib()
    n = int(args.gib * (1 << 30)) // 65536 * 65536
    d_in = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
    d_out = torch.empty_like(d_in)
    st = torch.cuda.current_stream().cuda_stream
    key = bytes(range(32))
    nc = bytes(range(12)) + bytes.fromhex("fffffff0")

    def rk(kb, mode=1):
        rc, ctx = (pkg.aes_setkey_enc if mode else pkg.aes_setkey_dec)(key[: kb // 8])
        return pkg.schedule_words(ctx), ctx.nr

    runs = []
    for kb in (128, 256):
        r, nr = rk(kb)
        runs.append((f"ctr{kb}", n, lambda r=r, nr=nr: pkg.check(L.bc_aes_ctr_dev(r.ctypes.data, nr, nc, 128, 0, d_in.data_ptr(), d_out.data_ptr(), n, None, st))))
        runs.append((f"ecb{kb}_enc", n, lambda r=r, nr=nr: pkg.check(L.bc_aes_ecb_dev(r.ctypes.data, nr, 1, d_in.data_ptr(), d_out.data_ptr(), n, st))))
        rd, _ = rk(kb, 0)
        runs.append((f"ecb{kb}_dec", n, lambda r=rd, nr=nr: pkg.check(L.bc_aes_ecb_dev(r.ctypes.data, nr, 0, d_in.data_ptr(), d_out.data_ptr(), n, st))))
        runs.append((f"cbc{kb}_dec", n, lambda r=rd, nr=nr: pkg.check(L.bc_aes_cbc_decrypt_dev(r.ctypes.data, nr, nc, d_in.data_ptr(), d_out.data_ptr(), n, st))))
        runs.append((f"cbc{kb}_dec_inplace", n, lambda r=rd, nr=nr: pkg.check(L.bc_aes_cbc_decrypt_dev(r.ctypes.data, nr, nc, d_out.data_ptr(), d_out.data_ptr(), n, st))))
    r, nr = rk(256)
    n_msgs = n // 65536
    d_iv = torch.randint(0, 256, (n_msgs * 16,), dtype=torch.uint8, device="cuda")
    runs.append(("cbc256_enc_batch_64k", n, lambda: pkg.check(L.bc_aes_cbc_encrypt_batch_dev(r.ctypes.data, nr, n_msgs, 65536, d_iv.data_ptr(), None, d_in.data_ptr(), d_out.data_ptr(), st))))
    n_streams = n // 4096
    d_keys = torch.randint(0, 256, (n_streams * 16,), dtype=torch.uint8, device="cuda")
    runs.append(("rc4_batch_4k", n, lambda: pkg.check(L.bc_arc4_batch_dev(n_streams, d_keys.data_ptr(), 16, 4096, d_in.data_ptr(), d_out.data_ptr(), None, None, st))))

    runs.append(("rc4_batch_4k_inplace", n, lambda: pkg.check(L.bc_arc4_batch_dev(n_streams, d_keys.data_ptr(), 16, 4096, d_out.data_ptr(), d_out.data_ptr(), None, None, st))))

    for name, nbytes, fn in runs:
        if args.only and args.only not in name:
            continue
        best, med = timeit(fn)
        print(f"{name:24s} {nbytes / best / 1e6:9.1f} GB/s best  {nbytes / med / 1e6:9.1f} GB/s median  ({best:.3f} ms)", flush=True)


if __name__ == "__main__":
    main()

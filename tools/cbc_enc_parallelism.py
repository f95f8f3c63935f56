#!/usr/bin/env python
"""Batched CBC-encrypt (AES-256) over 4 GiB cut into more and more independent messages: the chain
count is the only parallelism the mode has (aes.c:801-812), so throughput against messages-per-SM
shows where the kernel stops being latency-bound.  Prints a markdown table."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402


def main():
    pkg = g.load_package()
    L = pkg.lib()
    n = 4 << 30
    d_in = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
    d_out = torch.empty_like(d_in)
    rc, ctx = pkg.aes_setkey_enc(bytes(range(32)))
    rk = pkg.schedule_words(ctx)
    st = torch.cuda.current_stream().cuda_stream
    print("| messages x length | chains per SM | warps per SM | GB/s | of ECB-256 (630.6) |\n|---|---|---|---|---|")
    for msg_len in (1 << 20, 1 << 18, 1 << 16, 1 << 15, 1 << 14, 1 << 13, 1 << 12):
        n_msgs = n // msg_len
        d_iv = torch.randint(0, 256, (n_msgs * 16,), dtype=torch.uint8, device="cuda")
        fn = lambda: pkg.check(L.bc_aes_cbc_encrypt_batch_dev(rk.ctypes.data, ctx.nr, n_msgs, msg_len, d_iv.data_ptr(), None,  # noqa: E731
                                                              d_in.data_ptr(), d_out.data_ptr(), st))
        fn(); fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        gbs = n / best / 1e6
        print(f"| {n_msgs} x {msg_len >> 10} KiB | {n_msgs / 148:.0f} | {min(16, -(-n_msgs // 32 // 148))} | {gbs:.1f} | {gbs / 630.6:.2f} |")


if __name__ == "__main__":
    main()

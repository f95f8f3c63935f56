#!/usr/bin/env python
"""AES-128-CTR on small device-resident buffers: where the per-call floor goes.
Per size: one launch bracketed by two CUDA events (what a caller waits for: host launch path +
kernel), and 50 launches back to back between two events (launch latency pipelined away: the
kernel's own duration).  `--empty` adds the same two numbers for a 16-byte XOR kernel through the
same ctypes path = the floor of the measuring harness itself."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402


def measure(fn, reps=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    single = 1e9
    for _ in range(20):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        single = min(single, a.elapsed_time(b))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return single * 1e3, a.elapsed_time(b) / reps * 1e3


def main():
    pkg = g.load_package()
    L = pkg.lib()
    rc, ctx = pkg.aes_setkey_enc(bytes(range(16)))
    rk = pkg.schedule_words(ctx)
    nc = bytes(16)
    st = torch.cuda.current_stream().cuda_stream
    nmax = 1 << 28
    d_in = torch.randint(0, 256, (nmax,), dtype=torch.uint8, device="cuda")
    d_out = torch.empty_like(d_in)
    s, bb = measure(lambda: pkg.check(L.bc_xor_dev(d_in.data_ptr(), d_in.data_ptr(), d_out.data_ptr(), 16, st)))
    print(f"harness floor (16-byte xor kernel): single {s:.1f} us, back-to-back {bb:.1f} us")
    print("| bytes | single launch us | GB/s | back-to-back us | GB/s |")
    print("|---|---|---|---|---|")
    for sh in (12, 16, 18, 20, 22, 24, 26, 28):
        n = 1 << sh
        s, bb = measure(lambda: pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, 10, nc, 128, 0, d_in.data_ptr(), d_out.data_ptr(), n, None, st)))
        print(f"| {n} | {s:.1f} | {n / s / 1e3:.1f} | {bb:.1f} | {n / bb / 1e3:.1f} |")


if __name__ == "__main__":
    main()

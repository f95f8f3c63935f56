#!/usr/bin/env python
"""AES-128-CTR throughput vs buffer size: device-resident (CUDA events) and host-pinned
(bc_aes_ctr_host, wall clock).  Prints a markdown table."""
import ctypes
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402


def main():
    pkg = g.load_package()
    L = pkg.lib()
    rc, ctx = pkg.aes_setkey_enc(bytes(range(16)))
    rk = pkg.schedule_words(ctx)
    nc = bytes(16)
    st = torch.cuda.current_stream().cuda_stream
    nmax = 1 << 32
    d_in = torch.randint(0, 256, (nmax,), dtype=torch.uint8, device="cuda")
    d_out = torch.empty_like(d_in)
    hmax = 1 << 30
    h_in, h_out = L.bc_host_alloc(hmax), L.bc_host_alloc(hmax)
    np.ctypeslib.as_array((ctypes.c_ubyte * hmax).from_address(h_in))[:] = 3
    print("| bytes | device GB/s | device us | host-pinned GB/s | host us |")
    print("|---|---|---|---|---|")
    for sh in range(12, 33, 2):
        n = 1 << sh
        def dev():
            pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, 10, nc, 128, 0, d_in.data_ptr(), d_out.data_ptr(), n, None, st))
        for _ in range(3):
            dev()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(10 if sh < 28 else 4):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); dev(); b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        hb = ""
        if n <= hmax:
            ts = []
            for _ in range(5 if sh < 28 else 3):
                t0 = time.perf_counter()
                pkg.check(L.bc_aes_ctr_host(rk.ctypes.data, 10, nc, 128, 0, h_in, h_out, n, None))
                ts.append(time.perf_counter() - t0)
            hbest = min(ts[1:])
            hb = f"{n / hbest / 1e9:.2f} | {hbest * 1e6:.0f}"
        else:
            hb = "— | —"
        print(f"| {n} | {n / best / 1e6:.1f} | {best * 1e3:.1f} | {hb} |")


if __name__ == "__main__":
    main()

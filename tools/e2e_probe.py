#!/usr/bin/env python
"""Host-buffer (pinned) AES-128-CTR through bc_aes_ctr_host: GB/s for the current
BULKCIPHER_CHUNK_MIB / BULKCIPHER_SLOTS setting.  Development aid."""
import ctypes
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402


def main():
    pkg = g.load_package()
    L = pkg.lib()
    n = int(float(sys.argv[1]) * (1 << 30)) if len(sys.argv) > 1 else 2 << 30
    rc, ctx = pkg.aes_setkey_enc(bytes(range(16)))
    rk = pkg.schedule_words(ctx)
    pageable = len(sys.argv) > 2 and sys.argv[2] == "pageable"
    if pageable:
        a_in, a_out = np.full(n, 7, np.uint8), np.zeros(n, np.uint8)
        h_in, h_out = a_in.ctypes.data, a_out.ctypes.data
    else:
        h_in, h_out = L.bc_host_alloc(n), L.bc_host_alloc(n)
        np.ctypeslib.as_array((ctypes.c_ubyte * n).from_address(h_in))[:] = 7
    nc = bytes(16)
    ts = []
    zerocopy = len(sys.argv) > 2 and sys.argv[2] == "zerocopy"
    for _ in range(4):
        t0 = time.perf_counter()
        if zerocopy:      # the kernel reads / writes the pinned host buffers itself over PCIe (UVA), no staging
            pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, 10, nc, 128, 0, h_in, h_out, n, None, None))
            pkg.check(L.bc_device_sync())
        else:
            pkg.check(L.bc_aes_ctr_host(rk.ctypes.data, 10, nc, 128, 0, h_in, h_out, n, None))
        ts.append(time.perf_counter() - t0)
    print(("zerocopy " if zerocopy else "") + ("pageable " if pageable else "pinned ") + f"copy_threads={os.environ.get('BULKCIPHER_COPY_THREADS', 'auto')} " + f"chunk={os.environ.get('BULKCIPHER_CHUNK_MIB', '32')}MiB slots={os.environ.get('BULKCIPHER_SLOTS', '3')} "
          f"best {n / min(ts[1:]) / 1e9:.2f} GB/s  median {n / float(np.median(ts[1:])) / 1e9:.2f} GB/s")


if __name__ == "__main__":
    main()

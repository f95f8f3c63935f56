#!/usr/bin/env python
"""Soak: hammer every device kernel and host pipeline with random sizes for a while and require
bit-identical results from call to call and against the oracle on sampled windows.  Meant to
catch the intermittent kind of bug (stream ordering, TMA phase bookkeeping) that a single pass
of the parity tests can miss.  usage: soak.py [seconds]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402
import oracle  # noqa: E402


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    pkg = g.load_package()
    L = pkg.lib()
    orc = oracle.Oracle()
    rng = np.random.default_rng(int(time.time()))
    st = torch.cuda.current_stream().cuda_stream
    t0 = time.time()
    rounds = checks = 0
    while time.time() - t0 < budget:
        kb = int(rng.choice([128, 192, 256]))
        key = rng.integers(0, 256, kb // 8, dtype=np.uint8).tobytes()
        rc, e = pkg.aes_setkey_enc(key)
        rc, d = pkg.aes_setkey_dec(key)
        rke, rkd, nr = pkg.schedule_words(e), pkg.schedule_words(d), e.nr
        n = int(rng.integers(1, 1 << 22)) * 16
        data = rng.integers(0, 256, n, dtype=np.uint8)
        iv = rng.integers(0, 256, 16, dtype=np.uint8).tobytes()
        d_in = torch.from_numpy(data).cuda()
        outs = {}
        msg_len = int(rng.choice([16, 48, 64, 4096, 65536]))
        length = int(rng.choice([16, 64, 1024, 4096]))
        for rep in range(3):
            d_out = torch.empty_like(d_in)
            res = {}
            pkg.check(L.bc_aes_ctr_dev(rke.ctypes.data, nr, iv, 128, 9, d_in.data_ptr(), d_out.data_ptr(), n - 3, None, st))
            res["ctr"] = d_out[: n - 3].clone()
            pkg.check(L.bc_aes_ecb_dev(rkd.ctypes.data, nr, 0, d_in.data_ptr(), d_out.data_ptr(), n, st))
            res["ecbd"] = d_out.clone()
            pkg.check(L.bc_aes_cbc_decrypt_dev(rkd.ctypes.data, nr, iv, d_in.data_ptr(), d_out.data_ptr(), n, st))
            res["cbcd"] = d_out.clone()
            tmp = d_in.clone()
            pkg.check(L.bc_aes_cbc_decrypt_dev(rkd.ctypes.data, nr, iv, tmp.data_ptr(), tmp.data_ptr(), n, st))
            res["cbcd_inplace"] = tmp
            nm = n // msg_len
            if nm:
                d_iv = torch.from_numpy(np.resize(np.frombuffer(iv, np.uint8), nm * 16).copy()).cuda()
                pkg.check(L.bc_aes_cbc_encrypt_batch_dev(rke.ctypes.data, nr, nm, msg_len, d_iv.data_ptr(), None,
                                                         d_in.data_ptr(), d_out.data_ptr(), st))
                res["cbce"] = (msg_len, d_out[: nm * msg_len].clone())
            ns = n // length
            if ns:
                d_keys = d_in[: ns * 16] if ns * 16 <= n else None
                if d_keys is not None:
                    pkg.check(L.bc_arc4_batch_dev(ns, d_keys.data_ptr(), 16, length, d_in.data_ptr(), d_out.data_ptr(), None, None, st))
                    res["rc4"] = (length, d_out[: ns * length].clone())
                    tmp = d_in.clone()
                    pkg.check(L.bc_arc4_batch_dev(ns, d_keys.data_ptr(), 16, length, tmp.data_ptr(), tmp.data_ptr(), None, None, st))
                    res["rc4_inplace"] = (length, tmp[: ns * length])
            host = pkg.aes_ctr_host(rke, nr, iv, data[: n - 3], first_block=9)
            res["ctr_host"] = torch.from_numpy(host).cuda()
            torch.cuda.synchronize()
            outs[rep] = res
            if rep:
                for k, v in res.items():
                    a = v[1] if isinstance(v, tuple) else v
                    b = outs[0].get(k)
                    b = b[1] if isinstance(b, tuple) else b
                    if b is not None and a.shape == b.shape:
                        assert torch.equal(a, b), f"{k} differs between repetitions (n={n}, kb={kb})"
                        checks += 1
        r0 = outs[0]
        assert torch.equal(r0["ctr"], r0["ctr_host"]) and torch.equal(r0["cbcd"], r0["cbcd_inplace"])
        if "rc4" in r0:
            assert torch.equal(r0["rc4"][1], r0["rc4_inplace"][1])
        w = min(n, 8192)
        off = (int(rng.integers(0, n - w + 1)) // 16) * 16
        shifted = ((int.from_bytes(iv, "big") + 9 + off // 16) % (1 << 128)).to_bytes(16, "big")
        wlen = min(w, n - 3 - off)
        assert np.array_equal(r0["ctr"][off:off + wlen].cpu().numpy(), orc.aes_ctr(key, shifted, data[off:off + wlen])[0])
        assert np.array_equal(r0["ecbd"][off:off + w].cpu().numpy(), orc.aes_ecb(key, 0, data[off:off + w]))
        checks += 4
        rounds += 1
    print(f"soak: {rounds} rounds, {checks} comparisons, {time.time() - t0:.0f} s: all identical")


if __name__ == "__main__":
    main()

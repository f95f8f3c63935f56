"""Helpers for the -m gpu tests: device buffers via torch (plumbing only) and
thin callers of the device-pointer C-ABI."""
import numpy as np
import torch


def to_dev(a):
    return torch.from_numpy(np.array(a, dtype=np.uint8, copy=True)).cuda()


def stream():
    return torch.cuda.current_stream().cuda_stream


def ctr_dev(pkg, rk, nr, counter, d_in, d_out, nbytes, bits=128, first_block=0, d_last=None):
    pkg.check(pkg.lib().bc_aes_ctr_dev(rk.ctypes.data, nr, bytes(counter), bits, first_block, d_in.data_ptr(),
                                       d_out.data_ptr(), nbytes, d_last.data_ptr() if d_last is not None else None,
                                       stream()))


def ecb_dev(pkg, rk, nr, mode, d_in, d_out, nbytes):
    pkg.check(pkg.lib().bc_aes_ecb_dev(rk.ctypes.data, nr, mode, d_in.data_ptr(), d_out.data_ptr(), nbytes, stream()))


def cbc_dec_dev(pkg, rk, nr, iv, d_in, d_out, nbytes):
    pkg.check(pkg.lib().bc_aes_cbc_decrypt_dev(rk.ctypes.data, nr, bytes(iv), d_in.data_ptr(), d_out.data_ptr(),
                                               nbytes, stream()))


def cbc_enc_batch_dev(pkg, rk, nr, n_msgs, msg_len, d_ivs, d_ivs_out, d_in, d_out):
    pkg.check(pkg.lib().bc_aes_cbc_encrypt_batch_dev(rk.ctypes.data, nr, n_msgs, msg_len, d_ivs.data_ptr(),
                                                     d_ivs_out.data_ptr() if d_ivs_out is not None else None,
                                                     d_in.data_ptr(), d_out.data_ptr(), stream()))


def arc4_dev(pkg, n, d_keys, keylen, length, d_in, d_out, d_ctx_in=None, d_ctx_out=None):
    pkg.check(pkg.lib().bc_arc4_batch_dev(n, d_keys.data_ptr() if d_keys is not None else None, keylen, length,
                                          d_in.data_ptr() if d_in is not None else None,
                                          d_out.data_ptr() if d_out is not None else None,
                                          d_ctx_in.data_ptr() if d_ctx_in is not None else None,
                                          d_ctx_out.data_ptr() if d_ctx_out is not None else None, stream()))


def sched(pkg, key, mode=1):
    rc, ctx = (pkg.aes_setkey_enc if mode == 1 else pkg.aes_setkey_dec)(key)
    assert rc == 0
    return pkg.schedule_words(ctx), ctx.nr


def host(t):
    torch.cuda.synchronize()
    return t.cpu().numpy()

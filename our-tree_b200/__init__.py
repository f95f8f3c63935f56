"""our_tree_b200 — Python face of libbulkcipher.so (ctypes, no torch types).

This package is test/bench glue over the C-ABI in include/*.h; the product is
the shared library (CUDA kernels for sm_100a + the C host layer + the
reference's drop-in symbols).  Function names and argument meaning mirror the
reference's C interface (aes-modes/aes.h, aes-modes/aesni.h, arc4.h, rc4.h) so
the parity tests read like the reference's own self-tests.

There is NO CPU fallback: importing works anywhere, but `lib()` raises if the
library has not been built, and every compute call fails with a BC_ERR_* code
when no CUDA device is usable.

The on-disk directory is `our-tree_b200/` (not importable by name because of
the hyphen); `__graft_entry__.load_package()` / tests/conftest.py import it as
module `our_tree_b200`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BULKCIPHER_LIB") or os.path.join(_HERE, "libbulkcipher.so")   # override: A/B builds only

AES_ENCRYPT, AES_DECRYPT = 1, 0
ERR_INVALID_KEY_LENGTH = -0x20
ERR_INVALID_INPUT_LENGTH = -0x22
BC_OK, BC_ERR_CUDA, BC_ERR_BAD_ARG, BC_ERR_NO_DEVICE, BC_ERR_OVERLAP, BC_ERR_TOO_LARGE = (
    0, -0x7001, -0x7002, -0x7003, -0x7004, -0x7005)

# every symbol include/*.h declares (tests check the library exports all of them)
DROPIN_SYMBOLS = [
    "aes_setkey_enc", "aes_setkey_dec", "aes_crypt_ecb", "aes_crypt_cbc", "aes_crypt_ctr", "aes_crypt_cfb128",
    "aes_self_test",
    "CheckAESSupport", "AES_256_Key_Expansion", "AES_128_Key_Expansion", "AES_192_Key_Expansion",
    "AES_ECB_encrypt", "AES_ECB_decrypt", "AES_CTR_encrypt",
    "arc4_setup", "arc4_prep", "arc4_crypt", "arc4_self_test", "rc4_init", "rc4_crypt",
    "KEY_256_ASSIST_1", "KEY_256_ASSIST_2",       # exported by the reference's aesni.o, not in its header (aesni.c:13-36)
]
BC_SYMBOLS = [
    "bc_last_error", "bc_device_count", "bc_set_device", "bc_get_device", "bc_sm_count", "bc_shutdown",
    "bc_host_alloc", "bc_host_alloc_wc", "bc_host_free", "bc_dev_alloc", "bc_dev_free", "bc_memcpy_h2d", "bc_memcpy_d2h",
    "bc_device_sync",
    "bc_aes_ctr_dev", "bc_aes_ecb_dev", "bc_aes_cbc_decrypt_dev", "bc_aes_cbc_encrypt_batch_dev",
    "bc_arc4_batch_dev", "bc_xor_dev", "bc_aes_cbc_decrypt_inplace_noalloc_dev",
    "bc_aes_cfb128_decrypt_dev", "bc_aes_cfb128_encrypt_batch_dev", "bc_aes_cfb128_decrypt_host",
    "bc_aes_cfb128_encrypt_batch_host", "bc_aes_ecb_be32_dev", "bc_aes_ecb_be32_host",
    "bc_aes_expand_key", "bc_aes_invert_key",
    "bc_aes_ctr_host", "bc_aes_ecb_host", "bc_aes_cbc_decrypt_host", "bc_aes_cbc_encrypt_batch_host",
    "bc_arc4_batch_host", "bc_xor_host", "bc_aes_ctr_host_multi",
    "bc_aes_ecb_host_multi", "bc_aes_cbc_decrypt_host_multi", "bc_aes_cfb128_decrypt_host_multi",
    "bc_aes_cbc_encrypt_batch_host_multi", "bc_aes_cfb128_encrypt_batch_host_multi", "bc_arc4_batch_host_multi",
    "bc_device_numa_node", "bc_bind_host_thread_to_device", "bc_fill_splitmix64_dev", "bc_void_status",
    "bc_register_pageable",
]


class BulkCipherError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libbulkcipher error {code:#x}: {msg}")
        self.code = code


class aes_context(C.Structure):           # aes-modes/aes.h:41-47
    _fields_ = [("nr", C.c_int), ("rk", C.POINTER(C.c_ulong)), ("buf", C.c_ulong * 68)]


class arc4_context(C.Structure):          # arc4.h:35-41
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("m", C.c_ubyte * 256)]


class rc4_state(C.Structure):             # rc4.h:43-47
    _fields_ = [("perm", C.c_ubyte * 256), ("index1", C.c_int), ("index2", C.c_int)]   # char perm[256] in C


_lib = None


def lib() -> C.CDLL:
    """Load libbulkcipher.so; fail loudly if it is not built (no fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C our-tree_b200`.  There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, sz, u64, i32, u32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int, C.c_uint
    L.bc_last_error.restype = C.c_char_p
    L.bc_host_alloc.restype = vp
    L.bc_host_alloc.argtypes = [sz]
    L.bc_host_free.argtypes = [vp]
    L.bc_host_alloc_wc.restype = vp
    L.bc_host_alloc_wc.argtypes = [sz]
    L.bc_dev_alloc.restype = vp
    L.bc_dev_alloc.argtypes = [sz]
    L.bc_dev_free.argtypes = [vp]
    L.bc_memcpy_h2d.argtypes = [vp, vp, sz]
    L.bc_memcpy_d2h.argtypes = [vp, vp, sz]
    L.bc_aes_ctr_dev.argtypes = [vp, i32, vp, i32, u64, vp, vp, sz, vp, vp]
    L.bc_aes_ecb_dev.argtypes = [vp, i32, i32, vp, vp, sz, vp]
    L.bc_aes_cbc_decrypt_dev.argtypes = [vp, i32, vp, vp, vp, sz, vp]
    L.bc_aes_cbc_decrypt_inplace_noalloc_dev.argtypes = [vp, i32, vp, vp, sz, vp]
    L.bc_aes_cbc_encrypt_batch_dev.argtypes = [vp, i32, sz, sz, vp, vp, vp, vp, vp]
    L.bc_aes_cfb128_decrypt_dev.argtypes = [vp, i32, vp, vp, vp, sz, vp]
    L.bc_aes_cfb128_encrypt_batch_dev.argtypes = [vp, i32, sz, sz, vp, vp, vp, vp, vp]
    L.bc_aes_cfb128_decrypt_host.argtypes = [vp, i32, vp, vp, vp, sz]
    L.bc_aes_cfb128_encrypt_batch_host.argtypes = [vp, i32, sz, sz, vp, vp, vp, vp]
    L.aes_crypt_cfb128.argtypes = [vp, i32, sz, C.POINTER(C.c_int), vp, vp, vp]
    L.bc_arc4_batch_dev.argtypes = [sz, vp, u32, sz, vp, vp, vp, vp, vp]
    L.bc_xor_dev.argtypes = [vp, vp, vp, sz, vp]
    L.bc_aes_ctr_host.argtypes = [vp, i32, vp, i32, u64, vp, vp, sz, vp]
    L.bc_aes_ecb_host.argtypes = [vp, i32, i32, vp, vp, sz]
    L.bc_aes_cbc_decrypt_host.argtypes = [vp, i32, vp, vp, vp, sz]
    L.bc_aes_cbc_encrypt_batch_host.argtypes = [vp, i32, sz, sz, vp, vp, vp, vp]
    L.bc_arc4_batch_host.argtypes = [sz, vp, u32, sz, vp, vp, vp, vp]
    L.bc_xor_host.argtypes = [vp, vp, vp, sz]
    L.bc_aes_ctr_host_multi.argtypes = [i32, vp, i32, vp, i32, vp, vp, sz]
    L.bc_aes_ecb_host_multi.argtypes = [i32, vp, i32, i32, vp, vp, sz]
    L.bc_aes_cbc_decrypt_host_multi.argtypes = [i32, vp, i32, vp, vp, vp, sz]
    L.bc_aes_cfb128_decrypt_host_multi.argtypes = [i32, vp, i32, vp, vp, vp, sz]
    L.bc_aes_cbc_encrypt_batch_host_multi.argtypes = [i32, vp, i32, sz, sz, vp, vp, vp, vp]
    L.bc_aes_cfb128_encrypt_batch_host_multi.argtypes = [i32, vp, i32, sz, sz, vp, vp, vp, vp]
    L.bc_arc4_batch_host_multi.argtypes = [i32, sz, vp, u32, sz, vp, vp, vp, vp]
    L.bc_device_numa_node.argtypes = [i32]
    L.bc_bind_host_thread_to_device.argtypes = [i32]
    L.bc_fill_splitmix64_dev.argtypes = [vp, sz, u64, u64, vp]
    # every pointer-taking symbol gets argtypes (64-bit pointers must not be truncated)
    L.bc_aes_ecb_be32_dev.argtypes = [vp, i32, i32, vp, vp, sz, vp]
    L.bc_aes_ecb_be32_host.argtypes = [vp, i32, i32, vp, vp, sz]
    L.bc_aes_expand_key.argtypes = [vp, u32, vp]
    L.bc_aes_invert_key.argtypes = [vp, i32, vp]
    L.bc_aes_invert_key.restype = None
    L.bc_set_device.argtypes = [i32]
    L.bc_internal_reg_stats.argtypes = [vp]
    L.bc_register_pageable.argtypes = [i32]
    L.bc_internal_ecb_ahead_stats.argtypes = [vp]
    L.bc_internal_ecb_ahead_stats.restype = None
    L.aes_setkey_enc.argtypes = [vp, vp, u32]
    L.aes_setkey_dec.argtypes = [vp, vp, u32]
    L.aes_crypt_ecb.argtypes = [vp, i32, vp, vp]
    L.aes_crypt_cbc.argtypes = [vp, i32, sz, vp, vp, vp]
    L.aes_crypt_ctr.argtypes = [vp, i32, C.POINTER(C.c_int), vp, vp, vp, vp]
    for f in ("AES_256_Key_Expansion", "AES_128_Key_Expansion", "AES_192_Key_Expansion"):
        getattr(L, f).argtypes = [vp, vp]
        getattr(L, f).restype = None
    L.AES_ECB_encrypt.argtypes = [vp, vp, C.c_ulong, vp, i32]
    L.AES_ECB_decrypt.argtypes = [vp, vp, C.c_ulong, vp, i32]
    L.AES_CTR_encrypt.argtypes = [vp, vp, vp, vp, C.c_ulong, vp, i32]
    for f in ("AES_ECB_encrypt", "AES_ECB_decrypt", "AES_CTR_encrypt", "arc4_setup", "rc4_init", "rc4_crypt"):
        getattr(L, f).restype = None
    L.arc4_setup.argtypes = [vp, vp, u32]
    L.arc4_prep.argtypes = [vp, sz, vp]
    L.arc4_crypt.argtypes = [sz, vp, vp, vp]
    L.rc4_init.argtypes = [vp, vp, i32]
    L.rc4_crypt.argtypes = [vp, vp, vp, i32]
    _lib = L
    return L


def last_error() -> str:
    return (lib().bc_last_error() or b"").decode(errors="replace")


def check(rc: int) -> None:
    if rc != 0:
        raise BulkCipherError(rc, last_error())


def _u8(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(bytes(a), dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def _ptr(a: np.ndarray) -> int:
    return a.ctypes.data


# --------------------------------------------------------------- aes.h face

def aes_setkey_enc(key: bytes, keysize: int | None = None):
    ctx = aes_context()
    rc = lib().aes_setkey_enc(C.byref(ctx), bytes(key), len(key) * 8 if keysize is None else keysize)
    return rc, ctx


def aes_setkey_dec(key: bytes, keysize: int | None = None):
    ctx = aes_context()
    rc = lib().aes_setkey_dec(C.byref(ctx), bytes(key), len(key) * 8 if keysize is None else keysize)
    return rc, ctx


def schedule_words(ctx: aes_context) -> np.ndarray:
    """The 4*(nr+1) round-key words = low halves of ctx.rk[] (what bc_* calls take as `rk`)."""
    return np.array([ctx.buf[i] & 0xFFFFFFFF for i in range(4 * (ctx.nr + 1))], dtype=np.uint32)


def aes_crypt_ecb(ctx: aes_context, mode: int, block) -> tuple[int, bytes]:
    i = _u8(block)
    o = np.zeros(16, np.uint8)
    rc = lib().aes_crypt_ecb(C.byref(ctx), mode, _ptr(i), _ptr(o))
    return rc, o.tobytes()


def aes_crypt_cbc(ctx: aes_context, mode: int, iv: bytes, data, inplace: bool = False):
    d = np.array(_u8(data))
    o = d if inplace else np.empty_like(d)
    ivb = np.array(_u8(iv))
    rc = lib().aes_crypt_cbc(C.byref(ctx), mode, d.size, _ptr(ivb), _ptr(d), _ptr(o))
    return rc, o, ivb.tobytes()


def aes_crypt_ctr(ctx: aes_context, nonce_counter: bytes, data, nc_off: int = 0,
                  stream_block: bytes = bytes(16), inplace: bool = False):
    d = np.array(_u8(data))
    o = d if inplace else np.empty_like(d)
    nc = np.array(_u8(nonce_counter))
    sb = np.array(_u8(stream_block))
    off = C.c_int(nc_off)
    rc = lib().aes_crypt_ctr(C.byref(ctx), d.size, C.byref(off), _ptr(nc), _ptr(sb), _ptr(d), _ptr(o))
    return rc, o, off.value, nc.tobytes(), sb.tobytes()


def aes_crypt_cfb128(ctx: aes_context, mode: int, iv: bytes, data, iv_off: int = 0, inplace: bool = False):
    d = np.array(_u8(data))
    o = d if inplace else np.empty_like(d)
    ivb = np.array(_u8(iv))
    off = C.c_int(iv_off)
    rc = lib().aes_crypt_cfb128(C.byref(ctx), mode, d.size, C.byref(off), _ptr(ivb), _ptr(d), _ptr(o))
    return rc, o, off.value, ivb.tobytes()


# ------------------------------------------------------------- aesni.h face

def AES_Key_Expansion(key: bytes) -> bytes:
    out = np.zeros(16 * (len(key) // 4 + 7), np.uint8)
    f = {16: "AES_128_Key_Expansion", 24: "AES_192_Key_Expansion", 32: "AES_256_Key_Expansion"}[len(key)]
    getattr(lib(), f)(bytes(key), _ptr(out))
    return out.tobytes()


def AES_ECB_encrypt(data, sched: bytes, rounds: int) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    lib().AES_ECB_encrypt(_ptr(d), _ptr(o), d.size, bytes(sched), rounds)
    return o


def AES_ECB_decrypt(data, sched: bytes, rounds: int) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    lib().AES_ECB_decrypt(_ptr(d), _ptr(o), d.size, bytes(sched), rounds)
    return o


def AES_CTR_encrypt(data, ivec: bytes, nonce: bytes, sched: bytes, rounds: int) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    lib().AES_CTR_encrypt(_ptr(d), _ptr(o), bytes(ivec), bytes(nonce), d.size, bytes(sched), rounds)
    return o


# ------------------------------------------------------- arc4.h / rc4.h face

def arc4_setup(key: bytes) -> arc4_context:
    ctx = arc4_context()
    lib().arc4_setup(C.byref(ctx), bytes(key), len(key))
    return ctx


def arc4_prep(ctx: arc4_context, length: int) -> np.ndarray:
    ks = np.zeros(length, np.uint8)
    check(lib().arc4_prep(C.byref(ctx), length, _ptr(ks)))
    return ks


def arc4_crypt(data, keystream) -> np.ndarray:
    d, k = _u8(data), _u8(keystream)
    o = np.empty_like(d)
    check(lib().arc4_crypt(d.size, _ptr(d), _ptr(k), _ptr(o)))
    return o


def rc4_init(key: bytes) -> rc4_state:
    st = rc4_state()
    lib().rc4_init(C.byref(st), bytes(key), len(key))
    return st


def rc4_crypt(st: rc4_state, data) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    lib().rc4_crypt(C.byref(st), _ptr(d), _ptr(o), d.size)
    return o


# ------------------------------------------------ bulkcipher.h, host buffers

def _rk(rk) -> np.ndarray:
    return np.ascontiguousarray(rk, dtype=np.uint32)


def aes_ctr_host(rk, nr: int, counter: bytes, data, counter_bits: int = 128, first_block: int = 0,
                 want_last: bool = False, out=None):
    d = _u8(data)
    o = np.empty_like(d) if out is None else out
    r = _rk(rk)
    last = np.zeros(16, np.uint8)
    check(lib().bc_aes_ctr_host(_ptr(r), nr, bytes(counter), counter_bits, first_block, _ptr(d), _ptr(o), d.size,
                                _ptr(last) if want_last else None))
    return (o, last.tobytes()) if want_last else o


def aes_ecb_host(rk, nr: int, mode: int, data) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    r = _rk(rk)
    check(lib().bc_aes_ecb_host(_ptr(r), nr, mode, _ptr(d), _ptr(o), d.size))
    return o


def aes_cbc_decrypt_host(rk, nr: int, iv: bytes, data, inplace: bool = False) -> np.ndarray:
    d = np.array(_u8(data))
    o = d if inplace else np.empty_like(d)
    r = _rk(rk)
    check(lib().bc_aes_cbc_decrypt_host(_ptr(r), nr, bytes(iv), _ptr(d), _ptr(o), d.size))
    return o


def aes_cbc_encrypt_batch_host(rk, nr: int, ivs, data, msg_len: int):
    d = _u8(data)
    o = np.empty_like(d)
    iv = _u8(ivs)
    ivo = np.empty_like(iv)
    r = _rk(rk)
    check(lib().bc_aes_cbc_encrypt_batch_host(_ptr(r), nr, d.size // msg_len if msg_len else 0, msg_len,
                                              _ptr(iv), _ptr(ivo), _ptr(d), _ptr(o)))
    return o, ivo


def arc4_batch_host(keys, keylen: int, data, length: int, want_ctx: bool = False):
    d, k = _u8(data), _u8(keys)
    n = k.size // keylen
    o = np.empty_like(d)
    ctx = np.zeros(n * 264, np.uint8) if want_ctx else None
    check(lib().bc_arc4_batch_host(n, _ptr(k), keylen, length, _ptr(d) if d.size else None,
                                   _ptr(o) if d.size else None, None, _ptr(ctx) if want_ctx else None))
    return (o, ctx) if want_ctx else o


def aes_ctr_host_multi(n_gpus: int, rk, nr: int, counter: bytes, data, counter_bits: int = 128) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    r = _rk(rk)
    check(lib().bc_aes_ctr_host_multi(n_gpus, _ptr(r), nr, bytes(counter), counter_bits, _ptr(d), _ptr(o), d.size))
    return o


def aes_ecb_host_multi(n_gpus: int, rk, nr: int, mode: int, data) -> np.ndarray:
    d = _u8(data)
    o = np.empty_like(d)
    r = _rk(rk)
    check(lib().bc_aes_ecb_host_multi(n_gpus, _ptr(r), nr, mode, _ptr(d), _ptr(o), d.size))
    return o


def aes_chained_decrypt_host_multi(n_gpus: int, rk, nr: int, iv: bytes, data, inplace: bool = False, cfb: bool = False):
    d = np.array(_u8(data))
    o = d if inplace else np.empty_like(d)
    r = _rk(rk)
    fn = lib().bc_aes_cfb128_decrypt_host_multi if cfb else lib().bc_aes_cbc_decrypt_host_multi
    check(fn(n_gpus, _ptr(r), nr, bytes(iv), _ptr(d), _ptr(o), d.size))
    return o


def aes_chained_encrypt_batch_host_multi(n_gpus: int, rk, nr: int, ivs, data, msg_len: int, cfb: bool = False):
    d = _u8(data)
    o = np.empty_like(d)
    iv = _u8(ivs)
    ivo = np.empty_like(iv)
    r = _rk(rk)
    fn = lib().bc_aes_cfb128_encrypt_batch_host_multi if cfb else lib().bc_aes_cbc_encrypt_batch_host_multi
    check(fn(n_gpus, _ptr(r), nr, d.size // msg_len if msg_len else 0, msg_len, _ptr(iv), _ptr(ivo), _ptr(d), _ptr(o)))
    return o, ivo


def arc4_batch_host_multi(n_gpus: int, keys, keylen: int, data, length: int, want_ctx: bool = False):
    d, k = _u8(data), _u8(keys)
    n = k.size // keylen
    o = np.empty_like(d)
    ctx = np.zeros(n * 264, np.uint8) if want_ctx else None
    check(lib().bc_arc4_batch_host_multi(n_gpus, n, _ptr(k), keylen, length, _ptr(d) if d.size else None,
                                         _ptr(o) if d.size else None, None, _ptr(ctx) if want_ctx else None))
    return (o, ctx) if want_ctx else o

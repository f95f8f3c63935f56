"""ctypes front-ends for the two CPU checkers.  TEST INFRASTRUCTURE ONLY.

``Oracle``  — oracle/liboracle.so, our plain-C restatement (oracle.c).
``Ref``     — oracle/_ref/libref.so, the unmodified reference objects compiled
              by oracle/Makefile (present when built in a container that has
              /root/reference; the file travels to the GPU box).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's CPU-baseline legs may
import this package, and only as the checker.  The product
(our-tree_b200/libbulkcipher.so) never touches it.

Both classes expose the same methods so a test can run one body against both.
All data arguments/results are ``numpy.uint8`` arrays or ``bytes``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ENC, DEC = 1, 0  # aes-modes/aes.h:33-34


def build(verbose: bool = False) -> None:
    """Compile liboracle.so and, when the reference tree is present, _ref/libref.so."""
    out = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


def _u8(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(bytes(a), dtype=np.uint8)
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_ubyte))


class _OrcKey(C.Structure):
    _fields_ = [("nr", C.c_int), ("rk", C.c_uint32 * 60)]


class _OrcArc4(C.Structure):
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("m", C.c_ubyte * 256)]


class _OrcRc4(C.Structure):
    _fields_ = [("perm", C.c_ubyte * 256), ("index1", C.c_int), ("index2", C.c_int)]


class Oracle:
    """Our restatement (oracle.c).  kind == "port"."""

    kind = "port"

    def __init__(self):
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        self.lib = C.CDLL(path)
        L = self.lib
        L.orc_aes_ecb.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p]
        L.orc_aes_cbc.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_aes_ctr.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p]
        L.orc_aesni_ecb_encrypt.argtypes = [C.c_void_p, C.c_void_p, C.c_ulong, C.c_void_p, C.c_int]
        L.orc_aesni_ecb_decrypt.argtypes = [C.c_void_p, C.c_void_p, C.c_ulong, C.c_void_p, C.c_int]
        L.orc_aesni_ctr_encrypt.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_ulong,
                                            C.c_void_p, C.c_int]
        L.orc_arc4_prep.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_arc4_crypt.argtypes = [C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_arc4_batch.argtypes = [C.c_size_t, C.c_void_p, C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p]
        L.orc_aes_cbc_encrypt_batch.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
                                                C.c_void_p, C.c_void_p]
        L.orc_rc4_crypt.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]

    def harness_input(self, n_msg: int, n_key: int = 32, seed: int = 1337):
        """srand(seed); msg[i]=rand()%255; key[i]=rand()%255 (aes-modes/test.c:54-62, :354)."""
        msg = np.empty(n_msg, np.uint8)
        key = np.empty(n_key, np.uint8)
        self.lib.orc_harness_input.argtypes = [C.c_uint, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        self.lib.orc_harness_input(seed, _p(msg), n_msg, _p(key), n_key)
        return msg, key.tobytes()

    # -- key schedules ----------------------------------------------------
    def _key(self, key: bytes, mode: int) -> _OrcKey:
        k = _OrcKey()
        f = self.lib.orc_aes_setkey_enc if mode == ENC else self.lib.orc_aes_setkey_dec
        rc = f(C.byref(k), bytes(key), C.c_uint(len(key) * 8))
        if rc:
            raise ValueError(rc)
        return k

    def setkey_rc(self, key: bytes, keybits: int, mode: int = ENC) -> int:
        k = _OrcKey()
        f = self.lib.orc_aes_setkey_enc if mode == ENC else self.lib.orc_aes_setkey_dec
        return f(C.byref(k), bytes(key), C.c_uint(keybits))

    def schedule(self, key: bytes, mode: int = ENC) -> np.ndarray:
        k = self._key(key, mode)
        return np.array(k.rk[: 4 * (k.nr + 1)], dtype=np.uint32)

    def packed_schedule(self, key: bytes, mode: int = ENC) -> bytes:
        return self.schedule(key, mode).astype("<u4").tobytes()

    def aesni_key256(self, key: bytes) -> bytes:
        out = np.zeros(240, np.uint8)
        self.lib.orc_aesni_key256(bytes(key), _p(out))
        return out.tobytes()

    # -- aes.h path -------------------------------------------------------
    def aes_ecb(self, key: bytes, mode: int, data) -> np.ndarray:
        d = _u8(data)
        out = np.empty_like(d)
        k = self._key(key, mode)
        self.lib.orc_aes_ecb(C.byref(k), mode, d.size, _p(d), _p(out))
        return out

    def aes_cbc(self, key: bytes, mode: int, iv: bytes, data):
        d = _u8(data)
        out = np.empty_like(d)
        ivb = np.array(_u8(iv))
        k = self._key(key, mode)
        rc = self.lib.orc_aes_cbc(C.byref(k), mode, d.size, _p(ivb), _p(d), _p(out))
        return rc, out, ivb.tobytes()

    def aes_ctr(self, key: bytes, nonce_counter: bytes, data, nc_off: int = 0,
                stream_block: bytes = bytes(16)):
        d = _u8(data)
        out = np.empty_like(d)
        nc = np.array(_u8(nonce_counter))
        sb = np.array(_u8(stream_block))
        off = C.c_int(nc_off)
        k = self._key(key, ENC)
        self.lib.orc_aes_ctr(C.byref(k), d.size, C.byref(off), _p(nc), _p(sb), _p(d), _p(out))
        return out, off.value, nc.tobytes(), sb.tobytes()

    def aes_cfb128(self, key: bytes, mode: int, iv: bytes, data, iv_off: int = 0):
        d = _u8(data)
        out = np.empty_like(d)
        ivb = np.array(_u8(iv))
        off = C.c_int(iv_off)
        k = self._key(key, ENC)
        self.lib.orc_aes_cfb128.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.POINTER(C.c_int), C.c_void_p,
                                            C.c_void_p, C.c_void_p]
        self.lib.orc_aes_cfb128(C.byref(k), mode, d.size, C.byref(off), _p(ivb), _p(d), _p(out))
        return out, off.value, ivb.tobytes()

    def aes_cbc_encrypt_batch(self, key: bytes, ivs, data, msg_len: int):
        d = _u8(data)
        out = np.empty_like(d)
        iv = np.array(_u8(ivs))
        k = self._key(key, ENC)
        self.lib.orc_aes_cbc_encrypt_batch(C.byref(k), d.size // msg_len, msg_len, _p(iv), _p(d), _p(out))
        return out, iv

    # -- aesni.h path -----------------------------------------------------
    def aesni_ecb(self, sched: bytes, rounds: int, mode: int, data) -> np.ndarray:
        d = _u8(data)
        out = np.empty_like(d)
        f = self.lib.orc_aesni_ecb_encrypt if mode == ENC else self.lib.orc_aesni_ecb_decrypt
        f(_p(d), _p(out), d.size, bytes(sched), rounds)
        return out

    def aesni_ctr(self, sched: bytes, rounds: int, ivec: bytes, nonce: bytes, data) -> np.ndarray:
        d = _u8(data)
        out = np.empty_like(d)
        self.lib.orc_aesni_ctr_encrypt(_p(d), _p(out), bytes(ivec), bytes(nonce), d.size, bytes(sched), rounds)
        return out

    # -- arc4 / rc4 -------------------------------------------------------
    def arc4_keystream(self, key: bytes, lengths) -> tuple[np.ndarray, int, int, bytes]:
        """arc4_setup then one arc4_prep per entry of ``lengths`` (chained)."""
        c = _OrcArc4()
        self.lib.orc_arc4_setup(C.byref(c), bytes(key), C.c_uint(len(key)))
        parts = []
        for n in lengths:
            ks = np.empty(n, np.uint8)
            self.lib.orc_arc4_prep(C.byref(c), n, _p(ks))
            parts.append(ks)
        return np.concatenate(parts) if parts else np.empty(0, np.uint8), c.x, c.y, bytes(c.m)

    def arc4(self, key: bytes, data) -> np.ndarray:
        d = _u8(data)
        ks, _, _, _ = self.arc4_keystream(key, [d.size])
        out = np.empty_like(d)
        self.lib.orc_arc4_crypt(d.size, _p(d), _p(ks), _p(out))
        return out

    def arc4_batch(self, keys, keylen: int, data, length: int) -> np.ndarray:
        d = _u8(data)
        k = _u8(keys)
        out = np.empty_like(d)
        self.lib.orc_arc4_batch(d.size // length if length else 0, _p(k), keylen, length, _p(d), _p(out))
        return out

    def rc4(self, key: bytes, chunks) -> tuple[np.ndarray, int, int, bytes]:
        """rc4_init then one rc4_crypt per chunk (chained); returns state too."""
        s = _OrcRc4()
        self.lib.orc_rc4_init(C.byref(s), bytes(key), len(key))
        outs = []
        for ch in chunks:
            d = _u8(ch)
            o = np.empty_like(d)
            self.lib.orc_rc4_crypt(C.byref(s), _p(d), _p(o), d.size)
            outs.append(o)
        return (np.concatenate(outs) if outs else np.empty(0, np.uint8)), s.index1, s.index2, bytes(s.perm)


class Ref:
    """The unmodified reference objects (aes.c, aesni.c, arc4.c).  kind == "reference"."""

    kind = "reference"

    @staticmethod
    def path() -> str:
        return os.path.join(_HERE, "_ref", "libref.so")

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(cls.path())

    def __init__(self):
        if not self.available():
            try:
                build()
            except Exception:
                pass
        if not self.available():
            raise FileNotFoundError(self.path())
        self.lib = C.CDLL(self.path())
        L = self.lib
        L.ref_sizeof_aes_context.restype = C.c_size_t
        L.ref_sizeof_arc4_context.restype = C.c_size_t
        self.ctx_size = L.ref_sizeof_aes_context()
        L.ref_aes_ecb_bulk.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_aes_crypt_cbc.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_aes_ctr_bulk.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p]
        L.ref_AES_ECB_encrypt.argtypes = [C.c_void_p, C.c_void_p, C.c_ulong, C.c_void_p, C.c_int]
        L.ref_AES_ECB_decrypt.argtypes = [C.c_void_p, C.c_void_p, C.c_ulong, C.c_void_p, C.c_int]
        L.ref_AES_CTR_encrypt.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_ulong,
                                          C.c_void_p, C.c_int]
        L.ref_AES_ECB_encrypt.restype = None
        L.ref_AES_ECB_decrypt.restype = None
        L.ref_AES_CTR_encrypt.restype = None
        L.ref_arc4_prep.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_arc4_crypt.argtypes = [C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_arc4_batch.argtypes = [C.c_size_t, C.c_void_p, C.c_uint, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_aes_cbc_encrypt_batch.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
                                                C.c_void_p, C.c_void_p]
        L.ref_pack_schedule.argtypes = [C.c_void_p, C.c_void_p]

    def has_aesni(self) -> bool:
        return bool(self.lib.ref_CheckAESSupport())

    def self_tests(self) -> tuple[int, int]:
        return self.lib.ref_aes_self_test(0), self.lib.ref_arc4_self_test(0)

    def _ctx(self, key: bytes, mode: int):
        ctx = C.create_string_buffer(self.ctx_size)
        f = self.lib.ref_aes_setkey_enc if mode == ENC else self.lib.ref_aes_setkey_dec
        rc = f(ctx, bytes(key), C.c_uint(len(key) * 8))
        if rc:
            raise ValueError(rc)
        return ctx

    def setkey_rc(self, key: bytes, keybits: int, mode: int = ENC) -> int:
        ctx = C.create_string_buffer(self.ctx_size)
        f = self.lib.ref_aes_setkey_enc if mode == ENC else self.lib.ref_aes_setkey_dec
        return f(ctx, bytes(key), C.c_uint(keybits))

    def packed_schedule(self, key: bytes, mode: int = ENC) -> bytes:
        ctx = self._ctx(key, mode)
        nr = C.cast(ctx, C.POINTER(C.c_int))[0]
        out = np.zeros(16 * (nr + 1), np.uint8)
        self.lib.ref_pack_schedule(ctx, _p(out))
        return out.tobytes()

    def schedule(self, key: bytes, mode: int = ENC) -> np.ndarray:
        return np.frombuffer(self.packed_schedule(key, mode), dtype="<u4").astype(np.uint32)

    def aesni_key256(self, key: bytes) -> bytes:
        # the reference stores with aligned 128-bit moves (aesni.c:40): give it a 16-aligned buffer
        raw = np.zeros(240 + 16, np.uint8)
        off = (-raw.ctypes.data) % 16
        self.lib.ref_AES_256_Key_Expansion(bytes(key), C.c_void_p(raw.ctypes.data + off))
        return raw[off:off + 240].tobytes()

    def aes_ecb(self, key: bytes, mode: int, data) -> np.ndarray:
        d = _u8(data)
        out = np.empty_like(d)
        self.lib.ref_aes_ecb_bulk(self._ctx(key, mode), mode, d.size, _p(d), _p(out))
        return out

    def aes_cbc(self, key: bytes, mode: int, iv: bytes, data):
        d = _u8(data)
        out = np.empty_like(d)
        ivb = np.array(_u8(iv))
        rc = self.lib.ref_aes_crypt_cbc(self._ctx(key, mode), mode, d.size, _p(ivb), _p(d), _p(out))
        return rc, out, ivb.tobytes()

    def aes_ctr(self, key: bytes, nonce_counter: bytes, data, nc_off: int = 0,
                stream_block: bytes = bytes(16)):
        d = _u8(data)
        out = np.empty_like(d)
        nc = np.array(_u8(nonce_counter))
        sb = np.array(_u8(stream_block))
        off = C.c_int(nc_off)
        self.lib.ref_aes_ctr_bulk(self._ctx(key, ENC), d.size, C.byref(off), _p(nc), _p(sb), _p(d), _p(out))
        return out, off.value, nc.tobytes(), sb.tobytes()

    def aes_cfb128(self, key: bytes, mode: int, iv: bytes, data, iv_off: int = 0):
        d = _u8(data)
        out = np.empty_like(d)
        ivb = np.array(_u8(iv))
        off = C.c_int(iv_off)
        self.lib.ref_aes_crypt_cfb128.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.POINTER(C.c_int), C.c_void_p,
                                                  C.c_void_p, C.c_void_p]
        self.lib.ref_aes_crypt_cfb128(self._ctx(key, ENC), mode, d.size, C.byref(off), _p(ivb), _p(d), _p(out))
        return out, off.value, ivb.tobytes()

    def aes_cbc_encrypt_batch(self, key: bytes, ivs, data, msg_len: int):
        d = _u8(data)
        out = np.empty_like(d)
        iv = np.array(_u8(ivs))
        self.lib.ref_aes_cbc_encrypt_batch(self._ctx(key, ENC), d.size // msg_len, msg_len, _p(iv), _p(d), _p(out))
        return out, iv

    @staticmethod
    def _aligned(sched: bytes):
        raw = np.zeros(len(sched) + 16, np.uint8)
        off = (-raw.ctypes.data) % 16
        raw[off:off + len(sched)] = np.frombuffer(sched, np.uint8)
        return raw, C.c_void_p(raw.ctypes.data + off)

    def aesni_ecb(self, sched: bytes, rounds: int, mode: int, data) -> np.ndarray:
        d = _u8(data)
        out = np.empty_like(d)
        keep, kp = self._aligned(sched)   # aesni.c dereferences key[] as aligned __m128i
        f = self.lib.ref_AES_ECB_encrypt if mode == ENC else self.lib.ref_AES_ECB_decrypt
        f(_p(d), _p(out), d.size, kp, rounds)
        return out

    def aesni_ctr(self, sched: bytes, rounds: int, ivec: bytes, nonce: bytes, data) -> np.ndarray:
        d = _u8(data)
        out = np.empty_like(d)
        keep, kp = self._aligned(sched)
        nonce8 = bytes(nonce) + bytes(4)  # the reference reads 8 bytes through `*(long*)nonce` (aesni.c:136)
        self.lib.ref_AES_CTR_encrypt(_p(d), _p(out), bytes(ivec), nonce8, d.size, kp, rounds)
        return out

    def arc4_keystream(self, key: bytes, lengths):
        c = _OrcArc4()  # same layout as arc4_context (arc4.h:35-41)
        self.lib.ref_arc4_setup(C.byref(c), bytes(key), C.c_uint(len(key)))
        parts = []
        for n in lengths:
            ks = np.empty(n, np.uint8)
            self.lib.ref_arc4_prep(C.byref(c), n, _p(ks))
            parts.append(ks)
        return np.concatenate(parts) if parts else np.empty(0, np.uint8), c.x, c.y, bytes(c.m)

    def arc4(self, key: bytes, data) -> np.ndarray:
        d = _u8(data)
        ks, _, _, _ = self.arc4_keystream(key, [d.size])
        out = np.empty_like(d)
        self.lib.ref_arc4_crypt(d.size, _p(d), _p(ks), _p(out))
        return out

    def arc4_batch(self, keys, keylen: int, data, length: int) -> np.ndarray:
        d = _u8(data)
        k = _u8(keys)
        out = np.empty_like(d)
        self.lib.ref_arc4_batch(d.size // length if length else 0, _p(k), keylen, length, _p(d), _p(out))
        return out

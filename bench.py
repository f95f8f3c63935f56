#!/usr/bin/env python
"""bench.py — the contract benchmark.

Workload (BASELINE.json configs[1]; at N>1 configs[4]'s sharding): AES-128-CTR
over a 4 GiB synthetic buffer per GPU, sharded by counter offset across N GPUs
with no collective on the data path ("weak" scaling: per-GPU work is fixed).
A step = one pass of the hot path over the rank's resident 4 GiB shard.

  python bench.py [--gpus N] [--steps K] [--warmup W]          our CUDA path
  python bench.py --impl reference [...]                        the reference's own CPU path
  torchrun ... bench.py --gpus N ...                            one rank per GPU (N > 1)

Prints ONE JSON line on rank 0.  `value` is device-resident throughput in GB/s
(10^9 plaintext bytes per second, all ranks); `e2e` is the same metric through
the host-buffer C-ABI call (bc_aes_ctr_host: pinned host memory, H2D + kernel +
D2H inside the timed region); `roofline` is for the dominant kernel
(aes_ctr_kernel<10>) against the measured HBM copy bandwidth with 2 algorithmic
bytes of HBM traffic per plaintext byte; `cpu_baseline` is the reference's
aesni.c AES_CTR_encrypt (oracle/_ref, built from the unmodified sources) on all
host cores over a bounded sample.  At N > 1 the line also carries `config5`:
BASELINE configs[4] as written (256 GiB sharded by counter offset over the N
GPUs, filled on the device by a generator the host can replay, encrypted, and a
stratified sample of windows — shard ends, every 64 GiB seam, random — compared
with the reference's aes.c outside the timed region).  oracle/ is used only as
the CPU baseline and as that checker, never inside a timed region.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GIB = 1 << 30
SHARD_BYTES = 4 * GIB
KEY = bytes.fromhex("f9dfdf0c3cf14e4b9c296478904f979b")          # key[0..15] of the config-1 generator
COUNTER = bytes.fromhex("0f1e2d3c4b5a69788796a5b4fffffff0")      # bytes 12-15 = FF FF FF F0: the 2^32 carry is crossed
METRIC = "AES-128-CTR throughput"
UNIT = "GB/s"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s; MEASURED_PEAKS.json absent)"


def profile_traffic():
    """Per-launch dram bytes of aes_ctr_kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ctr128_ncu_summary.json")) as f:
            j = json.load(f)
        return j.get("dram_bytes_per_launch_4gib")
    except Exception:
        return None


SEED = 1337                                                     # srand(1337), aes-modes/test.c:354
M64 = (1 << 64) - 1


def splitmix_bytes(first_word: int, n_words: int) -> np.ndarray:
    """Host replay of bc_fill_splitmix64_dev (include/bulkcipher.h): bytes of words [first_word, +n_words)."""
    with np.errstate(over="ignore"):
        z = (np.arange(n_words, dtype=np.uint64) + np.uint64((SEED + first_word) & M64)) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z.view(np.uint8)


def checker():
    """The CPU checker for spot checks OUTSIDE timed regions: the reference's own aes.c
    (oracle/_ref/libref.so) when it travelled, else the oracle port."""
    import oracle

    return (oracle.Ref(), "oracle/_ref aes.c aes_crypt_ctr") if oracle.Ref.available() else (oracle.Oracle(), "oracle.c port")


def verify_ctr_windows(chk, d_cipher, shard_first_word: int, shard_first_block: int, offsets, win: int) -> int:
    """d_cipher (device tensor, this rank's shard) must equal splitmix plaintext XOR the reference's
    AES-128-CTR keystream at every window [off, off+win).  Returns the number of windows checked."""
    base = int.from_bytes(COUNTER, "big")
    for off in offsets:
        got = d_cipher[off:off + win].cpu().numpy()
        plain = splitmix_bytes(shard_first_word + off // 8, win // 8)
        nc = ((base + shard_first_block + off // 16) % (1 << 128)).to_bytes(16, "big")
        want = chk.aes_ctr(KEY, nc, plain)[0]
        if not np.array_equal(got, want):
            bad = int(np.flatnonzero(got != want)[0])
            raise SystemExit(f"PARITY FAILURE: window at shard byte {off} differs from the reference at byte {bad}")
    return len(offsets)


def host_facts(L, local: int):
    nodes = 0
    try:
        nodes = len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()])
    except OSError:
        pass
    mem_gib = None
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    mem_gib = int(ln.split()[1]) >> 20
    except OSError:
        pass
    return {"cpus": cpu_threads(), "numa_nodes": nodes, "gpu_numa_node": int(L.bc_device_numa_node(local)),
            "mem_available_gib": mem_gib}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [ln for (t, ln) in self.lines if t0 - 0.05 <= t <= t1 + 0.15] or [ln for (_, ln) in self.lines]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in rows:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------ reference arm

def cpu_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ReferenceCtr:
    """AES-128-CTR on the host cores with the reference's own code: aesni.c's AES_CTR_encrypt
    (oracle/_ref, unmodified sources) on equal contiguous slices, one thread per slice, like
    aes-modes/test.c:273-285 (each slice restarts the counter, as the reference harness does —
    fine for timing).  Falls back to aes.c's aes_crypt_ctr when the CPU lacks AES-NI and to the
    oracle port when libref.so did not travel."""

    def __init__(self, sample_bytes: int, threads: int):
        import ctypes as C
        from concurrent.futures import ThreadPoolExecutor

        import oracle

        self.threads = threads
        per = self.per = (sample_bytes // threads) // 16 * 16
        self.nbytes = per * threads
        buf = self.buf = np.resize(np.frombuffer(np.random.default_rng(1337).bytes(1 << 24), dtype=np.uint8), self.nbytes)
        out = self.out = np.empty_like(buf)
        use_ref = oracle.Ref.available()
        ref = oracle.Ref() if use_ref else None
        if use_ref and ref.has_aesni():
            self.kind, self.what = "reference", "aesni.c AES_CTR_encrypt (oracle/_ref/libref.so, gcc -O2)"
            sched = ref.packed_schedule(KEY, 1)
            self._keep, kp = ref._aligned(sched)
            nonce8 = COUNTER[0:4] + bytes(4)

            def work(t):
                off = t * per
                ref.lib.ref_AES_CTR_encrypt(C.c_void_p(buf.ctypes.data + off), C.c_void_p(out.ctypes.data + off),
                                            COUNTER[4:12], nonce8, per, kp, 10)
        else:
            impl = ref if use_ref else oracle.Oracle()
            self.kind, self.what = ("reference", "aes.c aes_crypt_ctr (oracle/_ref/libref.so, gcc -O2)") if use_ref else \
                                   ("port", "oracle.c orc_aes_ctr (restatement of aes.c, gcc -O2)")

            def work(t):
                off = t * per
                out[off:off + per] = impl.aes_ctr(KEY, COUNTER, buf[off:off + per])[0]

        self.work = work
        self.pool = ThreadPoolExecutor(threads)

    def one_pass(self) -> float:
        t0 = time.perf_counter()
        list(self.pool.map(self.work, range(self.threads)))
        return time.perf_counter() - t0

    def gbs(self, passes: int, warm: int = 1) -> float:
        for _ in range(warm):
            self.one_pass()
        return self.nbytes / float(np.median([self.one_pass() for _ in range(passes)])) / 1e9


def reference_extras(threads: int):
    """The reference's other CPU paths on the same host cores, bounded samples: aes.c's table-driven
    aes_crypt_ctr (the path the kernels restate) and arc4.c over independent 4 KiB streams
    (setup + prep + crypt each, BASELINE config 4's shape).  GB/s over all `threads` cores."""
    from concurrent.futures import ThreadPoolExecutor

    import oracle

    use_ref = oracle.Ref.available()
    impl = oracle.Ref() if use_ref else oracle.Oracle()
    tag = "oracle/_ref/libref.so" if use_ref else "oracle.c port"
    rng = np.random.default_rng(99)
    per = 8 << 20                                               # 8 MiB per thread per pass
    buf = np.frombuffer(rng.bytes(per), dtype=np.uint8)
    keys = np.frombuffer(rng.bytes((per // 4096) * 16), dtype=np.uint8)
    out = {}
    with ThreadPoolExecutor(threads) as ex:
        def ctr(_):
            impl.aes_ctr(KEY, COUNTER, buf)
        def rc4(_):
            impl.arc4_batch(keys, 16, buf, 4096)
        for name, fn, what in (("aes_c_ctr", ctr, "aes.c aes_crypt_ctr"), ("arc4_c_batched_4k", rc4, "arc4.c setup+prep+crypt per 4 KiB stream")):
            list(ex.map(fn, range(threads)))
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                list(ex.map(fn, range(threads)))
                ts.append(time.perf_counter() - t0)
            out[name] = {"value": round(per * threads / float(np.median(ts)) / 1e9, 3), "unit": UNIT, "cores": threads,
                         "sample": f"{what} ({tag}, gcc -O2); {per >> 20} MiB per thread x {threads} threads, median of 3"}
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = cpu_threads()
    t_all0 = time.perf_counter()
    # the whole 4 GiB-per-GPU workload per step (tests/test_bench_contract.py shrinks it through the environment)
    r = ReferenceCtr(int(os.environ.get("BULKCIPHER_BENCH_REF_MIB", SHARD_BYTES >> 20)) << 20, threads)
    for _ in range(args.warmup):
        r.one_pass()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r.one_pass()
    dt = time.perf_counter() - t0
    value = r.nbytes * args.steps / dt / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "AES-128-CTR encrypt, the reference's own CPU path on all host cores; each step a "
                               f"{r.nbytes >> 20} MiB = the whole 4 GiB-per-GPU workload", "key_bits": 128,
                   "bytes_per_gpu": r.nbytes, "sample_bytes": r.nbytes},
        "cpu_baseline": {"value": round(value, 3), "unit": UNIT, "cores": threads, "kind": r.kind,
                         "sample": f"{r.what}; {r.nbytes >> 20} MiB per step split over {threads} threads"},
        "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": round(time.perf_counter() - t_all0, 1),
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------ our arm

def link_probe(d_a, d_b, ranks, n_gpus: int):
    """What the host links allow for an in+out transfer: every rank copies 1 GiB H2D and 1 GiB D2H at
    the same time (pinned memory, monolithic copies on two streams).  Each trial starts behind a
    barrier on all N ranks and is timed as the MAX over ranks, so the figure is what the box gives
    when all N links are busy for the whole trial: aggregate GB/s per direction = N GiB / slowest
    rank's time.  The e2e figure cannot exceed it.  Returns (aggregate, slowest rank's own rate)."""
    import torch

    n = min(GIB, d_a.numel())
    h_a = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h_b = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    h_a.fill_(3)
    h_b.fill_(5)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    best = 1e9
    for _ in range(4):
        ranks.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_event(e0)
        s2.wait_event(e0)
        with torch.cuda.stream(s1):
            d_a[:n].copy_(h_a, non_blocking=True)
        with torch.cuda.stream(s2):
            h_b.copy_(d_b[:n], non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, ranks.reduce([e0.elapsed_time(e1)], "max")[0])
    return n_gpus * n / best / 1e6, n / best / 1e6


class Ranks:
    """The only cross-rank traffic of the bench: a barrier and MAX / MIN / SUM of a few floats
    (timings).  No collective touches cipher data."""

    def __init__(self, distributed: bool):
        self.on = distributed

    def barrier(self):
        import torch
        import torch.distributed as dist

        torch.cuda.synchronize()
        if self.on:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(self, values, op: str):
        import torch
        import torch.distributed as dist

        if not self.on:
            return [float(v) for v in values]
        t = torch.tensor([float(v) for v in values], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN, "sum": dist.ReduceOp.SUM}[op])
        return [float(x) for x in t.cpu()]


def other_workloads(pkg, L, d_in, d_out, st, hbm_peak, ranks: Ranks, n_gpus: int, rank: int):
    """BASELINE configs 3 and 4 (and the metric's "batched RC4 GB/s"), device resident on 4 GiB per
    GPU, EVERY rank on its own shard (messages / streams / blocks [rank*n, (rank+1)*n) of the job):
    K launches back to back between two CUDA events after a barrier, MAX over ranks; value = all
    ranks' plaintext bytes over that time, hbm_frac per GPU at the algorithmic 2 bytes of traffic
    per byte.  Reported next to the headline, not instead of it."""
    import torch

    n = d_in.numel()
    key32 = bytes(range(32))
    K = 5

    def sched(key, enc=True):
        rc, ctx = (pkg.aes_setkey_enc if enc else pkg.aes_setkey_dec)(key)
        return pkg.schedule_words(ctx), ctx.nr

    e256, nr = sched(key32)
    d256, _ = sched(key32, enc=False)
    iv = bytes(range(16))
    n_msgs = n // 65536
    n_streams = n // 4096
    g = torch.Generator(device="cuda")
    g.manual_seed(4242 + rank)
    d_iv = torch.randint(0, 256, (n_msgs * 16,), dtype=torch.uint8, device="cuda", generator=g)
    d_keys = torch.randint(0, 256, (n_streams * 16,), dtype=torch.uint8, device="cuda", generator=g)
    ip, op = d_in.data_ptr(), d_out.data_ptr()
    runs = {
        "aes256_ecb_encrypt_4gib": (lambda: pkg.check(L.bc_aes_ecb_dev(e256.ctypes.data, nr, 1, ip, op, n, st)), 2.0 * n),
        "aes256_cbc_decrypt_4gib": (lambda: pkg.check(L.bc_aes_cbc_decrypt_dev(d256.ctypes.data, nr, iv, ip, op, n, st)), 2.0 * n),
        "aes256_cbc_encrypt_65536x64k": (lambda: pkg.check(L.bc_aes_cbc_encrypt_batch_dev(
            e256.ctypes.data, nr, n_msgs, 65536, d_iv.data_ptr(), None, ip, op, st)), 2.0 * n + 16.0 * n_msgs),
        "rc4_batched_1m_keys_x_4k": (lambda: pkg.check(L.bc_arc4_batch_dev(
            n_streams, d_keys.data_ptr(), 16, 4096, ip, op, None, None, st)), 2.0 * n + 16.0 * n_streams),
        # in == out (how the reference itself calls rc4_crypt, test.c:163-166): the TMA unit XORs the keystream into HBM
        "rc4_batched_1m_keys_x_4k_in_place": (lambda: pkg.check(L.bc_arc4_batch_dev(
            n_streams, d_keys.data_ptr(), 16, 4096, op, op, None, None, st)), 2.0 * n + 16.0 * n_streams),
    }
    out = {}
    for name, (fn, traffic) in runs.items():
        fn()
        fn()
        ranks.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(K):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = ranks.reduce([a.elapsed_time(b) / K], "max")[0]
        out[name] = {"value": round(n_gpus * n / ms / 1e6, 1), "unit": UNIT, "kernel_ms": round(ms, 4),
                     "hbm_frac": round(traffic / ms / 1e6 / hbm_peak, 4), "per_gpu_bytes": n, "launches": K}
    # batched RC4 end to end: pinned host buffers through bc_arc4_batch_host (1 GiB = 2^18 streams x 4 KiB per GPU)
    hn = GIB
    h_in, h_out, h_keys = L.bc_host_alloc(hn), L.bc_host_alloc(hn), L.bc_host_alloc((hn // 4096) * 16)
    if h_in and h_out and h_keys:
        np.ctypeslib.as_array((ctypes.c_ubyte * hn).from_address(h_in))[:] = 0x5A
        np.ctypeslib.as_array((ctypes.c_ubyte * ((hn // 4096) * 16)).from_address(h_keys))[:] = np.resize(
            np.frombuffer(np.random.default_rng(5 + rank).bytes(1 << 20), dtype=np.uint8), (hn // 4096) * 16)
        pkg.check(L.bc_arc4_batch_host(hn // 4096, h_keys, 16, 4096, h_in, h_out, None, None))
        ranks.barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            pkg.check(L.bc_arc4_batch_host(hn // 4096, h_keys, 16, 4096, h_in, h_out, None, None))
        dt = ranks.reduce([time.perf_counter() - t0], "max")[0]
        out["rc4_batched_host_buffers_1gib"] = {"value": round(n_gpus * hn * 3 / dt / 1e9, 2), "unit": UNIT,
                                                 "api": "bc_arc4_batch_host (pinned buffers, H2D + kernel + D2H), 1 GiB per GPU"}
    for p in (h_in, h_out, h_keys):
        L.bc_host_free(p)
    return out


def run_config5(pkg, L, rk, nr, st, ranks: Ranks, n_gpus: int, rank: int, steps: int):
    """BASELINE configs[4] as written: AES-128-CTR over a 256 GiB logical buffer, GPU g of N owning
    bytes [g*256/N GiB, (g+1)*256/N GiB) with its counter offset by its first block index.  The
    buffer is generated in HBM (bc_fill_splitmix64_dev, seed 1337 + logical word index), so the host
    can replay any window; where input + output do not both fit (N = 2: 128 GiB per GPU) the shard
    is encrypted in place.  After the timed steps the shard is regenerated, encrypted once more and
    a stratified sample of 64 KiB windows — both shard ends, a window across every 64 GiB seam (the
    2^32-block carry into counter byte 11, aes.c:884-888), 32 random ones — is compared with the
    reference's aes.c, outside the timed region."""
    import torch

    total = 256 * GIB
    shard = total // n_gpus
    free_b, _ = torch.cuda.mem_get_info()
    free_b = int(ranks.reduce([free_b], "min")[0])                # one decision for all ranks (they meet at barriers below)
    in_place = 2 * shard > 0.92 * free_b
    if shard > 0.95 * free_b:
        return {"skipped": f"a {shard >> 30} GiB shard does not fit in {free_b >> 30} GiB of free HBM"}
    d_a = torch.empty(shard, dtype=torch.uint8, device="cuda")
    d_b = d_a if in_place else torch.empty(shard, dtype=torch.uint8, device="cuda")
    first_word = rank * (shard // 8)
    first_block = rank * (shard // 16)

    def fill():
        pkg.check(L.bc_fill_splitmix64_dev(d_a.data_ptr(), shard, SEED, first_word, st))

    def step():
        pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, nr, COUNTER, 128, first_block, d_a.data_ptr(), d_b.data_ptr(),
                                   shard, None, st))

    fill()
    for _ in range(2):
        step()
    ranks.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    ranks.barrier()
    ms = ranks.reduce([a.elapsed_time(b)], "max")[0]

    # ---- verification, outside the timed region
    fill()
    step()
    torch.cuda.synchronize()
    win = 65536
    offs = {0, shard - win}
    seam = 64 * GIB
    lo, hi = rank * shard, (rank + 1) * shard
    n_seams = 0
    for k in range(1, total // seam):
        pos = k * seam
        if lo < pos < hi:                                         # seam inside the shard: straddle it
            offs.add(pos - lo - win // 2)
            n_seams += 1
        elif pos == lo:                                           # seam at the shard start: the carry (byte pos + 256) is in window 0
            n_seams += 1
    rng = np.random.default_rng(1000 + rank)
    for r in rng.integers(0, (shard - win) // 16, 32):
        offs.add(int(r) * 16)
    chk, chk_name = checker()
    t0 = time.perf_counter()
    n_checked = verify_ctr_windows(chk, d_b, first_word, first_block, sorted(offs), win)
    verify_s = time.perf_counter() - t0
    tot_windows, tot_seams = ranks.reduce([n_checked, n_seams], "sum")
    del d_a, d_b
    torch.cuda.empty_cache()
    return {"workload": f"AES-128-CTR over a 256 GiB logical buffer sharded by counter offset over {n_gpus} GPUs "
                        "(BASELINE configs[4]); generated on the device, host-replayable",
            "value": round(total * steps / (ms * 1e-3) / 1e9, 1), "unit": UNIT, "ms_per_step": round(ms / steps, 3),
            "steps": steps, "bytes_per_gpu": shard, "in_place": bool(in_place),
            "verified_windows": int(tot_windows), "window_bytes": win, "seam_windows": int(tot_seams),
            "verified_against": chk_name, "verify_s_rank0": round(verify_s, 2),
            "hbm_frac_per_gpu": round(2.0 * shard * steps / (ms * 1e-3) / 1e9 / measured_peak()[0], 4)}


def run_ours(args):
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_gpus = world if distributed else 1
    ranks = Ranks(distributed)

    pkg = g.load_package()
    if not os.path.exists(pkg.LIB_PATH) and rank == 0:
        g.build()                                                 # fresh checkout: compile once (nvcc is in the image)
    if distributed:
        dist.barrier()
    L = pkg.lib()
    pkg.check(L.bc_set_device(local))
    # host-side locality before any pinned allocation: this rank's thread (and the pages it
    # touches) go to the NUMA node its GPU hangs off; a no-op on single-node boxes
    if not args.no_numa_bind:
        L.bc_bind_host_thread_to_device(local)
    facts = host_facts(L, local)
    rc, ctx = pkg.aes_setkey_enc(KEY)
    rk = pkg.schedule_words(ctx)
    nr = ctx.nr
    nbytes = int(args.shard_gib * GIB) // 65536 * 65536
    blocks = nbytes // 16
    first_block = rank * blocks                                   # shard by counter offset
    first_word = rank * (nbytes // 8)
    st = torch.cuda.current_stream().cuda_stream

    d_in = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    # synthetic plaintext generated in HBM, replayable on the host (checked below against the reference)
    pkg.check(L.bc_fill_splitmix64_dev(d_in.data_ptr(), nbytes, SEED, first_word, st))
    in_place = nbytes > 48 * GIB
    d_out = d_in if in_place else torch.empty_like(d_in)

    def step():
        pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, nr, COUNTER, 128, first_block, d_in.data_ptr(), d_out.data_ptr(),
                                   nbytes, None, st))

    barrier = ranks.barrier
    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0 = time.time()
    e_begin, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_begin.record()
    for a, b in ev:
        a.record()
        step()
        b.record()
    e_end.record()
    barrier()
    t1 = time.time()
    total_ms = e_begin.elapsed_time(e_end)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    # a longer soak for the clock record only (the timed region above lasts tens of ms,
    # shorter than nvidia-smi's sampling period)
    soak_t0 = time.time()
    if rank == 0:
        while time.time() - soak_t0 < 1.0:
            step()
        torch.cuda.synchronize()
    soak_t1 = time.time()
    clocks = sampler.stop(t0, soak_t1) if rank == 0 else None
    total_ms, kernel_ms = ranks.reduce([total_ms, kernel_ms], "max")

    # ---- parity of the timed output, outside the timed region: three 64 KiB windows of this rank's
    # ciphertext (first — it holds the 2^32 counter carry at block 16 —, middle, last) against the
    # reference's aes.c on the replayed plaintext
    parity = None
    if not in_place and (args.steps + args.warmup) > 0:
        chk, chk_name = checker()
        win = 65536
        offs = [0, (nbytes // 2) // win * win, nbytes - win]
        n_ok = verify_ctr_windows(chk, d_out, first_word, first_block, offs, win)
        parity = {"windows": int(ranks.reduce([n_ok], "sum")[0]), "window_bytes": win, "against": chk_name,
                  "what": "timed device-resident output vs the reference on the replayed splitmix plaintext"}

    # ---- e2e: the host-buffer C-ABI call, pinned host memory, H2D + kernel + D2H every step.
    # The 4 GiB shard streams through a pinned window: the whole shard when the host has room for
    # 2 x 4 GiB of pinned memory per rank, else 1 GiB walked four times with the counter offset
    # advancing (same bytes over the link per step).
    e2e_bytes = min(nbytes, SHARD_BYTES)
    roomy = (facts.get("mem_available_gib") or 0) >= 2 * n_gpus * (2 * e2e_bytes >> 30) + 16
    window = e2e_bytes if (n_gpus <= 2 or roomy) else GIB
    h_in = h_out = None
    while window >= (64 << 20):
        h_in, h_out = (L.bc_host_alloc_wc if args.wc_input else L.bc_host_alloc)(window), L.bc_host_alloc(window)
        if h_in and h_out:
            break
        L.bc_host_free(h_in)
        L.bc_host_free(h_out)
        h_in = h_out = None
        window //= 2
    if not h_in or not h_out:
        raise SystemExit("pinned host allocation failed: " + pkg.last_error())
    passes = e2e_bytes // window
    hin = np.ctypeslib.as_array((ctypes.c_ubyte * window).from_address(h_in))
    hin[:] = np.resize(np.frombuffer(np.random.default_rng(7 + rank).bytes(1 << 24), dtype=np.uint8), window)
    np.ctypeslib.as_array((ctypes.c_ubyte * window).from_address(h_out))[:] = 0      # first touch before timing
    last = np.zeros(16, np.uint8)

    def e2e_step():
        for p in range(passes):                                 # each call returns after its last D2H has landed
            pkg.check(L.bc_aes_ctr_host(rk.ctypes.data, nr, COUNTER, 128, first_block + p * (window // 16), h_in, h_out,
                                        window, last.ctypes.data))

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    w0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = ranks.reduce([time.perf_counter() - w0], "max")[0]
    # spot-check the e2e output (last window): the device-resident path on the same counter stream, and the reference
    d_chk = torch.from_numpy(hin[: 1 << 20].copy()).cuda()
    d_chk_out = torch.empty_like(d_chk)
    fb_last = first_block + (passes - 1) * (window // 16)
    pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, nr, COUNTER, 128, fb_last, d_chk.data_ptr(), d_chk_out.data_ptr(), 1 << 20, None, st))
    torch.cuda.synchronize()
    hout = np.ctypeslib.as_array((ctypes.c_ubyte * (1 << 20)).from_address(h_out))
    if not np.array_equal(hout, d_chk_out.cpu().numpy()):
        raise SystemExit("e2e output differs from the device-resident path")
    # ... and 1 MiB windows across the pipeline's chunk seams (the ramp starts with 4 / 8 / 16 / 32 MiB chunks) and the end
    hall = np.ctypeslib.as_array((ctypes.c_ubyte * window).from_address(h_out))
    for woff in (3 << 20, 11 << 20, 27 << 20, 59 << 20, window // 2, window - (1 << 20)):
        if woff + (1 << 20) > window:
            continue
        d_chk.copy_(torch.from_numpy(hin[woff:woff + (1 << 20)].copy()))
        pkg.check(L.bc_aes_ctr_dev(rk.ctypes.data, nr, COUNTER, 128, fb_last + woff // 16, d_chk.data_ptr(), d_chk_out.data_ptr(),
                                   1 << 20, None, st))
        torch.cuda.synchronize()
        if not np.array_equal(hall[woff:woff + (1 << 20)], d_chk_out.cpu().numpy()):
            raise SystemExit(f"e2e output differs from the device-resident path at host offset {woff}")
    chk, _ = checker()
    nc = ((int.from_bytes(COUNTER, "big") + fb_last) % (1 << 128)).to_bytes(16, "big")
    if not np.array_equal(hout[:65536], chk.aes_ctr(KEY, nc, hin[:65536])[0]):
        raise SystemExit("PARITY FAILURE: e2e output differs from the reference")
    L.bc_host_free(h_in)
    L.bc_host_free(h_out)

    # ---- what the host links allow with all N GPUs copying in and out at once
    link = None
    if not in_place:
        try:
            agg, each = link_probe(d_in, d_out, ranks, n_gpus)
            link = {"min": each, "sum": agg}
        except Exception as e:  # noqa: BLE001
            link = {"error": str(e)}

    # ---- the rest of BASELINE.json's metric / configs, device resident, every rank on its own GPU
    others = None
    if not args.no_others and not in_place and nbytes == SHARD_BYTES:
        others = other_workloads(pkg, L, d_in, d_out, st, measured_peak()[0], ranks, n_gpus, rank)

    # ---- BASELINE configs[4] as written (N > 1): 256 GiB over the N GPUs, verified
    config5 = None
    if n_gpus > 1 and not args.no_config5 and nbytes == SHARD_BYTES:
        del d_in, d_out
        torch.cuda.empty_cache()
        config5 = run_config5(pkg, L, rk, nr, st, ranks, n_gpus, rank, max(2, min(args.steps, 5)))

    if rank != 0:
        if distributed:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peak()
    value = n_gpus * nbytes * args.steps / (total_ms * 1e-3) / 1e9
    achieved = 2.0 * nbytes / (kernel_ms * 1e-3) / 1e9           # 2 algorithmic HBM bytes per plaintext byte
    e2e_value = n_gpus * e2e_bytes * e2e_steps / e2e_s / 1e9
    sm_clock_ghz = (clocks.get("sm_mhz") or 1965.0) / 1e3
    sms = L.bc_sm_count()
    lds_bound = 32.0 * sms * sm_clock_ghz / (140.0 / 16.0)       # GB/s plaintext: 132 lookups + 8 ld/st wavefronts per 16-byte block x 32 lanes
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"AES-128-CTR encrypt of a {nbytes / GIB:g} GiB synthetic buffer per GPU "
                               "(BASELINE configs[1] at 4 GiB); N GPUs shard by counter offset, no collective; "
                               "configs[4] (256 GiB over N GPUs) is the `config5` record at N > 1",
                   "key_bits": 128, "bytes_per_gpu": nbytes, "counter": COUNTER.hex(),
                   "input": "bc_fill_splitmix64_dev(seed 1337): generated in HBM, replayable on the host",
                   "l2": "inputs (>= 4 GiB) larger than L2 (126 MB); no flush needed", "in_place": in_place,
                   "host": facts},
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": profile_traffic(),
                     "kernel": "aes_ctr_kernel<10,1024>", "kernel_ms": round(kernel_ms, 4), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": 2 * nbytes,
                     "note": "the binding limit is shared-memory lookups, not HBM",
                     "issue_bound": {"what": "l1tex data pipe, 1 wavefront/clk/SM: 132 table lookups per block + 8 global ld/st wavefronts per 32 blocks",
                                     "gbs": round(lds_bound, 1), "frac": round(value / n_gpus / lds_bound, 4),
                                     "sm_ghz": sm_clock_ghz, "sms": sms}},
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": e2e_bytes, "d2h_bytes_per_step": e2e_bytes + 16,
                "steps": e2e_steps, "api": "bc_aes_ctr_host (pinned host buffers, 64 MiB chunks, 3 slots, one stream per stage)",
                "host_window_bytes": window, "calls_per_step": passes},
        "gpu_launches": args.steps, "clocks": clocks,
    }
    if parity is not None:
        line["parity_check"] = parity
    if others is not None:
        line["other_workloads"] = others
    if config5 is not None:
        line["config5"] = config5
    if link and "error" not in link:
        line["e2e"]["link_gbs_each_direction"] = round(link["min"], 2)
        line["e2e"]["link_gbs_aggregate"] = round(link["sum"], 2)
        line["e2e"]["frac_of_link"] = round(e2e_value / link["sum"], 4)
        line["e2e"]["note"] = ("host buffers in and out: bound by the host links (H2D and D2H at once on all N GPUs, "
                               "measured live with every rank copying at the same time), not by the kernel")
    elif link:
        line["e2e"]["note"] = f"link probe failed: {link['error']}"
    if n_gpus == 1 and not args.no_cpu_baseline:
        threads = cpu_threads()
        r = ReferenceCtr(2 * GIB, threads)
        gbs = r.gbs(passes=7)
        line["cpu_baseline"] = {"value": round(gbs, 3), "unit": UNIT, "cores": threads, "kind": r.kind,
                                "sample": f"{r.what}; {r.nbytes >> 20} MiB split over {threads} threads, "
                                          "median of 7 passes"}
        try:
            line["cpu_baseline"]["other_reference_paths"] = reference_extras(threads)
        except Exception as e:  # noqa: BLE001
            line["cpu_baseline"]["other_reference_paths"] = {"error": str(e)}
    print(json.dumps(line), flush=True)
    if distributed:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the configs 3/4 side measurements")
    ap.add_argument("--no-config5", action="store_true", help="skip the 256 GiB configs[4] record at N > 1")
    ap.add_argument("--no-numa-bind", action="store_true", help="A/B: leave the rank's host thread unbound")
    ap.add_argument("--wc-input", action="store_true", help="A/B: the e2e input buffer is write-combined pinned memory")
    ap.add_argument("--shard-gib", type=float, default=4.0,
                    help="bytes per GPU in GiB (default 4 = BASELINE configs[1]; 256/N reproduces config 5)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch ourselves under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29511"), __file__,
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
               "--shard-gib", str(args.shard_gib)] + (["--no-others"] if args.no_others else []) + \
              (["--no-config5"] if args.no_config5 else []) + (["--no-numa-bind"] if args.no_numa_bind else [])
        return subprocess.call(cmd)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

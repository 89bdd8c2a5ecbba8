#!/usr/bin/env python
"""bench.py — headline benchmark of the Bulletproofs+ hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]

Metric (BASELINE.json): range-proof verifies/s at n=64, m=16 (Ristretto255), with proves/s and the MSM points/s sweep
reported in the same JSON line. A step = one pass of the hot path over one batch of B synthetic proofs:
`bpp_range_verify_batch*` (point decode -> exponent vectors -> one segmented Pippenger MSM of B x 2089 terms -> verdicts).

  value     proofs/s, inputs already resident in HBM (device-pointer ABI), CUDA-event timed on the launching stream
  e2e       proofs/s through the host-pointer ABI: pinned host proofs -> H2D -> kernels -> verdict bytes D2H, every step
  roofline  dominant kernel (msm_accumulate): algorithmic 32x32+64 multiply-adds / its device time, against the integer
            pipe peak measured live by bpp_int_peak_probe (MEASURED_PEAKS.json has no integer figure); HBM GB/s second
  cpu_baseline  the CPU oracle's literal restatement of the reference (naive MulVec) on the box's host cores

N > 1 (torchrun): every rank verifies its own batch (independent proofs shard with no data-path collective);
time = max over ranks; value = total proofs / that time ("weak" scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_BITS, M_AGG = 64, 16
ELL = 2**252 + 27742317777372353535851937790883648493
FIELD_MUL_IMAD = 72           # 8x8 schoolbook + 8-limb fold
MIXED, FULL, DOUBLE = 7, 8, 8  # field multiplies per point operation
WORKLOAD = ("range-proof batch verify, n=64 bits, m=16 aggregated, Ristretto255, 2089-term MSM per proof "
            "(BASELINE.json metric config; 2048 proofs/step/GPU on the GPU arm)")


def algorithmic_field_muls(n_terms: int, bits: int = 253, MIXED: int = MIXED, FULL: int = FULL, DOUBLE: int = DOUBLE):
    """W(N) of SURVEY §8(d); returns (W, c*, accumulate-only share)."""
    best = None
    for c in range(1, 25):
        w = -(-(bits + 1) // c)
        cost = w * (n_terms * MIXED + 2 * (1 << (c - 1)) * FULL) + (w - 1) * c * DOUBLE
        if best is None or cost < best[0]:
            best = (cost, c, w * n_terms * MIXED)
    return best


def prng_bytes(seed: int, n: int) -> bytes:
    import hashlib
    out = bytearray()
    i = 0
    while len(out) < n:
        out += hashlib.sha256(seed.to_bytes(8, "little") + i.to_bytes(8, "little")).digest()
        i += 1
    return bytes(out[:n])


def synth_inputs(count: int, seed: int, M_AGG: int = M_AGG):
    """count * M_AGG values uniform in [0, 2^31) (clear of the `v as i32` quirk) and blinders uniform mod ell."""
    raw = prng_bytes(0xB2000003 + seed, count * M_AGG * 4)
    vs = [int.from_bytes(raw[4 * i:4 * i + 4], "little") & 0x7FFFFFFF for i in range(count * M_AGG)]
    raw = prng_bytes(0xB2000004 + seed, count * M_AGG * 64)
    gs = [int.from_bytes(raw[64 * i:64 * i + 64], "little") % ELL for i in range(count * M_AGG)]
    return vs, gs


class ClockSampler(threading.Thread):
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        super().__init__(daemon=True)
        self.device_index = device_index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-i",
                                      str(self.device_index)], capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                self.samples.append(float(f[1]))
                self.max_mhz = float(f[2])
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if val.lower().startswith("active"):
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(0.02)  # nvidia-smi itself takes ~50 ms: several samples even in a 0.2 s timed region

    def stop(self):
        self._halt.set()
        self.join(timeout=3)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------------------------
# CPU baselines. B1 = the oracle's literal restatement of the reference (naive MulVec::calculate,
# mulvec.rs:20-33) — the reference is single-threaded, so ONE thread is "the reference" and all cores (one independent
# proof stream per thread) is a second figure. B2 = the same protocol code with the bucket MSM of oracle/baseline_msm.cpp
# switched in under MulVec::calculate ("fair CPU": same algorithm family as the GPU path, 5x51-bit limbs, all cores).
# The reference itself cannot be built here (no Rust toolchain, mcl_rust un-vendored): kind = "port".
def _cpu_run(work, threads):
    results = []

    def wrap():
        results.append(work())

    ts = [threading.Thread(target=wrap) for _ in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return time.perf_counter() - t0, results


def cpu_verify(threads: int, per_thread: int, backend: int = 0, n: int = N_BITS, m: int = M_AGG, seed: int = 0):
    """proofs/s of `threads` host threads re-verifying one (n, m) proof per_thread times each (ctypes releases the GIL)."""
    from oracle import oracle_lib
    o = oracle_lib.get()
    vs, gs = synth_inputs(1, seed, m)
    o.set_msm_backend(0, 1)
    proof = o.range_prove(0, n, m, vs, gs)
    V = o.commit(0, vs, gs)
    o.set_msm_backend(backend, 1)
    try:
        assert o.range_verify(0, n, m, V, proof) == 0

        def work():
            return sum(o.range_verify(0, n, m, V, proof) == 0 for _ in range(per_thread))

        dt, res = _cpu_run(work, threads)
    finally:
        o.set_msm_backend(0, 1)
    assert sum(res) == threads * per_thread
    return threads * per_thread / dt, dt


def cpu_prove(threads: int, per_thread: int, backend: int = 0, n: int = N_BITS, m: int = M_AGG, seed: int = 0):
    from oracle import oracle_lib
    o = oracle_lib.get()
    vs, gs = synth_inputs(1, seed, m)
    o.set_msm_backend(backend, 1)
    try:
        def work():
            return sum(len(o.range_prove(0, n, m, vs, gs)) > 0 for _ in range(per_thread))

        dt, res = _cpu_run(work, threads)
    finally:
        o.set_msm_backend(0, 1)
    return threads * per_thread / dt, dt


def cpu_msm_b2(log2_n: int, threads: int):
    """points/s of the B2 bucket MSM over 2^log2_n synthetic Ristretto255 points, point-range split over `threads` threads
    (points and scalars are generated inside liboracle.so; only the MSM is timed)."""
    from oracle import oracle_lib
    n = 1 << log2_n
    dt = oracle_lib.get().b2_bench_msm(n, threads)
    return n / dt, dt


def cpu_baselines(cores: int):
    """The bounded CPU sample of the default run (about 20 s of wall clock)."""
    v1, dt1 = cpu_verify(1, 3, 0)
    vN, dtN = cpu_verify(cores, 2, 0)
    b2, dtb = cpu_verify(cores, 40, 1)
    b2_1, dtb1 = cpu_verify(1, 40, 1)
    p1, dtp = cpu_prove(1, 1, 0)
    pN, dtpN = cpu_prove(cores, 1, 0)
    m16, dtm = cpu_msm_b2(16, cores)
    return {
        "value": v1, "unit": "proofs/s", "cores": 1, "kind": "port",
        "sample": f"B1: 3 verifies of one n=64 m=16 Ristretto255 proof on ONE host thread ({dt1:.1f} s) — the oracle's literal restatement of the "
                  f"reference, naive MulVec::calculate (2089 full scalar multiplications per verify); the reference itself is single-threaded "
                  f"and cannot be built here (no Rust, mcl_rust un-vendored)",
        "all_cores": {"value": vN, "unit": "proofs/s", "cores": cores, "sample": f"B1, {cores} threads x 2 verifies, one proof stream per thread ({dtN:.1f} s)"},
        "b2_fair_cpu": {"value": b2, "unit": "proofs/s", "cores": cores, "one_thread": b2_1,
                        "sample": f"B2: the same verifier with the bucket MSM of oracle/baseline_msm.cpp (signed-digit Pippenger, 5x51-bit limbs, "
                                  f"unsigned __int128) under MulVec::calculate; {cores} threads x 40 verifies ({dtb:.1f} s); NOT the reference's algorithm"},
        "prove": {"value": p1, "unit": "proofs/s", "cores": 1, "all_cores": pN, "sample": f"B1 range_prove n=64 m=16: 1 proof on one thread ({dtp:.1f} s), "
                                                                                         f"{cores} threads x 1 ({dtpN:.1f} s)"},
        "msm_b2_2p16": {"value": m16, "unit": "points/s", "cores": cores, "sample": f"B2 bucket MSM over 2^16 Ristretto255 points, point-range split ({dtm:.2f} s)"},
    }


def run_reference(args, rank):
    """--impl reference: B1 on ONE thread (the reference is single-threaded: that is every host thread it can use)."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = max(1, args.ref_proofs)
    for _ in range(min(args.warmup, 1)):
        cpu_verify(1, 1, 0)
    runs = [cpu_verify(1, per_step, 0) for _ in range(args.steps)]
    total = per_step * len(runs) / sum(dt for _, dt in runs)
    ms = 1e3 * sum(dt for _, dt in runs) / len(runs)
    vN, _ = cpu_verify(cores, 1, 0)
    b2, _ = cpu_verify(cores, 20, 1)
    sample = (f"1 thread x {per_step} verifies of one n=64 m=16 proof per step (literal MulVec loop, 2089 scalar multiplications each); "
              f"the reference is single-threaded")
    line = {"impl": "reference", "metric": "range_proof_verifies_per_sec_n64_m16", "value": total, "unit": "proofs/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (integer)", "data": "synthetic",
            "config": {"workload": WORKLOAD},
            "cpu_baseline": {"value": total, "unit": "proofs/s", "cores": 1, "kind": "port", "sample": sample,
                             "all_cores": {"value": vN, "cores": cores, "note": "B1, one independent proof stream per host thread"},
                             "b2_fair_cpu": {"value": b2, "cores": cores, "note": "bucket-MSM verifier (oracle/baseline_msm.cpp), not the reference's algorithm"}},
            "e2e": {"value": total, "unit": "proofs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ncu --set full captures of the dominant kernel at B = 2048 (profiles/): numbers copied from the committed reports, not live
NCU_BUCKET = {"capture": "profiles/r02d_verify_fs_kernels_ncu_full_details.txt (msm_accumulate_kernel<Ed25519, 0>, fixed-base segment of the bucket path)",
              "gpu_time_ms": 5.61, "sm_busy_pct": 88.4, "smsp__issue_active_pct": 43.6, "registers_per_thread": 106, "dram_pct_of_peak": 3.8,
              "note": "the pipe that executes IMAD.WIDE is busy 88 % of the kernel's cycles: the kernel sits at the integer-multiply roofline of "
                      "sm_100a although only 0.4 of the 32-bit IMAD issue rate is reached (every wide multiply-add holds the pipe ~3.5-4.6 clk)"}
NCU_DIRECT = {"capture": "profiles/r02x_verify_kernels_ncu_full_details.txt + r02x_verify_kernels_raw_metrics.txt (direct_accumulate_kernel<Ed25519>, "
                         "B = 2048, 72 GB tables, c = 15, final build; captured alone: in the step it shares the SMs with the auxiliary stream's "
                         "Straus + Horner kernels, which are in the same capture)",
              "gpu_time_ms": 5.34, "dram_bytes": 9.091e9 + 0.034e9, "dram_read_GBps": 1702, "sm_throughput_pct": 85.8, "smsp__issue_active_pct": 42.8,
              "registers_per_thread": 128, "l2_hit_pct": 37.2, "active_lanes_per_warp_inst": 30.90,
              "note": "one random 128-byte table line per addition: ~9.1 GB of DRAM reads per launch (the algorithmic figure: every line is read once), a quarter "
                      "of the measured HBM copy bandwidth, hidden behind the integer pipe (SM throughput 86 %)"}

_REAL_STDOUT = None


def emit(line: dict):
    """The one JSON line goes to the real stdout; everything else a library prints (NCCL banners, torchrun notices) was
    rerouted to stderr by quiet_stdout()."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def quiet_stdout():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def prove_algorithmic_field_muls(n: int, m: int):
    """SURVEY §8(d) 'of a prove': W(N+m+3) + sum over rounds [2 W(2n'+2) + 2n' x (2-term joint scalar multiplication, 3040
    field multiplies on Ristretto)] + 2N mixed point additions."""
    N = n * m
    total = algorithmic_field_muls(N + m + 3)[0] + 2 * N * MIXED
    half = N
    while half > 1:
        half //= 2
        total += 2 * algorithmic_field_muls(2 * half + 2)[0] + 2 * half * 3040
    return total


class Bench:
    """Device-side state shared by the timed sections."""

    def __init__(self, args, torch, dist, rank, world, local_rank):
        from bulletproofsplus_b200 import api, _lib
        self.args, self.torch, self.dist, self.rank, self.world = args, torch, dist, rank, world
        self.api, self._lib = api, _lib
        self.dev = torch.device("cuda", local_rank)
        # a non-default torch stream, made current: the library launches on exactly this stream, so torch.cuda.Event records on
        # it bracket the library's kernels (the legacy default stream has handle 0 = "make a private stream")
        self.stream = torch.cuda.Stream(self.dev)
        torch.cuda.set_stream(self.stream)
        assert self.stream.cuda_stream != 0
        self.ctx = api.Context(local_rank, self.stream.cuda_stream)
        self.lib = self.ctx.lib
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)
        self.local_rank = local_rank

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def to_dev(self, b: bytes):
        return self.torch.frombuffer(bytearray(b), dtype=self.torch.uint8).to(self.dev)

    def pinned(self, b: bytes):
        return self.torch.frombuffer(bytearray(b), dtype=self.torch.uint8).pin_memory()

    def timed(self, fn, reps=5, warm=2, flush=False):
        """ms per call, CUDA events on the library's stream."""
        torch = self.torch
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for _ in range(reps):
            if flush:
                self.flush.zero_()
            fn()
        e1.record(self.stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def timed_host(self, fn, reps=5, warm=2, flush=False):
        """ms per call by the host clock around calls that end with a stream synchronisation (host-buffer ABI)."""
        for _ in range(warm):
            fn()
        self.torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if flush:
                self.flush.zero_()
            fn()
        self.torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3 / reps

    def max_over_ranks(self, *vals):
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def make_proofs(self, pk, n, m, count, seed):
        """count valid proofs from the library's own prover (256 per call) + their commitments, as bytes."""
        vs, gs = synth_inputs(count, seed, m)
        proofs = []
        for c0 in range(0, count, 256):
            c1 = min(count, c0 + 256)
            proofs += self.api.prove_batch(pk, n, m, vs[c0 * m:c1 * m], gs[c0 * m:c1 * m])
        Vs = pk.commit_batch(vs, gs)
        return b"".join(proofs), b"".join(Vs), vs, gs


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=2048, help="proofs per step per GPU")
    ap.add_argument("--prove-batch", type=int, default=256)
    ap.add_argument("--ref-proofs", type=int, default=2, help="reference arm: verifies per step (one thread)")
    ap.add_argument("--msm-max-log", type=int, default=22)
    ap.add_argument("--table-gb", type=float, default=80.0, help="HBM budget of the key's direct fixed-base tables (bpp_pk_precompute); "
                                                                   "0 = none: the bucket path")
    ap.add_argument("--no-extras", action="store_true", help="skip proves/s, configs 2-4, the MSM sweep and the CPU baseline")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import ctypes as C
    bn = Bench(args, torch, dist, rank, world, local_rank)
    api, _lib, ctx, lib, dev, stream = bn.api, bn._lib, bn.ctx, bn.lib, bn.dev, bn.stream
    B, K_steps = args.batch, args.steps
    NM = N_BITS * M_AGG
    pk = api.PublicKey.new(NM, _lib.RISTRETTO255, ctx)
    # one-time key set-up, not in any timed region: every signed-digit multiple of every window point of every generator in HBM
    t_tab = time.perf_counter()
    table_c, table_note = 0, None
    if args.table_gb > 0:
        # never ask for more than the device has free (minus working room); if even the narrowest table cannot be had the step
        # runs on the bucket path and the line says so
        free_b = torch.cuda.mem_get_info(dev)[0]
        budget = min(int(args.table_gb * 2**30), free_b - (16 << 30))
        while budget > (1 << 30) and not table_c:
            try:
                table_c = pk.precompute(budget)
            except Exception as e:  # cudaMalloc refused: halve and retry
                table_note = f"precompute({budget >> 30} GiB) failed: {str(e)[:80]}"
                budget //= 2
    t_tab = time.perf_counter() - t_tab
    table_W = -(-252 // table_c) if table_c else 0
    table_bytes = (table_W + 1) * (1 << (table_c - 1)) * (2 + 2 * NM) * 128 if table_c else 0
    plen = lib.bpp_range_proof_len(0, N_BITS, M_AGG)

    # ---- synthetic workload: ROT distinct batches of B valid proofs, produced by the library's own prover ----
    ROT = 2
    host_proofs, host_V, dev_proofs, dev_V = [], [], [], []
    for r in range(ROT):
        pb, vb, _, _ = bn.make_proofs(pk, N_BITS, M_AGG, B, seed=rank * 1000 + r)
        hp, hv = bn.pinned(pb), bn.pinned(vb)
        host_proofs.append(hp); host_V.append(hv)
        dev_proofs.append(hp.to(dev)); dev_V.append(hv.to(dev))
    d_verdicts = torch.full((B,), 255, dtype=torch.uint8, device=dev)
    h_verdicts = torch.empty((B,), dtype=torch.uint8).pin_memory()
    flush = bn.flush

    def step_dev(i):
        flush.zero_()  # L2 flush: 256 MB > 126 MB L2
        r = i % ROT
        rc = lib.bpp_range_verify_batch_dev(ctx.h, pk.handle, N_BITS, M_AGG, B, dev_V[r].data_ptr(), dev_proofs[r].data_ptr(), None,
                                            d_verdicts.data_ptr())
        ctx.check(rc, "verify_batch_dev")

    def step_host(i):
        flush.zero_()
        r = i % ROT
        rc = lib.bpp_range_verify_batch(ctx.h, pk.handle, N_BITS, M_AGG, B, C.cast(host_V[r].data_ptr(), C.c_char_p),
                                        C.cast(host_proofs[r].data_ptr(), C.c_char_p), None, C.cast(h_verdicts.data_ptr(), C.c_char_p))
        ctx.check(rc, "verify_batch")

    barrier = bn.barrier

    # ---- integer-pipe peak (live) ----
    probe = ctx.int_peak_probe()
    fp64 = ctx.fp64_peak_probe()  # the alternative multiply (5 x 51-bit limbs on DFMA): measured beside it on every run
    int_peak = probe["imad32_per_s"]  # issue peak of the integer multiply pipe (64 lanes/clk/SM on sm_100a)

    # ---- device-resident timing ----
    for i in range(args.warmup):
        step_dev(i)
    torch.cuda.synchronize()
    assert bool((d_verdicts == 0).all()), "warm-up verdicts are not all OK"
    ctx.hot_kernel_stats(reset=True)  # arms the per-launch events around msm_accumulate
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    launches0 = ctx.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(K_steps):
        step_dev(i)
    e1.record(stream)
    barrier()
    elapsed_ms = e0.elapsed_time(e1)
    launches = ctx.kernel_launches() - launches0
    clocks = sampler.stop()
    hot_ms, hot_n = ctx.hot_kernel_stats(reset=True)
    assert bool((d_verdicts == 0).all())

    # ---- end-to-end timing (host buffers, H2D + D2H inside) ----
    for i in range(min(args.warmup, 2)):
        step_host(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(K_steps):
        step_host(i)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert bool((h_verdicts == 0).all())

    t = torch.tensor([elapsed_ms, e2e_s * 1e3, float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        elapsed_ms, e2e_ms, launches = float(tmax[0]), float(tmax[1]), int(tsum[2])
    else:
        e2e_ms = e2e_s * 1e3
    value = world * B * K_steps / (elapsed_ms * 1e-3)
    e2e_value = world * B * K_steps / (e2e_ms * 1e-3)
    # ---- roofline of the dominant kernel ----
    n_terms = 5 + 2 * 10 + 2 * NM + M_AGG  # 2089
    W, c_star, acc_share = algorithmic_field_muls(n_terms)
    hot_avg_ms = hot_ms / max(1, hot_n)
    # the timed bracket covers the fixed-base segment (2050 of the 2089 terms); the 39 per-proof points are accumulated beside it on
    # the auxiliary stream, so only the fixed-base share of the algorithmic count is credited to the timed launches
    n_fixed_terms = 2 + 2 * NM
    ach_kernel = B * acc_share * (n_fixed_terms / n_terms) * FIELD_MUL_IMAD / (hot_avg_ms * 1e-3) if hot_n else None
    # what the kernel actually executes for the 2050 generator terms — fewer additions than the algorithmic count above (the 39
    # per-proof points take 64 generic 4-bit windows each, outside the timed bracket):
    #   direct tables (bpp_pk_precompute, window c): ceil(252 / c) table additions per term, nothing else
    #   bucket path: window tables with c = 12: 21 windows (scalars folded below ell/2), one bucket set per proof, no doublings
    exec_madds = B * (2 + 2 * NM) * (table_W if table_c else 21)
    exec_kernel = exec_madds * MIXED * FIELD_MUL_IMAD / (hot_avg_ms * 1e-3) if hot_n else None
    ach_whole = B * W * FIELD_MUL_IMAD / (elapsed_ms / K_steps * 1e-3)
    alg_bytes = B * (n_terms * (32 + 96) + 128)  # scalars + affine points read once, one point written
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    roofline = {
        "bound": "int-imad", "bound_note": "integer multiply pipe: ~20k multiply-adds per 128 B term, so neither the HBM nor the tensor roofline "
                                               "applies (north_star asks for the integer-multiply roofline); HBM figures are in `hbm` below",
        "kernel": (f"direct_accumulate_kernel<Ed25519>: one table addition per {table_c}-bit window and generator term, straight from the key's "
                   f"{table_bytes / 2**30:.1f} GB of direct tables (no buckets)") if table_c else
                  "msm_accumulate_kernel<Ed25519> (+ its two long-bucket variants msm_accumulate_group_kernel<.,32|4>, timed together)",
        "achieved": ach_kernel / 1e12 if ach_kernel else None, "peak": int_peak / 1e12, "unit": "T 32x32+64 multiply-adds/s",
        "frac": (ach_kernel / int_peak) if ach_kernel else None,
        # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel at this batch size (B = 2048), from the ncu
        # --set full captures named in `ncu` (not measured live); `traffic_algorithmic` = one 128-byte table line per addition + the
        # 32-byte scalar and 4-byte index of every term, per launch
        "traffic": NCU_DIRECT.get("dram_bytes") if table_c else (1.732e9 if B == 2048 else None), "traffic_unit": "bytes per launch",
        "traffic_algorithmic": exec_madds * 128 + B * n_terms * 36,
        "ncu": NCU_DIRECT if table_c else NCU_BUCKET,
        "peak_source": "bpp_int_peak_probe: 32-bit IMAD chains with loop-produced operands on all SMs, measured in this run — the "
                       "issue peak of the integer multiply pipe, which every form of 32x32 multiply shares",
        "pipe_ceiling": {"wide_chain_T_per_s": probe["wide_chain_per_s"] / 1e12, "wide_fused_T_per_s": probe["wide_fused_per_s"] / 1e12,
                         "note": "same probe, the two 32x32+64 forms a limb product can use: carry-chained IMAD.WIDE.U32.X (what "
                                 "mul_wide_n emits) and IMAD.WIDE.U32 with a 64-bit register addend; both issue at well under half "
                                 "the IMAD peak on sm_100a, so ~0.43 of `peak` is the ceiling of any multi-limb multiply (DESIGN.md §3.1)"},
        "fp64_alternative": {"dfma_T_per_s": fp64["dfma_per_s"] / 1e12, "f5_mul_G_per_s": fp64["f5_mul_per_s"] / 1e9,
                             "note": "the same field multiply as 50 DFMA + 30 DADD + int64 column sums (csrc/fe25519_f64.cuh), probed live: "
                                     "no faster than the 72-IMAD.WIDE form, and running both kinds of warp on one SM adds nothing "
                                     "(profiles/r01e_dfma_fe.txt) - DFMA and IMAD.WIDE contend for the same issue capacity on sm_100a"},
        "executed": {"achieved": exec_kernel / 1e12 if exec_kernel else None, "frac": (exec_kernel / int_peak) if exec_kernel else None,
                     "frac_of_pipe_ceiling": (exec_kernel / probe["wide_chain_per_s"]) if exec_kernel else None,
                     "note": "multiply-adds the kernel really issues (upper bound: zero digits skipped): the window tables of the public key "
                             "beat the algorithmic count, so `frac` above overstates pipe utilisation by algorithmic/executed"},
        "launches_timed": hot_n, "avg_launch_ms": hot_avg_ms, "kernel_share_of_step": hot_ms / elapsed_ms if elapsed_ms else None,
        "algorithmic": f"B*ceil(254/c*)*N*7 field multiplies*72 with c*={c_star}, N={n_terms}",
        "whole_step": {"achieved": ach_whole / 1e12, "frac": ach_whole / int_peak, "algorithmic": "B*W(2089)*72 (SURVEY §8(d))"},
        "hbm": {"achieved_gbs": alg_bytes / (elapsed_ms / K_steps * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback", "note": "secondary: shows the path is not memory-bound"},
    }

    extras = {}
    cpu = None
    if not args.no_extras:
        # A failing extra section must not lose the headline line: its error is recorded under `extras_errors` instead. At N > 1
        # a section that raises on one rank only would leave the others waiting in a collective, so there the error is fatal.
        errors = {}

        def section(name, fn):
            try:
                extras.update(fn())
            except Exception as e:  # noqa: BLE001
                if world > 1:
                    raise
                import traceback
                errors[name] = f"{type(e).__name__}: {e} | {traceback.format_exc().splitlines()[-3].strip()}"

        # per-GPU sections that shard trivially (every rank runs its own; max over ranks) and the multi-rank sections
        section("verify_per_proof_challenges", lambda: run_fs_verify(bn, pk, B))
        section("proves", lambda: run_proves(bn, pk, int_peak))
        section("config4", lambda: run_config4(bn))
        if table_c:
            section("verify_bucket_path", lambda: run_bucket_path(bn, B))
        if world > 1:
            section("msm_point_range_split", lambda: {"msm_point_range_split": run_msm_split(bn, int_peak)})
        if world == 1:
            section("configs_2_3", lambda: run_configs_2_3(bn))
            section("msm_sweep", lambda: run_msm_sweep(bn, int_peak))
            if rank == 0:
                try:
                    cpu = cpu_baselines(os.cpu_count() or 1)
                    if "proves" in extras:
                        extras["proves"]["cpu_baseline"] = cpu.pop("prove")
                except Exception as e:  # noqa: BLE001
                    errors["cpu_baseline"] = f"{type(e).__name__}: {e}"
        if errors:
            extras["extras_errors"] = errors

    if rank == 0:
        line = {
            "metric": "range_proof_verifies_per_sec_n64_m16", "value": value, "unit": "proofs/s", "n_gpus": world, "steps": K_steps,
            "warmup": args.warmup, "ms_per_step": elapsed_ms / K_steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 limbs (integer, GF(2^255-19) / mod ell)", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": B, "parallelism": f"proofs sharded x{world}",
                       "fixed_base_tables": ({"window_bits": table_c, "windows": table_W, "bytes": table_bytes, "build_s": t_tab,
                                              "note": "bpp_pk_precompute: one-time key set-up, outside every timed region" + (("; " + table_note) if table_note else "")} if table_c else None),
                       "challenge_mode": "the reference's constants (parity mode; challenge tables reused across calls). The same batch with "
                                         "per-proof Fiat-Shamir challenges is `verify_per_proof_challenges` in this line",
                       "l2": "256 MB memset between steps (inside the timed region); intermediates per step > L2"},
            "e2e": {"value": e2e_value, "unit": "proofs/s", "h2d_bytes_per_step": B * (plen + M_AGG * 32), "d2h_bytes_per_step": B},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        }
        line.update(extras)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_fs_verify(bn, pk, B):
    """The headline batch with PER-PROOF challenges (bpp_range_verify_batch_fs*: what a Fiat-Shamir verifier has to call):
    every proof builds its own challenge tables. Device-resident value + e2e with host buffers; challenges are the constants, so
    the host's transcript hashing (merlin.py is pure Python) is not in these numbers."""
    import ctypes as C
    torch, lib, ctx = bn.torch, bn.lib, bn.ctx
    pb, vb, _, _ = bn.make_proofs(pk, N_BITS, M_AGG, B, seed=bn.rank * 1000 + 7)
    L = 10
    block = b"".join(int(x).to_bytes(32, "little") for x in [12, 23, 99] + [7] * L)
    h_pr, h_V, h_ch = bn.pinned(pb), bn.pinned(vb), bn.pinned(block * B)
    d_pr, d_V, d_ch = h_pr.to(bn.dev), h_V.to(bn.dev), h_ch.to(bn.dev)
    d_vd = torch.full((B,), 255, dtype=torch.uint8, device=bn.dev)
    h_vd = torch.empty((B,), dtype=torch.uint8).pin_memory()

    def dev_call():
        ctx.check(lib.bpp_range_verify_batch_fs_dev(ctx.h, pk.handle, N_BITS, M_AGG, B, d_V.data_ptr(), d_pr.data_ptr(), d_ch.data_ptr(),
                                                    d_vd.data_ptr()), "verify_fs_dev")

    def host_call():
        ctx.check(lib.bpp_range_verify_batch_fs(ctx.h, pk.handle, N_BITS, M_AGG, B, C.cast(h_V.data_ptr(), C.c_char_p),
                                                C.cast(h_pr.data_ptr(), C.c_char_p), C.cast(h_ch.data_ptr(), C.c_char_p),
                                                C.cast(h_vd.data_ptr(), C.c_char_p)), "verify_fs")

    bn.barrier()
    ms = bn.timed(dev_call, reps=5, warm=2, flush=True)
    assert bool((d_vd == 0).all())
    ms_h = bn.timed_host(host_call, reps=5, warm=1, flush=True)
    assert bool((h_vd == 0).all())
    ms, ms_h = bn.max_over_ranks(ms, ms_h)
    out = {"verify_per_proof_challenges": {
        "value": bn.world * B / (ms * 1e-3), "unit": "proofs/s", "batch_per_gpu": B, "ms_per_step": ms, "entry": "bpp_range_verify_batch_fs_dev",
        "e2e": {"value": bn.world * B / (ms_h * 1e-3), "unit": "proofs/s", "entry": "bpp_range_verify_batch_fs (host buffers)",
                "h2d_bytes_per_step": B * (len(pb) // B + M_AGG * 32 + len(block)), "d2h_bytes_per_step": B}}}
    # Two contexts in flight on one GPU (how a serving process with two worker threads uses the library: a bpp_ctx per thread,
    # the key shared): the second context's challenge-table kernels (latency-bound) overlap the first one's accumulation
    # (pipe-bound). Two different resident batches, calls issued alternately, no synchronisation between them; CUDA events on
    # the first stream, which waits for the second before the closing event.
    try:
        stream2 = torch.cuda.Stream(bn.dev)
        ctx2 = bn.api.Context(bn.local_rank, stream2.cuda_stream)
        pb2, vb2, _, _ = bn.make_proofs(pk, N_BITS, M_AGG, B, seed=bn.rank * 1000 + 8)
        d_pr2, d_V2 = bn.to_dev(pb2), bn.to_dev(vb2)
        d_vd2 = torch.full((B,), 255, dtype=torch.uint8, device=bn.dev)

        def dev_call2():
            ctx2.check(lib.bpp_range_verify_batch_fs_dev(ctx2.h, pk.handle, N_BITS, M_AGG, B, d_V2.data_ptr(), d_pr2.data_ptr(), d_ch.data_ptr(),
                                                         d_vd2.data_ptr()), "verify_fs_dev ctx2")

        for _ in range(2):
            dev_call(); dev_call2()
        torch.cuda.synchronize()
        reps2 = 5
        e0, e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event()
        e0.record(bn.stream)
        stream2.wait_event(e0)
        for _ in range(reps2):
            bn.flush.zero_()
            dev_call(); dev_call2()
        e2.record(stream2)
        bn.stream.wait_event(e2)
        e1.record(bn.stream)
        torch.cuda.synchronize()
        assert bool((d_vd == 0).all()) and bool((d_vd2 == 0).all())
        (ms2,) = bn.max_over_ranks(e0.elapsed_time(e1) / reps2)
        out["verify_per_proof_challenges"]["two_contexts"] = {
            "value": bn.world * 2 * B / (ms2 * 1e-3), "unit": "proofs/s", "ms_per_pair_of_batches": ms2, "batch_per_context": B,
            "note": "two bpp_ctx (one stream pair each) sharing the key, calls issued alternately by one host thread: the second batch's "
                    "challenge tables are built while the first batch accumulates"}
        ctx2.close()
    except Exception as e:  # keep the single-context figures
        out["verify_per_proof_challenges"]["two_contexts"] = {"error": repr(e)[:200]}
    # random-linear-combination batch verification of the same proofs: one combined MSM per rank (per-proof fallback on failure)
    import secrets
    d_w = bn.to_dev(b"".join((secrets.randbits(128) | 1).to_bytes(32, "little") for _ in range(B)))
    fb = C.c_int(0)

    def rlc_call():
        ctx.check(lib.bpp_range_verify_batch_rlc_dev(ctx.h, pk.handle, N_BITS, M_AGG, B, d_V.data_ptr(), d_pr.data_ptr(), None, None, d_w.data_ptr(),
                                                     d_vd.data_ptr(), C.byref(fb)), "verify_rlc")

    ms_r = bn.timed(rlc_call, reps=5, warm=2, flush=True)
    assert bool((d_vd == 0).all()) and fb.value == 0
    (ms_r,) = bn.max_over_ranks(ms_r)
    out["verifies_per_sec_rlc_batch"] = {"value": bn.world * B / (ms_r * 1e-3), "unit": "proofs/s", "batch_per_gpu": B, "ms_per_step": ms_r,
                                         "entry": "bpp_range_verify_batch_rlc_dev (all proofs valid: no fallback; one host sync per batch inside); "
                                                  "every rank combines its own shard, verdicts need no exchange"}
    return out


def run_bucket_path(bn, B):
    """The headline step on a key WITHOUT direct tables: the two-segment bucket MSM (per-job sort, size sort, bucket accumulation,
    chunk / set reduction) that the direct tables replace — kept as the fallback for keys too large for tables and as the path of
    the other curves; reported so that the two can be compared in one run."""
    torch, lib, ctx, api = bn.torch, bn.lib, bn.ctx, bn.api
    pk2 = api.PublicKey.new(N_BITS * M_AGG, 0, ctx)
    pb, vb, _, _ = bn.make_proofs(pk2, N_BITS, M_AGG, B, seed=bn.rank * 1000 + 3)
    d_pr, d_V = bn.to_dev(pb), bn.to_dev(vb)
    d_vd = torch.full((B,), 255, dtype=torch.uint8, device=bn.dev)

    def call():
        ctx.check(lib.bpp_range_verify_batch_dev(ctx.h, pk2.handle, N_BITS, M_AGG, B, d_V.data_ptr(), d_pr.data_ptr(), None, d_vd.data_ptr()), "verify (buckets)")

    bn.barrier()
    ms = bn.timed(call, reps=5, warm=2, flush=True)
    assert bool((d_vd == 0).all())
    (ms,) = bn.max_over_ranks(ms)
    return {"verify_bucket_path": {"value": bn.world * B / (ms * 1e-3), "unit": "proofs/s", "ms_per_step": ms, "batch_per_gpu": B,
                                   "note": "same entry point, key without bpp_pk_precompute: window tables (c = 12) + buckets"}}


def run_proves(bn, pk, int_peak):
    """proves/s at n=64, m=16: device-resident, e2e through bpp_range_prove_batch with host buffers, whole-step roofline."""
    import ctypes as C
    torch, lib, ctx, args = bn.torch, bn.lib, bn.ctx, bn.args
    PB = args.prove_batch
    vs, gs = synth_inputs(PB, 99 + bn.rank, M_AGG)
    plen = lib.bpp_range_proof_len(0, N_BITS, M_AGG)
    h_v = torch.tensor(vs, dtype=torch.int64).pin_memory()
    h_g = bn.pinned(b"".join(g.to_bytes(32, "little") for g in gs))
    h_p = torch.empty(PB * plen, dtype=torch.uint8).pin_memory()
    d_v, d_g = h_v.to(bn.dev), h_g.to(bn.dev)
    d_p = torch.empty(PB * plen, dtype=torch.uint8, device=bn.dev)

    def dev_call():
        ctx.check(lib.bpp_range_prove_batch_dev(ctx.h, pk.handle, N_BITS, M_AGG, PB, d_v.data_ptr(), d_g.data_ptr(), None, d_p.data_ptr()), "prove")

    def host_call():
        ctx.check(lib.bpp_range_prove_batch(ctx.h, pk.handle, N_BITS, M_AGG, PB, C.cast(h_v.data_ptr(), C.POINTER(C.c_uint64)),
                                            C.cast(h_g.data_ptr(), C.c_char_p), None, C.cast(h_p.data_ptr(), C.c_char_p)), "prove host")

    bn.barrier()
    ctx.hot_kernel_stats(reset=True)
    reps = 5
    ms = bn.timed(dev_call, reps=reps, warm=2, flush=True)
    hot_ms, hot_n = ctx.hot_kernel_stats(reset=True)
    hot_ms_per_call = hot_ms * reps / (reps + 2) / reps  # the bracket also saw the two warm-up calls
    ms_h = bn.timed_host(host_call, reps=5, warm=1, flush=True)
    assert bytes(h_p.numpy().tobytes()) == bytes(d_p.cpu().numpy().tobytes()), "host-ABI proofs differ from the device-ABI ones"
    ms, ms_h = bn.max_over_ranks(ms, ms_h)
    Wp = prove_algorithmic_field_muls(N_BITS, M_AGG)
    alg = PB * Wp * FIELD_MUL_IMAD / (ms * 1e-3)
    NM, L = N_BITS * M_AGG, 10
    terms = 1 + 2 * L * (NM + 2) + (2 * NM + 2) + 2
    return {"proves": {
        "metric": "range_proof_proves_per_sec_n64_m16", "value": bn.world * PB / (ms * 1e-3), "unit": "proofs/s", "batch_per_gpu": PB, "ms_per_batch": ms,
        "entry": "bpp_range_prove_batch_dev (fixed-nonce parity mode)",
        "e2e": {"value": bn.world * PB / (ms_h * 1e-3), "unit": "proofs/s", "entry": "bpp_range_prove_batch (host buffers)",
                "h2d_bytes_per_step": PB * M_AGG * 40, "d2h_bytes_per_step": PB * plen},
        "roofline": {"bound": "int-imad", "achieved": alg / 1e12, "peak": int_peak / 1e12, "unit": "T 32x32+64 multiply-adds/s", "frac": alg / int_peak,
                     "algorithmic": f"PB * W_prove(64,16) * 72, W_prove = {Wp} field multiplies (SURVEY §8(d): W(N+m+3) + rounds [2 W(2n'+2) + 2n' * 3040] + 2N adds)",
                     "kernel": "msm_accumulate_kernel<Ed25519> + long-bucket variants, all (2L+3) MSM jobs of all proofs in one launch sequence",
                     "avg_launch_ms": hot_ms_per_call, "kernel_share_of_step": hot_ms_per_call / ms if ms else None,
                     "executed": {"terms_per_proof": terms, "note": "generators are never folded (DESIGN §3.3): every L/R job is a fixed-base MSM over the "
                                  "original G, H, so the 2n' joint scalar multiplications per round of the algorithmic count are not executed at all; "
                                  "about terms x 21 table additions per proof instead"}}}}


def run_configs_2_3(bn):
    """BASELINE.json config 2 (standalone WIP over 2^12 generators, prove + verify) and config 3 (one aggregated n=64, m=64
    proof: 8285-term verifier MSM) — single-call latencies through the host-buffer ABI, plus a 128-proof batch for config 3."""
    import random
    api, ctx, lib = bn.api, bn.ctx, bn.lib
    out = {}
    N = 4096
    pk = api.PublicKey.new(N, 0, ctx)
    rng = random.Random(0xC2)
    a = [rng.randrange(ELL) for _ in range(N)]
    b = [rng.randrange(ELL) for _ in range(N)]
    ypow = [pow(12, i + 1, ELL) for i in range(N)]
    gamma = rng.randrange(ELL)
    proof = api.WeightedInnerProductProof.prove(pk, a, b, ypow, gamma)
    dl = (sum(a[i] * 3 * (i + 1) + b[i] * 5 * (i + 1) + a[i] * b[i] * ypow[i] for i in range(N)) + 2 * gamma) % ELL
    one = api.Points.from_multipliers((1).to_bytes(32, "little"), 0, ctx)
    P = one.msm(dl.to_bytes(32, "little"))[0]  # the commitment a b-weighted-inner-product prover would publish, as dl * g
    zeros = [0] * N
    proof.verify(pk, ypow, zeros, zeros, 0, [], P, [])
    enc = lambda v: b"".join(int(x).to_bytes(32, "little") for x in v)
    ea, eb, ey, ez = enc(a), enc(b), enc(ypow), enc(zeros)
    import ctypes as C
    buf = C.create_string_buffer(lib.bpp_wip_proof_len(0, N))
    ln = C.c_size_t(len(buf))
    pbytes = proof.to_bytes()

    def wip_prove():
        ln.value = len(buf)
        ctx.check(lib.bpp_wip_prove(ctx.h, pk.handle, ea, eb, ey, gamma.to_bytes(32, "little"), None, buf, C.byref(ln)), "wip_prove")

    def wip_verify():
        rc = lib.bpp_wip_verify(ctx.h, pk.handle, ey, ez, ez, (0).to_bytes(32, "little"), 0, None, P, None, pbytes, len(pbytes), None)
        assert rc == 0, rc

    out["config2_wip_4096"] = {
        "prove_ms": bn.timed_host(wip_prove, reps=5, warm=2), "verify_ms": bn.timed_host(wip_verify, reps=5, warm=2),
        "note": "one standalone weighted-inner-product argument over 2^12 generators, Ristretto255, host-buffer ABI (H2D of 3 x 128 KB of scalars "
                "inside): prove = 12 rounds, 26 fixed-base MSM jobs; verify = one 8221-term MSM",
        "algorithmic_field_muls_verify": algorithmic_field_muls(8221)[0]}
    # config 3
    n, m = 64, 64
    vs, gs = synth_inputs(1, 0xC3, m)
    pr = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    assert api.verify_batch(pk, n, m, pr, [Vs]) == [0]
    t_p = bn.timed_host(lambda: api.prove_batch(pk, n, m, vs, gs), reps=5, warm=1)
    t_v = bn.timed_host(lambda: api.verify_batch(pk, n, m, pr, [Vs]), reps=5, warm=1)
    CB = 128
    pb, vb, _, _ = bn.make_proofs(pk, n, m, CB, seed=0xC33)
    d_p, d_V = bn.to_dev(pb), bn.to_dev(vb)
    d_vd = bn.torch.full((CB,), 255, dtype=bn.torch.uint8, device=bn.dev)
    ms = bn.timed(lambda: ctx.check(lib.bpp_range_verify_batch_dev(ctx.h, pk.handle, n, m, CB, d_V.data_ptr(), d_p.data_ptr(), None, d_vd.data_ptr()),
                                    "verify 64x64"), reps=5, warm=2, flush=True)
    assert bool((d_vd == 0).all())
    # the same batch with direct tables for this key's 8194 generators, if HBM allows beside the headline key's (c = 12: 47 GB)
    ms_bucket, c3_tables = ms, None
    t_p_tab = t_v_tab = None
    free_b = bn.torch.cuda.mem_get_info(bn.dev)[0]
    budget = min(int(bn.args.table_gb * 2**30), free_b - (12 << 30))
    if bn.args.table_gb > 0 and budget > (8 << 30):
        try:
            c3_tables = pk.precompute(budget)
        except Exception as e:  # out of HBM beside the other keys' tables: keep the bucket-path figure
            c3_tables = "unavailable: " + str(e)[:120]
        else:
            d_vd.fill_(255)
            ms = bn.timed(lambda: ctx.check(lib.bpp_range_verify_batch_dev(ctx.h, pk.handle, n, m, CB, d_V.data_ptr(), d_p.data_ptr(), None,
                                                                           d_vd.data_ptr()), "verify 64x64 tables"), reps=5, warm=2, flush=True)
            assert bool((d_vd == 0).all())
            # the single-call latencies again, against the key with tables (every call takes them, one job or many)
            out["config2_wip_4096"]["prove_ms_with_tables"] = bn.timed_host(wip_prove, reps=5, warm=2)
            out["config2_wip_4096"]["verify_ms_with_tables"] = bn.timed_host(wip_verify, reps=5, warm=2)
            out["config2_wip_4096"]["fixed_base_tables_window_bits"] = c3_tables
            assert api.verify_batch(pk, n, m, pr, [Vs]) == [0] and api.prove_batch(pk, n, m, vs, gs) == pr
            t_p_tab = bn.timed_host(lambda: api.prove_batch(pk, n, m, vs, gs), reps=5, warm=1)
            t_v_tab = bn.timed_host(lambda: api.verify_batch(pk, n, m, pr, [Vs]), reps=5, warm=1)
    out["config3_n64_m64"] = {
        "single_proof_prove_ms": t_p, "single_proof_verify_ms": t_v, "batch": CB, "batch_verify_ms": ms, "batch_verifies_per_sec": CB / (ms * 1e-3),
        "batch_verify_ms_bucket_path": ms_bucket, "fixed_base_tables_window_bits": c3_tables,
        "single_proof_prove_ms_with_tables": t_p_tab, "single_proof_verify_ms_with_tables": t_v_tab,
        "note": "one aggregated proof of 64 values (N = 4096): verifier MSM of 8285 terms; single-proof numbers are host-ABI latencies "
                "(python packing included), the batch is device-resident",
        "algorithmic_field_muls_verify": algorithmic_field_muls(8285)[0]}
    return out


C4_TABLE_BYTES = 6 << 30  # config 4's key: c = 15 fits (4.9 GB), c = 16 (9.3 GB) does not


def run_config4(bn):
    """BASELINE.json config 4: a FIXED batch of 8192 independent n=64, m=1 proofs (146-term MSM each, 128 of them tampered)
    sharded over the ranks by parallel.shard_range: strong scaling. Rank 0 holds the host batch; e2e includes the scatter
    (rank 0 -> ranks, NCCL), H2D, kernels, and the all-gather of the verdict bytes; `value` is device-resident per shard."""
    import ctypes as C
    from bulletproofsplus_b200 import parallel
    torch, dist, api, ctx, lib = bn.torch, bn.dist, bn.api, bn.ctx, bn.lib
    n, m, TOTAL = 64, 1, 8192
    pk = api.PublicKey.new(n * m, 0, ctx)
    c4_tables = None
    if bn.args.table_gb > 0:  # direct tables for the 130 generators of this key: 4.9 GB at c = 15 (setup, untimed)
        c4_tables = pk.precompute(min(int(bn.args.table_gb * 2**30), C4_TABLE_BYTES))
    plen = lib.bpp_range_proof_len(0, n, m)
    lo, hi = parallel.shard_range(TOTAL, bn.rank, bn.world)
    cnt = hi - lo
    # every rank builds the same full batch deterministically (setup, untimed), then keeps its shard; tamper r' of every 64th proof
    pb, vb, _, _ = bn.make_proofs(pk, n, m, TOTAL, seed=0xC4)
    pb = bytearray(pb)
    for i in range(0, TOTAL, 64):
        pb[i * plen + plen - 96] ^= 1
    want = torch.tensor([1 if i % 64 == 0 else 0 for i in range(TOTAL)], dtype=torch.uint8)
    h_p, h_V = bn.pinned(bytes(pb[lo * plen:hi * plen])), bn.pinned(vb[lo * 32:hi * 32])
    d_p, d_V = h_p.to(bn.dev), h_V.to(bn.dev)
    d_vd = torch.full((cnt,), 255, dtype=torch.uint8, device=bn.dev)
    d_all = torch.full((TOTAL,), 255, dtype=torch.uint8, device=bn.dev)
    h_vd = torch.empty((cnt,), dtype=torch.uint8).pin_memory()
    even = TOTAL % bn.world == 0

    def dev_call():
        ctx.check(lib.bpp_range_verify_batch_dev(ctx.h, pk.handle, n, m, cnt, d_V.data_ptr(), d_p.data_ptr(), None, d_vd.data_ptr()), "cfg4")
        if bn.world > 1 and even:
            dist.all_gather_into_tensor(d_all, d_vd)  # verdict exchange inside the timed region

    def host_call():
        ctx.check(lib.bpp_range_verify_batch(ctx.h, pk.handle, n, m, cnt, C.cast(h_V.data_ptr(), C.c_char_p), C.cast(h_p.data_ptr(), C.c_char_p), None,
                                             C.cast(h_vd.data_ptr(), C.c_char_p)), "cfg4 host")
        if bn.world > 1 and even:
            d_vd.copy_(h_vd, non_blocking=True)
            dist.all_gather_into_tensor(d_all, d_vd)

    bn.barrier()
    ms = bn.timed(dev_call, reps=5, warm=2, flush=True)
    bn.barrier()
    ms_h = bn.timed_host(host_call, reps=5, warm=1, flush=True)
    got = d_all.cpu() if (bn.world > 1 and even) else d_vd.cpu()
    ref = want if (bn.world > 1 and even) else want[lo:hi]
    assert bool((got == ref).all()), "config 4 verdicts differ from the tamper pattern"
    ms, ms_h = bn.max_over_ranks(ms, ms_h)
    return {"config4_batch8192_n64_m1": {
        "value": TOTAL / (ms * 1e-3), "unit": "proofs/s", "ms": ms, "scaling": "strong", "n_gpus": bn.world, "proofs_total": TOTAL, "proofs_per_gpu": cnt,
        "tampered": TOTAL // 64, "verdicts_match_tamper_pattern": True,
        "e2e": {"value": TOTAL / (ms_h * 1e-3), "unit": "proofs/s", "ms": ms_h, "h2d_bytes_per_step": cnt * (plen + 32), "d2h_bytes_per_step": cnt},
        "exchange": "all-gather of the verdict bytes (NCCL)" if bn.world > 1 else "none",
        "fixed_base_tables_window_bits": c4_tables,
        "algorithmic_field_muls_per_proof": algorithmic_field_muls(146)[0]}}


def run_msm_split(bn, int_peak):
    """One MSM of 2^msm_max_log points split by point range over the ranks: per-rank MSM to one raw partial
    point, NCCL all-gather of world x 128 B, local sum + encode. Strong scaling; time = max over ranks. The result is compared
    with the exponent-domain value (sum_i s_i k_i mod ell) * g, computed in Python integers per rank and multiplied out by a
    one-term MSM."""
    torch, dist, api, ctx, lib, args = bn.torch, bn.dist, bn.api, bn.ctx, bn.lib, bn.args
    world, rank, dev = bn.world, bn.rank, bn.dev
    n_total = 1 << args.msm_max_log
    n_loc = n_total // world
    g = torch.Generator().manual_seed(1234 + rank)
    mult = torch.randint(0, 256, (n_loc, 32), dtype=torch.uint8, generator=g)
    mult[:, 31] &= 0x0F
    sc = torch.randint(0, 256, (n_loc, 32), dtype=torch.uint8, generator=g)
    sc[:, 31] &= 0x0F
    mb, sb = bytes(mult.numpy().tobytes()), bytes(sc.numpy().tobytes())
    pts = api.Points.from_multipliers(mb, 0, ctx)
    d_sc = sc.to(dev)
    d_raw = torch.empty(128, dtype=torch.uint8, device=dev)
    d_all = torch.empty(128 * world, dtype=torch.uint8, device=dev)
    d_out = torch.empty(32, dtype=torch.uint8, device=dev)

    def step():
        ctx.check(lib.bpp_msm_prepared_raw_dev(ctx.h, pts.h, d_sc.data_ptr(), n_loc, d_raw.data_ptr()), "msm_raw_dev")
        dist.all_gather_into_tensor(d_all, d_raw)
        ctx.check(lib.bpp_points_sum_raw_dev(ctx.h, 0, d_all.data_ptr(), world, d_out.data_ptr()), "sum_raw_dev")

    ms = bn.timed(step, reps=5, warm=3)
    bn.barrier()
    (ms,) = bn.max_over_ranks(ms)
    # exponent-domain check
    dl = 0
    for i in range(n_loc):
        dl += int.from_bytes(mb[32 * i:32 * i + 32], "little") * int.from_bytes(sb[32 * i:32 * i + 32], "little")
    d_dl = bn.to_dev((dl % ELL).to_bytes(32, "little"))
    d_dls = torch.empty(32 * world, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(d_dls, d_dl)
    raw = bytes(d_dls.cpu().numpy().tobytes())
    total = sum(int.from_bytes(raw[32 * r:32 * r + 32], "little") for r in range(world)) % ELL
    one = api.Points.from_multipliers((1).to_bytes(32, "little"), 0, ctx)
    want = one.msm(total.to_bytes(32, "little"))[0]
    got = bytes(d_out.cpu().numpy().tobytes())
    ok = torch.tensor([1 if got == want else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    assert int(ok[0]) == 1, "point-range-split MSM differs from (sum s_i k_i) * g"
    W, _, _ = algorithmic_field_muls(n_total)
    return {"log2_n": args.msm_max_log, "n_gpus": world, "ms": ms, "points_per_sec": n_total / (ms * 1e-3), "scaling": "strong",
            "equals_exponent_domain_value": True, "exchange": f"all-gather of {world} x 128 B partial points + local sum",
            "int_roofline_frac_per_gpu": W * FIELD_MUL_IMAD / (ms * 1e-3) / int_peak / world}


def run_msm_sweep(bn, int_peak):
    """MSM points/s vs size, every power of two 2^10 .. 2^msm_max_log, three curves, device-resident, CUDA-event timed.
    SURVEY §8(d): field multiply = 72 IMAD for 2^255-19, 300 for BLS12-381; the secp256k1 field uses the special form of
    p = 2^256 - 2^32 - 977 (64 + 8 + 1 = 73 wide products executed; it was generic Montgomery, 128, until r02m), so its fraction
    is given for both the algorithmic 72 and the executed 73.
    Edwards mixed/full/double = 7/8/8 field multiplies, short Weierstrass 11/16/7."""
    torch, api, ctx, lib, _lib, args = bn.torch, bn.api, bn.ctx, bn.lib, bn._lib, bn.args
    out = {}
    curves = (("", _lib.RISTRETTO255, 253, 72, None, (7, 8, 8), 20),
              ("_secp256k1", _lib.SECP256K1, 256, 72, 73, (11, 16, 7), 16),
              ("_bls12_381_g1", _lib.BLS12_381_G1, 255, 300, None, (11, 16, 7), 16))
    for suffix, curve, bits, imad, imad_exec, costs, tables_max in curves:
        sweep = []
        max_log = args.msm_max_log
        n_max = 1 << max_log
        mult = torch.randint(0, 256, (n_max, 32), dtype=torch.uint8)
        mult[:, 31] &= 0x0F  # < 2^252 < every group order
        sc = torch.randint(0, 256, (n_max, 32), dtype=torch.uint8)
        sc[:, 31] &= 0x0F
        d_sc = sc.to(bn.dev)
        d_out = torch.empty(lib.bpp_point_bytes(curve), dtype=torch.uint8, device=bn.dev)
        mult_b = bytes(mult.numpy().tobytes())
        for lg in range(10, max_log + 1):
            n = 1 << lg
            # one point set per size: P_i = mult_i * g generated on the device
            pts = api.Points.from_multipliers(mult_b[:32 * n], curve, ctx)

            def msm():
                ctx.check(lib.bpp_msm_prepared_dev(ctx.h, pts.h, d_sc.data_ptr(), n, 1, d_out.data_ptr()), "msm")

            reps = 5 if lg >= 18 else 10
            generic_ms = bn.timed(msm, reps=reps, warm=2)  # Pippenger over the bare points (what a one-shot MulVec::calculate gets)
            first = bytes(d_out.cpu().numpy().tobytes())
            tables_ms = None
            if lg <= tables_max:  # window tables for a reused point set: no doublings, one bucket set; lose beyond ~2^18 points
                pts.precompute()
                tables_ms = bn.timed(msm, reps=reps, warm=2)
                assert bytes(d_out.cpu().numpy().tobytes()) == first, "window-table MSM disagrees with the generic one"
            ms, path = (tables_ms, "window tables") if tables_ms is not None and tables_ms < generic_ms else (generic_ms, "generic")
            direct_ms = direct_c = None
            if curve == _lib.RISTRETTO255 and lg <= 16:  # direct tables for a reused point set, as far as HBM allows beside the keys' tables
                budget = min(48 << 30, bn.torch.cuda.mem_get_info(bn.dev)[0] - (12 << 30))
                if budget > (1 << 30):
                    try:
                        direct_c = pts.precompute_direct(budget)
                        direct_ms = bn.timed(msm, reps=reps, warm=2)
                        assert bytes(d_out.cpu().numpy().tobytes()) == first, "direct-table MSM disagrees with the generic one"
                        if direct_ms < ms:
                            ms, path = direct_ms, "direct tables"
                    except bn.api.BppError as e:
                        direct_c = "unavailable: " + str(e)[:80]
            W, c_star, _ = algorithmic_field_muls(n, bits, *costs)
            row = {"log2_n": lg, "ms": ms, "points_per_sec": n / (ms * 1e-3), "int_roofline_frac": W * imad / (ms * 1e-3) / int_peak,
                   "path": path, "generic_ms": generic_ms, "tables_ms": tables_ms, "direct_tables_ms": direct_ms, "direct_tables_window_bits": direct_c,
                   "generic_int_roofline_frac": W * imad / (generic_ms * 1e-3) / int_peak}
            if imad_exec:
                row["int_roofline_frac_at_executed_imad_per_mul"] = W * imad_exec / (ms * 1e-3) / int_peak
            sweep.append(row)
            del pts
        out["msm_points_per_sec" + suffix] = sweep
    return out


if __name__ == "__main__":
    main()

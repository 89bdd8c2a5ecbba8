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


def synth_inputs(count: int, seed: int):
    """values uniform in [0, 2^31) (clear of the `v as i32` quirk), blinders uniform mod ell."""
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


def cpu_reference_verify(threads: int, proofs_per_thread: int, seed: int = 0):
    """The reference's CPU path for verify (n=64, m=16): oracle restatement, naive MulVec::calculate (mulvec.rs:20-33),
    one independent proof stream per host thread (the reference itself is single-threaded). Returns proofs/s."""
    from oracle import oracle_lib
    o = oracle_lib.get()
    vs, gs = synth_inputs(1, seed)
    proof = o.range_prove(0, N_BITS, M_AGG, vs, gs)
    V = o.commit(0, vs, gs)
    assert o.range_verify(0, N_BITS, M_AGG, V, proof) == 0
    results = []

    def work():
        ok = 0
        for _ in range(proofs_per_thread):
            ok += o.range_verify(0, N_BITS, M_AGG, V, proof) == 0  # ctypes releases the GIL
        results.append(ok)

    ts = [threading.Thread(target=work) for _ in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    assert sum(results) == threads * proofs_per_thread
    return threads * proofs_per_thread / dt, dt


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_thread = max(1, args.ref_proofs)
    for _ in range(min(args.warmup, 1)):
        cpu_reference_verify(cores, 1)
    vals = []
    for _ in range(args.steps):
        v, dt = cpu_reference_verify(cores, per_thread)
        vals.append((v, dt))
    total = sum(cores * per_thread for _ in vals) / sum(dt for _, dt in vals)
    ms = 1e3 * sum(dt for _, dt in vals) / len(vals)
    sample = f"{cores} threads x {per_thread} verifies of one n=64 m=16 proof per step (literal MulVec loop, 2089 scalar muls each)"
    line = {"impl": "reference", "metric": "range_proof_verifies_per_sec_n64_m16", "value": total, "unit": "proofs/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (integer)", "data": "synthetic",
            "config": {"workload": "range-proof verify n=64 m=16, Ristretto255, reference CPU algorithm (oracle port)"},
            "cpu_baseline": {"value": total, "unit": "proofs/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": total, "unit": "proofs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


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


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=2048, help="proofs per step per GPU")
    ap.add_argument("--prove-batch", type=int, default=256)
    ap.add_argument("--ref-proofs", type=int, default=2, help="reference arm: verifies per thread per step")
    ap.add_argument("--msm-max-log", type=int, default=22)
    ap.add_argument("--no-extras", action="store_true", help="skip proves/s, MSM sweep and CPU baseline")
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
    dev = torch.device("cuda", local_rank)
    from bulletproofsplus_b200 import api, _lib
    import ctypes as C
    # a non-default torch stream, made current: the library launches on exactly this stream, so torch.cuda.Event
    # records on it bracket the library's kernels (the legacy default stream has handle 0 = "make a private stream")
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx = api.Context(local_rank, stream.cuda_stream)
    lib, B, K_steps = ctx.lib, args.batch, args.steps
    NM = N_BITS * M_AGG
    pk = api.PublicKey.new(NM, _lib.RISTRETTO255, ctx)
    plen = lib.bpp_range_proof_len(0, N_BITS, M_AGG)

    # ---- synthetic workload: ROT distinct batches of B valid proofs, produced by the library's own prover ----
    ROT = 2
    host_proofs, host_V, dev_proofs, dev_V = [], [], [], []
    for r in range(ROT):
        vs, gs = synth_inputs(B, seed=rank * 1000 + r)
        proofs = []
        for c0 in range(0, B, 256):  # the library's own prover, 256 proofs per call
            c1 = min(B, c0 + 256)
            proofs += api.prove_batch(pk, N_BITS, M_AGG, vs[c0 * M_AGG:c1 * M_AGG], gs[c0 * M_AGG:c1 * M_AGG])
        Vs = pk.commit_batch(vs, gs)
        hp = torch.frombuffer(bytearray(b"".join(proofs)), dtype=torch.uint8).pin_memory()
        hv = torch.frombuffer(bytearray(b"".join(Vs)), dtype=torch.uint8).pin_memory()
        host_proofs.append(hp); host_V.append(hv)
        dev_proofs.append(hp.to(dev)); dev_V.append(hv.to(dev))
    d_verdicts = torch.full((B,), 255, dtype=torch.uint8, device=dev)
    h_verdicts = torch.empty((B,), dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

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

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    # what the kernel actually executes: fixed-base window tables (c = 12; 21 windows since the scalars are folded below ell/2, no
    # doublings) for the 2050 generator terms: fewer additions than the algorithmic count above (the 39 per-proof points take
    # 64 generic 4-bit windows each, outside the timed bracket)
    exec_madds = B * (2 + 2 * NM) * 21
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
        "kernel": "msm_accumulate_kernel<Ed25519> (+ its two long-bucket variants msm_accumulate_group_kernel<.,32|4>, timed together)",
        "achieved": ach_kernel / 1e12 if ach_kernel else None, "peak": int_peak / 1e12, "unit": "T 32x32+64 multiply-adds/s",
        "frac": (ach_kernel / int_peak) if ach_kernel else None,
        # dram__bytes_read.sum + dram__bytes_write.sum of ONE msm_accumulate_kernel launch at this batch size (B = 2048), from the ncu
        # --set full capture profiles/r01g_verify_kernels_ncu_full_details.txt: 1.220 GB + 0.511 GB. Not measured live.
        "traffic": 1.732e9 if B == 2048 else None, "traffic_unit": "bytes per launch",
        "ncu": {"capture": "profiles/r01g_verify_kernels_ncu_full_details.txt (msm_accumulate_kernel<Ed25519, 0>, fixed-base segment)",
                "gpu_time_ms": 5.63, "sm__pipe_fmaheavy_cycles_active_pct": 87.9, "smsp__issue_active_pct": 43.7,
                "active_lanes_per_warp_inst": 32.0, "registers_per_thread": 106, "dram_pct_of_peak": 3.8,
                "note": "the pipe that executes IMAD.WIDE is busy 87.9 % of the kernel's cycles: the kernel sits at the integer-multiply roofline "
                        "of sm_100a although only 0.4 of the 32-bit IMAD issue rate is reached (every wide multiply-add holds the pipe ~4.6 clk)"},
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
    split = None
    if not args.no_extras and world > 1:
        split = run_msm_split(args, torch, dist, api, _lib, ctx, dev, stream, rank, world, int_peak)
    if not args.no_extras and rank == 0:
        extras = run_extras(args, torch, api, _lib, ctx, pk, dev, stream, int_peak)
    if not args.no_extras and rank == 0 and world == 1:
        cores = os.cpu_count() or 1
        v, dt = cpu_reference_verify(cores, 4)
        cpu = {"value": v, "unit": "proofs/s", "cores": cores, "kind": "port",
               "sample": f"{cores} threads x 4 verifies of an n=64 m=16 proof ({dt:.1f} s wall, {cores * dt:.0f} CPU-s), oracle restatement "
                         f"of the reference (naive MulVec, one proof stream per thread; the reference itself is single-threaded)"}

    if rank == 0:
        line = {
            "metric": "range_proof_verifies_per_sec_n64_m16", "value": value, "unit": "proofs/s", "n_gpus": world, "steps": K_steps,
            "warmup": args.warmup, "ms_per_step": elapsed_ms / K_steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 limbs (integer, GF(2^255-19) / mod ell)", "data": "synthetic",
            "config": {"workload": f"range-proof batch verify, n=64 bits, m=16 aggregated, Ristretto255, {B} proofs/step/GPU, "
                                   f"2089-term MSM each, fixed-challenge (reference) mode", "batch_per_gpu": B, "parallelism": f"proofs sharded x{world}",
                       "l2": "256 MB memset between steps (inside the timed region); intermediates per step > L2"},
            "e2e": {"value": e2e_value, "unit": "proofs/s", "h2d_bytes_per_step": B * (plen + M_AGG * 32), "d2h_bytes_per_step": B},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        }
        line.update(extras)
        if split:
            line["msm_point_range_split"] = split
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_msm_split(args, torch, dist, api, _lib, ctx, dev, stream, rank, world, int_peak):
    """One MSM of 2^msm_max_log points split by point range over the ranks: per-rank MSM to one raw
    partial point, NCCL all-gather of world x 128 B, local sum + encode. Strong scaling; time = max over ranks."""
    lib = ctx.lib
    n_total = 1 << args.msm_max_log
    n_loc = n_total // world
    g = torch.Generator().manual_seed(1234 + rank)
    mult = torch.randint(0, 256, (n_loc, 32), dtype=torch.uint8, generator=g)
    mult[:, 31] &= 0x0F
    pts = api.Points.from_multipliers(bytes(mult.numpy().tobytes()), _lib.RISTRETTO255, ctx)
    sc = torch.randint(0, 256, (n_loc, 32), dtype=torch.uint8, generator=g)
    sc[:, 31] &= 0x0F
    d_sc = sc.to(dev)
    d_raw = torch.empty(128, dtype=torch.uint8, device=dev)
    d_all = torch.empty(128 * world, dtype=torch.uint8, device=dev)
    d_out = torch.empty(32, dtype=torch.uint8, device=dev)

    def step():
        ctx.check(lib.bpp_msm_prepared_raw_dev(ctx.h, pts.h, d_sc.data_ptr(), n_loc, d_raw.data_ptr()), "msm_raw_dev")
        dist.all_gather_into_tensor(d_all, d_raw)
        ctx.check(lib.bpp_points_sum_raw_dev(ctx.h, 0, d_all.data_ptr(), world, d_out.data_ptr()), "sum_raw_dev")

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    outs = [torch.empty(32, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(outs, d_out)
    same = all(bool((o == outs[0]).all()) for o in outs)
    reps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(reps):
        step()
    e1.record(stream)
    dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    W, _, _ = algorithmic_field_muls(n_total)
    return {"log2_n": args.msm_max_log, "n_gpus": world, "ms": ms, "points_per_sec": n_total / (ms * 1e-3), "scaling": "strong",
            "all_ranks_agree": same, "exchange": f"all-gather of {world} x 128 B partial points + local sum",
            "int_roofline_frac_per_gpu": W * FIELD_MUL_IMAD / (ms * 1e-3) / int_peak / world}


def run_extras(args, torch, api, _lib, ctx, pk, dev, stream, int_peak):
    """proves/s (n=64, m=16) and the MSM points/s sweep, device-resident, CUDA-event timed."""
    lib = ctx.lib
    out = {}
    # proves/s
    PB = args.prove_batch
    vs, gs = synth_inputs(PB, seed=99)
    d_v = torch.tensor(vs, dtype=torch.int64, device=dev)
    d_g = torch.frombuffer(bytearray(b"".join(g.to_bytes(32, "little") for g in gs)), dtype=torch.uint8).to(dev)
    plen = lib.bpp_range_proof_len(0, N_BITS, M_AGG)
    d_p = torch.empty(PB * plen, dtype=torch.uint8, device=dev)

    def prove():
        ctx.check(lib.bpp_range_prove_batch_dev(ctx.h, pk.handle, N_BITS, M_AGG, PB, d_v.data_ptr(), d_g.data_ptr(), None, d_p.data_ptr()), "prove")

    for _ in range(2):
        prove()
    torch.cuda.synchronize()
    reps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        prove()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out["proves_per_sec_n64_m16"] = {"value": PB / (ms * 1e-3), "unit": "proofs/s", "batch": PB, "ms_per_batch": ms}
    # real Fiat-Shamir, GPU side: per-proof challenges on the verifier, the interactive prover (12 host round trips per batch).
    # The challenges are the constants, so the transcript hashing of the host (merlin.py is pure Python) is not in these numbers.
    import ctypes as C
    B = args.batch
    vs2, gs2 = synth_inputs(B, seed=7)
    proofs = []
    for c0 in range(0, B, 256):
        c1 = min(B, c0 + 256)
        proofs += api.prove_batch(pk, N_BITS, M_AGG, vs2[c0 * M_AGG:c1 * M_AGG], gs2[c0 * M_AGG:c1 * M_AGG])
    d_pr = torch.frombuffer(bytearray(b"".join(proofs)), dtype=torch.uint8).to(dev)
    d_V = torch.frombuffer(bytearray(b"".join(pk.commit_batch(vs2, gs2))), dtype=torch.uint8).to(dev)
    L = 10
    block = b"".join(int(x).to_bytes(32, "little") for x in [12, 23, 99] + [7] * L)
    d_ch = torch.frombuffer(bytearray(block * B), dtype=torch.uint8).to(dev)
    d_vd = torch.full((B,), 255, dtype=torch.uint8, device=dev)

    def verify_fs():
        ctx.check(lib.bpp_range_verify_batch_fs_dev(ctx.h, pk.handle, N_BITS, M_AGG, B, d_V.data_ptr(), d_pr.data_ptr(), d_ch.data_ptr(), d_vd.data_ptr()),
                  "verify_fs")

    for _ in range(2):
        verify_fs()
    torch.cuda.synchronize()
    assert bool((d_vd == 0).all())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        verify_fs()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out["verifies_per_sec_per_proof_challenges"] = {"value": B / (ms * 1e-3), "unit": "proofs/s", "batch": B, "ms_per_batch": ms,
                                                    "entry": "bpp_range_verify_batch_fs_dev"}

    # random-linear-combination batch verification: same proofs, one combined MSM per batch (fallback to per-proof on failure)
    import secrets
    d_w = torch.frombuffer(bytearray(b"".join((secrets.randbits(128) | 1).to_bytes(32, "little") for _ in range(B))), dtype=torch.uint8).to(dev)
    fb = C.c_int(0)

    def verify_rlc():
        ctx.check(lib.bpp_range_verify_batch_rlc_dev(ctx.h, pk.handle, N_BITS, M_AGG, B, d_V.data_ptr(), d_pr.data_ptr(), None, None, d_w.data_ptr(),
                                                     d_vd.data_ptr(), C.byref(fb)), "verify_rlc")

    for _ in range(2):
        verify_rlc()
    torch.cuda.synchronize()
    assert bool((d_vd == 0).all()) and fb.value == 0
    e0.record(stream)
    for _ in range(reps):
        verify_rlc()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out["verifies_per_sec_rlc_batch"] = {"value": B / (ms * 1e-3), "unit": "proofs/s", "batch": B, "ms_per_batch": ms,
                                         "entry": "bpp_range_verify_batch_rlc_dev (all proofs valid: no fallback; one host sync per batch inside)"}

    def prove_fs():
        ses = api.ProveSession(pk, N_BITS, M_AGG, vs[:PB * M_AGG], gs[:PB * M_AGG])
        ch = [(12, 23)] * PB
        for _ in range(ses.rounds + 1):
            ses.step(ch)
            ch = [7] * PB
        return ses.finish([99] * PB)

    prove_fs()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        got = prove_fs()
    dt = (time.perf_counter() - t0) / 3
    out["proves_per_sec_interactive"] = {"value": PB / dt, "unit": "proofs/s", "batch": PB, "ms_per_batch": dt * 1e3, "host_round_trips": 12,
                                         "entry": "bpp_range_prove_begin/_step/_finish (host wall clock, host buffers)"}
    # MSM sweep, all three curves (field multiply = 72 IMAD for 2^255-19 and the secp256k1 prime, 300 for
    # BLS12-381; Edwards mixed/full/double = 7/8/8, short Weierstrass 11/16/7 field multiplies)
    curves = (("", _lib.RISTRETTO255, args.msm_max_log, 253, 72, (7, 8, 8)),
              ("_secp256k1", _lib.SECP256K1, min(args.msm_max_log, 20), 256, 72, (11, 16, 7)),
              ("_bls12_381_g1", _lib.BLS12_381_G1, min(args.msm_max_log, 20), 255, 300, (11, 16, 7)))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        reps = 5
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    for suffix, curve, max_log, bits, imad, costs in curves:
        sweep = []
        n_max = 1 << max_log
        mult = torch.randint(0, 256, (n_max, 32), dtype=torch.uint8)
        mult[:, 31] &= 0x0F  # < 2^252 < every group order
        sc = torch.randint(0, 256, (n_max, 32), dtype=torch.uint8)
        sc[:, 31] &= 0x0F
        d_sc = sc.to(dev)
        d_out = torch.empty(lib.bpp_point_bytes(curve), dtype=torch.uint8, device=dev)
        for lg in range(10, max_log + 1, 2):
            n = 1 << lg
            # one point set per size: P_i = mult_i * g generated on the device
            pts = api.Points.from_multipliers(bytes(mult[:n].numpy().tobytes()), curve, ctx)

            def msm():
                ctx.check(lib.bpp_msm_prepared_dev(ctx.h, pts.h, d_sc.data_ptr(), n, 1, d_out.data_ptr()), "msm")

            generic_ms = timed(msm)          # Pippenger over the bare points (what a one-shot MulVec::calculate gets)
            first = bytes(d_out.cpu().numpy().tobytes())
            pts.precompute()                 # window tables for a reused point set: no doublings, one bucket set
            tables_ms = timed(msm)
            assert bytes(d_out.cpu().numpy().tobytes()) == first, "window-table MSM disagrees with the generic one"
            # tables win while the Horner pass dominates (small N); beyond ~2^18 points the 16x larger gather footprint costs more
            ms, path = (tables_ms, "window tables") if tables_ms < generic_ms else (generic_ms, "generic")
            W, c_star, _ = algorithmic_field_muls(n, bits, *costs)
            sweep.append({"log2_n": lg, "ms": ms, "points_per_sec": n / (ms * 1e-3), "int_roofline_frac": W * imad / (ms * 1e-3) / int_peak,
                          "path": path, "generic_ms": generic_ms, "tables_ms": tables_ms})
            del pts
        out["msm_points_per_sec" + suffix] = sweep
    return out


if __name__ == "__main__":
    main()

"""World-size-2 NCCL parity (skipped with fewer than two GPUs): bulletproofsplus_b200.parallel over libbpp_b200.so on two
ranks, compared with the CPU oracle — not merely rank against rank. Covers the two ways the path shards:
independent proofs per rank (verdicts and proof bytes all-gathered) and one MSM split by point range (raw partial points
all-gathered, summed and encoded on every rank) against the oracle's (sum_i s_i k_i) * g."""
import os
import random
import socket

import pytest

pytestmark = pytest.mark.gpu

ELL = 2**252 + 27742317777372353535851937790883648493


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _le32(x):
    return int(x).to_bytes(32, "little")


def _inputs(count, n, m, seed):
    rng = random.Random(seed)
    return [rng.randrange(2**min(n, 31)) for _ in range(count * m)], [rng.randrange(ELL) for _ in range(count * m)]


def _worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from bulletproofsplus_b200 import api, parallel
        ctx = api.Context(rank)
        dev = torch.device("cuda", rank)
        n, m, count = 64, 2, 9  # 9 proofs over 2 ranks: blocks of 5 and 4
        pk = api.PublicKey.new(n * m, 0, ctx)
        vs, gs = _inputs(count, n, m, 21)
        plen = ctx.lib.bpp_range_proof_len(0, n, m)
        proofs = parallel.prove_batch_sharded(lambda v, g: api.prove_batch(pk, n, m, list(v), list(g)), m, vs, gs, plen, device=dev)
        Vs = pk.commit_batch(vs, gs)
        Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
        tampered = list(proofs)
        for i in (2, 7):  # one in each rank's block
            bad = bytearray(tampered[i])
            bad[-96] ^= 1  # r' +- 1
            tampered[i] = bytes(bad)
        verdicts = parallel.verify_batch_sharded(lambda p, c: api.verify_batch(pk, n, m, list(p), list(c)), tampered, Vl, device=dev)
        # one MSM split by point range
        k = 3001
        rng = random.Random(33)
        mult = [rng.randrange(ELL) for _ in range(k)]
        sc = [rng.randrange(ELL) for _ in range(k)]

        def partial_fn(lo, hi):
            pts = api.Points.from_multipliers(b"".join(_le32(x) for x in mult[lo:hi]), 0, ctx)
            return pts.msm_raw(b"".join(_le32(x) for x in sc[lo:hi]))

        msm = parallel.msm_sharded(partial_fn, lambda raws, cnt: api.points_sum_raw(raws, 0, ctx), k, ctx.lib.bpp_raw_point_bytes(0), device=dev)
        out.put((rank, proofs, verdicts, msm))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_nccl_match_the_oracle(oracle):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mpc = mp.get_context("spawn")
    out = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=600) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    n, m, count = 64, 2, 9
    vs, gs = _inputs(count, n, m, 21)
    want_proofs = [oracle.range_prove(0, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m]) for i in range(count)]
    want_verdicts = []
    for i in range(count):
        p = bytearray(want_proofs[i])
        if i in (2, 7):
            p[-96] ^= 1
        want_verdicts.append(oracle.range_verify(0, n, m, oracle.commit(0, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m]), bytes(p)))
    assert want_verdicts == [1 if i in (2, 7) else 0 for i in range(count)]
    rng = random.Random(33)
    k = 3001
    mult = [rng.randrange(ELL) for _ in range(k)]
    sc = [rng.randrange(ELL) for _ in range(k)]
    want_msm = oracle.scalar_mul(0, sum(a * b for a, b in zip(mult, sc)) % ELL)
    for rank, proofs, verdicts, msm in results:
        assert proofs == want_proofs, f"rank {rank}: sharded proofs differ from the oracle"
        assert verdicts == want_verdicts, f"rank {rank}: {verdicts}"
        assert msm == want_msm, f"rank {rank}: point-range-split MSM differs from (sum s_i k_i) g"

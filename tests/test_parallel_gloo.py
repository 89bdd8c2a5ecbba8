"""N > 1 host logic on CPU: world_size-2 `gloo` process groups driving bulletproofsplus_b200.parallel with the CPU oracle
as the compute (the oracle is the checker here; on the GPU box the same functions wrap libbpp_b200.so — see
tests/test_gpu_parity.py::test_sharded_* and bench.py). Checks that every sharded result is identical to the unsharded
one: verdict bytes, proof bytes and the MSM point."""
import os
import random
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bulletproofsplus_b200 import parallel

ELL = 2**252 + 27742317777372353535851937790883648493


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs(count, n, m, seed):
    rng = random.Random(seed)
    return [rng.randrange(2**n) for _ in range(count * m)], [rng.randrange(ELL) for _ in range(count * m)]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle_lib
        o = oracle_lib.get()
        n, m, count = 8, 2, 5  # 5 proofs over 2 ranks: blocks of 3 and 2
        vs, gs = _inputs(count, n, m, 11)
        plen = len(o.range_prove(0, n, m, vs[:m], gs[:m]))

        def prove_fn(v, g):
            return [o.range_prove(0, n, m, v[i * m:(i + 1) * m], g[i * m:(i + 1) * m]) for i in range(len(v) // m)]

        proofs = parallel.prove_batch_sharded(prove_fn, m, vs, gs, plen)
        Vs = [[o.commit(0, [vs[i * m + j]], [gs[i * m + j]]) for j in range(m)] for i in range(count)]
        bad = bytearray(proofs[3])
        bad[-1] ^= 1  # delta' changed (still canonical or not: either way not OK)
        proofs_t = list(proofs)
        proofs_t[3] = bytes(bad)

        def verify_fn(ps, cs):
            return [o.range_verify(0, n, m, b"".join(c), p) for p, c in zip(ps, cs)]

        verdicts = parallel.verify_batch_sharded(verify_fn, proofs_t, Vs)

        # MSM split by point range: the "raw" partial is simply the encoded point on the CPU
        k = 37
        rng = random.Random(5)
        mult = [rng.randrange(ELL) for _ in range(k)]
        sc = [rng.randrange(ELL) for _ in range(k)]
        pts = [o.scalar_mul(0, x) for x in mult]

        def partial_fn(lo, hi):
            if hi == lo:
                return o.scalar_mul(0, 0)
            return o.msm_naive(0, b"".join(s.to_bytes(32, "little") for s in sc[lo:hi]), b"".join(pts[lo:hi]))

        def sum_fn(raws, cnt):
            acc = raws[:32]
            for i in range(1, cnt):
                acc = o.point_add(0, acc, raws[32 * i:32 * (i + 1)])
            return acc

        msm = parallel.msm_sharded(partial_fn, sum_fn, k, 32)
        out.put((rank, proofs, verdicts, msm))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_shard_range_covers_everything():
    for count in (0, 1, 5, 8, 8191, 8192):
        for world in (1, 2, 3, 8):
            blocks = [parallel.shard_range(count, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == count
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        parallel.shard_range(4, 2, 2)


def test_world_size_2_gloo_matches_single_process():
    from oracle import oracle_lib
    o = oracle_lib.get()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=180) for _ in range(2)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process expectations
    n, m, count = 8, 2, 5
    vs, gs = _inputs(count, n, m, 11)
    want_proofs = [o.range_prove(0, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m]) for i in range(count)]
    rng = random.Random(5)
    k = 37
    mult = [rng.randrange(ELL) for _ in range(k)]
    sc = [rng.randrange(ELL) for _ in range(k)]
    want_msm = o.scalar_mul(0, sum(a * b for a, b in zip(mult, sc)) % ELL)
    for rank, proofs, verdicts, msm in results:
        assert proofs == want_proofs, f"rank {rank}: sharded proofs differ"
        assert verdicts[:3] == [0, 0, 0] and verdicts[4] == 0 and verdicts[3] in (1, 2), verdicts
        assert msm == want_msm
    assert results[0][2] == results[1][2]

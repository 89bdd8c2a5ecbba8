"""The engine's A/B switches must not change a single byte: the same small workload (commitments, an aggregated proof,
a 40-proof batch with tampered members, window-table and generic MSMs over 4096 points, short split-digit jobs) runs in a
fresh process per setting — default, BPP_HYBRID=5 / f (half / all accumulation blocks on the FP64 pipe, csrc/fe25519_f64.cuh),
BPP_NO_OVERLAP=1 (generic segment on the caller's stream), BPP_NO_SPLIT=1, BPP_NO_STRAUS=1 (per-proof points through the
bucket pipeline instead of msm_straus_kernel), BPP_FUSE=1 (exponents inside the direct kernel; the workload's second batch runs
against direct tables), BPP_NO_CLUSTER=1 (one block per proof in wip_prove_kernel instead of a thread-block cluster; the
workload proves one n = 64, m = 16 proof alone for it), BPP_DIRECT_MIN_JOBS=32 (few jobs keep the bucket path although the key
has tables) — and every output is compared with the CPU oracle."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKLOAD = r'''
import random, sys
sys.path.insert(0, %(root)r)
from bulletproofsplus_b200 import api
from oracle import oracle_lib
o = oracle_lib.get()
ctx = api.Arith.init(0)
q = api.ORDERS[0]
rng = random.Random(2024)
le32 = lambda x: x.to_bytes(32, "little")
n, m, count = 64, 2, 40
pk = api.PublicKey.new(n * m, 0, ctx)
vs = [rng.randrange(2**31) for _ in range(count * m)]
gs = [rng.randrange(q) for _ in range(count * m)]
Vs = pk.commit_batch(vs, gs)
assert b"".join(Vs) == o.commit(0, vs, gs), "commitments"
proofs = api.prove_batch(pk, n, m, vs, gs)
for i in (0, 17, 39):
    assert proofs[i] == o.range_prove(0, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m]), ("proof", i)
bad = {3: -96, 11: 5, 29: -2}  # r' + 1, a byte of A, a byte of delta'
batch = []
for i, p in enumerate(proofs):
    b = bytearray(p)
    if i in bad:
        b[bad[i]] ^= 1
    batch.append(bytes(b))
Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
got = api.verify_batch(pk, n, m, batch, Vl)
want = [o.range_verify(0, n, m, b"".join(Vl[i]), batch[i]) for i in range(count)]
assert got == want, (got, want)
assert [i for i, v in enumerate(got) if v] == sorted(bad), got
assert pk.precompute(1 << 30) >= 8  # direct tables: the same batch again (>= 32 jobs take them)
assert api.verify_batch(pk, n, m, batch, Vl) == want, "verify with direct tables"
assert api.verify_batch(pk, n, m, batch[:5], Vl[:5]) == want[:5], "five proofs with direct tables"
assert api.verify_batch(pk, n, m, batch[3:4], Vl[3:4]) == want[3:4] == [1], "one (tampered) proof with direct tables"
assert api.prove_batch(pk, n, m, vs, gs) == proofs, "prove with direct tables"
pk16 = api.PublicKey.new(1024, 0, ctx)  # one large proof alone: a cluster of 8 blocks in the WIP prover
v16 = [rng.randrange(2**31) for _ in range(16)]
g16 = [rng.randrange(q) for _ in range(16)]
assert api.prove_batch(pk16, 64, 16, v16, g16)[0] == o.range_prove(0, 64, 16, v16, g16), "single n=64 m=16 proof"
N = 4096
mult = [rng.randrange(q) for _ in range(N)]
sc = [rng.randrange(q) for _ in range(N)]
pts = api.Points.from_multipliers(b"".join(le32(x) for x in mult), 0, ctx)
want = o.scalar_mul(0, sum(a * b for a, b in zip(mult, sc)) %% q)
assert pts.msm(b"".join(le32(x) for x in sc))[0] == want, "generic msm"
assert pts.precompute().msm(b"".join(le32(x) for x in sc))[0] == want, "window-table msm"
for k, jobs in ((7, 2), (150, 2), (900, 1)):
    ss = [rng.randrange(q) for _ in range(k * jobs)]
    res = pts.msm(b"".join(le32(x) for x in ss), k=k, n_jobs=jobs)
    for j in range(jobs):
        assert res[j] == o.scalar_mul(0, sum(a * b for a, b in zip(mult[:k], ss[j * k:(j + 1) * k])) %% q), (k, j)
print("switch workload ok", ctx.kernel_launches())
'''


@pytest.mark.parametrize("env", [{}, {"BPP_HYBRID": "5"}, {"BPP_HYBRID": "f"}, {"BPP_NO_OVERLAP": "1"}, {"BPP_NO_SPLIT": "1"}, {"BPP_NO_STRAUS": "1"},
                                 {"BPP_FUSE": "1"}, {"BPP_NO_CLUSTER": "1"}, {"BPP_DIRECT_MIN_JOBS": "32"}],
                         ids=["default", "hybrid5", "hybridf", "no_overlap", "no_split", "no_straus", "fuse", "no_cluster", "direct_min_32"])
def test_switch_does_not_change_results(env):
    e = dict(os.environ)
    for k in ("BPP_HYBRID", "BPP_NO_OVERLAP", "BPP_NO_SPLIT", "BPP_NO_STRAUS", "BPP_FUSE", "BPP_NO_CLUSTER", "BPP_DIRECT_MIN_JOBS"):
        e.pop(k, None)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", WORKLOAD % {"root": ROOT}], env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "switch workload ok" in r.stdout

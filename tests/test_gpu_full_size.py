"""GPU parity at BASELINE.json's full sizes (configs 2-5), through the C ABI.

Where the CPU oracle finishes in seconds (one N = 4096 proof) the comparison is byte for byte; where it does not (8192
proofs, 2^20 / 2^22-term MSMs) the check is the size-independent property the domain offers: an MSM over P_i = k_i g must
equal (sum s_i k_i) g — one CPU scalar multiplication — and a batch's verdict vector must be non-zero exactly at the
tampered positions, with the oracle's own verdicts on the tampered proofs and on a sample of the valid ones."""
import random

import pytest

from tests import pyref
from tests.test_gpu_parity import _tampers, le32

pytestmark = pytest.mark.gpu

R = 0
ELL = pyref.ELL


@pytest.fixture(scope="module")
def api():
    from bulletproofsplus_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    return api.Arith.init()


def _scalars(seed, count, bound=ELL):
    """`count` scalars below `bound` from one seeded byte stream (the SHA-256 counter stream of pyref is too slow for 2^22)."""
    raw = random.Random(seed).randbytes(40 * count)
    return [int.from_bytes(raw[40 * i:40 * i + 40], "little") % bound for i in range(count)]


def _enc(v):
    return b"".join(le32(x) for x in v)


# ---------------------------------------------------------------------------------- config 2: standalone WIP, 2^12 generators
def test_config2_wip_4096(api, ctx, oracle):
    N, y = 4096, 12
    a, b = _scalars(0xC2A, N), _scalars(0xC2B, N)
    ypow = [pow(y, i + 1, ELL) for i in range(N)]
    gamma = _scalars(0xC2C, 1)[0]
    pk = api.PublicKey.new(N, R, ctx)
    proof = api.WeightedInnerProductProof.prove(pk, a, b, ypow, gamma)
    want = oracle.wip_prove(R, N, _enc(a), _enc(b), _enc(ypow), gamma)
    assert proof.to_bytes() == want
    assert len(want) == (2 * 12 + 2) * 32 + 96
    dl = (sum(a[i] * 3 * (i + 1) + b[i] * 5 * (i + 1) + a[i] * b[i] * ypow[i] for i in range(N)) + 2 * gamma) % ELL
    P = oracle.scalar_mul(R, dl)
    zeros = [0] * N
    proof.verify(pk, ypow, zeros, zeros, 0, [], P, [])  # 8221-term verifier MSM
    with pytest.raises(api.VerificationError):
        proof.verify(pk, ypow, zeros, zeros, 0, [], oracle.scalar_mul(R, (dl + 1) % ELL), [])


# ---------------------------------------------------------------------------------- config 3: n = 64, m = 64 (8285-term verifier MSM)
def test_config3_aggregated_64x64(api, ctx, oracle):
    n, m = 64, 64
    vs, gs = _scalars(0xC3A, m, 2**31), _scalars(0xC3B, m)
    pk = api.PublicKey.new(n * m, R, ctx)
    proof = api.prove_batch(pk, n, m, vs, gs)[0]
    assert proof == oracle.range_prove(R, n, m, vs, gs)
    V = oracle.commit(R, vs, gs)
    assert b"".join(pk.commit_batch(vs, gs)) == V
    cases = [("valid", proof, V)] + _tampers(oracle, n, m, proof, V)
    got = api.verify_batch(pk, n, m, [c[1] for c in cases], [[c[2][i * 32:(i + 1) * 32] for i in range(m)] for c in cases])
    assert got[0] == 0 and all(v == 1 for v in got[1:]), got
    assert oracle.range_verify(R, n, m, V, proof) == 0
    assert oracle.range_verify(R, n, m, cases[1][2], cases[1][1]) == 1


# ---------------------------------------------------------------------------------- config 4: 8192 independent n = 64, m = 1 proofs
def test_config4_batch_8192(api, ctx, oracle):
    n, m, count, bad = 64, 1, 8192, 128
    pk = api.PublicKey.new(n, R, ctx)
    vs, gs = _scalars(0xC4A, count, 2**31), _scalars(0xC4B, count)
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    rng = random.Random(0xB2000005)
    picked = sorted(rng.sample(range(count), bad))
    sample = sorted(rng.sample(sorted(set(range(count)) - set(picked)), 64))
    for i in sample[:8]:
        assert proofs[i] == oracle.range_prove(R, n, m, [vs[i]], [gs[i]])
    for t, j in enumerate(picked):
        kinds = _tampers(oracle, n, m, proofs[j], Vs[j])
        _, proofs[j], Vs[j] = kinds[t % len(kinds)]
    got = api.verify_batch(pk, n, m, proofs, [[v] for v in Vs])
    assert [i for i, v in enumerate(got) if v != 0] == picked
    for i in picked + sample:
        assert got[i] == oracle.range_verify(R, n, m, Vs[i], proofs[i])


# ---------------------------------------------------------------------------------- config 5: raw MSM sweep, top sizes and distributions
@pytest.mark.parametrize("logn", [20, 22])
def test_config5_msm_large(api, ctx, oracle, logn):
    n = 1 << logn
    ks, ss = _scalars(0xC5A + logn, n), _scalars(0xC5B + logn, n)
    pts = api.Points.from_multipliers(_enc(ks), R, ctx)
    for i in (0, n - 1, n // 3):  # spot-check the device-generated table: a unit scalar selects one point
        unit = bytearray(32 * n)
        unit[32 * i] = 1
        assert pts.msm(bytes(unit))[0] == oracle.scalar_mul(R, ks[i])
    want = oracle.scalar_mul(R, sum(s * k for s, k in zip(ss, ks)) % ELL)
    assert pts.msm(_enc(ss))[0] == want
    assert pts.precompute().msm(_enc(ss))[0] == want  # window tables of the same points: same point


@pytest.mark.parametrize("logn", [16, 20])
def test_config5_msm_distributions(api, ctx, oracle, logn):
    n = 1 << logn
    rng = random.Random(logn)
    ks = _scalars(0xC5C + logn, n)
    pts = api.Points.from_multipliers(_enc(ks), R, ctx)
    s_eq = rng.randrange(ELL)
    full = _scalars(0xC5D + logn, n)
    dists = {
        "all equal": [s_eq] * n,
        "below 2^16": _scalars(0xC5E + logn, n, 2**16),
        "quarter zero": [0 if i % 4 == 0 else s for i, s in enumerate(full)],
    }
    same = api.Points.from_multipliers(le32(ks[0]) * n, R, ctx)
    alt = [ks[1] if i % 2 == 0 else ELL - ks[1] for i in range(n)]  # P, -P, P, -P, ...
    pm = api.Points.from_multipliers(_enc(alt), R, ctx)
    gens = [3 * (i + 1) if i < n // 2 else 5 * (i - n // 2 + 1) for i in range(n)]  # the reference's structured generators
    pg = api.Points.from_multipliers(_enc(gens), R, ctx)
    want = {name: oracle.scalar_mul(R, sum(s * k for s, k in zip(ss, ks)) % ELL) for name, ss in dists.items()}
    for tables in (False, True):
        if tables:
            for p in (pts, same, pm, pg):
                p.precompute()
        for name, ss in dists.items():
            assert pts.msm(_enc(ss))[0] == want[name], (name, tables)
        assert same.msm(_enc(full))[0] == oracle.scalar_mul(R, sum(full) * ks[0] % ELL)
        assert pm.msm(_enc(full))[0] == oracle.scalar_mul(R, sum(s * k for s, k in zip(full, alt)) % ELL)
        assert pg.msm(_enc(full))[0] == oracle.scalar_mul(R, sum(s * k for s, k in zip(full, gens)) % ELL)

"""GPU parity tests (run on the B200 box: `pytest -m gpu`). Every test drives the CUDA path through the C ABI
(libbpp_b200.so via bulletproofsplus_b200.api) and compares bit-for-bit with the CPU oracle / golden
fixtures on the same seeded inputs. Bar: byte-identical encodings, scalars, proofs and verdicts."""
import json
import os
import random

import pytest

from oracle import exponent_checker as ec
from oracle import oracle_lib
from tests import pyref

pytestmark = pytest.mark.gpu

R = 0
ELL = pyref.ELL
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def api():
    from bulletproofsplus_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    return api.Arith.init()


def le32(x):
    return int(x).to_bytes(32, "little")


# ---------------------------------------------------------------------------------- unit level (debug ops)
def test_field_ops(ctx):
    p = pyref.P25519
    rng = random.Random(1)
    edge = [0, 1, 2, 19, 38, p - 1, p, p + 1, 2 * p - 1, 2 * p, 2 * p + 1, 2**255, 2**256 - 1, 2**256 - 38, 2**256 - 39, 2**128, 2**224 - 1]
    xs = edge + [rng.randrange(2**256) for _ in range(300)]
    ys = list(reversed(edge)) + [rng.randrange(2**256) for _ in range(300)]
    a, b = b"".join(le32(x) for x in xs), b"".join(le32(y) for y in ys)
    for op, fn in ((0, lambda x, y: x * y), (1, lambda x, y: x + y), (2, lambda x, y: x - y), (3, lambda x, y: pow(x % p, p - 2, p)),
                   (4, lambda x, y: x * x)):
        out, _ = ctx.debug_op(R, op, a, b, 32)
        for i, (x, y) in enumerate(zip(xs, ys)):
            got = int.from_bytes(out[32 * i:32 * i + 32], "little")
            assert got == fn(x, y) % p, (op, i, hex(x), hex(y))


def test_scalar_field_ops(ctx):
    rng = random.Random(2)
    edge = [0, 1, 2, ELL - 1, ELL - 2, 2**252, 7, 99]
    xs = edge + [rng.randrange(ELL) for _ in range(100)]
    ys = list(reversed(edge)) + [rng.randrange(ELL) for _ in range(100)]
    a, b = b"".join(le32(x) for x in xs), b"".join(le32(y) for y in ys)
    for op, fn in ((20, lambda x, y: x * y), (21, lambda x, y: pow(x, ELL - 2, ELL)), (22, lambda x, y: x + y), (23, lambda x, y: x - y)):
        out, _ = ctx.debug_op(R, op, a, b, 32)
        for i, (x, y) in enumerate(zip(xs, ys)):
            assert int.from_bytes(out[32 * i:32 * i + 32], "little") == fn(x, y) % ELL, (op, i)


def test_group_ops(ctx, oracle):
    rng = random.Random(3)
    ks = [0, 1, 2, 3, 5, 15, ELL - 1] + [rng.randrange(ELL) for _ in range(20)]
    js = [1, 1, 2, ELL - 3, 5, 0, 1] + [rng.randrange(ELL) for _ in range(20)]
    A = [oracle.scalar_mul(R, k) for k in ks]
    B = [oracle.scalar_mul(R, j) for j in js]
    a, b = b"".join(A), b"".join(B)
    exp = {10: A, 11: [oracle.scalar_mul(R, (k + j) % ELL) for k, j in zip(ks, js)], 12: [oracle.scalar_mul(R, 2 * k % ELL) for k in ks],
           14: [oracle.scalar_mul(R, (k - j) % ELL) for k, j in zip(ks, js)]}
    exp[13] = exp[11]
    for op, want in exp.items():
        out, ok = ctx.debug_op(R, op, a, b, 32)
        assert all(ok), op
        assert [out[32 * i:32 * i + 32] for i in range(len(ks))] == want, op
    # invalid encodings are flagged exactly where the oracle flags them
    p = pyref.P25519
    cands = [p.to_bytes(32, "little"), (1).to_bytes(32, "little"), (2**256 - 1).to_bytes(32, "little")]
    cands += [(rng.randrange(p) & ~1).to_bytes(32, "little") for _ in range(61)]
    _, ok = ctx.debug_op(R, 10, b"".join(cands), b"".join(A[:1] * len(cands)), 32)
    assert [bool(o) for o in ok] == [oracle.point_decode_ok(R, c) for c in cands]


# ---------------------------------------------------------------------------------- PublicKey / commit
def test_public_key(api, ctx, oracle):
    for length in (1, 8, 64):
        pk = api.PublicKey.new(length, R, ctx)
        g, h, G, H = oracle.pk_new(R, length)
        assert pk.g == g and pk.h == h
        assert b"".join(pk.G_vec) == G and b"".join(pk.H_vec) == H


def test_commit_batch(api, ctx, oracle):
    pk = api.PublicKey.new(4, R, ctx)
    rng = random.Random(5)
    vs = [0, 1, 2, 5, 2**31 - 1, 2**31, 2**32 + 9, 2**64 - 1] + [rng.randrange(2**31) for _ in range(40)]
    gs = [0, 1, 3, 7, ELL - 1, 11, 13, 17] + [rng.randrange(ELL) for _ in range(40)]
    got = b"".join(pk.commit_batch(vs, gs))
    assert got == oracle.commit(R, vs, gs)
    prover = api.RangeProver()
    prover.commit(pk, 2, 3)
    assert prover.commitment_vec[0] == oracle.scalar_mul(R, 8)  # main.rs: V0 = 8*g


# ---------------------------------------------------------------------------------- MulVec::calculate
def _msm_case(api, ctx, oracle, scalars, mults):
    pts = [oracle.scalar_mul(R, k % ELL) for k in mults]
    mv = api.MulVec(R, ctx)
    mv.add_scalars(scalars)
    mv.add_points(pts)
    want = oracle.scalar_mul(R, sum(s * k for s, k in zip(scalars, mults)) % ELL)
    assert mv.calculate() == want
    return pts, want


def test_msm_small_and_adversarial(api, ctx, oracle):
    rng = random.Random(6)
    for k in (1, 2, 3, 17, 100, 531):
        sc = [rng.randrange(ELL) for _ in range(k)]
        ml = [rng.randrange(1, ELL) for _ in range(k)]
        pts, want = _msm_case(api, ctx, oracle, sc, ml)
        if k <= 100:  # literal MulVec restatement agrees too
            assert oracle.msm_naive(R, b"".join(le32(s) for s in sc), b"".join(pts)) == want
    k = 64
    base = [rng.randrange(1, ELL) for _ in range(k)]
    cases = {
        "duplicate bases": ([rng.randrange(ELL) for _ in range(k)], [base[0]] * k),
        "negated bases": ([5] * k, [base[i // 2] * (1 if i % 2 == 0 else -1) for i in range(k)]),
        "zero scalars": ([0] * k, base),
        "order minus one": ([ELL - 1] * k, base),
        "all equal scalars": ([rng.randrange(ELL)] * k, base),
        "small scalars": ([rng.randrange(2**16) for _ in range(k)], base),
        "quarter zero": ([0 if i % 4 == 0 else rng.randrange(ELL) for i in range(k)], base),
        "identity points": ([rng.randrange(ELL) for _ in range(k)], [0 if i % 3 == 0 else base[i] for i in range(k)]),
        "reference generators": ([rng.randrange(ELL) for _ in range(k)], [3 * (i + 1) if i < 32 else 5 * (i - 31) for i in range(k)]),
    }
    for name, (sc, ml) in cases.items():
        _msm_case(api, ctx, oracle, sc, ml)
    assert api.MulVec(R, ctx).calculate() == bytes(32)  # empty MulVec = Point::zero()
    mv = api.MulVec(R, ctx)
    mv.add_scalar(1)
    with pytest.raises(RuntimeError):
        mv.calculate()  # mulvec.rs:23-25
    mv = api.MulVec(R, ctx)
    mv.add_scalar(1)
    mv.add_point((1).to_bytes(32, "little"))  # invalid encoding
    with pytest.raises(api.BppError) as ei:
        mv.calculate()
    assert ei.value.rc == 2


@pytest.mark.parametrize("logn", [10, 14, 17])
def test_msm_large_device_generated_points(api, ctx, oracle, logn):
    """Config 5 shape: P_i = k_i*g generated on device, result must equal (sum s_i k_i)*g."""
    n = 1 << logn
    ks = [pyref.prng_scalar(0xB2000002, i, ELL) for i in range(n)]
    ss = [pyref.prng_scalar(0xB2000001, i, ELL) for i in range(n)]
    pts = api.Points.from_multipliers(b"".join(le32(k) for k in ks), R, ctx)
    assert len(pts) == n
    s2 = [pyref.prng_scalar(0xB2000011, i, ELL) for i in range(3 * 256)]
    for tables in (False, True):  # generic Pippenger, then the same calls against window tables of the same points
        if tables:
            pts.precompute()
        got = pts.msm(b"".join(le32(s) for s in ss))[0]
        assert got == oracle.scalar_mul(R, sum(s * k for s, k in zip(ss, ks)) % ELL)
        # several jobs against the same table; and the raw partial-sum path used for the multi-GPU split
        jobs = 3
        got = pts.msm(b"".join(le32(s) for s in s2), k=256, n_jobs=jobs)
        for j in range(jobs):
            assert got[j] == oracle.scalar_mul(R, sum(s * k for s, k in zip(s2[j * 256:(j + 1) * 256], ks)) % ELL)
        raw = pts.msm_raw(b"".join(le32(s) for s in ss[:512]), 512) + pts.msm_raw(b"".join(le32(s) for s in ss[:7]), 7)
        want = (sum(s * k for s, k in zip(ss[:512], ks)) + sum(s * k for s, k in zip(ss[:7], ks))) % ELL
        assert api.points_sum_raw(raw, R, ctx) == oracle.scalar_mul(R, want)


def test_msm_degenerate_distributions(api, ctx, oracle):
    n = 1 << 12
    ks = [pyref.prng_scalar(0xB2000002, i, ELL) for i in range(n)]
    pts = api.Points.from_multipliers(b"".join(le32(k) for k in ks), R, ctx)
    rng = random.Random(8)
    s_eq = rng.randrange(ELL)
    dists = ([s_eq] * n, [rng.randrange(2**16) for _ in range(n)], [0 if i % 4 == 0 else rng.randrange(ELL) for i in range(n)],
             [ELL - 1] * n, [0] * n)
    same = api.Points.from_multipliers(le32(ks[0]) * n, R, ctx)
    with_identity = api.Points.from_multipliers(b"".join(le32(0 if i % 5 == 0 else k) for i, k in enumerate(ks)), R, ctx)
    s_full = [rng.randrange(ELL) for _ in range(n)]
    for tables in (False, True):
        if tables:
            pts.precompute(); same.precompute(); with_identity.precompute()
        for ss in dists:
            assert pts.msm(b"".join(le32(s) for s in ss))[0] == oracle.scalar_mul(R, sum(s * k for s, k in zip(ss, ks)) % ELL)
        assert same.msm(b"".join(le32(s) for s in s_full))[0] == oracle.scalar_mul(R, sum(s_full) * ks[0] % ELL)
        want = sum(s * k for i, (s, k) in enumerate(zip(s_full, ks)) if i % 5) % ELL
        assert with_identity.msm(b"".join(le32(s) for s in s_full))[0] == oracle.scalar_mul(R, want)


# ---------------------------------------------------------------------------------- RangeProof::prove / verify
def _inputs(n, m, seed):
    rng = random.Random(seed)
    return [rng.randrange(0, 2**min(n, 31)) for _ in range(m)], [rng.randrange(ELL) for _ in range(m)]


# the reference's intended grid (range/mod.rs:520-566): n in {32, 64} x m in {1, 2, 4, 8} — all eight cells — plus small shapes and
# the BASELINE.json shapes (32, 16), (64, 16); (64, 64) is in test_gpu_full_size.py
@pytest.mark.parametrize("n,m", [(8, 1), (8, 4), (32, 1), (32, 2), (32, 4), (32, 8), (64, 1), (64, 2), (64, 4), (64, 8), (32, 16), (64, 16)])
def test_range_prove_is_byte_identical(api, ctx, oracle, n, m):
    pk = api.PublicKey.new(n * m, R, ctx)
    vs, gs = ([2, 5], [3, 7]) if (n, m) == (64, 2) else _inputs(n, m, 1000 + n * 100 + m)  # (64,2) = src/main.rs
    prover = api.RangeProver()
    for v, g in zip(vs, gs):
        prover.commit(pk, v, g)
    proof = api.RangeProof.prove(pk, n, prover)
    want = oracle.range_prove(R, n, m, vs, gs)
    assert proof.to_bytes() == want
    assert b"".join(prover.commitment_vec) == oracle.commit(R, vs, gs)
    proof.verify(pk, n, prover.commitment_vec)  # main.rs:56 — Ok(())
    assert oracle.range_verify(R, n, m, b"".join(prover.commitment_vec), proof.to_bytes()) == 0


def test_main_rs_golden_fixture(api, ctx):
    """Committed fixture (tests/golden/make_fixtures.py): the src/main.rs proof on Ristretto255."""
    fx = json.load(open(os.path.join(GOLDEN, "ristretto_proofs.json")))["main_rs"]
    pk = api.PublicKey.new(128, R, ctx)
    proof = api.prove_batch(pk, 64, 2, [2, 5], [3, 7])[0]
    assert proof.hex() == fx["proof"]
    V = pk.commit_batch([2, 5], [3, 7])
    assert b"".join(V).hex() == fx["V"]
    assert api.verify_batch(pk, 64, 2, [proof], [V]) == [0]
    # the proof points are dlog*g for the discrete logs of SURVEY KAT set 2
    exp = ec.main_rs_kat(R)
    assert exp.kat_digest() == fx["dlog_digest"]


def _tampers(oracle, n, m, proof, V, pb=32):
    """The tamper cycle of SURVEY §8(d) config 4."""
    rounds = (n * m).bit_length() - 1
    g = oracle.scalar_mul(R, 1)

    def bump_point(buf, off):
        return buf[:off] + oracle.point_add(R, buf[off:off + pb], g) + buf[off + pb:]

    def bump_scalar(buf, off):
        v = (int.from_bytes(buf[off:off + 32], "little") + 1) % ELL
        return buf[:off] + le32(v) + buf[off + 32:]

    sc0 = (2 * rounds + 3) * pb
    out = [
        ("A+g", bump_point(proof, 0), V), ("L0+g", bump_point(proof, pb), V), ("R_last+g", bump_point(proof, 2 * rounds * pb), V),
        ("wipA+g", bump_point(proof, (2 * rounds + 1) * pb), V), ("wipB+g", bump_point(proof, (2 * rounds + 2) * pb), V),
        ("r'+1", bump_scalar(proof, sc0), V), ("s'+1", bump_scalar(proof, sc0 + 32), V), ("d'+1", bump_scalar(proof, sc0 + 64), V),
        ("V+g", proof, bump_point(V, 0)),
    ]
    return out


@pytest.mark.parametrize("n,m", [(8, 1), (8, 4), (64, 1), (64, 2)])
def test_range_verify_verdicts_match_oracle(api, ctx, oracle, n, m):
    pk = api.PublicKey.new(n * m, R, ctx)
    vs, gs = _inputs(n, m, 2000 + n * 100 + m)
    proof = oracle.range_prove(R, n, m, vs, gs)
    V = oracle.commit(R, vs, gs)
    cases = [("valid", proof, V)] + _tampers(oracle, n, m, proof, V)
    # format errors: non-canonical scalar, invalid point encodings
    sc0 = (2 * ((n * m).bit_length() - 1) + 3) * 32
    cases.append(("r' = ell", proof[:sc0] + le32(ELL) + proof[sc0 + 32:], V))
    cases.append(("A invalid", (1).to_bytes(32, "little") + proof[32:], V))
    cases.append(("V invalid", proof, pyref.P25519.to_bytes(32, "little") + V[32:]))
    if n < 31:  # out-of-range value: honest prover, proof must not verify
        vs2 = [2**n] + vs[1:]
        cases.append(("out of range", oracle.range_prove(R, n, m, vs2, gs), oracle.commit(R, vs2, gs)))
    got = api.verify_batch(pk, n, m, [c[1] for c in cases], [[c[2][i * 32:(i + 1) * 32] for i in range(m)] for c in cases])
    want = [oracle.range_verify(R, n, m, c[2], c[1]) for c in cases]
    assert got == want, [(c[0], g_, w) for c, g_, w in zip(cases, got, want) if g_ != w]
    assert got[0] == 0 and all(v != 0 for v in got[1:])
    # the single-proof API raises like the reference returns Err
    rp = api.RangeProof.from_bytes(cases[1][1], 32)
    with pytest.raises(api.VerificationError):
        rp.verify(pk, n, [V[i * 32:(i + 1) * 32] for i in range(m)])


def test_custom_challenges_and_nonces(api, ctx, oracle):
    """Caller-supplied Fiat-Shamir values (the host keeps the transcript): same bytes as the oracle."""
    n, m = 16, 4
    rounds = 6
    rng = random.Random(77)
    vals = dict(y=rng.randrange(1, ELL), z=rng.randrange(1, ELL), e_final=rng.randrange(1, ELL), alpha=rng.randrange(ELL),
                r=rng.randrange(ELL), s=rng.randrange(ELL), delta=rng.randrange(ELL), eta=rng.randrange(ELL),
                e_round=[rng.randrange(1, ELL) for _ in range(rounds)], d_L=[rng.randrange(ELL) for _ in range(rounds)],
                d_R=[rng.randrange(ELL) for _ in range(rounds)])
    bp = api.make_params(**vals)
    op = oracle_lib.make_params(oracle_lib.OrcParams, **vals)
    pk = api.PublicKey.new(n * m, R, ctx)
    vs, gs = _inputs(n, m, 4242)
    proof = api.prove_batch(pk, n, m, vs, gs, bp)[0]
    assert proof == oracle.range_prove(R, n, m, vs, gs, op)
    V = pk.commit_batch(vs, gs)
    assert api.verify_batch(pk, n, m, [proof], [V], bp) == [0]
    assert api.verify_batch(pk, n, m, [proof], [V]) == [1]  # wrong challenges
    assert oracle.range_verify(R, n, m, b"".join(V), proof, op) == 0
    # exponent-domain checker agrees on the scalars
    exp = ec.Checker(ELL).range_prove(n, m, vs, gs, ec.Params(**vals))
    rp = api.RangeProof.from_bytes(proof, 32)
    assert (rp.proof.r_prime, rp.proof.s_prime, rp.proof.d_prime) == (exp.r_prime, exp.s_prime, exp.d_prime)


def test_challenge_table_reuse_across_calls(api, ctx, oracle):
    """The library keeps the challenge-only tables of the previous call when parameters and shape repeat. Interleave
    repeats, parameter changes, shape changes and prove/verify on ONE context: every result must still match the oracle."""
    rng = random.Random(99)
    rounds = 6
    vals = dict(y=rng.randrange(1, ELL), z=rng.randrange(1, ELL), e_final=rng.randrange(1, ELL), alpha=rng.randrange(ELL),
                r=rng.randrange(ELL), s=rng.randrange(ELL), delta=rng.randrange(ELL), eta=rng.randrange(ELL),
                e_round=[rng.randrange(1, ELL) for _ in range(rounds)], d_L=[rng.randrange(ELL) for _ in range(rounds)],
                d_R=[rng.randrange(ELL) for _ in range(rounds)])
    bp, op = api.make_params(**vals), oracle_lib.make_params(oracle_lib.OrcParams, **vals)
    shapes = {(16, 4): api.PublicKey.new(64, R, ctx), (8, 8): api.PublicKey.new(64, R, ctx), (32, 1): api.PublicKey.new(32, R, ctx)}
    data = {}
    for (n, m) in shapes:
        vs, gs = _inputs(n, m, 31 * n + m)
        data[(n, m)] = (vs, gs, oracle.commit(R, vs, gs))
    script = [((16, 4), None), ((16, 4), None), ((16, 4), "custom"), ((16, 4), "custom"), ((8, 8), "custom"), ((8, 8), None),
              ((16, 4), None), ((32, 1), None), ((32, 1), None), ((16, 4), "custom"), ((16, 4), None)]
    for step, (shape, which) in enumerate(script):
        n, m = shape
        if m == 1 and which:
            continue
        pk = shapes[shape]
        vs, gs, V = data[shape]
        b_par, o_par = (bp, op) if which else (None, None)
        want = oracle.range_prove(R, n, m, vs, gs, o_par) if o_par else oracle.range_prove(R, n, m, vs, gs)
        Vl = [[V[i * 32:(i + 1) * 32] for i in range(m)]]
        if step % 3 != 1:  # prove on most steps: it overwrites the verifier's tables
            assert api.prove_batch(pk, n, m, vs, gs, b_par)[0] == want, (step, shape, which)
        assert api.verify_batch(pk, n, m, [want], Vl, b_par) == [0], (step, shape, which)
        assert api.verify_batch(pk, n, m, [want], Vl, b_par) == [0], (step, shape, which)  # immediate repeat: tables reused
        other = None if which else bp
        if m > 1:
            assert api.verify_batch(pk, n, m, [want], Vl, other) == [1], (step, shape, which)  # other challenges must reject


def test_batch_prove_and_verify_many(api, ctx, oracle):
    """Config 4 shape at reduced count: independent n=64, m=1 proofs, some tampered, verdict per proof."""
    n, m, count = 64, 1, 96
    pk = api.PublicKey.new(n, R, ctx)
    vs = [pyref.prng_scalar(0xB2000003, i, 2**31) for i in range(count)]
    gs = [pyref.prng_scalar(0xB2000004, i, ELL) for i in range(count)]
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    for i in (0, 1, count - 1):
        assert proofs[i] == oracle.range_prove(R, n, m, [vs[i]], [gs[i]])
    tampered = {}
    kinds = None
    for j in range(0, count, 7):
        kinds = _tampers(oracle, n, m, proofs[j], Vs[j])
        name, p2, v2 = kinds[(j // 7) % len(kinds)]
        tampered[j] = name
        proofs[j], Vs[j] = p2, v2
    got = api.verify_batch(pk, n, m, proofs, [[v] for v in Vs])
    want = [oracle.range_verify(R, n, m, Vs[i], proofs[i]) for i in range(count)]
    assert got == want
    assert [i for i, v in enumerate(got) if v != 0] == sorted(tampered)


# ---------------------------------------------------------------------------------- standalone WIP (config 2 shape)
@pytest.mark.parametrize("N", [2, 64, 1024])
def test_wip_prove_verify(api, ctx, oracle, N):
    rng = random.Random(N)
    y = 12
    a = [rng.randrange(ELL) for _ in range(N)]
    b = [rng.randrange(ELL) for _ in range(N)]
    ypow = [pow(y, i + 1, ELL) for i in range(N)]
    gamma = rng.randrange(ELL)
    pk = api.PublicKey.new(N, R, ctx)
    proof = api.WeightedInnerProductProof.prove(pk, a, b, ypow, gamma)
    enc = lambda v: b"".join(le32(x) for x in v)
    want = oracle.wip_prove(R, N, enc(a), enc(b), enc(ypow), gamma)
    assert proof.to_bytes() == want
    # commitment P = sum a_i G_i + sum b_i H_i + (sum a_i b_i y^(i+1)) g + gamma h  (comment at wip.rs:229-237)
    dl = (sum(a[i] * 3 * (i + 1) + b[i] * 5 * (i + 1) + a[i] * b[i] * ypow[i] for i in range(N)) + 2 * gamma) % ELL
    P = oracle.scalar_mul(R, dl)
    zeros = [0] * N
    proof.verify(pk, ypow, zeros, zeros, 0, [], P, [])
    assert oracle.wip_verify(R, N, enc(ypow), enc(zeros), enc(zeros), 0, b"", P, b"", want) == 0
    with pytest.raises(api.VerificationError):
        proof.verify(pk, ypow, zeros, zeros, 0, [], oracle.scalar_mul(R, (dl + 1) % ELL), [])
    with pytest.raises(AssertionError):
        api.WeightedInnerProductProof.prove(pk, a[:-1], b, ypow, gamma)  # wip.rs:61-64


def test_reference_panics_are_errors(api, ctx):
    pk = api.PublicKey.new(24, R, ctx)  # not a power of two
    with pytest.raises(api.BppError) as ei:
        api.prove_batch(pk, 8, 3, [1, 2, 3], [1, 2, 3])
    assert ei.value.rc == 4  # wip.rs:67
    pk = api.PublicKey.new(16, R, ctx)
    with pytest.raises(AssertionError):
        api.prove_batch(pk, 8, 4, [1, 2, 3, 4], [1, 2, 3, 4])  # range/mod.rs:252-253


@pytest.mark.parametrize("name", ["n64_m16", "n32_m16", "n64_m1"])
def test_golden_fixture_proofs(api, ctx, name):
    fx = json.load(open(os.path.join(GOLDEN, "ristretto_proofs.json")))[name]
    n, m = fx["n"], fx["m"]
    pk = api.PublicKey.new(n * m, R, ctx)
    gs = [int(g, 16) for g in fx["gamma"]]
    assert api.prove_batch(pk, n, m, fx["v"], gs)[0].hex() == fx["proof"]
    V = pk.commit_batch(fx["v"], gs)
    assert b"".join(V).hex() == fx["V"]
    assert api.verify_batch(pk, n, m, [bytes.fromhex(fx["proof"])], [V]) == [0]


# ---------------------------------------------------------------------------------- multi-GPU decomposition, one GPU
@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_msm_point_range_split(api, ctx, oracle, world):
    """The N-range split of one MSM: per-"rank" partial sums in the raw device form + bpp_points_sum_raw equal the
    unsplit result and the oracle's (sum s_i k_i) * g. The ranks are emulated as slices on one GPU."""
    from bulletproofsplus_b200 import parallel
    rng = random.Random(world)
    k = 1500
    mult = [rng.randrange(ELL) for _ in range(k)]
    sc = [rng.randrange(ELL) for _ in range(k)]
    want = oracle.scalar_mul(R, sum(a * b for a, b in zip(mult, sc)) % ELL)
    raws = b""
    for r in range(world):
        lo, hi = parallel.shard_range(k, r, world)
        pts = api.Points.from_multipliers(b"".join(le32(x) for x in mult[lo:hi]), R, ctx)
        raws += pts.msm_raw(b"".join(le32(x) for x in sc[lo:hi]))
    assert api.points_sum_raw(raws, R, ctx) == want
    whole = api.Points.from_multipliers(b"".join(le32(x) for x in mult), R, ctx)
    assert whole.msm(b"".join(le32(x) for x in sc))[0] == want


def test_sharded_verify_blocks_match_whole_batch(api, ctx, oracle):
    """Block-sharding a batch of proofs (parallel.verify_batch_sharded without a process group = world 1, and manual
    2-way blocks) gives the verdicts of the whole batch."""
    from bulletproofsplus_b200 import parallel
    n, m, count = 16, 2, 7
    pk = api.PublicKey.new(n * m, R, ctx)
    rng = random.Random(9)
    vs = [rng.randrange(2**n) for _ in range(count * m)]
    gs = [rng.randrange(ELL) for _ in range(count * m)]
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    bad = bytearray(proofs[5]); bad[40] ^= 4
    proofs[5] = bytes(bad)
    whole = api.verify_batch(pk, n, m, proofs, Vl)
    assert whole == [oracle.range_verify(R, n, m, b"".join(Vl[i]), proofs[i]) for i in range(count)]
    assert parallel.verify_batch_sharded(lambda p, c: api.verify_batch(pk, n, m, list(p), list(c)), proofs, Vl) == whole
    blocks = []
    for r in range(2):
        lo, hi = parallel.shard_range(count, r, 2)
        blocks += api.verify_batch(pk, n, m, proofs[lo:hi], Vl[lo:hi])
    assert blocks == whole

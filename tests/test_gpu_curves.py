"""GPU parity for the two short-Weierstrass instantiations of the same kernel templates (run on the B200 box):
secp256k1 (curve 1) and BLS12-381 G1 (curve 2 — the group the reference snapshot actually runs on).
Every case goes through the C ABI and is compared bit-for-bit with the CPU oracle; the curve constants are pinned
against the reference's own golden vectors (secp256k1.rs:151-160, point.rs:16) in tests/test_oracle_kat.py."""
import random

import pytest

pytestmark = pytest.mark.gpu

ORD = {
    1: 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141,
    2: 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001,
}
FP = {
    1: 2**256 - 2**32 - 977,
    2: 0x1A0111EA397FE69A4B1BA7B6434BACD764774B84F38512BF6730D2A0F6B0F6241EABFFFEB153FFFFB9FEFFFFFFFFAAAB,
}
FB = {1: 32, 2: 48}
CURVES = [1, 2]


@pytest.fixture(scope="module")
def api():
    from bulletproofsplus_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    return api.Arith.init()


def le32(x):
    return int(x).to_bytes(32, "little")


@pytest.mark.parametrize("R", CURVES)
def test_field_ops(ctx, R):
    p, fb = FP[R], FB[R]
    rng = random.Random(R)
    edge = [0, 1, 2, 3, p - 1, p - 2, 2**32, 2**(8 * fb - 8) % p, (p + 1) // 2]
    # operands that push the reductions' carries: just below p and 2^k, all-ones halves (secp256k1 folds the high half of the
    # product with 2^256 = 2^32 + 977: fp_mont.cuh k1_reduce)
    edge += [p - 977, p - 2**32, p - 2**32 - 977, (2**(8 * fb) - 1) % p, 2**(8 * fb - 1) % p, 2**(4 * fb) - 1, ((2**(4 * fb) - 1) << (4 * fb - 1)) % p,
             2**(8 * fb - 32) % p, (p - 1) // 2, (p - 1) // 3]
    xs = edge + [rng.randrange(p) for _ in range(200)]
    ys = list(reversed(edge)) + [rng.randrange(p) for _ in range(200)]
    a, b = b"".join(x.to_bytes(fb, "little") for x in xs), b"".join(y.to_bytes(fb, "little") for y in ys)
    for op, fn in ((0, lambda x, y: x * y), (1, lambda x, y: x + y), (2, lambda x, y: x - y), (3, lambda x, y: pow(x, p - 2, p)),
                   (4, lambda x, y: x * x)):
        out, _ = ctx.debug_op(R, op, a, b, fb)
        for i, (x, y) in enumerate(zip(xs, ys)):
            assert int.from_bytes(out[fb * i:fb * i + fb], "little") == fn(x, y) % p, (R, op, i)


@pytest.mark.parametrize("R", CURVES)
def test_scalar_field_ops(ctx, R):
    q = ORD[R]
    rng = random.Random(10 + R)
    edge = [0, 1, 2, q - 1, q - 2, 7, 99, 2**255 % q]
    xs = edge + [rng.randrange(q) for _ in range(100)]
    ys = list(reversed(edge)) + [rng.randrange(q) for _ in range(100)]
    a, b = b"".join(le32(x) for x in xs), b"".join(le32(y) for y in ys)
    for op, fn in ((20, lambda x, y: x * y), (21, lambda x, y: pow(x, q - 2, q)), (22, lambda x, y: x + y), (23, lambda x, y: x - y)):
        out, _ = ctx.debug_op(R, op, a, b, 32)
        for i, (x, y) in enumerate(zip(xs, ys)):
            assert int.from_bytes(out[32 * i:32 * i + 32], "little") == fn(x, y) % q, (R, op, i)


@pytest.mark.parametrize("R", CURVES)
def test_group_ops_with_exceptional_cases(ctx, oracle, R):
    """P+Q, 2P, P-Q through the mixed and the full addition, including P+P, P+(-P), O+P, P+O (the branches the
    XYZZ formulas need and the reference's generators force)."""
    q, pb = ORD[R], 2 * FB[R]
    rng = random.Random(20 + R)
    ks = [0, 1, 2, 3, 5, 15, q - 1, 7, 7, 0] + [rng.randrange(q) for _ in range(20)]
    js = [1, 1, 2, q - 3, 5, 0, 1, 7, q - 7, 0] + [rng.randrange(q) for _ in range(20)]
    A = [oracle.scalar_mul(R, k) for k in ks]
    B = [oracle.scalar_mul(R, j) for j in js]
    a, b = b"".join(A), b"".join(B)
    exp = {10: A, 11: [oracle.scalar_mul(R, (k + j) % q) for k, j in zip(ks, js)], 12: [oracle.scalar_mul(R, 2 * k % q) for k in ks],
           14: [oracle.scalar_mul(R, (k - j) % q) for k, j in zip(ks, js)]}
    exp[13] = exp[11]
    for op, want in exp.items():
        out, ok = ctx.debug_op(R, op, a, b, pb)
        assert all(ok), (R, op)
        assert [out[pb * i:pb * i + pb] for i in range(len(ks))] == want, (R, op)
    # invalid encodings (off-curve, non-canonical coordinate) are flagged exactly where the oracle flags them
    good = A[3]
    cands = [good[:-1] + bytes([good[-1] ^ 1]), b"\xff" * pb, good[:FB[R]] + bytes(FB[R]), bytes(pb), good]
    _, ok = ctx.debug_op(R, 10, b"".join(cands), b"".join([good] * len(cands)), pb)
    assert [bool(o) for o in ok] == [oracle.point_decode_ok(R, c) for c in cands]


@pytest.mark.parametrize("R", CURVES)
def test_public_key_and_commitments(api, ctx, oracle, R):
    q = ORD[R]
    for length in (1, 8):
        pk = api.PublicKey.new(length, R, ctx)
        g, h, G, H = oracle.pk_new(R, length)
        assert pk.g == g and pk.h == h
        assert b"".join(pk.G_vec) == G and b"".join(pk.H_vec) == H
    pk = api.PublicKey.new(4, R, ctx)
    rng = random.Random(30 + R)
    vs = [0, 1, 2, 5, 2**31 - 1, 2**31, 2**32 + 9, 2**64 - 1] + [rng.randrange(2**31) for _ in range(24)]
    gs = [0, 1, 3, 7, q - 1, 11, 13, 17] + [rng.randrange(q) for _ in range(24)]
    assert b"".join(pk.commit_batch(vs, gs)) == oracle.commit(R, vs, gs)


@pytest.mark.parametrize("R", CURVES)
def test_msm(api, ctx, oracle, R):
    q = ORD[R]
    rng = random.Random(40 + R)

    def case(scalars, mults):
        pts = [oracle.scalar_mul(R, k % q) for k in mults]
        mv = api.MulVec(R, ctx)
        mv.add_scalars(scalars)
        mv.add_points(pts)
        assert mv.calculate() == oracle.scalar_mul(R, sum(s * k for s, k in zip(scalars, mults)) % q)

    for k in (1, 2, 17, 300):
        case([rng.randrange(q) for _ in range(k)], [rng.randrange(1, q) for _ in range(k)])
    k = 48
    base = [rng.randrange(1, q) for _ in range(k)]
    case([rng.randrange(q) for _ in range(k)], [base[0]] * k)                                   # duplicate bases
    case([5] * k, [base[i // 2] * (1 if i % 2 == 0 else -1) for i in range(k)])                  # P + (-P)
    case([0] * k, base)                                                                          # zero scalars
    case([q - 1] * k, base)                                                                      # order - 1
    case([rng.randrange(q)] * k, base)                                                           # all equal
    case([rng.randrange(q) for _ in range(k)], [3 * (i + 1) for i in range(k // 2)] + [5 * (i + 1) for i in range(k // 2)])  # F5 generators
    # a device-generated table, 4096 points, against the exponent identity (sum s_i k_i) * g
    n = 4096
    mult = [rng.randrange(q) for _ in range(n)]
    sc = [rng.randrange(q) for _ in range(n)]
    pts = api.Points.from_multipliers(b"".join(le32(x) for x in mult), R, ctx)
    want = oracle.scalar_mul(R, sum(a * b for a, b in zip(mult, sc)) % q)
    assert pts.msm(b"".join(le32(x) for x in sc))[0] == want
    assert pts.precompute().msm(b"".join(le32(x) for x in sc))[0] == want  # window tables: same point, no doublings
    # short jobs against the 4096-point window tables: the engine splits every window digit into 3 (k <= N_B/64) or 2
    # (k <= N_B/9) signed sub-digits that share the table point (msm.cuh, msm_split_digit); scalars above order/2 are folded
    for k, jobs in ((5, 3), (40, 2), (200, 3), (450, 1), (700, 2)):
        ss = [rng.randrange(q) for _ in range(k * jobs)]
        ss[0], ss[k - 1] = q - 1, (q + 1) // 2
        got = pts.msm(b"".join(le32(x) for x in ss), k=k, n_jobs=jobs)
        for j in range(jobs):
            assert got[j] == oracle.scalar_mul(R, sum(a * b for a, b in zip(mult[:k], ss[j * k:(j + 1) * k])) % q), (k, j)
    dup = api.Points.from_multipliers(b"".join(le32(mult[i // 2] if i % 2 == 0 else (q - mult[i // 2])) for i in range(256)), R, ctx).precompute()
    s256 = sc[:256]
    want = sum(s * (mult[i // 2] if i % 2 == 0 else q - mult[i // 2]) for i, s in enumerate(s256)) % q
    assert dup.msm(b"".join(le32(x) for x in s256))[0] == oracle.scalar_mul(R, want)  # P, -P pairs: exceptional additions in the buckets


@pytest.mark.parametrize("R", CURVES)
@pytest.mark.parametrize("n,m", [(8, 1), (8, 4), (64, 2), (32, 8), (64, 8)])
def test_range_prove_and_verify(api, ctx, oracle, R, n, m):
    """Byte-identical proofs and identical verdicts (valid + one tamper per proof field) on the reference's own group.
    N = 256 and 512 take the WIP prover's thread-block clusters (2 and 4 blocks per proof) on every curve's scalar field."""
    q, pb = ORD[R], 2 * FB[R]
    rng = random.Random(100 * R + n + m)
    vs = [rng.randrange(2**min(n, 31)) for _ in range(m)]
    gs = [rng.randrange(q) for _ in range(m)]
    pk = api.PublicKey.new(n * m, R, ctx)
    proof = api.prove_batch(pk, n, m, vs, gs)[0]
    assert proof == oracle.range_prove(R, n, m, vs, gs)
    V = pk.commit_batch(vs, gs)
    Vb = b"".join(V)
    assert api.verify_batch(pk, n, m, [proof], [V]) == [0] == [oracle.range_verify(R, n, m, Vb, proof)]
    L = (n * m).bit_length() - 1
    g1 = oracle.scalar_mul(R, 1)
    tampered = []
    for k in (0, 1, 1 + L, 2 * L + 1, 2 * L + 2):  # A, L[0], R[0], wipA, wipB  += g
        t = bytearray(proof)
        t[k * pb:(k + 1) * pb] = oracle.point_add(R, bytes(t[k * pb:(k + 1) * pb]), g1)
        tampered.append(bytes(t))
    for j in range(3):  # r', s', delta' += 1
        t = bytearray(proof)
        off = (2 * L + 3) * pb + 32 * j
        t[off:off + 32] = le32((int.from_bytes(t[off:off + 32], "little") + 1) % q)
        tampered.append(bytes(t))
    got = api.verify_batch(pk, n, m, tampered, [V] * len(tampered))
    assert got == [oracle.range_verify(R, n, m, Vb, t) for t in tampered] == [1] * len(tampered)
    # wrong commitment
    V2 = [oracle.point_add(R, V[0], g1)] + V[1:]
    assert api.verify_batch(pk, n, m, [proof], [V2]) == [1]


def test_main_rs_on_the_reference_group(api, ctx, oracle):
    """src/main.rs:6-57 (n=64, m=2, v={2,5}, gamma={3,7}) on BLS12-381 G1 — the snapshot's configuration end to end."""
    R = 2
    pk = api.PublicKey.new(128, R, ctx)
    prover = api.RangeProver()
    prover.commit(pk, 2, 3)
    prover.commit(pk, 5, 7)
    assert prover.commitment_vec == [oracle.scalar_mul(R, 8), oracle.scalar_mul(R, 19)]  # SURVEY §8(c) KAT set 2
    proof = api.RangeProof.prove(pk, 64, prover)
    assert proof.to_bytes() == oracle.range_prove(R, 64, 2, [2, 5], [3, 7])
    proof.verify(pk, 64, prover.commitment_vec)  # assert_eq!(result, Ok(())) main.rs:56
    assert proof.proof.B == oracle.scalar_mul(R, 0x4506)  # wip.B = 0x4506 * g for every modulus


@pytest.mark.parametrize("R", CURVES)
def test_wip_prove_verify(api, ctx, oracle, R):
    q = ORD[R]
    N = 16
    rng = random.Random(50 + R)
    pk = api.PublicKey.new(N, R, ctx)
    a = [rng.randrange(q) for _ in range(N)]
    b = [rng.randrange(q) for _ in range(N)]
    ypow = [pow(12, i + 1, q) for i in range(N)]
    gamma = rng.randrange(q)
    enc = lambda v: b"".join(le32(x) for x in v)
    proof = api.WeightedInnerProductProof.prove(pk, a, b, ypow, gamma)
    assert proof.to_bytes() == oracle.wip_prove(R, N, enc(a), enc(b), enc(ypow), gamma)

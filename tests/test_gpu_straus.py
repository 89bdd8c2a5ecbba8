"""Per-proof points through msm_straus_kernel (csrc/msm_straus.cuh): taken for batches of >= 16 proofs whose per-call point
count is <= 48 (3 + 2 log2(n m) + m points per range proof). Verdicts of such batches must equal the CPU oracle's verdict for
every proof AND the library's own verdict for the same proof verified alone (one job: the bucket pipeline) — for every shape of
the reference's test grid that fits, with commitments that are the identity point (v = 0, gamma = 0), an identity point placed
in the proof, scalars 0 and ell - 1 in the proof, and the tamper cycle of SURVEY §8(d) config 4. Shapes with more than 48
per-call points (m = 32: 57) keep the bucket pipeline at any batch size and serve as the cross-check of the dispatch."""
import pytest

from tests import pyref

pytestmark = pytest.mark.gpu

ELL = 2**252 + 27742317777372353535851937790883648493
R = 0


@pytest.fixture(scope="module")
def env():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from bulletproofsplus_b200 import api
    return api, api.Arith.init()


def le32(x):
    return int(x).to_bytes(32, "little")


def _variants(oracle, n, m, proof, V):
    """(name, proof, commitments) variants of one valid proof; the oracle decides what each of them is worth."""
    rounds = (n * m).bit_length() - 1
    pb = 32
    g = oracle.scalar_mul(R, 1)
    ident = bytes(32)  # the canonical Ristretto255 encoding of the identity
    sc0 = (2 * rounds + 3) * pb

    def put(buf, off, val):
        return buf[:off] + val + buf[off + len(val):]

    out = [("valid", proof, V),
           ("A+g", put(proof, 0, oracle.point_add(R, proof[:pb], g)), V),
           ("L0 = identity", put(proof, pb, ident), V),
           ("R_last = identity", put(proof, 2 * rounds * pb, ident), V),
           ("wipB = identity", put(proof, (2 * rounds + 2) * pb, ident), V),
           ("r' = 0", put(proof, sc0, le32(0)), V),
           ("s' = ell - 1", put(proof, sc0 + 32, le32(ELL - 1)), V),
           ("d' + 1", put(proof, sc0 + 64, le32((int.from_bytes(proof[sc0 + 64:sc0 + 96], "little") + 1) % ELL)), V),
           ("V0 = identity", proof, [ident] + list(V[1:])),
           ("V_last + g", proof, list(V[:-1]) + [oracle.point_add(R, V[-1], g)])]
    return out


@pytest.mark.parametrize("n,m", [(8, 1), (8, 4), (32, 2), (32, 8), (64, 1), (64, 4), (64, 16), (16, 32)])
def test_batches_through_straus_match_oracle_and_single_proof_calls(env, oracle, n, m):
    api, ctx = env
    pk = api.PublicKey.new(n * m, R, ctx)
    base = 24  # distinct witnesses; >= 16 jobs with the variants below
    vs, gs = [], []
    for i in range(base):
        for j in range(m):
            if i == 3:  # every commitment of this proof is the identity point
                vs.append(0); gs.append(0)
            elif i == 4 and j == 0:
                vs.append(2**min(n, 31) - 1); gs.append(ELL - 1)  # below 2^31: clear of the reference's `v as i32` (prover.rs:37)
            else:
                vs.append(pyref.prng_scalar(0xB2000003, 1000 * n + 10 * i + j, 2**min(n, 31)))
                gs.append(pyref.prng_scalar(0xB2000004, 1000 * n + 10 * i + j, ELL))
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(base)]
    assert Vl[3] == [bytes(32)] * m
    batch_p, batch_V, names = [], [], []
    for i in range(base):
        var = _variants(oracle, n, m, proofs[i], Vl[i])
        name, p2, v2 = var[i % len(var)]
        batch_p.append(p2); batch_V.append(list(v2)); names.append(name)
    want = [oracle.range_verify(R, n, m, b"".join(batch_V[i]), batch_p[i]) for i in range(base)]
    assert any(w == 0 for w in want) and any(w != 0 for w in want)
    got = api.verify_batch(pk, n, m, batch_p, batch_V)
    assert got == want, [(i, names[i], got[i], want[i]) for i in range(base) if got[i] != want[i]]
    # the same proofs one at a time: one job each, i.e. the bucket pipeline for the per-proof points
    single = [api.verify_batch(pk, n, m, [batch_p[i]], [batch_V[i]])[0] for i in range(0, base, 5)]
    assert single == want[0:base:5]
    # per-proof challenges: the same dispatch behind bpp_range_verify_batch_fs
    L = (n * m).bit_length() - 1
    y, z = (7, 7) if m == 1 else (12, 23)
    chal = [[y, z, 99] + [7] * L] * base
    assert api.verify_batch_fs(pk, n, m, batch_p, batch_V, challenges=chal) == want
    # and with the key's direct tables (>= 32 jobs take them; the per-proof points still go through Straus)
    if n * m <= 256:
        assert pk.precompute(1 << 30) >= 8
        assert api.verify_batch(pk, n, m, batch_p * 2, batch_V * 2) == want * 2


def test_large_batch_per_proof_challenges_mixed_challenge_sets(env, oracle):
    """bpp_range_verify_batch_fs over 1100 proofs against a key with direct tables, where the proofs of the second part are
    produced with OTHER Fiat-Shamir values than those of the first (so every per-proof table really is per proof), with
    tampered proofs on both sides of the cut and of the middle: every verdict must match the tamper pattern, three times in a
    row on one context, and equal the verdicts of the same batch verified in pieces of 200. (Written for an in-library split
    of such batches over two stream pairs, which was measured no faster and dropped — DESIGN §9; the test stays.)"""
    import random
    api, ctx = env
    n, m, count, cut = 8, 2, 1100, 470
    L = 4
    rng = random.Random(5)
    vals = dict(y=rng.randrange(1, ELL), z=rng.randrange(1, ELL), e_final=rng.randrange(1, ELL), alpha=rng.randrange(ELL),
                r=rng.randrange(ELL), s=rng.randrange(ELL), delta=rng.randrange(ELL), eta=rng.randrange(ELL),
                e_round=[rng.randrange(1, ELL) for _ in range(L)], d_L=[rng.randrange(ELL) for _ in range(L)],
                d_R=[rng.randrange(ELL) for _ in range(L)])
    bp = api.make_params(**vals)
    pk = api.PublicKey.new(n * m, R, ctx)
    assert pk.precompute(1 << 28) >= 8
    vs = [rng.randrange(2**n) for _ in range(count * m)]
    gs = [rng.randrange(ELL) for _ in range(count * m)]
    proofs = []
    for c0 in range(0, cut, 256):
        c1 = min(cut, c0 + 256)
        proofs += api.prove_batch(pk, n, m, vs[c0 * m:c1 * m], gs[c0 * m:c1 * m])
    for c0 in range(cut, count, 256):
        c1 = min(count, c0 + 256)
        proofs += api.prove_batch(pk, n, m, vs[c0 * m:c1 * m], gs[c0 * m:c1 * m], bp)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    chal = [[12, 23, 99] + [7] * L] * cut + [[vals["y"], vals["z"], vals["e_final"]] + vals["e_round"]] * (count - cut)
    want = [0] * count
    for i in list(range(3, count, 97)) + [cut - 1, cut, count // 2 - 1, count // 2, count - 1]:
        p = bytearray(proofs[i])
        p[-1 - 32 * (i % 3)] ^= 1  # r', s' or delta' (stays canonical)
        proofs[i] = bytes(p)
        want[i] = 1
    # spot-check the expectation itself against the oracle on both sides of the cut
    from oracle import oracle_lib
    op = oracle_lib.make_params(oracle_lib.OrcParams, **vals)
    for i in (0, 3, cut - 1, cut - 2):
        assert oracle.range_verify(R, n, m, b"".join(Vl[i]), proofs[i]) == want[i]
    for i in (cut, cut + 1, count - 1, count - 2):
        assert oracle.range_verify(R, n, m, b"".join(Vl[i]), proofs[i], op) == want[i]
    for _ in range(3):
        assert api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=chal) == want
    pieces = []
    for c0 in range(0, count, 200):
        pieces += api.verify_batch_fs(pk, n, m, proofs[c0:c0 + 200], Vl[c0:c0 + 200], challenges=chal[c0:c0 + 200])
    assert pieces == want
    # swapped challenge sets must reject everything that was valid
    swapped = chal[cut:] + chal[:cut]
    if count - cut != cut:
        got = api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=swapped)
        assert all(g == 1 for g in got[:min(cut, count - cut)])

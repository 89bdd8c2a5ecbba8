"""Real Fiat-Shamir on the GPU path: interactive proving (bpp_range_prove_begin/_step/_finish) and batch verification with
per-proof challenges (bpp_range_verify_batch_fs), host-owned Merlin transcripts. The reference hard-codes its challenges, so
parity is anchored two ways: (1) with the reference's constants fed back, the session's bytes equal the one-shot prover's and
the oracle's; (2) with transcript challenges, every proof equals what the CPU oracle produces when handed that proof's
challenges and nonces as explicit parameters."""
import random

import pytest

from bulletproofsplus_b200 import merlin
from oracle import oracle_lib
from tests import pyref

pytestmark = pytest.mark.gpu

ELL = pyref.ELL
ORD = {0: ELL, 1: 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141,
       2: 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001}


@pytest.fixture(scope="module")
def api():
    from bulletproofsplus_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    return api.Arith.init()


def _inputs(n, m, count, seed, q):
    rng = random.Random(seed)
    return [rng.randrange(0, 2**min(n, 31)) for _ in range(count * m)], [rng.randrange(q) for _ in range(count * m)]


@pytest.mark.parametrize("n,m", [(8, 1), (8, 4), (64, 2), (64, 16)])
def test_session_with_reference_constants_equals_one_shot(api, ctx, oracle, n, m):
    count = 3
    L = (n * m).bit_length() - 1
    pk = api.PublicKey.new(n * m, 0, ctx)
    vs, gs = _inputs(n, m, count, 5 * n + m, ELL)
    ses = api.ProveSession(pk, n, m, vs, gs, api.REFERENCE_NONCES)
    y, z = (7, 7) if m == 1 else (12, 23)
    ch = [(y, z)] * count
    pts = [ses.A]
    for _ in range(L):
        pts.append(ses.step(ch))
        ch = [7] * count
    pts.append(ses.step(ch))
    proofs = ses.finish([99] * count)
    assert proofs == api.prove_batch(pk, n, m, vs, gs)
    for i in range(count):
        assert proofs[i] == oracle.range_prove(0, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m])
        assert proofs[i][:32] == pts[0][i] and proofs[i][32:64] == pts[1][i][0]  # what the host was shown is what the proof holds
    with pytest.raises(api.BppError):
        ses.step([7] * count)  # no step left
    # a second batch of the same shape reuses the parked device state of the first: same bytes for the same inputs, fresh ones otherwise
    del ses
    vs2 = [v ^ 1 for v in vs]
    for values in (vs, vs2):
        ses = api.ProveSession(pk, n, m, values, gs, api.REFERENCE_NONCES)
        ch = [(y, z)] * count
        for _ in range(L + 1):
            ses.step(ch)
            ch = [7] * count
        assert ses.finish([99] * count) == api.prove_batch(pk, n, m, values, gs)
        del ses


@pytest.mark.parametrize("curve,n,m,count", [(0, 8, 2, 5), (0, 64, 16, 4), (0, 32, 1, 6), (1, 8, 4, 2), (2, 8, 2, 2)])
def test_transcript_proofs_match_oracle_and_verify(api, ctx, oracle, curve, n, m, count):
    q = ORD[curve]
    L = (n * m).bit_length() - 1
    pk = api.PublicKey.new(n * m, curve, ctx)
    pb = pk.pb
    vs, gs = _inputs(n, m, count, 1000 + 7 * n + m + curve, q)
    rng = random.Random(77 + curve)
    nonces = [dict(alpha=rng.randrange(q), r=rng.randrange(q), s=rng.randrange(q), delta=rng.randrange(q), eta=rng.randrange(q),
                   d_L=[rng.randrange(q) for _ in range(L)], d_R=[rng.randrange(q) for _ in range(L)]) for _ in range(count)]
    label = lambda i: b"RangeProof Test %d" % i  # README.md:39 uses one label; distinct transcripts per proof here
    proofs = api.prove_batch_fs(pk, n, m, vs, gs, [merlin.Transcript(label(i)) for i in range(count)], nonces)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    for i in range(count):
        ch = merlin.verifier_challenges(merlin.Transcript(label(i)), q, n, m, Vl[i], proofs[i], pb)
        vals = dict(y=ch[0], z=ch[1], e_final=ch[2], e_round=ch[3:], **nonces[i])
        want = oracle.range_prove(curve, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m], oracle_lib.make_params(oracle_lib.OrcParams, **vals))
        assert proofs[i] == want, i
        assert api.prove_batch(pk, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m], api.make_params(**vals))[0] == want
        assert oracle.range_verify(curve, n, m, b"".join(Vl[i]), proofs[i], oracle_lib.make_params(oracle_lib.OrcParams, **vals)) == 0
    fresh = lambda: [merlin.Transcript(label(i)) for i in range(count)]
    assert api.verify_batch_fs(pk, n, m, proofs, Vl, fresh()) == [0] * count
    # a proof under another proof's transcript, a tampered proof, a tampered commitment
    swapped = fresh()
    swapped[0], swapped[1] = swapped[1], swapped[0]
    assert api.verify_batch_fs(pk, n, m, proofs, Vl, swapped) == [1, 1] + [0] * (count - 2)
    bad = list(proofs)
    sc0 = (2 * L + 3) * pb
    bad[1] = bad[1][:sc0] + ((int.from_bytes(bad[1][sc0:sc0 + 32], "little") + 1) % q).to_bytes(32, "little") + bad[1][sc0 + 32:]
    assert api.verify_batch_fs(pk, n, m, bad, Vl, fresh()) == [0, 1] + [0] * (count - 2)
    Vbad = [list(v) for v in Vl]
    Vbad[count - 1][0] = Vl[0][0]
    assert api.verify_batch_fs(pk, n, m, proofs, Vbad, fresh())[count - 1] == 1


def test_fs_verify_with_constant_blocks_equals_shared_challenge_verify(api, ctx, oracle):
    n, m, count = 16, 4, 9
    L = 6
    pk = api.PublicKey.new(n * m, 0, ctx)
    vs, gs = _inputs(n, m, count, 31, ELL)
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    proofs[2] = proofs[2][:40] + bytes([proofs[2][40] ^ 1]) + proofs[2][41:]  # breaks L_0 (FORMAT or VERIFICATION, same in both paths)
    proofs[5] = proofs[5][:-1] + bytes([0xFF])                                  # delta' not canonical: FORMAT
    const = [12, 23, 99] + [7] * L
    assert api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=[const] * count) == api.verify_batch(pk, n, m, proofs, Vl)
    blocks = [list(const) for _ in range(count)]
    blocks[1][0] = 0        # y = 0: not invertible
    blocks[3][4] = ELL      # e_1 not canonical
    blocks[4][1] = 0        # z = 0 is a legal challenge value: the proof was made with z = 23, so it must simply fail
    got = api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=blocks)
    assert got[1] == 2 and got[3] == 2 and got[4] == 1 and got[0] == 0 and got[5] == 2


def test_session_rejects_bad_host_input(api, ctx):
    n, m = 8, 2
    pk = api.PublicKey.new(n * m, 0, ctx)
    ses = api.ProveSession(pk, n, m, [1, 2], [3, 4])
    with pytest.raises(api.BppError) as ei:
        ses.finish([5])  # out of order
    assert ei.value.rc == 3
    with pytest.raises(api.BppError) as ei:
        ses.step([(0, 5)])  # y = 0
    assert ei.value.rc == 2


def test_default_nonces_are_fresh_and_proofs_verify(api, ctx):
    """with transcript challenges and the reference's public constant blinders a verifier could solve the proof
    scalars for gamma and v. The Fiat-Shamir prover therefore draws alpha, r, s, delta, eta, d_L, d_R per proof from the CSPRNG
    unless told otherwise: two proofs of one witness differ in every blinded element, and both verify."""
    from bulletproofsplus_b200 import merlin
    n, m, count = 32, 2, 2
    pk = api.PublicKey.new(n * m, 0, ctx)
    vs, gs = [5, 6] * count, [7, 8] * count  # the same witness twice
    label = lambda: [merlin.Transcript(b"zk test") for _ in range(count)]
    proofs = api.prove_batch_fs(pk, n, m, vs, gs, label())
    again = api.prove_batch_fs(pk, n, m, vs, gs, label())
    parity = api.prove_batch_fs(pk, n, m, vs, gs, label(), api.REFERENCE_NONCES)
    assert parity[0] == parity[1]  # constants: same witness, same transcript, same bytes
    L = 6
    for a, b in ((proofs[0], proofs[1]), (proofs[0], again[0]), (proofs[0], parity[0])):
        pa = [a[i * 32:(i + 1) * 32] for i in range(2 * L + 3 + 3)]
        pb_ = [b[i * 32:(i + 1) * 32] for i in range(2 * L + 3 + 3)]
        assert all(x != y for x, y in zip(pa, pb_)), "a proof element repeats across proofs made with fresh nonces"
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    assert api.verify_batch_fs(pk, n, m, proofs, Vl, transcripts=label()) == [0] * count
    assert api.verify_batch_fs(pk, n, m, again, Vl, transcripts=label()) == [0] * count


def test_fs_flow_proves_values_above_2_pow_31(api, ctx, oracle):
    """bpp_commit_batch mirrors the reference's `v as i32` (prover.rs:37), so with it no n = 64 proof of
    v >= 2^31 can verify. bpp_commit_batch_u64 commits to all 64 bits; the Fiat-Shamir helpers use it."""
    from bulletproofsplus_b200 import merlin
    n, m = 64, 2
    pk = api.PublicKey.new(n * m, 0, ctx)
    vs, gs = [2**63 + 12345, 2**31], [17, 19]
    V64 = pk.commit_batch(vs, gs, full_u64=True)
    q = api.ORDERS[0]
    assert V64 == [oracle.scalar_mul(0, (v + 2 * g) % q) for v, g in zip(vs, gs)]          # g, h = 1g, 2g
    assert pk.commit_batch(vs, gs) == [oracle.commit(0, [v], [g]) for v, g in zip(vs, gs)]  # the quirk stays where parity needs it
    assert pk.commit_batch(vs, gs) != V64
    proofs = api.prove_batch_fs(pk, n, m, vs, gs, [merlin.Transcript(b"u64")])
    assert api.verify_batch_fs(pk, n, m, proofs, [V64], transcripts=[merlin.Transcript(b"u64")]) == [0]
    assert api.verify_batch_fs(pk, n, m, proofs, [pk.commit_batch(vs, gs)], transcripts=[merlin.Transcript(b"u64")]) == [1]

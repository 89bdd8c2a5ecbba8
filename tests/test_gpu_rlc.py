"""Random-linear-combination batch verification (bpp_range_verify_batch_rlc): the verdict vector must be the per-proof path's
for every batch — all valid (one combined MSM, no fallback), some invalid (fallback names them), undecodable proofs (FORMAT,
left out of the combination) — with one challenge set and with per-proof challenges, on all three curves."""
import random

import pytest

from bulletproofsplus_b200 import merlin
from tests import pyref
from tests.test_gpu_parity import _tampers

pytestmark = pytest.mark.gpu

ORD = {0: pyref.ELL, 1: 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141,
       2: 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001}


@pytest.fixture(scope="module")
def api():
    from bulletproofsplus_b200 import api as a
    return a


@pytest.fixture(scope="module")
def ctx(api):
    return api.Arith.init()


def _batch(api, ctx, curve, n, m, count, seed):
    q = ORD[curve]
    rng = random.Random(seed)
    pk = api.PublicKey.new(n * m, curve, ctx)
    vs = [rng.randrange(0, 2**min(n, 31)) for _ in range(count * m)]
    gs = [rng.randrange(q) for _ in range(count * m)]
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    return pk, proofs, [Vs[i * m:(i + 1) * m] for i in range(count)]


@pytest.mark.parametrize("curve,n,m,count", [(0, 64, 16, 40), (0, 64, 1, 300), (0, 8, 4, 33), (1, 8, 2, 9), (2, 8, 2, 7)])
def test_all_valid_batch_passes_without_fallback(api, ctx, curve, n, m, count):
    pk, proofs, Vl = _batch(api, ctx, curve, n, m, count, 10 * n + m + curve)
    got, fallback = api.verify_batch_rlc(pk, n, m, proofs, Vl)
    assert got == [0] * count and not fallback
    # fixed weights, including huge ones
    q = ORD[curve]
    got, fallback = api.verify_batch_rlc(pk, n, m, proofs, Vl, weights=[q - 1 - i for i in range(count)])
    assert got == [0] * count and not fallback


def test_bad_proofs_are_named_by_the_fallback(api, ctx, oracle):
    n, m, count = 64, 1, 64
    pk, proofs, Vl = _batch(api, ctx, 0, n, m, count, 5)
    Vflat = [v[0] for v in Vl]
    bad = {}
    for t, j in enumerate(range(3, count, 9)):
        kinds = _tampers(oracle, n, m, proofs[j], Vflat[j])
        name, proofs[j], Vflat[j] = kinds[t % len(kinds)]
        bad[j] = name
    Vl = [[v] for v in Vflat]
    want = api.verify_batch(pk, n, m, proofs, Vl)
    assert [i for i, v in enumerate(want) if v] == sorted(bad)
    got, fallback = api.verify_batch_rlc(pk, n, m, proofs, Vl)
    assert got == want and fallback
    # a single bad proof among many, with adversarially simple weights (all ones): still caught
    pk2, proofs2, Vl2 = _batch(api, ctx, 0, 8, 4, 50, 6)
    proofs2[17] = _tampers(oracle, 8, 4, proofs2[17], b"".join(Vl2[17]))[5][1]  # r' + 1
    got, fallback = api.verify_batch_rlc(pk2, 8, 4, proofs2, Vl2, weights=[1] * 50)
    assert got == [1 if i == 17 else 0 for i in range(50)] and fallback


def test_undecodable_proofs_get_format_and_do_not_spoil_the_batch(api, ctx):
    n, m, count = 16, 2, 20
    pk, proofs, Vl = _batch(api, ctx, 0, n, m, count, 8)
    proofs[4] = (1).to_bytes(32, "little") + proofs[4][32:]            # A is not a Ristretto encoding
    proofs[11] = proofs[11][:-32] + pyref.ELL.to_bytes(32, "little")   # delta' = ell: not canonical
    Vl[15] = [pyref.P25519.to_bytes(32, "little"), Vl[15][1]]          # a commitment that does not decode
    want = api.verify_batch(pk, n, m, proofs, Vl)
    assert want == [2 if i in (4, 11, 15) else 0 for i in range(count)]
    got, fallback = api.verify_batch_rlc(pk, n, m, proofs, Vl)
    assert got == want and not fallback  # FORMAT proofs leave the combination: the rest passes in one MSM


def test_rlc_with_per_proof_challenges(api, ctx):
    n, m, count = 32, 2, 12
    q = ORD[0]
    pk = api.PublicKey.new(n * m, 0, ctx)
    rng = random.Random(3)
    vs = [rng.randrange(2**31) for _ in range(count * m)]
    gs = [rng.randrange(q) for _ in range(count * m)]
    label = lambda i: b"rlc test %d" % i
    proofs = api.prove_batch_fs(pk, n, m, vs, gs, [merlin.Transcript(label(i)) for i in range(count)])
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    ch = [merlin.verifier_challenges(merlin.Transcript(label(i)), q, n, m, Vl[i], proofs[i], 32) for i in range(count)]
    got, fallback = api.verify_batch_rlc(pk, n, m, proofs, Vl, challenges=ch)
    assert got == [0] * count and not fallback
    ch[7], ch[8] = ch[8], ch[7]
    got, fallback = api.verify_batch_rlc(pk, n, m, proofs, Vl, challenges=ch)
    assert got == [1 if i in (7, 8) else 0 for i in range(count)] and fallback


def test_rlc_rejects_bad_weights(api, ctx):
    pk, proofs, Vl = _batch(api, ctx, 0, 8, 2, 4, 9)
    for w in ([1, 0, 1, 1], [1, 1, pyref.ELL, 1]):
        with pytest.raises(api.BppError) as ei:
            api.verify_batch_rlc(pk, 8, 2, proofs, Vl, weights=w)
        assert ei.value.rc == 2

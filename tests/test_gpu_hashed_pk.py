"""bpp_pk_new_hashed: generators derived by hashing (RFC 9496 §4.3.4 element derivation of SHA-512 output) instead of the
reference's placeholder multiples of g (publickey.rs:21-48). Parity against tests/pyref.py, which is pinned to the RFC 9496
A.3 vector by tests/test_oracle_kat.py; then the whole protocol over such a key: prove -> verify, tamper -> reject, transcript
flow, RLC batch."""
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from bulletproofsplus_b200 import api
    return api, api.Arith.init()


def test_generators_match_python_hash_to_ristretto(env):
    api, ctx = env
    from tests import pyref
    dom = b"bulletproofsplus_b200 test key"
    pk = api.PublicKey.new_hashed(16, dom, 0, ctx)
    assert pk.g == pyref.hashed_generator(dom, b"g", 0)
    assert pk.h == pyref.hashed_generator(dom, b"h", 0)
    assert pk.G_vec == [pyref.hashed_generator(dom, b"G", i) for i in range(16)]
    assert pk.H_vec == [pyref.hashed_generator(dom, b"H", i) for i in range(16)]
    other = api.PublicKey.new_hashed(16, dom + b"!", 0, ctx)
    assert other.g != pk.g and set(other.G_vec).isdisjoint(pk.G_vec)
    assert len(set(pk.G_vec + pk.H_vec + [pk.g, pk.h])) == 34  # no collisions (the reference rule has G_4 == H_2)
    empty = api.PublicKey.new_hashed(4, b"", 0, ctx)
    assert empty.g == pyref.hashed_generator(b"", b"g", 0)


def test_commitment_is_v_g_plus_gamma_h(env, oracle):
    api, ctx = env
    pk = api.PublicKey.new_hashed(8, b"k", 0, ctx)
    v, gm = 12345, 987654321
    want = oracle.point_add(0, oracle.scalar_mul(0, v, pk.g), oracle.scalar_mul(0, gm, pk.h))
    assert pk.commit_batch([v], [gm]) == [want]


@pytest.mark.parametrize("n,m", [(8, 2), (64, 1), (64, 16)])
def test_protocol_over_a_hashed_key(env, n, m):
    api, ctx = env
    from bulletproofsplus_b200 import merlin
    import random
    rng = random.Random(n * m)
    q = api.ORDERS[0]
    pk = api.PublicKey.new_hashed(n * m, b"protocol", 0, ctx)
    count = 3
    vs = [rng.randrange(2**min(n, 31)) for _ in range(count * m)]
    gs = [rng.randrange(q) for _ in range(count * m)]
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    bad = bytearray(proofs[1]); bad[-96] ^= 1
    assert api.verify_batch(pk, n, m, [proofs[0], bytes(bad), proofs[2]], Vl) == [0, 1, 0]
    assert api.verify_batch(pk, n, m, proofs, [Vl[1], Vl[0], Vl[2]]) == [1, 1, 0]
    # a proof made against the reference-rule key does not verify against the hashed key
    ref_pk = api.PublicKey.new(n * m, 0, ctx)
    ref_proofs = api.prove_batch(ref_pk, n, m, vs[:m], gs[:m])
    assert api.verify_batch(pk, n, m, ref_proofs, [Vl[0]]) == [1]
    # transcript flow + random-linear-combination batch
    tr = lambda: [merlin.Transcript(b"hashed") for _ in range(count)]
    fs = api.prove_batch_fs(pk, n, m, vs, gs, tr())
    V64 = pk.commit_batch(vs, gs, full_u64=True)
    Vl64 = [V64[i * m:(i + 1) * m] for i in range(count)]
    assert api.verify_batch_fs(pk, n, m, fs, Vl64, transcripts=tr()) == [0] * count
    assert api.verify_batch_rlc(pk, n, m, proofs, Vl, weights=[rng.getrandbits(128) | 1 for _ in range(count)]) == ([0] * count, False)


def test_unsupported_on_other_curves(env):
    api, ctx = env
    with pytest.raises(api.BppError) as ei:
        api.PublicKey.new_hashed(4, b"x", 1, ctx)
    assert ei.value.rc == 8

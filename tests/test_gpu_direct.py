"""Direct fixed-base tables (bpp_pk_precompute, csrc/pk_direct.cuh): with them every call (one job or thousands) evaluates the generator terms
by one table addition per window instead of the bucket method. The results must be the SAME bytes: against the CPU oracle, against
the exponent-domain value (sum_i s_i k_i) g, and against the bucket path of the same library on the same inputs — for adversarial
scalars (0, 1, ell-1, the values around ell/2 where half-order folding switches, 2^252-1, all-equal), for window widths where the
unsigned last window is full (c divides 252: 12, 9) and where it is partial (c = 8, 13), and through every protocol entry point."""
import random

import pytest

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


def _table_bytes(np_, c):
    W = (252 + c - 1) // c
    return (W + 1) * (1 << (c - 1)) * np_ * 128


@pytest.mark.parametrize("length,c", [(3, 12), (3, 9), (3, 13), (16, 8)])
def test_pk_msm_direct_equals_bucket_path_and_exponent_value(env, oracle, length, c):
    api, ctx = env
    np_ = 2 + 2 * length
    dlog = [1, 2] + [3 * (i + 1) for i in range(length)] + [5 * (i + 1) for i in range(length)]  # publickey.rs:21-48
    rng = random.Random(100 * length + c)
    edge = [0, 1, 2, ELL - 1, ELL - 2, (ELL - 1) // 2, (ELL + 1) // 2, (ELL + 3) // 2, 2**252 - 1, 2**252, 2**251, 2**(c - 1), 2**(c - 1) + 1,
            2**c - 1, (1 << 252) - (1 << (252 - c)), ELL // 3]
    jobs = 40  # the direct path is taken once the tables exist
    scal = []
    for j in range(jobs):
        if j == 0:
            row = [edge[i % len(edge)] for i in range(np_)]
        elif j == 1:
            row = [edge[(i + 5) % len(edge)] for i in range(np_)]
        elif j == 2:
            row = [(ELL - 1) // 2] * np_
        elif j == 3:
            row = [0] * np_
        else:
            row = [rng.randrange(ELL) for _ in range(np_)]
        scal += row
    want = [oracle.scalar_mul(R, sum(s * k for s, k in zip(scal[j * np_:(j + 1) * np_], dlog)) % ELL) for j in range(jobs)]
    pk = api.PublicKey.new(length, R, ctx)
    bucket = pk.msm(scal, jobs)
    assert bucket == want
    got_c = pk.precompute(_table_bytes(np_, c))
    assert got_c == c
    assert pk.precompute(1) == c  # idempotent: the tables exist
    assert pk.msm(scal, jobs) == want
    assert pk.msm(scal[:np_ * 3], 3) == want[:3]  # few jobs: up to 64 slices per job and the tree join


@pytest.mark.parametrize("n,m,count", [(64, 1, 70), (32, 4, 33), (64, 16, 34)])
def test_protocol_bytes_and_verdicts_do_not_change(env, oracle, n, m, count):
    api, ctx = env
    rng = random.Random(n * m + count)
    vs = [rng.randrange(2**min(n, 31)) for _ in range(count * m)]
    gs = [rng.randrange(ELL) for _ in range(count * m)]
    pk = api.PublicKey.new(n * m, R, ctx)
    proofs0 = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    tampered = list(proofs0)
    for i in range(0, count, 7):
        b = bytearray(tampered[i]); b[-96 + (i % 3) * 32] ^= 1
        tampered[i] = bytes(b)
    verdicts0 = api.verify_batch(pk, n, m, tampered, Vl)
    assert verdicts0 == [1 if i % 7 == 0 else 0 for i in range(count)]
    c = pk.precompute(3 << 30)
    assert 8 <= c <= 16
    proofs1 = api.prove_batch(pk, n, m, vs, gs)
    assert proofs1 == proofs0
    for i in (0, count - 1):
        assert proofs1[i] == oracle.range_prove(R, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m])
    assert api.verify_batch(pk, n, m, tampered, Vl) == verdicts0
    chal = [[12, 23, 99] + [7] * ((n * m).bit_length() - 1)] * count if m > 1 else [[7, 7, 99] + [7] * ((n * m).bit_length() - 1)] * count
    assert api.verify_batch_fs(pk, n, m, tampered, Vl, challenges=chal) == verdicts0
    ws = [rng.getrandbits(128) | 1 for _ in range(count)]
    assert api.verify_batch_rlc(pk, n, m, proofs1, Vl, weights=ws) == ([0] * count, False)
    # the interactive prover (two jobs per proof and step) over the tables
    ses = api.ProveSession(pk, n, m, vs, gs, api.REFERENCE_NONCES)
    ch = [(12, 23)] * count if m > 1 else [(7, 7)] * count
    for _ in range(ses.rounds + 1):
        ses.step(ch)
        ch = [7] * count
    assert ses.finish([99] * count) == proofs0


def test_precompute_errors(env):
    api, ctx = env
    pk = api.PublicKey.new(8, R, ctx)
    with pytest.raises(api.BppError) as ei:
        pk.precompute(1 << 10)  # not even c = 8 fits
    assert ei.value.rc == 3
    with pytest.raises(api.BppError) as ei:
        api.PublicKey.new(8, 1, ctx).precompute(1 << 30)  # secp256k1: not implemented
    assert ei.value.rc == 8


def test_precompute_that_cannot_be_allocated_fails_cleanly(env, oracle):
    """A table budget above what the device can give (c = 16 for 8194 generators: 583 GB) must fail with BPP_ERR_CUDA, leave the
    key usable on the bucket path, leave no pending CUDA error in the context, and a later request that fits must succeed."""
    api, ctx = env
    n, m = 64, 64
    pk = api.PublicKey.new(n * m, R, ctx)
    with pytest.raises(api.BppError):
        pk.precompute(10**12)
    rng = random.Random(9)
    vs = [rng.randrange(2**31) for _ in range(m)]
    gs = [rng.randrange(ELL) for _ in range(m)]
    proof = api.prove_batch(pk, n, m, vs, gs)[0]
    V = pk.commit_batch(vs, gs)
    assert api.verify_batch(pk, n, m, [proof], [V]) == [0]
    assert proof == oracle.range_prove(R, n, m, vs, gs)
    c = pk.precompute(_table_bytes(2 + 2 * n * m, 9))
    assert c == 9
    assert api.verify_batch(pk, n, m, [proof] * 33, [V] * 33) == [0] * 33  # through the tables


@pytest.mark.parametrize("n,m", [(8, 1), (32, 2), (64, 4), (64, 16)])
def test_few_jobs_take_the_tables_too(env, oracle, n, m):
    """One proof, a handful of proofs, 1..31 key MSMs against a key WITH tables: up to 64 slices per job, tree join, the
    per-proof points through Straus beside them. Proof bytes and verdicts against the oracle, MSMs against (sum s_i k_i) g."""
    api, ctx = env
    N = n * m
    np_ = 2 + 2 * N
    pk = api.PublicKey.new(N, R, ctx)
    assert pk.precompute(_table_bytes(np_, 10 if N <= 256 else 8)) >= 8
    rng = random.Random(7000 + N)
    dlog = [1, 2] + [3 * (i + 1) for i in range(N)] + [5 * (i + 1) for i in range(N)]
    for jobs in (1, 2, 5, 31):
        scal = [rng.randrange(ELL) for _ in range(jobs * np_)]
        if jobs == 2:
            scal[:np_] = [0] * np_                                       # the identity as a result
            scal[np_:np_ + 3] = [ELL - 1, (ELL + 1) // 2, 2**252 - 1]
        want = [oracle.scalar_mul(R, sum(s * k for s, k in zip(scal[j * np_:(j + 1) * np_], dlog)) % ELL) for j in range(jobs)]
        assert pk.msm(scal, jobs) == want, jobs
    for count in (1, 3, 15):
        vs = [rng.randrange(2**min(n, 31)) for _ in range(count * m)]
        gs = [rng.randrange(ELL) for _ in range(count * m)]
        proofs = api.prove_batch(pk, n, m, vs, gs)
        for i in range(count):
            assert proofs[i] == oracle.range_prove(R, n, m, vs[i * m:(i + 1) * m], gs[i * m:(i + 1) * m]), (count, i)
        Vs = pk.commit_batch(vs, gs)
        Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
        bad = bytearray(proofs[-1]); bad[-1] ^= 1
        batch = proofs[:-1] + [bytes(bad)]
        want = [0] * (count - 1) + [1]
        assert api.verify_batch(pk, n, m, batch, Vl) == want
        L = N.bit_length() - 1
        y, z = (7, 7) if m == 1 else (12, 23)
        assert api.verify_batch_fs(pk, n, m, batch, Vl, challenges=[[y, z, 99] + [7] * L] * count) == want


@pytest.mark.parametrize("log2n,jobs_k", [(6, [(1, 64), (3, 20)]), (10, [(1, 1024), (5, 200), (40, 25)]), (13, [(1, 8192), (2, 4096)])])
def test_point_set_direct_tables(env, oracle, log2n, jobs_k):
    """bpp_points_precompute_direct: a caller's point set with direct tables. Every MSM against it (one job or many, all points
    or a prefix) must give the bytes of the generic path and of the exponent-domain value (sum s_i k_i) g."""
    api, ctx = env
    N = 1 << log2n
    rng = random.Random(500 + log2n)
    mult = [rng.randrange(ELL) for _ in range(N)]
    mult[0], mult[1] = 0, ELL - 1  # the identity and -g among the points
    le32 = lambda x: int(x).to_bytes(32, "little")
    pts = api.Points.from_multipliers(b"".join(le32(x) for x in mult), R, ctx)
    cases = []
    for jobs, k in jobs_k:
        sc = [rng.randrange(ELL) for _ in range(jobs * k)]
        sc[:4] = [0, 1, ELL - 1, (ELL + 1) // 2][:min(4, len(sc))]
        blob = b"".join(le32(x) for x in sc)
        want = [oracle.scalar_mul(R, sum(s * m for s, m in zip(sc[j * k:(j + 1) * k], mult[:k])) % ELL) for j in range(jobs)]
        assert pts.msm(blob, k=k, n_jobs=jobs) == want  # generic path
        cases.append((blob, k, jobs, want))
    budget = (252 // 9 + 2) * (1 << 8) * N * 128  # room for c = 9 at least
    c = pts.precompute_direct(min(budget, 8 << 30))
    assert 8 <= c <= 16 and pts.precompute_direct(1) == c  # idempotent
    for blob, k, jobs, want in cases:
        assert pts.msm(blob, k=k, n_jobs=jobs) == want, (k, jobs)


def test_point_set_direct_tables_unsupported_curve(env):
    api, ctx = env
    pts = api.Points.from_multipliers((7).to_bytes(32, "little") * 4, 1, ctx)
    with pytest.raises(api.BppError):
        pts.precompute_direct(1 << 30)

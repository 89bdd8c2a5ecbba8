"""Two contexts on one GPU sharing one public key (include/bpp.h: a bpp_ctx belongs to one host thread, a bpp_pk is immutable
after set-up and may be used by several contexts of its device at once): two host threads verify and prove different batches
at the same time, with and without the key's direct tables. Every verdict must equal the tamper pattern of its own batch, every
proof the bytes the single-context call produces (those are compared with the CPU oracle in test_gpu_parity.py)."""
import random
import threading

import pytest

pytestmark = pytest.mark.gpu

ELL = 2**252 + 27742317777372353535851937790883648493


@pytest.fixture(scope="module")
def api():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from bulletproofsplus_b200 import api as a
    return a


def _batch(api, pk, n, m, count, seed, tamper_every):
    rng = random.Random(seed)
    vs = [rng.randrange(2**n) for _ in range(count * m)]
    gs = [rng.randrange(ELL) for _ in range(count * m)]
    proofs = api.prove_batch(pk, n, m, vs, gs)
    Vs = pk.commit_batch(vs, gs)
    Vl = [Vs[i * m:(i + 1) * m] for i in range(count)]
    want = []
    for i in range(count):
        if tamper_every and i % tamper_every == 1:
            p = bytearray(proofs[i])
            p[-1 - 32 * (i % 3)] ^= 1  # one of r', s', delta' (still canonical: the top byte's low bit)
            proofs[i] = bytes(p)
            want.append(1)
        else:
            want.append(0)
    return vs, gs, proofs, Vl, want


@pytest.mark.parametrize("tables", [False, True])
def test_two_threads_two_contexts_one_key(api, tables):
    n, m, count, rounds = 16, 2, 96, 6
    ctx_a, ctx_b = api.Context(0), api.Context(0)
    pk = api.PublicKey.new(n * m, 0, ctx_a)
    if tables:
        assert pk.precompute(1 << 30) >= 8
    ba = _batch(api, pk, n, m, count, 11, 5)
    bb = _batch(api, pk, n, m, count, 12, 7)
    L = (n * m).bit_length() - 1
    chal = [[12, 23, 99] + [7] * L] * count
    ref_proofs_a = api.prove_batch(pk, n, m, ba[0], ba[1])
    # the same key object seen from the second context (the handle is shared; only the calling context differs)
    pk_b = api.PublicKey.__new__(api.PublicKey)
    pk_b.__dict__.update(pk.__dict__)
    pk_b.ctx = ctx_b
    errors = []

    def worker(pkx, batch, also_prove):
        try:
            vs, gs, proofs, Vl, want = batch
            for r in range(rounds):
                assert api.verify_batch(pkx, n, m, proofs, Vl) == want, "verify_batch"
                assert api.verify_batch_fs(pkx, n, m, proofs, Vl, challenges=chal) == want, "verify_batch_fs"
                if also_prove:
                    assert api.prove_batch(pkx, n, m, vs, gs) == ref_proofs_a, "prove_batch"
        except Exception as e:  # noqa: BLE001 - reported to the main thread
            errors.append(repr(e))

    ta = threading.Thread(target=worker, args=(pk, ba, True))
    tb = threading.Thread(target=worker, args=(pk_b, bb, False))
    ta.start(); tb.start()
    ta.join(); tb.join()
    assert not errors, errors
    pk_b.__dict__.clear()  # the handle is owned (and released) by pk alone

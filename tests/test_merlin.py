"""Host-side transcript (bulletproofsplus_b200/merlin.py): Keccak-f[1600] against hashlib's SHA3-256, STROBE/Merlin against
the merlin crate's `equivalence_simple` test value, and the determinism properties the prover/verifier pair relies on."""
import hashlib

from bulletproofsplus_b200 import merlin

ELL = 2**252 + 27742317777372353535851937790883648493


def _sha3_256(msg: bytes) -> bytes:
    rate, st = 136, bytearray(200)
    m = bytearray(msg) + b"\x06"
    while len(m) % rate:
        m += b"\x00"
    m[-1] |= 0x80
    for off in range(0, len(m), rate):
        for i in range(rate):
            st[i] ^= m[off + i]
        merlin.keccak_f1600(st)
    return bytes(st[:32])


def test_keccak_f1600_matches_hashlib():
    for msg in (b"", b"abc", bytes(range(135)), bytes(range(136)), bytes(200) * 3):
        assert _sha3_256(msg) == hashlib.sha3_256(msg).digest()


def test_merlin_equivalence_simple_vector():
    t = merlin.Transcript(b"test protocol")
    t.append_message(b"some label", b"some data")
    assert t.challenge_bytes(b"challenge", 32).hex() == "d5a21972d0d5fe320c0d263fac7fffb8145aa640af6e9bca177c03c7efcf0615"


def test_transcript_binding_and_clone():
    a, b = merlin.Transcript(b"p"), merlin.Transcript(b"p")
    a.append_message(b"x", b"1")
    b.append_message(b"x", b"2")
    assert a.challenge_bytes(b"c", 16) != b.challenge_bytes(b"c", 16)
    c = merlin.Transcript(b"p")
    d = c.clone()
    c.append_u64(b"n", 7)
    d.append_u64(b"n", 7)
    assert c.challenge_scalar(b"e", ELL) == d.challenge_scalar(b"e", ELL)
    long = merlin.Transcript(b"p")
    long.append_message(b"big", bytes(1000))  # crosses several rate blocks
    assert len(long.challenge_bytes(b"c", 400)) == 400


def test_verifier_recomputes_the_provers_challenges():
    pb, n, m, rounds = 32, 8, 2, 4
    Vs = [bytes([i + 1]) * 32 for i in range(m)]
    pts = [bytes([0x40 + i]) * 32 for i in range(2 * rounds + 3)]
    proof = b"".join(pts) + bytes(96)
    rt = merlin.RangeProofTranscript(merlin.Transcript(b"RangeProof Test"), ELL, n, m, Vs)
    y, z = rt.after_A(pts[0])
    es = [rt.after_round(pts[1 + k], pts[1 + rounds + k]) for k in range(rounds)]
    e = rt.after_final(pts[1 + 2 * rounds], pts[2 + 2 * rounds])
    got = merlin.verifier_challenges(merlin.Transcript(b"RangeProof Test"), ELL, n, m, Vs, proof, pb)
    assert got == [y, z, e] + es and all(0 < x < ELL for x in got) and len(set(got)) == len(got)
    other = merlin.verifier_challenges(merlin.Transcript(b"another label"), ELL, n, m, Vs, proof, pb)
    assert other != got

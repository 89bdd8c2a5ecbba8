"""Host-side Fiat-Shamir: Merlin transcripts (STROBE-128 over Keccak-f[1600]) and the range-proof transcript protocol.

`north_star` keeps the transcript on the host; the library only ever sees the challenges (bpp_range_prove_begin/_step/_finish,
bpp_range_verify_batch_fs). The reference snapshot has no transcript (every challenge is a constant), so there is
nothing of its own to be bit-exact with here; this module is pinned instead to Merlin's published test vector
(merlin/src/transcript.rs `equivalence_simple`: protocol "test protocol", message ("some label", "some data"), 32 challenge
bytes under "challenge" = d5a21972...0615) and Keccak-f to hashlib's SHA3-256 (tests/test_merlin.py). include/bpp_merlin.hpp is
the same thing for the C++ host.

Protocol `bulletproofsplus_b200 rangeproof v1` (labels; the wip part uses the labels of the commented-out block at
weighted_inner_product_proof.rs:339-347 of the reference, the range part follows the README's usage at :39-50):

    dom-sep  "rangeproof v1"      n (u64)   m (u64)      V  x m        A        -> y, z     (challenge_scalar)
    dom-sep  "wipp v1"            n (u64, = n*m generators)            [the weights are y's powers: bound through y]
    per round k:   L, R  -> e   ;   finally   A, B (wip) -> e
Challenge scalars are 64 transcript bytes reduced mod the group order; a zero challenge is re-drawn (next 64 bytes).
"""
from __future__ import annotations

from typing import List, Sequence

_RC = [0x0000000000000001, 0x0000000000008082, 0x800000000000808A, 0x8000000080008000, 0x000000000000808B, 0x0000000080000001,
       0x8000000080008081, 0x8000000000008009, 0x000000000000008A, 0x0000000000000088, 0x0000000080008009, 0x000000008000000A,
       0x000000008000808B, 0x800000000000008B, 0x8000000000008089, 0x8000000000008003, 0x8000000000008002, 0x8000000000000080,
       0x000000000000800A, 0x800000008000000A, 0x8000000080008081, 0x8000000000008080, 0x0000000080000001, 0x8000000080008008]
_ROT = [[0, 36, 3, 41, 18], [1, 44, 10, 45, 2], [62, 6, 43, 15, 61], [28, 55, 25, 21, 56], [27, 20, 39, 8, 14]]
_M64 = (1 << 64) - 1


def _rol(x, n):
    n %= 64
    return ((x << n) | (x >> (64 - n))) & _M64 if n else x


def keccak_f1600(state: bytearray) -> None:
    a = [[int.from_bytes(state[8 * (x + 5 * y):8 * (x + 5 * y) + 8], "little") for y in range(5)] for x in range(5)]
    for rc in _RC:
        c = [a[x][0] ^ a[x][1] ^ a[x][2] ^ a[x][3] ^ a[x][4] for x in range(5)]
        d = [c[(x - 1) % 5] ^ _rol(c[(x + 1) % 5], 1) for x in range(5)]
        a = [[a[x][y] ^ d[x] for y in range(5)] for x in range(5)]
        b = [[0] * 5 for _ in range(5)]
        for x in range(5):
            for y in range(5):
                b[y][(2 * x + 3 * y) % 5] = _rol(a[x][y], _ROT[x][y])
        a = [[b[x][y] ^ ((~b[(x + 1) % 5][y]) & b[(x + 2) % 5][y]) for y in range(5)] for x in range(5)]
        a[0][0] ^= rc
    for x in range(5):
        for y in range(5):
            state[8 * (x + 5 * y):8 * (x + 5 * y) + 8] = a[x][y].to_bytes(8, "little")


class Strobe128:
    R = 166
    FLAG_I, FLAG_A, FLAG_C, FLAG_T, FLAG_M, FLAG_K = 1, 2, 4, 8, 16, 32

    def __init__(self, protocol_label: bytes):
        st = bytearray(200)
        st[0:6] = bytes([1, self.R + 2, 1, 0, 1, 96])
        st[6:18] = b"STROBEv1.0.2"
        keccak_f1600(st)
        self.state, self.pos, self.pos_begin, self.cur_flags = st, 0, 0, 0
        self.meta_ad(protocol_label, False)

    def clone(self) -> "Strobe128":
        c = Strobe128.__new__(Strobe128)
        c.state, c.pos, c.pos_begin, c.cur_flags = bytearray(self.state), self.pos, self.pos_begin, self.cur_flags
        return c

    def _run_f(self):
        self.state[self.pos] ^= self.pos_begin
        self.state[self.pos + 1] ^= 0x04
        self.state[self.R + 1] ^= 0x80
        keccak_f1600(self.state)
        self.pos, self.pos_begin = 0, 0

    def _absorb(self, data: bytes):
        for byte in data:
            self.state[self.pos] ^= byte
            self.pos += 1
            if self.pos == self.R:
                self._run_f()

    def _overwrite(self, data: bytes):
        for byte in data:
            self.state[self.pos] = byte
            self.pos += 1
            if self.pos == self.R:
                self._run_f()

    def _squeeze(self, n: int) -> bytes:
        out = bytearray()
        for _ in range(n):
            out.append(self.state[self.pos])
            self.state[self.pos] = 0
            self.pos += 1
            if self.pos == self.R:
                self._run_f()
        return bytes(out)

    def _begin_op(self, flags: int, more: bool):
        if more:
            assert self.cur_flags == flags
            return
        assert not flags & self.FLAG_T
        old_begin = self.pos_begin
        self.pos_begin = self.pos + 1
        self.cur_flags = flags
        self._absorb(bytes([old_begin, flags]))
        if flags & (self.FLAG_C | self.FLAG_K) and self.pos != 0:
            self._run_f()

    def meta_ad(self, data: bytes, more: bool):
        self._begin_op(self.FLAG_M | self.FLAG_A, more)
        self._absorb(data)

    def ad(self, data: bytes, more: bool):
        self._begin_op(self.FLAG_A, more)
        self._absorb(data)

    def prf(self, n: int, more: bool) -> bytes:
        self._begin_op(self.FLAG_I | self.FLAG_A | self.FLAG_C, more)
        return self._squeeze(n)

    def key(self, data: bytes, more: bool):
        self._begin_op(self.FLAG_A | self.FLAG_C, more)
        self._overwrite(data)


class Transcript:
    """merlin::Transcript"""

    def __init__(self, label: bytes):
        self.strobe = Strobe128(b"Merlin v1.0")
        self.append_message(b"dom-sep", label)

    def clone(self) -> "Transcript":
        t = Transcript.__new__(Transcript)
        t.strobe = self.strobe.clone()
        return t

    def append_message(self, label: bytes, message: bytes):
        self.strobe.meta_ad(label, False)
        self.strobe.meta_ad(len(message).to_bytes(4, "little"), True)
        self.strobe.ad(message, False)

    def append_u64(self, label: bytes, x: int):
        self.append_message(label, int(x).to_bytes(8, "little"))

    def challenge_bytes(self, label: bytes, n: int) -> bytes:
        self.strobe.meta_ad(label, False)
        self.strobe.meta_ad(int(n).to_bytes(4, "little"), True)
        return self.strobe.prf(n, False)

    def challenge_scalar(self, label: bytes, order: int) -> int:
        while True:
            x = int.from_bytes(self.challenge_bytes(label, 64), "little") % order
            if x:
                return x


class RangeProofTranscript:
    """The protocol in the module docstring, one instance per proof; the same calls on the prover's and the verifier's side."""

    def __init__(self, transcript: Transcript, order: int, n: int, m: int, commitments: Sequence[bytes]):
        self.t, self.q = transcript, order
        self.t.append_message(b"dom-sep", b"rangeproof v1")
        self.t.append_u64(b"n", n)
        self.t.append_u64(b"m", m)
        for V in commitments:
            self.t.append_message(b"V", V)
        self.N = n * m

    def after_A(self, A: bytes):
        self.t.append_message(b"A", A)
        y, z = self.t.challenge_scalar(b"y", self.q), self.t.challenge_scalar(b"z", self.q)
        self.t.append_message(b"dom-sep", b"wipp v1")
        self.t.append_u64(b"n", self.N)
        return y, z

    def after_round(self, L: bytes, R: bytes) -> int:
        self.t.append_message(b"L", L)
        self.t.append_message(b"R", R)
        return self.t.challenge_scalar(b"e", self.q)

    def after_final(self, A: bytes, B: bytes) -> int:
        self.t.append_message(b"A", A)
        self.t.append_message(b"B", B)
        return self.t.challenge_scalar(b"e", self.q)


def verifier_challenges(transcript: Transcript, order: int, n: int, m: int, commitments: Sequence[bytes], proof: bytes, pb: int) -> List[int]:
    """[y, z, e, e_0 .. e_{L-1}] of one proof (wire layout A | L[L] | R[L] | wipA | wipB | scalars): the block bpp_range_verify_batch_fs takes."""
    rounds = ((len(proof) - 96) // pb - 3) // 2
    pt = lambda i: proof[i * pb:(i + 1) * pb]
    rt = RangeProofTranscript(transcript, order, n, m, commitments)
    y, z = rt.after_A(pt(0))
    es = [rt.after_round(pt(1 + k), pt(1 + rounds + k)) for k in range(rounds)]
    e = rt.after_final(pt(1 + 2 * rounds), pt(2 + 2 * rounds))
    return [y, z, e] + es

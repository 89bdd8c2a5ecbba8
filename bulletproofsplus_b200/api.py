"""Host-side mirror of the reference crate's public API over the C ABI (include/bpp.h).

Names, argument meaning and error behaviour follow the reference (paths relative to its src/):
    Arith.init                          bls12_381/building_block/arith.rs:9-17
    PublicKey.new / .commitment         publickey.rs:21-52
    RangeProver.commit                  range/prover.rs:28-42
    RangeProof.prove / .verify          range/mod.rs:31-78
    WeightedInnerProductProof.prove / .verify   weighted_inner_product_proof.rs:36-43, :238-248
    MulVec                              bls12_381/building_block/mulvec.rs:13-53
Everything is computed by libbpp_b200.so on the GPU; this file only marshals bytes. Scalars are Python
ints (reduced mod the group order like PrimeFieldElem::new), points are their boundary encodings (bytes).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

from . import _lib
from ._lib import BppParams

ORDERS = {
    0: 2**252 + 27742317777372353535851937790883648493,
    1: 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141,
    2: 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001,
}


class ProofError(Exception):
    """errors.rs:14-50; only VerificationError is ever produced by the reference path."""


class VerificationError(ProofError):
    pass


class FormatError(ProofError):
    pass


class BppError(RuntimeError):
    def __init__(self, rc, msg=""):
        super().__init__(f"bpp error {rc} ({_lib.ERROR_NAMES[rc] if 0 <= rc < len(_lib.ERROR_NAMES) else '?'}) {msg}")
        self.rc = rc


def _sc(x: int) -> bytes:
    return int(x).to_bytes(32, "little")


def make_params(y, z, e_final, alpha, r, s, delta, eta, e_round, d_L, d_R) -> BppParams:
    p = BppParams()
    for k, v in (("y", y), ("z", z), ("e_final", e_final), ("alpha", alpha), ("r", r), ("s", s), ("delta", delta), ("eta", eta)):
        C.memmove(getattr(p, k), _sc(v), 32)
    p.n_rounds = len(e_round)
    p._keep = [b"".join(_sc(x) for x in lst) for lst in (e_round, d_L, d_R)]
    p.e_round, p.d_L, p.d_R = p._keep
    return p


class Arith:
    """One context per process and device (Arith::init is a process-global Once in the reference)."""
    _ctx: dict = {}

    @classmethod
    def init(cls, device: int = 0, stream: int | None = None) -> "Context":
        key = (device, stream)
        if key not in cls._ctx:
            cls._ctx[key] = Context(device, stream)
        return cls._ctx[key]


class Context:
    def __init__(self, device: int = 0, stream: int | None = None):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.bpp_init(device, C.c_void_p(stream) if stream else None, C.byref(h))
        if rc != _lib.OK:
            raise BppError(rc, "bpp_init failed: no usable CUDA device (this package has no CPU fallback)")  # arith.rs:12-14
        self.h = h
        self.device = device

    def check(self, rc, what=""):
        if rc != _lib.OK:
            raise BppError(rc, f"{what}: {self.lib.bpp_last_error(self.h).decode()}")

    def close(self):
        """bpp_destroy: waits for the context's streams and frees its device scratch. Keys and point sets created through the
        context stay valid (they own their memory); nothing else may use the context afterwards."""
        if self.h:
            self.lib.bpp_destroy(self.h)
            self.h = None

    def kernel_launches(self) -> int:
        return int(self.lib.bpp_kernel_launches(self.h))

    def sync(self):
        self.check(self.lib.bpp_sync(self.h), "sync")

    def int_peak_probe(self):
        rates, ms = (C.c_double * 3)(), C.c_double()
        self.check(self.lib.bpp_int_peak_probe(self.h, rates, C.byref(ms)), "int_peak_probe")
        return dict(imad32_per_s=rates[0], wide_chain_per_s=rates[1], wide_fused_per_s=rates[2], elapsed_ms=ms.value)

    def fp64_peak_probe(self):
        rates = (C.c_double * 2)()
        self.check(self.lib.bpp_fp64_peak_probe(self.h, rates), "fp64_peak_probe")
        return dict(dfma_per_s=rates[0], f5_mul_per_s=rates[1])

    def trace(self, on: bool):
        self.check(self.lib.bpp_trace(self.h, 1 if on else 0), "trace")

    def trace_dump(self):
        """[(stream, start_us, end_us, kernel)] of the launches recorded since trace(True), over both streams."""
        buf = C.create_string_buffer(1 << 22)
        self.lib.bpp_trace_dump(self.h, buf, len(buf))
        out = []
        for line in buf.value.decode().splitlines():
            s_, a, b, name = line.split(" ", 3)
            out.append((int(s_), float(a), float(b), name))
        return out

    def hot_kernel_stats(self, reset=False):
        ms, n = C.c_double(), C.c_uint64()
        self.check(self.lib.bpp_hot_kernel_stats(self.h, C.byref(ms), C.byref(n), 1 if reset else 0), "hot_kernel_stats")
        return ms.value, int(n.value)

    def debug_op(self, curve, op, a: bytes, b: bytes, elem: int):
        n = len(a) // elem
        out, ok = C.create_string_buffer(n * elem), C.create_string_buffer(n)
        self.check(self.lib.bpp_debug_op(self.h, curve, op, n, a, b, out, ok), "debug_op")
        return out.raw, ok.raw


class PublicKey:
    """publickey.rs:13-18 — g, h, G_vec, H_vec live on the device; `.g .h .G_vec .H_vec` export them."""

    def __init__(self, ctx: Context, curve: int, length: int, domain_sep: bytes | None = None):
        self.ctx, self.curve, self.length = ctx, curve, length
        self.pb = ctx.lib.bpp_point_bytes(curve)
        h = C.c_void_p()
        if domain_sep is None:
            ctx.check(ctx.lib.bpp_pk_new(ctx.h, curve, length, C.byref(h)), "pk_new")
        else:
            ctx.check(ctx.lib.bpp_pk_new_hashed(ctx.h, curve, length, domain_sep, len(domain_sep), C.byref(h)), "pk_new_hashed")
        self.handle = h
        self._exported = None

    @staticmethod
    def new(length: int, curve: int = _lib.RISTRETTO255, ctx: Context | None = None) -> "PublicKey":
        return PublicKey(ctx or Arith.init(), curve, length)

    @staticmethod
    def new_hashed(length: int, domain_sep: bytes, curve: int = _lib.RISTRETTO255, ctx: Context | None = None) -> "PublicKey":
        """Hash-derived generators (no known discrete-log relations): the constructor to use outside parity tests."""
        return PublicKey(ctx or Arith.init(), curve, length, domain_sep)

    def precompute(self, max_bytes: int = 48 << 30) -> int:
        """Build the direct fixed-base tables (bpp_pk_precompute) within max_bytes of HBM; returns the window width used."""
        c = C.c_int(0)
        self.ctx.check(self.ctx.lib.bpp_pk_precompute(self.ctx.h, self.handle, max_bytes, C.byref(c)), "pk_precompute")
        return c.value

    def msm(self, scalars, n_jobs: int = 1) -> list:
        """n_jobs MSMs against [g, h, G.., H..] (bpp_pk_msm); scalars: flat list of n_jobs * (2 + 2 len) ints."""
        q = ORDERS[self.curve]
        out = C.create_string_buffer(n_jobs * self.pb)
        self.ctx.check(self.ctx.lib.bpp_pk_msm(self.ctx.h, self.handle, b"".join(_sc(x % q) for x in scalars), n_jobs, out), "pk_msm")
        return [out.raw[i * self.pb:(i + 1) * self.pb] for i in range(n_jobs)]

    def _export(self):
        if self._exported is None:
            pb, n = self.pb, self.length
            g, h = C.create_string_buffer(pb), C.create_string_buffer(pb)
            G, H = C.create_string_buffer(max(1, pb * n)), C.create_string_buffer(max(1, pb * n))
            self.ctx.check(self.ctx.lib.bpp_pk_export(self.ctx.h, self.handle, g, h, G, H), "pk_export")
            self._exported = (g.raw, h.raw, [G.raw[i * pb:(i + 1) * pb] for i in range(n)], [H.raw[i * pb:(i + 1) * pb] for i in range(n)])
        return self._exported

    g = property(lambda self: self._export()[0])
    h = property(lambda self: self._export()[1])
    G_vec = property(lambda self: self._export()[2])
    H_vec = property(lambda self: self._export()[3])

    def commit_batch(self, vs, gammas, full_u64: bool = False) -> list:
        """V_i = Fr(v_i as i32) g + gamma_i h as in the reference (prover.rs:37), or with all 64 bits of v_i (full_u64)."""
        n = len(vs)
        arr = (C.c_uint64 * n)(*[v & (2**64 - 1) for v in vs])
        out = C.create_string_buffer(max(1, n * self.pb))
        gam = b"".join(_sc(g % ORDERS[self.curve]) for g in gammas)
        fn = self.ctx.lib.bpp_commit_batch_u64 if full_u64 else self.ctx.lib.bpp_commit_batch
        self.ctx.check(fn(self.ctx.h, self.handle, arr, gam, n, out), "commit_batch")
        return [out.raw[i * self.pb:(i + 1) * self.pb] for i in range(n)]

    def __del__(self):
        try:
            self.ctx.lib.bpp_pk_free(self.handle)
        except Exception:
            pass


class RangeProver:
    """range/prover.rs:13-42 — holds the witness; commit() appends (v, gamma, V)."""

    def __init__(self):
        self.v_vec, self.gamma_vec, self.commitment_vec = [], [], []

    def commit(self, pk: PublicKey, v: int, gamma: int):
        self.v_vec.append(v)
        self.gamma_vec.append(gamma)
        self.commitment_vec.append(pk.commit_batch([v], [gamma])[0])


@dataclass
class WeightedInnerProductProof:
    """weighted_inner_product_proof.rs:25-33"""
    L_vec: list
    R_vec: list
    A: bytes
    B: bytes
    r_prime: int
    s_prime: int
    d_prime: int

    def to_bytes(self) -> bytes:
        return b"".join(self.L_vec) + b"".join(self.R_vec) + self.A + self.B + _sc(self.r_prime) + _sc(self.s_prime) + _sc(self.d_prime)

    @staticmethod
    def from_bytes(b: bytes, pb: int) -> "WeightedInnerProductProof":
        rounds = ((len(b) - 96) // pb - 2) // 2
        pts = [b[i * pb:(i + 1) * pb] for i in range(2 * rounds + 2)]
        sc = [int.from_bytes(b[(2 * rounds + 2) * pb + 32 * i:(2 * rounds + 2) * pb + 32 * (i + 1)], "little") for i in range(3)]
        return WeightedInnerProductProof(pts[:rounds], pts[rounds:2 * rounds], pts[2 * rounds], pts[2 * rounds + 1], *sc)

    @staticmethod
    def prove(pk: PublicKey, a_vec, b_vec, power_of_y_vec, gamma: int, commitment: bytes | None = None,
              params: BppParams | None = None) -> "WeightedInnerProductProof":
        """wip.rs:36-43. `commitment` is accepted and ignored, like the reference (dead update, :137-142)."""
        n = pk.length
        if not (len(a_vec) == len(b_vec) == len(power_of_y_vec) == n):
            raise AssertionError("wip: all input vectors must have the same length")  # wip.rs:61-64
        if n & (n - 1):
            raise AssertionError("wip: length must be a power of two")  # wip.rs:67
        q = ORDERS[pk.curve]
        ln = C.c_size_t(pk.ctx.lib.bpp_wip_proof_len(pk.curve, n))
        out = C.create_string_buffer(ln.value)
        enc = lambda v: b"".join(_sc(x % q) for x in v)
        pk.ctx.check(pk.ctx.lib.bpp_wip_prove(pk.ctx.h, pk.handle, enc(a_vec), enc(b_vec), enc(power_of_y_vec), _sc(gamma % q), params, out,
                                              C.byref(ln)), "wip_prove")
        return WeightedInnerProductProof.from_bytes(out.raw[:ln.value], pk.pb)

    def verify(self, pk: PublicKey, power_of_y_vec, G_exp_of_commitment, H_exp_of_commitment, g_exp_of_commitment: int,
               V_exp_of_commitment, A_prime: bytes, V, params: BppParams | None = None) -> None:
        """wip.rs:238-248. Returns None for Ok(()), raises VerificationError otherwise."""
        q = ORDERS[pk.curve]
        enc = lambda v: b"".join(_sc(x % q) for x in v)
        proof = self.to_bytes()
        rc = pk.ctx.lib.bpp_wip_verify(pk.ctx.h, pk.handle, enc(power_of_y_vec), enc(G_exp_of_commitment), enc(H_exp_of_commitment),
                                       _sc(g_exp_of_commitment % q), len(V), enc(V_exp_of_commitment) or None, A_prime,
                                       b"".join(V) or None, proof, len(proof), params)
        _raise_verdict(pk.ctx, rc)


def _raise_verdict(ctx, rc):
    if rc == _lib.OK:
        return
    if rc == _lib.ERR_VERIFICATION:
        raise VerificationError("Proof verification failed.")
    if rc == _lib.ERR_FORMAT:
        raise FormatError("Proof data could not be parsed.")
    ctx.check(rc, "verify")


@dataclass
class RangeProof:
    """range/mod.rs:24-28"""
    A: bytes
    proof: WeightedInnerProductProof

    def to_bytes(self) -> bytes:
        return self.A + self.proof.to_bytes()

    @staticmethod
    def from_bytes(b: bytes, pb: int) -> "RangeProof":
        return RangeProof(b[:pb], WeightedInnerProductProof.from_bytes(b[pb:], pb))

    @staticmethod
    def prove(pk: PublicKey, n: int, prover: RangeProver, params: BppParams | None = None) -> "RangeProof":
        """range/mod.rs:31-55 (dispatches on m == 1 inside the library, like the reference)."""
        m = len(prover.v_vec)
        return RangeProof.from_bytes(prove_batch(pk, n, m, prover.v_vec, prover.gamma_vec, params)[0], pk.pb)

    def verify(self, pk: PublicKey, n: int, commitment_vec, params: BppParams | None = None) -> None:
        """range/mod.rs:57-78. Returns None for Ok(()), raises VerificationError otherwise."""
        m = len(commitment_vec)
        _raise_verdict(pk.ctx, verify_batch(pk, n, m, [self.to_bytes()], [list(commitment_vec)], params)[0])


def _shape_asserts(pk, n, m):
    if pk.length != n * m:
        raise AssertionError("pk.G_vec.len() != n*m")  # range/mod.rs:90-91, :252-253


def prove_batch(pk: PublicKey, n: int, m: int, values, gammas, params: BppParams | None = None) -> list:
    """`count` independent aggregated proofs in one call; values/gammas are flat lists of count*m entries."""
    _shape_asserts(pk, n, m)
    count = len(values) // m
    q = ORDERS[pk.curve]
    plen = pk.ctx.lib.bpp_range_proof_len(pk.curve, n, m)
    arr = (C.c_uint64 * (count * m))(*[v & (2**64 - 1) for v in values])
    out = C.create_string_buffer(max(1, count * plen))
    pk.ctx.check(pk.ctx.lib.bpp_range_prove_batch(pk.ctx.h, pk.handle, n, m, count, arr, b"".join(_sc(g % q) for g in gammas), params, out),
                 "range_prove_batch")
    return [out.raw[i * plen:(i + 1) * plen] for i in range(count)]


def verify_batch(pk: PublicKey, n: int, m: int, proofs, commitments, params: BppParams | None = None) -> list:
    """Verdict code per proof (0 ok, 1 verification, 2 format). proofs: list of bytes; commitments: list of m-lists."""
    _shape_asserts(pk, n, m)
    count = len(proofs)
    plen = pk.ctx.lib.bpp_range_proof_len(pk.curve, n, m)
    if any(len(p) != plen for p in proofs):
        raise FormatError("wrong proof length")
    out = C.create_string_buffer(max(1, count))
    V = b"".join(b"".join(c) for c in commitments)
    pk.ctx.check(pk.ctx.lib.bpp_range_verify_batch(pk.ctx.h, pk.handle, n, m, count, V, b"".join(proofs), params, out), "range_verify_batch")
    return list(out.raw[:count])


REFERENCE_NONCES = "reference-nonces"
"""Pass as `nonces` to get the reference snapshot's PUBLIC constant blinding scalars (alpha = 7 / 33, r, s, delta, eta = 33, 44, 88,
123, d_L = 4, d_R = 5: range/mod.rs:94,256, wip.rs:94-95,175-178). Proofs made with them are byte-comparable with the CPU restatement of the reference and
NOT zero-knowledge: anyone can solve r', s', delta' for gamma and the bits of v. Parity tests and benchmarks only."""


def random_nonces(q: int, count: int, rounds: int) -> list:
    """Fresh uniform non-zero blinding scalars per proof from the operating system's CSPRNG."""
    import secrets
    draw = lambda: 1 + secrets.randbelow(q - 1)
    return [dict(alpha=draw(), r=draw(), s=draw(), delta=draw(), eta=draw(), d_L=[draw() for _ in range(rounds)], d_R=[draw() for _ in range(rounds)])
            for _ in range(count)]


class ProveSession:
    """Interactive batch proving for a transcript-owning host (bpp_range_prove_begin / _step / _finish). Scalars are ints.
    nonces: None = fresh random blinding scalars per proof (random_nonces); a list of dicts = the caller's; REFERENCE_NONCES = the
    reference's public constants (parity only)."""

    def __init__(self, pk: PublicKey, n: int, m: int, values, gammas, nonces=None):
        _shape_asserts(pk, n, m)
        self.pk, self.n, self.m, self.count = pk, n, m, len(values) // m
        self.q = ORDERS[pk.curve]
        self.plen = pk.ctx.lib.bpp_range_proof_len(pk.curve, n, m)
        arr = (C.c_uint64 * (self.count * m))(*[v & (2**64 - 1) for v in values])
        nb = None
        if nonces is None:
            nonces = random_nonces(self.q, self.count, (n * m).bit_length() - 1)
        if nonces != REFERENCE_NONCES:  # per proof: dict(alpha, r, s, delta, eta, d_L[L], d_R[L])
            nb = b"".join(b"".join(_sc(x % self.q) for x in [d["alpha"], d["r"], d["s"], d["delta"], d["eta"], *d["d_L"], *d["d_R"]]) for d in nonces)
        h = C.c_void_p()
        out = C.create_string_buffer(self.count * pk.pb)
        pk.ctx.check(pk.ctx.lib.bpp_range_prove_begin(pk.ctx.h, pk.handle, n, m, self.count, arr, b"".join(_sc(g % self.q) for g in gammas), nb,
                                                      C.byref(h), out), "range_prove_begin")
        self.h = h
        self.rounds = pk.ctx.lib.bpp_prove_session_rounds(h)
        self.A = [out.raw[i * pk.pb:(i + 1) * pk.pb] for i in range(self.count)]

    def step(self, challenges) -> list:
        """challenges: per proof (y, z) for the first call, then one scalar. Returns per proof the pair of points produced."""
        flat = b"".join(b"".join(_sc(x) for x in (c if isinstance(c, (tuple, list)) else (c,))) for c in challenges)
        out = C.create_string_buffer(self.count * 2 * self.pk.pb)
        self.pk.ctx.check(self.pk.ctx.lib.bpp_range_prove_step(self.pk.ctx.h, self.h, flat, out), "range_prove_step")
        pb = self.pk.pb
        return [(out.raw[(2 * i) * pb:(2 * i + 1) * pb], out.raw[(2 * i + 1) * pb:(2 * i + 2) * pb]) for i in range(self.count)]

    def finish(self, e_final) -> list:
        out = C.create_string_buffer(self.count * self.plen)
        self.pk.ctx.check(self.pk.ctx.lib.bpp_range_prove_finish(self.pk.ctx.h, self.h, b"".join(_sc(x) for x in e_final), out), "range_prove_finish")
        return [out.raw[i * self.plen:(i + 1) * self.plen] for i in range(self.count)]

    def __del__(self):
        try:
            self.pk.ctx.lib.bpp_prove_session_free(self.h)
        except Exception:
            pass


def prove_batch_fs(pk: PublicKey, n: int, m: int, values, gammas, transcripts, nonces=None) -> list:
    """Real Fiat-Shamir: `transcripts[i]` is proof i's merlin.Transcript (consumed). Returns the proofs."""
    from . import merlin
    count = len(values) // m
    Vs = pk.commit_batch(values, gammas, full_u64=True)  # the statement proved is about all 64 bits of v
    ses = ProveSession(pk, n, m, values, gammas, nonces)
    rts = [merlin.RangeProofTranscript(transcripts[i], ORDERS[pk.curve], n, m, Vs[i * m:(i + 1) * m]) for i in range(count)]
    ch = [rts[i].after_A(ses.A[i]) for i in range(count)]
    for _ in range(ses.rounds):
        LR = ses.step(ch)
        ch = [rts[i].after_round(*LR[i]) for i in range(count)]
    AB = ses.step(ch)
    return ses.finish([rts[i].after_final(*AB[i]) for i in range(count)])


def verify_batch_fs(pk: PublicKey, n: int, m: int, proofs, commitments, transcripts=None, challenges=None) -> list:
    """Per-proof challenges, derived here from `transcripts` (one merlin.Transcript per proof) unless given explicitly as lists
    [y, z, e, e_0..e_{L-1}]. Verdict code per proof."""
    from . import merlin
    _shape_asserts(pk, n, m)
    count = len(proofs)
    plen = pk.ctx.lib.bpp_range_proof_len(pk.curve, n, m)
    if any(len(p) != plen for p in proofs):
        raise FormatError("wrong proof length")
    if challenges is None:
        challenges = [merlin.verifier_challenges(transcripts[i], ORDERS[pk.curve], n, m, commitments[i], proofs[i], pk.pb) for i in range(count)]
    out = C.create_string_buffer(max(1, count))
    V = b"".join(b"".join(c) for c in commitments)
    blob = b"".join(b"".join(_sc(x) for x in ch) for ch in challenges)
    pk.ctx.check(pk.ctx.lib.bpp_range_verify_batch_fs(pk.ctx.h, pk.handle, n, m, count, V, b"".join(proofs), blob, out), "range_verify_batch_fs")
    return list(out.raw[:count])


def verify_batch_rlc(pk: PublicKey, n: int, m: int, proofs, commitments, params: BppParams | None = None, challenges=None, weights=None):
    """Random-linear-combination batch verification: (verdict per proof, used_fallback). weights default to 128 fresh random bits each."""
    import secrets
    _shape_asserts(pk, n, m)
    count = len(proofs)
    plen = pk.ctx.lib.bpp_range_proof_len(pk.curve, n, m)
    if any(len(p) != plen for p in proofs):
        raise FormatError("wrong proof length")
    if weights is None:
        weights = [secrets.randbits(128) | 1 for _ in range(count)]
    out = C.create_string_buffer(max(1, count))
    fb = C.c_int(0)
    V = b"".join(b"".join(c) for c in commitments)
    blob = b"".join(b"".join(_sc(x) for x in ch) for ch in challenges) if challenges is not None else None
    pk.ctx.check(pk.ctx.lib.bpp_range_verify_batch_rlc(pk.ctx.h, pk.handle, n, m, count, V, b"".join(proofs), params, blob,
                                                       b"".join(_sc(w) for w in weights), out, C.byref(fb)), "range_verify_batch_rlc")
    return list(out.raw[:count]), bool(fb.value)


class MulVec:
    """bls12_381/building_block/mulvec.rs:13-53 — accumulate (scalar, point) pairs, calculate() = sum s_i * P_i."""

    def __init__(self, curve: int = _lib.RISTRETTO255, ctx: Context | None = None):
        self.curve, self.ctx = curve, ctx or Arith.init()
        self.scalars, self.points = [], []

    def add_scalar(self, s: int):
        self.scalars.append(s)

    def add_scalars(self, ss):
        self.scalars.extend(ss)

    def add_point(self, p: bytes):
        self.points.append(p)

    def add_points(self, ps):
        self.points.extend(ps)

    def calculate(self) -> bytes:
        if len(self.scalars) != len(self.points):
            raise RuntimeError("mulvec: lengths of scalars and points must match")  # mulvec.rs:23-25
        pb = self.ctx.lib.bpp_point_bytes(self.curve)
        out = C.create_string_buffer(pb)
        q = ORDERS[self.curve]
        k = len(self.scalars)
        self.ctx.check(self.ctx.lib.bpp_msm(self.ctx.h, self.curve, b"".join(_sc(s % q) for s in self.scalars) or None,
                                            b"".join(self.points) or None, k, out), "msm")
        return out.raw


class Points:
    """A decoded point vector resident in HBM (bpp_points) for repeated / large MSMs."""

    def __init__(self, ctx: Context, curve: int, h):
        self.ctx, self.curve, self.h = ctx, curve, h
        self.pb = ctx.lib.bpp_point_bytes(curve)

    @staticmethod
    def upload(points: bytes, curve: int = _lib.RISTRETTO255, ctx: Context | None = None) -> "Points":
        ctx = ctx or Arith.init()
        h = C.c_void_p()
        k = len(points) // ctx.lib.bpp_point_bytes(curve)
        ctx.check(ctx.lib.bpp_points_upload(ctx.h, curve, points, k, C.byref(h)), "points_upload")
        return Points(ctx, curve, h)

    @staticmethod
    def from_multipliers(mults: bytes, curve: int = _lib.RISTRETTO255, ctx: Context | None = None) -> "Points":
        ctx = ctx or Arith.init()
        h = C.c_void_p()
        ctx.check(ctx.lib.bpp_points_from_multipliers(ctx.h, curve, mults, len(mults) // 32, C.byref(h)), "points_from_multipliers")
        return Points(ctx, curve, h)

    def __len__(self):
        return int(self.ctx.lib.bpp_points_len(self.h))

    def precompute(self) -> "Points":
        """Build window tables for this point set (bpp_points_precompute): same results, no doublings in later MSMs."""
        self.ctx.check(self.ctx.lib.bpp_points_precompute(self.ctx.h, self.h), "points_precompute")
        return self

    def precompute_direct(self, max_bytes: int = 48 << 30) -> int:
        """Direct tables for this point set within max_bytes of HBM (bpp_points_precompute_direct); returns the window width."""
        c = C.c_int(0)
        self.ctx.check(self.ctx.lib.bpp_points_precompute_direct(self.ctx.h, self.h, max_bytes, C.byref(c)), "points_precompute_direct")
        return c.value

    def msm(self, scalars: bytes, k: int | None = None, n_jobs: int = 1) -> list:
        k = k if k is not None else len(scalars) // 32 // n_jobs
        out = C.create_string_buffer(n_jobs * self.pb)
        self.ctx.check(self.ctx.lib.bpp_msm_prepared(self.ctx.h, self.h, scalars, k, n_jobs, out), "msm_prepared")
        return [out.raw[i * self.pb:(i + 1) * self.pb] for i in range(n_jobs)]

    def msm_raw(self, scalars: bytes, k: int | None = None) -> bytes:
        k = k if k is not None else len(scalars) // 32
        out = C.create_string_buffer(self.ctx.lib.bpp_raw_point_bytes(self.curve))
        self.ctx.check(self.ctx.lib.bpp_msm_prepared_raw(self.ctx.h, self.h, scalars, k, out), "msm_prepared_raw")
        return out.raw

    def __del__(self):
        try:
            self.ctx.lib.bpp_points_free(self.h)
        except Exception:
            pass


def points_sum_raw(raw: bytes, curve: int = _lib.RISTRETTO255, ctx: Context | None = None) -> bytes:
    ctx = ctx or Arith.init()
    rb = ctx.lib.bpp_raw_point_bytes(curve)
    out = C.create_string_buffer(ctx.lib.bpp_point_bytes(curve))
    ctx.check(ctx.lib.bpp_points_sum_raw(ctx.h, curve, raw, len(raw) // rb, out), "points_sum_raw")
    return out.raw

"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the C++ CPU restatement of the reference). May be
imported only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs — never from bulletproofsplus_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "liboracle.so")

RISTRETTO, SECP256K1, BLS12_381 = 0, 1, 2
POINT_BYTES = {0: 32, 1: 64, 2: 96}


def build(force: bool = False) -> str:
    srcs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cpp", ".hpp"))]
    if force or not os.path.exists(SO) or any(os.path.getmtime(s) > os.path.getmtime(SO) for s in srcs):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return SO


class OrcParams(C.Structure):
    _fields_ = [(k, C.c_uint8 * 32) for k in ("y", "z", "e_final", "alpha", "r", "s", "delta", "eta")] + [
        ("n_rounds", C.c_uint32), ("e_round", C.c_char_p), ("d_L", C.c_char_p), ("d_R", C.c_char_p)]


def sc_bytes(x: int) -> bytes:
    return int(x).to_bytes(32, "little")


def make_params(cls, y, z, e_final, alpha, r, s, delta, eta, e_round, d_L, d_R):
    """Build an OrcParams/BppParams-layout struct from Python ints. Keeps byte buffers alive on the object."""
    p = cls()
    for k, v in (("y", y), ("z", z), ("e_final", e_final), ("alpha", alpha), ("r", r), ("s", s), ("delta", delta), ("eta", eta)):
        C.memmove(getattr(p, k), sc_bytes(v), 32)
    p.n_rounds = len(e_round)
    p._keep = [b"".join(sc_bytes(x) for x in lst) for lst in (e_round, d_L, d_R)]
    p.e_round, p.d_L, p.d_R = p._keep
    return p


class Oracle:
    def __init__(self):
        build()
        self.lib = C.CDLL(SO)
        L = self.lib
        sz, u8p, u64p = C.c_size_t, C.c_char_p, C.POINTER(C.c_uint64)
        pp = C.POINTER(OrcParams)
        L.orc_pk_new.argtypes = [C.c_int, sz, u8p, u8p, u8p, u8p]
        L.orc_set_msm_backend.argtypes = [C.c_int, C.c_int]
        L.orc_set_msm_backend.restype = None
        L.orc_b2_bench_msm.argtypes = [sz, C.c_int]
        L.orc_b2_bench_msm.restype = C.c_double
        L.orc_commit.argtypes = [C.c_int, sz, u64p, u8p, u8p]
        L.orc_msm_naive.argtypes = [C.c_int, sz, u8p, u8p, u8p]
        L.orc_scalar_mul.argtypes = [C.c_int, u8p, u8p, u8p]
        L.orc_point_add.argtypes = [C.c_int, u8p, u8p, u8p]
        L.orc_point_neg.argtypes = [C.c_int, u8p, u8p]
        L.orc_sc_op.argtypes = [C.c_int, C.c_int, u8p, u8p, u8p]
        L.orc_fq_op.argtypes = [C.c_int, C.c_int, u8p, u8p, u8p]
        L.orc_sc_reduce_wide.argtypes = [C.c_int, u8p, sz, u8p]
        L.orc_range_prove.argtypes = [C.c_int, sz, sz, u64p, u8p, pp, u8p, C.POINTER(sz)]
        L.orc_range_verify.argtypes = [C.c_int, sz, sz, u8p, u8p, sz, pp]
        L.orc_wip_prove.argtypes = [C.c_int, sz, u8p, u8p, u8p, u8p, pp, u8p, C.POINTER(sz)]
        L.orc_wip_verify.argtypes = [C.c_int, sz, u8p, u8p, u8p, u8p, sz, u8p, u8p, u8p, u8p, sz, pp]

    def set_msm_backend(self, backend: int, threads: int = 1):
        """CPU baselines only: 0 = the reference's per-term MulVec loop (B1), 1 = bucket MSM (B2, Ristretto255)."""
        self.lib.orc_set_msm_backend(backend, threads)

    def b2_bench_msm(self, n: int, threads: int) -> float:
        """seconds of one B2 bucket MSM over n synthetic Ristretto255 points (CPU baseline timing only)."""
        return float(self.lib.orc_b2_bench_msm(n, threads))

    @staticmethod
    def _chk(rc, what):
        if rc != 0:
            raise RuntimeError(f"oracle {what} failed rc={rc}")

    def pk_new(self, curve, length):
        pb = POINT_BYTES[curve]
        g, h = C.create_string_buffer(pb), C.create_string_buffer(pb)
        G, H = C.create_string_buffer(max(1, pb * length)), C.create_string_buffer(max(1, pb * length))
        self._chk(self.lib.orc_pk_new(curve, length, g, h, G, H), "pk_new")
        return g.raw, h.raw, G.raw[: pb * length], H.raw[: pb * length]

    def commit(self, curve, vs, gammas):
        pb = POINT_BYTES[curve]
        n = len(vs)
        arr = (C.c_uint64 * n)(*vs)
        out = C.create_string_buffer(pb * n)
        self._chk(self.lib.orc_commit(curve, n, arr, b"".join(sc_bytes(g) for g in gammas), out), "commit")
        return out.raw

    def msm_naive(self, curve, scalars: bytes, points: bytes):
        pb = POINT_BYTES[curve]
        k = len(scalars) // 32
        assert len(points) == k * pb
        out = C.create_string_buffer(pb)
        self._chk(self.lib.orc_msm_naive(curve, k, scalars, points, out), "msm_naive")
        return out.raw

    def scalar_mul(self, curve, k: int, point: bytes | None = None):
        out = C.create_string_buffer(POINT_BYTES[curve])
        self._chk(self.lib.orc_scalar_mul(curve, sc_bytes(k), point, out), "scalar_mul")
        return out.raw

    def point_add(self, curve, a, b):
        out = C.create_string_buffer(POINT_BYTES[curve])
        self._chk(self.lib.orc_point_add(curve, a, b, out), "point_add")
        return out.raw

    def point_neg(self, curve, a):
        out = C.create_string_buffer(POINT_BYTES[curve])
        self._chk(self.lib.orc_point_neg(curve, a, out), "point_neg")
        return out.raw

    def point_decode_ok(self, curve, a) -> bool:
        out = C.create_string_buffer(POINT_BYTES[curve])
        return self.lib.orc_point_neg(curve, a, out) == 0

    def sc_op(self, curve, op, a: int, b: int | None = None) -> int:
        out = C.create_string_buffer(32)
        self._chk(self.lib.orc_sc_op(curve, op, sc_bytes(a), None if b is None else sc_bytes(b), out), "sc_op")
        return int.from_bytes(out.raw, "little")

    def fq_op(self, curve, op, a: int, b: int | None = None) -> int:
        fb = 48 if curve == 2 else 32
        out = C.create_string_buffer(fb)
        bb = None if b is None else int(b).to_bytes(fb, "little")
        self._chk(self.lib.orc_fq_op(curve, op, int(a).to_bytes(fb, "little"), bb, out), "fq_op")
        return int.from_bytes(out.raw, "little")

    def range_prove(self, curve, n, m, vs, gammas, params=None) -> bytes:
        arr = (C.c_uint64 * m)(*vs)
        ln = C.c_size_t(1 << 20)
        out = C.create_string_buffer(ln.value)
        gam = gammas if isinstance(gammas, (bytes, bytearray)) else b"".join(sc_bytes(g) for g in gammas)
        self._chk(self.lib.orc_range_prove(curve, n, m, arr, gam, params, out, C.byref(ln)), "range_prove")
        return out.raw[: ln.value]

    def range_verify(self, curve, n, m, V: bytes, proof: bytes, params=None) -> int:
        return self.lib.orc_range_verify(curve, n, m, V, proof, len(proof), params)

    def wip_prove(self, curve, n, a: bytes, b: bytes, ypow: bytes, gamma: int, params=None) -> bytes:
        ln = C.c_size_t(1 << 20)
        out = C.create_string_buffer(ln.value)
        self._chk(self.lib.orc_wip_prove(curve, n, a, b, ypow, sc_bytes(gamma), params, out, C.byref(ln)), "wip_prove")
        return out.raw[: ln.value]

    def wip_verify(self, curve, n, ypow, G_exp_c, H_exp_c, g_exp_c: int, V_exp_c: bytes, A_prime: bytes, V: bytes, proof: bytes,
                   params=None) -> int:
        nV = len(V_exp_c) // 32
        return self.lib.orc_wip_verify(curve, n, ypow, G_exp_c, H_exp_c, sc_bytes(g_exp_c), nV, V_exp_c or None, A_prime,
                                       V or None, proof, len(proof), params)


_ORACLE = None


def get() -> Oracle:
    global _ORACLE
    if _ORACLE is None:
        _ORACLE = Oracle()
    return _ORACLE

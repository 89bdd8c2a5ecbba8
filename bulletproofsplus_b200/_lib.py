"""ctypes binding of libbpp_b200.so (include/bpp.h). Loading fails loudly when the library has not been
built; there is no CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.environ.get("BPP_LIB") or os.path.join(HERE, "libbpp_b200.so")  # BPP_LIB: kernel-variant experiments only

RISTRETTO255, SECP256K1, BLS12_381_G1 = 0, 1, 2
OK, ERR_VERIFICATION, ERR_FORMAT, ERR_LENGTH, ERR_NOT_POW2, ERR_GENERATORS, ERR_NOT_INIT, ERR_CUDA, ERR_UNSUPPORTED = range(9)
ERROR_NAMES = ["OK", "VERIFICATION", "FORMAT", "LENGTH", "NOT_POW2", "GENERATORS", "NOT_INIT", "CUDA", "UNSUPPORTED"]

# every symbol include/bpp.h declares (checked by tests/test_abi.py against the header and the .so)
SYMBOLS = [
    "bpp_init", "bpp_destroy", "bpp_last_error", "bpp_point_bytes", "bpp_range_proof_len", "bpp_wip_proof_len",
    "bpp_kernel_launches", "bpp_sync", "bpp_pk_new", "bpp_pk_new_hashed", "bpp_pk_precompute", "bpp_pk_msm", "bpp_pk_export", "bpp_pk_len", "bpp_pk_free", "bpp_commit_batch", "bpp_commit_batch_u64", "bpp_msm",
    "bpp_points_upload", "bpp_points_from_multipliers", "bpp_points_len", "bpp_points_precompute", "bpp_points_precompute_direct", "bpp_points_free", "bpp_msm_prepared",
    "bpp_msm_prepared_dev", "bpp_raw_point_bytes", "bpp_msm_prepared_raw", "bpp_points_sum_raw", "bpp_msm_prepared_raw_dev", "bpp_points_sum_raw_dev", "bpp_points_sum_raw_noalloc_dev", "bpp_range_prove_batch",
    "bpp_range_verify_batch", "bpp_range_prove_batch_dev", "bpp_range_verify_batch_dev", "bpp_range_prove", "bpp_range_verify",
    "bpp_wip_prove", "bpp_wip_verify", "bpp_range_verify_batch_fs", "bpp_range_verify_batch_rlc", "bpp_range_verify_batch_rlc_dev", "bpp_range_verify_batch_fs_dev", "bpp_range_prove_begin", "bpp_range_prove_step",
    "bpp_range_prove_finish", "bpp_prove_session_rounds", "bpp_prove_session_free", "bpp_int_peak_probe", "bpp_fp64_peak_probe", "bpp_fe_mul_probe", "bpp_hot_kernel_stats", "bpp_trace", "bpp_trace_dump", "bpp_debug_op",
]


class BppParams(C.Structure):
    _fields_ = [(k, C.c_uint8 * 32) for k in ("y", "z", "e_final", "alpha", "r", "s", "delta", "eta")] + [
        ("n_rounds", C.c_uint32), ("e_round", C.c_char_p), ("d_L", C.c_char_p), ("d_R", C.c_char_p)]


def build(force: bool = False) -> str:
    """Compile the CUDA library in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    srcs = [os.path.join(HERE, "csrc", f) for f in os.listdir(os.path.join(HERE, "csrc"))]
    srcs.append(os.path.join(HERE, "..", "include", "bpp.h"))
    if force or not os.path.exists(SO) or any(os.path.getmtime(s) > os.path.getmtime(SO) for s in srcs):
        subprocess.check_call(["make", f"-j{min(10, os.cpu_count() or 4)}", "-C", HERE, "libbpp_b200.so"])
    return SO


_LIB = None


def load() -> C.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(SO):
        raise RuntimeError(f"{SO} is missing: build it with `make -C bulletproofsplus_b200` (or __graft_entry__.build()); "
                           "bulletproofsplus_b200 has no CPU fallback")
    L = C.CDLL(SO)
    vp, sz, u8p, i = C.c_void_p, C.c_size_t, C.c_char_p, C.c_int
    pp = C.POINTER(BppParams)
    L.bpp_init.argtypes = [i, vp, C.POINTER(vp)]
    L.bpp_destroy.argtypes = [vp]; L.bpp_destroy.restype = None
    L.bpp_last_error.argtypes = [vp]; L.bpp_last_error.restype = C.c_char_p
    L.bpp_point_bytes.argtypes = [i]
    L.bpp_raw_point_bytes.argtypes = [i]
    L.bpp_range_proof_len.argtypes = [i, sz, sz]; L.bpp_range_proof_len.restype = sz
    L.bpp_wip_proof_len.argtypes = [i, sz]; L.bpp_wip_proof_len.restype = sz
    L.bpp_kernel_launches.argtypes = [vp]; L.bpp_kernel_launches.restype = C.c_uint64
    L.bpp_sync.argtypes = [vp]
    L.bpp_pk_new.argtypes = [vp, i, sz, C.POINTER(vp)]
    L.bpp_pk_new_hashed.argtypes = [vp, i, sz, u8p, sz, C.POINTER(vp)]
    L.bpp_pk_precompute.argtypes = [vp, vp, sz, C.POINTER(C.c_int)]
    L.bpp_pk_msm.argtypes = [vp, vp, u8p, sz, u8p]
    L.bpp_pk_export.argtypes = [vp, vp, u8p, u8p, u8p, u8p]
    L.bpp_pk_len.argtypes = [vp]; L.bpp_pk_len.restype = sz
    L.bpp_pk_free.argtypes = [vp]; L.bpp_pk_free.restype = None
    L.bpp_commit_batch.argtypes = [vp, vp, C.POINTER(C.c_uint64), u8p, sz, u8p]
    L.bpp_commit_batch_u64.argtypes = [vp, vp, C.POINTER(C.c_uint64), u8p, sz, u8p]
    L.bpp_msm.argtypes = [vp, i, u8p, u8p, sz, u8p]
    L.bpp_points_upload.argtypes = [vp, i, u8p, sz, C.POINTER(vp)]
    L.bpp_points_from_multipliers.argtypes = [vp, i, u8p, sz, C.POINTER(vp)]
    L.bpp_points_len.argtypes = [vp]; L.bpp_points_len.restype = sz
    L.bpp_points_precompute.argtypes = [vp, vp]
    L.bpp_points_precompute_direct.argtypes = [vp, vp, sz, C.POINTER(C.c_int)]
    L.bpp_points_free.argtypes = [vp]; L.bpp_points_free.restype = None
    L.bpp_msm_prepared.argtypes = [vp, vp, u8p, sz, sz, u8p]
    L.bpp_msm_prepared_dev.argtypes = [vp, vp, vp, sz, sz, vp]
    L.bpp_msm_prepared_raw.argtypes = [vp, vp, u8p, sz, u8p]
    L.bpp_points_sum_raw.argtypes = [vp, i, u8p, sz, u8p]
    L.bpp_msm_prepared_raw_dev.argtypes = [vp, vp, vp, sz, vp]
    L.bpp_points_sum_raw_dev.argtypes = [vp, i, vp, sz, vp]
    L.bpp_points_sum_raw_noalloc_dev.argtypes = [vp, i, vp, sz, vp]
    L.bpp_range_prove_batch.argtypes = [vp, vp, sz, sz, sz, C.POINTER(C.c_uint64), u8p, pp, u8p]
    L.bpp_range_verify_batch.argtypes = [vp, vp, sz, sz, sz, u8p, u8p, pp, u8p]
    L.bpp_range_prove_batch_dev.argtypes = [vp, vp, sz, sz, sz, vp, vp, pp, vp]
    L.bpp_range_verify_batch_dev.argtypes = [vp, vp, sz, sz, sz, vp, vp, pp, vp]
    L.bpp_range_prove.argtypes = [vp, vp, sz, sz, C.POINTER(C.c_uint64), u8p, pp, u8p, C.POINTER(sz)]
    L.bpp_range_verify.argtypes = [vp, vp, sz, sz, u8p, u8p, sz, pp]
    L.bpp_wip_prove.argtypes = [vp, vp, u8p, u8p, u8p, u8p, pp, u8p, C.POINTER(sz)]
    L.bpp_wip_verify.argtypes = [vp, vp, u8p, u8p, u8p, u8p, sz, u8p, u8p, u8p, u8p, sz, pp]
    L.bpp_range_verify_batch_fs.argtypes = [vp, vp, sz, sz, sz, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p]
    L.bpp_range_verify_batch_fs_dev.argtypes = [vp, vp, sz, sz, sz, vp, vp, vp, vp]
    L.bpp_range_verify_batch_rlc.argtypes = [vp, vp, sz, sz, sz, C.c_char_p, C.c_char_p, pp, C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_int)]
    L.bpp_range_verify_batch_rlc_dev.argtypes = [vp, vp, sz, sz, sz, vp, vp, pp, vp, vp, vp, C.POINTER(C.c_int)]
    L.bpp_range_prove_begin.argtypes = [vp, vp, sz, sz, sz, C.POINTER(C.c_uint64), C.c_char_p, C.c_char_p, C.POINTER(vp), C.c_char_p]
    L.bpp_range_prove_step.argtypes = [vp, vp, C.c_char_p, C.c_char_p]
    L.bpp_range_prove_finish.argtypes = [vp, vp, C.c_char_p, C.c_char_p]
    L.bpp_prove_session_rounds.argtypes = [vp]
    L.bpp_prove_session_free.argtypes = [vp]; L.bpp_prove_session_free.restype = None
    L.bpp_int_peak_probe.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.bpp_fp64_peak_probe.argtypes = [vp, C.POINTER(C.c_double)]
    L.bpp_fe_mul_probe.argtypes = [vp, i, i, i, C.POINTER(C.c_double)]
    L.bpp_hot_kernel_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64), i]
    L.bpp_trace.argtypes = [vp, i]
    L.bpp_trace_dump.argtypes = [vp, C.c_char_p, sz]
    L.bpp_debug_op.argtypes = [vp, i, i, sz, u8p, u8p, u8p, u8p]
    _LIB = L
    return L

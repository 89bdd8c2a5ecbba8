"""CPU-only checks of the boundary: the C-ABI library loads, exports every symbol include/bpp.h declares, and the
product path fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes as C
import json
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "bpp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bpp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from bulletproofsplus_b200 import _lib
    _lib.build()
    lib = _lib.load()
    declared = header_symbols()
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/bpp.h but not exported"
    assert sorted(_lib.SYMBOLS) == declared


def test_static_queries_need_no_gpu():
    from bulletproofsplus_b200 import _lib
    lib = _lib.load()
    assert [lib.bpp_point_bytes(c) for c in (0, 1, 2, 3)] == [32, 64, 96, -1]
    assert lib.bpp_range_proof_len(0, 64, 16) == (2 * 10 + 3) * 32 + 96  # 23 points + 3 scalars
    assert lib.bpp_range_proof_len(0, 64, 1) == 15 * 32 + 96
    assert lib.bpp_range_proof_len(0, 24, 1) == 0
    assert lib.bpp_wip_proof_len(0, 4096) == 26 * 32 + 96


def test_params_struct_layout_matches_oracle():
    from bulletproofsplus_b200._lib import BppParams
    from oracle.oracle_lib import OrcParams
    assert C.sizeof(BppParams) == C.sizeof(OrcParams) == 8 * 32 + 8 + 3 * 8
    for (n1, _), (n2, _) in zip(BppParams._fields_, OrcParams._fields_):
        assert n1 == n2 and getattr(BppParams, n1).offset == getattr(OrcParams, n2).offset


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from bulletproofsplus_b200 import api
    with pytest.raises(api.BppError) as ei:
        api.Context(0)
    assert ei.value.rc == 7


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "bulletproofsplus_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.lower() or f == "__init__.py" and False, f"{f} mentions the oracle"


def test_golden_fixtures_match_oracle(oracle):
    fx = json.load(open(os.path.join(ROOT, "tests", "golden", "ristretto_proofs.json")))
    for name in ("main_rs", "n64_m1"):
        e = fx[name]
        gs = [int(g, 16) if isinstance(g, str) else g for g in e["gamma"]]
        assert oracle.range_prove(0, e["n"], e["m"], e["v"], gs).hex() == e["proof"]
        assert oracle.range_verify(0, e["n"], e["m"], bytes.fromhex(e["V"]), bytes.fromhex(e["proof"])) == 0

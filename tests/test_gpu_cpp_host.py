"""The C++ host mirror (include/bpp.hpp) on the GPU: the reference's main.rs and the crate-style test program, built by
`make -C bulletproofsplus_b200 host` (g++ over libbpp_b200.so). Their printed encodings are compared with the CPU oracle
and with the committed main.rs golden fixture."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "bulletproofsplus_b200", "bin")
ORD = {0: 2**252 + 27742317777372353535851937790883648493,
       1: 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141,
       2: 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001}


SOURCES = {"main": os.path.join(ROOT, "examples", "main.cpp"), "host_mirror_test": os.path.join(ROOT, "tests", "cpp", "host_mirror_test.cpp")}


def _build(force=False):
    """g++ directly (what `make -C bulletproofsplus_b200 host` runs), never make: on the GPU box the prebuilt library and
    programs travel with the snapshot and make would rebuild the CUDA objects there."""
    os.makedirs(BIN, exist_ok=True)
    for name, src in SOURCES.items():
        exe = os.path.join(BIN, name)
        if force or not os.access(exe, os.X_OK):
            subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), "-o", exe, src,
                                   "-L" + os.path.join(ROOT, "bulletproofsplus_b200"), "-lbpp_b200", "-Wl,-rpath,$ORIGIN/.."])


def _run(name, *args):
    p = subprocess.run([os.path.join(BIN, name), *map(str, args)], capture_output=True, text=True, timeout=600)
    return p.returncode, p.stdout


def _fields(out):
    return dict(line.split("=", 1) for line in out.splitlines() if "=" in line and not line.startswith(("PASS", "FAIL")))


def test_host_programs_build():
    """CPU side: the header compiles as C++17 and the programs link against the library (no device needed to link)."""
    _build(force=True)
    assert os.access(os.path.join(BIN, "main"), os.X_OK) and os.access(os.path.join(BIN, "host_mirror_test"), os.X_OK)


@pytest.mark.gpu
@pytest.mark.parametrize("curve", [0, 1, 2])
def test_main_rs_through_cpp_host(oracle, curve):
    _build()
    rc, out = _run("main", curve)
    assert rc == 0, out
    assert "result=Ok(())" in out
    proof = bytes.fromhex(_fields(out)["proof"])
    assert proof == oracle.range_prove(curve, 64, 2, [2, 5], [3, 7])  # main.rs:6-57 inputs
    if curve == 0:
        fx = json.load(open(os.path.join(ROOT, "tests", "golden", "ristretto_proofs.json")))["main_rs"]
        assert proof.hex() == fx["proof"]


@pytest.mark.gpu
@pytest.mark.parametrize("curve", [0, 1, 2])
def test_crate_style_checks_through_cpp_host(oracle, curve):
    _build()
    rc, out = _run("host_mirror_test", curve)
    lines = [l for l in out.splitlines() if l.startswith(("PASS", "FAIL"))]
    assert rc == 0 and lines and all(l.startswith("PASS") for l in lines), out
    assert len(lines) >= 30
    f = _fields(out)
    q = ORD[curve]
    for n, m in ((64, 2), (32, 1), (8, 4)):
        vs = [2 + 3 * j for j in range(m)]
        gs = [3 + 4 * j for j in range(m)]
        assert bytes.fromhex(f[f"proof_n{n}_m{m}"]) == oracle.range_prove(curve, n, m, vs, gs), (n, m)
    N = 16
    enc = lambda v: b"".join(int(x % q).to_bytes(32, "little") for x in v)
    a, b, yp = [i + 1 for i in range(N)], [2 * i + 3 for i in range(N)], [2 ** (i + 1) for i in range(N)]
    assert bytes.fromhex(f["wip_N16"]) == oracle.wip_prove(curve, N, enc(a), enc(b), enc(yp), 9)
    # the C++ host's transcript proof equals the Python host's for the same label and statement (two Merlin implementations,
    # one protocol), and the oracle's proof under the challenges the transcript yields
    from bulletproofsplus_b200 import api, merlin
    n, m = 32, 4
    pk = api.PublicKey.new(n * m, curve, api.Arith.init())
    vs, gs = [31 + j for j in range(m)], [11 + j for j in range(m)]
    py = api.prove_batch_fs(pk, n, m, vs, gs, [merlin.Transcript(b"RangeProof Test")], api.REFERENCE_NONCES)[0]
    assert bytes.fromhex(f["transcript_proof_n32_m4"]) == py

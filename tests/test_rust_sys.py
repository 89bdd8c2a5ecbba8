"""The Rust crates (bpp-sys + bulletproofs-plus-gpu) cannot be compiled here (no Rust toolchain), so they are
checked mechanically: bpp-sys/src/lib.rs is exactly what tools/gen_bpp_sys.py derives from include/bpp.h, its `extern "C"` block
binds every function the header declares — same names, same argument counts — and every symbol the wrapper crate calls exists.
CPU only."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_bpp_sys  # noqa: E402


def _rust_fns(src):
    return {m.group(1): m.group(2) for m in re.finditer(r"pub fn (bpp_\w+)\((.*?)\)( -> [^;]+)?;", src)}


def test_committed_bindings_are_what_the_header_generates():
    assert open(os.path.join(ROOT, "bpp-sys", "src", "lib.rs")).read() == gen_bpp_sys.generate()


def test_extern_block_covers_every_header_function():
    decls = gen_bpp_sys.parse_header()
    rust = _rust_fns(open(os.path.join(ROOT, "bpp-sys", "src", "lib.rs")).read())
    assert set(rust) == {d[0] for d in decls}
    for name, _, args in decls:
        assert len([a for a in rust[name].split(",") if a.strip()]) == len(args), name
    from bulletproofsplus_b200 import _lib
    assert set(rust) == set(_lib.SYMBOLS)  # the ctypes binding and the Rust binding name the same ABI


def test_wrapper_only_calls_bound_symbols_and_has_no_stubs():
    rust = _rust_fns(open(os.path.join(ROOT, "bpp-sys", "src", "lib.rs")).read())
    src = open(os.path.join(ROOT, "bulletproofs-plus-gpu", "src", "lib.rs")).read()
    used = set(re.findall(r"\b(bpp_[a-z0-9_]+)\(", src))
    assert used and used <= set(rust), used - set(rust)
    assert "unimplemented!" not in src and "todo!" not in src
    for name in ("PublicKey", "RangeProver", "RangeVerifier", "RangeProof", "WeightedInnerProductProof", "MulVec", "ProofError", "Arith"):
        assert re.search(rf"pub (struct|enum) {name}\b", src), name  # the crate surface of the reference's src/lib.rs:11-13
    for fn in ("pub fn new(len: usize)", "pub fn commit(&mut self", "pub fn allocate(&mut self", "pub fn prove(pk: &PublicKey, n: usize, prover: &RangeProver)",
               "pub fn verify(&self, pk: &PublicKey, n: usize, commitment_vec: &[Point])", "pub fn prove_with_transcript", "pub fn verify_with_transcript"):
        assert fn in src, fn

"""Generates bpp-sys/src/lib.rs (the raw Rust bindings of include/bpp.h) from the header itself, so that the `extern "C"` block
cannot drift from the ABI. `python tools/gen_bpp_sys.py` rewrites the file; tests/test_rust_sys.py checks that the committed file
equals what this script produces and that every exported symbol of libbpp_b200.so is bound."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OPAQUE = ["bpp_ctx", "bpp_pk", "bpp_points", "bpp_prove_session"]
SCALAR = {"int": "c_int", "size_t": "usize", "uint64_t": "u64", "uint32_t": "u32", "double": "f64", "char": "c_char", "uint8_t": "u8",
          "void": "c_void", "bpp_params": "bpp_params", **{o: o for o in OPAQUE}}


def parse_header(path=None):
    """[(name, return C type, [(C type, arg name)])] for every function include/bpp.h declares."""
    h = open(path or os.path.join(ROOT, "include", "bpp.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    out = []
    for full, name, args in re.findall(r"\n\s*([A-Za-z_][\w\s\*]*?\b(bpp_\w+))\s*\(([^;{]*?)\)\s*;", h):
        ret = " ".join(full.replace(name, "").split())
        alist = []
        for a in " ".join(args.split()).split(","):
            a = a.strip()
            m = re.match(r"(.*?)(\w+)$", a)
            alist.append((m.group(1).strip().replace(" *", "*"), m.group(2)))
        out.append((name, ret, alist))
    return out


def rust_type(c: str) -> str:
    c = c.strip()
    const = c.startswith("const ")
    base = c[6:] if const else c
    stars = base.count("*")
    base = base.replace("*", "").strip()
    t = SCALAR[base]
    for i in range(stars):
        t = ("*const " if (const and i == 0) else "*mut ") + t
    return t


def generate() -> str:
    lines = ['// bpp-sys: raw bindings to libbpp_b200.so. GENERATED from include/bpp.h by tools/gen_bpp_sys.py — do not edit.',
             '// The reference (gogoex/BulletProofsPlus) reaches its arithmetic through per-operation mcl_rust FFI calls (call sites listed in',
             '// SURVEY.md §8(b)); this crate is what replaces that dependency: one call per batch of commitments / proofs / MSMs.',
             '// Uncompiled here (no Rust toolchain in the build image); checked against the header and the .so by tests/test_rust_sys.py.',
             '#![allow(non_camel_case_types, non_snake_case)]',
             'use std::os::raw::{c_char, c_int, c_void};', '']
    for o in OPAQUE:
        lines.append(f'#[repr(C)] pub struct {o} {{ _private: [u8; 0] }}')
    h = open(os.path.join(ROOT, "include", "bpp.h")).read()
    lines.append('')
    for enum in re.findall(r"enum\s*\{(.*?)\};", re.sub(r"/\*.*?\*/", "", h, flags=re.S), flags=re.S):
        for item in enum.split(","):
            item = item.strip()
            if item:
                k, v = [x.strip() for x in item.split("=")]
                lines.append(f'pub const {k}: c_int = {v};')
    m = re.search(r"#define BPP_NOALLOC_MAX_POINTS (\d+)", h)
    lines.append(f'pub const BPP_NOALLOC_MAX_POINTS: usize = {m.group(1)};')
    lines += ['', '/// Fiat-Shamir challenges and prover nonces of one call; NULL = the reference snapshot\'s constants (parity mode).',
              '#[repr(C)]', 'pub struct bpp_params {',
              '    pub y: [u8; 32], pub z: [u8; 32], pub e_final: [u8; 32], pub alpha: [u8; 32],',
              '    pub r: [u8; 32], pub s: [u8; 32], pub delta: [u8; 32], pub eta: [u8; 32],',
              '    pub n_rounds: u32,', '    pub e_round: *const u8, pub d_L: *const u8, pub d_R: *const u8,', '}', '',
              '#[link(name = "bpp_b200")]', 'extern "C" {']
    for name, ret, args in parse_header():
        a = ", ".join(f"{n}: {rust_type(t)}" for t, n in args)
        r = "" if ret == "void" else f" -> {rust_type(ret)}"
        lines.append(f"    pub fn {name}({a}){r};")
    lines += ['}', '']
    return "\n".join(lines)


if __name__ == "__main__":
    path = os.path.join(ROOT, "bpp-sys", "src", "lib.rs")
    open(path, "w").write(generate())
    print("wrote", path)

// Finds libbpp_b200.so: BPP_B200_LIB_DIR, else ../bulletproofsplus_b200 (the in-tree build: `make -C bulletproofsplus_b200`).
// The library itself is built by nvcc for sm_100a (bulletproofsplus_b200/Makefile); cargo only links it.
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("BPP_B200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("..").join("bulletproofsplus_b200")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=bpp_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=BPP_B200_LIB_DIR");
    println!("cargo:rerun-if-changed=../include/bpp.h");
}

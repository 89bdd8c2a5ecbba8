// bpp-sys: raw bindings to libbpp_b200.so. GENERATED from include/bpp.h by tools/gen_bpp_sys.py — do not edit.
// The reference (gogoex/BulletProofsPlus) reaches its arithmetic through per-operation mcl_rust FFI calls (call sites listed in
// SURVEY.md §8(b)); this crate is what replaces that dependency: one call per batch of commitments / proofs / MSMs.
// Uncompiled here (no Rust toolchain in the build image); checked against the header and the .so by tests/test_rust_sys.py.
#![allow(non_camel_case_types, non_snake_case)]
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)] pub struct bpp_ctx { _private: [u8; 0] }
#[repr(C)] pub struct bpp_pk { _private: [u8; 0] }
#[repr(C)] pub struct bpp_points { _private: [u8; 0] }
#[repr(C)] pub struct bpp_prove_session { _private: [u8; 0] }

pub const BPP_RISTRETTO255: c_int = 0;
pub const BPP_SECP256K1: c_int = 1;
pub const BPP_BLS12_381_G1: c_int = 2;
pub const BPP_OK: c_int = 0;
pub const BPP_ERR_VERIFICATION: c_int = 1;
pub const BPP_ERR_FORMAT: c_int = 2;
pub const BPP_ERR_LENGTH: c_int = 3;
pub const BPP_ERR_NOT_POW2: c_int = 4;
pub const BPP_ERR_GENERATORS: c_int = 5;
pub const BPP_ERR_NOT_INIT: c_int = 6;
pub const BPP_ERR_CUDA: c_int = 7;
pub const BPP_ERR_UNSUPPORTED: c_int = 8;
pub const BPP_NOALLOC_MAX_POINTS: usize = 65536;

/// Fiat-Shamir challenges and prover nonces of one call; NULL = the reference snapshot's constants (parity mode).
#[repr(C)]
pub struct bpp_params {
    pub y: [u8; 32], pub z: [u8; 32], pub e_final: [u8; 32], pub alpha: [u8; 32],
    pub r: [u8; 32], pub s: [u8; 32], pub delta: [u8; 32], pub eta: [u8; 32],
    pub n_rounds: u32,
    pub e_round: *const u8, pub d_L: *const u8, pub d_R: *const u8,
}

#[link(name = "bpp_b200")]
extern "C" {
    pub fn bpp_init(device: c_int, cuda_stream: *mut c_void, out: *mut *mut bpp_ctx) -> c_int;
    pub fn bpp_destroy(ctx: *mut bpp_ctx);
    pub fn bpp_last_error(ctx: *const bpp_ctx) -> *const c_char;
    pub fn bpp_point_bytes(curve: c_int) -> c_int;
    pub fn bpp_range_proof_len(curve: c_int, n: usize, m: usize) -> usize;
    pub fn bpp_wip_proof_len(curve: c_int, len: usize) -> usize;
    pub fn bpp_kernel_launches(ctx: *const bpp_ctx) -> u64;
    pub fn bpp_sync(ctx: *mut bpp_ctx) -> c_int;
    pub fn bpp_pk_new(ctx: *mut bpp_ctx, curve: c_int, len: usize, out: *mut *mut bpp_pk) -> c_int;
    pub fn bpp_pk_new_hashed(ctx: *mut bpp_ctx, curve: c_int, len: usize, domain_sep: *const u8, domain_sep_len: usize, out: *mut *mut bpp_pk) -> c_int;
    pub fn bpp_pk_precompute(ctx: *mut bpp_ctx, pk: *mut bpp_pk, max_bytes: usize, window_bits: *mut c_int) -> c_int;
    pub fn bpp_pk_msm(ctx: *mut bpp_ctx, pk: *const bpp_pk, scalars: *const u8, n_jobs: usize, out_points: *mut u8) -> c_int;
    pub fn bpp_pk_export(ctx: *mut bpp_ctx, pk: *const bpp_pk, g: *mut u8, h: *mut u8, G: *mut u8, H: *mut u8) -> c_int;
    pub fn bpp_pk_len(pk: *const bpp_pk) -> usize;
    pub fn bpp_pk_free(pk: *mut bpp_pk);
    pub fn bpp_commit_batch(ctx: *mut bpp_ctx, pk: *const bpp_pk, v: *const u64, gamma: *const u8, count: usize, out_points: *mut u8) -> c_int;
    pub fn bpp_commit_batch_u64(ctx: *mut bpp_ctx, pk: *const bpp_pk, v: *const u64, gamma: *const u8, count: usize, out_points: *mut u8) -> c_int;
    pub fn bpp_msm(ctx: *mut bpp_ctx, curve: c_int, scalars: *const u8, points: *const u8, k: usize, out_point: *mut u8) -> c_int;
    pub fn bpp_points_upload(ctx: *mut bpp_ctx, curve: c_int, points: *const u8, k: usize, out: *mut *mut bpp_points) -> c_int;
    pub fn bpp_points_from_multipliers(ctx: *mut bpp_ctx, curve: c_int, mult_scalars: *const u8, k: usize, out: *mut *mut bpp_points) -> c_int;
    pub fn bpp_points_len(pts: *const bpp_points) -> usize;
    pub fn bpp_points_free(pts: *mut bpp_points);
    pub fn bpp_points_precompute(ctx: *mut bpp_ctx, pts: *mut bpp_points) -> c_int;
    pub fn bpp_points_precompute_direct(ctx: *mut bpp_ctx, pts: *mut bpp_points, max_bytes: usize, window_bits: *mut c_int) -> c_int;
    pub fn bpp_msm_prepared(ctx: *mut bpp_ctx, pts: *const bpp_points, scalars: *const u8, k: usize, n_jobs: usize, out_points: *mut u8) -> c_int;
    pub fn bpp_msm_prepared_dev(ctx: *mut bpp_ctx, pts: *const bpp_points, d_scalars: *const c_void, k: usize, n_jobs: usize, d_out_points: *mut c_void) -> c_int;
    pub fn bpp_raw_point_bytes(curve: c_int) -> c_int;
    pub fn bpp_msm_prepared_raw(ctx: *mut bpp_ctx, pts: *const bpp_points, scalars: *const u8, k: usize, out_raw: *mut u8) -> c_int;
    pub fn bpp_points_sum_raw(ctx: *mut bpp_ctx, curve: c_int, raw_points: *const u8, count: usize, out_point: *mut u8) -> c_int;
    pub fn bpp_msm_prepared_raw_dev(ctx: *mut bpp_ctx, pts: *const bpp_points, d_scalars: *const c_void, k: usize, d_out_raw: *mut c_void) -> c_int;
    pub fn bpp_points_sum_raw_dev(ctx: *mut bpp_ctx, curve: c_int, d_raw_points: *const c_void, count: usize, d_out_point: *mut c_void) -> c_int;
    pub fn bpp_points_sum_raw_noalloc_dev(ctx: *mut bpp_ctx, curve: c_int, d_raw_points: *const c_void, count: usize, d_out_point: *mut c_void) -> c_int;
    pub fn bpp_range_prove_batch(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, v: *const u64, gamma: *const u8, params: *const bpp_params, proofs: *mut u8) -> c_int;
    pub fn bpp_range_verify_batch(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, V: *const u8, proofs: *const u8, params: *const bpp_params, verdicts: *mut u8) -> c_int;
    pub fn bpp_range_prove_batch_dev(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, d_v: *const c_void, d_gamma: *const c_void, params: *const bpp_params, d_proofs: *mut c_void) -> c_int;
    pub fn bpp_range_verify_batch_dev(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, d_V: *const c_void, d_proofs: *const c_void, params: *const bpp_params, d_verdicts: *mut c_void) -> c_int;
    pub fn bpp_range_prove(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, v: *const u64, gamma: *const u8, params: *const bpp_params, proof: *mut u8, proof_len: *mut usize) -> c_int;
    pub fn bpp_range_verify(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, V: *const u8, proof: *const u8, proof_len: usize, params: *const bpp_params) -> c_int;
    pub fn bpp_range_verify_batch_fs(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, V: *const u8, proofs: *const u8, challenges: *const u8, verdicts: *mut u8) -> c_int;
    pub fn bpp_range_verify_batch_fs_dev(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, d_V: *const c_void, d_proofs: *const c_void, d_challenges: *const c_void, d_verdicts: *mut c_void) -> c_int;
    pub fn bpp_range_verify_batch_rlc(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, V: *const u8, proofs: *const u8, params: *const bpp_params, challenges: *const u8, weights: *const u8, verdicts: *mut u8, used_fallback: *mut c_int) -> c_int;
    pub fn bpp_range_verify_batch_rlc_dev(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, d_V: *const c_void, d_proofs: *const c_void, params: *const bpp_params, d_challenges: *const c_void, d_weights: *const c_void, d_verdicts: *mut c_void, used_fallback: *mut c_int) -> c_int;
    pub fn bpp_range_prove_begin(ctx: *mut bpp_ctx, pk: *const bpp_pk, n: usize, m: usize, count: usize, v: *const u64, gamma: *const u8, nonces: *const u8, out: *mut *mut bpp_prove_session, out_A: *mut u8) -> c_int;
    pub fn bpp_range_prove_step(ctx: *mut bpp_ctx, session: *mut bpp_prove_session, challenges: *const u8, out_points: *mut u8) -> c_int;
    pub fn bpp_range_prove_finish(ctx: *mut bpp_ctx, session: *mut bpp_prove_session, e_final: *const u8, proofs: *mut u8) -> c_int;
    pub fn bpp_prove_session_rounds(session: *const bpp_prove_session) -> c_int;
    pub fn bpp_prove_session_free(session: *mut bpp_prove_session);
    pub fn bpp_wip_prove(ctx: *mut bpp_ctx, pk: *const bpp_pk, a: *const u8, b: *const u8, power_of_y: *const u8, gamma: *const u8, params: *const bpp_params, proof: *mut u8, proof_len: *mut usize) -> c_int;
    pub fn bpp_wip_verify(ctx: *mut bpp_ctx, pk: *const bpp_pk, power_of_y: *const u8, G_exp_of_commitment: *const u8, H_exp_of_commitment: *const u8, g_exp_of_commitment: *const u8, n_V: usize, V_exp_of_commitment: *const u8, A_prime: *const u8, V: *const u8, proof: *const u8, proof_len: usize, params: *const bpp_params) -> c_int;
    pub fn bpp_int_peak_probe(ctx: *mut bpp_ctx, rates3: *mut f64, elapsed_ms: *mut f64) -> c_int;
    pub fn bpp_fe_mul_probe(ctx: *mut bpp_ctx, ilp: c_int, threads_per_block: c_int, blocks_per_sm: c_int, muls_per_sec: *mut f64) -> c_int;
    pub fn bpp_fp64_peak_probe(ctx: *mut bpp_ctx, rates2: *mut f64) -> c_int;
    pub fn bpp_hot_kernel_stats(ctx: *mut bpp_ctx, ms: *mut f64, launches: *mut u64, reset: c_int) -> c_int;
    pub fn bpp_trace(ctx: *mut bpp_ctx, on: c_int) -> c_int;
    pub fn bpp_trace_dump(ctx: *mut bpp_ctx, buf: *mut c_char, cap: usize) -> c_int;
    pub fn bpp_debug_op(ctx: *mut bpp_ctx, curve: c_int, op: c_int, n: usize, a: *const u8, b: *const u8, out: *mut u8, ok: *mut u8) -> c_int;
}

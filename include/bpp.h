/* bpp.h — C ABI of the B200-native Bulletproofs+ hot path (libbpp_b200.so).
 *
 * Drop-in boundary for gogoex/BulletProofsPlus: each entry point replaces one function of the
 * reference crate (paths relative to the reference's src/). A Rust `bpp-sys` crate binds exactly
 * these symbols. No torch / C++ types cross this boundary; every buffer is
 * caller-owned; the library owns only the opaque handles. Nothing here falls back to the CPU: with
 * no CUDA device bpp_init fails with BPP_ERR_CUDA.
 *
 * Encodings
 *   scalar           32 bytes little-endian, canonical (< group order), else BPP_ERR_FORMAT
 *   point, curve 0   Ristretto255: 32-byte canonical encoding (RFC 9496)
 *   point, curve 1   secp256k1: x||y, 32+32 bytes little-endian; infinity = 64 zero bytes
 *   point, curve 2   BLS12-381 G1: x||y, 48+48 bytes little-endian; infinity = 96 zero bytes
 *   range proof      A | L[0..L) | R[0..L) | wipA | wipB | r' | s' | delta'   (L = log2(n*m))
 *                    = RangeProof{A, proof: WeightedInnerProductProof{L_vec,R_vec,A,B,r_prime,s_prime,d_prime}}
 *                      (range/mod.rs:24-28, weighted_inner_product_proof.rs:25-33)
 *   WIP proof        the same without the leading A
 */
#ifndef BPP_H
#define BPP_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bpp_ctx bpp_ctx;       /* one per host thread / GPU: stream + scratch arena. Not thread-safe.     */
typedef struct bpp_pk bpp_pk;         /* PublicKey (publickey.rs:13-18): device-resident generator table         */
typedef struct bpp_points bpp_points; /* a decoded, device-resident point vector for repeated MSMs               */
typedef struct bpp_prove_session bpp_prove_session; /* device state of an interactive batch prove (host-owned transcript) */

enum { BPP_RISTRETTO255 = 0, BPP_SECP256K1 = 1, BPP_BLS12_381_G1 = 2 };

enum {
  BPP_OK = 0,
  BPP_ERR_VERIFICATION = 1, /* ProofError::VerificationError (errors.rs:17; range/mod.rs:508, wip.rs:326,336) */
  BPP_ERR_FORMAT = 2,       /* non-canonical scalar / invalid point / wrong proof length                        */
  BPP_ERR_LENGTH = 3,       /* reference panics: mulvec.rs:23-25, wip.rs:61-64                                   */
  BPP_ERR_NOT_POW2 = 4,     /* reference panic: wip.rs:67                                                        */
  BPP_ERR_GENERATORS = 5,   /* reference panics: range/mod.rs:90-91, :252-253                                    */
  BPP_ERR_NOT_INIT = 6,     /* reference panic: arith.rs:12-14                                                   */
  BPP_ERR_CUDA = 7,         /* see bpp_last_error                                                                */
  BPP_ERR_UNSUPPORTED = 8
};

/* Fiat-Shamir challenges and prover nonces. NULL everywhere below = the constants hard-coded in the
 * reference snapshot (parity mode): y,z = 7,7 (m == 1) or 12,23 (range/mod.rs:109-110, :278-279);
 * alpha = 7 / 33 (:94, :256); round e = 7, d_L = 4, d_R = 5 (wip.rs:131, :94-95); r,s,delta,eta =
 * 33,44,88,123 (:175-178); final e = 99 (:211). The host (Merlin transcript) owns challenge derivation. */
typedef struct bpp_params {
  uint8_t y[32], z[32], e_final[32], alpha[32], r[32], s[32], delta[32], eta[32];
  uint32_t n_rounds;      /* entries in each array below, >= log2(n*m) */
  const uint8_t* e_round; /* n_rounds * 32 bytes */
  const uint8_t* d_L;
  const uint8_t* d_R;
} bpp_params;

/* ---- context: Arith::init (bls12_381/building_block/arith.rs:9-17) ------------------------------- */
/* cuda_stream: a cudaStream_t to launch on (e.g. the caller's current stream), or NULL for a private one. */
int bpp_init(int device, void* cuda_stream, bpp_ctx** out);
void bpp_destroy(bpp_ctx* ctx);
const char* bpp_last_error(const bpp_ctx* ctx);
int bpp_point_bytes(int curve);
size_t bpp_range_proof_len(int curve, size_t n, size_t m);
size_t bpp_wip_proof_len(int curve, size_t len);
uint64_t bpp_kernel_launches(const bpp_ctx* ctx); /* kernels launched by this context so far */
int bpp_sync(bpp_ctx* ctx);

/* ---- PublicKey::new / fields (publickey.rs:21-48) ------------------------------------------------ */
int bpp_pk_new(bpp_ctx* ctx, int curve, size_t len, bpp_pk** out);
/* The same key shape with HASH-DERIVED generators (no counterpart in the snapshot, whose generators are the
 * placeholder multiples g, 2g, 3(i+1)g, 5(i+1)g of publickey.rs:21-48 with every discrete logarithm known — fine for parity,
 * unsound for anything else). Generator X_i = RFC 9496 §4.3.4 element derivation of
 * SHA-512("bpp-b200 generators v1" | le32(domain_sep_len) | domain_sep | label | le64(i)), label = 'g', 'h' (i = 0), 'G', 'H'.
 * BPP_RISTRETTO255 only (BPP_ERR_UNSUPPORTED otherwise). Everything else (commit, prove, verify, MSM) takes the key as usual. */
int bpp_pk_new_hashed(bpp_ctx* ctx, int curve, size_t len, const uint8_t* domain_sep, size_t domain_sep_len, bpp_pk** out);
/* Optional, once per key: DIRECT fixed-base tables — every signed-digit multiple of every window point of every generator, in
 * HBM (no counterpart in the reference, whose PublicKey is four vectors of points, publickey.rs:13-18). With them every call
 * against the key (one proof or thousands) evaluates its generator terms with one table addition per window and no bucket sort or reduction;
 * results are the same group elements (bit-identical encodings). max_bytes bounds the table: the widest window c in [8, 16]
 * that fits is used and returned in *window_bits ((ceil(252/c) + 1) * 2^(c-1) * 128 B per generator: 40 GB for n*m = 1024 at
 * c = 14, 8.7 GB for n*m = 64 at c = 16). BPP_RISTRETTO255 only (BPP_ERR_UNSUPPORTED otherwise); BPP_ERR_LENGTH if not even
 * c = 8 fits. The key must not be in use by another context during the call. */
int bpp_pk_precompute(bpp_ctx* ctx, bpp_pk* pk, size_t max_bytes, int* window_bits);
/* n_jobs multi-scalar multiplications against the key's own generators (MulVec::calculate, mulvec.rs:20-33, with the points
 * fixed to [g, h, G_0..G_{len-1}, H_0..H_{len-1}]): scalars = n_jobs x (2 + 2 len) canonical scalars in that order,
 * out_points = n_jobs encodings. Vector Pedersen commitments; fixed-base paths only (window tables, or the direct tables of
 * bpp_pk_precompute). */
int bpp_pk_msm(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* scalars, size_t n_jobs, uint8_t* out_points);
int bpp_pk_export(bpp_ctx* ctx, const bpp_pk* pk, uint8_t* g, uint8_t* h, uint8_t* G, uint8_t* H);
size_t bpp_pk_len(const bpp_pk* pk);
void bpp_pk_free(bpp_pk* pk);

/* ---- RangeProver::commit / PublicKey::commitment (range/prover.rs:28-42, publickey.rs:50-52) ----- */
/* V_i = Fr(v_i as i32)*g + gamma_i*h  (the i32 cast of prover.rs:37 is mirrored) */
int bpp_commit_batch(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out_points);
/* V_i = Fr(v_i)*g + gamma_i*h with all 64 bits of v_i: the commitment an n = 64 range proof is about. The reference's cast
 * (prover.rs:37) wraps v >= 2^31, for which its own prover — which uses every bit of v, range/mod.rs:260-277 — can then never
 * produce a verifying proof; the Fiat-Shamir helpers of the host mirrors use this entry. */
int bpp_commit_batch_u64(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out_points);

/* ---- MulVec::calculate (bls12_381/building_block/mulvec.rs:20-33) -------------------------------- */
int bpp_msm(bpp_ctx* ctx, int curve, const uint8_t* scalars, const uint8_t* points, size_t k, uint8_t* out_point);
/* decode once, reuse: points stay in HBM as an affine table */
int bpp_points_upload(bpp_ctx* ctx, int curve, const uint8_t* points, size_t k, bpp_points** out);
/* the table {mult[i] * g}: synthetic MSM inputs generated on device */
int bpp_points_from_multipliers(bpp_ctx* ctx, int curve, const uint8_t* mult_scalars, size_t k, bpp_points** out);
size_t bpp_points_len(const bpp_points* pts);
void bpp_points_free(bpp_points* pts);
/* Optional, for a point set that many MSMs will reuse (what the builder pattern of mulvec.rs:35-53 cannot express: the
 * reference clones its points into every MulVec): window tables 2^(c*w) * P_i, as the public key carries for its generators.
 * Costs about 254 doublings and ceil(254/c) affine conversions per point once and ceil(254/c) x the table memory (refused
 * above 48 GB); later bpp_msm_prepared* calls against `pts` run without doublings, per-window bucket sets or Horner pass.
 * Results are identical with and without it. */
int bpp_points_precompute(bpp_ctx* ctx, bpp_points* pts);
/* The same step further, as bpp_pk_precompute does for a key's generators: DIRECT tables — every signed-digit multiple of every
 * window point of every point of the set, (ceil(252/c) + 1) * 2^(c-1) * 128 B per point for the widest window c in [8, 16] that
 * fits max_bytes (returned in *window_bits; 38 GB for 2^10 points at c = 15, 62 GB for 2^16 points at c = 9). Builds the window
 * tables of bpp_points_precompute first if they are missing. Later bpp_msm_prepared* calls against `pts` then cost one table
 * addition per window and term, with no bucket sort or reduction, whatever the number of jobs. No counterpart in the reference
 * (mulvec.rs:35-53 clones its points per MulVec). BPP_RISTRETTO255 only (BPP_ERR_UNSUPPORTED otherwise); BPP_ERR_LENGTH if not
 * even c = 8 fits; results are identical with and without it. */
int bpp_points_precompute_direct(bpp_ctx* ctx, bpp_points* pts, size_t max_bytes, int* window_bits);
/* n_jobs independent MSMs over consecutive slices of `scalars`, all against pts[0..k): out = n_jobs points */
int bpp_msm_prepared(bpp_ctx* ctx, const bpp_points* pts, const uint8_t* scalars, size_t k, size_t n_jobs, uint8_t* out_points);
/* same with device pointers (no host copies, no synchronisation): d_scalars = n_jobs*k*32 B, d_out = n_jobs encodings */
int bpp_msm_prepared_dev(bpp_ctx* ctx, const bpp_points* pts, const void* d_scalars, size_t k, size_t n_jobs, void* d_out_points);
/* partial sum over pts[first, first+count) in uncompressed device form (for the multi-GPU point-range split):
 * out_raw = bpp_raw_point_bytes(curve) bytes; combine the partials with bpp_points_sum_raw */
int bpp_raw_point_bytes(int curve);
int bpp_msm_prepared_raw(bpp_ctx* ctx, const bpp_points* pts, const uint8_t* scalars, size_t k, uint8_t* out_raw);
int bpp_points_sum_raw(bpp_ctx* ctx, int curve, const uint8_t* raw_points, size_t count, uint8_t* out_point);
/* device-pointer forms of the two above (asynchronous on the context's stream): the NCCL all-gather of the per-rank
 * partial sums sits between them (point addition is not an NCCL reduction operator) */
int bpp_msm_prepared_raw_dev(bpp_ctx* ctx, const bpp_points* pts, const void* d_scalars, size_t k, void* d_out_raw);
int bpp_points_sum_raw_dev(bpp_ctx* ctx, int curve, const void* d_raw_points, size_t count, void* d_out_point);
/* bpp_points_sum_raw_dev for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc). No counterpart in the reference. Same inputs, same output
 * bytes; calls no cudaMalloc / cudaFree / allocating routine. Its intermediate block sums live in a file-scope __device__
 * array of 65 slots x 192 B (12 480 B, held by every GPU that loads the library until the process exits), so
 * count <= BPP_NOALLOC_MAX_POINTS, beyond which it returns BPP_ERR_LENGTH with a message in bpp_last_error. All calls on
 * a device share that array: calls that overlap on the GPU (two streams / contexts) race; calls on one stream do not. */
#define BPP_NOALLOC_MAX_POINTS 65536
int bpp_points_sum_raw_noalloc_dev(bpp_ctx* ctx, int curve, const void* d_raw_points, size_t count, void* d_out_point);

/* ---- RangeProof::prove / verify (range/mod.rs:31-78) ---------------------------------------------- */
/* `count` independent proofs, each aggregating m values of n bits against pk (bpp_pk_len(pk) == n*m).
 * v: count*m values; gamma: count*m scalars; proofs: count * bpp_range_proof_len bytes. */
int bpp_range_prove_batch(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint64_t* v,
                          const uint8_t* gamma, const bpp_params* params, uint8_t* proofs);
/* V: count*m commitments. verdicts[i] = BPP_OK | BPP_ERR_VERIFICATION | BPP_ERR_FORMAT. Returns BPP_OK if the
 * batch ran (whatever the verdicts). */
int bpp_range_verify_batch(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V,
                           const uint8_t* proofs, const bpp_params* params, uint8_t* verdicts);
/* device-pointer variants: inputs already in HBM, outputs left in HBM, asynchronous on the context's stream */
int bpp_range_prove_batch_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_v,
                              const void* d_gamma, const bpp_params* params, void* d_proofs);
int bpp_range_verify_batch_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V,
                               const void* d_proofs, const bpp_params* params, void* d_verdicts);
/* single-proof forms; verify returns the verdict */
int bpp_range_prove(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, const uint64_t* v, const uint8_t* gamma,
                    const bpp_params* params, uint8_t* proof, size_t* proof_len);
int bpp_range_verify(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, const uint8_t* V, const uint8_t* proof,
                     size_t proof_len, const bpp_params* params);

/* ---- real Fiat-Shamir: the host owns the transcript (task statement; the snapshot hard-codes every challenge) ---- */
/* Batch verification with PER-PROOF challenges: challenges = count blocks of (3 + L) scalars [y, z, e, e_round[0..L)],
 * L = log2(n*m), which the host derives from each proof's own transcript (A -> y, z; L_k, R_k -> e_k; wip A, B -> e). A zero
 * or non-canonical challenge gives that proof the verdict BPP_ERR_FORMAT. Same checks, same MSM as bpp_range_verify_batch;
 * with every block set to the reference's constants the verdicts are identical to it. */
int bpp_range_verify_batch_fs(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V,
                              const uint8_t* proofs, const uint8_t* challenges, uint8_t* verdicts);
int bpp_range_verify_batch_fs_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V,
                                  const void* d_proofs, const void* d_challenges, void* d_verdicts);
/* Random-linear-combination batch verification: sum_p w_p * (equation of proof p) = 0 is checked with
 * ONE MSM of 2 + 2*n*m + count*(2L + 3 + m) terms — the generator terms of all proofs collapse — instead of `count` MSMs.
 * weights = count canonical, non-zero scalars the prover could not predict (the host's CSPRNG; BPP_ERR_FORMAT otherwise).
 * challenges = NULL (one set: `params`, NULL = the reference's constants) or per-proof blocks as in bpp_range_verify_batch_fs.
 * The verdicts are always the per-proof ones: proofs that do not decode get BPP_ERR_FORMAT and leave the combination; if the
 * combination fails, the batch is verified again proof by proof (*used_fallback = 1) to name the bad ones. A wrong accept
 * needs a bad proof whose error the random weights cancel: probability about 1 / group order per batch.
 * Both forms synchronise the context's stream (the fallback decision is taken on the host). */
int bpp_range_verify_batch_rlc(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                               const bpp_params* params, const uint8_t* challenges, const uint8_t* weights, uint8_t* verdicts,
                               int* used_fallback);
int bpp_range_verify_batch_rlc_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V, const void* d_proofs,
                                   const bpp_params* params, const void* d_challenges, const void* d_weights, void* d_verdicts,
                                   int* used_fallback);
/* Interactive proving of a batch, cut where the challenges enter (range/mod.rs:31-55 and wip.rs:36-226 with the constants
 * replaced by the host's answers). All scalars 32-byte canonical; every array is per proof, in batch order.
 *   begin      nonces = NULL (the reference's constants) or count x (5 + 2L) scalars [alpha, r, s, delta, eta, d_L[L], d_R[L]];
 *              writes A (count points): the host absorbs it and answers (y, z)
 *   step 0     challenges = count x [y, z]        -> out_points = count x [L_0, R_0]
 *   step k     challenges = count x e_{k-1}       -> out_points = count x [L_k, R_k]      (k = 1 .. L-1)
 *   step L     challenges = count x e_{L-1}       -> out_points = count x [wip A, wip B]
 *   finish     e_final    = count x e             -> proofs = count complete proofs (wire layout above)
 * Calls must come in this order (BPP_ERR_LENGTH otherwise); a zero / non-canonical challenge or nonce is BPP_ERR_FORMAT.
 * The session holds device memory until bpp_prove_session_free (which may park it in the context for the next batch of the
 * same shape); `pk` and `ctx` must outlive it. */
int bpp_range_prove_begin(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint64_t* v, const uint8_t* gamma,
                          const uint8_t* nonces, bpp_prove_session** out, uint8_t* out_A);
int bpp_range_prove_step(bpp_ctx* ctx, bpp_prove_session* session, const uint8_t* challenges, uint8_t* out_points);
int bpp_range_prove_finish(bpp_ctx* ctx, bpp_prove_session* session, const uint8_t* e_final, uint8_t* proofs);
int bpp_prove_session_rounds(const bpp_prove_session* session); /* L */
void bpp_prove_session_free(bpp_prove_session* session);

/* ---- WeightedInnerProductProof::prove / verify (weighted_inner_product_proof.rs:36-43, :238-248) -- */
/* a, b, power_of_y: len scalars each (len = bpp_pk_len(pk)); gamma: one scalar. The reference's `commitment`
 * argument only feeds a dead update (wip.rs:137-142) and is not part of this ABI. */
int bpp_wip_prove(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* a, const uint8_t* b, const uint8_t* power_of_y,
                  const uint8_t* gamma, const bpp_params* params, uint8_t* proof, size_t* proof_len);
int bpp_wip_verify(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* power_of_y, const uint8_t* G_exp_of_commitment,
                   const uint8_t* H_exp_of_commitment, const uint8_t* g_exp_of_commitment, size_t n_V,
                   const uint8_t* V_exp_of_commitment, const uint8_t* A_prime, const uint8_t* V, const uint8_t* proof,
                   size_t proof_len, const bpp_params* params);

/* ---- measurement support --------------------------------------------------------------------------- */
/* Integer-pipe microbenchmark, register-only chains on every SM with loop-produced multiplicands. rates3[0] = 32-bit
 * multiply-adds/s (IMAD: the issue peak of the integer multiply pipe), rates3[1] = carry-chained 32x32+64 multiply-adds/s
 * (IMAD.WIDE.U32.X, the form a multi-limb product is made of), rates3[2] = 32x32+64 multiply-adds/s with a plain 64-bit
 * register addend (IMAD.WIDE.U32). */
int bpp_int_peak_probe(bpp_ctx* ctx, double* rates3, double* elapsed_ms);
/* Field-multiply probe: `ilp` independent chains of GF(2^255-19) multiplications per thread on blocks_per_sm x threads_per_block
 * threads per SM, no memory traffic: field multiplications/s (x 72 = multiply-adds/s the carry-chain multiply reaches alone). */
int bpp_fe_mul_probe(bpp_ctx* ctx, int ilp, int threads_per_block, int blocks_per_sm, double* muls_per_sec);
/* FP64-pipe microbenchmark (the Ristretto bucket accumulators multiply there, csrc/fe25519_f64.cuh). rates2[0] = DFMA
 * lane-operations/s (fma.rz.f64 chains with loop-produced operands: the issue peak of the FP64 pipe), rates2[1] =
 * GF(2^255-19) multiplications/s of the 5 x 51-bit double-limb multiply alone (50 DFMA + 30 DADD each, no memory traffic). */
int bpp_fp64_peak_probe(bpp_ctx* ctx, double* rates2);
/* device time (ms, CUDA events on the context's stream) spent in the bucket-accumulation kernel, and its launch
 * count, since the last reset */
int bpp_hot_kernel_stats(bpp_ctx* ctx, double* ms, uint64_t* launches, int reset);

/* launch tracer (development / measurement aid): bpp_trace(ctx, 1) starts recording one CUDA-event pair per kernel launch
 * on whichever of the context's two streams it goes to, bpp_trace(ctx, 0) stops. bpp_trace_dump synchronises and writes
 * one line per recorded launch, "<stream 0|1> <start_us> <end_us> <kernel>", times relative to the first record, into buf
 * (NUL-terminated, truncated to cap); returns the number of records and clears them. */
int bpp_trace(bpp_ctx* ctx, int on);
int bpp_trace_dump(bpp_ctx* ctx, char* buf, size_t cap);

/* test surface: element-wise device arithmetic (field ops 0-4, group ops 10-14, scalar-field ops 20-23) */
int bpp_debug_op(bpp_ctx* ctx, int curve, int op, size_t n, const uint8_t* a, const uint8_t* b, uint8_t* out, uint8_t* ok);

#ifdef __cplusplus
}
#endif
#endif /* BPP_H */

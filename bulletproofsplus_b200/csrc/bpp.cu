// libbpp_b200.so — extern "C" surface (include/bpp.h). Dispatches on the curve id to the templated
// launch sequences in api_impl.cuh. No CPU fallback exists: every entry point needs a live bpp_ctx,
// and bpp_init fails when there is no CUDA device.
#include "fe25519.cuh"
#include "fe25519_f64.cuh"
#include "engine.cuh"
#include "proto_misc.cuh"

using namespace bpp;

namespace bpp {
struct CurveVtbl;
const CurveVtbl* vtbl_ristretto255();
const CurveVtbl* vtbl_secp256k1();
const CurveVtbl* vtbl_bls12_381_g1();
}  // namespace bpp
#include "curve_vtbl.cuh"

namespace {

const CurveVtbl* vtbl(int curve) {
  switch (curve) {
    case BPP_RISTRETTO255: return vtbl_ristretto255();
    case BPP_SECP256K1: return vtbl_secp256k1();
    case BPP_BLS12_381_G1: return vtbl_bls12_381_g1();
    default: return nullptr;
  }
}

int point_bytes(int curve) { return curve == 0 ? 32 : curve == 1 ? 64 : curve == 2 ? 96 : -1; }
bool pow2(size_t x) { return x && !(x & (x - 1)); }

int set_device(bpp_ctx* ctx) {
  if (!ctx) return BPP_ERR_NOT_INIT;
  BPP_CK(cudaSetDevice(ctx->device));
  return BPP_OK;
}

// drain the event pairs recorded around the hot kernel
int drain_hot(bpp_ctx* ctx) {
  for (auto& pr : ctx->hot_pending) {
    BPP_CK(cudaEventSynchronize(pr.second));
    float ms = 0;
    BPP_CK(cudaEventElapsedTime(&ms, pr.first, pr.second));
    ctx->hot_ms += ms;
    cudaEventDestroy(pr.first);
    cudaEventDestroy(pr.second);
  }
  ctx->hot_pending.clear();
  return BPP_OK;
}

}  // namespace

#define BPP_ENTER(ctx)            \
  do {                            \
    int _rc = set_device(ctx);    \
    if (_rc != BPP_OK) return _rc; \
  } while (0)

// `call` is a member call on the curve's table
#define BPP_DISPATCH(curve, call)                                                          \
  do {                                                                                     \
    const CurveVtbl* vt_ = vtbl(curve);                                                    \
    if (!vt_) { ctx->err = "unknown curve id"; return BPP_ERR_UNSUPPORTED; }                 \
    return vt_->call;                                                                      \
  } while (0)

// Field-multiply throughput probe: ILP independent dependent-chains of fe_mul per thread, no memory traffic. Tells how much
// of the IMAD.WIDE peak the 72-IMAD carry-chain multiply can reach by itself at a given occupancy.
template <int ILP>
__global__ void fe_mul_probe_kernel(uint32_t iters, uint32_t seed, uint32_t* __restrict__ sink) {
  Fe25519 x[ILP], y;
  for (int k = 0; k < ILP; k++)
    for (int i = 0; i < 8; i++) x[k].v[i] = seed * (k + 3) + threadIdx.x * 2654435761u + i * 40503u;
  for (int i = 0; i < 8; i++) y.v[i] = (seed + threadIdx.x) ^ (0x9e3779b9u * (i + 1));
  for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < ILP; k++) fe_mul(x[k], x[k], y);
  }
  uint32_t acc = 0;
  for (int k = 0; k < ILP; k++)
    for (int i = 0; i < 8; i++) acc ^= x[k].v[i];
  if (acc == 0x12345u) sink[0] = acc;
}

extern "C" {

int bpp_init(int device, void* cuda_stream, bpp_ctx** out) {
  if (!out) return BPP_ERR_LENGTH;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return BPP_ERR_CUDA;  // no device: fail, never fall back
  if (cudaSetDevice(device) != cudaSuccess) return BPP_ERR_CUDA;
  bpp_ctx* ctx = new bpp_ctx();
  ctx->device = device;
  if (cuda_stream) {
    ctx->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
  } else {
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return BPP_ERR_CUDA; }
    ctx->own_stream = true;
  }
  *out = ctx;
  return BPP_OK;
}

void bpp_destroy(bpp_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  drain_hot(ctx);
  for (bpp_prove_session* ps : ctx->session_pool) { ps->release(); delete ps; }
  ctx->session_pool.clear();
  for (auto& b : ctx->buf) b.release();
  if (ctx->aux_stream) {
    cudaStreamSynchronize(ctx->aux_stream);
    cudaStreamDestroy(ctx->aux_stream);
    cudaEventDestroy(ctx->aux_fork);
    cudaEventDestroy(ctx->aux_join);
  }
  for (auto& b : ctx->buf2) b.release();
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* bpp_last_error(const bpp_ctx* ctx) { return ctx ? ctx->err.c_str() : "no context"; }
int bpp_point_bytes(int curve) { return point_bytes(curve); }
int bpp_raw_point_bytes(int curve) { const CurveVtbl* V = vtbl(curve); return V ? V->raw_bytes : -1; }
size_t bpp_range_proof_len(int curve, size_t n, size_t m) {
  int pb = point_bytes(curve);
  if (pb < 0 || !pow2(n * m) || n * m < 2) return 0;
  return (2 * (size_t)ilog2(n * m) + 3) * pb + 96;
}
size_t bpp_wip_proof_len(int curve, size_t len) {
  int pb = point_bytes(curve);
  if (pb < 0 || !pow2(len) || len < 2) return 0;
  return (2 * (size_t)ilog2(len) + 2) * pb + 96;
}
uint64_t bpp_kernel_launches(const bpp_ctx* ctx) { return ctx ? ctx->launches : 0; }
int bpp_sync(bpp_ctx* ctx) {
  BPP_ENTER(ctx);
  BPP_CK(cudaStreamSynchronize(ctx->stream));
  return BPP_OK;
}

int bpp_pk_new(bpp_ctx* ctx, int curve, size_t len, bpp_pk** out) {
  BPP_ENTER(ctx);
  if (!out) return BPP_ERR_LENGTH;
  BPP_DISPATCH(curve, pk_new(ctx, len, out));
}
int bpp_pk_new_hashed(bpp_ctx* ctx, int curve, size_t len, const uint8_t* domain_sep, size_t domain_sep_len, bpp_pk** out) {
  BPP_ENTER(ctx);
  if (!out) return BPP_ERR_LENGTH;
  BPP_DISPATCH(curve, pk_new_hashed(ctx, len, domain_sep, domain_sep_len, out));
}
int bpp_pk_precompute(bpp_ctx* ctx, bpp_pk* pk, size_t max_bytes, int* window_bits) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, pk_precompute(ctx, pk, max_bytes, window_bits));
}
int bpp_pk_msm(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* scalars, size_t n_jobs, uint8_t* out_points) {
  BPP_ENTER(ctx);
  if (!pk || !scalars || !out_points) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, pk_msm(ctx, pk, scalars, n_jobs, out_points));
}
int bpp_pk_export(bpp_ctx* ctx, const bpp_pk* pk, uint8_t* g, uint8_t* h, uint8_t* G, uint8_t* H) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, pk_export(ctx, pk, g, h, G, H));
}
size_t bpp_pk_len(const bpp_pk* pk) { return pk ? pk->len : 0; }
void bpp_pk_free(bpp_pk* pk) {
  if (!pk) return;
  cudaSetDevice(pk->device);
  if (pk->table) cudaFree(pk->table);
  if (pk->wtab) cudaFree(pk->wtab);
  if (pk->wtab_f) cudaFree(pk->wtab_f);
  if (pk->comb) cudaFree(pk->comb);
  delete pk;
}

int bpp_commit_batch(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out_points) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, commit_batch(ctx, pk, v, gamma, count, out_points));
}

int bpp_commit_batch_u64(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out_points) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, commit_batch_u64(ctx, pk, v, gamma, count, out_points));
}

int bpp_msm(bpp_ctx* ctx, int curve, const uint8_t* scalars, const uint8_t* points, size_t k, uint8_t* out_point) {
  BPP_ENTER(ctx);
  BPP_DISPATCH(curve, msm(ctx, scalars, points, k, out_point));
}
int bpp_points_upload(bpp_ctx* ctx, int curve, const uint8_t* points, size_t k, bpp_points** out) {
  BPP_ENTER(ctx);
  BPP_DISPATCH(curve, points_upload(ctx, points, k, false, out));
}
int bpp_points_from_multipliers(bpp_ctx* ctx, int curve, const uint8_t* mult, size_t k, bpp_points** out) {
  BPP_ENTER(ctx);
  BPP_DISPATCH(curve, points_upload(ctx, mult, k, true, out));
}
size_t bpp_points_len(const bpp_points* pts) { return pts ? pts->len : 0; }
void bpp_points_free(bpp_points* pts) {
  if (!pts) return;
  cudaFree(pts->table);
  if (pts->wtab) cudaFree(pts->wtab);
  if (pts->table_f) cudaFree(pts->table_f);
  if (pts->wtab_f) cudaFree(pts->wtab_f);
  if (pts->dtab) cudaFree(pts->dtab);
  delete pts;
}
int bpp_points_precompute(bpp_ctx* ctx, bpp_points* pts) {
  BPP_ENTER(ctx);
  if (!pts) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pts->curve, points_precompute(ctx, pts));
}
int bpp_points_precompute_direct(bpp_ctx* ctx, bpp_points* pts, size_t max_bytes, int* window_bits) {
  BPP_ENTER(ctx);
  if (!pts) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pts->curve, points_precompute_direct(ctx, pts, max_bytes, window_bits));
}
int bpp_msm_prepared(bpp_ctx* ctx, const bpp_points* pts, const uint8_t* scalars, size_t k, size_t n_jobs, uint8_t* out_points) {
  BPP_ENTER(ctx);
  if (!pts) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pts->curve, msm_prepared(ctx, pts, scalars, k, n_jobs, out_points, false));
}
int bpp_msm_prepared_dev(bpp_ctx* ctx, const bpp_points* pts, const void* d_scalars, size_t k, size_t n_jobs, void* d_out_points) {
  BPP_ENTER(ctx);
  if (!pts) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pts->curve, msm_prepared_dev(ctx, pts, d_scalars, k, n_jobs, d_out_points));
}
int bpp_msm_prepared_raw(bpp_ctx* ctx, const bpp_points* pts, const uint8_t* scalars, size_t k, uint8_t* out_raw) {
  BPP_ENTER(ctx);
  if (!pts) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pts->curve, msm_prepared(ctx, pts, scalars, k, 1, out_raw, true));
}
int bpp_points_sum_raw(bpp_ctx* ctx, int curve, const uint8_t* raw_points, size_t count, uint8_t* out_point) {
  BPP_ENTER(ctx);
  BPP_DISPATCH(curve, points_sum_raw(ctx, raw_points, count, out_point));
}

int bpp_msm_prepared_raw_dev(bpp_ctx* ctx, const bpp_points* pts, const void* d_scalars, size_t k, void* d_out_raw) {
  BPP_ENTER(ctx);
  if (!pts) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pts->curve, msm_prepared_raw_dev(ctx, pts, d_scalars, k, d_out_raw));
}
int bpp_points_sum_raw_dev(bpp_ctx* ctx, int curve, const void* d_raw_points, size_t count, void* d_out_point) {
  BPP_ENTER(ctx);
  BPP_DISPATCH(curve, points_sum_raw_dev(ctx, d_raw_points, count, d_out_point));
}

int bpp_range_prove_batch(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint64_t* v, const uint8_t* gamma,
                          const bpp_params* params, uint8_t* proofs) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_prove_batch(ctx, pk, n, m, count, v, gamma, params, proofs));
}
int bpp_range_verify_batch(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                           const bpp_params* params, uint8_t* verdicts) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_verify_batch(ctx, pk, n, m, count, V, proofs, params, verdicts));
}
int bpp_range_prove_batch_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_v, const void* d_gamma,
                              const bpp_params* params, void* d_proofs) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_prove_batch_dev(ctx, pk, n, m, count, d_v, d_gamma, params, d_proofs));
}
int bpp_range_verify_batch_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V, const void* d_proofs,
                               const bpp_params* params, void* d_verdicts) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_verify_batch_dev(ctx, pk, n, m, count, d_V, d_proofs, params, d_verdicts));
}
int bpp_range_verify_batch_fs(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                              const uint8_t* challenges, uint8_t* verdicts) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_verify_batch_fs(ctx, pk, n, m, count, V, proofs, challenges, verdicts));
}
int bpp_range_verify_batch_fs_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V, const void* d_proofs,
                                  const void* d_challenges, void* d_verdicts) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_verify_batch_fs_dev(ctx, pk, n, m, count, d_V, d_proofs, d_challenges, d_verdicts));
}
int bpp_range_verify_batch_rlc(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                               const bpp_params* params, const uint8_t* challenges, const uint8_t* weights, uint8_t* verdicts, int* used_fallback) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_verify_batch_rlc(ctx, pk, n, m, count, V, proofs, params, challenges, weights, verdicts, used_fallback));
}
int bpp_range_verify_batch_rlc_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V, const void* d_proofs,
                                   const bpp_params* params, const void* d_challenges, const void* d_weights, void* d_verdicts,
                                   int* used_fallback) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, range_verify_batch_rlc_dev(ctx, pk, n, m, count, d_V, d_proofs, params, d_challenges, d_weights, d_verdicts, used_fallback));
}
int bpp_range_prove_begin(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint64_t* v, const uint8_t* gamma,
                          const uint8_t* nonces, bpp_prove_session** out, uint8_t* out_A) {
  BPP_ENTER(ctx);
  if (!pk || !out) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, prove_begin(ctx, pk, n, m, count, v, gamma, nonces, out, out_A));
}
int bpp_range_prove_step(bpp_ctx* ctx, bpp_prove_session* session, const uint8_t* challenges, uint8_t* out_points) {
  BPP_ENTER(ctx);
  if (!session) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(session->curve, prove_step(ctx, session, challenges, out_points));
}
int bpp_range_prove_finish(bpp_ctx* ctx, bpp_prove_session* session, const uint8_t* e_final, uint8_t* proofs) {
  BPP_ENTER(ctx);
  if (!session) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(session->curve, prove_finish(ctx, session, e_final, proofs));
}
int bpp_prove_session_rounds(const bpp_prove_session* session) { return session ? (int)session->L : -1; }
void bpp_prove_session_free(bpp_prove_session* session) {
  if (!session) return;
  if (session->owner && session->owner->session_pool.size() < 2) {
    session->owner->session_pool.push_back(session);  // keeps its device memory for the next batch of the same shape
    return;
  }
  session->release();
  delete session;
}
int bpp_range_prove(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, const uint64_t* v, const uint8_t* gamma, const bpp_params* params,
                    uint8_t* proof, size_t* proof_len) {
  if (!pk || !proof_len) return BPP_ERR_NOT_INIT;
  size_t need = bpp_range_proof_len(pk->curve, n, m);
  if (need == 0) return pow2(n * m) ? BPP_ERR_LENGTH : BPP_ERR_NOT_POW2;
  if (*proof_len < need) { *proof_len = need; return BPP_ERR_LENGTH; }
  *proof_len = need;
  return bpp_range_prove_batch(ctx, pk, n, m, 1, v, gamma, params, proof);
}
int bpp_range_verify(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, const uint8_t* V, const uint8_t* proof, size_t proof_len,
                     const bpp_params* params) {
  if (!pk) return BPP_ERR_NOT_INIT;
  size_t need = bpp_range_proof_len(pk->curve, n, m);
  if (need == 0) return pow2(n * m) ? BPP_ERR_LENGTH : BPP_ERR_NOT_POW2;
  if (proof_len != need) return BPP_ERR_FORMAT;
  uint8_t verdict = BPP_ERR_CUDA;
  int rc = bpp_range_verify_batch(ctx, pk, n, m, 1, V, proof, params, &verdict);
  return rc != BPP_OK ? rc : verdict;
}

int bpp_wip_prove(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* a, const uint8_t* b, const uint8_t* power_of_y, const uint8_t* gamma,
                  const bpp_params* params, uint8_t* proof, size_t* proof_len) {
  BPP_ENTER(ctx);
  if (!pk || !proof_len) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, wip_prove(ctx, pk, a, b, power_of_y, gamma, params, proof, proof_len));
}
int bpp_wip_verify(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* power_of_y, const uint8_t* G_exp_of_commitment,
                   const uint8_t* H_exp_of_commitment, const uint8_t* g_exp_of_commitment, size_t n_V, const uint8_t* V_exp_of_commitment,
                   const uint8_t* A_prime, const uint8_t* V, const uint8_t* proof, size_t proof_len, const bpp_params* params) {
  BPP_ENTER(ctx);
  if (!pk) return BPP_ERR_NOT_INIT;
  BPP_DISPATCH(pk->curve, wip_verify(ctx, pk, power_of_y, G_exp_of_commitment, H_exp_of_commitment, g_exp_of_commitment, n_V,
                                           V_exp_of_commitment, A_prime, V, proof, proof_len, params));
}

int bpp_int_peak_probe(bpp_ctx* ctx, double* rates3, double* elapsed_ms) {
  BPP_ENTER(ctx);
  cudaDeviceProp prop;
  BPP_CK(cudaGetDeviceProperties(&prop, ctx->device));
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  const uint32_t iters = 4096;  // 32 multiply-adds per iteration per thread
  BPP_CK(ctx->buf[B_OUT0].ensure(64));
  cudaEvent_t e0, e1;
  BPP_CK(cudaEventCreate(&e0));
  BPP_CK(cudaEventCreate(&e1));
  double total_ms = 0;
  for (int mode = 0; mode < 3; mode++) {
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {  // first repetition is the warm-up
      uint64_t* sink = ctx->buf[B_OUT0].as<uint64_t>();
      BPP_CK(cudaEventRecord(e0, ctx->stream));
      if (mode == 0) int_peak_kernel<0><<<blocks, threads, 0, ctx->stream>>>(iters, 12345u + rep, sink);
      else if (mode == 1) int_peak_kernel<1><<<blocks, threads, 0, ctx->stream>>>(iters, 12345u + rep, sink);
      else int_peak_kernel<2><<<blocks, threads, 0, ctx->stream>>>(iters, 12345u + rep, sink);
      ctx->launches++;
      BPP_CK(cudaEventRecord(e1, ctx->stream));
      BPP_CK(cudaEventSynchronize(e1));
      float ms = 0;
      BPP_CK(cudaEventElapsedTime(&ms, e0, e1));
      total_ms += ms;
      if (rep > 0 && ms < best) best = ms;
    }
    if (rates3) rates3[mode] = (double)blocks * threads * iters * 32.0 / (best * 1e-3);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (elapsed_ms) *elapsed_ms = total_ms;
  return BPP_OK;
}

int bpp_fp64_peak_probe(bpp_ctx* ctx, double* rates2) {
  BPP_ENTER(ctx);
  cudaDeviceProp prop;
  BPP_CK(cudaGetDeviceProperties(&prop, ctx->device));
  const int blocks = prop.multiProcessorCount * 4, threads = 256;
  BPP_CK(ctx->buf[B_OUT0].ensure(64));
  cudaEvent_t e0, e1;
  BPP_CK(cudaEventCreate(&e0));
  BPP_CK(cudaEventCreate(&e1));
  for (int mode = 0; mode < 2; mode++) {
    const uint32_t iters = mode == 0 ? 8192 : 2048;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {  // first repetition is the warm-up
      double* sink = ctx->buf[B_OUT0].as<double>();
      BPP_CK(cudaEventRecord(e0, ctx->stream));
      if (mode == 0) fp64_peak_kernel<0><<<blocks, threads, 0, ctx->stream>>>(iters, 1.0000001 + rep, sink);
      else fp64_peak_kernel<1><<<blocks, threads, 0, ctx->stream>>>(iters, 1.0, sink);
      ctx->launches++;
      BPP_CK(cudaGetLastError());
      BPP_CK(cudaEventRecord(e1, ctx->stream));
      BPP_CK(cudaEventSynchronize(e1));
      float ms = 0;
      BPP_CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep > 0 && ms < best) best = ms;
    }
    if (rates2) rates2[mode] = (double)blocks * threads * iters * (mode == 0 ? 32.0 : 1.0) / (best * 1e-3);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return BPP_OK;
}

int bpp_fe_mul_probe(bpp_ctx* ctx, int ilp, int threads_per_block, int blocks_per_sm, double* muls_per_sec) {
  BPP_ENTER(ctx);
  cudaDeviceProp prop;
  BPP_CK(cudaGetDeviceProperties(&prop, ctx->device));
  const int blocks = prop.multiProcessorCount * blocks_per_sm;
  const uint32_t iters = 2048;
  BPP_CK(ctx->buf[B_OUT0].ensure(64));
  cudaEvent_t e0, e1;
  BPP_CK(cudaEventCreate(&e0));
  BPP_CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    BPP_CK(cudaEventRecord(e0, ctx->stream));
    uint32_t* sink = ctx->buf[B_OUT0].as<uint32_t>();
    switch (ilp) {
      case 1: fe_mul_probe_kernel<1><<<blocks, threads_per_block, 0, ctx->stream>>>(iters, 77u + rep, sink); break;
      case 2: fe_mul_probe_kernel<2><<<blocks, threads_per_block, 0, ctx->stream>>>(iters, 77u + rep, sink); break;
      case 3: fe_mul_probe_kernel<3><<<blocks, threads_per_block, 0, ctx->stream>>>(iters, 77u + rep, sink); break;
      default: ilp = 4; fe_mul_probe_kernel<4><<<blocks, threads_per_block, 0, ctx->stream>>>(iters, 77u + rep, sink); break;
    }
    ctx->launches++;
    BPP_CK(cudaGetLastError());
    BPP_CK(cudaEventRecord(e1, ctx->stream));
    BPP_CK(cudaEventSynchronize(e1));
    float ms = 0;
    BPP_CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (muls_per_sec) *muls_per_sec = (double)blocks * threads_per_block * iters * ilp / (best * 1e-3);
  return BPP_OK;
}

// test surface: element-wise device arithmetic (see debug_ops.cuh). elem = field bytes (ops 0-4), point bytes (10-14), 32 (20-23).
int bpp_debug_op(bpp_ctx* ctx, int curve, int op, size_t n, const uint8_t* a, const uint8_t* b, uint8_t* out, uint8_t* ok) {
  BPP_ENTER(ctx);
  BPP_DISPATCH(curve, debug_op(ctx, op, n, a, b, out, ok));
}

int bpp_hot_kernel_stats(bpp_ctx* ctx, double* ms, uint64_t* launches, int reset) {
  BPP_ENTER(ctx);
  BPP_TRY(drain_hot(ctx));
  if (ms) *ms = ctx->hot_ms;
  if (launches) *launches = ctx->hot_launches;
  if (reset) { ctx->hot_ms = 0; ctx->hot_launches = 0; }
  ctx->hot_timing = true;  // timing is armed by the first call
  return BPP_OK;
}

int bpp_trace(bpp_ctx* ctx, int on) {
  BPP_ENTER(ctx);
  ctx->trace_on = on != 0;
  return BPP_OK;
}
int bpp_trace_dump(bpp_ctx* ctx, char* buf, size_t cap) {
  if (!ctx) return -1;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->aux_stream) cudaStreamSynchronize(ctx->aux_stream);
  std::string out;
  const int n = (int)ctx->trace.size();
  for (auto& r : ctx->trace) {
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, ctx->trace[0].e0, r.e0);
    cudaEventElapsedTime(&b, ctx->trace[0].e0, r.e1);
    char line[256];
    snprintf(line, sizeof line, "%d %.1f %.1f %s\n", r.aux, a * 1e3, b * 1e3, r.name);
    out += line;
  }
  for (auto& r : ctx->trace) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  ctx->trace.clear();
  if (buf && cap) {
    size_t k = out.size() < cap - 1 ? out.size() : cap - 1;
    memcpy(buf, out.data(), k);
    buf[k] = 0;
  }
  return n;
}

}  // extern "C"

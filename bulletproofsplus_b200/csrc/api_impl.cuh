// Launch sequences for each C-ABI entry point, templated on (curve policy C, scalar modulus M).
#pragma once
#include "engine.cuh"
#include "proto_misc.cuh"
#include "pk_hashed.cuh"
#include "sha512.hpp"

namespace bpp {

struct HostOrder { uint8_t le[32]; };
inline bool host_scalar_ok(const uint8_t* s, const uint8_t* order_le) {
  for (int i = 31; i >= 0; i--) {
    if (s[i] != order_le[i]) return s[i] < order_le[i];
  }
  return false;
}

static __global__ void prove_jobstart_kernel(uint32_t n_proofs, uint32_t L, uint32_t N, uint32_t has_A, uint32_t terms_per_proof,
                                      uint32_t* __restrict__ job_start) {
  const uint32_t J = has_A + 2 * L + 2;
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > n_proofs * J) return;
  if (t == n_proofs * J) { job_start[t] = n_proofs * terms_per_proof; return; }
  uint32_t p = t / J, k = t % J, off = 0;
  if (has_A) { if (k == 0) { job_start[t] = p * terms_per_proof; return; } off = 1; k -= 1; }
  if (k < 2 * L) off += k * (N + 2);
  else if (k == 2 * L) off += 2 * L * (N + 2);
  else off += 2 * L * (N + 2) + 2 * N + 2;
  job_start[t] = p * terms_per_proof + off;
}

template <class C, class M>
struct Api {
  static constexpr size_t PB = C::POINT_BYTES;

  // ---- parameters -----------------------------------------------------------------------------------
  // packs (and validates) the parameter block into B_PARAMS; NULL = the reference's constants
  static int load_params(bpp_ctx* ctx, const bpp_params* prm, bool single, uint32_t L, const uint8_t* order_le) {
    std::vector<uint8_t> h((size_t)(P_FIXED + 3 * L) * 32, 0);
    auto put_small = [&](size_t slot, uint32_t v) { memset(&h[slot * 32], 0, 32); memcpy(&h[slot * 32], &v, 4); };
    if (!prm) {
      put_small(P_Y, single ? 7 : 12); put_small(P_Z, single ? 7 : 23); put_small(P_EFINAL, 99);
      put_small(P_ALPHA, single ? 7 : 33); put_small(P_R, 33); put_small(P_S, 44); put_small(P_DELTA, 88); put_small(P_ETA, 123);
      for (uint32_t k = 0; k < L; k++) { put_small(P_FIXED + k, 7); put_small(P_FIXED + L + k, 4); put_small(P_FIXED + 2 * L + k, 5); }
    } else {
      if (prm->n_rounds < L || (L && (!prm->e_round || !prm->d_L || !prm->d_R))) { ctx->err = "params: too few rounds"; return BPP_ERR_LENGTH; }
      const uint8_t* fixed[8] = {prm->y, prm->z, prm->e_final, prm->alpha, prm->r, prm->s, prm->delta, prm->eta};
      for (int i = 0; i < 8; i++) memcpy(&h[(size_t)i * 32], fixed[i], 32);
      memcpy(&h[(size_t)P_FIXED * 32], prm->e_round, (size_t)L * 32);
      memcpy(&h[(size_t)(P_FIXED + L) * 32], prm->d_L, (size_t)L * 32);
      memcpy(&h[(size_t)(P_FIXED + 2 * L) * 32], prm->d_R, (size_t)L * 32);
      for (size_t i = 0; i < P_FIXED + 3 * (size_t)L; i++)
        if (!host_scalar_ok(&h[i * 32], order_le)) { ctx->err = "params: non-canonical scalar"; return BPP_ERR_FORMAT; }
      auto is_zero = [&](size_t slot) { for (int b = 0; b < 32; b++) if (h[slot * 32 + b]) return false; return true; };
      bool zero_challenge = is_zero(P_Y) || is_zero(P_EFINAL);
      for (uint32_t k = 0; k < L; k++) zero_challenge = zero_challenge || is_zero(P_FIXED + k);
      if (zero_challenge) { ctx->err = "params: a challenge is zero (not invertible)"; return BPP_ERR_FORMAT; }
    }
    if (ctx->params_gen && h == ctx->params_host) return BPP_OK;  // same block as the previous call: already on the device
    BPP_CK(ctx->buf[B_PARAMS].ensure(h.size()));
    ctx->params_gen++;
    ctx->tab_gen = 0;
    ctx->params_host.swap(h);  // staging buffer with context lifetime: the copy below needs no synchronisation
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_PARAMS].p, ctx->params_host.data(), ctx->params_host.size(), cudaMemcpyHostToDevice, ctx->stream));
    return BPP_OK;
  }
  // true when B_TAB* already hold the tables for (kind, current parameter block, shape, pk); otherwise records that they will
  static bool tables_cached(bpp_ctx* ctx, uint32_t kind, const bpp_pk* pk, uint32_t N, uint32_t n, uint32_t m, uint32_t mode) {
    const bool hit = mode != 0 && ctx->tab_gen == ctx->params_gen && ctx->tab_kind == kind && ctx->tab_pk == pk && ctx->tab_shape[0] == N &&
                     ctx->tab_shape[1] == n && ctx->tab_shape[2] == m && ctx->tab_shape[3] == mode && ctx->tab_shape[4] == (uint32_t)C::CURVE_ID;
    ctx->tab_gen = mode != 0 ? ctx->params_gen : 0;  // mode 0 (standalone WIP) tables depend on caller vectors: never reused
    ctx->tab_kind = kind; ctx->tab_pk = pk;
    ctx->tab_shape[0] = N; ctx->tab_shape[1] = n; ctx->tab_shape[2] = m; ctx->tab_shape[3] = mode; ctx->tab_shape[4] = (uint32_t)C::CURVE_ID;
    return hit;
  }
  static int read_flag(bpp_ctx* ctx, const uint8_t* d_flag, uint8_t* out) {
    BPP_CK(cudaMemcpyAsync(out, d_flag, 1, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }
  static int clear_flags(bpp_ctx* ctx) {
    BPP_CK(ctx->buf[B_FLAGS].ensure(16));
    BPP_CK(cudaMemsetAsync(ctx->buf[B_FLAGS].p, 0, 16, ctx->stream));
    return BPP_OK;
  }

  // ---- PublicKey::new ---------------------------------------------------------------------------------
  static void pk_release(bpp_pk* pk) {
    if (pk->table) cudaFree(pk->table);
    if (pk->wtab) cudaFree(pk->wtab);
    if (pk->wtab_f) cudaFree(pk->wtab_f);
    if (pk->comb) cudaFree(pk->comb);
    if (pk->dtab) cudaFree(pk->dtab);
    delete pk;
  }
  // Direct fixed-base tables (pk_direct.cuh): the widest window c in [8, 16] whose table fits in max_bytes. One-time cost: a
  // few tens of milliseconds and (W + 1) * 2^(c-1) * 128 B per generator (40 GB for a 1024-generator key at c = 14).
  // the tables of `np` affine points at `table` (the public key's generators, or a caller's point set)
  static int build_direct(bpp_ctx* ctx, const uint32_t* table, uint32_t np, size_t max_bytes, const char* what, uint32_t** dtab_out,
                          uint32_t* c_out, uint32_t* W_out) {
    if constexpr (!C::HAS_COOP4) {
      ctx->err = std::string(what) + ": direct tables are implemented for Ristretto255 only";
      return BPP_ERR_UNSUPPORTED;
    } else {
      uint32_t c = 0, W = 0;
      for (uint32_t cc = 16; cc >= 8; cc--) {
        const uint32_t WW = dtab_windows(C::SCALAR_BITS, cc);
        if (dtab_entries(np, cc, WW) * C::AFF_WORDS * 4 <= max_bytes) { c = cc; W = WW; break; }
      }
      if (!c) { ctx->err = std::string(what) + ": max_bytes is too small for any window width >= 8"; return BPP_ERR_LENGTH; }
      const size_t entries = dtab_entries(np, c, W);
      if ((uint64_t)np * W >= (1ull << 31)) { ctx->err = std::string(what) + ": too many points"; return BPP_ERR_LENGTH; }
      uint32_t *dtab = nullptr, *base = nullptr;
      cudaError_t e = cudaMalloc(&dtab, entries * C::AFF_WORDS * 4);
      if (e == cudaSuccess) e = cudaMalloc(&base, (size_t)np * W * C::AFF_WORDS * 4);
      if (e == cudaSuccess) e = ctx->buf[B_EXT].ensure((size_t)np * W * C::EXT_WORDS * 4);
      if (e == cudaSuccess) {
        wtab_build_kernel<C><<<ceil_div(np, 32), 32, 0, ctx->stream>>>(np, 0u, np, c, W, table, base, ctx->buf[B_EXT].as<uint32_t>());
        dtab_fill_kernel<C><<<ceil_div((uint64_t)np * W, 64), 64, 0, ctx->stream>>>(np, c, W, base, dtab);
        dtab_normalise_kernel<C><<<(unsigned)((entries / DTAB_NORM + 128) / 128), 128, 0, ctx->stream>>>(entries, dtab);
        ctx->launches += 3;
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      }
      if (base) cudaFree(base);
      if (e != cudaSuccess) {
        if (dtab) cudaFree(dtab);
        cudaGetLastError();  // an allocation failure is not sticky, but it would surface at the next launch check of this context
        ctx->err = cudaGetErrorString(e);
        return BPP_ERR_CUDA;
      }
      *dtab_out = dtab; *c_out = c; *W_out = W;
      return BPP_OK;
    }
  }
  static int pk_precompute(bpp_ctx* ctx, bpp_pk* pk, size_t max_bytes, int* c_out) {
    if (pk->dtab) { if (c_out) *c_out = (int)pk->cD; return BPP_OK; }
    uint32_t* dtab = nullptr;
    uint32_t c = 0, W = 0;
    BPP_TRY(build_direct(ctx, pk->table, pk->n_points(), max_bytes, "pk_precompute", &dtab, &c, &W));
    pk->dtab = dtab; pk->cD = c; pk->WD = W;
    if (c_out) *c_out = (int)c;
    return BPP_OK;
  }
  // the same for a caller's point set (bpp_points_precompute_direct): window tables first (the engine's plan needs them), then
  // every signed-digit multiple. Later bpp_msm_prepared* calls against `pts` take them whatever their number of jobs.
  static int points_precompute_direct(bpp_ctx* ctx, bpp_points* pts, size_t max_bytes, int* c_out) {
    if (pts->dtab) { if (c_out) *c_out = (int)pts->cD; return BPP_OK; }
    if constexpr (!C::HAS_COOP4) {
      ctx->err = "points_precompute_direct: direct tables are implemented for Ristretto255 only";
      return BPP_ERR_UNSUPPORTED;
    } else {
      BPP_TRY(points_precompute(ctx, pts));
      uint32_t* dtab = nullptr;
      uint32_t c = 0, W = 0;
      BPP_TRY(build_direct(ctx, pts->table, (uint32_t)pts->len, max_bytes, "points_precompute_direct", &dtab, &c, &W));
      pts->dtab = dtab; pts->cD = c; pts->WD = W;
      if (c_out) *c_out = (int)c;
      return BPP_OK;
    }
  }
  static int pk_new(bpp_ctx* ctx, size_t len, bpp_pk** out) { return pk_new_impl(ctx, len, nullptr, 0, false, out); }
  // Generators with no known relation between them: each is the RFC 9496 element derivation of SHA-512 output under the
  // caller's domain separator (pk_hashed.cuh). Ristretto255 only. publickey.rs:21-48 stays the parity constructor (pk_new).
  static int pk_new_hashed(bpp_ctx* ctx, size_t len, const uint8_t* domain_sep, size_t domain_len, bpp_pk** out) {
    if (C::CURVE_ID != 0) { ctx->err = "pk_new_hashed: hash-to-group is implemented for Ristretto255 only"; return BPP_ERR_UNSUPPORTED; }
    if (domain_len && !domain_sep) return BPP_ERR_LENGTH;
    return pk_new_impl(ctx, len, domain_sep, domain_len, true, out);
  }
  static int pk_new_impl(bpp_ctx* ctx, size_t len, const uint8_t* domain_sep, size_t domain_len, bool hashed, bpp_pk** out) {
    if (len > (1u << 22)) return BPP_ERR_LENGTH;
    bpp_pk* pk = new bpp_pk();
    pk->curve = C::CURVE_ID; pk->len = len; pk->device = ctx->device;
    const uint32_t np = pk->n_points();
    // window width for jobs of about 1.5 * len terms (L/R jobs have len + 2, the verifier and the final A 2 * len + 2)
    pk->cA = choose_window_fixed(len + len / 2 + 2, C::SCALAR_BITS);
    pk->WA = (C::SCALAR_BITS + 1 + pk->cA - 1) / pk->cA;
    pk->comb_w = (C::SCALAR_BITS + 1 + COMB_C - 1) / COMB_C;
    cudaError_t e = cudaMalloc(&pk->table, (size_t)np * C::AFF_WORDS * 4);
    if (e == cudaSuccess) e = cudaMalloc(&pk->wtab, (size_t)np * pk->WA * C::AFF_WORDS * 4);
    if (e == cudaSuccess) e = cudaMalloc(&pk->comb, (size_t)2 * pk->comb_w * COMB_D * C::AFF_WORDS * 4);
    if (e != cudaSuccess) { pk_release(pk); ctx->err = cudaGetErrorString(e); return BPP_ERR_CUDA; }
    if constexpr (C::CURVE_ID == 0) {
      if (hashed) {  // table order g, h, G[len], H[len]: labels 'g', 'h', 'G' + i, 'H' + i
        std::vector<uint8_t> seeds((size_t)np * 64);
        hashed_generator_seed(domain_sep, domain_len, 'g', 0, &seeds[0]);
        hashed_generator_seed(domain_sep, domain_len, 'h', 0, &seeds[64]);
        for (size_t i = 0; i < len; i++) {
          hashed_generator_seed(domain_sep, domain_len, 'G', i, &seeds[(2 + i) * 64]);
          hashed_generator_seed(domain_sep, domain_len, 'H', i, &seeds[(2 + len + i) * 64]);
        }
        cudaError_t e2 = ctx->buf[B_IN0].ensure(seeds.size());
        if (e2 == cudaSuccess) e2 = cudaMemcpyAsync(ctx->buf[B_IN0].p, seeds.data(), seeds.size(), cudaMemcpyHostToDevice, ctx->stream);
        if (e2 == cudaSuccess) e2 = cudaStreamSynchronize(ctx->stream);  // `seeds` is pageable host memory going out of scope
        if (e2 != cudaSuccess) { pk_release(pk); ctx->err = cudaGetErrorString(e2); return BPP_ERR_CUDA; }
        pk_gen_hashed_kernel<<<ceil_div(np, 64), 64, 0, ctx->stream>>>(np, ctx->buf[B_IN0].as<uint8_t>(), pk->table);
      } else {
        pk_gen_kernel<C><<<ceil_div(np, 64), 64, 0, ctx->stream>>>((uint32_t)len, pk->table);
      }
    } else {
      pk_gen_kernel<C><<<ceil_div(np, 64), 64, 0, ctx->stream>>>((uint32_t)len, pk->table);
    }
    {
      cudaError_t e3 = ctx->buf[B_EXT].ensure((size_t)np * pk->WA * C::EXT_WORDS * 4);
      if (e3 != cudaSuccess || pk->WA > WTAB_MAX_W) { pk_release(pk); ctx->err = e3 != cudaSuccess ? cudaGetErrorString(e3) : "window count"; return BPP_ERR_CUDA; }
    }
    wtab_build_kernel<C><<<ceil_div(np, 32), 32, 0, ctx->stream>>>(np, 0u, np, pk->cA, pk->WA, pk->table, pk->wtab, ctx->buf[B_EXT].as<uint32_t>());
    comb_build_kernel<C><<<ceil_div(2 * pk->comb_w * COMB_D, 64), 64, 0, ctx->stream>>>(pk->comb_w, pk->table, pk->comb);
    ctx->launches += 3;
    if (C::HAS_F64 && hybrid_pattern()) {  // the window tables once more, in the slot format of the FP64 accumulators (hybrid accumulation kernel)
      e = cudaMalloc(&pk->wtab_f, (size_t)np * pk->WA * C::AFF_WORDS * 4);
      if (e != cudaSuccess) { pk_release(pk); ctx->err = cudaGetErrorString(e); return BPP_ERR_CUDA; }
      msm_table_format_kernel<C><<<ceil_div(np * pk->WA, 128), 128, 0, ctx->stream>>>(np * pk->WA, pk->wtab, pk->wtab_f);
      ctx->launches++;
    }
    e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { pk_release(pk); ctx->err = cudaGetErrorString(e); return BPP_ERR_CUDA; }
    *out = pk;
    return BPP_OK;
  }
  // n_jobs MSMs against the key's own generators: out[j] = sum_i scalars[j][i] * table[i], table = g, h, G[len], H[len]
  // (vector Pedersen commitments; also the test surface of the fixed-base paths). Fixed-base segment only.
  static int pk_msm(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* scalars, size_t n_jobs, uint8_t* out) {
    const size_t np = pk->n_points();
    if (n_jobs == 0 || n_jobs * np > (1ull << 27)) return BPP_ERR_LENGTH;
    const uint32_t T = (uint32_t)(n_jobs * np), J = (uint32_t)n_jobs;
    BPP_TRY(clear_flags(ctx));
    BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)T * 32));
    BPP_CK(ctx->buf[B_PIDX].ensure((size_t)T * 4));
    BPP_CK(ctx->buf[B_OUT0].ensure((size_t)J * PB));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_SCALARS].p, scalars, (size_t)T * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_LAUNCH(scalars_check_kernel<M>, ceil_div(T, 256), 256, 0, T, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_FLAGS].as<uint8_t>());
    BPP_LAUNCH(iota_kernel, ceil_div(T, 256), 256, 0, T, (uint32_t)np, ctx->buf[B_PIDX].as<uint32_t>());
    BPP_TRY(Msm<C>::run(ctx, T, J, (uint32_t)np, nullptr, pk->wtab, (uint32_t)np, pk->cA, pk->WA, nullptr, 0, pk->wtab_f, nullptr, pk->dtab, pk->cD, pk->WD));
    BPP_LAUNCH(points_encode_kernel<C>, ceil_div(J, 64), 64, 0, J, ctx->buf[B_RESULTS].as<uint32_t>(), ctx->buf[B_OUT0].as<uint8_t>());
    BPP_CK(cudaMemcpyAsync(out, ctx->buf[B_OUT0].p, (size_t)J * PB, cudaMemcpyDeviceToHost, ctx->stream));
    uint8_t bad = 0;
    BPP_TRY(read_flag(ctx, ctx->buf[B_FLAGS].as<uint8_t>(), &bad));
    return bad ? BPP_ERR_FORMAT : BPP_OK;
  }
  // re-encode an affine table
  static int export_table(bpp_ctx* ctx, const uint32_t* table, uint32_t n, uint8_t* host_out) {
    BPP_CK(ctx->buf[B_OUT0].ensure((size_t)n * PB));
    BPP_LAUNCH(points_encode_aff_kernel<C>, ceil_div(n, 64), 64, 0, n, table, ctx->buf[B_OUT0].as<uint8_t>());
    BPP_CK(cudaMemcpyAsync(host_out, ctx->buf[B_OUT0].p, (size_t)n * PB, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }
  static int pk_export(bpp_ctx* ctx, const bpp_pk* pk, uint8_t* g, uint8_t* h, uint8_t* G, uint8_t* H) {
    std::vector<uint8_t> all((size_t)pk->n_points() * PB);
    BPP_TRY(export_table(ctx, pk->table, pk->n_points(), all.data()));
    memcpy(g, &all[0], PB); memcpy(h, &all[PB], PB);
    if (pk->len) { memcpy(G, &all[2 * PB], pk->len * PB); memcpy(H, &all[(2 + pk->len) * PB], pk->len * PB); }
    return BPP_OK;
  }

  // ---- RangeProver::commit ----------------------------------------------------------------------------
  static int commit_batch_u64(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out) {
    return commit_batch(ctx, pk, v, gamma, count, out, true);
  }
  static int commit_batch_i32(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out) {
    return commit_batch(ctx, pk, v, gamma, count, out, false);
  }
  static int commit_batch(bpp_ctx* ctx, const bpp_pk* pk, const uint64_t* v, const uint8_t* gamma, size_t count, uint8_t* out, bool full_u64) {
    if (count == 0) return BPP_OK;
    if (count > (1u << 26)) return BPP_ERR_LENGTH;
    const uint32_t n = (uint32_t)count;
    BPP_TRY(clear_flags(ctx));
    BPP_CK(ctx->buf[B_IN0].ensure((size_t)n * 8));
    BPP_CK(ctx->buf[B_IN1].ensure((size_t)n * 32));
    BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)n * 64));
    BPP_CK(ctx->buf[B_PIDX].ensure((size_t)n * 8));
    BPP_CK(ctx->buf[B_OUT0].ensure((size_t)n * PB));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN0].p, v, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, gamma, (size_t)n * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_LAUNCH(commit_terms_kernel<M>, ceil_div(n, 128), 128, 0, n, full_u64 ? 1u : 0u, ctx->buf[B_IN0].as<uint64_t>(), ctx->buf[B_IN1].as<uint32_t>(),
               ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_PIDX].as<uint32_t>(), ctx->buf[B_FLAGS].as<uint8_t>());
    BPP_CK(ctx->buf[B_RESULTS].ensure((size_t)n * C::EXT_WORDS * 4));
    BPP_LAUNCH(comb_eval_kernel<C>, ceil_div(n, 128), 128, 0, n, pk->comb_w, ctx->buf[B_SCALARS].as<uint32_t>(), pk->comb,
               ctx->buf[B_RESULTS].as<uint32_t>());
    BPP_LAUNCH(points_encode_kernel<C>, ceil_div(n, 64), 64, 0, n, ctx->buf[B_RESULTS].as<uint32_t>(), ctx->buf[B_OUT0].as<uint8_t>());
    BPP_CK(cudaMemcpyAsync(out, ctx->buf[B_OUT0].p, (size_t)n * PB, cudaMemcpyDeviceToHost, ctx->stream));
    uint8_t bad = 0;
    BPP_TRY(read_flag(ctx, ctx->buf[B_FLAGS].as<uint8_t>(), &bad));
    return bad ? BPP_ERR_FORMAT : BPP_OK;
  }

  // ---- points / MSM -----------------------------------------------------------------------------------
  static int points_upload(bpp_ctx* ctx, const uint8_t* points, size_t k, bool from_multipliers, bpp_points** out) {
    if (k == 0 || k > (1u << 26)) return BPP_ERR_LENGTH;
    const uint32_t n = (uint32_t)k;
    BPP_TRY(clear_flags(ctx));
    const size_t in_bytes = from_multipliers ? 32 : PB;
    BPP_CK(ctx->buf[B_IN0].ensure((size_t)n * in_bytes));
    BPP_CK(ctx->buf[B_DYN_OK].ensure(n));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN0].p, points, (size_t)n * in_bytes, cudaMemcpyHostToDevice, ctx->stream));
    bpp_points* pts = new bpp_points();
    pts->curve = C::CURVE_ID; pts->len = k;
    cudaError_t e = cudaMalloc(&pts->table, (size_t)n * C::AFF_WORDS * 4);
    if (e != cudaSuccess) { delete pts; ctx->err = cudaGetErrorString(e); return BPP_ERR_CUDA; }
    int rc = BPP_OK;
    uint8_t bad = 0;
    do {
      if (from_multipliers) {
        base_mul_kernel<C, M><<<ceil_div(n, 64), 64, 0, ctx->stream>>>(n, ctx->buf[B_IN0].as<uint32_t>(), pts->table, ctx->buf[B_FLAGS].as<uint8_t>());
      } else {
        points_decode_kernel<C><<<ceil_div(n, 64), 64, 0, ctx->stream>>>(n, ctx->buf[B_IN0].as<uint8_t>(), pts->table, ctx->buf[B_DYN_OK].as<uint8_t>());
        any_zero_kernel<<<ceil_div(n, 256), 256, 0, ctx->stream>>>(n, ctx->buf[B_DYN_OK].as<uint8_t>(), ctx->buf[B_FLAGS].as<uint8_t>());
        ctx->launches++;
      }
      ctx->launches++;
      rc = read_flag(ctx, ctx->buf[B_FLAGS].as<uint8_t>(), &bad);
    } while (0);
    if (rc == BPP_OK && bad) rc = BPP_ERR_FORMAT;
    if (rc != BPP_OK) { points_release(pts); return rc; }
    *out = pts;
    return BPP_OK;
  }
  // Window tables for a point set that will be multiplied many times (the same construction as the public key's): about
  // SCALAR_BITS doublings and WA affine conversions per point once, WA x the memory; MSMs against it then need no doublings,
  // no per-window bucket sets and no Horner pass.
  static int points_precompute(bpp_ctx* ctx, bpp_points* pts) {
    if (pts->wtab) return BPP_OK;
    const uint32_t n = (uint32_t)pts->len;
    const uint32_t c = choose_window_fixed(n, C::SCALAR_BITS), W = (C::SCALAR_BITS + 1 + c - 1) / c;
    const size_t bytes = (size_t)n * W * C::AFF_WORDS * 4;
    if (bytes > ((size_t)48 << 30)) { ctx->err = "points_precompute: window tables would exceed 48 GB"; return BPP_ERR_LENGTH; }
    uint32_t* wtab = nullptr;
    cudaError_t e = cudaMalloc(&wtab, bytes);
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return BPP_ERR_CUDA; }
    // in slices of 2^16 points: the scratch for the extended points of a slice stays at W * 8 MB (Ristretto255)
    const uint32_t SL = 1u << 16;
    e = W > WTAB_MAX_W ? cudaErrorInvalidValue : ctx->buf[B_EXT].ensure((size_t)(n < SL ? n : SL) * W * C::EXT_WORDS * 4);
    for (uint32_t i0 = 0; e == cudaSuccess && i0 < n; i0 += SL) {
      const uint32_t cnt = n - i0 < SL ? n - i0 : SL;
      wtab_build_kernel<C><<<ceil_div(cnt, 64), 64, 0, ctx->stream>>>(n, i0, cnt, c, W, pts->table, wtab, ctx->buf[B_EXT].as<uint32_t>());
      ctx->launches++;
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { cudaFree(wtab); ctx->err = cudaGetErrorString(e); return BPP_ERR_CUDA; }
    pts->wtab = wtab; pts->cA = c; pts->WA = W;
    if (C::HAS_F64 && hybrid_pattern() && bytes <= ((size_t)8 << 30)) {
      if (cudaMalloc(&pts->wtab_f, bytes) == cudaSuccess) {
        msm_table_format_kernel<C><<<ceil_div((uint64_t)n * W, 128), 128, 0, ctx->stream>>>(n * W, wtab, pts->wtab_f);
        ctx->launches++;
        if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { cudaFree(pts->wtab_f); pts->wtab_f = nullptr; }
      } else {
        pts->wtab_f = nullptr;
        cudaGetLastError();
      }
    }
    return BPP_OK;
  }
  static void points_release(bpp_points* pts) {
    if (pts->table) cudaFree(pts->table);
    if (pts->wtab) cudaFree(pts->wtab);
    if (pts->table_f) cudaFree(pts->table_f);
    if (pts->wtab_f) cudaFree(pts->wtab_f);
    if (pts->dtab) cudaFree(pts->dtab);
    delete pts;
  }
  // scalars already on device (n_jobs*k); result: n_jobs extended points in B_RESULTS
  static int msm_prepared_core(bpp_ctx* ctx, const bpp_points* pts, const uint32_t* d_scalars, size_t k, size_t n_jobs) {
    if (k == 0 || k > pts->len || n_jobs == 0 || k * n_jobs > (1ull << 27)) return BPP_ERR_LENGTH;
    const uint32_t T = (uint32_t)(k * n_jobs);
    BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)T * 32));
    BPP_CK(ctx->buf[B_PIDX].ensure((size_t)T * 4));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_SCALARS].p, d_scalars, (size_t)T * 32, cudaMemcpyDeviceToDevice, ctx->stream));
    BPP_LAUNCH(scalars_check_kernel<M>, ceil_div(T, 256), 256, 0, T, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_FLAGS].as<uint8_t>());
    BPP_LAUNCH(iota_kernel, ceil_div(T, 256), 256, 0, T, (uint32_t)k, ctx->buf[B_PIDX].as<uint32_t>());
    if (pts->wtab)  // point index i < len: fixed-base segment only
      return Msm<C>::run(ctx, T, (uint32_t)n_jobs, (uint32_t)k, nullptr, pts->wtab, (uint32_t)pts->len, pts->cA, pts->WA, nullptr, 0, pts->wtab_f, nullptr,
                         pts->dtab, pts->cD, pts->WD);
    if (C::HAS_F64 && hybrid_pattern()) {  // the same points in the FP64 accumulators' slot format, built once per point set
      bpp_points* mp = const_cast<bpp_points*>(pts);
      if (!mp->table_f) {
        BPP_CK(cudaMalloc(&mp->table_f, pts->len * C::AFF_WORDS * 4));
        BPP_LAUNCH(msm_table_format_kernel<C>, ceil_div(pts->len, 128), 128, 0, (uint32_t)pts->len, pts->table, mp->table_f);
      }
    }
    return Msm<C>::run(ctx, T, (uint32_t)n_jobs, (uint32_t)k, nullptr, nullptr, 0, 0, 0, pts->table, (uint32_t)k, nullptr, pts->table_f);
  }
  static int msm_prepared(bpp_ctx* ctx, const bpp_points* pts, const uint8_t* scalars, size_t k, size_t n_jobs, uint8_t* out, bool raw) {
    BPP_TRY(clear_flags(ctx));
    BPP_CK(ctx->buf[B_IN1].ensure(k * n_jobs * 32));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, scalars, k * n_jobs * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_TRY(msm_prepared_core(ctx, pts, ctx->buf[B_IN1].as<uint32_t>(), k, n_jobs));
    const uint32_t J = (uint32_t)n_jobs;
    if (raw) {
      BPP_CK(cudaMemcpyAsync(out, ctx->buf[B_RESULTS].p, (size_t)J * C::EXT_WORDS * 4, cudaMemcpyDeviceToHost, ctx->stream));
    } else {
      BPP_CK(ctx->buf[B_OUT0].ensure((size_t)J * PB));
      BPP_LAUNCH(points_encode_kernel<C>, ceil_div(J, 64), 64, 0, J, ctx->buf[B_RESULTS].as<uint32_t>(), ctx->buf[B_OUT0].as<uint8_t>());
      BPP_CK(cudaMemcpyAsync(out, ctx->buf[B_OUT0].p, (size_t)J * PB, cudaMemcpyDeviceToHost, ctx->stream));
    }
    uint8_t bad = 0;
    BPP_TRY(read_flag(ctx, ctx->buf[B_FLAGS].as<uint8_t>(), &bad));
    return bad ? BPP_ERR_FORMAT : BPP_OK;
  }
  static int msm_prepared_dev(bpp_ctx* ctx, const bpp_points* pts, const void* d_scalars, size_t k, size_t n_jobs, void* d_out) {
    BPP_TRY(clear_flags(ctx));
    BPP_TRY(msm_prepared_core(ctx, pts, reinterpret_cast<const uint32_t*>(d_scalars), k, n_jobs));
    const uint32_t J = (uint32_t)n_jobs;
    BPP_LAUNCH(points_encode_kernel<C>, ceil_div(J, 64), 64, 0, J, ctx->buf[B_RESULTS].as<uint32_t>(), reinterpret_cast<uint8_t*>(d_out));
    return BPP_OK;
  }
  static int msm_prepared_raw_dev(bpp_ctx* ctx, const bpp_points* pts, const void* d_scalars, size_t k, void* d_out_raw) {
    BPP_TRY(clear_flags(ctx));
    BPP_TRY(msm_prepared_core(ctx, pts, reinterpret_cast<const uint32_t*>(d_scalars), k, 1));
    BPP_CK(cudaMemcpyAsync(d_out_raw, ctx->buf[B_RESULTS].p, (size_t)C::EXT_WORDS * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    return BPP_OK;
  }
  static int points_sum_raw_dev(bpp_ctx* ctx, const void* d_raw, size_t count, void* d_out) {
    if (count == 0 || count > (1u << 20)) return BPP_ERR_LENGTH;
    BPP_CK(ctx->buf[B_RESULTS].ensure(C::EXT_WORDS * 4));
    BPP_LAUNCH(ext_sum_kernel<C>, 1, 128, 0, (uint32_t)count, reinterpret_cast<const uint32_t*>(d_raw), ctx->buf[B_RESULTS].as<uint32_t>());
    BPP_LAUNCH(points_encode_kernel<C>, 1, 32, 0, 1u, ctx->buf[B_RESULTS].as<uint32_t>(), reinterpret_cast<uint8_t*>(d_out));
    return BPP_OK;
  }
  static int msm(bpp_ctx* ctx, const uint8_t* scalars, const uint8_t* points, size_t k, uint8_t* out) {
    if (k == 0) {  // empty MulVec: Point::zero()
      bpp_points* one = nullptr;
      std::vector<uint8_t> z(32, 0), idb(PB);
      // encode the identity by running a 1-term MSM with scalar 0 against the base point table
      std::vector<uint8_t> mult(32, 0); mult[0] = 1;
      BPP_TRY(points_upload(ctx, mult.data(), 1, true, &one));
      int rc = msm_prepared(ctx, one, z.data(), 1, 1, out, false);
      points_release(one);
      return rc;
    }
    bpp_points* pts = nullptr;
    BPP_TRY(points_upload(ctx, points, k, false, &pts));
    int rc = msm_prepared(ctx, pts, scalars, k, 1, out, false);
    points_release(pts);
    return rc;
  }
  static int points_sum_raw(bpp_ctx* ctx, const uint8_t* raw, size_t count, uint8_t* out) {
    if (count == 0 || count > (1u << 20)) return BPP_ERR_LENGTH;
    BPP_CK(ctx->buf[B_EXT].ensure(count * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_RESULTS].ensure(C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_OUT0].ensure(PB));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_EXT].p, raw, count * C::EXT_WORDS * 4, cudaMemcpyHostToDevice, ctx->stream));
    BPP_LAUNCH(ext_sum_kernel<C>, 1, 128, 0, (uint32_t)count, ctx->buf[B_EXT].as<uint32_t>(), ctx->buf[B_RESULTS].as<uint32_t>());
    BPP_LAUNCH(points_encode_kernel<C>, 1, 32, 0, 1u, ctx->buf[B_RESULTS].as<uint32_t>(), ctx->buf[B_OUT0].as<uint8_t>());
    BPP_CK(cudaMemcpyAsync(out, ctx->buf[B_OUT0].p, PB, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }

  // ---- RangeProof::verify (batch) ---------------------------------------------------------------------
  // mode 1/2: range single / multiple. d_V: count*m encoded commitments; d_proofs: count proofs.
  // mode 0: standalone WIP verify (count = 1): d_proofs = A_prime | wip proof (same layout as a range proof),
  //         ext_* = the caller's commitment exponents (device, canonical).
  static int verify_core(bpp_ctx* ctx, const bpp_pk* pk, uint32_t mode, uint32_t n, uint32_t m, uint32_t count, const uint8_t* d_V,
                         const uint8_t* d_proofs, size_t proof_len, const uint32_t* ext_Gc, const uint32_t* ext_Hc,
                         const uint32_t* ext_gc, const uint32_t* ext_Vc, const uint32_t* ext_y, uint8_t* d_verdicts,
                         const uint8_t* d_challenges = nullptr, const uint8_t* d_weights = nullptr, uint8_t* rlc_pass = nullptr) {
    VerifyShape vs;
    vs.N = (uint32_t)pk->len; vs.L = ilog2(vs.N); vs.n = n; vs.m = m; vs.nV = m; vs.mode = mode; vs.n_proofs = count;
    vs.pk_G0 = 2; vs.pk_H0 = 2 + vs.N; vs.dyn0 = pk->n_points();
    const uint32_t N = vs.N, L = vs.L, ppp = 2 * L + 3, K = 5 + 2 * L + 2 * N + vs.nV;
    if ((uint64_t)count * K > (1ull << 27) || (uint64_t)count * ppp * PB >= (1ull << 31)) { ctx->err = "verify: batch too large"; return BPP_ERR_LENGTH; }
    const uint32_t n_dyn = count * (ppp + vs.nV);
    BPP_CK(ctx->buf[B_IN0].ensure((size_t)count * ppp * PB));
    BPP_CK(ctx->buf[B_DYN].ensure((size_t)n_dyn * C::AFF_WORDS * 4));
    BPP_CK(ctx->buf[B_DYN_OK].ensure(n_dyn));
    BPP_CK(ctx->buf[B_FLAGS2].ensure(count));
    BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)count * K * 32));
    BPP_CK(ctx->buf[B_PIDX].ensure((size_t)count * K * 4));
    // d_challenges: every proof has its own challenge set (count x (3 + L) scalars) and therefore its own tables
    const size_t per = d_challenges ? count : 1;
    BPP_CK(ctx->buf[B_TAB0].ensure(per * K * 32));
    BPP_CK(ctx->buf[B_TAB1].ensure(per * K * 32));
    BPP_CK(ctx->buf[B_TAB2].ensure(per * ((size_t)N + 64) * 32));
    BPP_CK(ctx->buf[B_TAB3].ensure(per * ((size_t)vs.m + 2) * 32));
    if (d_challenges) BPP_CK(ctx->buf[B_TAB5].ensure((size_t)count * (P_FIXED + 3 * L) * 32));
    BPP_CK(ctx->buf[B_TAB4].ensure((size_t)count * 128));
    // decode the per-proof points and the commitments into the dynamic table
    // Only the generic MSM segment reads the decoded points: when that segment will run on the auxiliary stream (per-proof
    // MSMs, >= 16 proofs) the decoding starts there too, beside the exponent kernels and the fixed-base sort on the main stream.
    const bool aux_decode = !d_weights && Msm<C>::overlaps(count);
    struct AuxGuard {
      bpp_ctx* c; bool on;
      ~AuxGuard() { if (on) Msm<C>::aux_leave(c); }
    } aux_guard{ctx, false};
    if (aux_decode) { BPP_TRY(Msm<C>::aux_enter(ctx)); aux_guard.on = true; }
    const uint32_t gb = count * ppp * (uint32_t)PB;
    BPP_LAUNCH(gather_proof_points_kernel, ceil_div(gb, 256), 256, 0, count, ppp * (uint32_t)PB, d_proofs, proof_len, ctx->buf[B_IN0].as<uint8_t>());
    BPP_LAUNCH(points_decode_kernel<C>, ceil_div(count * ppp, 64), 64, 0, count * ppp, ctx->buf[B_IN0].as<uint8_t>(),
               ctx->buf[B_DYN].as<uint32_t>(), ctx->buf[B_DYN_OK].as<uint8_t>());
    if (vs.nV)
      BPP_LAUNCH(points_decode_kernel<C>, ceil_div(count * vs.nV, 64), 64, 0, count * vs.nV, d_V,
                 ctx->buf[B_DYN].as<uint32_t>() + (size_t)count * ppp * C::AFF_WORDS, ctx->buf[B_DYN_OK].as<uint8_t>() + (size_t)count * ppp);
    if (aux_guard.on) { Msm<C>::aux_leave(ctx); aux_guard.on = false; }
    VerifyTables T;
    T.Tmul = ctx->buf[B_TAB0].as<uint32_t>(); T.Tadd = ctx->buf[B_TAB1].as<uint32_t>(); T.tmp = ctx->buf[B_TAB2].as<uint32_t>();
    T.zpow = ctx->buf[B_TAB3].as<uint32_t>();
    T.stride_T = T.stride_tmp = T.stride_zpow = T.stride_params = 0;
    T.pre = nullptr; T.stride_pre = 0;
    uint32_t* xs = ctx->buf[B_TAB4].as<uint32_t>();
    BPP_LAUNCH(verify_proof_scalars_kernel<M>, ceil_div(count, 128), 128, 0, count, d_proofs, proof_len, (size_t)ppp * PB, xs,
               ctx->buf[B_FLAGS2].as<uint8_t>());
    if (d_challenges) {
      T.stride_T = K * 8; T.stride_tmp = (N + 64) * 8; T.stride_zpow = (vs.m + 2) * 8; T.stride_params = (P_FIXED + 3 * L) * 8;
      ctx->tab_gen = 0;  // the shared tables of an earlier call are overwritten
      BPP_LAUNCH(verify_expand_challenges_kernel<M>, ceil_div(count, 128), 128, 0, count, L, d_challenges, ctx->buf[B_TAB5].as<uint32_t>(),
                 ctx->buf[B_FLAGS2].as<uint8_t>());
      // the L + 2 inversions of every proof, all proofs side by side, then one 128-thread block per proof (256 threads measured 1.50 ms against 1.05 ms per 2048 proofs: r02k trace) for its tables
      BPP_CK(ctx->buf[B_TAB6].ensure((size_t)count * (L + 2) * 32));
      T.pre = ctx->buf[B_TAB6].as<uint32_t>(); T.stride_pre = (L + 2) * 8;
      BPP_LAUNCH(verify_inverses_kernel<M>, ceil_div(count, 32), 32, 0, count, L, ctx->buf[B_TAB5].as<uint32_t>(), T.stride_params,
                 ctx->buf[B_TAB6].as<uint32_t>(), T.stride_pre);
      BPP_LAUNCH((verify_tables_kernel<M, 128>), count, 128, 0, vs, ctx->buf[B_TAB5].as<uint32_t>(), ext_Gc, ext_Hc, ext_gc, ext_Vc, ext_y, T);
    } else {
      const size_t cap0 = ctx->buf[B_TAB0].cap + ctx->buf[B_TAB1].cap + ctx->buf[B_TAB2].cap + ctx->buf[B_TAB3].cap;
      if (!tables_cached(ctx, 1, pk, N, n, m, mode) || cap0 != ctx->tab_caps)
        BPP_LAUNCH((verify_tables_kernel<M, VT_THREADS>), 1, VT_THREADS, 0, vs, ctx->buf[B_PARAMS].as<uint32_t>(), ext_Gc, ext_Hc, ext_gc, ext_Vc, ext_y, T);
      ctx->tab_caps = cap0;
    }
    dim3 grid(ceil_div(K, 256), count);
    // with the key's direct tables the generator terms' exponents are computed inside the accumulation kernel (DirectVerifySrc)
    // (measured slower — 7.08 vs 6.72 ms per 2048 proofs: the exponent arithmetic costs the accumulation kernel 6 registers and
    // 144 B of spills, more than the 0.25 ms it saves here; off unless BPP_FUSE is set, profiles/r02h_bench_{fused,nofuse}_*.json)
    static const bool want_fuse = getenv("BPP_FUSE") != nullptr;  // A/B switch
    const bool fuse = want_fuse && !d_weights && C::CURVE_ID == 0 && Msm<C>::uses_direct(pk->dtab, count);
    BPP_LAUNCH(verify_scalars_kernel<M>, grid, 256, 0, vs, T, xs, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_PIDX].as<uint32_t>(), fuse ? 1u : 0u);
    DirectVerifySrc vsrc;
    vsrc.enabled = 1; vsrc.K0 = 5 + 2 * L; vsrc.N = N; vsrc.stride_T = T.stride_T; vsrc.Tmul = T.Tmul; vsrc.Tadd = T.Tadd; vsrc.xs = xs;
    if (d_weights) {
      // random linear combination: one MSM of 2 + 2N + count * (ppp + nV) terms for the whole batch
      const uint32_t per = ppp + vs.nV, T2 = 2 + 2 * N + count * per;
      BPP_CK(ctx->buf[B_WORK0].ensure((size_t)T2 * 32));
      BPP_CK(ctx->buf[B_WORK1].ensure((size_t)T2 * 4));
      BPP_CK(ctx->buf[B_WORK2].ensure((size_t)count * 32));
      BPP_CK(ctx->buf[B_IN0].ensure((size_t)count * ppp * PB > count ? (size_t)count * ppp * PB : count));
      BPP_CK(ctx->buf[B_FLAGS].ensure(16));
      BPP_CK(cudaMemsetAsync(ctx->buf[B_FLAGS].p, 0, 16, ctx->stream));
      uint32_t* w_eff = ctx->buf[B_WORK2].as<uint32_t>();
      uint8_t* fmt = ctx->buf[B_IN0].as<uint8_t>();  // the gathered proof bytes are no longer needed
      uint8_t* flags = ctx->buf[B_FLAGS].as<uint8_t>();
      BPP_LAUNCH(rlc_weights_kernel<M>, ceil_div(count, 128), 128, 0, count, ppp, vs.nV, d_weights, ctx->buf[B_DYN_OK].as<uint8_t>(),
                 ctx->buf[B_FLAGS2].as<uint8_t>(), w_eff, fmt, flags + 1);
      BPP_CK(ctx->buf[B_EXT].ensure((size_t)RLC_GROUPS * (2 + 2 * N) * 32));
      BPP_LAUNCH(rlc_fold_fixed_kernel<M>, dim3(ceil_div(2 + 2 * N, 32), RLC_GROUPS), 256, 0, vs, w_eff, ctx->buf[B_SCALARS].as<uint32_t>(),
                 ctx->buf[B_EXT].as<uint32_t>());
      BPP_LAUNCH(rlc_fold_finish_kernel<M>, ceil_div(2 + 2 * N, 128), 128, 0, 2 + 2 * N, RLC_GROUPS, ctx->buf[B_EXT].as<uint32_t>(),
                 ctx->buf[B_WORK0].as<uint32_t>(), ctx->buf[B_WORK1].as<uint32_t>());
      BPP_LAUNCH(rlc_scale_dyn_kernel<M>, ceil_div(count * per, 256), 256, 0, vs, w_eff, ctx->buf[B_SCALARS].as<uint32_t>(),
                 ctx->buf[B_PIDX].as<uint32_t>(), ctx->buf[B_WORK0].as<uint32_t>(), ctx->buf[B_WORK1].as<uint32_t>());
      BPP_CK(cudaMemcpyAsync(ctx->buf[B_SCALARS].p, ctx->buf[B_WORK0].p, (size_t)T2 * 32, cudaMemcpyDeviceToDevice, ctx->stream));
      BPP_CK(cudaMemcpyAsync(ctx->buf[B_PIDX].p, ctx->buf[B_WORK1].p, (size_t)T2 * 4, cudaMemcpyDeviceToDevice, ctx->stream));
      BPP_TRY(Msm<C>::run(ctx, T2, 1, T2, nullptr, pk->wtab, pk->n_points(), pk->cA, pk->WA, ctx->buf[B_DYN].as<uint32_t>(), count * per, pk->wtab_f, nullptr, pk->dtab, pk->cD, pk->WD));
      BPP_LAUNCH(rlc_verdict_kernel<C>, ceil_div(count, 256) < 64 ? ceil_div(count, 256) : 64, 256, 0, count, ctx->buf[B_RESULTS].as<uint32_t>(), fmt,
                 d_verdicts, flags);
      uint8_t res[2] = {0, 0};
      BPP_CK(cudaMemcpyAsync(res, flags, 2, cudaMemcpyDeviceToHost, ctx->stream));
      BPP_CK(cudaStreamSynchronize(ctx->stream));
      if (res[1]) { ctx->err = "rlc verify: a weight is zero or not canonical"; return BPP_ERR_FORMAT; }
      *rlc_pass = res[0];
      return BPP_OK;
    }
    BPP_TRY(Msm<C>::run(ctx, count * K, count, K, nullptr, pk->wtab, pk->n_points(), pk->cA, pk->WA, ctx->buf[B_DYN].as<uint32_t>(), ppp + vs.nV, pk->wtab_f, nullptr, pk->dtab, pk->cD, pk->WD, fuse ? &vsrc : nullptr));
    BPP_LAUNCH(verdict_kernel<C>, ceil_div(count, 128), 128, 0, count, ppp, vs.nV, ctx->buf[B_RESULTS].as<uint32_t>(),
               ctx->buf[B_DYN_OK].as<uint8_t>(), ctx->buf[B_FLAGS2].as<uint8_t>(), d_verdicts);
    return BPP_OK;
  }
  static int check_range_shape(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count) {
    if (n == 0 || m == 0 || count == 0 || count > (1u << 24)) return BPP_ERR_LENGTH;
    const size_t N = n * m;
    if (N & (N - 1)) return BPP_ERR_NOT_POW2;  // wip.rs:67
    if (N < 2) return BPP_ERR_LENGTH;
    if (pk->len != N) { ctx->err = "generator count != n*m"; return BPP_ERR_GENERATORS; }  // range/mod.rs:90-91,252-253
    return BPP_OK;
  }
  static int range_verify_batch_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V,
                                    const void* d_proofs, const bpp_params* prm, void* d_verdicts, const uint8_t* order_le) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    const uint32_t L = ilog2(n * m);
    BPP_TRY(load_params(ctx, prm, m == 1, L, order_le));
    const size_t plen = (2 * (size_t)L + 3) * PB + 96;
    return verify_core(ctx, pk, m == 1 ? 1 : 2, (uint32_t)n, (uint32_t)m, (uint32_t)count, reinterpret_cast<const uint8_t*>(d_V),
                       reinterpret_cast<const uint8_t*>(d_proofs), plen, nullptr, nullptr, nullptr, nullptr, nullptr,
                       reinterpret_cast<uint8_t*>(d_verdicts));
  }
  static int range_verify_batch(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                                const bpp_params* prm, uint8_t* verdicts, const uint8_t* order_le) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    const size_t plen = (2 * (size_t)ilog2(n * m) + 3) * PB + 96;
    BPP_CK(ctx->buf[B_IN1].ensure(count * m * PB));
    BPP_CK(ctx->buf[B_IN2].ensure(count * plen));
    BPP_CK(ctx->buf[B_IN3].ensure(count));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, V, count * m * PB, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN2].p, proofs, count * plen, cudaMemcpyHostToDevice, ctx->stream));
    BPP_TRY(range_verify_batch_dev(ctx, pk, n, m, count, ctx->buf[B_IN1].p, ctx->buf[B_IN2].p, prm, ctx->buf[B_IN3].p, order_le));
    BPP_CK(cudaMemcpyAsync(verdicts, ctx->buf[B_IN3].p, count, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }

  // per-proof Fiat-Shamir challenges: challenges = count x (3 + L) scalars [y, z, e_final, e_round[0..L)]
  static int range_verify_batch_fs_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V, const void* d_proofs,
                                       const void* d_challenges, void* d_verdicts) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    if (!d_challenges) return BPP_ERR_LENGTH;
    const size_t plen = (2 * (size_t)ilog2(n * m) + 3) * PB + 96;
    return verify_core(ctx, pk, m == 1 ? 1 : 2, (uint32_t)n, (uint32_t)m, (uint32_t)count, reinterpret_cast<const uint8_t*>(d_V),
                       reinterpret_cast<const uint8_t*>(d_proofs), plen, nullptr, nullptr, nullptr, nullptr, nullptr,
                       reinterpret_cast<uint8_t*>(d_verdicts), reinterpret_cast<const uint8_t*>(d_challenges));
  }
  static int range_verify_batch_fs(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                                   const uint8_t* challenges, uint8_t* verdicts) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    if (!challenges) return BPP_ERR_LENGTH;
    const size_t L = ilog2(n * m), plen = (2 * L + 3) * PB + 96, clen = (3 + L) * 32;
    BPP_CK(ctx->buf[B_IN1].ensure(count * m * PB));
    BPP_CK(ctx->buf[B_IN2].ensure(count * plen));
    BPP_CK(ctx->buf[B_IN3].ensure(count));
    BPP_CK(ctx->buf[B_OUT0].ensure(count * clen));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, V, count * m * PB, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN2].p, proofs, count * plen, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_OUT0].p, challenges, count * clen, cudaMemcpyHostToDevice, ctx->stream));
    BPP_TRY(range_verify_batch_fs_dev(ctx, pk, n, m, count, ctx->buf[B_IN1].p, ctx->buf[B_IN2].p, ctx->buf[B_OUT0].p, ctx->buf[B_IN3].p));
    BPP_CK(cudaMemcpyAsync(verdicts, ctx->buf[B_IN3].p, count, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }

  // Random-linear-combination batch verification. weights: count canonical non-zero scalars the prover could not predict (the
  // host's CSPRNG). challenges: NULL (one challenge set from `prm`) or per-proof blocks as in range_verify_batch_fs. Verdicts are
  // the per-proof ones: when the combined check fails the batch is re-verified proof by proof. Synchronises the stream.
  static int range_verify_batch_rlc_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_V, const void* d_proofs,
                                        const bpp_params* prm, const void* d_challenges, const void* d_weights, void* d_verdicts,
                                        const uint8_t* order_le, int* used_fallback) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    if (!d_weights) return BPP_ERR_LENGTH;
    const uint32_t L = ilog2(n * m);
    if (!d_challenges) BPP_TRY(load_params(ctx, prm, m == 1, L, order_le));
    const size_t plen = (2 * (size_t)L + 3) * PB + 96;
    uint8_t pass = 0;
    BPP_TRY(verify_core(ctx, pk, m == 1 ? 1 : 2, (uint32_t)n, (uint32_t)m, (uint32_t)count, reinterpret_cast<const uint8_t*>(d_V),
                        reinterpret_cast<const uint8_t*>(d_proofs), plen, nullptr, nullptr, nullptr, nullptr, nullptr,
                        reinterpret_cast<uint8_t*>(d_verdicts), reinterpret_cast<const uint8_t*>(d_challenges),
                        reinterpret_cast<const uint8_t*>(d_weights), &pass));
    if (used_fallback) *used_fallback = pass ? 0 : 1;
    if (pass) return BPP_OK;
    return verify_core(ctx, pk, m == 1 ? 1 : 2, (uint32_t)n, (uint32_t)m, (uint32_t)count, reinterpret_cast<const uint8_t*>(d_V),
                       reinterpret_cast<const uint8_t*>(d_proofs), plen, nullptr, nullptr, nullptr, nullptr, nullptr,
                       reinterpret_cast<uint8_t*>(d_verdicts), reinterpret_cast<const uint8_t*>(d_challenges));
  }
  static int range_verify_batch_rlc(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint8_t* V, const uint8_t* proofs,
                                    const bpp_params* prm, const uint8_t* challenges, const uint8_t* weights, uint8_t* verdicts,
                                    const uint8_t* order_le, int* used_fallback) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    if (!weights) return BPP_ERR_LENGTH;
    const size_t L = ilog2(n * m), plen = (2 * L + 3) * PB + 96, clen = (3 + L) * 32;
    BPP_CK(ctx->buf[B_IN1].ensure(count * m * PB));
    BPP_CK(ctx->buf[B_IN2].ensure(count * plen));
    BPP_CK(ctx->buf[B_IN3].ensure(count));
    BPP_CK(ctx->buf[B_OUT0].ensure(count * (clen + 32)));
    uint8_t* d_w = ctx->buf[B_OUT0].as<uint8_t>();
    uint8_t* d_c = d_w + count * 32;
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, V, count * m * PB, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN2].p, proofs, count * plen, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(d_w, weights, count * 32, cudaMemcpyHostToDevice, ctx->stream));
    if (challenges) BPP_CK(cudaMemcpyAsync(d_c, challenges, count * clen, cudaMemcpyHostToDevice, ctx->stream));
    BPP_TRY(range_verify_batch_rlc_dev(ctx, pk, n, m, count, ctx->buf[B_IN1].p, ctx->buf[B_IN2].p, prm, challenges ? d_c : nullptr, d_w,
                                       ctx->buf[B_IN3].p, order_le, used_fallback));
    BPP_CK(cudaMemcpyAsync(verdicts, ctx->buf[B_IN3].p, count, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }

  // ---- RangeProof::prove (batch) / WeightedInnerProductProof::prove -------------------------------------
  // mode 1/2: d_v (count*m u64), d_gamma (count*m scalars). mode 0 (count = 1): d_a, d_b, d_ypow (N scalars), d_gamma (1).
  // d_proofs: count proofs of proof_len bytes; for mode 0 the WIP proof is written at d_proofs (no leading A).
  static int prove_core(bpp_ctx* ctx, const bpp_pk* pk, uint32_t mode, uint32_t n, uint32_t m, uint32_t count, const uint64_t* d_v,
                        const uint32_t* d_gamma, const uint32_t* d_a, const uint32_t* d_b, const uint32_t* d_ypow, uint8_t* d_proofs,
                        size_t proof_len, uint8_t* format_bad) {
    ProveShape ps;
    ps.N = (uint32_t)pk->len; ps.L = ilog2(ps.N); ps.n = n; ps.m = m; ps.mode = mode; ps.n_proofs = count;
    ps.pk_G0 = 2; ps.pk_H0 = 2 + ps.N; ps.dyn0 = pk->n_points();
    ps.has_A = mode != 0;
    const uint32_t N = ps.N, L = ps.L, J = ps.has_A + 2 * L + 2;
    ps.terms_per_proof = ps.has_A + 2 * L * (N + 2) + (2 * N + 2) + 2;
    if ((uint64_t)count * ps.terms_per_proof > (1ull << 27)) { ctx->err = "prove: batch too large"; return BPP_ERR_LENGTH; }
    const uint32_t T_all = count * ps.terms_per_proof;
    BPP_TRY(clear_flags(ctx));
    BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)T_all * 32));
    BPP_CK(ctx->buf[B_PIDX].ensure((size_t)T_all * 4));
    BPP_CK(ctx->buf[B_JOBSTART].ensure(((size_t)count * J + 1) * 4));
    BPP_CK(ctx->buf[B_WORK0].ensure((size_t)count * N * 32));
    BPP_CK(ctx->buf[B_WORK1].ensure((size_t)count * N * 32));
    BPP_CK(ctx->buf[B_WORK2].ensure((size_t)count * 32));
    BPP_CK(ctx->buf[B_EXT].ensure((size_t)count * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_TAB0].ensure((size_t)N * 32));
    BPP_CK(ctx->buf[B_TAB1].ensure((size_t)N * 32));
    BPP_CK(ctx->buf[B_TAB2].ensure((size_t)(L + 1) * N * 32));
    BPP_CK(ctx->buf[B_TAB3].ensure((size_t)(L + 1) * N * 32));
    BPP_CK(ctx->buf[B_TAB4].ensure((size_t)(4 * L + 4) * 32));
    BPP_CK(ctx->buf[B_TAB5].ensure((size_t)(PM_HDR + m + 4) * 32));
    ProveTables T;
    T.ypow = ctx->buf[B_TAB0].as<uint32_t>(); T.Hexp = ctx->buf[B_TAB1].as<uint32_t>(); T.gcoef = ctx->buf[B_TAB2].as<uint32_t>();
    T.hcoef = ctx->buf[B_TAB3].as<uint32_t>(); T.rnd = ctx->buf[B_TAB4].as<uint32_t>(); T.misc = ctx->buf[B_TAB5].as<uint32_t>();
    const uint32_t* params = ctx->buf[B_PARAMS].as<uint32_t>();
    uint32_t* a = ctx->buf[B_WORK0].as<uint32_t>();
    uint32_t* b = ctx->buf[B_WORK1].as<uint32_t>();
    uint32_t* alpha = ctx->buf[B_WORK2].as<uint32_t>();
    const size_t cap0 = ctx->buf[B_TAB0].cap + ctx->buf[B_TAB1].cap + ctx->buf[B_TAB2].cap + ctx->buf[B_TAB3].cap + ctx->buf[B_TAB4].cap +
                        ctx->buf[B_TAB5].cap;
    if (!tables_cached(ctx, 2, pk, N, n, m, mode) || cap0 != ctx->tab_caps)
      BPP_LAUNCH(prove_tables_kernel<M>, 1, 256, 0, ps, params, d_ypow, T);
    ctx->tab_caps = cap0;
    if (mode == 0) {
      BPP_LAUNCH(wip_load_kernel<M>, ceil_div(N, 128), 128, 0, N, d_a, d_b, d_gamma, a, b, alpha, ctx->buf[B_FLAGS].as<uint8_t>());
      BPP_LAUNCH(scalars_check_kernel<M>, ceil_div(N, 128), 128, 0, N, d_ypow, ctx->buf[B_FLAGS].as<uint8_t>());
    } else {
      BPP_LAUNCH(range_prove_init_kernel<M>, count, 256, 0, ps, T, params, d_v, d_gamma, a, b, alpha, ctx->buf[B_FLAGS].as<uint8_t>());
      BPP_LAUNCH(range_bits_sum_kernel<C>, count, 128, 0, ps, d_v, pk->table, ctx->buf[B_EXT].as<uint32_t>());
    }
    const size_t scalar_off = (size_t)J * PB;
    {
      // few proofs: a thread-block cluster per proof (proto_prove.cuh) - as many blocks as leave every block >= 128 vector
      // elements and the whole launch within one wave of two blocks per SM
      static const bool no_cluster = getenv("BPP_NO_CLUSTER") != nullptr;
      uint32_t cs = 1;
      while (!no_cluster && cs < 8 && N / (2 * cs) >= 128 && (uint64_t)count * 2 * cs <= 2 * 148) cs <<= 1;
      uint32_t* o_sc = ctx->buf[B_SCALARS].as<uint32_t>();
      uint32_t* o_pi = ctx->buf[B_PIDX].as<uint32_t>();
      const uint32_t* c_params = params;
      const uint32_t* c_alpha = alpha;
      BPP_LAUNCH_CLUSTER(wip_prove_kernel<M>, count * cs, 256, 256 * 32, cs, ps, T, c_params, a, b, c_alpha, o_sc, o_pi, d_proofs, (size_t)proof_len, (size_t)scalar_off);
    }
    BPP_LAUNCH(prove_jobstart_kernel, ceil_div(count * J + 1, 256), 256, 0, count, L, N, ps.has_A, ps.terms_per_proof,
               ctx->buf[B_JOBSTART].as<uint32_t>());
    BPP_TRY(Msm<C>::run(ctx, T_all, count * J, 0, ctx->buf[B_JOBSTART].as<uint32_t>(), pk->wtab, pk->n_points(), pk->cA, pk->WA, nullptr, 0, pk->wtab_f, nullptr, pk->dtab, pk->cD, pk->WD));
    // job 0 of a range proof is alpha * h; RangeProof.A = that + sum(bit ? G_i : -H_i) (range/mod.rs:98-106 / :260-277)
    BPP_LAUNCH(encode_into_proofs_kernel<C>, ceil_div(count * J, 64), 64, 0, count, J, ctx->buf[B_RESULTS].as<uint32_t>(),
               ps.has_A ? ctx->buf[B_EXT].as<uint32_t>() : nullptr, d_proofs, proof_len);
    // flag [0] set = some input scalar (a, b, y-powers, gamma) was not canonical
    return read_flag(ctx, ctx->buf[B_FLAGS].as<uint8_t>(), format_bad);
  }
  static int range_prove_batch_dev(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const void* d_v, const void* d_gamma,
                                   const bpp_params* prm, void* d_proofs, const uint8_t* order_le) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    const uint32_t L = ilog2(n * m);
    BPP_TRY(load_params(ctx, prm, m == 1, L, order_le));
    const size_t plen = (2 * (size_t)L + 3) * PB + 96;
    uint8_t bad = 0;
    BPP_TRY(prove_core(ctx, pk, m == 1 ? 1 : 2, (uint32_t)n, (uint32_t)m, (uint32_t)count, reinterpret_cast<const uint64_t*>(d_v),
                       reinterpret_cast<const uint32_t*>(d_gamma), nullptr, nullptr, nullptr, reinterpret_cast<uint8_t*>(d_proofs), plen, &bad));
    return bad ? BPP_ERR_FORMAT : BPP_OK;
  }
  static int range_prove_batch(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint64_t* v, const uint8_t* gamma,
                               const bpp_params* prm, uint8_t* proofs, const uint8_t* order_le) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    const size_t plen = (2 * (size_t)ilog2(n * m) + 3) * PB + 96;
    BPP_CK(ctx->buf[B_IN1].ensure(count * m * 8));
    BPP_CK(ctx->buf[B_IN2].ensure(count * m * 32));
    BPP_CK(ctx->buf[B_IN3].ensure(count * plen));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, v, count * m * 8, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN2].p, gamma, count * m * 32, cudaMemcpyHostToDevice, ctx->stream));
    int rc = range_prove_batch_dev(ctx, pk, n, m, count, ctx->buf[B_IN1].p, ctx->buf[B_IN2].p, prm, ctx->buf[B_IN3].p, order_le);
    if (rc != BPP_OK) return rc;
    BPP_CK(cudaMemcpyAsync(proofs, ctx->buf[B_IN3].p, count * plen, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    return BPP_OK;
  }

  // ---- interactive proving for a transcript-owning host (proto_session.cuh) ----------------------------------
  static FsState session_state(bpp_prove_session* s) {
    FsState st;
    st.params = s->params.as<uint32_t>(); st.a = s->a.as<uint32_t>(); st.b = s->b.as<uint32_t>(); st.alpha = s->alpha.as<uint32_t>();
    st.ypow = s->ypow.as<uint32_t>(); st.gcoef = s->gcoef.as<uint32_t>(); st.hcoef = s->hcoef.as<uint32_t>(); st.rnd = s->rnd.as<uint32_t>();
    st.PW = (P_FIXED + 3 * s->L) * 8;
    return st;
  }
  static ProveShape session_shape(const bpp_prove_session* s) {
    ProveShape ps;
    memset(&ps, 0, sizeof(ps));
    ps.N = s->N; ps.L = s->L; ps.n = s->n; ps.m = s->m; ps.mode = s->m == 1 ? 1 : 2; ps.n_proofs = s->count;
    ps.pk_G0 = 2; ps.pk_H0 = 2 + s->N; ps.dyn0 = s->pk->n_points(); ps.has_A = 1;
    return ps;
  }
  // runs the MSM over the terms in B_SCALARS / B_PIDX, encodes `jobs` points per proof into the proofs and into s->out, copies
  // them to the host and reports a pending FORMAT flag
  static int session_emit(bpp_ctx* ctx, bpp_prove_session* s, uint32_t n_terms, uint32_t jobs, uint32_t uniform, const uint32_t* d_job_start,
                          const uint32_t* add_to_job0, uint32_t slot0, uint32_t slot1, uint8_t* host_out) {
    const size_t plen = (2 * (size_t)s->L + 3) * PB + 96;
    BPP_TRY(Msm<C>::run(ctx, n_terms, s->count * jobs, uniform, d_job_start, s->pk->wtab, s->pk->n_points(), s->pk->cA, s->pk->WA, nullptr, 0, s->pk->wtab_f, nullptr, s->pk->dtab, s->pk->cD, s->pk->WD));
    BPP_CK(s->out.ensure((size_t)s->count * jobs * PB));
    BPP_LAUNCH(fs_encode_kernel<C>, ceil_div(s->count * jobs, 64), 64, 0, s->count, jobs, ctx->buf[B_RESULTS].as<uint32_t>(), add_to_job0, slot0, slot1,
               s->proofs.as<uint8_t>(), plen, s->out.as<uint8_t>());
    BPP_CK(cudaMemcpyAsync(host_out, s->out.p, (size_t)s->count * jobs * PB, cudaMemcpyDeviceToHost, ctx->stream));
    uint8_t bad = 0;
    BPP_TRY(read_flag(ctx, s->flag.as<uint8_t>(), &bad));
    if (bad) { ctx->err = "prove session: a scalar supplied by the caller is not canonical, or a challenge is zero"; return BPP_ERR_FORMAT; }
    return BPP_OK;
  }
  static int prove_begin(bpp_ctx* ctx, const bpp_pk* pk, size_t n, size_t m, size_t count, const uint64_t* v, const uint8_t* gamma,
                         const uint8_t* nonces, bpp_prove_session** out, uint8_t* out_A) {
    BPP_TRY(check_range_shape(ctx, pk, n, m, count));
    const uint32_t N = (uint32_t)(n * m), L = ilog2(N), cnt = (uint32_t)count;
    if ((uint64_t)count * (2 * N + 4) > (1ull << 27)) { ctx->err = "prove session: batch too large"; return BPP_ERR_LENGTH; }
    bpp_prove_session* s = nullptr;
    for (size_t i = 0; i < ctx->session_pool.size(); i++) {
      bpp_prove_session* c = ctx->session_pool[i];
      if (c->curve == C::CURVE_ID && c->n == n && c->m == m && c->count == cnt) {
        s = c;
        ctx->session_pool.erase(ctx->session_pool.begin() + i);
        break;
      }
    }
    if (!s) s = new bpp_prove_session();
    s->owner = ctx; s->step = 0;
    s->curve = C::CURVE_ID; s->pk = pk; s->n = (uint32_t)n; s->m = (uint32_t)m; s->count = cnt; s->N = N; s->L = L;
    const size_t plen = (2 * (size_t)L + 3) * PB + 96, vec = (size_t)cnt * N * 32;
    cudaError_t e = s->params.ensure((size_t)cnt * (P_FIXED + 3 * L) * 32);
    for (DevBuf* d : {&s->a, &s->b, &s->ypow, &s->gcoef, &s->hcoef})
      if (e == cudaSuccess) e = d->ensure(vec);
    if (e == cudaSuccess) e = s->alpha.ensure((size_t)cnt * 32);
    if (e == cudaSuccess) e = s->rnd.ensure((size_t)cnt * 64);
    if (e == cudaSuccess) e = s->values.ensure((size_t)cnt * m * 8);
    if (e == cudaSuccess) e = s->gammas.ensure((size_t)cnt * m * 32);
    if (e == cudaSuccess) e = s->ext.ensure((size_t)cnt * C::EXT_WORDS * 4);
    if (e == cudaSuccess) e = s->proofs.ensure((size_t)cnt * plen);
    if (e == cudaSuccess) e = s->flag.ensure(16);
    if (e == cudaSuccess && nonces) e = s->out.ensure((size_t)cnt * (5 + 2 * L) * 32);
    auto fail = [&](int rc) { s->release(); delete s; return rc; };
    if (e != cudaSuccess) { ctx->err = cudaGetErrorString(e); return fail(BPP_ERR_CUDA); }
    int rc = [&]() -> int {
      BPP_CK(cudaMemsetAsync(s->flag.p, 0, 16, ctx->stream));
      BPP_CK(cudaMemsetAsync(s->proofs.p, 0, (size_t)cnt * plen, ctx->stream));
      BPP_CK(cudaMemcpyAsync(s->values.p, v, (size_t)cnt * m * 8, cudaMemcpyHostToDevice, ctx->stream));
      BPP_CK(cudaMemcpyAsync(s->gammas.p, gamma, (size_t)cnt * m * 32, cudaMemcpyHostToDevice, ctx->stream));
      if (nonces) BPP_CK(cudaMemcpyAsync(s->out.p, nonces, (size_t)cnt * (5 + 2 * L) * 32, cudaMemcpyHostToDevice, ctx->stream));
      FsState st = session_state(s);
      ProveShape ps = session_shape(s);
      BPP_LAUNCH(fs_params_init_kernel<M>, ceil_div(cnt, 128), 128, 0, cnt, L, (uint32_t)(m == 1), nonces ? s->out.as<uint8_t>() : nullptr, st,
                 s->flag.as<uint8_t>());
      BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)cnt * 32));
      BPP_CK(ctx->buf[B_PIDX].ensure((size_t)cnt * 4));
      BPP_LAUNCH(fs_alpha_terms_kernel<M>, ceil_div(cnt, 128), 128, 0, cnt, st, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_PIDX].as<uint32_t>());
      BPP_LAUNCH(range_bits_sum_kernel<C>, cnt, 128, 0, ps, s->values.as<uint64_t>(), pk->table, s->ext.as<uint32_t>());
      ctx->tab_gen = 0;
      return session_emit(ctx, s, cnt, 1, 1, nullptr, s->ext.as<uint32_t>(), 0, 0, out_A);
    }();
    if (rc != BPP_OK) return fail(rc);
    *out = s;
    return BPP_OK;
  }
  // step 0: challenges = count x [y, z]; step k in 1..L: count x e_{k-1}. out: count x [L_k, R_k] (steps < L) or count x [A, B] (step L)
  static int prove_step(bpp_ctx* ctx, bpp_prove_session* s, const uint8_t* challenges, uint8_t* out_points) {
    if (s->step > s->L) { ctx->err = "prove session: no step left, call finish"; return BPP_ERR_LENGTH; }
    const uint32_t k = s->step, cnt = s->count, N = s->N, L = s->L;
    FsState st = session_state(s);
    ProveShape ps = session_shape(s);
    const uint32_t n_ch = k == 0 ? 2 : 1;
    BPP_CK(ctx->buf[B_IN0].ensure((size_t)cnt * n_ch * 32));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN0].p, challenges, (size_t)cnt * n_ch * 32, cudaMemcpyHostToDevice, ctx->stream));
    if (k == 0) {
      BPP_LAUNCH(fs_store_challenges_kernel<M>, ceil_div(cnt, 128), 128, 0, cnt, 2u, (uint32_t)P_Y, (uint32_t)P_Z, 1u, ctx->buf[B_IN0].as<uint8_t>(), st,
                 s->flag.as<uint8_t>());
      BPP_LAUNCH(fs_init_kernel<M>, cnt, 256, 0, ps, st, s->values.as<uint64_t>(), s->gammas.as<uint32_t>(), s->flag.as<uint8_t>());
    } else {
      BPP_LAUNCH(fs_store_challenges_kernel<M>, ceil_div(cnt, 128), 128, 0, cnt, 1u, (uint32_t)(P_FIXED + k - 1), 0u, 1u,
                 ctx->buf[B_IN0].as<uint8_t>(), st, s->flag.as<uint8_t>());
      BPP_LAUNCH(fs_fold_kernel<M>, cnt, 256, 0, ps, st, k - 1);
    }
    ctx->tab_gen = 0;
    int rc;
    if (k < L) {
      const uint32_t T = cnt * 2 * (N + 2);
      BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)T * 32));
      BPP_CK(ctx->buf[B_PIDX].ensure((size_t)T * 4));
      BPP_LAUNCH(fs_round_kernel<M>, cnt, 256, 256 * 32, ps, st, k, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_PIDX].as<uint32_t>());
      rc = session_emit(ctx, s, T, 2, N + 2, nullptr, nullptr, 1 + k, 1 + L + k, out_points);
    } else {
      const uint32_t T = cnt * (2 * N + 4);
      BPP_CK(ctx->buf[B_SCALARS].ensure((size_t)T * 32));
      BPP_CK(ctx->buf[B_PIDX].ensure((size_t)T * 4));
      BPP_CK(ctx->buf[B_JOBSTART].ensure(((size_t)cnt * 2 + 1) * 4));
      BPP_LAUNCH(fs_final_points_kernel<M>, cnt, 256, 0, ps, st, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_PIDX].as<uint32_t>(),
                 ctx->buf[B_JOBSTART].as<uint32_t>());
      rc = session_emit(ctx, s, T, 2, 0, ctx->buf[B_JOBSTART].as<uint32_t>(), nullptr, 1 + 2 * L, 2 + 2 * L, out_points);
    }
    if (rc == BPP_OK) s->step++;
    return rc;
  }
  static int prove_finish(bpp_ctx* ctx, bpp_prove_session* s, const uint8_t* e_final, uint8_t* proofs) {
    if (s->step != s->L + 1) { ctx->err = "prove session: finish comes after step L"; return BPP_ERR_LENGTH; }
    const uint32_t cnt = s->count;
    const size_t plen = (2 * (size_t)s->L + 3) * PB + 96;
    FsState st = session_state(s);
    ProveShape ps = session_shape(s);
    BPP_CK(ctx->buf[B_IN0].ensure((size_t)cnt * 32));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN0].p, e_final, (size_t)cnt * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_LAUNCH(fs_store_challenges_kernel<M>, ceil_div(cnt, 128), 128, 0, cnt, 1u, (uint32_t)P_EFINAL, 0u, 1u, ctx->buf[B_IN0].as<uint8_t>(), st,
               s->flag.as<uint8_t>());
    BPP_LAUNCH(fs_finish_kernel<M>, ceil_div(cnt, 128), 128, 0, ps, st, s->proofs.as<uint8_t>(), plen, (2 * (size_t)s->L + 3) * PB);
    BPP_CK(cudaMemcpyAsync(proofs, s->proofs.p, (size_t)cnt * plen, cudaMemcpyDeviceToHost, ctx->stream));
    uint8_t bad = 0;
    BPP_TRY(read_flag(ctx, s->flag.as<uint8_t>(), &bad));
    if (bad) { ctx->err = "prove session: a challenge is zero or not canonical"; return BPP_ERR_FORMAT; }
    s->step++;
    return BPP_OK;
  }

  // ---- WeightedInnerProductProof::prove / verify (standalone) -----------------------------------------
  static int check_wip_shape(bpp_ctx* ctx, const bpp_pk* pk) {
    const size_t N = pk->len;
    if (N < 2) return BPP_ERR_LENGTH;
    if (N & (N - 1)) return BPP_ERR_NOT_POW2;  // wip.rs:67
    return BPP_OK;
  }
  static int wip_prove(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* a, const uint8_t* b, const uint8_t* ypow, const uint8_t* gamma,
                       const bpp_params* prm, uint8_t* proof, size_t* proof_len, const uint8_t* order_le) {
    BPP_TRY(check_wip_shape(ctx, pk));
    const size_t N = pk->len;
    const uint32_t L = ilog2(N);
    const size_t plen = (2 * (size_t)L + 2) * PB + 96;
    if (*proof_len < plen) { *proof_len = plen; return BPP_ERR_LENGTH; }
    BPP_TRY(load_params(ctx, prm, false, L, order_le));
    BPP_CK(ctx->buf[B_IN0].ensure(N * 32));
    BPP_CK(ctx->buf[B_IN1].ensure(N * 32));
    BPP_CK(ctx->buf[B_IN2].ensure(N * 32));
    BPP_CK(ctx->buf[B_IN3].ensure(32 + plen));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN0].p, a, N * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN1].p, b, N * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN2].p, ypow, N * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(ctx->buf[B_IN3].p, gamma, 32, cudaMemcpyHostToDevice, ctx->stream));
    uint8_t bad = 0;
    uint8_t* d_proof = ctx->buf[B_IN3].as<uint8_t>() + 32;
    BPP_TRY(prove_core(ctx, pk, 0, 0, 0, 1, nullptr, ctx->buf[B_IN3].as<uint32_t>(), ctx->buf[B_IN0].as<uint32_t>(),
                       ctx->buf[B_IN1].as<uint32_t>(), ctx->buf[B_IN2].as<uint32_t>(), d_proof, plen, &bad));
    if (bad) return BPP_ERR_FORMAT;
    BPP_CK(cudaMemcpyAsync(proof, d_proof, plen, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    *proof_len = plen;
    return BPP_OK;
  }
  static int wip_verify(bpp_ctx* ctx, const bpp_pk* pk, const uint8_t* ypow, const uint8_t* Gc, const uint8_t* Hc, const uint8_t* gc, size_t nV,
                        const uint8_t* Vc, const uint8_t* A_prime, const uint8_t* V, const uint8_t* proof, size_t proof_len,
                        const bpp_params* prm, const uint8_t* order_le) {
    BPP_TRY(check_wip_shape(ctx, pk));
    const size_t N = pk->len;
    const uint32_t L = ilog2(N);
    const size_t wlen = (2 * (size_t)L + 2) * PB + 96;
    if (proof_len != wlen) return BPP_ERR_FORMAT;  // wrong number of rounds: wip.rs:334-337 reports VerificationError for n != 2^|L_vec|
    if (nV > (1u << 16)) return BPP_ERR_LENGTH;
    BPP_TRY(load_params(ctx, prm, false, L, order_le));
    BPP_TRY(clear_flags(ctx));
    // scalars: [Gc N][Hc N][Vc nV][gc][y]; bytes: [A_prime | proof][V nV][verdict]
    const size_t n_sc = 2 * N + nV + 2;
    BPP_CK(ctx->buf[B_WORK0].ensure(n_sc * 32));
    BPP_CK(ctx->buf[B_WORK1].ensure(PB + wlen + nV * PB + 16));
    uint8_t* d_sc = ctx->buf[B_WORK0].as<uint8_t>();
    uint8_t* d_by = ctx->buf[B_WORK1].as<uint8_t>();
    BPP_CK(cudaMemcpyAsync(d_sc, Gc, N * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(d_sc + N * 32, Hc, N * 32, cudaMemcpyHostToDevice, ctx->stream));
    if (nV) BPP_CK(cudaMemcpyAsync(d_sc + 2 * N * 32, Vc, nV * 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(d_sc + (2 * N + nV) * 32, gc, 32, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(d_sc + (2 * N + nV + 1) * 32, ypow, 32, cudaMemcpyHostToDevice, ctx->stream));  // y = power_of_y_vec[0], wip.rs:253
    BPP_CK(cudaMemcpyAsync(d_by, A_prime, PB, cudaMemcpyHostToDevice, ctx->stream));
    BPP_CK(cudaMemcpyAsync(d_by + PB, proof, wlen, cudaMemcpyHostToDevice, ctx->stream));
    if (nV) BPP_CK(cudaMemcpyAsync(d_by + PB + wlen, V, nV * PB, cudaMemcpyHostToDevice, ctx->stream));
    BPP_LAUNCH(scalars_check_kernel<M>, ceil_div((uint32_t)n_sc, 128), 128, 0, (uint32_t)n_sc, reinterpret_cast<const uint32_t*>(d_sc),
               ctx->buf[B_FLAGS].as<uint8_t>());
    uint8_t* d_verdict = d_by + PB + wlen + nV * PB;
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(d_sc);
    BPP_TRY(verify_core(ctx, pk, 0, 0, (uint32_t)nV, 1, d_by + PB + wlen, d_by, PB + wlen, s32, s32 + N * 8, s32 + (2 * N + nV) * 8,
                        s32 + 2 * N * 8, s32 + (2 * N + nV + 1) * 8, d_verdict));
    uint8_t res[2] = {0, 0};
    BPP_CK(cudaMemcpyAsync(&res[0], d_verdict, 1, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaMemcpyAsync(&res[1], ctx->buf[B_FLAGS].p, 1, cudaMemcpyDeviceToHost, ctx->stream));
    BPP_CK(cudaStreamSynchronize(ctx->stream));
    if (res[1]) return BPP_ERR_FORMAT;
    return res[0];
  }
};

}  // namespace bpp

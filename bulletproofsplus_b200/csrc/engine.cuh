// Host-side engine behind the C ABI (include/bpp.h): stream, scratch arena, and the launch sequences of
// the kernels in msm.cuh / proto_*.cuh. Templated on (curve policy C, scalar modulus M).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/bpp.h"
#include "msm.cuh"
#include "pk_direct.cuh"
#include "msm_straus.cuh"
#include "proto_prove.cuh"
#include "proto_session.cuh"

#ifndef BPP_HYBRID_DEFAULT
#define BPP_HYBRID_DEFAULT 0x0u  // block roles of the hybrid accumulation kernel (bit i: blocks with index % 4 == i use the FP64 pipe)
#endif

namespace bpp {

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) { cudaError_t e = cudaFree(p); if (e != cudaSuccess) return e; p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

enum BufId {
  B_SCALARS, B_PIDX, B_JOBSTART, B_COUNTS, B_OFFSETS, B_CURSOR, B_TILES, B_SORTED, B_PERM, B_BINS, B_BUCKETS, B_CHUNK_S, B_CHUNK_T,
  B_WINDOWS, B_FIXEDRES, B_RESULTS, B_DYN, B_DYN_OK, B_IN0, B_IN1, B_IN2, B_IN3, B_OUT0, B_PARAMS, B_TAB0, B_TAB1, B_TAB2, B_TAB3, B_TAB4,
  B_TAB5, B_TAB6, B_WORK0, B_WORK1, B_WORK2, B_FLAGS, B_FLAGS2, B_EXT, B_GIANT_CTL, B_GIANT_ITEMS, B_GIANT_KEYS, B_GIANT_PART, B_DYNF, B_CHUNK_S2, B_CHUNK_T2, B_COUNT_
};

}  // namespace bpp

struct bpp_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  bpp::DevBuf buf[bpp::B_COUNT_];
  // second scratch set + stream for the generic segment of a two-segment MSM, which runs beside the fixed-base segment
  // (Msm::run): its Horner pass is a chain of ~SCALAR_BITS dependent doublings per job, latency-bound on a few warps
  bpp::DevBuf buf2[bpp::B_COUNT_];
  cudaStream_t aux_stream = nullptr;
  bool aux_swapped = false;  // ctx->stream currently names the auxiliary stream (Msm::aux_enter / swap_scratch)
  cudaEvent_t aux_fork = nullptr, aux_join = nullptr;
  std::string err;
  uint64_t launches = 0;
  // hot-kernel timing (bucket accumulation)
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool hot_timing = false;
  double hot_ms = 0;
  uint64_t hot_launches = 0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> hot_pending;
  std::vector<uint32_t> host_jobstart;
  // The challenge-only tables of the prover / verifier (B_TAB0..5) depend on the parameter block and the proof shape, not on
  // the proofs: a call that repeats the previous call's parameters and shape reuses them (one single-block kernel and one
  // upload + sync less per batch). params_gen counts distinct uploads; tab_* describe what B_TAB* currently holds.
  std::vector<uint8_t> params_host;
  uint64_t params_gen = 0;
  uint64_t tab_gen = 0;      // params_gen the tables were built from; 0 = none
  uint32_t tab_kind = 0;     // 1 verifier tables, 2 prover tables
  uint32_t tab_shape[5] = {0, 0, 0, 0, 0};  // N, n, m, mode, curve
  const void* tab_pk = nullptr;
  // launch tracer (bpp_trace): one event pair per BPP_LAUNCH on whichever stream it went to; the timeline of a call over
  // both streams is read back by bpp_trace_dump. Off by default (two event records per launch when on).
  struct TraceRec { const char* name; cudaEvent_t e0, e1; int aux; };
  bool trace_on = false;
  std::vector<TraceRec> trace;
  std::vector<struct bpp_prove_session*> session_pool;  // at most two parked sessions (cudaMalloc/cudaFree cost milliseconds each)
  size_t tab_caps = 0;       // capacities of the table buffers when they were filled (a re-allocation invalidates them)
};

struct bpp_pk {
  int curve = 0;
  size_t len = 0;
  int device = 0;
  uint32_t* table = nullptr;  // affine: g, h, G[len], H[len]
  uint32_t* wtab = nullptr;   // window tables: wtab[w * n_points + i] = 2^(cA * w) * table[i]
  uint32_t* wtab_f = nullptr; // the same in the FP64 accumulators' slot format (curves with HAS_F64; hybrid accumulation)
  uint32_t* comb = nullptr;   // comb tables of g and h (K5): [(p * comb_w + w) * 128 + d - 1] = d * 2^(8w) * {g, h}
  uint32_t cA = 0, WA = 0, comb_w = 0;
  // direct fixed-base tables (bpp_pk_precompute, pk_direct.cuh): every signed-digit multiple of every window point; nullptr
  // until built. With them the fixed-base segment of an MSM needs no buckets (Msm::run_direct).
  uint32_t* dtab = nullptr;
  uint32_t cD = 0, WD = 0;
  uint32_t n_points() const { return (uint32_t)(2 + 2 * len); }
};

struct bpp_points {
  int curve = 0;
  size_t len = 0;
  uint32_t* table = nullptr;  // affine
  uint32_t* wtab = nullptr;   // optional window tables (bpp_points_precompute): wtab[w * len + i] = 2^(cA * w) * table[i]
  uint32_t* table_f = nullptr, *wtab_f = nullptr;  // `table` / `wtab` in the FP64 accumulators' slot format (HAS_F64; built on first use)
  uint32_t cA = 0, WA = 0;
  uint32_t* dtab = nullptr;   // optional direct tables (bpp_points_precompute_direct), laid out like bpp_pk::dtab
  uint32_t cD = 0, WD = 0;
};

// interactive proving (bpp_range_prove_begin / _step / _finish): device state of a batch between the host's challenges
struct bpp_prove_session {
  int curve = 0;
  const bpp_pk* pk = nullptr;
  uint32_t n = 0, m = 0, count = 0, N = 0, L = 0;
  uint32_t step = 0;  // next expected: 0..L = bpp_range_prove_step, L + 1 = bpp_range_prove_finish, L + 2 = done
  bpp_ctx* owner = nullptr;  // freed sessions are parked in the owner's pool: a same-shaped begin reuses the device memory
  bpp::DevBuf params, a, b, alpha, ypow, gcoef, hcoef, rnd, values, gammas, ext, proofs, out, flag;
  void release() {
    for (bpp::DevBuf* d : {&params, &a, &b, &alpha, &ypow, &gcoef, &hcoef, &rnd, &values, &gammas, &ext, &proofs, &out, &flag}) d->release();
  }
};

namespace bpp {

#define BPP_CK(call)                                                                              \
  do {                                                                                            \
    cudaError_t _e = (call);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(_e);                              \
      return BPP_ERR_CUDA;                                                                        \
    }                                                                                             \
  } while (0)
#define BPP_TRY(expr)            \
  do {                           \
    int _rc = (expr);            \
    if (_rc != BPP_OK) return _rc; \
  } while (0)
#define BPP_LAUNCH(kernel, grid, block, smem, ...)                         \
  do {                                                                     \
    cudaEvent_t _t0 = nullptr, _t1 = nullptr;                              \
    if (ctx->trace_on) { cudaEventCreate(&_t0); cudaEventCreate(&_t1); cudaEventRecord(_t0, ctx->stream); } \
    kernel<<<(grid), (block), (smem), ctx->stream>>>(__VA_ARGS__);         \
    ctx->launches++;                                                       \
    BPP_CK(cudaGetLastError());                                            \
    if (ctx->trace_on) {                                                   \
      cudaEventRecord(_t1, ctx->stream);                                   \
      ctx->trace.push_back({#kernel, _t0, _t1, ctx->aux_swapped ? 1 : 0}); \
    }                                                                      \
  } while (0)

// The same for a kernel that takes a thread-block cluster of `cs` blocks (1, 2, 4 or 8; grid must be a multiple of it).
#define BPP_LAUNCH_CLUSTER(kernel, grid, block, smem, cs, ...)                 \
  do {                                                                     \
    cudaEvent_t _t0 = nullptr, _t1 = nullptr;                              \
    if (ctx->trace_on) { cudaEventCreate(&_t0); cudaEventCreate(&_t1); cudaEventRecord(_t0, ctx->stream); } \
    cudaLaunchConfig_t _cfg = {};                                          \
    _cfg.gridDim = dim3(grid); _cfg.blockDim = dim3(block); _cfg.dynamicSmemBytes = (smem); _cfg.stream = ctx->stream; \
    cudaLaunchAttribute _at[1];                                            \
    _at[0].id = cudaLaunchAttributeClusterDimension;                       \
    _at[0].val.clusterDim.x = (cs); _at[0].val.clusterDim.y = 1; _at[0].val.clusterDim.z = 1; \
    _cfg.attrs = _at; _cfg.numAttrs = 1;                                   \
    BPP_CK(cudaLaunchKernelEx(&_cfg, kernel, __VA_ARGS__));                \
    ctx->launches++;                                                       \
    BPP_CK(cudaGetLastError());                                            \
    if (ctx->trace_on) {                                                   \
      cudaEventRecord(_t1, ctx->stream);                                   \
      ctx->trace.push_back({#kernel, _t0, _t1, ctx->aux_swapped ? 1 : 0}); \
    }                                                                      \
  } while (0)

// block roles of the hybrid accumulation kernel: bit i set = blocks with (index % 4) == i run on the FP64 pipe. 0 (the
// default: measured, mixing the two multiplies gains nothing on sm_100a — DESIGN.md §3.2) = integer pipe only.
inline uint32_t hybrid_pattern() {
  static const uint32_t v = (getenv("BPP_HYBRID") ? (uint32_t)strtoul(getenv("BPP_HYBRID"), nullptr, 16) : BPP_HYBRID_DEFAULT) & 0xfu;
  return v;
}
inline uint32_t ceil_div(uint64_t a, uint64_t b) { return (uint32_t)((a + b - 1) / b); }
inline uint32_t ilog2(uint64_t x) { uint32_t l = 0; while ((1ull << (l + 1)) <= x) l++; return l; }

// Cost model in units of one mixed addition (7 field multiplies); a full addition costs 9/7.
// fixed-base: W * K additions + one reduction of 2^(c-1) buckets (2 full additions per bucket) per job
inline uint32_t choose_window_fixed(uint64_t terms_per_job, uint32_t scalar_bits) {
  uint32_t best = 4;
  double best_cost = 1e300;
  for (uint32_t c = 4; c <= 16; c++) {
    uint64_t W = (scalar_bits + 1 + c - 1) / c, B = 1ull << (c - 1);
    double cost = (double)W * (double)terms_per_job + 2.0 * (double)B * 9.0 / 7.0;
    if (cost < best_cost) { best_cost = cost; best = c; }
  }
  return best;
}
// generic: W * (K additions + one reduction) + Horner
inline uint32_t choose_window_generic(uint64_t terms_per_job, uint32_t scalar_bits, uint64_t n_jobs) {
  uint32_t best = 2;
  double best_cost = 1e300;
  for (uint32_t c = 2; c <= 18; c++) {
    uint64_t W = (scalar_bits + 1 + c - 1) / c, B = 1ull << (c - 1);
    if (n_jobs * W * B > (1ull << 26)) continue;  // keep the key space (and bucket memory) bounded
    double cost = (double)W * ((double)terms_per_job + 2.0 * (double)B * 9.0 / 7.0);
    if (cost < best_cost) { best_cost = cost; best = c; }
  }
  return best;
}

template <class C>
struct Msm {
  // a two-segment MSM over this many jobs runs its generic segment on the auxiliary stream (run() below)
  static bool overlaps(uint32_t n_jobs) {
    static const bool off = getenv("BPP_NO_OVERLAP") != nullptr;
    return n_jobs >= 16 && !off;
  }
  static int aux_ensure(bpp_ctx* ctx) {
    if (!ctx->aux_stream) {
      // highest priority: its many small kernels take SM slots as blocks of the big accumulation grid retire, instead of
      // queueing behind that grid (measured without it: no overlap at all)
      int prio_lo = 0, prio_hi = 0;
      BPP_CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
      BPP_CK(cudaStreamCreateWithPriority(&ctx->aux_stream, cudaStreamNonBlocking, prio_hi));
      BPP_CK(cudaEventCreateWithFlags(&ctx->aux_fork, cudaEventDisableTiming));
      BPP_CK(cudaEventCreateWithFlags(&ctx->aux_join, cudaEventDisableTiming));
    }
    return BPP_OK;
  }
  // Work that only the generic segment needs (decoding the per-call points) can start on the auxiliary stream before run():
  // aux_enter forks from the caller's stream and redirects launches until aux_leave. The caller must go on to a run() for
  // which overlaps() holds — that run joins the auxiliary stream back.
  static int aux_enter(bpp_ctx* ctx) {
    BPP_TRY(aux_ensure(ctx));
    BPP_CK(cudaEventRecord(ctx->aux_fork, ctx->stream));
    BPP_CK(cudaStreamWaitEvent(ctx->aux_stream, ctx->aux_fork, 0));
    std::swap(ctx->stream, ctx->aux_stream);
    ctx->aux_swapped = !ctx->aux_swapped;
    return BPP_OK;
  }
  static void aux_leave(bpp_ctx* ctx) { std::swap(ctx->stream, ctx->aux_stream); ctx->aux_swapped = !ctx->aux_swapped; }

  // Runs the segmented MSM over terms already in ctx->buf[B_SCALARS] / [B_PIDX]. Terms with point index < n_fixed use the
  // window tables `wtab` (cA, WA from the public key); the others index `dyn` (per-call affine table) at idx - n_fixed and
  // number about `dyn_per_job` per job (0 = none). job_start: device array (n_jobs+1) or nullptr when `uniform` > 0.
  // Leaves n_jobs extended points in ctx->buf[B_RESULTS].
  static int run(bpp_ctx* ctx, uint32_t n_terms, uint32_t n_jobs, uint32_t uniform, const uint32_t* d_job_start,
                 const uint32_t* wtab, uint32_t n_fixed, uint32_t cA, uint32_t WA, const uint32_t* dyn, uint32_t dyn_per_job,
                 const uint32_t* wtab_f = nullptr, const uint32_t* dyn_f = nullptr, const uint32_t* dtab = nullptr, uint32_t cD = 0,
                 uint32_t WD = 0, const DirectVerifySrc* vsrc = nullptr) {
    if (n_jobs == 0) return BPP_OK;
    // direct tables exist for the Ed25519 policy only. Any number of jobs takes them: run_direct cuts a job into up to 64 slices,
    // so that one proof's generator terms are 8192 threads' work for a few dozen additions each (one n = 64, m = 16
    // verification 886 -> 682 us, one proof 1022 -> 702 us: profiles/r02ar_ab.txt; BPP_DIRECT_MIN_JOBS=32 restores the old rule)
    const bool direct = uses_direct(dtab, n_jobs) && n_fixed && wtab;
    if (vsrc && !direct) { ctx->err = "msm: fused exponents need the direct tables"; return BPP_ERR_UNSUPPORTED; }
    // with direct tables the two segments are different kernels anyway: run them side by side for any number of jobs
    const bool two = n_fixed && wtab && dyn_per_job && (overlaps(n_jobs) || (direct && overlaps(16)));
    if (!two) {
      if (!direct) return run_seg(ctx, n_terms, n_jobs, uniform, d_job_start, wtab, n_fixed, cA, WA, dyn, dyn_per_job, 3, wtab_f, dyn_f);
      BPP_CK(ctx->buf[B_RESULTS].ensure((size_t)n_jobs * C::EXT_WORDS * 4));
      if (!dyn_per_job) return run_direct(ctx, n_terms, n_jobs, uniform, d_job_start, n_fixed, dtab, cD, WD, ctx->buf[B_RESULTS].as<uint32_t>(), vsrc);
      // few jobs with both kinds of terms: generic segment, then the fixed-base terms from the direct tables, one stream
      BPP_TRY(run_seg(ctx, n_terms, n_jobs, uniform, d_job_start, wtab, n_fixed, cA, WA, dyn, dyn_per_job, 2, wtab_f, dyn_f));
      BPP_CK(ctx->buf[B_FIXEDRES].ensure((size_t)n_jobs * C::EXT_WORDS * 4));
      BPP_TRY(run_direct(ctx, n_terms, n_jobs, uniform, d_job_start, n_fixed, dtab, cD, WD, ctx->buf[B_FIXEDRES].as<uint32_t>(), vsrc));
      BPP_LAUNCH(msm_join_kernel<C>, ceil_div(n_jobs, 128), 128, 0, n_jobs, ctx->buf[B_RESULTS].as<uint32_t>(), ctx->buf[B_FIXEDRES].as<uint32_t>());
      return BPP_OK;
    }
    // Two segments, many jobs: the generic segment (sort, buckets, reduction, Horner) goes to the auxiliary stream with its
    // own scratch and overlaps the fixed-base segment on the main stream; one addition per job joins them.
    BPP_TRY(aux_ensure(ctx));
    BPP_CK(ctx->buf[B_RESULTS].ensure((size_t)n_jobs * C::EXT_WORDS * 4));
    BPP_CK(cudaEventRecord(ctx->aux_fork, ctx->stream));
    BPP_CK(cudaStreamWaitEvent(ctx->aux_stream, ctx->aux_fork, 0));
    static const BufId scratch[] = {B_COUNTS, B_OFFSETS, B_CURSOR, B_TILES, B_SORTED, B_PERM, B_BINS, B_BUCKETS, B_CHUNK_S, B_CHUNK_T, B_WINDOWS,
                                    B_FIXEDRES, B_RESULTS, B_GIANT_CTL, B_GIANT_ITEMS, B_GIANT_KEYS, B_GIANT_PART, B_DYNF, B_CHUNK_S2, B_CHUNK_T2};
    auto swap_scratch = [&]() {
      for (BufId id : scratch) std::swap(ctx->buf[id], ctx->buf2[id]);
      std::swap(ctx->stream, ctx->aux_stream);
      ctx->aux_swapped = !ctx->aux_swapped;
    };
    swap_scratch();
    const bool timing = ctx->hot_timing;
    ctx->hot_timing = false;  // the hot-kernel bracket stays on the main stream
    int rc = run_seg(ctx, n_terms, n_jobs, uniform, d_job_start, wtab, n_fixed, cA, WA, dyn, dyn_per_job, 2, wtab_f, dyn_f);
    ctx->hot_timing = timing;
    const uint32_t* part_b = ctx->buf[B_RESULTS].as<uint32_t>();
    if (rc == BPP_OK && cudaEventRecord(ctx->aux_join, ctx->stream) != cudaSuccess) { ctx->err = "cudaEventRecord(aux_join)"; rc = BPP_ERR_CUDA; }
    swap_scratch();
    // whatever was queued on the auxiliary stream shares B_SCALARS / B_DYN with the next call: on any failure from here on
    // wait for it before returning
    if (rc != BPP_OK) { cudaStreamSynchronize(ctx->aux_stream); return rc; }
    rc = direct ? run_direct(ctx, n_terms, n_jobs, uniform, d_job_start, n_fixed, dtab, cD, WD, ctx->buf[B_RESULTS].as<uint32_t>(), vsrc)
                : run_seg(ctx, n_terms, n_jobs, uniform, d_job_start, wtab, n_fixed, cA, WA, dyn, 0, 1, wtab_f, dyn_f);
    if (rc != BPP_OK) { cudaStreamSynchronize(ctx->aux_stream); return rc; }
    BPP_CK(cudaStreamWaitEvent(ctx->stream, ctx->aux_join, 0));
    BPP_LAUNCH(msm_join_kernel<C>, ceil_div(n_jobs, 128), 128, 0, n_jobs, ctx->buf[B_RESULTS].as<uint32_t>(), part_b);
    return BPP_OK;
  }

  // The fixed-base terms (point index < n_fixed) of every job straight from the key's direct tables: ONE kernel, no sort, no
  // buckets, no reduction (pk_direct.cuh). Terms with other indices are skipped (the generic segment takes them).
  // whether a launch sequence over n_jobs jobs takes the direct tables (callers that fuse work into that kernel ask beforehand)
  static bool uses_direct(const uint32_t* dtab, uint32_t n_jobs) {
    static const uint32_t min_jobs = getenv("BPP_DIRECT_MIN_JOBS") ? (uint32_t)atoi(getenv("BPP_DIRECT_MIN_JOBS")) : 1u;
    return dtab && C::HAS_COOP4 && n_jobs >= min_jobs;
  }
  static int run_direct(bpp_ctx* ctx, uint32_t n_terms, uint32_t n_jobs, uint32_t uniform, const uint32_t* d_job_start, uint32_t n_fixed,
                        const uint32_t* dtab, uint32_t cD, uint32_t WD, uint32_t* out, const DirectVerifySrc* vsrc = nullptr) {
    DirectVerifySrc src;
    memset(&src, 0, sizeof(src));
    if (vsrc) src = *vsrc;
    MsmPlan pl;
    memset(&pl, 0, sizeof(pl));
    pl.n_terms = n_terms; pl.n_jobs = n_jobs; pl.uniform = uniform; pl.n_fixed = n_fixed; pl.cA = cD; pl.WA = WD;
    for (int i = 0; i < 8; i++) pl.order[i] = C::order_word(i);
    // threads per slice: at least ~8 terms per thread for small jobs, never fewer than 4 threads (the tree is cheap); slices per
    // job: enough blocks for ~8 waves of the 592 resident ones, while every thread keeps at least two whole terms
    const uint32_t avg = n_terms / n_jobs;
    uint32_t tpj = 128;
    while (tpj > 4 && tpj * 8 > avg) tpj >>= 1;
    uint32_t S = 1;
    while (S < 8 && (uint64_t)ceil_div(n_jobs, 128 / tpj) * S < 8 * 592 && (uint64_t)tpj * S * 2 * 2 <= avg) S <<= 1;
    // few jobs (a single proof, a handful): keep doubling while the launch stays within one wave of the chip and every lane
    // keeps at least one (term, window) pair - one job then runs on up to 64 x tpj threads instead of 8 x tpj
    while (S < 64 && (uint64_t)n_jobs * tpj * S * 2 <= 592 * 128 && (uint64_t)tpj * S * 2 <= (uint64_t)avg * WD) S <<= 1;
    uint32_t* partial = out;
    if (S > 1) {
      BPP_CK(ctx->buf[B_CHUNK_S2].ensure((size_t)n_jobs * S * C::EXT_WORDS * 4));
      partial = ctx->buf[B_CHUNK_S2].as<uint32_t>();
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->hot_timing) {
      BPP_CK(cudaEventCreate(&e0));
      BPP_CK(cudaEventCreate(&e1));
      BPP_CK(cudaEventRecord(e0, ctx->stream));
    }
    if constexpr (C::HAS_COOP4) {
      if (src.enabled)
        BPP_LAUNCH((direct_accumulate_kernel<C, true>), ceil_div((uint64_t)n_jobs * S, 128 / tpj), 128, 0, pl, tpj, S, ctx->buf[B_SCALARS].as<uint32_t>(),
                   ctx->buf[B_PIDX].as<uint32_t>(), d_job_start, dtab, partial, src);
      else
        BPP_LAUNCH((direct_accumulate_kernel<C, false>), ceil_div((uint64_t)n_jobs * S, 128 / tpj), 128, 0, pl, tpj, S, ctx->buf[B_SCALARS].as<uint32_t>(),
                   ctx->buf[B_PIDX].as<uint32_t>(), d_job_start, dtab, partial, src);
      if (S > 8) BPP_LAUNCH(direct_join_tree_kernel<C>, ceil_div(n_jobs, 4), 128, 0, n_jobs, S, partial, out);
      else if (S > 1) BPP_LAUNCH(direct_join_kernel<C>, ceil_div(n_jobs, 128), 128, 0, n_jobs, S, partial, out);
    } else {
      ctx->err = "direct tables: unsupported curve"; return BPP_ERR_UNSUPPORTED;
    }
    if (ctx->hot_timing) {
      BPP_CK(cudaEventRecord(e1, ctx->stream));
      ctx->hot_pending.push_back({e0, e1});
      ctx->hot_launches++;
    }
    return BPP_OK;
  }

  // Generic segment of jobs with at most STRAUS_MAX_D per-call points each (msm_straus.cuh): window sums, then the Horner kernel
  // of the bucket path. Result in B_RESULTS like run_seg(seg = 2).
  static int run_straus(bpp_ctx* ctx, MsmPlan pl, const uint32_t* d_job_start, const uint32_t* dyn, uint32_t dyn_per_job) {
    if constexpr (C::HAS_COOP4) {
      pl.cA = pl.WA = pl.NBA = 0;
      pl.cB = 4; pl.WB = STRAUS_WINDOWS; pl.NBB = 8;
      BPP_CK(ctx->buf[B_WINDOWS].ensure((size_t)pl.n_jobs * STRAUS_WINDOWS * C::EXT_WORDS * 4));
      BPP_CK(ctx->buf[B_RESULTS].ensure((size_t)pl.n_jobs * C::EXT_WORDS * 4));
      uint32_t* windows = ctx->buf[B_WINDOWS].as<uint32_t>();
      auto k = msm_straus_kernel<C>;
      const size_t smem = straus_smem_bytes(dyn_per_job);
      if (smem > 40 * 1024) BPP_CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      BPP_LAUNCH(k, pl.n_jobs, 64, smem, pl, dyn_per_job, ctx->buf[B_SCALARS].as<uint32_t>(), ctx->buf[B_PIDX].as<uint32_t>(), d_job_start, dyn, windows);
      BPP_LAUNCH(msm_combine4_kernel<C>, ceil_div(pl.n_jobs, 32), 128, 0, pl, windows, ctx->buf[B_FIXEDRES].as<uint32_t>(), ctx->buf[B_RESULTS].as<uint32_t>());
      return BPP_OK;
    } else {
      ctx->err = "straus: unsupported curve"; return BPP_ERR_UNSUPPORTED;
    }
  }

  // seg: bit 0 = fixed-base segment, bit 1 = generic segment; terms of a segment that is left out are skipped by the sort
  // (their window count is zero).
  static int run_seg(bpp_ctx* ctx, uint32_t n_terms, uint32_t n_jobs, uint32_t uniform, const uint32_t* d_job_start,
                     const uint32_t* wtab, uint32_t n_fixed, uint32_t cA, uint32_t WA, const uint32_t* dyn, uint32_t dyn_per_job, int seg,
                     const uint32_t* wtab_f, const uint32_t* dyn_f) {
    MsmPlan pl;
    memset(&pl, 0, sizeof(pl));
    pl.n_terms = n_terms; pl.n_jobs = n_jobs; pl.uniform = uniform;
    if (n_fixed && wtab) pl.n_fixed = n_fixed;
    if (n_fixed && wtab && (seg & 1)) {
      pl.cA = cA; pl.WA = WA; pl.NBA = 1u << (cA - 1);
      pl.qA = pl.NBA < 16 ? pl.NBA : 16;
    }
    if (!(seg & 2)) dyn_per_job = 0;
    // the generic segment alone, a few per-call points per job: Straus window sums in one kernel instead of the bucket pipeline
    if constexpr (C::HAS_COOP4) {
      static const bool no_straus = getenv("BPP_NO_STRAUS") != nullptr;
      if (seg == 2 && pl.n_fixed && dyn && dyn_per_job && dyn_per_job <= STRAUS_MAX_D && !no_straus)
        return run_straus(ctx, pl, d_job_start, dyn, dyn_per_job);
    }
    if (dyn_per_job) {
      pl.cB = choose_window_generic(dyn_per_job, C::SCALAR_BITS, n_jobs);
      pl.WB = (C::SCALAR_BITS + 1 + pl.cB - 1) / pl.cB;
      pl.NBB = 1u << (pl.cB - 1);
      pl.qB = pl.NBB < 16 ? pl.NBB : 16;
    }
    pl.wtab_points = pl.WA * pl.n_fixed;
    const uint32_t KPJ = pl.NBA + pl.WB * pl.NBB;
    // per-job sort in shared memory (one block per job) when the job's key space fits AND there are enough jobs to fill the
    // chip or the jobs are small; one or a few large jobs (a raw MSM of 2^12 points took longer than one of 2^14 when a
    // single 512-thread block sorted it) go through the grid-wide histogram / scan / scatter instead
    const uint32_t Wmax = pl.WA > pl.WB ? pl.WA : pl.WB;
    const bool job_sort = KPJ <= 12288 && (n_jobs >= 16 || (uint64_t)n_terms * Wmax / n_jobs <= 16384);
    pl.sort_mult = 1;
    if (pl.NBA) {
      for (int i = 0; i < 8; i++) pl.order[i] = C::order_word(i);
      // split digits (msm.cuh): only on the per-job shared-memory sort, and only where some job can be small enough
      if (job_sort && pl.cA >= 8 && getenv("BPP_NO_SPLIT") == nullptr) {
        const uint32_t k2 = pl.NBA / 9, k3 = pl.NBA / 64;
        if (!uniform || uniform <= k2) {
          pl.k2 = k2; pl.k3 = k3;
          pl.sub2 = (pl.cA + 1) / 2; pl.sub3 = (pl.cA + 2) / 3;
          pl.sort_mult = (!uniform || uniform <= k3) ? 3 : 2;
        }
      }
    }
    const uint64_t keysA64 = (uint64_t)n_jobs * pl.NBA, keysB64 = (uint64_t)n_jobs * pl.WB * pl.NBB;
    if (keysA64 + keysB64 >= (1ull << 30) || keysA64 + keysB64 == 0) { ctx->err = "msm: key space out of range"; return BPP_ERR_LENGTH; }
    pl.keysA = (uint32_t)keysA64;
    pl.n_keys = (uint32_t)(keysA64 + keysB64);
    pl.chunksA = pl.NBA ? pl.keysA / pl.qA : 0;
    const uint32_t chunksB = pl.WB ? (uint32_t)(keysB64 / pl.qB) : 0;
    pl.n_chunks = pl.chunksA + chunksB;
    const uint32_t n_keys = pl.n_keys;
    const uint64_t n_entries = pl.WA ? (uint64_t)n_terms * pl.WA * pl.sort_mult + (uint64_t)n_jobs * dyn_per_job * (pl.WB > pl.WA ? pl.WB - pl.WA : 0)
                                      : pl.n_fixed ? (uint64_t)n_jobs * dyn_per_job * pl.WB  // generic segment alone: fixed-base terms are skipped
                                      : (uint64_t)n_terms * pl.WB;  // upper bound; the per-job slices of the shared-memory sort
    const uint32_t n_setsB = n_jobs * pl.WB;
    BPP_CK(ctx->buf[B_COUNTS].ensure((size_t)n_keys * 4));
    BPP_CK(ctx->buf[B_OFFSETS].ensure(((size_t)n_keys + 1) * 4));
    BPP_CK(ctx->buf[B_CURSOR].ensure((size_t)n_keys * 4));
    BPP_CK(ctx->buf[B_PERM].ensure((size_t)n_keys * 4));
    BPP_CK(ctx->buf[B_BINS].ensure((256 + 8) * 4));
    const uint32_t n_tiles = ceil_div(n_keys, SCAN_TILE);
    BPP_CK(ctx->buf[B_TILES].ensure((size_t)n_tiles * 4));
    BPP_CK(ctx->buf[B_SORTED].ensure((size_t)(n_entries ? n_entries : 1) * 4));
    BPP_CK(ctx->buf[B_BUCKETS].ensure((size_t)n_keys * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_CHUNK_S].ensure((size_t)pl.n_chunks * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_CHUNK_T].ensure((size_t)pl.n_chunks * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_WINDOWS].ensure((size_t)(n_setsB ? n_setsB : 1) * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_FIXEDRES].ensure((size_t)n_jobs * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_RESULTS].ensure((size_t)n_jobs * C::EXT_WORDS * 4));
    // giant buckets: sum over buckets of ceil(count / GIANT_SEG) <= entries / GIANT_SEG + (number of giant buckets)
    const size_t max_items = 2 * (size_t)(n_entries / GIANT_SEG) + 64;
    BPP_CK(ctx->buf[B_GIANT_CTL].ensure(16));
    BPP_CK(ctx->buf[B_GIANT_ITEMS].ensure(max_items * sizeof(uint2)));
    BPP_CK(ctx->buf[B_GIANT_KEYS].ensure(max_items * sizeof(uint4)));
    BPP_CK(ctx->buf[B_GIANT_PART].ensure(max_items * C::EXT_WORDS * 4));
    uint32_t* gctl = ctx->buf[B_GIANT_CTL].as<uint32_t>();
    uint2* gitems = ctx->buf[B_GIANT_ITEMS].as<uint2>();
    uint4* gkeys = ctx->buf[B_GIANT_KEYS].as<uint4>();
    uint32_t* gpart = ctx->buf[B_GIANT_PART].as<uint32_t>();
    uint32_t* counts = ctx->buf[B_COUNTS].as<uint32_t>();
    uint32_t* offsets = ctx->buf[B_OFFSETS].as<uint32_t>();
    uint32_t* cursor = ctx->buf[B_CURSOR].as<uint32_t>();
    uint32_t* tiles = ctx->buf[B_TILES].as<uint32_t>();
    uint32_t* sorted = ctx->buf[B_SORTED].as<uint32_t>();
    uint32_t* perm = ctx->buf[B_PERM].as<uint32_t>();
    uint32_t* bins = ctx->buf[B_BINS].as<uint32_t>();
    uint32_t* buckets = ctx->buf[B_BUCKETS].as<uint32_t>();
    uint32_t* chunk_S = ctx->buf[B_CHUNK_S].as<uint32_t>();
    uint32_t* chunk_T = ctx->buf[B_CHUNK_T].as<uint32_t>();
    uint32_t* windows = ctx->buf[B_WINDOWS].as<uint32_t>();
    uint32_t* fixedres = ctx->buf[B_FIXEDRES].as<uint32_t>();
    uint32_t* results = ctx->buf[B_RESULTS].as<uint32_t>();
    const uint32_t* scalars = ctx->buf[B_SCALARS].as<uint32_t>();
    const uint32_t* pidx = ctx->buf[B_PIDX].as<uint32_t>();
    // hybrid accumulation (msm_accumulate_kernel<C, true>): blocks chosen by `pattern` run on the FP64 pipe and need the
    // tables in its slot format; per-call points are converted here
    uint32_t pattern = C::HAS_F64 ? hybrid_pattern() : 0u;
    if (pl.NBA && !wtab_f) pattern = 0;
    if (pattern && dyn && dyn_per_job && !dyn_f) {
      const uint32_t n_dyn = n_jobs * dyn_per_job;
      BPP_CK(ctx->buf[B_DYNF].ensure((size_t)n_dyn * C::AFF_WORDS * 4));
      BPP_LAUNCH(msm_table_format_kernel<C>, ceil_div(n_dyn, 128), 128, 0, n_dyn, dyn, ctx->buf[B_DYNF].as<uint32_t>());
      dyn_f = ctx->buf[B_DYNF].as<uint32_t>();
    }
    BPP_CK(cudaMemsetAsync(bins, 0, 256 * 4, ctx->stream));
    BPP_CK(cudaMemsetAsync(gctl, 0, 16, ctx->stream));
    if (n_entries >= (1ull << 32)) { ctx->err = "msm: too many bucket entries for one call"; return BPP_ERR_LENGTH; }
    if (job_sort) {
      // small key space per job: one block sorts a whole job in shared memory
      auto k = msm_sort_job_kernel<0>;
      const size_t smem = (size_t)KPJ * 4;
      // the kernel also has 64 B of static shared memory: opt in before dynamic + static can pass the 48 KB default
      if (smem > 40 * 1024) BPP_CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      BPP_LAUNCH(k, n_jobs, SJ_THREADS, smem, pl, dyn_per_job, scalars, pidx, d_job_start, offsets, counts, sorted);
    } else {
      BPP_CK(cudaMemsetAsync(counts, 0, (size_t)n_keys * 4, ctx->stream));
      // window groups (msm_digits_kernel): one generic job whose references do not fit in L2 is scattered a few windows at a
      // time, each pass filling the GPU, so that the slice being written (<= 32 MB) stays in L2 until it is complete
      uint32_t n_groups = 1, gw = Wmax;
      if (n_jobs == 1 && !pl.NBA && pl.WB && (uint64_t)n_terms * pl.WB * 4 > (48ull << 20)) {
        gw = (uint32_t)((32ull << 20) / ((uint64_t)n_terms * 4));
        gw = gw < 1 ? 1 : gw;
        n_groups = ceil_div(pl.WB, gw);
      }
      uint32_t dg = n_terms ? (ceil_div(n_terms, 256) < 148 * 16 ? ceil_div(n_terms, 256) : 148 * 16) : 1;
      if (n_groups > 1) dg = 148 * 8 * n_groups;  // 8 blocks of 256 threads per SM = one full wave per group
      BPP_LAUNCH(msm_digits_kernel<0>, dg, 256, 0, pl, n_groups, gw, scalars, pidx, d_job_start, counts, sorted);
      BPP_LAUNCH(scan_tile_sums, n_tiles, SCAN_THREADS, 0, counts, n_keys, tiles);
      BPP_LAUNCH(scan_tile_offsets, 1, SCAN_THREADS, 0, tiles, n_tiles);
      BPP_LAUNCH(scan_apply, n_tiles, SCAN_THREADS, 0, counts, n_keys, tiles, offsets, cursor);
      BPP_LAUNCH(msm_digits_kernel<1>, dg, 256, 0, pl, n_groups, gw, scalars, pidx, d_job_start, cursor, sorted);
    }
    const uint32_t ss_tiles = ceil_div(n_keys, SS_TILE);
    BPP_LAUNCH(sizesort_hist_kernel, ss_tiles, SS_THREADS, 0, counts, n_keys, bins);
    // few buckets in all (one small MSM): spread buckets over lanes much earlier — the longest one-thread chain is the kernel
    const bool lat = n_keys <= LAT_MAX_KEYS;
    const uint32_t acc_cap = lat ? ACC_CAP_LAT : ACC_CAP, warp_cap = lat ? WARP_CAP_LAT : WARP_CAP;
    BPP_LAUNCH(sizesort_bins_scan_kernel, 1, 256, 0, bins, bins + 256, acc_cap, warp_cap);
    BPP_LAUNCH(sizesort_scatter_kernel, ss_tiles, SS_THREADS, 0, counts, n_keys, bins, perm);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->hot_timing) {
      BPP_CK(cudaEventCreate(&e0));
      BPP_CK(cudaEventCreate(&e1));
      BPP_CK(cudaEventRecord(e0, ctx->stream));
    }
    {  // the long buckets first: a warp each for count > warp_cap, four lanes each for acc_cap < count <= warp_cap
      const uint32_t g32 = ceil_div(n_keys, 4) < 1184 ? ceil_div(n_keys, 4) : 1184, g4 = ceil_div(n_keys, 32) < 1184 ? ceil_div(n_keys, 32) : 1184;
      auto k32 = msm_accumulate_group_kernel<C, 32>;
      auto k4 = msm_accumulate_group_kernel<C, 4>;
      if (n_entries > GIANT_SEG) {  // a bucket can only be giant if there are that many entries at all
        BPP_LAUNCH(msm_giant_enum_kernel, g32 < 148 ? g32 : 148, 256, 0, bins + 256, perm, counts, gctl, gitems, gkeys);
        BPP_LAUNCH(msm_giant_accumulate_kernel<C>, g32, 128, 0, gctl, gitems, offsets, counts, sorted, wtab, pl.wtab_points, dyn, gpart);
        BPP_LAUNCH(msm_giant_merge_kernel<C>, g32 < 296 ? g32 : 296, 128, 0, gctl, gkeys, gpart, buckets);
      }
      BPP_LAUNCH(k32, g32, 128, 0, bins + 256, 0u, perm, offsets, counts, sorted, wtab, pl.wtab_points, dyn, buckets);
      BPP_LAUNCH(k4, g4, 128, 0, bins + 256, 1u, perm, offsets, counts, sorted, wtab, pl.wtab_points, dyn, buckets);
    }
    if (pattern)
      BPP_LAUNCH((msm_accumulate_kernel<C, true>), ceil_div(n_keys, 128), 128, 0, n_keys, perm, offsets, counts, sorted, wtab, pl.wtab_points, dyn, buckets,
                 wtab_f, dyn_f, pattern, acc_cap);
    else
      BPP_LAUNCH((msm_accumulate_kernel<C, false>), ceil_div(n_keys, 128), 128, 0, n_keys, perm, offsets, counts, sorted, wtab, pl.wtab_points, dyn, buckets,
                 nullptr, nullptr, 0u, acc_cap);
    if (ctx->hot_timing) {
      BPP_CK(cudaEventRecord(e1, ctx->stream));
      ctx->hot_pending.push_back({e0, e1});
      ctx->hot_launches++;
    }
    static const bool chunks4 = getenv("BPP_CHUNKS_BLOCKS") && atoi(getenv("BPP_CHUNKS_BLOCKS")) == 4;  // A/B: 128 registers, 4 blocks per SM
    if (chunks4) BPP_LAUNCH((msm_reduce_chunks_kernel<C, 4>), ceil_div(pl.n_chunks, 128), 128, 0, pl, counts, buckets, chunk_S, chunk_T);
    else BPP_LAUNCH((msm_reduce_chunks_kernel<C, 1>), ceil_div(pl.n_chunks, 128), 128, 0, pl, counts, buckets, chunk_S, chunk_T);
    auto sets = [&](uint32_t n_sets, uint32_t first_chunk, uint32_t q, uint32_t NB, uint32_t* out) -> int {
      uint32_t cps = NB / q;  // chunks per set
      const uint32_t* cS = chunk_S + (size_t)first_chunk * C::EXT_WORDS;
      const uint32_t* cT = chunk_T + (size_t)first_chunk * C::EXT_WORDS;
      // few sets with many chunks each (one large MSM): merge groups of chunks into heavier chunks with whole-chip kernels
      // until a set is 128 chunks, so that the set kernel below has no serial walk left (it was 0.36 ms of a 2^22-point MSM)
      if (n_sets < 256 && cps > 128) {
        BPP_CK(ctx->buf[B_CHUNK_S2].ensure((size_t)n_sets * (cps / 2) * C::EXT_WORDS * 4));
        BPP_CK(ctx->buf[B_CHUNK_T2].ensure((size_t)n_sets * (cps / 2) * C::EXT_WORDS * 4));
        uint32_t* pingS = ctx->buf[B_CHUNK_S2].as<uint32_t>();
        uint32_t* pingT = ctx->buf[B_CHUNK_T2].as<uint32_t>();
        uint32_t* pongS = chunk_S + (size_t)first_chunk * C::EXT_WORDS;  // the consumed input of the previous level
        uint32_t* pongT = chunk_T + (size_t)first_chunk * C::EXT_WORDS;
        while (cps > 128) {
          const uint32_t per = cps / 128 < 8 ? cps / 128 : 8, groups = n_sets * (cps / per);
          BPP_LAUNCH(msm_reduce_level_kernel<C>, ceil_div(groups, 128), 128, 0, groups, q, per, cS, cT, pingS, pingT);
          cS = pingS; cT = pingT;
          std::swap(pingS, pongS); std::swap(pingT, pongT);
          cps /= per; q *= per;
        }
      }
      // threads per set: the group's suffix scan costs log2(tps) additions in EVERY thread, so many sets get few threads
      // each (serial running sums over `per` chunks, work-efficient) and few sets get many (latency); ~32 K threads in all
      uint32_t tps = 1;
      while (tps < 128 && tps < cps && (uint64_t)tps * n_sets < 32768) tps <<= 1;
      const uint32_t per = cps / tps;
      BPP_LAUNCH(msm_reduce_sets_kernel<C>, ceil_div(n_sets, 128 / tps), 128, 0, n_sets, 0u, q, per, tps, cS, cT, out);
      return BPP_OK;
    };
    if (pl.NBA) BPP_TRY(sets(n_jobs, 0, pl.qA, pl.NBA, pl.WB ? fixedres : results));
    if (pl.WB) {
      BPP_TRY(sets(n_setsB, pl.chunksA, pl.qB, pl.NBB, windows));
      if constexpr (C::HAS_COOP4) {
        // four warps per 32 jobs, one per SM sub-partition, each computing one of the four products of every level
        BPP_LAUNCH(msm_combine4_kernel<C>, ceil_div(n_jobs, 32), 128, 0, pl, windows, fixedres, results);
      } else {
        // jobs per warp: one warp per SM at most. More single-warp blocks than that (tried: ~8 per SM) fill every register
        // file with this kernel's 255-register threads and hold up the other stream's kernels for the whole Horner chain.
        uint32_t lanes = ceil_div(n_jobs, 148);
        lanes = lanes > 32 ? 32 : lanes;
        BPP_LAUNCH(msm_combine_kernel<C>, ceil_div(n_jobs, lanes), 32, 0, pl, lanes, windows, fixedres, results);
      }
    }
    return BPP_OK;
  }
};

}  // namespace bpp

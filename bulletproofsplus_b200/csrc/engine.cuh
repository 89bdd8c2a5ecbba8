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
#include "proto_prove.cuh"

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
  B_SCALARS, B_PIDX, B_JOBSTART, B_COUNTS, B_OFFSETS, B_CURSOR, B_TILES, B_SORTED, B_BUCKETS, B_CHUNK_S, B_CHUNK_T, B_WINDOWS,
  B_RESULTS, B_DYN, B_DYN_OK, B_IN0, B_IN1, B_IN2, B_IN3, B_OUT0, B_PARAMS, B_TAB0, B_TAB1, B_TAB2, B_TAB3, B_TAB4, B_TAB5, B_TAB6,
  B_WORK0, B_WORK1, B_WORK2, B_FLAGS, B_FLAGS2, B_EXT, B_COUNT_
};

}  // namespace bpp

struct bpp_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  bpp::DevBuf buf[bpp::B_COUNT_];
  std::string err;
  uint64_t launches = 0;
  // hot-kernel timing (bucket accumulation)
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool hot_timing = false;
  double hot_ms = 0;
  uint64_t hot_launches = 0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> hot_pending;
  std::vector<uint32_t> host_jobstart;
};

struct bpp_pk {
  int curve = 0;
  size_t len = 0;
  int device = 0;
  uint32_t* table = nullptr;  // affine: g, h, G[len], H[len]
  uint32_t n_points() const { return (uint32_t)(2 + 2 * len); }
};

struct bpp_points {
  int curve = 0;
  size_t len = 0;
  uint32_t* table = nullptr;  // affine
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
    kernel<<<(grid), (block), (smem), ctx->stream>>>(__VA_ARGS__);         \
    ctx->launches++;                                                       \
    BPP_CK(cudaGetLastError());                                            \
  } while (0)

inline uint32_t ceil_div(uint64_t a, uint64_t b) { return (uint32_t)((a + b - 1) / b); }
inline uint32_t ilog2(uint64_t x) { uint32_t l = 0; while ((1ull << (l + 1)) <= x) l++; return l; }

// Window width minimising the algorithmic cost of SURVEY §8(d): W * (K*madd + 2*2^(c-1)*add) for K terms per job
inline uint32_t choose_window(uint64_t terms_per_job, uint32_t scalar_bits, uint64_t n_jobs, double madd_cost, double add_cost) {
  uint32_t best = 2;
  double best_cost = 1e300;
  for (uint32_t c = 2; c <= 16; c++) {
    uint64_t W = (scalar_bits + 1 + c - 1) / c, B = 1ull << (c - 1);
    if (n_jobs * W * B > (1ull << 27)) continue;  // keep the key space (and bucket memory) bounded
    double cost = (double)W * ((double)terms_per_job * madd_cost + 2.0 * (double)B * add_cost);
    if (cost < best_cost) { best_cost = cost; best = c; }
  }
  return best;
}

template <class C>
struct Msm {
  // Runs the segmented MSM over terms already in ctx->buf[B_SCALARS] / [B_PIDX]. job_start: device array (n_jobs+1)
  // or nullptr when `uniform` > 0. Leaves n_jobs extended points in ctx->buf[B_RESULTS].
  static int run(bpp_ctx* ctx, uint32_t n_terms, uint32_t n_jobs, uint32_t uniform, const uint32_t* d_job_start,
                 const uint32_t* tab0, uint32_t n0, const uint32_t* tab1, uint32_t window_override = 0) {
    if (n_jobs == 0) return BPP_OK;
    MsmShape sh;
    sh.n_terms = n_terms;
    sh.n_jobs = n_jobs;
    sh.uniform = uniform;
    uint64_t per_job = n_jobs ? (n_terms + n_jobs - 1) / n_jobs : 0;
    sh.c = window_override ? window_override : choose_window(per_job ? per_job : 1, C::SCALAR_BITS, n_jobs, 7.0, 9.0);
    sh.n_windows = (C::SCALAR_BITS + 1 + sh.c - 1) / sh.c;
    sh.n_buckets = 1u << (sh.c - 1);
    sh.chunk = sh.n_buckets < 64 ? sh.n_buckets : (sh.n_buckets >= 4096 ? 64 : 32);
    const uint64_t n_keys64 = (uint64_t)n_jobs * sh.n_windows * sh.n_buckets;
    if (n_keys64 >= (1ull << 31)) { ctx->err = "msm: key space too large"; return BPP_ERR_LENGTH; }
    const uint32_t n_keys = (uint32_t)n_keys64;
    const uint32_t n_jw = n_jobs * sh.n_windows;
    const uint32_t n_chunks = n_jw * (sh.n_buckets / sh.chunk);
    const uint64_t n_entries = (uint64_t)n_terms * sh.n_windows;
    BPP_CK(ctx->buf[B_COUNTS].ensure((size_t)n_keys * 4));
    BPP_CK(ctx->buf[B_OFFSETS].ensure(((size_t)n_keys + 1) * 4));
    BPP_CK(ctx->buf[B_CURSOR].ensure((size_t)n_keys * 4));
    const uint32_t n_tiles = ceil_div(n_keys, SCAN_TILE);
    BPP_CK(ctx->buf[B_TILES].ensure((size_t)n_tiles * 4));
    BPP_CK(ctx->buf[B_SORTED].ensure((size_t)(n_entries ? n_entries : 1) * 4));
    BPP_CK(ctx->buf[B_BUCKETS].ensure((size_t)n_keys * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_CHUNK_S].ensure((size_t)n_chunks * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_CHUNK_T].ensure((size_t)n_chunks * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_WINDOWS].ensure((size_t)n_jw * C::EXT_WORDS * 4));
    BPP_CK(ctx->buf[B_RESULTS].ensure((size_t)n_jobs * C::EXT_WORDS * 4));
    uint32_t* counts = ctx->buf[B_COUNTS].as<uint32_t>();
    uint32_t* offsets = ctx->buf[B_OFFSETS].as<uint32_t>();
    uint32_t* cursor = ctx->buf[B_CURSOR].as<uint32_t>();
    uint32_t* tiles = ctx->buf[B_TILES].as<uint32_t>();
    uint32_t* sorted = ctx->buf[B_SORTED].as<uint32_t>();
    uint32_t* buckets = ctx->buf[B_BUCKETS].as<uint32_t>();
    const uint32_t* scalars = ctx->buf[B_SCALARS].as<uint32_t>();
    const uint32_t* pidx = ctx->buf[B_PIDX].as<uint32_t>();
    BPP_CK(cudaMemsetAsync(counts, 0, (size_t)n_keys * 4, ctx->stream));
    const uint32_t dg = n_terms ? (ceil_div(n_terms, 256) < 148 * 16 ? ceil_div(n_terms, 256) : 148 * 16) : 1;
    BPP_LAUNCH(msm_digits_kernel<0>, dg, 256, 0, sh, scalars, pidx, d_job_start, counts, sorted);
    BPP_LAUNCH(scan_tile_sums, n_tiles, SCAN_THREADS, 0, counts, n_keys, tiles);
    BPP_LAUNCH(scan_tile_offsets, 1, SCAN_THREADS, 0, tiles, n_tiles);
    BPP_LAUNCH(scan_apply, n_tiles, SCAN_THREADS, 0, counts, n_keys, tiles, offsets, cursor);
    BPP_LAUNCH(msm_digits_kernel<1>, dg, 256, 0, sh, scalars, pidx, d_job_start, cursor, sorted);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ctx->hot_timing) {
      BPP_CK(cudaEventCreate(&e0));
      BPP_CK(cudaEventCreate(&e1));
      BPP_CK(cudaEventRecord(e0, ctx->stream));
    }
    BPP_LAUNCH(msm_accumulate_kernel<C>, ceil_div(n_keys, 128), 128, 0, n_keys, offsets, sorted, tab0, n0, tab1, buckets);
    if (ctx->hot_timing) {
      BPP_CK(cudaEventRecord(e1, ctx->stream));
      ctx->hot_pending.push_back({e0, e1});
      ctx->hot_launches++;
    }
    BPP_LAUNCH(msm_reduce_chunks_kernel<C>, ceil_div(n_chunks, 128), 128, 0, sh, buckets, ctx->buf[B_CHUNK_S].as<uint32_t>(),
               ctx->buf[B_CHUNK_T].as<uint32_t>());
    BPP_LAUNCH(msm_reduce_windows_kernel<C>, ceil_div(n_jw, 128), 128, 0, sh, ctx->buf[B_CHUNK_S].as<uint32_t>(),
               ctx->buf[B_CHUNK_T].as<uint32_t>(), ctx->buf[B_WINDOWS].as<uint32_t>());
    BPP_LAUNCH(msm_combine_kernel<C>, ceil_div(n_jobs, 128), 128, 0, sh, ctx->buf[B_WINDOWS].as<uint32_t>(),
               ctx->buf[B_RESULTS].as<uint32_t>());
    return BPP_OK;
  }
};

}  // namespace bpp

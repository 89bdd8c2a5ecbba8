// K1/K2 — segmented signed-digit Pippenger multi-scalar multiplication for sm_100a.
//
// Replaces MulVec::calculate (reference src/bls12_381/building_block/mulvec.rs:20-33: one full
// G1::mul + one G1::add per term) for every caller on the path. One launch sequence evaluates J
// independent MSMs ("jobs"): J = 1 for a large MSM (config 5), J = one per proof for batch
// verification (weighted_inner_product_proof.rs:297-320, range/mod.rs:480-503), J = 2L+3 per proof
// for proving. Terms are (32-byte canonical scalar, point reference) pairs; a point reference
// indexes an affine table in HBM (bit 31 = negate).
//
// Pipeline (all kernels templated on the curve policy C):
//   msm_digits_count   signed-digit recode, c-bit windows; histogram of (job, window, bucket) keys
//   scan_*             exclusive prefix sum of the histogram (3-phase, hand-written)
//   msm_digits_scatter counting-sort scatter of point references by key
//   msm_accumulate     HOT: one thread per bucket walks its list with mixed additions (IMAD-bound)
//   msm_reduce_chunks / msm_reduce_windows   running-sum bucket reduction, chunked
//   msm_combine        Horner over windows (c doublings per window)
// Group addition is commutative, so the (unstable) atomic scatter order cannot change the result;
// outputs are compared only after canonical encoding.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace bpp {

struct MsmShape {
  uint32_t n_terms;    // total terms over all jobs
  uint32_t n_jobs;
  uint32_t c;          // window bits
  uint32_t n_windows;  // ceil((scalar_bits + 1) / c)
  uint32_t n_buckets;  // 2^(c-1)
  uint32_t chunk;      // buckets per reduce chunk (power of two, divides n_buckets)
  uint32_t uniform;    // > 0: every job has exactly this many terms (job = t / uniform)
};

__device__ __forceinline__ uint32_t msm_find_job(const uint32_t* __restrict__ job_start, uint32_t n_jobs, uint32_t uniform, uint32_t t) {
  if (uniform) return t / uniform;
  uint32_t lo = 0, hi = n_jobs;  // largest j with job_start[j] <= t
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (job_start[mid] <= t) lo = mid; else hi = mid;
  }
  return lo;
}

// c-bit field at bit offset `off` of a 256-bit little-endian scalar (zero beyond bit 255)
__device__ __forceinline__ uint32_t msm_bits(const uint32_t* s, uint32_t off, uint32_t c) {
  uint32_t w = off >> 5, sh = off & 31;
  if (w >= 8) return 0;
  uint64_t lo = s[w];
  uint64_t hi = (w + 1 < 8) ? s[w + 1] : 0;
  return (uint32_t)(((hi << 32) | lo) >> sh) & ((1u << c) - 1u);
}

// MODE 0: histogram; MODE 1: scatter
template <int MODE>
__global__ void msm_digits_kernel(MsmShape sh, const uint32_t* __restrict__ scalars, const uint32_t* __restrict__ pidx,
                                  const uint32_t* __restrict__ job_start, uint32_t* __restrict__ counts_or_cursor,
                                  uint32_t* __restrict__ sorted) {
  const uint32_t half = 1u << (sh.c - 1);
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < sh.n_terms; t += gridDim.x * blockDim.x) {
    uint32_t s[8];
    const uint4* sp = reinterpret_cast<const uint4*>(scalars + 8ull * t);
    uint4 a = __ldg(sp), b = __ldg(sp + 1);
    s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
    uint32_t job = msm_find_job(job_start, sh.n_jobs, sh.uniform, t);
    uint32_t ref = MODE ? pidx[t] : 0;
    uint32_t carry = 0;
    for (uint32_t w = 0; w < sh.n_windows; w++) {
      uint32_t raw = msm_bits(s, w * sh.c, sh.c) + carry;
      uint32_t neg = raw > half;
      uint32_t mag = neg ? (1u << sh.c) - raw : raw;
      carry = neg;
      if (mag) {
        uint32_t key = (job * sh.n_windows + w) * sh.n_buckets + (mag - 1);
        if (MODE == 0) atomicAdd(&counts_or_cursor[key], 1u);
        else sorted[atomicAdd(&counts_or_cursor[key], 1u)] = ref ^ (neg << 31);
      }
    }
  }
}

// ---- exclusive scan over n uint32 (3 phases). out has n+1 entries; out2 (optional) gets a copy of out[0..n) ----
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total_out) {
  __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
  __shared__ uint32_t total;
  uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) warp_sums[wid] = x;
  __syncthreads();
  if (wid == 0) {
    uint32_t s = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < SCAN_THREADS / 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += y;
    }
    if (lane < SCAN_THREADS / 32) warp_sums[lane] = s;
    if (lane == SCAN_THREADS / 32 - 1) total = s;
  }
  __syncthreads();
  uint32_t base = wid ? warp_sums[wid - 1] : 0;
  *total_out = total;
  __syncthreads();
  return base + x - v;
}

__global__ void scan_tile_sums(const uint32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ tile_sums) {
  uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) if (base + i < n) s += in[base + i];
  uint32_t tot;
  block_exclusive_scan(s, &tot);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}
__global__ void scan_tile_offsets(uint32_t* tile_sums, uint32_t n_tiles) {  // single block, in place -> exclusive
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n_tiles; base += SCAN_THREADS) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = i < n_tiles ? tile_sums[i] : 0, tot;
    uint32_t ex = block_exclusive_scan(v, &tot);
    if (i < n_tiles) tile_sums[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry += tot;
    __syncthreads();
  }
}
__global__ void scan_apply(const uint32_t* __restrict__ in, uint32_t n, const uint32_t* __restrict__ tile_offsets,
                           uint32_t* __restrict__ out, uint32_t* __restrict__ out2) {
  uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { v[i] = (base + i < n) ? in[base + i] : 0; s += v[i]; }
  uint32_t tot;
  uint32_t ex = block_exclusive_scan(s, &tot) + tile_offsets[blockIdx.x];
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) {
    if (base + i < n) { out[base + i] = ex; if (out2) out2[base + i] = ex; }
    ex += v[i];
    if (base + i + 1 == n) out[n] = ex;
  }
}

// ---- HOT: bucket accumulation -----------------------------------------------------------------------
// One thread per (job, window, bucket). tab0 holds the n0 resident points of the public key, tab1 the
// per-call points; both are AoS affine, read with 128-bit loads.
template <class C>
__global__ void __launch_bounds__(128)
msm_accumulate_kernel(uint32_t n_keys, const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ sorted,
                      const uint32_t* __restrict__ tab0, uint32_t n0, const uint32_t* __restrict__ tab1,
                      uint32_t* __restrict__ buckets) {
  uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
  if (key >= n_keys) return;
  uint32_t beg = offsets[key], end = offsets[key + 1];
  typename C::Ext acc;
  C::set_identity(acc);
  for (uint32_t e = beg; e < end; e++) {
    uint32_t ref = __ldg(sorted + e);
    uint32_t idx = ref & 0x7fffffffu;
    const uint32_t* src = idx < n0 ? tab0 + (size_t)idx * C::AFF_WORDS : tab1 + (size_t)(idx - n0) * C::AFF_WORDS;
    typename C::Aff q;
    load_words<C::AFF_WORDS>(q, src);
    C::madd(acc, q, ref >> 31);
  }
  store_words<C::EXT_WORDS>(buckets + (size_t)key * C::EXT_WORDS, acc);
}

// ---- bucket reduction: sum_b (b+1) * bucket[b], chunked ---------------------------------------------
// thread per (job*window, chunk): S = sum of the chunk's buckets, T = sum_{b in chunk} (b - lo + 1) * bucket[b]
template <class C>
__global__ void __launch_bounds__(128)
msm_reduce_chunks_kernel(MsmShape sh, const uint32_t* __restrict__ buckets, uint32_t* __restrict__ chunk_S,
                         uint32_t* __restrict__ chunk_T) {
  uint32_t n_chunks = sh.n_buckets / sh.chunk;
  uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= sh.n_jobs * sh.n_windows * n_chunks) return;
  uint32_t jw = id / n_chunks, ci = id % n_chunks;
  const uint32_t* base = buckets + ((size_t)jw * sh.n_buckets + (size_t)ci * sh.chunk) * C::EXT_WORDS;
  typename C::Ext run, sum, b;
  C::set_identity(run);
  C::set_identity(sum);
  for (int k = (int)sh.chunk - 1; k >= 0; k--) {
    load_words<C::EXT_WORDS>(b, base + (size_t)k * C::EXT_WORDS);
    C::add(run, b);
    C::add(sum, run);
  }
  store_words<C::EXT_WORDS>(chunk_S + (size_t)id * C::EXT_WORDS, run);
  store_words<C::EXT_WORDS>(chunk_T + (size_t)id * C::EXT_WORDS, sum);
}
// thread per (job*window): window = sum_ci T_ci + chunk * sum_ci ci * S_ci
template <class C>
__global__ void __launch_bounds__(128)
msm_reduce_windows_kernel(MsmShape sh, const uint32_t* __restrict__ chunk_S, const uint32_t* __restrict__ chunk_T,
                          uint32_t* __restrict__ windows) {
  uint32_t n_chunks = sh.n_buckets / sh.chunk;
  uint32_t jw = blockIdx.x * blockDim.x + threadIdx.x;
  if (jw >= sh.n_jobs * sh.n_windows) return;
  typename C::Ext run, sum, tsum, p;
  C::set_identity(run);
  C::set_identity(sum);
  C::set_identity(tsum);
  for (int ci = (int)n_chunks - 1; ci >= 0; ci--) {
    size_t id = (size_t)jw * n_chunks + ci;
    load_words<C::EXT_WORDS>(p, chunk_T + id * C::EXT_WORDS);
    C::add(tsum, p);
    if (ci > 0) {
      load_words<C::EXT_WORDS>(p, chunk_S + id * C::EXT_WORDS);
      C::add(run, p);
      C::add(sum, run);
    }
  }
  for (uint32_t s = sh.chunk; s > 1; s >>= 1) C::dbl(sum);
  C::add(tsum, sum);
  store_words<C::EXT_WORDS>(windows + (size_t)jw * C::EXT_WORDS, tsum);
}
// thread per job: Horner over windows
template <class C>
__global__ void __launch_bounds__(128)
msm_combine_kernel(MsmShape sh, const uint32_t* __restrict__ windows, uint32_t* __restrict__ results) {
  uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= sh.n_jobs) return;
  typename C::Ext acc, p;
  const uint32_t* base = windows + (size_t)job * sh.n_windows * C::EXT_WORDS;
  load_words<C::EXT_WORDS>(acc, base + (size_t)(sh.n_windows - 1) * C::EXT_WORDS);
  for (int w = (int)sh.n_windows - 2; w >= 0; w--) {
    for (uint32_t k = 0; k < sh.c; k++) C::dbl(acc);
    load_words<C::EXT_WORDS>(p, base + (size_t)w * C::EXT_WORDS);
    C::add(acc, p);
  }
  store_words<C::EXT_WORDS>(results + (size_t)job * C::EXT_WORDS, acc);
}

// ---- boundary kernels ---------------------------------------------------------------------------------
template <class C>
__global__ void points_decode_kernel(uint32_t n, const uint8_t* __restrict__ bytes, uint32_t* __restrict__ aff, uint8_t* __restrict__ ok) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename C::Aff q;
  bool good = C::decode(q, bytes + (size_t)i * C::POINT_BYTES);
  store_words<C::AFF_WORDS>(aff + (size_t)i * C::AFF_WORDS, q);
  ok[i] = good ? 1 : 0;
}
template <class C>
__global__ void points_encode_kernel(uint32_t n, const uint32_t* __restrict__ ext, uint8_t* __restrict__ bytes) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename C::Ext p;
  load_words<C::EXT_WORDS>(p, ext + (size_t)i * C::EXT_WORDS);
  C::encode(bytes + (size_t)i * C::POINT_BYTES, p);
}
template <class C>
__global__ void points_encode_aff_kernel(uint32_t n, const uint32_t* __restrict__ aff, uint8_t* __restrict__ bytes) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename C::Aff q;
  typename C::Ext p;
  load_words<C::AFF_WORDS>(q, aff + (size_t)i * C::AFF_WORDS);
  C::aff_to_ext(p, q);
  C::encode(bytes + (size_t)i * C::POINT_BYTES, p);
}
template <class C>
__global__ void points_is_identity_kernel(uint32_t n, const uint32_t* __restrict__ ext, uint8_t* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename C::Ext p;
  load_words<C::EXT_WORDS>(p, ext + (size_t)i * C::EXT_WORDS);
  out[i] = C::is_identity(p) ? 1 : 0;
}
template <class C>
__global__ void points_ext_to_aff_kernel(uint32_t n, const uint32_t* __restrict__ ext, uint32_t* __restrict__ aff) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename C::Ext p;
  typename C::Aff q;
  load_words<C::EXT_WORDS>(p, ext + (size_t)i * C::EXT_WORDS);
  C::to_aff(q, p);
  store_words<C::AFF_WORDS>(aff + (size_t)i * C::AFF_WORDS, q);
}

}  // namespace bpp

// K1/K2 — two-segment signed-digit bucket multi-scalar multiplication for sm_100a.
//
// Replaces MulVec::calculate (reference src/bls12_381/building_block/mulvec.rs:20-33: one full
// G1::mul + one G1::add per term) for every caller on the path. One launch sequence evaluates J
// independent MSMs ("jobs"): J = 1 for a large MSM (config 5), J = one per proof for batch
// verification (weighted_inner_product_proof.rs:297-320, range/mod.rs:480-503), J = 2L+3 per proof
// for proving. A term is (32-byte canonical scalar, point index).
//
// Segment A — FIXED-BASE terms (point index < n_fixed: the public key's g, h, G_i, H_i). The key
//   carries a window table wtab[w][i] = 2^(cA*w) * P_i (affine, built once in bpp_pk_new), so all
//   WA windows of a job drop into ONE set of 2^(cA-1) buckets: WA*K mixed additions and a single
//   bucket reduction per job, no doublings at all.
// Segment B — GENERIC terms (per-call points: proof elements, commitments, raw MSM inputs):
//   classic Pippenger, one bucket set per (job, window), Horner over windows at the end.
//
// Pipeline (all kernels templated on the curve policy C where they touch points):
//   msm_sort_job       small key spaces (batched proofs): per-job counting sort entirely in shared memory; else:
//   msm_digits<0>      signed-digit recode; histogram of bucket keys (global atomics)
//   scan_*             exclusive prefix sum of the histogram
//   msm_digits<1>      counting-sort scatter of point references by key
//   sizesort_*         permutation of the keys by descending bucket size: every warp of the hot
//                      kernel then owns 32 buckets of (almost) the same length -> no lane idles
//   msm_accumulate     HOT: one thread per bucket walks its list with mixed additions (IMAD-bound)
//   msm_reduce_chunks  running sums over chunks of 16 buckets
//   msm_reduce_sets    per bucket set: suffix scan + tree in shared memory -> sum_b (b+1) * B_b
//   msm_combine        Horner over the generic windows, plus the fixed-base part
// Group addition is commutative, so the (unstable) atomic scatter order cannot change the result;
// outputs are compared only after canonical encoding.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace bpp {

struct MsmPlan {
  uint32_t n_terms;  // total terms over all jobs
  uint32_t n_jobs;
  uint32_t uniform;  // > 0: every job has exactly this many terms (job = t / uniform)
  // segment A (fixed-base); NBA == 0 disables it
  uint32_t n_fixed;  // points with index < n_fixed have window tables
  uint32_t cA, WA, NBA;
  // segment B (generic); WB == 0 disables it
  uint32_t cB, WB, NBB;
  uint32_t keysA;        // n_jobs * NBA
  uint32_t n_keys;       // keysA + n_jobs * WB * NBB
  uint32_t wtab_points;  // WA * n_fixed: references >= this index the per-call table
  uint32_t qA, qB;       // reduce chunk sizes (powers of two dividing NBA / NBB)
  uint32_t chunksA;      // keysA / qA
  uint32_t n_chunks;     // chunksA + n_jobs * WB * NBB / qB
  // fixed-base scalars above order/2 are replaced by order - s with every digit sign flipped (the table points are
  // torsion-free multiples of the base point): the top partial window, which would otherwise collect ~K/2 carries per job
  // in ONE bucket, stays empty. order[0] == 0 && order[7] == 0 disables it.
  uint32_t order[8];
  // small jobs split every cA-bit digit into 2 or 3 signed sub-digits of `sub2` / `sub3` bits that reference the SAME
  // table point from buckets of weight d * 2^(i*sub): 2-3x the entries, but only ~2^(sub-1) * levels distinct buckets per
  // job to reduce instead of 2^(cA-1). Jobs with <= k3 terms use 3 levels, <= k2 terms 2 levels. sort_mult = capacity
  // multiplier of the per-job slices of `sorted` (1 when no job can split).
  uint32_t k2, k3, sub2, sub3, sort_mult;
};

// s (canonical, < order) -> min(s, order - s); returns 1 when negated
__device__ __forceinline__ uint32_t msm_fold_half(const MsmPlan& pl, uint32_t* s) {
  uint32_t t[8], bw = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint64_t d = (uint64_t)pl.order[i] - s[i] - bw;
    t[i] = (uint32_t)d;
    bw = (uint32_t)(d >> 63);
  }
  if (bw) return 0;  // s > order: not canonical, left alone (the callers reject such scalars before they get here)
  bool lt = false, decided = false;
#pragma unroll
  for (int i = 7; i >= 0; i--)
    if (!decided && t[i] != s[i]) { lt = t[i] < s[i]; decided = true; }
  if (!lt) return 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s[i] = t[i];
  return 1;
}

// Calls emit(bucket weight - 1, negative) for the entries of one window digit of magnitude mag (1..2^(cA-1)).
// levels == 1: one entry of weight mag. levels 2/3: signed base-2^sub digits d_i of mag, entry i has weight |d_i| * 2^(i*sub).
template <class F>
__device__ __forceinline__ void msm_split_digit(uint32_t mag, uint32_t neg, uint32_t levels, uint32_t sub, F emit) {
  if (levels == 1) { emit(mag - 1, neg); return; }
  const uint32_t base = 1u << sub, hb = base >> 1;
  uint32_t rest = mag;
  for (uint32_t i = 0; i + 1 < levels; i++) {
    uint32_t d = rest & (base - 1);
    rest >>= sub;
    uint32_t dn = d > hb;
    if (dn) { d = base - d; rest += 1; }
    if (d) emit((d << (i * sub)) - 1, neg ^ dn);
  }
  if (rest) emit((rest << ((levels - 1) * sub)) - 1, neg);
}

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

// MODE 0: histogram; MODE 1: scatter. Reference bit 31 = subtract the point.
// The windows are cut into `n_groups` groups of `gw` consecutive windows and the grid into n_groups equal parts: part g makes
// one pass over all terms for the windows of group g only (the signed-digit carry into the group's first window is recomputed
// from the lower bits, a few integer instructions per skipped window). With one group this is the plain kernel. With several,
// the blocks that run at the same time touch only that group's keys: for a single large generic MSM (one bucket set per
// window) the counters and the slice of `sorted` they scatter into — 2^22 points: 16 MB per window against 252 MB in all —
// stay in L2 until they are complete, instead of every 4-byte reference going to DRAM as a partial sector write
// (msm_digits<1> was 1.3 ms of a 7.4 ms MSM of 2^22 points).
template <int MODE>
__global__ void msm_digits_kernel(MsmPlan pl, uint32_t n_groups, uint32_t gw, const uint32_t* __restrict__ scalars, const uint32_t* __restrict__ pidx,
                                  const uint32_t* __restrict__ job_start, uint32_t* __restrict__ counts_or_cursor,
                                  uint32_t* __restrict__ sorted) {
  const uint32_t bpg = gridDim.x / n_groups;  // blocks per group (the host makes gridDim.x a multiple of n_groups)
  const uint32_t grp = blockIdx.x / bpg, blk = blockIdx.x - grp * bpg;
  const uint32_t w_lo = grp * gw;
  for (uint32_t t = blk * blockDim.x + threadIdx.x; t < pl.n_terms; t += bpg * blockDim.x) {
    const uint32_t p = pidx[t];
    const bool fixed = p < pl.n_fixed;
    const uint32_t c = fixed ? pl.cA : pl.cB, W = fixed ? pl.WA : pl.WB;
    if (w_lo >= W) continue;
    const uint32_t w_hi = w_lo + gw < W ? w_lo + gw : W;
    uint32_t s[8];
    const uint4* sp = reinterpret_cast<const uint4*>(scalars + 8ull * t);
    // several passes over the scalars: streaming loads, so that they do not push the group's slice of `sorted` out of L2
    uint4 a = n_groups > 1 ? __ldcs(sp) : __ldg(sp), b = n_groups > 1 ? __ldcs(sp + 1) : __ldg(sp + 1);
    s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
    const uint32_t job = msm_find_job(job_start, pl.n_jobs, pl.uniform, t);
    const uint32_t half = 1u << (c - 1);
    const uint32_t keybase = fixed ? job * pl.NBA : pl.keysA + job * pl.WB * pl.NBB;
    const uint32_t refbase = fixed ? p : pl.wtab_points + (p - pl.n_fixed);
    const uint32_t flip = (fixed && (pl.order[0] | pl.order[7])) ? msm_fold_half(pl, s) : 0u;
    uint32_t carry = 0;
    for (uint32_t w = 0; w < w_lo; w++) carry = (msm_bits(s, w * c, c) + carry) > half;
    for (uint32_t w = w_lo; w < w_hi; w++) {
      uint32_t raw = msm_bits(s, w * c, c) + carry;
      uint32_t neg = raw > half;
      uint32_t mag = neg ? (1u << c) - raw : raw;
      carry = neg;
      if (mag) {
        uint32_t key = keybase + (fixed ? 0u : w * pl.NBB) + (mag - 1);
        if (MODE == 0) atomicAdd(&counts_or_cursor[key], 1u);
        else sorted[atomicAdd(&counts_or_cursor[key], 1u)] = (refbase + (fixed ? w * pl.n_fixed : 0u)) | ((neg ^ flip) << 31);
      }
    }
  }
}

// ---- per-job counting sort in shared memory (warp-cooperative bucket sorting) ---------------------------------
// One block per job, for jobs whose key space (NBA + WB * NBB counters) fits in shared memory: histogram with shared
// atomics, block scan, scatter with shared cursors. Only the sorted references and the per-key (offset, count) leave
// the SM; the global histogram, its memset and the three scan kernels are not needed on this path. Job j owns the
// slice [ebase_j, ebase_j + cap_j) of `sorted`, cap_j = sort_mult * K_j * WA + dyn_per_job * max(WB - WA, 0).
constexpr int SJ_THREADS = 512;
// point index first: terms of a segment this pass leaves out (window count 0) never load their scalar
__device__ __forceinline__ bool msm_job_digits(const MsmPlan& pl, const uint32_t* __restrict__ scalars, const uint32_t* __restrict__ pidx,
                                               uint32_t t, uint32_t* s, uint32_t& p) {
  p = pidx[t];
  if ((p < pl.n_fixed ? pl.WA : pl.WB) == 0) return false;
  const uint4* sp = reinterpret_cast<const uint4*>(scalars + 8ull * t);
  uint4 a = __ldg(sp), b = __ldg(sp + 1);
  s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
  return true;
}
// Four consecutive window digits of one term: shared-memory keys (0xffffffff = no entry) and sorted references. The kernel
// issues the four shared atomics back to back and only then uses their results: a warp otherwise spends most of its time
// waiting for ONE atomic's return at a time (short-scoreboard stalls were 56 % of the issue cycles, r01g capture).
__device__ __forceinline__ void msm_digits4(const MsmPlan& pl, const uint32_t* s, bool fixed, uint32_t c, uint32_t W, uint32_t w0,
                                            uint32_t refbase, uint32_t flip, uint32_t& carry, uint32_t* key, uint32_t* ref) {
  const uint32_t half = 1u << (c - 1);
#pragma unroll
  for (int u = 0; u < 4; u++) {
    const uint32_t w = w0 + u;
    key[u] = 0xffffffffu;
    ref[u] = 0;
    if (w < W) {
      const uint32_t raw = msm_bits(s, w * c, c) + carry;
      const uint32_t neg = raw > half;
      const uint32_t mag = neg ? (1u << c) - raw : raw;
      carry = neg;
      if (mag) {
        key[u] = fixed ? mag - 1 : pl.NBA + w * pl.NBB + mag - 1;
        ref[u] = (fixed ? refbase + w * pl.n_fixed : refbase) | ((neg ^ flip) << 31);
      }
    }
  }
}

template <int DUMMY>
__global__ void __launch_bounds__(SJ_THREADS)
msm_sort_job_kernel(MsmPlan pl, uint32_t dyn_per_job, const uint32_t* __restrict__ scalars, const uint32_t* __restrict__ pidx,
                    const uint32_t* __restrict__ job_start, uint32_t* __restrict__ offsets, uint32_t* __restrict__ counts,
                    uint32_t* __restrict__ sorted) {
  extern __shared__ uint32_t sm[];  // cnt[KPJ] (becomes the cursor)
  __shared__ uint32_t warp_sums[SJ_THREADS / 32];
  const uint32_t job = blockIdx.x, tid = threadIdx.x;
  const uint32_t KPJ = pl.NBA + pl.WB * pl.NBB;
  const uint32_t t0 = pl.uniform ? job * pl.uniform : job_start[job];
  const uint32_t t1 = pl.uniform ? t0 + pl.uniform : job_start[job + 1];
  const uint32_t extraW = pl.WB > pl.WA ? pl.WB - pl.WA : 0;
  const uint32_t ebase = t0 * pl.WA * pl.sort_mult + job * dyn_per_job * extraW;
  const uint32_t K = t1 - t0;
  const uint32_t levels = K <= pl.k3 ? 3u : K <= pl.k2 ? 2u : 1u, sub = levels == 3 ? pl.sub3 : pl.sub2;
  const bool fold = (pl.order[0] | pl.order[7]) != 0;
  for (uint32_t k = tid; k < KPJ; k += SJ_THREADS) sm[k] = 0;
  __syncthreads();
  // pass 1: histogram
  for (uint32_t t = t0 + tid; t < t1; t += SJ_THREADS) {
    uint32_t s[8], p;
    if (!msm_job_digits(pl, scalars, pidx, t, s, p)) continue;
    const bool fixed = p < pl.n_fixed;
    const uint32_t c = fixed ? pl.cA : pl.cB, W = fixed ? pl.WA : pl.WB, half = 1u << (c - 1);
    if (fixed && fold) msm_fold_half(pl, s);
    uint32_t carry = 0;
    if (levels == 1 || !fixed) {
      for (uint32_t w0 = 0; w0 < W; w0 += 4) {
        uint32_t key[4], ref[4];
        msm_digits4(pl, s, fixed, c, W, w0, 0u, 0u, carry, key, ref);
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (key[u] != 0xffffffffu) atomicAdd(&sm[key[u]], 1u);
      }
      continue;
    }
    for (uint32_t w = 0; w < W; w++) {
      uint32_t raw = msm_bits(s, w * c, c) + carry;
      uint32_t neg = raw > half;
      uint32_t mag = neg ? (1u << c) - raw : raw;
      carry = neg;
      if (!mag) continue;
      if (fixed) msm_split_digit(mag, 0u, levels, sub, [&](uint32_t k, uint32_t) { atomicAdd(&sm[k], 1u); });
      else atomicAdd(&sm[pl.NBA + w * pl.NBB + mag - 1], 1u);
    }
  }
  __syncthreads();
  // exclusive scan of cnt[0..KPJ): contiguous chunk per thread
  const uint32_t ch = (KPJ + SJ_THREADS - 1) / SJ_THREADS;
  const uint32_t k0 = tid * ch, k1 = k0 + ch < KPJ ? k0 + ch : KPJ;
  uint32_t mysum = 0;
  for (uint32_t k = k0; k < k1; k++) mysum += sm[k];
  uint32_t x = mysum;
  const uint32_t lane = tid & 31, wid = tid >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) warp_sums[wid] = x;
  __syncthreads();
  if (wid == 0) {
    uint32_t v = lane < SJ_THREADS / 32 ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < SJ_THREADS / 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += y;
    }
    if (lane < SJ_THREADS / 32) warp_sums[lane] = v;
  }
  __syncthreads();
  uint32_t run = (wid ? warp_sums[wid - 1] : 0) + x - mysum;
  const uint32_t gA = job * pl.NBA, gB = pl.keysA + job * pl.WB * pl.NBB;
  for (uint32_t k = k0; k < k1; k++) {
    const uint32_t cnt = sm[k];
    const uint32_t gkey = k < pl.NBA ? gA + k : gB + (k - pl.NBA);
    offsets[gkey] = ebase + run;
    counts[gkey] = cnt;
    sm[k] = run;  // cursor
    run += cnt;
  }
  __syncthreads();
  // pass 2: scatter
  for (uint32_t t = t0 + tid; t < t1; t += SJ_THREADS) {
    uint32_t s[8], p;
    if (!msm_job_digits(pl, scalars, pidx, t, s, p)) continue;
    const bool fixed = p < pl.n_fixed;
    const uint32_t c = fixed ? pl.cA : pl.cB, W = fixed ? pl.WA : pl.WB, half = 1u << (c - 1);
    const uint32_t refbase = fixed ? p : pl.wtab_points + (p - pl.n_fixed);
    const uint32_t flip = (fixed && fold) ? msm_fold_half(pl, s) : 0u;
    uint32_t carry = 0;
    if (levels == 1 || !fixed) {
      for (uint32_t w0 = 0; w0 < W; w0 += 4) {
        uint32_t key[4], ref[4], pos[4];
        msm_digits4(pl, s, fixed, c, W, w0, refbase, flip, carry, key, ref);
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (key[u] != 0xffffffffu) pos[u] = atomicAdd(&sm[key[u]], 1u);
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (key[u] != 0xffffffffu) sorted[ebase + pos[u]] = ref[u];
      }
      continue;
    }
    for (uint32_t w = 0; w < W; w++) {
      uint32_t raw = msm_bits(s, w * c, c) + carry;
      uint32_t neg = raw > half;
      uint32_t mag = neg ? (1u << c) - raw : raw;
      carry = neg;
      if (!mag) continue;
      if (fixed) {
        const uint32_t ref = refbase + w * pl.n_fixed;
        msm_split_digit(mag, neg ^ flip, levels, sub, [&](uint32_t k, uint32_t ng) {
          const uint32_t pos = atomicAdd(&sm[k], 1u);
          sorted[ebase + pos] = ref | (ng << 31);
        });
      } else {
        const uint32_t pos = atomicAdd(&sm[pl.NBA + w * pl.NBB + mag - 1], 1u);
        sorted[ebase + pos] = refbase | (neg << 31);
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

static __global__ void scan_tile_sums(const uint32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ tile_sums) {
  uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) if (base + i < n) s += in[base + i];
  uint32_t tot;
  block_exclusive_scan(s, &tot);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}
static __global__ void scan_tile_offsets(uint32_t* tile_sums, uint32_t n_tiles) {  // single block, in place -> exclusive
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
static __global__ void scan_apply(const uint32_t* __restrict__ in, uint32_t n, const uint32_t* __restrict__ tile_offsets,
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

// ---- keys by descending bucket size (counting sort on min(count, 255)) ---------------------------------
// bins[0..256): global bin totals (zeroed by the caller); after sizesort_bins_scan they hold the bin starts.
constexpr int SS_THREADS = 256;
constexpr int SS_ITEMS = 16;
constexpr int SS_TILE = SS_THREADS * SS_ITEMS;
__device__ __forceinline__ uint32_t ss_bin(uint32_t count) { return 255u - (count < 255u ? count : 255u); }

static __global__ void __launch_bounds__(SS_THREADS) sizesort_hist_kernel(const uint32_t* __restrict__ counts, uint32_t n, uint32_t* __restrict__ bins) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  uint32_t base = blockIdx.x * SS_TILE;
  for (uint32_t i = threadIdx.x; i < SS_TILE && base + i < n; i += SS_THREADS) atomicAdd(&h[ss_bin(counts[base + i])], 1u);
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(&bins[threadIdx.x], h[threadIdx.x]);
}
// one block: bin totals -> exclusive starts; tiers[0..2] = first perm index of the buckets with count <= warp_cap / <= acc_cap
// (32 <= warp_cap <= 254, 4 <= acc_cap < warp_cap: the group kernels give every lane an entry)
static __global__ void __launch_bounds__(256) sizesort_bins_scan_kernel(uint32_t* __restrict__ bins, uint32_t* __restrict__ tiers, uint32_t acc_cap,
                                                                        uint32_t warp_cap) {
  uint32_t tot;
  uint32_t ex = block_exclusive_scan(bins[threadIdx.x], &tot);
  bins[threadIdx.x] = ex;
  if (threadIdx.x == 0) tiers[0] = 0;
  if (threadIdx.x == 255 - warp_cap) tiers[1] = ex;
  if (threadIdx.x == 255 - acc_cap) tiers[2] = ex;
}
static __global__ void __launch_bounds__(SS_THREADS) sizesort_scatter_kernel(const uint32_t* __restrict__ counts, uint32_t n, uint32_t* __restrict__ bins,
                                                                      uint32_t* __restrict__ perm) {
  __shared__ uint32_t h[256], start[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  uint32_t base = blockIdx.x * SS_TILE;
  uint32_t local[SS_ITEMS], bin[SS_ITEMS];
#pragma unroll
  for (int k = 0; k < SS_ITEMS; k++) {
    uint32_t i = threadIdx.x + k * SS_THREADS;
    bin[k] = 0xffffffffu;
    if (base + i < n) { bin[k] = ss_bin(counts[base + i]); local[k] = atomicAdd(&h[bin[k]], 1u); }
  }
  __syncthreads();
  start[threadIdx.x] = h[threadIdx.x] ? atomicAdd(&bins[threadIdx.x], h[threadIdx.x]) : 0;  // reserve this tile's range of every bin
  __syncthreads();
#pragma unroll
  for (int k = 0; k < SS_ITEMS; k++) {
    uint32_t i = threadIdx.x + k * SS_THREADS;
    if (bin[k] != 0xffffffffu) perm[start[bin[k]] + local[k]] = base + i;
  }
}

// ---- HOT: bucket accumulation ---------------------------------------------------------------------------
// One thread per bucket, buckets taken in descending-size order. wtab: window tables of the public key
// (wtab_points entries), dyn: per-call points; both AoS affine, C::AFF_WORDS words per point, read with
// 128-bit loads. Empty buckets are not written (the reduce kernels test the counts). Buckets longer than
// acc_cap (the partial top window always produces some; adversarial scalars produce more) are left to
// msm_accumulate_group_kernel, which spreads each of them over 4 or 32 lanes. Throughput setting (many buckets: the chip is
// full of one-thread chains anyway): one thread up to ACC_CAP entries, four lanes up to 254, a warp beyond. Latency setting
// (few buckets — one small MSM, where a 96-entry chain IS the kernel's duration): ACC_CAP_LAT / WARP_CAP_LAT.
constexpr uint32_t ACC_CAP = 96, WARP_CAP = 254;
constexpr uint32_t ACC_CAP_LAT = 8, WARP_CAP_LAT = 64;
constexpr uint32_t LAT_MAX_KEYS = 262144;  // at most this many buckets in all: latency setting (one thread per bucket would leave most of the chip idle)
// Buckets longer than GIANT_SEG entries are cut into segments of GIANT_SEG, one warp per segment, and merged afterwards:
// skewed scalar sets (all equal, P/-P pairs, a top window of one or two bits) put a large share of ALL terms into a few
// buckets, which would otherwise serialise on one warp each.
constexpr uint32_t GIANT_SEG = 2048;

template <class C>
__device__ __forceinline__ const uint32_t* msm_point_ptr(uint32_t ref, const uint32_t* __restrict__ wtab, uint32_t wtab_points,
                                                         const uint32_t* __restrict__ dyn) {
  uint32_t idx = ref & 0x7fffffffu;
  return idx < wtab_points ? wtab + (size_t)idx * C::AFF_WORDS : dyn + (size_t)(idx - wtab_points) * C::AFF_WORDS;
}

// HYB: blocks whose (index mod 4) bit is set in `pattern` keep their buckets on the FP64 pipe (C::AccF, tables wtab_f / dyn_f
// in the store_table_point_f slot format), the others on the integer pipe. A block's four warps sit on the four
// sub-partitions of its SM, so every sub-partition runs both kinds of warp and both pipes at once.
template <class C, bool HYB>
__global__ void __launch_bounds__(128)
msm_accumulate_kernel(uint32_t n_keys, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ offsets,
                      const uint32_t* __restrict__ counts, const uint32_t* __restrict__ sorted, const uint32_t* __restrict__ wtab, uint32_t wtab_points,
                      const uint32_t* __restrict__ dyn, uint32_t* __restrict__ buckets, const uint32_t* __restrict__ wtab_f,
                      const uint32_t* __restrict__ dyn_f, uint32_t pattern, uint32_t acc_cap) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_keys) return;
  const uint32_t key = perm[t];
  const uint32_t beg = offsets[key], cnt = counts[key], end = beg + cnt;
  if (cnt == 0 || cnt > acc_cap) return;
  typename C::Ext out;
  uint32_t ref = __ldg(sorted + beg);
  if (HYB && ((pattern >> (blockIdx.x & 3)) & 1u)) {
    typename C::AccF acc;
    C::accf_init(acc, msm_point_ptr<C>(ref, wtab_f, wtab_points, dyn_f), ref >> 31);
    for (uint32_t e = beg + 1; e < end; e++) {
      ref = __ldg(sorted + e);
      C::accf_madd(acc, msm_point_ptr<C>(ref, wtab_f, wtab_points, dyn_f), ref >> 31);
    }
    C::accf_to_ext(out, acc);
  } else {
    C::init_from_ptr(out, msm_point_ptr<C>(ref, wtab, wtab_points, dyn), ref >> 31);
    for (uint32_t e = beg + 1; e < end; e++) {
      ref = __ldg(sorted + e);
      C::madd_ptr(out, msm_point_ptr<C>(ref, wtab, wtab_points, dyn), ref >> 31);
    }
  }
  store_words<C::EXT_WORDS>(buckets + (size_t)key * C::EXT_WORDS, out);
}

// Buckets longer than acc_cap (the first entries of `perm`): G lanes per bucket — lane-strided walk, then a log2(G)-level
// tree through shared memory. Tier 0 (count > warp_cap) runs with G = 32, tier 1 (acc_cap < count <= warp_cap) with G = 4.
template <class C, int G>
__global__ void __launch_bounds__(128)
msm_accumulate_group_kernel(const uint32_t* __restrict__ tiers, uint32_t tier, const uint32_t* __restrict__ perm,
                            const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ counts, const uint32_t* __restrict__ sorted,
                            const uint32_t* __restrict__ wtab, uint32_t wtab_points, const uint32_t* __restrict__ dyn,
                            uint32_t* __restrict__ buckets) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t tid = threadIdx.x, lane = tid % G;
  typename C::Ext* mine = reinterpret_cast<typename C::Ext*>(sh + tid * C::EXT_WORDS);
  const uint32_t lo = tiers[tier], hi = tiers[tier + 1];
  const uint32_t n_groups = gridDim.x * (128 / G);
  for (uint32_t t = lo + blockIdx.x * (128 / G) + tid / G; t < hi; t += n_groups) {
    const uint32_t key = perm[t];
    if (counts[key] > GIANT_SEG) continue;  // split over several warps by the msm_giant_* kernels (warp-uniform test)
    const uint32_t beg = offsets[key], end = beg + counts[key];  // count > cap >= G: every lane has an entry
    typename C::Ext acc;
    uint32_t e = beg + lane;
    uint32_t ref = __ldg(sorted + e);
    C::init_from_ptr(acc, msm_point_ptr<C>(ref, wtab, wtab_points, dyn), ref >> 31);
    for (e += G; e < end; e += G) {
      ref = __ldg(sorted + e);
      C::madd_ptr(acc, msm_point_ptr<C>(ref, wtab, wtab_points, dyn), ref >> 31);
    }
    *mine = acc;
    __syncwarp();
    for (uint32_t s = G / 2; s > 0; s >>= 1) {
      if (lane < s) {
        typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (tid + s) * C::EXT_WORDS);
        C::add(acc, o);
        *mine = acc;
      }
      __syncwarp();
    }
    if (lane == 0) store_words<C::EXT_WORDS>(buckets + (size_t)key * C::EXT_WORDS, acc);
    __syncwarp();
  }
}

// ---- giant buckets: enumerate (bucket, segment) work items, accumulate a segment per warp, merge per bucket -------------
// ctl[0] = number of work items, ctl[1] = number of giant buckets (both zeroed by the host before the launch)
static __global__ void msm_giant_enum_kernel(const uint32_t* __restrict__ tiers, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ counts,
                                             uint32_t* __restrict__ ctl, uint2* __restrict__ items, uint4* __restrict__ gkeys) {
  const uint32_t lo = tiers[0], hi = tiers[1];
  for (uint32_t t = lo + blockIdx.x * blockDim.x + threadIdx.x; t < hi; t += gridDim.x * blockDim.x) {
    const uint32_t key = perm[t], cnt = counts[key];
    if (cnt <= GIANT_SEG) continue;
    const uint32_t n_seg = (cnt + GIANT_SEG - 1) / GIANT_SEG;
    const uint32_t base = atomicAdd(&ctl[0], n_seg), g = atomicAdd(&ctl[1], 1u);
    gkeys[g] = make_uint4(key, base, n_seg, 0u);
    for (uint32_t sgm = 0; sgm < n_seg; sgm++) items[base + sgm] = make_uint2(key, sgm);
  }
}

// warp-wide sum of the lanes' points through shared memory; the total ends in lane 0's `acc`
template <class C>
__device__ __forceinline__ void msm_warp_tree(typename C::Ext& acc, uint32_t* sh, uint32_t tid, uint32_t lane) {
  typename C::Ext* mine = reinterpret_cast<typename C::Ext*>(sh + tid * C::EXT_WORDS);
  *mine = acc;
  __syncwarp();
  for (uint32_t s = 16; s > 0; s >>= 1) {
    if (lane < s) {
      typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (tid + s) * C::EXT_WORDS);
      C::add(acc, o);
      *mine = acc;
    }
    __syncwarp();
  }
}

template <class C>
__global__ void __launch_bounds__(128)
msm_giant_accumulate_kernel(const uint32_t* __restrict__ ctl, const uint2* __restrict__ items, const uint32_t* __restrict__ offsets,
                            const uint32_t* __restrict__ counts, const uint32_t* __restrict__ sorted, const uint32_t* __restrict__ wtab,
                            uint32_t wtab_points, const uint32_t* __restrict__ dyn, uint32_t* __restrict__ partial) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t tid = threadIdx.x, lane = tid & 31, n_items = ctl[0];
  for (uint32_t it = blockIdx.x * 4 + tid / 32; it < n_items; it += gridDim.x * 4) {
    const uint2 w = items[it];
    const uint32_t beg = offsets[w.x] + w.y * GIANT_SEG, stop = offsets[w.x] + counts[w.x];
    const uint32_t end = beg + GIANT_SEG < stop ? beg + GIANT_SEG : stop;
    typename C::Ext acc;
    uint32_t e = beg + lane;
    if (e < end) {  // the last segment of a bucket may be shorter than a warp
      uint32_t ref = __ldg(sorted + e);
      C::init_from_ptr(acc, msm_point_ptr<C>(ref, wtab, wtab_points, dyn), ref >> 31);
      for (e += 32; e < end; e += 32) {
        ref = __ldg(sorted + e);
        C::madd_ptr(acc, msm_point_ptr<C>(ref, wtab, wtab_points, dyn), ref >> 31);
      }
    } else {
      C::set_identity(acc);
    }
    msm_warp_tree<C>(acc, sh, tid, lane);
    if (lane == 0) store_words<C::EXT_WORDS>(partial + (size_t)it * C::EXT_WORDS, acc);
    __syncwarp();
  }
}

template <class C>
__global__ void __launch_bounds__(128)
msm_giant_merge_kernel(const uint32_t* __restrict__ ctl, const uint4* __restrict__ gkeys, const uint32_t* __restrict__ partial,
                       uint32_t* __restrict__ buckets) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t tid = threadIdx.x, lane = tid & 31, n_g = ctl[1];
  for (uint32_t g = blockIdx.x * 4 + tid / 32; g < n_g; g += gridDim.x * 4) {
    const uint4 k = gkeys[g];
    typename C::Ext acc, p;
    C::set_identity(acc);
    for (uint32_t sgm = lane; sgm < k.z; sgm += 32) {
      load_words<C::EXT_WORDS>(p, partial + (size_t)(k.y + sgm) * C::EXT_WORDS);
      C::add(acc, p);
    }
    msm_warp_tree<C>(acc, sh, tid, lane);
    if (lane == 0) store_words<C::EXT_WORDS>(buckets + (size_t)k.x * C::EXT_WORDS, acc);
    __syncwarp();
  }
}

// ---- bucket reduction: sum_b (b+1) * bucket[b] per bucket set ---------------------------------------------
// level 1 — thread per chunk of q consecutive buckets: S = sum of the chunk, T = sum_k (k+1) * bucket[lo+k]
// MINB: resident blocks per SM the register allocation is held to (1 = unconstrained: 139 registers, 3 blocks on Ed25519)
template <class C, int MINB>
__global__ void __launch_bounds__(128, MINB)
msm_reduce_chunks_kernel(MsmPlan pl, const uint32_t* __restrict__ counts, const uint32_t* __restrict__ buckets,
                         uint32_t* __restrict__ chunk_S, uint32_t* __restrict__ chunk_T) {
  uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= pl.n_chunks) return;
  uint32_t q, lo;
  if (id < pl.chunksA) { q = pl.qA; lo = id * q; }
  else { q = pl.qB; lo = pl.keysA + (id - pl.chunksA) * q; }
  typename C::Ext run, sum, b;
  bool have_run = false, have_sum = false;
  // sparse chunks (the split digits of small jobs leave one bucket per chunk, at weight |d| * 2^(i*sub)): S = B, T = (k+1) * B
  // by double-and-add instead of k+1 running-sum additions
  uint32_t n_live = 0, last = 0;
  for (uint32_t k = 0; k < q; k++)
    if (counts[lo + k]) { n_live++; last = k; }
  if (n_live == 1) {
    load_words<C::EXT_WORDS>(b, buckets + (size_t)(lo + last) * C::EXT_WORDS);
    store_words<C::EXT_WORDS>(chunk_S + (size_t)id * C::EXT_WORDS, b);
    const uint32_t wgt = last + 1;
    sum = b;
    for (int bit = 30 - (int)__clz(wgt); bit >= 0; bit--) {
      C::dbl(sum);
      if ((wgt >> bit) & 1u) C::add(sum, b);
    }
    store_words<C::EXT_WORDS>(chunk_T + (size_t)id * C::EXT_WORDS, sum);
    return;
  }
  for (int k = (int)q - 1; k >= 0 && n_live; k--) {
    uint32_t key = lo + k;
    if (counts[key]) {
      load_words<C::EXT_WORDS>(b, buckets + (size_t)key * C::EXT_WORDS);
      if (have_run) C::add(run, b); else { run = b; have_run = true; }
    }
    if (have_run) {
      if (have_sum) C::add(sum, run); else { sum = run; have_sum = true; }
    }
  }
  if (!have_run) { C::set_identity(run); C::set_identity(sum); }
  store_words<C::EXT_WORDS>(chunk_S + (size_t)id * C::EXT_WORDS, run);
  store_words<C::EXT_WORDS>(chunk_T + (size_t)id * C::EXT_WORDS, sum);
}

// level 1b — thread per group of `per` consecutive chunks of weight q: (S', T') of the group as one chunk of weight q * per,
//   S' = sum_k S_k,  T' = sum_k T_k + q * sum_k k * S_k.
// Run before the set reduction when there are few sets with many chunks each (one large MSM): the set kernel then has no
// serial walk left and the work of this level spreads over the whole chip.
template <class C>
__global__ void __launch_bounds__(128)
msm_reduce_level_kernel(uint32_t n_groups, uint32_t q, uint32_t per, const uint32_t* __restrict__ in_S, const uint32_t* __restrict__ in_T,
                        uint32_t* __restrict__ out_S, uint32_t* __restrict__ out_T) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_groups) return;
  const size_t c0 = (size_t)g * per;
  typename C::Ext S, T, acc, p;
  load_words<C::EXT_WORDS>(S, in_S + (c0 + per - 1) * C::EXT_WORDS);
  load_words<C::EXT_WORDS>(T, in_T + (c0 + per - 1) * C::EXT_WORDS);
  acc = S;  // sum_k k * S_k as running suffix sums: after the loop acc = sum_{t >= 1} sum_{u >= t} S_u
  for (int k = (int)per - 2; k >= 0; k--) {
    load_words<C::EXT_WORDS>(p, in_T + (c0 + k) * C::EXT_WORDS);
    C::add(T, p);
    load_words<C::EXT_WORDS>(p, in_S + (c0 + k) * C::EXT_WORDS);
    C::add(S, p);
    if (k > 0) C::add(acc, S);
  }
  C::dbl_n(acc, 31u - __clz(q));
  C::add(T, acc);
  store_words<C::EXT_WORDS>(out_S + (size_t)g * C::EXT_WORDS, S);
  store_words<C::EXT_WORDS>(out_T + (size_t)g * C::EXT_WORDS, T);
}

// level 2 — a group of `tps` threads per bucket set (128-thread blocks, 128 / tps sets per block); each thread owns
// `per` consecutive chunks, then the group does a suffix scan and a tree sum in shared memory:
//   set total = sum_t [ T'_t + (q * per) * t * S'_t ],   sum_t t * S'_t = sum_{t >= 1} (sum_{u >= t} S'_u)
// first_chunk/first_set select the segment (A: sets = jobs, B: sets = job * WB + window). q, per, tps powers of two.
template <class C>
__global__ void __launch_bounds__(128)
msm_reduce_sets_kernel(uint32_t n_sets, uint32_t first_chunk, uint32_t q, uint32_t per, uint32_t tps,
                       const uint32_t* __restrict__ chunk_S, const uint32_t* __restrict__ chunk_T, uint32_t* __restrict__ out) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t tid = threadIdx.x, g = tid / tps, lane = tid % tps;
  const uint32_t set = blockIdx.x * (128 / tps) + g;
  const bool live = set < n_sets;
  typename C::Ext S, T, p;
  C::set_identity(S);
  C::set_identity(T);
  if (live) {
    const size_t c0 = (size_t)first_chunk + ((size_t)set * tps + lane) * per;
    // local: S' = sum_k S_k, T' = sum_k T_k + q * sum_k k * S_k
    typename C::Ext acc;
    C::set_identity(acc);
    for (int k = (int)per - 1; k >= 0; k--) {
      load_words<C::EXT_WORDS>(p, chunk_T + (c0 + k) * C::EXT_WORDS);
      C::add(T, p);
      load_words<C::EXT_WORDS>(p, chunk_S + (c0 + k) * C::EXT_WORDS);
      if (k > 0) { C::add(S, p); C::add(acc, S); }
      else {
        if (per > 1) C::dbl_n(acc, 31u - __clz(q));
        if (per > 1) C::add(T, acc);
        C::add(S, p);
      }
    }
  }
  typename C::Ext* mine = reinterpret_cast<typename C::Ext*>(sh + tid * C::EXT_WORDS);
  // inclusive suffix scan of S' over the group
  *mine = S;
  __syncthreads();
  for (uint32_t off = 1; off < tps; off <<= 1) {
    const bool have = lane + off < tps;
    if (have) p = *reinterpret_cast<typename C::Ext*>(sh + (tid + off) * C::EXT_WORDS);
    __syncthreads();
    if (have) { C::add(S, p); *mine = S; }
    __syncthreads();
  }
  // V_t = T'_t + (q * per) * R_t for t >= 1
  if (lane >= 1) {
    C::dbl_n(S, 31u - __clz(q * per));
    C::add(T, S);
  }
  *mine = T;
  __syncthreads();
  for (uint32_t s = tps >> 1; s > 0; s >>= 1) {
    if (lane < s) {
      p = *reinterpret_cast<typename C::Ext*>(sh + (tid + s) * C::EXT_WORDS);
      C::add(T, p);
      *mine = T;
    }
    __syncthreads();
  }
  if (live && lane == 0) store_words<C::EXT_WORDS>(out + (size_t)set * C::EXT_WORDS, T);
}

// Horner over the generic windows, plus the fixed-base part: one job per thread, but only `lanes` threads of each
// 32-thread block are used. The chain is ~SCALAR_BITS dependent doublings; a warp instruction costs its sub-partition's
// multiply pipe the same whether 1 or 32 lanes are active, so few jobs per warp on many warps (spread over all 592
// sub-partitions) finish sooner than the same jobs packed 32 to a warp on a few SMs.
template <class C>
__global__ void __launch_bounds__(32)
msm_combine_kernel(MsmPlan pl, uint32_t lanes, const uint32_t* __restrict__ windows, const uint32_t* __restrict__ fixed_res, uint32_t* __restrict__ results) {
  if (threadIdx.x >= lanes) return;
  uint32_t job = blockIdx.x * lanes + threadIdx.x;
  if (job >= pl.n_jobs) return;
  typename C::Ext acc, p;
  if (pl.WB) {
    const uint32_t* base = windows + (size_t)job * pl.WB * C::EXT_WORDS;
    load_words<C::EXT_WORDS>(acc, base + (size_t)(pl.WB - 1) * C::EXT_WORDS);
    for (int w = (int)pl.WB - 2; w >= 0; w--) {
      C::dbl_n(acc, pl.cB);
      load_words<C::EXT_WORDS>(p, base + (size_t)w * C::EXT_WORDS);
      C::add(acc, p);
    }
    if (pl.NBA) {
      load_words<C::EXT_WORDS>(p, fixed_res + (size_t)job * C::EXT_WORDS);
      C::add(acc, p);
    }
  } else {
    load_words<C::EXT_WORDS>(acc, fixed_res + (size_t)job * C::EXT_WORDS);
  }
  store_words<C::EXT_WORDS>(results + (size_t)job * C::EXT_WORDS, acc);
}

// The same Horner pass for curves whose policy has the four-role point operations (C::HAS_COOP4: Ed25519). ONE warp already
// saturates the multiply pipe of its SM sub-partition whatever its number of active lanes (72 IMAD.WIDE at ~4.6 issue cycles
// each per field multiply), so the serial chain of ~SCALAR_BITS doublings cannot be shortened inside a warp — spreading the
// four independent multiplies of a doubling over four LANES was measured slower. Here they go to four WARPS of a
// block, i.e. to the four sub-partitions of an SM: warp r computes the r-th product of each of the two levels of a doubling /
// addition, lane l of every warp works on job 32 * blockIdx.x + l, and the operands cross warps through shared memory (two
// barriers per point operation). Role r keeps coordinate r of the accumulator (X, Y, Z, T) in registers; X and Y are also
// published in shared memory because three roles read them.
template <class C>
__global__ void __launch_bounds__(128)
msm_combine4_kernel(MsmPlan pl, const uint32_t* __restrict__ windows, const uint32_t* __restrict__ fixed_res, uint32_t* __restrict__ results) {
  typedef typename C::F F;
  __shared__ uint32_t sh_l1[4][8][32];  // level-1 products of the operation in flight
  __shared__ uint32_t sh_xy[2][8][32];  // X, Y of the accumulator
  const uint32_t lane = threadIdx.x & 31, role = threadIdx.x >> 5;
  const uint32_t job = blockIdx.x * 32 + lane;
  const bool live = job < pl.n_jobs;
  const uint32_t jj = live ? job : 0;  // idle lanes shadow job 0 (every thread has to reach the barriers)
  auto put = [&](uint32_t (*buf)[8][32], uint32_t slot, const F& v) {
#pragma unroll
    for (int i = 0; i < 8; i++) buf[slot][i][lane] = v.v[i];
  };
  auto get = [&](uint32_t (*buf)[8][32], uint32_t slot, F& v) {
#pragma unroll
    for (int i = 0; i < 8; i++) v.v[i] = buf[slot][i][lane];
  };
  // level 2 of both operations: E F G H from the four level-1 values, then this role's product
  auto finish = [&](F& mine, const F& e, const F& f, const F& g, const F& h) {
    if (role == 0) fe_mul(mine, e, f);
    else if (role == 1) fe_mul(mine, g, h);
    else if (role == 2) fe_mul(mine, f, g);
    else fe_mul(mine, e, h);
    if (role < 2) put(sh_xy, role, mine);
    __syncthreads();
  };
  // acc += q (q: extended point in global memory); add-2008-hwcd-3 with k = 2d
  auto add_point = [&](F& mine, const uint32_t* q) {
    F v, x, y, qx, qy;
    if (role < 2) {
      get(sh_xy, 0, x); get(sh_xy, 1, y);
      C::load_fe(qx, q); C::load_fe(qy, q + 8);
      F t0, t1;
      if (role == 0) { fe_sub(t0, y, x); fe_sub(t1, qy, qx); } else { fe_add(t0, y, x); fe_add(t1, qy, qx); }
      fe_mul(v, t0, t1);                       // A = (Y1-X1)(Y2-X2)   |   B = (Y1+X1)(Y2+X2)
    } else if (role == 2) {
      C::load_fe(qx, q + 16);
      fe_mul(v, mine, qx); fe_dbl(v, v);       // D = 2 Z1 Z2
    } else {
      C::load_fe(qx, q + 24);
      fe_mul(v, mine, qx); fe_mul(v, v, C::k_d2());  // C = 2d T1 T2
    }
    put(sh_l1, role, v);
    __syncthreads();
    F a, b, c, d, e, f, g, h;
    get(sh_l1, 0, a); get(sh_l1, 1, b); get(sh_l1, 3, c); get(sh_l1, 2, d);
    fe_sub(e, b, a); fe_sub(f, d, c); fe_add(g, d, c); fe_add(h, b, a);
    finish(mine, e, f, g, h);
  };
  auto dbl_point = [&](F& mine) {  // dbl-2008-hwcd, a = -1
    F v;
    if (role == 3) { F x, y; get(sh_xy, 0, x); get(sh_xy, 1, y); fe_add(v, x, y); fe_sqr(v, v); }  // (X+Y)^2
    else fe_sqr(v, mine);                                                                          // X^2 | Y^2 | Z^2
    put(sh_l1, role, v);
    __syncthreads();
    F a, b, c, e, f, g, h;
    get(sh_l1, 0, a); get(sh_l1, 1, b); get(sh_l1, 2, c); get(sh_l1, 3, e);
    fe_dbl(c, c);
    fe_sub(e, e, a); fe_sub(e, e, b);
    fe_sub(g, b, a);
    fe_sub(f, g, c);
    fe_add(h, a, b); fe_neg(h, h);
    finish(mine, e, f, g, h);
  };
  F mine;
  const uint32_t* base = pl.WB ? windows + (size_t)jj * pl.WB * C::EXT_WORDS : fixed_res + (size_t)jj * C::EXT_WORDS;
  C::load_fe(mine, base + (pl.WB ? (size_t)(pl.WB - 1) * C::EXT_WORDS : 0) + 8 * role);
  if (role < 2) put(sh_xy, role, mine);
  __syncthreads();
  if (pl.WB) {
    for (int w = (int)pl.WB - 2; w >= 0; w--) {
      for (uint32_t k = 0; k < pl.cB; k++) dbl_point(mine);
      add_point(mine, base + (size_t)w * C::EXT_WORDS);
    }
    if (pl.NBA) add_point(mine, fixed_res + (size_t)jj * C::EXT_WORDS);
  }
  if (live) {
    uint4* o = reinterpret_cast<uint4*>(results + (size_t)job * C::EXT_WORDS + 8 * role);
    o[0] = make_uint4(mine.v[0], mine.v[1], mine.v[2], mine.v[3]);
    o[1] = make_uint4(mine.v[4], mine.v[5], mine.v[6], mine.v[7]);
  }
}

// thread per job: results[job] += part[job] (joins the two segments of Msm::run)
template <class C>
__global__ void __launch_bounds__(128) msm_join_kernel(uint32_t n_jobs, uint32_t* __restrict__ results, const uint32_t* __restrict__ part) {
  uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= n_jobs) return;
  typename C::Ext a, b;
  load_words<C::EXT_WORDS>(a, results + (size_t)job * C::EXT_WORDS);
  load_words<C::EXT_WORDS>(b, part + (size_t)job * C::EXT_WORDS);
  C::add(a, b);
  store_words<C::EXT_WORDS>(results + (size_t)job * C::EXT_WORDS, a);
}

// ---- window tables of a public key: wtab[w * n + i] = 2^(c*w) * P_i, ------------------------------------------
// One thread per point: W - 1 runs of c doublings, then ALL W - 1 normalisations with ONE field inversion (Montgomery's trick
// over the windows of the point: prefix products of the denominators forward, one Fermat chain, inverses peeled off backward).
// The version before did a full inversion per window per point: 22 x 2050 chains of ~265 multiplications for a 1024-generator
// key and 16 per point for bpp_points_precompute. `ext` is scratch for the W - 1 extended points of each thread
// (cnt * (W - 1) * C::EXT_WORDS words; large point sets are built slice by slice).
constexpr uint32_t WTAB_MAX_W = 72;  // cA >= 4 -> W <= 65 for a 256-bit order
template <class C>
__global__ void __launch_bounds__(64) wtab_build_kernel(uint32_t n, uint32_t i0, uint32_t cnt, uint32_t c, uint32_t W, const uint32_t* __restrict__ aff,
                                                        uint32_t* __restrict__ wtab, uint32_t* __restrict__ ext) {
  typedef typename C::F F;
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;  // index inside the slice [i0, i0 + cnt) of the n points
  if (j >= cnt) return;
  const uint32_t i = i0 + j;
  typename C::Aff q;
  typename C::Ext p;
  load_words<C::AFF_WORDS>(q, aff + (size_t)i * C::AFF_WORDS);
  store_words<C::AFF_WORDS>(wtab + (size_t)i * C::AFF_WORDS, q);
  C::aff_to_ext(p, q);
  F prefix[WTAB_MAX_W];  // prefix[w] = product of the denominators of windows 1 .. w
  F run = C::fone();
  for (uint32_t w = 1; w < W; w++) {
    C::dbl_n(p, c);
    store_words<C::EXT_WORDS>(ext + ((size_t)(w - 1) * cnt + j) * C::EXT_WORDS, p);
    run = C::fmul(run, C::aff_den(p));
    prefix[w] = run;
  }
  if (W < 2) return;
  F inv = C::finv(run);  // 1 / (den_1 ... den_{W-1})
  for (uint32_t w = W - 1; w >= 1; w--) {
    load_words<C::EXT_WORDS>(p, ext + ((size_t)(w - 1) * cnt + j) * C::EXT_WORDS);
    const F den_inv = w > 1 ? C::fmul(inv, prefix[w - 1]) : inv;
    C::to_aff_inv(q, p, den_inv);
    store_words<C::AFF_WORDS>(wtab + ((size_t)w * n + i) * C::AFF_WORDS, q);
    inv = C::fmul(inv, C::aff_den(p));
  }
}

// affine points -> the slot format of the FP64 accumulators (curves with HAS_F64; hybrid accumulation kernel)
template <class C>
__global__ void __launch_bounds__(128) msm_table_format_kernel(uint32_t n, const uint32_t* __restrict__ aff, uint32_t* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  typename C::Aff q;
  load_words<C::AFF_WORDS>(q, aff + (size_t)i * C::AFF_WORDS);
  C::store_table_point_f(out + (size_t)i * C::AFF_WORDS, q);
}

// ---- K5: comb tables for the two commitment bases: comb[(p * CW + w) * 128 + d - 1] = d * 2^(8w) * P_p -------
// (built by comb_build_kernel in proto_misc.cuh). out[j] = a_j * P_0 + b_j * P_1 (extended); scalars = (a_j, b_j)
// pairs of canonical 256-bit values; at most 2 * CW table additions per commitment, no buckets, no doublings.
constexpr int COMB_C = 8, COMB_D = 128;
template <class C>
__global__ void __launch_bounds__(128) comb_eval_kernel(uint32_t n, uint32_t CW, const uint32_t* __restrict__ scalars,
                                                        const uint32_t* __restrict__ comb, uint32_t* __restrict__ out) {
  uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  typename C::Ext acc;
  C::set_identity(acc);
  for (uint32_t pi = 0; pi < 2; pi++) {
    uint32_t s[8];
    for (int i = 0; i < 8; i++) s[i] = scalars[(size_t)(2 * j + pi) * 8 + i];
    uint32_t carry = 0;
    for (uint32_t w = 0; w < CW; w++) {
      uint32_t raw = (w < 32 ? (s[w >> 2] >> (8 * (w & 3))) & 0xffu : 0u) + carry;
      uint32_t neg = raw > 128u;
      uint32_t mag = neg ? 256u - raw : raw;
      carry = neg;
      if (mag) C::madd_ptr(acc, comb + ((size_t)(pi * CW + w) * COMB_D + mag - 1) * C::AFF_WORDS, neg);
    }
  }
  store_words<C::EXT_WORDS>(out + (size_t)j * C::EXT_WORDS, acc);
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

}  // namespace bpp

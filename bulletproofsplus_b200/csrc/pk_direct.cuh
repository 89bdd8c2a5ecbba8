// Direct fixed-base tables for the public key's generators (bpp_pk_precompute) — the B200-first form of the fixed-base
// segment of the MSM that replaces MulVec::calculate (reference src/bls12_381/building_block/mulvec.rs:20-33) for every
// caller on the path: with 180 GB of HBM per GPU the key can hold EVERY signed-digit multiple of every window point,
//     dtab[dtab_offset(w, i) + j - 1] = j * 2^(c w) * P_i     (affine Niels, 128 B; j = 1 .. 2^(c-1); last window: 1 .. 2^c),
// so a term costs one table addition per window with no buckets at all: no per-job sort, no size sort, no bucket
// reduction (together 2.4 ms of the 8.4 ms verify step with the bucket path) and wider windows than buckets allow
// (c = 14: 19 additions per term instead of 21, 40 GB for a 1024-generator key; the bucket path stops at c = 12 because
// every job pays 2^(c-1) bucket reductions). The price is memory traffic: every addition reads its own 128-byte line at a
// random place of the table (10 GB per 2048 proofs, ~2 TB/s), which HBM3e delivers beside the integer pipe's work.
// Ristretto255 only (the tuned primary curve); the other curves and keys without these tables keep the bucket path.
#pragma once
#include "msm.cuh"
#include "scalar.cuh"

namespace bpp {

// Scalars are folded below order/2 < 2^(B), B = SCALAR_BITS - 1, before they are recoded, so W = ceil(B / c) windows hold every
// bit. Windows 0 .. W-2 take signed digits (|d| <= 2^(c-1), carry into the next window); the LAST window takes its digit
// unsigned (raw bits + carry <= 2^c), so nothing carries out of it and no (W+1)-th window exists: its regions hold 2^c
// entries instead of 2^(c-1). Entries per point: (W + 1) * 2^(c-1).
__host__ __device__ inline uint32_t dtab_windows(uint32_t scalar_bits, uint32_t c) { return (scalar_bits - 1 + c - 1) / c; }
__host__ __device__ inline size_t dtab_entries(uint32_t n, uint32_t c, uint32_t W) { return ((size_t)(W + 1) * n) << (c - 1); }
__host__ __device__ inline size_t dtab_offset(uint32_t w, uint32_t i, uint32_t n, uint32_t c, uint32_t W) {
  return w + 1 < W ? ((size_t)w * n + i) << (c - 1) : (((size_t)(W - 1) * n) << (c - 1)) + ((size_t)i << c);
}

// ---- table build -------------------------------------------------------------------------------------------------------
// step 1 (wtab_build_kernel, msm.cuh): window bases 2^(c w) P_i -> base[w * n + i]
// step 2: thread per (w, i): running sums j * base as EXTENDED points written into the table slots (an extended point and an
//         affine table point are both 128 B on this curve)
template <class C>
__global__ void __launch_bounds__(64) dtab_fill_kernel(uint32_t n, uint32_t c, uint32_t W, const uint32_t* __restrict__ base, uint32_t* __restrict__ dtab) {
  static_assert(C::EXT_WORDS == C::AFF_WORDS, "the fill pass stores extended points in the table slots");
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;  // region = w * n + i
  if (r >= n * W) return;
  const uint32_t w = r / n, i = r - w * n;
  const uint32_t half = w + 1 < W ? 1u << (c - 1) : 1u << c;
  const uint32_t* b = base + (size_t)r * C::AFF_WORDS;
  uint32_t* out = dtab + dtab_offset(w, i, n, c, W) * C::AFF_WORDS;
  typename C::Ext acc;
  C::init_from_ptr(acc, b, 0);
  for (uint32_t j = 0; j < half; j++) {
    store_words<C::EXT_WORDS>(out + (size_t)j * C::EXT_WORDS, acc);
    C::madd_ptr(acc, b, 0);
  }
}
// step 3: thread per run of DTAB_NORM entries: extended -> affine in place with one inversion per run (Montgomery's trick)
constexpr uint32_t DTAB_NORM = 32;
template <class C>
__global__ void __launch_bounds__(128) dtab_normalise_kernel(uint64_t n_entries, uint32_t* __restrict__ dtab) {
  typedef typename C::F F;
  const uint64_t run = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t e0 = run * DTAB_NORM;
  if (e0 >= n_entries) return;
  const uint32_t cnt = n_entries - e0 < DTAB_NORM ? (uint32_t)(n_entries - e0) : DTAB_NORM;
  uint32_t* p0 = dtab + e0 * C::EXT_WORDS;
  F prefix[DTAB_NORM];
  F acc = C::fone();
  typename C::Ext p;
  for (uint32_t k = 0; k < cnt; k++) {
    load_words<C::EXT_WORDS>(p, p0 + (size_t)k * C::EXT_WORDS);
    acc = C::fmul(acc, C::aff_den(p));
    prefix[k] = acc;
  }
  F inv = C::finv(acc);
  typename C::Aff q;
  for (int k = (int)cnt - 1; k >= 0; k--) {
    load_words<C::EXT_WORDS>(p, p0 + (size_t)k * C::EXT_WORDS);
    const F den_inv = k > 0 ? C::fmul(inv, prefix[k - 1]) : inv;
    C::to_aff_inv(q, p, den_inv);
    inv = C::fmul(inv, C::aff_den(p));
    store_words<C::AFF_WORDS>(p0 + (size_t)k * C::AFF_WORDS, q);
  }
}

// ---- HOT: accumulation straight from the tables ------------------------------------------------------------------------
// A job is cut into S slices of TPJ threads (TPJ a power of two <= 128, 128 / TPJ slices per block): S * TPJ "lanes" per job.
//   phase 1  whole rounds: lane l takes terms l, l + S*TPJ, ... and adds one table point per non-zero digit of each (half-order
//            folding + signed c-bit digits as in msm_digits_kernel; the next digit's line is prefetched into L2 meanwhile)
//   phase 2  the K mod (S*TPJ) leftover terms are dealt out as (term, window) PAIRS, so that no lane gets a whole extra term
//            (1026 = 8 * 128 + 2 terms per prover job would otherwise make two lanes work 9/8 as long as the others)
// then the TPJ partial sums of a slice are added in a shared-memory tree -> partial[job * S + slice] (extended points; S == 1:
// the job's result). Slices exist for wave quantisation: 2048 one-block jobs are 3.46 waves of the 592 resident blocks of a
// B200 (86 % of four), 8192 quarter-job blocks are 13.8. Terms whose point is not in the key (index >= n_fixed: per-call
// points) are left to the generic segment.
// Exponents of the generator terms of a verification MSM computed where they are used (proto_verify.cuh: exponent_k =
// Tmul_k * x + Tadd_k, x one of the proof's r', s', delta', r's'): term k of job p is g (k = 3), h (k = 4), G_i (k = K0 + i) or
// H_i (k = K0 + N + i); every other term is a per-call point and belongs to the generic segment. enabled == 0: plain
// (scalar, point index) arrays as for every other caller.
struct DirectVerifySrc {
  uint32_t enabled, K0, N, stride_T;       // stride_T in words: 0 = one challenge set for all proofs
  const uint32_t* Tmul;                    // Montgomery form
  const uint32_t* Tadd;                    // canonical
  const uint32_t* xs;                      // [proof][r', s', delta', r' s'] canonical
};
// returns false when term k is not a generator term; s = canonical exponent, p = table index (g, h, G.., H..)
static __device__ __noinline__ bool direct_verify_term(const DirectVerifySrc& v, uint32_t job, uint32_t k, uint32_t* s, uint32_t& p) {
  typedef Sc<ModEll> S;
  uint32_t sel;
  if (k == 3) { p = 0; sel = 3; }
  else if (k == 4) { p = 1; sel = 2; }
  else if (k >= v.K0 && k < v.K0 + v.N) { p = 2 + (k - v.K0); sel = 0; }
  else if (k >= v.K0 + v.N && k < v.K0 + 2 * v.N) { p = 2 + v.N + (k - v.K0 - v.N); sel = 1; }
  else return false;
  const uint32_t* tm = v.Tmul + (size_t)job * v.stride_T + 8 * k;
  const uint32_t* ta = v.Tadd + (size_t)job * v.stride_T + 8 * k;
  S val = S::raw(tm) * S::raw(v.xs + (size_t)job * 32 + 8 * sel);  // Montgomery x canonical = canonical
  if (k != 4) val = val + S::raw(ta);                              // h carries no constant part (wip.rs:292)
#pragma unroll
  for (int i = 0; i < 8; i++) s[i] = val.v[i];
  return true;
}

__device__ __forceinline__ void direct_load_scalar(const uint32_t* __restrict__ scalars, uint32_t t, uint32_t* s) {
  const uint4* sp = reinterpret_cast<const uint4*>(scalars + 8ull * t);
  const uint4 a = __ldg(sp), b = __ldg(sp + 1);
  s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
}

// FUSED: the exponents come from DirectVerifySrc (a second instantiation, so that the plain one keeps its register allocation)
template <class C, bool FUSED>
__global__ void __launch_bounds__(128, 4)
direct_accumulate_kernel(MsmPlan pl, uint32_t tpj, uint32_t S, const uint32_t* __restrict__ scalars, const uint32_t* __restrict__ pidx,
                         const uint32_t* __restrict__ job_start, const uint32_t* __restrict__ dtab, uint32_t* __restrict__ partial,
                         DirectVerifySrc vsrc) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t tid = threadIdx.x, lane = tid & (tpj - 1);
  const uint32_t unit = blockIdx.x * (128 / tpj) + tid / tpj;  // (job, slice)
  const uint32_t job = unit / S, slice = unit - job * S;
  const bool live = job < pl.n_jobs;
  const uint32_t c = pl.cA, W = pl.WA, half = 1u << (c - 1), n = pl.n_fixed;
  typename C::Ext acc;
  C::set_identity(acc);  // complete formulas: adding to the identity is an ordinary addition
  if (live) {
    const uint32_t t0 = pl.uniform ? job * pl.uniform : job_start[job];
    const uint32_t t1 = pl.uniform ? t0 + pl.uniform : job_start[job + 1];
    const uint32_t GL = tpj * S, gl = slice * tpj + lane, K = t1 - t0, full = K / GL;
    // Digit loop: the scalar is SHIFTED down by c bits per window (eight funnel shifts on registers; indexing s[] by a runtime
    // window number put it in local memory), the table region of (window, point) is a running pointer advanced by one window
    // stride, and the accumulator starts as the identity so that the loop holds one addition and no first-entry case. Measured
    // (profiles/r02_sass_direct_loop.txt): 1342 -> 1109 instructions per addition bought under 1 % — the loop is bound by the
    // IMAD.WIDE issue rate, not by instruction count — while a variant with three inlined additions lost 13 % to the
    // instruction cache. Keep ONE madd_ptr per phase.
    const size_t wstride = ((size_t)n << (c - 1)) * C::AFF_WORDS;  // words from window w to w + 1 (signed windows)
    const uint32_t cmask = (1u << c) - 1u;
    for (uint32_t r = 0; r < full; r++) {  // phase 1
      const uint32_t t = t0 + r * GL + gl;
      uint32_t s[8], p;
      if (FUSED) {
        if (!direct_verify_term(vsrc, job, t - t0, s, p)) continue;
      } else {
        p = pidx[t];
        if (p >= n) continue;
        direct_load_scalar(scalars, t, s);
      }
      const uint32_t flip = msm_fold_half(pl, s);
      const uint32_t* wp = dtab + ((size_t)p << (c - 1)) * C::AFF_WORDS;                    // region of (window 0, point p)
      const uint32_t* wlast = dtab + dtab_offset(W - 1, p, n, c, W) * C::AFF_WORDS;         // region of the last window (2^c entries)
      uint32_t carry = 0, cur_neg = 0;
      const uint32_t* cur = nullptr;  // the addition in flight: (cur) is added while the next digit is recoded and prefetched
      // ONE call site of madd_ptr for the whole term: a version with the last window peeled out of the loop had three inlined
      // copies (51 KB of hot code) and ran 13 % slower although its loop was 260 instructions shorter — instruction cache.
      for (uint32_t w = 0; w <= W; w++) {
        const uint32_t* nxt = nullptr;
        uint32_t nxt_neg = 0;
        if (w < W) {
          const bool last = w + 1 == W;  // taken unsigned: what is left of the scalar (< 2^c; masked: a non-canonical scalar stays in range)
          const uint32_t raw = (s[0] & cmask) + carry;
#pragma unroll
          for (int q = 0; q < 7; q++) s[q] = __funnelshift_r(s[q], s[q + 1], c);
          s[7] >>= c;
          const uint32_t neg = !last && raw > half;
          const uint32_t mag = neg ? (1u << c) - raw : raw;
          carry = neg;
          if (mag) {
            nxt = (last ? wlast : wp) + (size_t)(mag - 1) * C::AFF_WORDS;
            nxt_neg = neg ^ flip;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nxt));
          }
          wp += wstride;
        }
        if (cur) C::madd_ptr(acc, cur, cur_neg);
        cur = nxt;
        cur_neg = nxt_neg;
      }
    }
    const uint32_t tb = t0 + full * GL, R = K - full * GL;  // phase 2
    for (uint32_t idx = gl; idx < R * W; idx += GL) {
      const uint32_t t = tb + idx / W, w = idx % W;
      uint32_t s[8], p;
      if (FUSED) {
        if (!direct_verify_term(vsrc, job, t - t0, s, p)) continue;
      } else {
        p = pidx[t];
        if (p >= n) continue;
        direct_load_scalar(scalars, t, s);
      }
      const uint32_t flip = msm_fold_half(pl, s);
      uint32_t carry = 0;
      for (uint32_t v = 0; v < w; v++) carry = (msm_bits(s, v * c, c) + carry) > half;  // windows below w are all signed
      const uint32_t raw = msm_bits(s, w * c, c) + carry;
      const uint32_t neg = (w + 1 < W) && raw > half;
      const uint32_t mag = neg ? (1u << c) - raw : raw;
      if (mag) C::madd_ptr(acc, dtab + (dtab_offset(w, p, n, c, W) + (mag - 1)) * C::AFF_WORDS, neg ^ flip);
    }
  }
  typename C::Ext* mine = reinterpret_cast<typename C::Ext*>(sh + tid * C::EXT_WORDS);
  *mine = acc;
  __syncthreads();
  for (uint32_t sft = tpj >> 1; sft > 0; sft >>= 1) {
    if (lane < sft) {
      typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (tid + sft) * C::EXT_WORDS);
      C::add(acc, o);
      *mine = acc;
    }
    __syncthreads();
  }
  if (live && lane == 0) store_words<C::EXT_WORDS>(partial + (size_t)unit * C::EXT_WORDS, acc);
}

// thread per job: results[job] = sum of its S slice sums
template <class C>
__global__ void __launch_bounds__(128) direct_join_kernel(uint32_t n_jobs, uint32_t S, const uint32_t* __restrict__ partial, uint32_t* __restrict__ results) {
  const uint32_t job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= n_jobs) return;
  typename C::Ext a, b;
  load_words<C::EXT_WORDS>(a, partial + (size_t)job * S * C::EXT_WORDS);
  for (uint32_t k = 1; k < S; k++) {
    load_words<C::EXT_WORDS>(b, partial + ((size_t)job * S + k) * C::EXT_WORDS);
    C::add(a, b);
  }
  store_words<C::EXT_WORDS>(results + (size_t)job * C::EXT_WORDS, a);
}

// The same for many slices per job (few jobs spread over the whole chip): a warp per job, lane-strided partial sums, then a
// five-level tree through shared memory.
template <class C>
__global__ void __launch_bounds__(128) direct_join_tree_kernel(uint32_t n_jobs, uint32_t S, const uint32_t* __restrict__ partial,
                                                               uint32_t* __restrict__ results) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t tid = threadIdx.x, lane = tid & 31, job = blockIdx.x * 4 + (tid >> 5);
  typename C::Ext acc, b;
  C::set_identity(acc);
  if (job < n_jobs)
    for (uint32_t k = lane; k < S; k += 32) {
      load_words<C::EXT_WORDS>(b, partial + ((size_t)job * S + k) * C::EXT_WORDS);
      C::add(acc, b);
    }
  typename C::Ext* mine = reinterpret_cast<typename C::Ext*>(sh + tid * C::EXT_WORDS);
  *mine = acc;
  __syncthreads();
  for (uint32_t sft = 16; sft > 0; sft >>= 1) {
    if (lane < sft) {
      typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (tid + sft) * C::EXT_WORDS);
      C::add(acc, o);
      *mine = acc;
    }
    __syncthreads();
  }
  if (job < n_jobs && lane == 0) store_words<C::EXT_WORDS>(results + (size_t)job * C::EXT_WORDS, acc);
}

}  // namespace bpp

// Small boundary kernels: generator table (K6), commitment terms (K5), format checks, verdicts,
// proof assembly, and the integer-pipe microbenchmark.
#pragma once
#include "proto_prove.cuh"
#include "msm.cuh"
#include "fe25519_f64.cuh"

namespace bpp {

// k * P by MSB-first double-and-add over a 256-bit little-endian scalar
template <class C>
__device__ void scalar_mul_vartime(typename C::Ext& out, const typename C::Ext& P, const uint32_t* k) {
  typename C::Ext acc;
  C::set_identity(acc);
  bool started = false;
  for (int i = 7; i >= 0; i--) {
    if (!started && k[i] == 0) continue;
    for (int b = 31; b >= 0; b--) {
      if (started) C::dbl(acc);
      if ((k[i] >> b) & 1) {
        if (started) C::add(acc, P); else { acc = P; started = true; }
      }
    }
  }
  out = acc;
}

// K6 — PublicKey::new (publickey.rs:21-48): g, h = 2g, G_i = 3(i+1) g, H_i = 5(i+1) g; the multipliers are computed
// in i32 like the reference (`(i as i32 + 1) * 3`), negative values meaning modulus - |n| i.e. the negated point.
template <class C>
__global__ void pk_gen_kernel(uint32_t len, uint32_t* __restrict__ table) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 2 + 2 * len) return;
  int32_t mult;
  if (t == 0) mult = 1;
  else if (t == 1) mult = 2;
  else if (t < 2 + len) mult = (int32_t)((uint32_t)(t - 2 + 1) * 3u);
  else mult = (int32_t)((uint32_t)(t - 2 - len + 1) * 5u);
  bool neg = mult < 0;
  uint32_t k[8] = {neg ? (uint32_t)(-(int64_t)mult) : (uint32_t)mult, 0, 0, 0, 0, 0, 0, 0};
  typename C::Ext b, r;
  C::base(b);
  scalar_mul_vartime<C>(r, b, k);
  if (neg) C::neg(r);
  typename C::Aff q;
  C::to_aff(q, r);
  store_words<C::AFF_WORDS>(table + (size_t)t * C::AFF_WORDS, q);
}

// table[i] = mult[i] * g  (synthetic MSM inputs)
template <class C, class M>
__global__ void base_mul_kernel(uint32_t n, const uint32_t* __restrict__ mult, uint32_t* __restrict__ table, uint8_t* __restrict__ bad) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  uint32_t k[8];
  for (int i = 0; i < 8; i++) k[i] = mult[8 * (size_t)t + i];
  if (!Sc<M>::canonical_ok(k)) { bad[0] = 1; for (int i = 0; i < 8; i++) k[i] = 0; }
  typename C::Ext b, r;
  C::base(b);
  scalar_mul_vartime<C>(r, b, k);
  typename C::Aff q;
  C::to_aff(q, r);
  store_words<C::AFF_WORDS>(table + (size_t)t * C::AFF_WORDS, q);
}

// K5 terms — V = Fr(v as i32) * g + gamma * h  (range/prover.rs:36-39, publickey.rs:50-52)
// full_u64: V = Fr(v) * g + gamma * h with all 64 bits of v (bpp_commit_batch_u64) — what an n = 64 prover needs, since the
// prover itself uses every bit of v (range/mod.rs:260-277) and the reference's cast makes v >= 2^31 unprovable.
template <class M>
__global__ void commit_terms_kernel(uint32_t count, uint32_t full_u64, const uint64_t* __restrict__ v, const uint32_t* __restrict__ gamma,
                                    uint32_t* __restrict__ scalars, uint32_t* __restrict__ pidx, uint8_t* __restrict__ bad) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= count) return;
  sc_store<M>(scalars + 16 * (size_t)t, full_u64 ? Sc<M>::from_u64(v[t]) : Sc<M>::from_i32((int32_t)(uint32_t)v[t]));
  bool ok = Sc<M>::canonical_ok(gamma + 8 * (size_t)t);
  for (int i = 0; i < 8; i++) scalars[16 * (size_t)t + 8 + i] = ok ? gamma[8 * (size_t)t + i] : 0;
  if (!ok) bad[0] = 1;
  pidx[2 * (size_t)t] = 0;
  pidx[2 * (size_t)t + 1] = 1;
}

template <class M>
__global__ void scalars_check_kernel(uint32_t n, const uint32_t* __restrict__ s, uint8_t* __restrict__ bad) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n && !Sc<M>::canonical_ok(s + 8 * (size_t)t)) bad[0] = 1;
}
static __global__ void iota_kernel(uint32_t n, uint32_t period, uint32_t* __restrict__ out) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = period ? t % period : t;
}
static __global__ void any_zero_kernel(uint32_t n, const uint8_t* __restrict__ ok, uint8_t* __restrict__ bad) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n && !ok[t]) bad[0] = 1;
}

// standalone WIP prove inputs: canonical a, b -> Montgomery work arrays; gamma -> alpha
template <class M>
__global__ void wip_load_kernel(uint32_t N, const uint32_t* __restrict__ a, const uint32_t* __restrict__ b, const uint32_t* __restrict__ gamma,
                                uint32_t* __restrict__ a_out, uint32_t* __restrict__ b_out, uint32_t* __restrict__ alpha_out,
                                uint8_t* __restrict__ bad) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < N) {
    if (!Sc<M>::canonical_ok(a + 8 * (size_t)t) || !Sc<M>::canonical_ok(b + 8 * (size_t)t)) bad[0] = 1;
    st_mont<M>(a_out, t, sc_load<M>(a + 8 * (size_t)t));
    st_mont<M>(b_out, t, sc_load<M>(b + 8 * (size_t)t));
  }
  if (t == 0) {
    if (!Sc<M>::canonical_ok(gamma)) bad[0] = 1;
    st_mont<M>(alpha_out, 0, sc_load<M>(gamma));
  }
}

// verdict per proof: FORMAT if any of its points or scalars failed to decode, else identity test of the MSM
template <class C>
__global__ void verdict_kernel(uint32_t n_proofs, uint32_t pts_per_proof, uint32_t nV, const uint32_t* __restrict__ results,
                               const uint8_t* __restrict__ dyn_ok, const uint8_t* __restrict__ bad_scalar, uint8_t* __restrict__ verdicts) {
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_proofs) return;
  bool fmt = bad_scalar[p] != 0;
  for (uint32_t k = 0; k < pts_per_proof; k++) fmt = fmt || !dyn_ok[(size_t)p * pts_per_proof + k];
  for (uint32_t k = 0; k < nV; k++) fmt = fmt || !dyn_ok[(size_t)n_proofs * pts_per_proof + (size_t)p * nV + k];
  typename C::Ext r;
  load_words<C::EXT_WORDS>(r, results + (size_t)p * C::EXT_WORDS);
  verdicts[p] = fmt ? BPP_ERR_FORMAT : (C::is_identity(r) ? BPP_OK : BPP_ERR_VERIFICATION);
}

// random-linear-combination check: the one combined MSM result decides the whole batch. pass[0] = 1 and per-proof verdicts
// (FORMAT for proofs that did not decode, OK for the rest) if it is the identity; pass[0] = 0 otherwise (verdicts untouched).
template <class C>
__global__ void rlc_verdict_kernel(uint32_t n_proofs, const uint32_t* __restrict__ result, const uint8_t* __restrict__ fmt,
                                   uint8_t* __restrict__ verdicts, uint8_t* __restrict__ pass) {
  typename C::Ext r;
  load_words<C::EXT_WORDS>(r, result);
  const bool ok = C::is_identity(r);
  if (blockIdx.x == 0 && threadIdx.x == 0) pass[0] = ok ? 1 : 0;
  if (!ok) return;
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n_proofs; p += gridDim.x * blockDim.x)
    verdicts[p] = fmt[p] ? BPP_ERR_FORMAT : BPP_OK;
}

// proof assembly: job results (A, L.., R.., wipA, wipB per proof) -> encoded points inside each proof
template <class C>
__global__ void encode_into_proofs_kernel(uint32_t n_proofs, uint32_t jobs_per_proof, const uint32_t* __restrict__ results,
                                          const uint32_t* __restrict__ add_to_job0, uint8_t* __restrict__ proofs, size_t proof_stride) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_proofs * jobs_per_proof) return;
  uint32_t p = t / jobs_per_proof, k = t % jobs_per_proof;
  typename C::Ext r;
  load_words<C::EXT_WORDS>(r, results + (size_t)t * C::EXT_WORDS);
  if (k == 0 && add_to_job0) {
    typename C::Ext s;
    load_words<C::EXT_WORDS>(s, add_to_job0 + (size_t)p * C::EXT_WORDS);
    C::add(r, s);
  }
  C::encode(proofs + (size_t)p * proof_stride + (size_t)k * C::POINT_BYTES, r);
}

// gather the point bytes of every proof (first pts_per_proof points of each) into one contiguous array
static __global__ void gather_proof_points_kernel(uint32_t n_proofs, uint32_t bytes_per_proof, const uint8_t* __restrict__ proofs,
                                           size_t proof_stride, uint8_t* __restrict__ out) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)n_proofs * bytes_per_proof) return;
  out[t] = proofs[(t / bytes_per_proof) * proof_stride + (t % bytes_per_proof)];
}

// sum of `count` extended points, one block
template <class C>
__global__ void __launch_bounds__(128) ext_sum_kernel(uint32_t count, const uint32_t* __restrict__ pts, uint32_t* __restrict__ out) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  typename C::Ext acc, p;
  C::set_identity(acc);
  for (uint32_t i = threadIdx.x; i < count; i += blockDim.x) {
    load_words<C::EXT_WORDS>(p, pts + (size_t)i * C::EXT_WORDS);
    C::add(acc, p);
  }
  uint32_t* mine = sh + threadIdx.x * C::EXT_WORDS;
  *reinterpret_cast<typename C::Ext*>(mine) = acc;
  __syncthreads();
  for (uint32_t s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (threadIdx.x + s) * C::EXT_WORDS);
      C::add(acc, o);
      *reinterpret_cast<typename C::Ext*>(mine) = acc;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) store_words<C::EXT_WORDS>(out, acc);
}

// K5 tables: comb[(p * CW + w) * 128 + d - 1] = d * 2^(8w) * P_p for the two commitment bases P_0 = g, P_1 = h
// (publickey.rs:50-52). One thread per entry; built once per public key.
template <class C>
__global__ void comb_build_kernel(uint32_t CW, const uint32_t* __restrict__ aff2, uint32_t* __restrict__ comb) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 2 * CW * COMB_D) return;
  uint32_t d = t % COMB_D + 1, pw = t / COMB_D, w = pw % CW, pi = pw / CW;
  typename C::Aff q;
  typename C::Ext base, r;
  load_words<C::AFF_WORDS>(q, aff2 + (size_t)pi * C::AFF_WORDS);
  C::aff_to_ext(base, q);
  for (uint32_t k = 0; k < w * COMB_C; k++) C::dbl(base);
  uint32_t k[8] = {d, 0, 0, 0, 0, 0, 0, 0};
  scalar_mul_vartime<C>(r, base, k);
  C::to_aff(q, r);
  store_words<C::AFF_WORDS>(comb + (size_t)t * C::AFF_WORDS, q);
}

// ---- integer-pipe roofline probe: ILP-8 independent chains per thread, no memory traffic -------------
// Every multiplicand is a value the loop itself produces, so ptxas can neither hoist the product nor turn the chain into
// additions (it does both when the operands are loop-invariant: a first version of this probe measured 64-bit adds).
// MODE 0: 32-bit multiply-add (IMAD)                        — the issue-rate peak of the integer multiply pipe
// MODE 1: carry-chained 32x32+64 multiply-add (IMAD.WIDE.U32.X rows of 4, as in mul_wide_n) — what a limb product can reach
// MODE 2: 32x32+64 multiply-add with a plain 64-bit register addend (IMAD.WIDE.U32 Rd, Ra, Rb, Rc)
template <int MODE>
__global__ void __launch_bounds__(256) int_peak_kernel(uint32_t iters, uint32_t seed, uint64_t* __restrict__ sink) {
  uint32_t a = seed + threadIdx.x * 2654435761u + 1u, b = a ^ 0x9e3779b9u;
  uint64_t acc[8];
  uint32_t lo[8], hi[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { acc[i] = a + i; lo[i] = b + i; hi[i] = a * (i + 3); }
  for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (MODE == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(lo[i]) : "r"(a), "r"(b));
        if (MODE == 1) {
          if ((i & 3) == 0)
            asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(lo[i]), "+r"(hi[i]) : "r"(lo[(i + 1) & 7]), "r"(b));
          else
            asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(lo[i]), "+r"(hi[i]) : "r"(lo[(i + 1) & 7]), "r"(b));
          if ((i & 3) == 3) asm volatile("addc.u32 %0, %0, 0;" : "+r"(hi[(i + 1) & 7]));
        }
        if (MODE == 2) acc[i] += (uint64_t)(uint32_t)acc[i] * b;
      }
    }
  }
  uint64_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += acc[i] + lo[i] + hi[i];
  if (s == 0x1234567ull) sink[0] = s;  // never true in practice; keeps the chains alive
}

// ---- FP64-pipe probes (bpp_fp64_peak_probe) -------------------------------------------------------------------------
// MODE 0: eight interleaved fma.rz.f64 chains whose operands the loop itself produces; MODE 1: the 5 x 51-bit double-limb
// field multiply of fe25519_f64.cuh, x <- x * y, y <- old x
template <int MODE>
__global__ void __launch_bounds__(256) fp64_peak_kernel(uint32_t iters, double seed, double* __restrict__ sink) {
  if (MODE == 0) {
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = seed + (double)(threadIdx.x + i) * 0x1p-20;
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
      for (int u = 0; u < 4; u++) {
#pragma unroll
        for (int i = 0; i < 8; i++) x[i] = __fma_rz(x[i], x[(i + 1) & 7], x[(i + 2) & 7]);
      }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i];
    if (s == 0.123456789) sink[0] = s;
  } else {
    F5 x, y, z;
#pragma unroll
    for (int i = 0; i < 5; i++) { x.v[i] = (double)(long long)(0x2345678abcdell + 977 * i + threadIdx.x); y.v[i] = -(double)(long long)(0x1fedcba987654ll - 31 * i); }
    for (uint32_t it = 0; it < iters; it++) {
      f5_mul(z, x, y);
      y = x;
      x = z;
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 5; i++) s += x.v[i];
    if (s == 0.123456789) sink[0] = s;
  }
}

}  // namespace bpp

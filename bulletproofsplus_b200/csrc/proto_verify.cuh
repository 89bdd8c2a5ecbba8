// K8 (verifier side) — exponent vectors of the single verification MSM, on device.
//
// Follows (reference paths relative to /root/reference/src):
//   WeightedInnerProductProof::verification_scalars  weighted_inner_product_proof.rs:330-382
//   WeightedInnerProductProof::verify                weighted_inner_product_proof.rs:238-328
//   RangeProof::verify_single                        range/mod.rs:189-238
//   RangeProof::verify_multiple                      range/mod.rs:405-510
// verify_multiple's exponents are exactly verify()'s multiplied by e^-2 (it divides the whole check
// by e^2), so both paths use one kernel with a scale factor `lam` (1 or e^-2).
//
// Split in two phases: `verify_tables_kernel` computes everything that depends only on the
// challenges (s-vector, powers of y, commitment exponents) once per challenge set;
// `verify_scalars_kernel` combines it with each proof's (r', s', delta') and writes the MSM terms.
#pragma once
#include "scalar.cuh"

namespace bpp {

// packed parameter block (canonical limbs, 8 words each): y z e_final alpha r s delta eta, then
// e_round[L], d_L[L], d_R[L]   (constants of the reference: SURVEY §0 F3/F4 when the caller passes NULL)
enum { P_Y = 0, P_Z = 1, P_EFINAL = 2, P_ALPHA = 3, P_R = 4, P_S = 5, P_DELTA = 6, P_ETA = 7, P_FIXED = 8 };

struct VerifyShape {
  uint32_t N;        // n*m, power of two
  uint32_t L;        // log2 N
  uint32_t n, m;     // bit width, aggregation (range mode only)
  uint32_t nV;       // number of V terms
  uint32_t mode;     // 0: caller supplies commitment exponents (standalone WIP verify); 1: range single; 2: range multiple
  uint32_t n_proofs;
  uint32_t pk_G0, pk_H0;  // table indices of G_0 and H_0 (g = 0, h = 1)
  uint32_t dyn0;          // table index of the first per-call point
};

// challenge-only tables, Montgomery form
struct VerifyTables {
  uint32_t* S;      // [N]  s_i * y^-(i+1)
  uint32_t* Srev;   // [N]  s_{N-1-i}
  uint32_t* Gc;     // [N]  commitment exponents of G
  uint32_t* Hc;     // [N]
  uint32_t* Vc;     // [nV]
  uint32_t* misc;   // [8 + 2L] : gc, e, e^2, lam, y, (3 spare), ch_sqr[L], ch_inv_sqr[L]
  uint32_t* tmp;    // [N] scratch (s-vector)
};
enum { VM_GC = 0, VM_E = 1, VM_E2 = 2, VM_LAM = 3, VM_Y = 4, VM_HDR = 8 };

template <class M>
BPP_DEV Sc<M> ld_mont(const uint32_t* p, size_t i) {
  Sc<M> r;
#pragma unroll
  for (int k = 0; k < 8; k++) r.v[k] = p[8 * i + k];
  return r;
}
template <class M>
BPP_DEV void st_mont(uint32_t* p, size_t i, const Sc<M>& s) {
#pragma unroll
  for (int k = 0; k < 8; k++) p[8 * i + k] = s.v[k];
}

// sum_{i=1..n} x^i for n a power of two (util.rs:81-106 sum_of_powers_type2; same value as the slow path)
template <class M>
__device__ Sc<M> sum_powers_from1(const Sc<M>& x, uint32_t n) {
  Sc<M> sum = Sc<M>::zero(), cur = x;
  if (n & (n - 1)) {  // not a power of two: plain loop
    for (uint32_t i = 0; i < n; i++) { sum = sum + cur; cur = cur * x; }
    return sum;
  }
  if (n == 0) return sum;
  Sc<M> result = x, factor = x;  // result = sum_{1..k} x^i, factor = x^k
  for (uint32_t k = 1; k < n; k <<= 1) {
    result = result + factor * result;
    factor = factor * factor;
  }
  return result;
}

// one block; ext_* are canonical inputs used only when mode == 0
template <class M>
__global__ void verify_tables_kernel(VerifyShape vs, const uint32_t* __restrict__ params, const uint32_t* __restrict__ ext_Gc,
                                     const uint32_t* __restrict__ ext_Hc, const uint32_t* __restrict__ ext_gc,
                                     const uint32_t* __restrict__ ext_Vc, const uint32_t* __restrict__ ext_y, VerifyTables T) {
  typedef Sc<M> S;
  const uint32_t N = vs.N, L = vs.L, tid = threadIdx.x, nt = blockDim.x;
  __shared__ uint32_t sh_einv[32 * 8];
  __shared__ uint32_t sh_misc[4 * 8];  // yinv, s0
  S y = vs.mode == 0 ? sc_load<M>(ext_y) : sc_load<M>(params + 8 * P_Y);
  S z = sc_load<M>(params + 8 * P_Z);
  S e = sc_load<M>(params + 8 * P_EFINAL);
  // inverses of the round challenges (batch_invert, prime_field_elem.rs:239-248) and of y
  if (tid < L) st_mont<M>(sh_einv, tid, sc_load<M>(params + 8 * (P_FIXED + tid)).inv());
  if (tid == L) st_mont<M>(sh_misc, 0, y.inv());
  __syncthreads();
  if (tid == 0) {
    S s0 = S::one();
    for (uint32_t k = 0; k < L; k++) s0 = s0 * ld_mont<M>(sh_einv, k);
    st_mont<M>(T.tmp, 0, s0);
    S e2 = e * e;
    S lam = vs.mode == 2 ? e2.inv() : S::one();
    st_mont<M>(T.misc, VM_E, e);
    st_mont<M>(T.misc, VM_E2, e2);
    st_mont<M>(T.misc, VM_LAM, lam);
    st_mont<M>(T.misc, VM_Y, y);
  }
  if (tid < L) {
    S ek = sc_load<M>(params + 8 * (P_FIXED + tid)), ei = ld_mont<M>(sh_einv, tid);
    st_mont<M>(T.misc, VM_HDR + tid, ek * ek);
    st_mont<M>(T.misc, VM_HDR + L + tid, ei * ei);
  }
  __syncthreads();
  // s_i = s_{i-k} * e^2_{L-1-log2 i}  (weighted_inner_product_proof.rs:372-380), level by level
  for (uint32_t t = 0; t < L; t++) {
    S f = ld_mont<M>(T.misc, VM_HDR + (L - 1 - t));
    uint32_t half = 1u << t;
    for (uint32_t j = tid; j < half; j += nt) st_mont<M>(T.tmp, half + j, ld_mont<M>(T.tmp, j) * f);
    __syncthreads();
  }
  S yinv = ld_mont<M>(sh_misc, 0);
  for (uint32_t i = tid; i < N; i += nt) {
    st_mont<M>(T.S, i, ld_mont<M>(T.tmp, i) * yinv.pow_u64(i + 1));
    st_mont<M>(T.Srev, i, ld_mont<M>(T.tmp, N - 1 - i));
  }
  // commitment exponents
  if (vs.mode == 0) {
    for (uint32_t i = tid; i < N; i += nt) {
      st_mont<M>(T.Gc, i, sc_load<M>(ext_Gc + 8 * i));
      st_mont<M>(T.Hc, i, sc_load<M>(ext_Hc + 8 * i));
    }
    for (uint32_t j = tid; j < vs.nV; j += nt) st_mont<M>(T.Vc, j, sc_load<M>(ext_Vc + 8 * j));
    if (tid == 0) st_mont<M>(T.misc, VM_GC, sc_load<M>(ext_gc));
  } else {
    const bool single = vs.mode == 1;
    S zz = z * z, mz = -z;
    S y_N1 = y.pow_u64((uint64_t)N + 1);
    for (uint32_t i = tid; i < N; i += nt) {
      // d_i = 2^(i mod n) [* z^(2(j+1))]  (range/mod.rs:214-218 single, :435-438 multiple)
      S d = S::from_u64(2).pow_u64(i % vs.n);
      if (!single) d = d * zz.pow_u64(i / vs.n + 1);
      st_mont<M>(T.Gc, i, mz);                                // range/mod.rs:213 / :456-459
      st_mont<M>(T.Hc, i, d * y.pow_u64(N - i) + z);          // range/mod.rs:214-218 / :461-465
    }
    for (uint32_t j = tid; j < vs.nV; j += nt)                // range/mod.rs:219 / :475-477
      st_mont<M>(T.Vc, j, single ? y_N1 : zz.pow_u64(j + 1) * y_N1);
    if (tid == 0) {
      S sum_y = sum_powers_from1<M>(y, N);
      S sum_2 = S::from_u64(2).pow_u64(vs.n) - S::one();
      S gc;
      if (single) gc = sum_y * (z - zz) - sum_2 * y_N1 * z;                       // range/mod.rs:221-226
      else gc = sum_y * (z - zz) - y_N1 * z * sum_2 * sum_powers_from1<M>(zz, vs.m);  // range/mod.rs:467-471
      st_mont<M>(T.misc, VM_GC, gc);
    }
  }
}

// grid: (ceil(K / blockDim), n_proofs). Writes K = 5 + 2L + 2N + nV terms per proof:
//   [wipB, wipA, A', g, h, L.., R.., G.., H.., V..] with exponents lam * [1, e, e^2, g_exp, h_exp, ...]
// proofs: wire layout A | L[L] | R[L] | wipA | wipB | r' | s' | d'   (include/bpp.h)
template <class M>
__global__ void verify_scalars_kernel(VerifyShape vs, VerifyTables T, const uint8_t* __restrict__ proofs, size_t proof_stride,
                                      size_t scalar_off, uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx,
                                      uint8_t* __restrict__ bad_scalar) {
  typedef Sc<M> S;
  const uint32_t N = vs.N, L = vs.L, K = 5 + 2 * L + 2 * N + vs.nV;
  const uint32_t p = blockIdx.y;
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const uint8_t* sb = proofs + (size_t)p * proof_stride + scalar_off;
  uint32_t raw[3][8];
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 3; j++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint8_t* q = sb + 32 * j + 4 * i;
      raw[j][i] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    }
    ok = ok && S::canonical_ok(raw[j]);
  }
  if (k == 0) bad_scalar[p] = ok ? 0 : 1;
  if (!ok) {  // keep the MSM well-formed; the verdict is FORMAT regardless of its result
#pragma unroll
    for (int j = 0; j < 3; j++)
      for (int i = 0; i < 8; i++) raw[j][i] = 0;
  }
  S rp = sc_load<M>(raw[0]), sp = sc_load<M>(raw[1]), dp = sc_load<M>(raw[2]);
  S e = ld_mont<M>(T.misc, VM_E), e2 = ld_mont<M>(T.misc, VM_E2), lam = ld_mont<M>(T.misc, VM_LAM), y = ld_mont<M>(T.misc, VM_Y);
  const uint32_t dyn = vs.dyn0 + p * (2 * L + 3);  // this proof's decoded points: A, L.., R.., wipA, wipB
  S val;
  uint32_t idx;
  if (k == 0) { val = lam; idx = dyn + 2 * L + 2; }                   // wipB
  else if (k == 1) { val = lam * e; idx = dyn + 2 * L + 1; }          // wipA
  else if (k == 2) { val = lam * e2; idx = dyn; }                     // A' (RangeProof.A)
  else if (k == 3) { val = lam * ((-rp) * y * sp + ld_mont<M>(T.misc, VM_GC) * e2); idx = 0; }  // wip.rs:290
  else if (k == 4) { val = lam * (-dp); idx = 1; }                                              // wip.rs:292
  else if (k < 5 + L) { uint32_t j = k - 5; val = lam * ld_mont<M>(T.misc, VM_HDR + j) * e2; idx = dyn + 1 + j; }          // wip.rs:265-268
  else if (k < 5 + 2 * L) { uint32_t j = k - 5 - L; val = lam * ld_mont<M>(T.misc, VM_HDR + L + j) * e2; idx = dyn + 1 + L + j; }  // :270-273
  else if (k < 5 + 2 * L + N) {  // wip.rs:275-282
    uint32_t i = k - 5 - 2 * L;
    val = lam * ((-ld_mont<M>(T.S, i)) * (rp * e * y) + ld_mont<M>(T.Gc, i) * e2);
    idx = vs.pk_G0 + i;
  } else if (k < 5 + 2 * L + 2 * N) {  // wip.rs:284-288
    uint32_t i = k - 5 - 2 * L - N;
    val = lam * ((-ld_mont<M>(T.Srev, i)) * (sp * e) + ld_mont<M>(T.Hc, i) * e2);
    idx = vs.pk_H0 + i;
  } else {  // wip.rs:294-297
    uint32_t j = k - 5 - 2 * L - 2 * N;
    val = lam * ld_mont<M>(T.Vc, j) * e2;
    idx = vs.dyn0 + vs.n_proofs * (2 * L + 3) + p * vs.nV + j;
  }
  size_t o = (size_t)p * K + k;
  sc_store<M>(out_scalars + 8 * o, val);
  out_pidx[o] = idx;
}

}  // namespace bpp

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
// Every exponent of the MSM [wipB, wipA, A', g, h, L.., R.., G.., H.., V..] is affine in one of the
// proof's scalars:    exponent_k = Tmul_k * x_k + Tadd_k,   x_k in {r', s', delta', r'*s'},
// with Tmul, Tadd depending on the challenges only. `verify_tables_kernel` builds Tmul (Montgomery
// form) and Tadd (canonical) once per challenge set — one batched inversion, running products, no
// per-element exponentiation; `verify_scalars_kernel` then spends one Montgomery multiplication and
// one modular addition per term and proof (Montgomery x canonical = canonical, so no conversions).
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

// challenge-only tables
struct VerifyTables {
  uint32_t* Tmul;   // [K]  Montgomery form (used for k = 3, 4 and the G, H ranges)
  uint32_t* Tadd;   // [K]  canonical
  uint32_t* tmp;    // [N + 64] scratch: s-vector (Montgomery), then small values
  uint32_t* zpow;   // [m + 1] scratch: z^(2(j+1)) (Montgomery)
  // per-proof challenges (a transcript-owning host): block p of the tables kernel works on proof p's parameter block and its
  // own table slices; all strides are in 32-bit words and are 0 when one challenge set serves the whole batch
  uint32_t stride_T, stride_tmp, stride_zpow, stride_params;
  // per-proof challenges only: inverses of [e_0..e_{L-1}, y, e] (Montgomery) computed beforehand by verify_inverses_kernel,
  // one thread per proof; nullptr = the tables kernel inverts them itself (one challenge set: a single block)
  const uint32_t* pre;
  uint32_t stride_pre;
};

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

// Per-proof challenges: thread per proof, the one batched inversion of its L + 2 challenges (batch_invert,
// prime_field_elem.rs:239-248). Inside the tables kernel this chain of ~400 dependent multiplications ran on ONE thread of
// every proof's block while the other 127 waited: 2.9 ms per 2048 proofs. Here all proofs' chains run side by side.
template <class M>
__global__ void __launch_bounds__(32)
verify_inverses_kernel(uint32_t n_proofs, uint32_t L, const uint32_t* __restrict__ params_all, uint32_t stride_params, uint32_t* __restrict__ pre,
                       uint32_t stride_pre) {
  typedef Sc<M> S;
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_proofs) return;
  const uint32_t* params = params_all + (size_t)p * stride_params;
  uint32_t x[34 * 8], scr[34 * 8];
  for (uint32_t k = 0; k < L; k++) st_mont<M>(x, k, sc_load<M>(params + 8 * (P_FIXED + k)));
  st_mont<M>(x, L, sc_load<M>(params + 8 * P_Y));
  st_mont<M>(x, L + 1, sc_load<M>(params + 8 * P_EFINAL));
  sc_batch_invert<M>(x, scr, L + 2);
  uint32_t* o = pre + (size_t)p * stride_pre;
  for (uint32_t i = 0; i < (L + 2) * 8; i++) o[i] = x[i];
}

// one block of VT_THREADS threads; ext_* are canonical inputs used only when mode == 0
constexpr int VT_THREADS = 512;
// NT = block size: 512 for the one block of a shared challenge set, 256 for one block per proof. (Holding the per-proof instance
// to 80 registers for six 128-thread blocks per SM instead of four was measured slower — 1.20 vs 1.05 ms per 2048 proofs: 384 B of
// spills on the scalar arithmetic.)
template <class M, int NT>
__global__ void __launch_bounds__(NT)
verify_tables_kernel(VerifyShape vs, const uint32_t* __restrict__ params, const uint32_t* __restrict__ ext_Gc,
                     const uint32_t* __restrict__ ext_Hc, const uint32_t* __restrict__ ext_gc,
                     const uint32_t* __restrict__ ext_Vc, const uint32_t* __restrict__ ext_y, VerifyTables T) {
  typedef Sc<M> S;
  const uint32_t N = vs.N, L = vs.L, tid = threadIdx.x, nt = blockDim.x;
  const uint32_t K0 = 5 + 2 * L;  // first G term
  params += (size_t)blockIdx.x * T.stride_params;
  T.Tmul += (size_t)blockIdx.x * T.stride_T; T.Tadd += (size_t)blockIdx.x * T.stride_T;
  T.tmp += (size_t)blockIdx.x * T.stride_tmp; T.zpow += (size_t)blockIdx.x * T.stride_zpow;
  // shared: inverses of e_0..e_{L-1}, y, e  (slots 0..L+1), scratch for the batch inversion, and a few broadcast values
  __shared__ uint32_t sh_inv[34 * 8];
  __shared__ uint32_t sh_scr[34 * 8];
  __shared__ uint32_t sh_val[8 * 8];  // 0: lam*e2  1: -lam*e  2: y^N  3: lam
  enum { V_LE2 = 0, V_NLE = 1, V_YN = 2, V_LAM = 3 };
  const S y = vs.mode == 0 ? sc_load<M>(ext_y) : sc_load<M>(params + 8 * P_Y);
  const S z = sc_load<M>(params + 8 * P_Z);
  const S e = sc_load<M>(params + 8 * P_EFINAL);
  // ---- phase 0: one inversion for all of e_k^-1 (batch_invert, prime_field_elem.rs:239-248), y^-1, e^-1 ----
  if (T.pre) {
    const uint32_t* pre = T.pre + (size_t)blockIdx.x * T.stride_pre;
    for (uint32_t i = tid; i < (L + 2) * 8; i += nt) sh_inv[i] = pre[i];
  } else {
    if (tid < L) st_mont<M>(sh_inv, tid, sc_load<M>(params + 8 * (P_FIXED + tid)));
    if (tid == L) st_mont<M>(sh_inv, L, y);
    if (tid == L + 1) st_mont<M>(sh_inv, L + 1, e);
    __syncthreads();
    if (tid == 0) sc_batch_invert<M>(sh_inv, sh_scr, L + 2);
  }
  __syncthreads();
  const S yinv = ld_mont<M>(sh_inv, L), einv = ld_mont<M>(sh_inv, L + 1);
  // ---- phase 1: s_0, per-round squares, broadcast values ----
  if (tid == 0) {
    // tmp[i] = -lam e s_i: the factor every use of s_i carries (Tmul of G_i and H_i) is folded into s_0
    S s0 = vs.mode == 2 ? -(einv * einv * e) : -e;
    for (uint32_t k = 0; k < L; k++) s0 = s0 * ld_mont<M>(sh_inv, k);  // wip.rs:372
    st_mont<M>(T.tmp, 0, s0);
  }
  if (tid == 32) {
    S e2 = e * e;
    S lam = vs.mode == 2 ? einv * einv : S::one();
    st_mont<M>(sh_val, V_LE2, lam * e2);
    st_mont<M>(sh_val, V_NLE, -(lam * e));
    st_mont<M>(sh_val, V_LAM, lam);
    // constant terms: exponents of wipB, wipA, A' are lam * [1, e, e^2]  (wip.rs:257-262 scaled)
    sc_store<M>(T.Tadd + 8 * 0, lam);
    sc_store<M>(T.Tadd + 8 * 1, lam * e);
    sc_store<M>(T.Tadd + 8 * 2, lam * e2);
    // g: lam * (-(r' s' y) + gc e^2) -> Tmul = -lam*y (times r'*s');  h: lam * (-delta') -> Tmul = -lam, Tadd = 0
    st_mont<M>(T.Tmul, 3, -(lam * y));
    st_mont<M>(T.Tmul, 4, -lam);
    sc_store<M>(T.Tadd + 8 * 4, S::zero());
  }
  if (tid == 64) st_mont<M>(sh_val, V_YN, y.pow_u64(N));
  __syncthreads();
  const S le2 = ld_mont<M>(sh_val, V_LE2), nle = ld_mont<M>(sh_val, V_NLE);
  if (tid < L) {  // L_j, R_j exponents: lam * e^2 * e_j^(+-2)  (wip.rs:265-273)
    S ek = sc_load<M>(params + 8 * (P_FIXED + tid)), ei = ld_mont<M>(sh_inv, tid);
    S ek2 = ek * ek;
    st_mont<M>(sh_scr, tid, ek2);
    sc_store<M>(T.Tadd + 8 * (5 + tid), le2 * ek2);
    sc_store<M>(T.Tadd + 8 * (5 + L + tid), le2 * (ei * ei));
  }
  __syncthreads();
  // s_i = s_{i-k} * e^2_{L-1-log2 i}  (weighted_inner_product_proof.rs:372-380), level by level
  for (uint32_t t = 0; t < L; t++) {
    S f = ld_mont<M>(sh_scr, L - 1 - t);
    uint32_t half = 1u << t;
    for (uint32_t j = tid; j < half; j += nt) st_mont<M>(T.tmp, half + j, ld_mont<M>(T.tmp, j) * f);
    __syncthreads();
  }
  // ---- phase 2: G and H terms, a contiguous chunk of indices per thread, running products ----
  const bool range = vs.mode != 0, single = vs.mode == 1;
  const S zz = z * z;
  if (range) {  // z^(2(j+1)) (range/mod.rs:287 / :430-433)
    for (uint32_t j = tid; j < vs.m; j += nt) st_mont<M>(T.zpow, j, zz.pow_u64(j + 1));
  }
  __syncthreads();
  {
    const uint32_t ch = (N + nt - 1) / nt;
    const uint32_t i0 = tid * ch, i1 = (i0 + ch < N) ? i0 + ch : N;
    if (i0 < N) {
      S yi = yinv.pow_u64(i0);  // y^-i
      const S le2yN = le2 * ld_mont<M>(sh_val, V_YN), le2z = le2 * z, two_yinv = yinv + yinv;
      uint32_t gconst[8];
      sc_store<M>(gconst, le2 * (-z));
      S q = S::zero();  // lam e^2 d_i y^(N-i) with d_i = 2^(i mod n) [* z^(2(j+1))]; q_(i+1) = q_i * 2 / y inside a block of n
      for (uint32_t i = i0; i < i1; i++) {
        // G_i: lam * (-(s_i y^-(i+1)) (r' e y) + Gc_i e^2) = [-lam e s_i y^-i] r' + lam e^2 Gc_i   (wip.rs:275-282)
        st_mont<M>(T.Tmul, K0 + i, ld_mont<M>(T.tmp, i) * yi);
        // H_i: lam * (-(s' e) s_{N-1-i} + Hc_i e^2)                                              (wip.rs:284-288)
        st_mont<M>(T.Tmul, K0 + N + i, ld_mont<M>(T.tmp, N - 1 - i));
        if (range) {
          // d_i = 2^(i mod n) [* z^(2(j+1))]  (range/mod.rs:214-218 single, :435-438 multiple)
          if (i == i0 || i % vs.n == 0) {
            const S base = single ? le2yN : le2yN * ld_mont<M>(T.zpow, i / vs.n);
            const uint32_t k = i % vs.n;  // n <= 64 (u64 values): 2^k fits a word pair
            q = base * yi * (k < 64 ? S::from_u64(1ull << k) : S::from_u64(2).pow_u64(k));
          }
#pragma unroll
          for (int w = 0; w < 8; w++) T.Tadd[8 * (K0 + i) + w] = gconst[w];    // Gc_i = -z  (range/mod.rs:213 / :456-459)
          sc_store<M>(T.Tadd + 8 * (K0 + N + i), q + le2z);                    // Hc_i = d_i y^(N-i) + z  (:214-218 / :461-465)
          q = q * two_yinv;
        } else {
          sc_store<M>(T.Tadd + 8 * (K0 + i), le2 * sc_load<M>(ext_Gc + 8 * i));
          sc_store<M>(T.Tadd + 8 * (K0 + N + i), le2 * sc_load<M>(ext_Hc + 8 * i));
        }
        yi = yi * yinv;
      }
    }
  }
  // ---- phase 3: V terms and the g constant ----
  if (range) {
    const S y_N1 = ld_mont<M>(sh_val, V_YN) * y;
    for (uint32_t j = tid; j < vs.nV; j += nt)  // range/mod.rs:219 / :475-477
      sc_store<M>(T.Tadd + 8 * (K0 + 2 * N + j), le2 * (single ? y_N1 : ld_mont<M>(T.zpow, j) * y_N1));
    if (tid == nt - 1) {
      S sum_y = sum_powers_from1<M>(y, N);
      S sum_2 = S::from_u64(2).pow_u64(vs.n) - S::one();
      S gc;
      if (single) gc = sum_y * (z - zz) - sum_2 * y_N1 * z;                           // range/mod.rs:221-226
      else gc = sum_y * (z - zz) - y_N1 * z * sum_2 * sum_powers_from1<M>(zz, vs.m);  // range/mod.rs:467-471
      sc_store<M>(T.Tadd + 8 * 3, le2 * gc);
    }
  } else {
    for (uint32_t j = tid; j < vs.nV; j += nt) sc_store<M>(T.Tadd + 8 * (K0 + 2 * N + j), le2 * sc_load<M>(ext_Vc + 8 * j));
    if (tid == nt - 1) sc_store<M>(T.Tadd + 8 * 3, le2 * sc_load<M>(ext_gc));
  }
}

// Per-proof challenges: thread per proof expands the caller's compact block [y, z, e_final, e_round[0..L)] into the packed
// parameter layout the tables kernel reads. A non-canonical or zero challenge (not invertible) marks the proof FORMAT and is
// replaced by 1 so that the arithmetic stays defined.
template <class M>
__global__ void verify_expand_challenges_kernel(uint32_t n_proofs, uint32_t L, const uint8_t* __restrict__ challenges,
                                                uint32_t* __restrict__ params_all, uint8_t* __restrict__ bad) {
  typedef Sc<M> S;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_proofs) return;
  const uint8_t* src = challenges + (size_t)p * (3 + L) * 32;
  uint32_t* dst = params_all + (size_t)p * (P_FIXED + 3 * L) * 8;
  bool any_bad = false;
  for (uint32_t j = 0; j < 3 + L; j++) {
    uint32_t w[8], nz = 0;
    for (int i = 0; i < 8; i++) {
      const uint8_t* q = src + 32 * j + 4 * i;
      w[i] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
      nz |= w[i];
    }
    const bool must_invert = j != 1;  // y, e_final, e_k are inverted; z is not
    if (!S::canonical_ok(w) || (must_invert && !nz)) {
      any_bad = true;
      for (int i = 0; i < 8; i++) w[i] = i == 0;
    }
    const uint32_t slot = j == 0 ? P_Y : j == 1 ? P_Z : j == 2 ? P_EFINAL : P_FIXED + (j - 3);
    for (int i = 0; i < 8; i++) dst[8 * slot + i] = w[i];
  }
  if (any_bad) bad[p] = 1;
}

// thread per proof: parse and check (r', s', delta'); per-proof scalars x = [r', s', delta', r'*s'] (canonical) -> xs[p][4][8]
// proofs: wire layout A | L[L] | R[L] | wipA | wipB | r' | s' | d'   (include/bpp.h)
template <class M>
__global__ void verify_proof_scalars_kernel(uint32_t n_proofs, const uint8_t* __restrict__ proofs, size_t proof_stride, size_t scalar_off,
                                            uint32_t* __restrict__ xs, uint8_t* __restrict__ bad_scalar) {
  typedef Sc<M> S;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_proofs) return;
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
  bad_scalar[p] = ok ? 0 : 1;
  if (!ok) {  // keep the MSM well-formed; the verdict is FORMAT regardless of its result
#pragma unroll
    for (int j = 0; j < 3; j++)
      for (int i = 0; i < 8; i++) raw[j][i] = 0;
  }
  uint32_t* o = xs + (size_t)p * 32;
  for (int j = 0; j < 3; j++)
    for (int i = 0; i < 8; i++) o[8 * j + i] = raw[j][i];
  S rs = sc_load<M>(raw[0]) * S::raw(raw[1]);  // Montgomery x canonical = canonical r' * s'
  for (int i = 0; i < 8; i++) o[24 + i] = rs.v[i];
}

// grid: (ceil(K / blockDim), n_proofs). Writes K = 5 + 2L + 2N + nV terms per proof:
//   [wipB, wipA, A', g, h, L.., R.., G.., H.., V..] with exponents lam * [1, e, e^2, g_exp, h_exp, ...]  (wip.rs:297-318)
// skip_fixed: the generator terms (g, h, G_i, H_i) get their point index only — their exponents are computed where they are
// used, inside direct_accumulate_kernel (pk_direct.cuh, DirectVerifySrc), which saves this kernel's multiplication and the
// round trip of 2 + 2N scalars per proof through HBM.
template <class M>
__global__ void verify_scalars_kernel(VerifyShape vs, VerifyTables T, const uint32_t* __restrict__ xs,
                                      uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx, uint32_t skip_fixed) {
  typedef Sc<M> S;
  const uint32_t N = vs.N, L = vs.L, K = 5 + 2 * L + 2 * N + vs.nV;
  const uint32_t p = blockIdx.y;
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  if (skip_fixed && (k == 3 || k == 4 || (k >= 5 + 2 * L && k < 5 + 2 * L + 2 * N))) {
    out_pidx[(size_t)p * K + k] = k == 3 ? 0u : k == 4 ? 1u : k < 5 + 2 * L + N ? vs.pk_G0 + (k - 5 - 2 * L) : vs.pk_H0 + (k - 5 - 2 * L - N);
    return;
  }
  const uint32_t dyn = vs.dyn0 + p * (2 * L + 3);  // this proof's decoded points: A, L.., R.., wipA, wipB
  const uint32_t* x = xs + (size_t)p * 32;
  T.Tadd += (size_t)p * T.stride_T; T.Tmul += (size_t)p * T.stride_T;
  S val = S::raw(T.Tadd + 8 * k);
  uint32_t idx;
  if (k == 0) idx = dyn + 2 * L + 2;                                                   // wipB
  else if (k == 1) idx = dyn + 2 * L + 1;                                              // wipA
  else if (k == 2) idx = dyn;                                                          // A' (RangeProof.A)
  else if (k == 3) { val = val + S::raw(T.Tmul + 8 * 3) * S::raw(x + 24); idx = 0; }   // g  (wip.rs:290)
  else if (k == 4) { val = S::raw(T.Tmul + 8 * 4) * S::raw(x + 16); idx = 1; }         // h  (wip.rs:292)
  else if (k < 5 + L) idx = dyn + 1 + (k - 5);                                         // L_j (wip.rs:265-268)
  else if (k < 5 + 2 * L) idx = dyn + 1 + L + (k - 5 - L);                             // R_j (:270-273)
  else if (k < 5 + 2 * L + N) { val = val + S::raw(T.Tmul + 8 * k) * S::raw(x); idx = vs.pk_G0 + (k - 5 - 2 * L); }              // G_i (:275-282)
  else if (k < 5 + 2 * L + 2 * N) { val = val + S::raw(T.Tmul + 8 * k) * S::raw(x + 8); idx = vs.pk_H0 + (k - 5 - 2 * L - N); }  // H_i (:284-288)
  else idx = vs.dyn0 + vs.n_proofs * (2 * L + 3) + p * vs.nV + (k - 5 - 2 * L - 2 * N);                                           // V_j (:294-297)
  const size_t o = (size_t)p * K + k;
#pragma unroll
  for (int i = 0; i < 8; i++) out_scalars[8 * o + i] = val.v[i];
  out_pidx[o] = idx;
}

// ---- random-linear-combination batch check ---------------------------------------------------------
// sum_p w_p * (verification equation of proof p) = 0 for weights the prover could not predict: the generator terms of all
// proofs collapse into ONE term per generator, sum_p w_p * exponent_{p,k}; only the 2L + 3 + nV per-proof points stay apart.
// One MSM of 2 + 2N + count * (2L + 3 + nV) terms replaces `count` MSMs of K terms. A failing combination says only "some proof
// is bad": the caller falls back to the per-proof path, so verdicts are always the per-proof ones.

// thread per proof: effective weight (Montgomery; 0 for a proof that failed to decode) and its FORMAT flag
template <class M>
__global__ void rlc_weights_kernel(uint32_t n_proofs, uint32_t pts_per_proof, uint32_t nV, const uint8_t* __restrict__ weights,
                                   const uint8_t* __restrict__ dyn_ok, const uint8_t* __restrict__ bad_scalar, uint32_t* __restrict__ w_eff,
                                   uint8_t* __restrict__ fmt, uint8_t* __restrict__ bad_weight) {
  typedef Sc<M> S;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_proofs) return;
  bool f = bad_scalar[p] != 0;
  for (uint32_t k = 0; k < pts_per_proof; k++) f = f || !dyn_ok[(size_t)p * pts_per_proof + k];
  for (uint32_t k = 0; k < nV; k++) f = f || !dyn_ok[(size_t)n_proofs * pts_per_proof + (size_t)p * nV + k];
  uint32_t w[8], nz = 0;
  for (int i = 0; i < 8; i++) {
    const uint8_t* q = weights + (size_t)p * 32 + 4 * i;
    w[i] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    nz |= w[i];
  }
  if (!S::canonical_ok(w) || !nz) { bad_weight[0] = 1; f = true; }
  fmt[p] = f ? 1 : 0;
  st_mont<M>(w_eff, p, f ? S::zero() : sc_load<M>(w));
}

// generator term t in [0, 2 + 2N) (g, h, G_i, H_i)  ->  sum_p w_p * exponent_{p,k(t)}. Block = 32 consecutive terms x 8 proof lanes,
// gridDim.y proof groups: every thread adds up n_proofs / (8 * gridDim.y) products (1 KB contiguous per proof and warp), the
// eight lanes meet in shared memory, the groups in rlc_fold_finish_kernel. partial: [gridDim.y][2 + 2N] canonical scalars.
constexpr uint32_t RLC_GROUPS = 16;
template <class M>
__global__ void __launch_bounds__(256)
rlc_fold_fixed_kernel(VerifyShape vs, const uint32_t* __restrict__ w_eff, const uint32_t* __restrict__ in_scalars, uint32_t* __restrict__ partial) {
  typedef Sc<M> S;
  __shared__ uint32_t sh[256 * 8];
  const uint32_t N = vs.N, L = vs.L, K = 5 + 2 * L + 2 * N + vs.nV, nT = 2 + 2 * N;
  const uint32_t tx = threadIdx.x & 31, ly = threadIdx.x >> 5, t = blockIdx.x * 32 + tx;
  const uint32_t k = t < 2 ? 3 + t : 5 + 2 * L + (t - 2);
  S acc = S::zero();
  if (t < nT)
    for (uint32_t p = blockIdx.y * 8 + ly; p < vs.n_proofs; p += 8 * gridDim.y)
      acc = acc + ld_mont<M>(w_eff, p) * S::raw(in_scalars + ((size_t)p * K + k) * 8);
  st_mont<M>(sh, threadIdx.x, acc);
  __syncthreads();
  if (ly == 0 && t < nT) {
    for (uint32_t j = 1; j < 8; j++) acc = acc + ld_mont<M>(sh, j * 32 + tx);
    st_mont<M>(partial, (size_t)blockIdx.y * nT + t, acc);
  }
}
template <class M>
__global__ void rlc_fold_finish_kernel(uint32_t nT, uint32_t groups, const uint32_t* __restrict__ partial, uint32_t* __restrict__ out_scalars,
                                       uint32_t* __restrict__ out_pidx) {
  typedef Sc<M> S;
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nT) return;
  S acc = ld_mont<M>(partial, t);
  for (uint32_t g = 1; g < groups; g++) acc = acc + ld_mont<M>(partial, (size_t)g * nT + t);
#pragma unroll
  for (int i = 0; i < 8; i++) out_scalars[8 * (size_t)t + i] = acc.v[i];
  out_pidx[t] = t;  // table order of the public key: g, h, G[N], H[N]
}

// thread per (proof, per-proof point): w_p * exponent; the point index is the one verify_scalars_kernel assigned
template <class M>
__global__ void rlc_scale_dyn_kernel(VerifyShape vs, const uint32_t* __restrict__ w_eff, const uint32_t* __restrict__ in_scalars,
                                     const uint32_t* __restrict__ in_pidx, uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx) {
  typedef Sc<M> S;
  const uint32_t N = vs.N, L = vs.L, K = 5 + 2 * L + 2 * N + vs.nV, per = 2 * L + 3 + vs.nV;
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= vs.n_proofs * per) return;
  const uint32_t p = t / per, j = t % per;
  const uint32_t k = j < 3 ? j : j < 3 + 2 * L ? 5 + (j - 3) : 5 + 2 * L + 2 * N + (j - 3 - 2 * L);
  const S v = ld_mont<M>(w_eff, p) * S::raw(in_scalars + ((size_t)p * K + k) * 8);
  const size_t o = 2 + 2 * (size_t)N + t;
#pragma unroll
  for (int i = 0; i < 8; i++) out_scalars[8 * o + i] = v.v[i];
  out_pidx[o] = in_pidx[(size_t)p * K + k];
}

}  // namespace bpp

// K4/K8 (prover side) + K7 — scalar side of proving, on device.
//
// Follows (reference paths relative to /root/reference/src):
//   RangeProof::prove_single    range/mod.rs:80-187
//   RangeProof::prove_multiple  range/mod.rs:240-403
//   WeightedInnerProductProof::prove  weighted_inner_product_proof.rs:36-226
//
// Design choice (option (c)): the generators are never folded. The reference
// folds G,H with a 2-term MulVec per element per round (weighted_inner_product_proof.rs:147-164);
// here the coefficient each ORIGINAL generator carries after k rounds,
//     gcoef_k[j] = prod_{t<=k} (bit_t(j) ? yhat_t^-1 * e_t : e_t^-1),   hcoef_k[j] = prod (bit_t(j) ? e_t^-1 : e_t),
// is tracked as a scalar, and every L_k, R_k, A is one MSM job over the original G,H with the folded
// scalars. The group elements are identical to the reference's; all point work becomes MSM (K1/K2).
// A_hat (range/mod.rs:330-343) and the P update (wip.rs:137-142) are dead in the reference
// (never read) and are not computed.
#pragma once
#include <cooperative_groups.h>
#include "proto_verify.cuh"

namespace bpp {
namespace cg = cooperative_groups;

struct ProveShape {
  uint32_t N, L, n, m;
  uint32_t mode;     // 0: standalone WIP (caller supplies a, b, y-powers, gamma); 1: range single; 2: range multiple
  uint32_t n_proofs;
  uint32_t pk_G0, pk_H0, dyn0;
  uint32_t terms_per_proof;  // hasA + 2L(N+2) + (2N+2) + 2
  uint32_t has_A;
};

// challenge-only tables, Montgomery form
struct ProveTables {
  uint32_t* ypow;   // [N]      y^(i+1)            (mode 0: the caller's power_of_y_vec)
  uint32_t* Hexp;   // [N]      d_i*y^(N-i) + z    (range modes)
  uint32_t* gcoef;  // [(L+1)*N]
  uint32_t* hcoef;  // [(L+1)*N]
  uint32_t* rnd;    // [4*L]    per round: yhat, yhat^-1, e, e^-1
  uint32_t* misc;   // [4 + m]  z, y^(N+1), spare, spare, zpow[m]
};
enum { PM_Z = 0, PM_YN1 = 1, PM_HDR = 4 };

// one block
template <class M>
__global__ void prove_tables_kernel(ProveShape ps, const uint32_t* __restrict__ params, const uint32_t* __restrict__ ext_ypow,
                                    ProveTables T) {
  typedef Sc<M> S;
  const uint32_t N = ps.N, L = ps.L, tid = threadIdx.x, nt = blockDim.x;
  S y = sc_load<M>(params + 8 * P_Y), z = sc_load<M>(params + 8 * P_Z);
  for (uint32_t i = tid; i < N; i += nt)
    st_mont<M>(T.ypow, i, ps.mode == 0 ? sc_load<M>(ext_ypow + 8 * i) : y.pow_u64(i + 1));  // util.rs:34-37
  __syncthreads();
  if (tid < L) {  // weighted_inner_product_proof.rs:98-99,131-134
    uint32_t nk = N >> (tid + 1);
    S yh = ld_mont<M>(T.ypow, nk - 1), e = sc_load<M>(params + 8 * (P_FIXED + tid));
    st_mont<M>(T.rnd, 4 * tid + 0, yh);
    st_mont<M>(T.rnd, 4 * tid + 1, yh.inv());
    st_mont<M>(T.rnd, 4 * tid + 2, e);
    st_mont<M>(T.rnd, 4 * tid + 3, e.inv());
  }
  __syncthreads();
  for (uint32_t j = tid; j < N; j += nt) {
    S gc = S::one(), hc = S::one();
    st_mont<M>(T.gcoef, j, gc);
    st_mont<M>(T.hcoef, j, hc);
    for (uint32_t k = 0; k < L; k++) {  // weighted_inner_product_proof.rs:144-164 as scalars
      uint32_t nk = N >> (k + 1);
      bool upper = (j / nk) & 1;
      S yhi = ld_mont<M>(T.rnd, 4 * k + 1), e = ld_mont<M>(T.rnd, 4 * k + 2), ei = ld_mont<M>(T.rnd, 4 * k + 3);
      gc = gc * (upper ? yhi * e : ei);
      hc = hc * (upper ? ei : e);
      st_mont<M>(T.gcoef, (size_t)(k + 1) * N + j, gc);
      st_mont<M>(T.hcoef, (size_t)(k + 1) * N + j, hc);
    }
  }
  if (ps.mode != 0) {
    const bool single = ps.mode == 1;
    S zz = z * z;
    for (uint32_t i = tid; i < N; i += nt) {  // range/mod.rs:125-129 / :288-302
      S d = S::from_u64(2).pow_u64(i % ps.n);
      if (!single) d = d * zz.pow_u64(i / ps.n + 1);
      st_mont<M>(T.Hexp, i, d * y.pow_u64(N - i) + z);
    }
    for (uint32_t j = tid; j < ps.m; j += nt) st_mont<M>(T.misc, PM_HDR + j, zz.pow_u64(j + 1));  // range/mod.rs:287
    if (tid == 0) {
      st_mont<M>(T.misc, PM_Z, z);
      st_mont<M>(T.misc, PM_YN1, y.pow_u64((uint64_t)N + 1));  // range/mod.rs:131 / :304
    }
  }
}

// one block per proof: a_vec, b_vec, alpha_hat  (range/mod.rs:156-172 / :347-376)
template <class M>
__global__ void range_prove_init_kernel(ProveShape ps, ProveTables T, const uint32_t* __restrict__ params,
                                        const uint64_t* __restrict__ values, const uint32_t* __restrict__ gammas,
                                        uint32_t* __restrict__ a_out, uint32_t* __restrict__ b_out, uint32_t* __restrict__ alpha_out,
                                        uint8_t* __restrict__ bad) {
  typedef Sc<M> S;
  const uint32_t N = ps.N, p = blockIdx.x;
  S z = ld_mont<M>(T.misc, PM_Z), one = S::one();
  S nz = -z, omz = one - z;
  for (uint32_t i = threadIdx.x; i < N; i += blockDim.x) {
    uint32_t i1 = i % ps.n, i2 = i / ps.n;
    uint64_t v = values[(size_t)p * ps.m + i2];
    bool bit = i1 < 64 ? ((v >> i1) & 1) : false;
    S he = ld_mont<M>(T.Hexp, i);
    st_mont<M>(a_out, (size_t)p * N + i, bit ? omz : nz);
    st_mont<M>(b_out, (size_t)p * N + i, bit ? he : he - one);
  }
  if (threadIdx.x == 0) {
    S alpha = sc_load<M>(params + 8 * P_ALPHA), yN1 = ld_mont<M>(T.misc, PM_YN1), acc = S::zero();
    bool ok = true;
    for (uint32_t j = 0; j < ps.m; j++) {
      const uint32_t* g = gammas + 8 * ((size_t)p * ps.m + j);
      ok = ok && S::canonical_ok(g);
      S gm = sc_load<M>(g);
      acc = acc + (ps.mode == 1 ? gm : ld_mont<M>(T.misc, PM_HDR + j) * gm);
    }
    st_mont<M>(alpha_out, p, alpha + acc * yN1);
    if (!ok) bad[0] = 1;
  }
}

template <class M>
__device__ Sc<M> block_sum(Sc<M> v, uint32_t* sh) {  // sh: blockDim.x * 8 words
  st_mont<M>(sh, threadIdx.x, v);
  __syncthreads();
  for (uint32_t s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (threadIdx.x < s) st_mont<M>(sh, threadIdx.x, ld_mont<M>(sh, threadIdx.x) + ld_mont<M>(sh, threadIdx.x + s));
    __syncthreads();
  }
  Sc<M> r = ld_mont<M>(sh, 0);
  __syncthreads();
  return r;
}

// The two block totals x, y summed over the blocks of a thread-block cluster: every block publishes its pair in its own shared
// memory, reads every rank's pair through distributed shared memory (same order in every block) and waits until all have
// read before the slots can be reused. cs == 1: nothing to exchange.
template <class M>
__device__ void cluster_sum2(Sc<M>& x, Sc<M>& y, uint32_t* slot, uint32_t cs) {  // slot: 16 words of static shared memory
  if (cs == 1) return;
  cg::cluster_group cl = cg::this_cluster();
  if (threadIdx.x == 0) { st_mont<M>(slot, 0, x); st_mont<M>(slot, 1, y); }
  cl.sync();
  Sc<M> sx = Sc<M>::zero(), sy = Sc<M>::zero();
  for (uint32_t r = 0; r < cs; r++) {
    const uint32_t* rs = cl.map_shared_rank(slot, r);
    sx = sx + ld_mont<M>(rs, 0);
    sy = sy + ld_mont<M>(rs, 1);
  }
  cl.sync();
  x = sx; y = sy;
}

// one block per proof; blockDim.x must be a power of two. a, b: this proof's vectors (Montgomery, N each,
// overwritten). Emits the MSM terms of every job of the proof and the three final scalars.
// job order per proof: [A], L_0..L_{L-1}, R_0..R_{L-1}, wipA, wipB   (= wire order of include/bpp.h)
// 256 threads, two blocks per SM (<= 128 registers): at 153 registers one block fitted per SM and a batch of 256 proofs ran as
// two waves, the second 27 % full.
// A proof may be given a thread-block CLUSTER instead of one block (launch attribute; 2, 4 or 8 blocks): every loop over the
// vector is then dealt out over the cluster's threads, the two inner products of a round are summed through distributed shared
// memory, and the block barriers between the phases of a round become cluster barriers (release/acquire at cluster scope: the
// folded a, b written by one block are read by the others in the next phase). One or a few large proofs otherwise keep a single
// SM busy for all L rounds of O(N) scalar work: 1.5 ms of the 2.7 ms of one n = 64, m = 64 proof.
template <class M>
__global__ void __launch_bounds__(256, 2) wip_prove_kernel(ProveShape ps, ProveTables T, const uint32_t* __restrict__ params, uint32_t* __restrict__ a_all,
                                 uint32_t* __restrict__ b_all, const uint32_t* __restrict__ alpha_all,
                                 uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx, uint8_t* __restrict__ proofs,
                                 size_t proof_stride, size_t scalar_off) {
  typedef Sc<M> S;
  extern __shared__ uint32_t sh[];
  cg::cluster_group cl = cg::this_cluster();
  const uint32_t cs = cl.num_blocks(), rank = cl.block_rank();  // 1, 0 when launched without a cluster
  __shared__ uint32_t cl_slot[16];
  const uint32_t N = ps.N, L = ps.L, p = blockIdx.x / cs;
  const uint32_t tid = rank * blockDim.x + threadIdx.x, nt = cs * blockDim.x;  // this thread among the proof's threads
  auto phase_sync = [&]() { if (cs > 1) cl.sync(); else __syncthreads(); };
  uint32_t* a = a_all + (size_t)p * N * 8;
  uint32_t* b = b_all + (size_t)p * N * 8;
  uint32_t* osc = out_scalars + (size_t)p * ps.terms_per_proof * 8;
  uint32_t* opi = out_pidx + (size_t)p * ps.terms_per_proof;
  S alpha = ld_mont<M>(alpha_all, p);
  uint32_t base = 0;
  if (ps.has_A) {  // A = alpha*h + sum(bit ? G_i : -H_i)  (range/mod.rs:98-106 / :260-277); the sum is the K7 point, added when the proof is encoded
    if (tid == 0) { sc_store<M>(osc + 0, sc_load<M>(params + 8 * P_ALPHA)); opi[0] = 1; }
    base = 1;
  }
  const uint32_t offL = base, offR = base + L * (N + 2), offA = base + 2 * L * (N + 2), offB = offA + 2 * N + 2;
  for (uint32_t k = 0; k < L; k++) {
    const uint32_t nk = N >> (k + 1);
    S yh = ld_mont<M>(T.rnd, 4 * k), yhi = ld_mont<M>(T.rnd, 4 * k + 1), e = ld_mont<M>(T.rnd, 4 * k + 2), ei = ld_mont<M>(T.rnd, 4 * k + 3);
    // c_L, c_R  (weighted_inner_product_proof.rs:90-91, util.rs:117-127)
    S pl = S::zero(), pr = S::zero();
    for (uint32_t i = tid; i < nk; i += nt) {
      S a1 = ld_mont<M>(a, i), a2 = ld_mont<M>(a, nk + i), b1 = ld_mont<M>(b, i), b2 = ld_mont<M>(b, nk + i);
      pl = pl + a1 * b2 * ld_mont<M>(T.ypow, i);
      pr = pr + a2 * b1 * ld_mont<M>(T.ypow, nk + i);
    }
    S cL = block_sum<M>(pl, sh), cR = block_sum<M>(pr, sh);
    cluster_sum2<M>(cL, cR, cl_slot, cs);
    // L_k, R_k as MSM jobs over the original generators (wip.rs:100-125)
    uint32_t* lsc = osc + (size_t)(offL + k * (N + 2)) * 8;
    uint32_t* lpi = opi + offL + k * (N + 2);
    uint32_t* rsc = osc + (size_t)(offR + k * (N + 2)) * 8;
    uint32_t* rpi = opi + offR + k * (N + 2);
    for (uint32_t j = tid; j < N; j += nt) {
      uint32_t i = j & (nk - 1);
      bool upper = (j / nk) & 1;
      S gc = ld_mont<M>(T.gcoef, (size_t)k * N + j), hc = ld_mont<M>(T.hcoef, (size_t)k * N + j);
      if (upper) {
        sc_store<M>(lsc + 8 * j, yhi * ld_mont<M>(a, i) * gc); lpi[j] = ps.pk_G0 + j;  // G2 with y_nhat^-1 * a1
        sc_store<M>(rsc + 8 * j, ld_mont<M>(b, i) * hc);       rpi[j] = ps.pk_H0 + j;  // H2 with b1
      } else {
        sc_store<M>(lsc + 8 * j, ld_mont<M>(b, nk + i) * hc);      lpi[j] = ps.pk_H0 + j;  // H1 with b2
        sc_store<M>(rsc + 8 * j, yh * ld_mont<M>(a, nk + i) * gc); rpi[j] = ps.pk_G0 + j;  // G1 with y_nhat * a2
      }
    }
    S dL = sc_load<M>(params + 8 * (P_FIXED + L + k)), dR = sc_load<M>(params + 8 * (P_FIXED + 2 * L + k));
    if (tid == 0) {
      sc_store<M>(lsc + 8 * N, cL); lpi[N] = 0; sc_store<M>(lsc + 8 * (N + 1), dL); lpi[N + 1] = 1;
      sc_store<M>(rsc + 8 * N, cR); rpi[N] = 0; sc_store<M>(rsc + 8 * (N + 1), dR); rpi[N + 1] = 1;
    }
    phase_sync();
    // fold a, b  (wip.rs:144-149)
    S yhei = yh * ei;
    for (uint32_t i = tid; i < nk; i += nt) {
      S a1 = ld_mont<M>(a, i), a2 = ld_mont<M>(a, nk + i), b1 = ld_mont<M>(b, i), b2 = ld_mont<M>(b, nk + i);
      st_mont<M>(a, i, a1 * e + a2 * yhei);
      st_mont<M>(b, i, b1 * ei + b2 * e);
    }
    alpha = alpha + (e * e * dL + ei * ei * dR);  // wip.rs:171
    phase_sync();
  }
  // final step (wip.rs:174-216)
  S r = sc_load<M>(params + 8 * P_R), s = sc_load<M>(params + 8 * P_S), delta = sc_load<M>(params + 8 * P_DELTA),
    eta = sc_load<M>(params + 8 * P_ETA), ef = sc_load<M>(params + 8 * P_EFINAL);
  S a0 = ld_mont<M>(a, 0), b0 = ld_mont<M>(b, 0), y0 = ld_mont<M>(T.ypow, 0);
  uint32_t* asc = osc + (size_t)offA * 8;
  uint32_t* api = opi + offA;
  for (uint32_t j = tid; j < N; j += nt) {  // A = r*G[0] + s*H[0] + ... with the fully folded generators
    sc_store<M>(asc + 8 * j, r * ld_mont<M>(T.gcoef, (size_t)L * N + j)); api[j] = ps.pk_G0 + j;
    sc_store<M>(asc + 8 * (N + j), s * ld_mont<M>(T.hcoef, (size_t)L * N + j)); api[N + j] = ps.pk_H0 + j;
  }
  if (tid == 0) {
    sc_store<M>(asc + 8 * (2 * N), r * y0 * b0 + s * y0 * a0); api[2 * N] = 0;  // wip.rs:181
    sc_store<M>(asc + 8 * (2 * N + 1), delta); api[2 * N + 1] = 1;
    uint32_t* bsc = osc + (size_t)offB * 8;
    sc_store<M>(bsc, r * y0 * s); opi[offB] = 0;                               // wip.rs:182,202-208
    sc_store<M>(bsc + 8, eta); opi[offB + 1] = 1;
    uint32_t fin[3][8];
    sc_store<M>(fin[0], r + a0 * ef);                             // wip.rs:214
    sc_store<M>(fin[1], s + b0 * ef);                             // wip.rs:215
    sc_store<M>(fin[2], eta + delta * ef + alpha * ef * ef);      // wip.rs:216
    uint8_t* o = proofs + (size_t)p * proof_stride + scalar_off;
    for (int q = 0; q < 3; q++)
      for (int i = 0; i < 8; i++)
        for (int bb = 0; bb < 4; bb++) o[32 * q + 4 * i + bb] = (uint8_t)(fin[q][i] >> (8 * bb));
  }
}

// K7: S_p = sum_i (bit_i ? G_i : -H_i), one block per proof, written as an extended point
template <class C>
__global__ void __launch_bounds__(128)
range_bits_sum_kernel(ProveShape ps, const uint64_t* __restrict__ values, const uint32_t* __restrict__ tab0,
                      uint32_t* __restrict__ out_ext) {
  __shared__ uint32_t sh[128 * C::EXT_WORDS];
  const uint32_t p = blockIdx.x, tid = threadIdx.x;
  typename C::Ext acc;
  C::set_identity(acc);
  for (uint32_t i = tid; i < ps.N; i += blockDim.x) {
    uint32_t i1 = i % ps.n, i2 = i / ps.n;
    uint64_t v = values[(size_t)p * ps.m + i2];
    bool bit = i1 < 64 ? ((v >> i1) & 1) : false;
    C::madd_ptr(acc, tab0 + (size_t)((bit ? ps.pk_G0 : ps.pk_H0) + i) * C::AFF_WORDS, bit ? 0u : 1u);
  }
  uint32_t* mine = sh + tid * C::EXT_WORDS;
  *reinterpret_cast<typename C::Ext*>(mine) = acc;
  __syncthreads();
  for (uint32_t s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (tid < s) {
      typename C::Ext o = *reinterpret_cast<typename C::Ext*>(sh + (tid + s) * C::EXT_WORDS);
      C::add(acc, o);
      *reinterpret_cast<typename C::Ext*>(mine) = acc;
    }
    __syncthreads();
  }
  if (tid == 0) store_words<C::EXT_WORDS>(out_ext + (size_t)p * C::EXT_WORDS, acc);
}

}  // namespace bpp

// Interactive proving for a transcript-owning host: the same arithmetic as proto_prove.cuh, cut at the
// points where the Fiat-Shamir challenges enter, with every proof of the batch carrying its own challenges and nonces.
//
//   begin            A_p = alpha_p h + sum(bit ? G_i : -H_i)                       range/mod.rs:98-106 / :260-277
//   step 0           host -> (y_p, z_p); a, b, alpha_hat; L_0, R_0                 range/mod.rs:125-172 / :283-376, wip.rs:79-125
//   step k           host -> e_{k-1}; fold a, b, generator coefficients; L_k, R_k  wip.rs:144-171, :90-125
//   step L           host -> e_{L-1}; fold; wip A, B                               wip.rs:174-208
//   finish           host -> e; r', s', delta'                                     wip.rs:214-216
//
// The reference fixes all challenges as constants, so it has no such cut; with the constants fed back in, the bytes are the
// one-shot prover's (tests/test_gpu_session.py checks that, and each proof against the CPU checker under its own challenges).
// State per proof: packed parameter block (layout of proto_verify.cuh: y z e alpha r s delta eta, e_round[L], d_L[L], d_R[L]),
// a, b (Montgomery, N each, folded in place), alpha_hat, y-powers, and the coefficient every ORIGINAL generator carries after
// the rounds folded so far (generators are never folded, see proto_prove.cuh).
#pragma once
#include "proto_prove.cuh"

namespace bpp {

struct FsState {
  uint32_t* params;  // [count][(P_FIXED + 3L) * 8]  canonical
  uint32_t* a;       // [count][N * 8]   Montgomery
  uint32_t* b;       // [count][N * 8]
  uint32_t* alpha;   // [count][8]
  uint32_t* ypow;    // [count][N * 8]   y_p^(i+1)
  uint32_t* gcoef;   // [count][N * 8]
  uint32_t* hcoef;   // [count][N * 8]
  uint32_t* rnd;     // [count][2 * 8]   yhat, yhat^-1 of the round in flight
  uint32_t PW;       // words per parameter block
};

// parameter blocks: nonces = count x (5 + 2L) canonical scalars [alpha, r, s, delta, eta, d_L[0..L), d_R[0..L)], or NULL for the
// reference's constants (alpha 7 / 33, r s delta eta = 33 44 88 123, d_L = 4, d_R = 5). Challenge slots start at 1.
template <class M>
__global__ void fs_params_init_kernel(uint32_t count, uint32_t L, uint32_t single, const uint8_t* __restrict__ nonces, FsState st,
                                      uint8_t* __restrict__ bad) {
  typedef Sc<M> S;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= count) return;
  uint32_t* dst = st.params + (size_t)p * st.PW;
  for (uint32_t i = 0; i < st.PW; i++) dst[i] = 0;
  auto put_small = [&](uint32_t slot, uint32_t v) { dst[8 * slot] = v; };
  put_small(P_Y, 1); put_small(P_Z, 1); put_small(P_EFINAL, 1);
  for (uint32_t k = 0; k < L; k++) put_small(P_FIXED + k, 1);
  if (!nonces) {
    put_small(P_ALPHA, single ? 7 : 33); put_small(P_R, 33); put_small(P_S, 44); put_small(P_DELTA, 88); put_small(P_ETA, 123);
    for (uint32_t k = 0; k < L; k++) { put_small(P_FIXED + L + k, 4); put_small(P_FIXED + 2 * L + k, 5); }
    return;
  }
  const uint8_t* src = nonces + (size_t)p * (5 + 2 * L) * 32;
  bool any_bad = false;
  for (uint32_t j = 0; j < 5 + 2 * L; j++) {
    uint32_t w[8];
    for (int i = 0; i < 8; i++) {
      const uint8_t* q = src + 32 * j + 4 * i;
      w[i] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    }
    if (!S::canonical_ok(w)) { any_bad = true; for (int i = 0; i < 8; i++) w[i] = 0; }
    const uint32_t slot = j < 5 ? P_ALPHA + j : P_FIXED + L + (j - 5);
    for (int i = 0; i < 8; i++) dst[8 * slot + i] = w[i];
  }
  if (any_bad) bad[0] = 1;
}

// challenges from the host: src = count x n_slots canonical scalars, written to slot0 (and slot1). Zero or non-canonical
// challenges are a caller error (FORMAT); 1 is stored instead so that the inversions stay defined.
template <class M>
__global__ void fs_store_challenges_kernel(uint32_t count, uint32_t n_slots, uint32_t slot0, uint32_t slot1, uint32_t invert_mask,
                                           const uint8_t* __restrict__ src, FsState st, uint8_t* __restrict__ bad) {
  typedef Sc<M> S;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= count) return;
  for (uint32_t j = 0; j < n_slots; j++) {
    uint32_t w[8], nz = 0;
    for (int i = 0; i < 8; i++) {
      const uint8_t* q = src + ((size_t)p * n_slots + j) * 32 + 4 * i;
      w[i] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
      nz |= w[i];
    }
    if (!S::canonical_ok(w) || (((invert_mask >> j) & 1) && !nz)) {
      bad[0] = 1;
      for (int i = 0; i < 8; i++) w[i] = i == 0;
    }
    uint32_t* dst = st.params + (size_t)p * st.PW + 8 * (j == 0 ? slot0 : slot1);
    for (int i = 0; i < 8; i++) dst[i] = w[i];
  }
}

// the one MSM term of job A: alpha_p * h
template <class M>
__global__ void fs_alpha_terms_kernel(uint32_t count, FsState st, uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx) {
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= count) return;
  for (int i = 0; i < 8; i++) out_scalars[8 * (size_t)p + i] = st.params[(size_t)p * st.PW + 8 * P_ALPHA + i];
  out_pidx[p] = 1;
}

// block per proof, after (y, z): y-powers, unit coefficients, a, b, alpha_hat
template <class M>
__global__ void __launch_bounds__(256)
fs_init_kernel(ProveShape ps, FsState st, const uint64_t* __restrict__ values, const uint32_t* __restrict__ gammas, uint8_t* __restrict__ bad) {
  typedef Sc<M> S;
  const uint32_t N = ps.N, p = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
  const uint32_t* prm = st.params + (size_t)p * st.PW;
  uint32_t* ypow = st.ypow + (size_t)p * N * 8;
  const S y = sc_load<M>(prm + 8 * P_Y), z = sc_load<M>(prm + 8 * P_Z), one = S::one();
  {
    const uint32_t ch = (N + nt - 1) / nt, i0 = tid * ch, i1 = i0 + ch < N ? i0 + ch : N;
    if (i0 < N) {
      S cur = y.pow_u64(i0 + 1);  // util.rs:34-37
      for (uint32_t i = i0; i < i1; i++) { st_mont<M>(ypow, i, cur); cur = cur * y; }
    }
  }
  for (uint32_t i = tid; i < N; i += nt) {
    st_mont<M>(st.gcoef + (size_t)p * N * 8, i, one);
    st_mont<M>(st.hcoef + (size_t)p * N * 8, i, one);
  }
  __syncthreads();
  const bool single = ps.mode == 1;
  const S zz = z * z, nz = -z, omz = one - z;
  for (uint32_t i = tid; i < N; i += nt) {  // range/mod.rs:125-129, :156-163 / :288-302, :347-364
    const uint32_t i1 = i % ps.n, i2 = i / ps.n;
    S d = S::from_u64(2).pow_u64(i1);
    if (!single) d = d * zz.pow_u64(i2 + 1);
    const S he = d * ld_mont<M>(ypow, N - i - 1) + z;  // d_i y^(N-i) + z
    const uint64_t v = values[(size_t)p * ps.m + i2];
    const bool bit = i1 < 64 ? ((v >> i1) & 1) : false;
    st_mont<M>(st.a + (size_t)p * N * 8, i, bit ? omz : nz);
    st_mont<M>(st.b + (size_t)p * N * 8, i, bit ? he : he - one);
  }
  if (tid == 0) {  // alpha_hat (range/mod.rs:164-172 / :366-376)
    S acc = S::zero();
    bool ok = true;
    for (uint32_t j = 0; j < ps.m; j++) {
      const uint32_t* g = gammas + 8 * ((size_t)p * ps.m + j);
      ok = ok && S::canonical_ok(g);
      const S gm = sc_load<M>(g);
      acc = acc + (single ? gm : zz.pow_u64(j + 1) * gm);
    }
    const S yN1 = ld_mont<M>(ypow, N - 1) * y;
    st_mont<M>(st.alpha, p, sc_load<M>(prm + 8 * P_ALPHA) + acc * yN1);
    if (!ok) bad[0] = 1;
  }
}

// block per proof: fold with e_k (wip.rs:144-171). nk = N >> (k + 1) is the length after the fold.
template <class M>
__global__ void __launch_bounds__(256) fs_fold_kernel(ProveShape ps, FsState st, uint32_t k) {
  typedef Sc<M> S;
  const uint32_t N = ps.N, L = ps.L, p = blockIdx.x, tid = threadIdx.x, nt = blockDim.x, nk = N >> (k + 1);
  const uint32_t* prm = st.params + (size_t)p * st.PW;
  uint32_t* a = st.a + (size_t)p * N * 8;
  uint32_t* b = st.b + (size_t)p * N * 8;
  __shared__ uint32_t sh_ei[8];
  const S e = sc_load<M>(prm + 8 * (P_FIXED + k));
  if (tid == 0) st_mont<M>(sh_ei, 0, e.inv());
  __syncthreads();
  const S ei = ld_mont<M>(sh_ei, 0), yh = ld_mont<M>(st.rnd, (size_t)p * 2), yhi = ld_mont<M>(st.rnd, (size_t)p * 2 + 1);
  const S yhei = yh * ei, yhie = yhi * e;
  // read the pairs first, then write: element i of the folded vectors overwrites half 1
  for (uint32_t i = tid; i < nk; i += nt) {
    S a1 = ld_mont<M>(a, i), a2 = ld_mont<M>(a, nk + i), b1 = ld_mont<M>(b, i), b2 = ld_mont<M>(b, nk + i);
    st_mont<M>(a, i, a1 * e + a2 * yhei);
    st_mont<M>(b, i, b1 * ei + b2 * e);
  }
  for (uint32_t j = tid; j < N; j += nt) {  // wip.rs:151-163 as scalars on the original generators
    const bool upper = (j / nk) & 1;
    uint32_t* gc = st.gcoef + (size_t)p * N * 8;
    uint32_t* hc = st.hcoef + (size_t)p * N * 8;
    st_mont<M>(gc, j, ld_mont<M>(gc, j) * (upper ? yhie : ei));
    st_mont<M>(hc, j, ld_mont<M>(hc, j) * (upper ? ei : e));
  }
  if (tid == 0) {
    const S dL = sc_load<M>(prm + 8 * (P_FIXED + L + k)), dR = sc_load<M>(prm + 8 * (P_FIXED + 2 * L + k));
    st_mont<M>(st.alpha, p, ld_mont<M>(st.alpha, p) + (e * e * dL + ei * ei * dR));  // wip.rs:171
  }
}

// block per proof (blockDim.x a power of two): the two MSM jobs L_k, R_k of N + 2 terms each (wip.rs:90-125)
template <class M>
__global__ void __launch_bounds__(256)
fs_round_kernel(ProveShape ps, FsState st, uint32_t k, uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx) {
  typedef Sc<M> S;
  extern __shared__ uint32_t sh[];
  const uint32_t N = ps.N, L = ps.L, p = blockIdx.x, tid = threadIdx.x, nt = blockDim.x, nk = N >> (k + 1);
  const uint32_t* prm = st.params + (size_t)p * st.PW;
  const uint32_t* a = st.a + (size_t)p * N * 8;
  const uint32_t* b = st.b + (size_t)p * N * 8;
  const uint32_t* ypow = st.ypow + (size_t)p * N * 8;
  const uint32_t* gcoef = st.gcoef + (size_t)p * N * 8;
  const uint32_t* hcoef = st.hcoef + (size_t)p * N * 8;
  __shared__ uint32_t sh_y[16];
  if (tid == 0) {  // weighted_inner_product_proof.rs:98-99
    const S yh = ld_mont<M>(ypow, nk - 1);
    st_mont<M>(sh_y, 0, yh);
    st_mont<M>(sh_y, 1, yh.inv());
    st_mont<M>(st.rnd, (size_t)p * 2, yh);
    st_mont<M>(st.rnd, (size_t)p * 2 + 1, ld_mont<M>(sh_y, 1));
  }
  __syncthreads();
  const S yh = ld_mont<M>(sh_y, 0), yhi = ld_mont<M>(sh_y, 1);
  S pl = S::zero(), pr = S::zero();
  for (uint32_t i = tid; i < nk; i += nt) {
    S a1 = ld_mont<M>(a, i), a2 = ld_mont<M>(a, nk + i), b1 = ld_mont<M>(b, i), b2 = ld_mont<M>(b, nk + i);
    pl = pl + a1 * b2 * ld_mont<M>(ypow, i);
    pr = pr + a2 * b1 * ld_mont<M>(ypow, nk + i);
  }
  const S cL = block_sum<M>(pl, sh), cR = block_sum<M>(pr, sh);
  uint32_t* lsc = out_scalars + (size_t)p * 2 * (N + 2) * 8;
  uint32_t* lpi = out_pidx + (size_t)p * 2 * (N + 2);
  uint32_t* rsc = lsc + (size_t)(N + 2) * 8;
  uint32_t* rpi = lpi + (N + 2);
  for (uint32_t j = tid; j < N; j += nt) {
    const uint32_t i = j & (nk - 1);
    const bool upper = (j / nk) & 1;
    const S gc = ld_mont<M>(gcoef, j), hc = ld_mont<M>(hcoef, j);
    if (upper) {
      sc_store<M>(lsc + 8 * j, yhi * ld_mont<M>(a, i) * gc); lpi[j] = ps.pk_G0 + j;
      sc_store<M>(rsc + 8 * j, ld_mont<M>(b, i) * hc);       rpi[j] = ps.pk_H0 + j;
    } else {
      sc_store<M>(lsc + 8 * j, ld_mont<M>(b, nk + i) * hc);      lpi[j] = ps.pk_H0 + j;
      sc_store<M>(rsc + 8 * j, yh * ld_mont<M>(a, nk + i) * gc); rpi[j] = ps.pk_G0 + j;
    }
  }
  if (tid == 0) {
    const S dL = sc_load<M>(prm + 8 * (P_FIXED + L + k)), dR = sc_load<M>(prm + 8 * (P_FIXED + 2 * L + k));
    sc_store<M>(lsc + 8 * N, cL); lpi[N] = 0; sc_store<M>(lsc + 8 * (N + 1), dL); lpi[N + 1] = 1;
    sc_store<M>(rsc + 8 * N, cR); rpi[N] = 0; sc_store<M>(rsc + 8 * (N + 1), dR); rpi[N + 1] = 1;
  }
}

// block per proof: the jobs of wip A (2N + 2 terms) and wip B (2 terms) (wip.rs:174-208)
template <class M>
__global__ void __launch_bounds__(256)
fs_final_points_kernel(ProveShape ps, FsState st, uint32_t* __restrict__ out_scalars, uint32_t* __restrict__ out_pidx,
                       uint32_t* __restrict__ job_start) {
  typedef Sc<M> S;
  const uint32_t N = ps.N, p = blockIdx.x, tid = threadIdx.x, nt = blockDim.x, per = 2 * N + 4;
  const uint32_t* prm = st.params + (size_t)p * st.PW;
  const S r = sc_load<M>(prm + 8 * P_R), s = sc_load<M>(prm + 8 * P_S), delta = sc_load<M>(prm + 8 * P_DELTA), eta = sc_load<M>(prm + 8 * P_ETA);
  const S a0 = ld_mont<M>(st.a + (size_t)p * N * 8, 0), b0 = ld_mont<M>(st.b + (size_t)p * N * 8, 0), y0 = ld_mont<M>(st.ypow + (size_t)p * N * 8, 0);
  uint32_t* asc = out_scalars + (size_t)p * per * 8;
  uint32_t* api = out_pidx + (size_t)p * per;
  for (uint32_t j = tid; j < N; j += nt) {
    sc_store<M>(asc + 8 * j, r * ld_mont<M>(st.gcoef + (size_t)p * N * 8, j)); api[j] = ps.pk_G0 + j;
    sc_store<M>(asc + 8 * (N + j), s * ld_mont<M>(st.hcoef + (size_t)p * N * 8, j)); api[N + j] = ps.pk_H0 + j;
  }
  if (tid == 0) {
    sc_store<M>(asc + 8 * (2 * N), r * y0 * b0 + s * y0 * a0); api[2 * N] = 0;  // wip.rs:181
    sc_store<M>(asc + 8 * (2 * N + 1), delta); api[2 * N + 1] = 1;
    sc_store<M>(asc + 8 * (2 * N + 2), r * y0 * s); api[2 * N + 2] = 0;         // wip.rs:182, :202-208
    sc_store<M>(asc + 8 * (2 * N + 3), eta); api[2 * N + 3] = 1;
    job_start[2 * p] = p * per;
    job_start[2 * p + 1] = p * per + 2 * N + 2;
    if (p == gridDim.x - 1) job_start[2 * p + 2] = (p + 1) * per;
  }
}

// thread per proof: r', s', delta' (wip.rs:214-216) into the proof
template <class M>
__global__ void fs_finish_kernel(ProveShape ps, FsState st, uint8_t* __restrict__ proofs, size_t proof_stride, size_t scalar_off) {
  typedef Sc<M> S;
  uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= ps.n_proofs) return;
  const uint32_t N = ps.N;
  const uint32_t* prm = st.params + (size_t)p * st.PW;
  const S r = sc_load<M>(prm + 8 * P_R), s = sc_load<M>(prm + 8 * P_S), delta = sc_load<M>(prm + 8 * P_DELTA), eta = sc_load<M>(prm + 8 * P_ETA),
          ef = sc_load<M>(prm + 8 * P_EFINAL);
  const S a0 = ld_mont<M>(st.a + (size_t)p * N * 8, 0), b0 = ld_mont<M>(st.b + (size_t)p * N * 8, 0), alpha = ld_mont<M>(st.alpha, p);
  uint32_t fin[3][8];
  sc_store<M>(fin[0], r + a0 * ef);
  sc_store<M>(fin[1], s + b0 * ef);
  sc_store<M>(fin[2], eta + delta * ef + alpha * ef * ef);
  uint8_t* o = proofs + (size_t)p * proof_stride + scalar_off;
  for (int q = 0; q < 3; q++)
    for (int i = 0; i < 8; i++)
      for (int bb = 0; bb < 4; bb++) o[32 * q + 4 * i + bb] = (uint8_t)(fin[q][i] >> (8 * bb));
}

// job results -> encoded points, into the proofs (job k of proof p at point slot slot0 / slot1) and into a compact array for the host
template <class C>
__global__ void fs_encode_kernel(uint32_t n_proofs, uint32_t jobs_per_proof, const uint32_t* __restrict__ results, const uint32_t* __restrict__ add_to_job0,
                                 uint32_t slot0, uint32_t slot1, uint8_t* __restrict__ proofs, size_t proof_stride, uint8_t* __restrict__ compact) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_proofs * jobs_per_proof) return;
  const uint32_t p = t / jobs_per_proof, k = t % jobs_per_proof;
  typename C::Ext r;
  load_words<C::EXT_WORDS>(r, results + (size_t)t * C::EXT_WORDS);
  if (k == 0 && add_to_job0) {
    typename C::Ext s;
    load_words<C::EXT_WORDS>(s, add_to_job0 + (size_t)p * C::EXT_WORDS);
    C::add(r, s);
  }
  uint8_t* dst = proofs + (size_t)p * proof_stride + (size_t)(k == 0 ? slot0 : slot1) * C::POINT_BYTES;
  C::encode(dst, r);
  for (int i = 0; i < C::POINT_BYTES; i++) compact[(size_t)t * C::POINT_BYTES + i] = dst[i];
}

}  // namespace bpp

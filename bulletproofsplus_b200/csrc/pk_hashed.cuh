// Hash-derived generators for Ristretto255: bpp_pk_new_hashed.
//
// The reference's PublicKey::new (publickey.rs:21-48) sets g, h = 1g, 2g and G_i, H_i = 3(i+1)g, 5(i+1)g: every discrete
// logarithm is known, which is what lets the exponent-domain checker work and what makes the key unusable outside parity
// tests (a prover who knows the relations between the generators can open a commitment to any value). Here every generator
// is the RFC 9496 §4.3.4 element derivation of 64 bytes of SHA-512 output (hashed on the host, sha512.hpp), so no relation
// between them is known to anybody. Parity: tests/pyref.py (pinned to the RFC 9496 appendix A.3 vector).
#pragma once
#include "ed25519.cuh"

namespace bpp {

// MAP of RFC 9496 §4.3.4 -> extended coordinates
static __device__ __noinline__ void ristretto_map(Ed25519::Ext& out, const Fe25519& t) {
  typedef Fe25519 F;
  // 1 - d^2, (d - 1)^2, sqrt(a d - 1): little-endian 32-bit words of the constants of RFC 9496 §4.1
  const uint32_t c_one_minus_d_sq[8] = {0x945fc176u, 0xe27c09c1u, 0xcd5e350fu, 0x2c81a138u, 0xbe70dfe4u, 0x9994abddu, 0xb2b3e0d7u, 0x029072a8u};
  const uint32_t c_d_minus_one_sq[8] = {0x44ed4d20u, 0x31ad5aaau, 0xb01e1999u, 0xd29e4a2cu, 0x529b4eebu, 0x4cdcd32fu, 0xf66c2241u, 0x5968b37au};
  const uint32_t c_sqrt_ad_minus_one[8] = {0x497b2e1bu, 0x7e97f6a0u, 0x1b7854bdu, 0xaf9d8e0cu, 0x31f5d1fdu, 0x0f3cfcc9u, 0x2b8348acu, 0x376931bfu};
  const F one_minus_d_sq = Ed25519::constant(c_one_minus_d_sq), d_minus_one_sq = Ed25519::constant(c_d_minus_one_sq);
  const F sqrt_ad_minus_one = Ed25519::constant(c_sqrt_ad_minus_one), d = Ed25519::k_d(), sqrt_m1 = Ed25519::k_sqrt_m1();
  F one, r, u, v, s, sp, c, N, w0, w1, w2, w3, t0, t1;
  fe_one(one);
  fe_sqr(t0, t); fe_mul(r, sqrt_m1, t0);                 // r = sqrt(-1) t^2
  fe_add(t0, r, one); fe_mul(u, t0, one_minus_d_sq);     // u = (r + 1)(1 - d^2)
  fe_mul(t0, r, d); fe_add(t0, t0, one); fe_neg(t0, t0); // -1 - r d
  fe_add(t1, r, d); fe_mul(v, t0, t1);                   // v = (-1 - r d)(r + d)
  const bool was_square = Ed25519::sqrt_ratio_m1(s, u, v);
  fe_mul(sp, s, t);
  fe_cneg(sp, !fe_is_neg(sp));                           // s' = -|s t|
  fe_cmov(s, sp, !was_square);
  fe_neg(c, one);
  fe_cmov(c, r, !was_square);                            // c = -1 or r
  fe_sub(t0, r, one); fe_mul(t0, c, t0); fe_mul(t0, t0, d_minus_one_sq); fe_sub(N, t0, v);  // N = c (r - 1)(d - 1)^2 - v
  fe_mul(t0, s, v); fe_dbl(w0, t0);                      // w0 = 2 s v
  fe_mul(w1, N, sqrt_ad_minus_one);
  fe_sqr(t0, s); fe_sub(w2, one, t0); fe_add(w3, one, t0);
  fe_mul(out.X, w0, w3); fe_mul(out.Y, w2, w1); fe_mul(out.Z, w1, w3); fe_mul(out.T, w0, w2);
}

// seeds: n x 64 uniform bytes -> n affine table points
static __global__ void pk_gen_hashed_kernel(uint32_t n, const uint8_t* __restrict__ seeds, uint32_t* __restrict__ table) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Fe25519 t[2];
  for (int h = 0; h < 2; h++) {
    const uint8_t* b = seeds + (size_t)i * 64 + 32 * h;
    for (int k = 0; k < 8; k++)
      t[h].v[k] = (uint32_t)b[4 * k] | ((uint32_t)b[4 * k + 1] << 8) | ((uint32_t)b[4 * k + 2] << 16) | ((uint32_t)b[4 * k + 3] << 24);
    t[h].v[7] &= 0x7fffffffu;  // the top bit is masked; the loose representation takes any value below 2^255 as it is
  }
  Ed25519::Ext p, q;
  ristretto_map(p, t[0]);
  ristretto_map(q, t[1]);
  Ed25519::add(p, q);
  Ed25519::Aff a;
  Ed25519::to_aff(a, p);
  store_words<Ed25519::AFF_WORDS>(table + (size_t)i * Ed25519::AFF_WORDS, a);
}

}  // namespace bpp

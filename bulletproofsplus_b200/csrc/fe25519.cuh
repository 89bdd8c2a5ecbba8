// GF(2^255 - 19) for sm_100a: 8 x 32-bit saturated limbs, "loose" representation
// (any value in [0, 2^256) congruent to the element; 2^256 == 38). The 8x8 schoolbook product is
// written as mad.lo.cc / madc.hi.cc pairs on even/odd columns, which ptxas fuses into one
// IMAD.WIDE.U32(.X) per 32x32 product with the carry in a predicate (64 wide IMADs per product,
// checked with cuobjdump -sass). No tensor cores: this is an integer carry chain, not a contraction.
//
// This is the B200 stand-in for the reference's L0 field layer (mcl Fp reached through mcl_rust;
// call sites listed in SURVEY.md §8(b)).
#pragma once
#include <cstdint>
#include "common.cuh"

namespace bpp {

struct Fe25519 {
  uint32_t v[8];
};

// ---- fold a 16-limb value to 8 loose limbs: lo + 38*hi ----------------------------------
BPP_DEV void fe25519_reduce_wide(Fe25519& out, const uint32_t* t) {
  uint32_t r[8], q[8], top;
#pragma unroll
  for (int i = 0; i < 8; i++) r[i] = t[i];
  const uint32_t k38 = 38;
  // even hi limbs: accumulate 38*t[8+j] into (r[j], r[j+1]) as one carry chain
  asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[0]), "+r"(r[1]) : "r"(t[8]), "r"(k38));
  asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[2]), "+r"(r[3]) : "r"(t[10]), "r"(k38));
  asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[4]), "+r"(r[5]) : "r"(t[12]), "r"(k38));
  asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;" : "+r"(r[6]), "+r"(r[7]) : "r"(t[14]), "r"(k38));
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(top));
  // odd hi limbs: independent wide products on columns 1..8
#pragma unroll
  for (int j = 0; j < 4; j++)
    asm volatile("mul.lo.u32 %0, %2, %3; mul.hi.u32 %1, %2, %3;" : "=r"(q[2 * j]), "=r"(q[2 * j + 1]) : "r"(t[9 + 2 * j]), "r"(k38));
  asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(r[1]) : "r"(q[0]));
#pragma unroll
  for (int i = 2; i < 8; i++) asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(q[i - 1]));
  asm volatile("addc.u32 %0, %0, %1;" : "+r"(top) : "r"(q[7]));
  // second fold: top < 2^7
  uint32_t f = top * 38u, c2;
  asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(r[0]) : "r"(f));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("addc.cc.u32 %0, %0, 0;" : "+r"(r[i]));
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(c2));
  r[0] += c2 * 38u;  // after a wrap the value is < 2^13: cannot carry again
#pragma unroll
  for (int i = 0; i < 8; i++) out.v[i] = r[i];
}

BPP_DEV void fe_mul(Fe25519& r, const Fe25519& a, const Fe25519& b) {
  uint32_t t[16];
#ifdef BPP_FE_KARATSUBA
  mul_wide_8_karatsuba(t, a.v, b.v);
#else
  mul_wide_n<8>(t, a.v, b.v);
#endif
  fe25519_reduce_wide(r, t);
}
// 36 wide products (28 off-diagonal, doubled by a shift, + 8 squares) instead of 64, then the same fold
BPP_DEV void fe_sqr(Fe25519& r, const Fe25519& a) {
  uint32_t t[16];
  sqr_wide_n<8>(t, a.v);
  fe25519_reduce_wide(r, t);
}

BPP_DEV void fe_add(Fe25519& r, const Fe25519& a, const Fe25519& b) {
  uint32_t t[8], c;
  asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(t[0]) : "r"(a.v[0]), "r"(b.v[0]));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(t[i]) : "r"(a.v[i]), "r"(b.v[i]));
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(c));
  uint32_t f = c * 38u, c2;
  asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(t[0]) : "r"(f));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("addc.cc.u32 %0, %0, 0;" : "+r"(t[i]));
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(c2));
  t[0] += c2 * 38u;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = t[i];
}

BPP_DEV void fe_sub(Fe25519& r, const Fe25519& a, const Fe25519& b) {
  uint32_t t[8], bw;
  asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(t[0]) : "r"(a.v[0]), "r"(b.v[0]));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(t[i]) : "r"(a.v[i]), "r"(b.v[i]));
  asm volatile("subc.u32 %0, 0, 0;" : "=r"(bw));  // 0 or 0xffffffff
  uint32_t f = bw & 38u, bw2;
  asm volatile("sub.cc.u32 %0, %0, %1;" : "+r"(t[0]) : "r"(f));
#pragma unroll
  for (int i = 1; i < 8; i++) asm volatile("subc.cc.u32 %0, %0, 0;" : "+r"(t[i]));
  asm volatile("subc.u32 %0, 0, 0;" : "=r"(bw2));
  t[0] -= bw2 & 38u;  // after a second wrap the value is within 38 of 2^256: cannot borrow again
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = t[i];
}

BPP_DEV void fe_zero(Fe25519& r) {
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = 0;
}
BPP_DEV void fe_one(Fe25519& r) { fe_zero(r); r.v[0] = 1; }
BPP_DEV void fe_neg(Fe25519& r, const Fe25519& a) { Fe25519 z; fe_zero(z); fe_sub(r, z, a); }
BPP_DEV void fe_dbl(Fe25519& r, const Fe25519& a) { fe_add(r, a, a); }

// canonical representative in [0, p)
BPP_DEV void fe_canon(Fe25519& r, const Fe25519& a) {
  uint32_t t[8];
#pragma unroll
  for (int i = 0; i < 8; i++) t[i] = a.v[i];
  // fold bit 255: 2^255 == 19
  uint32_t hi = t[7] >> 31;
  t[7] &= 0x7fffffffu;
  uint32_t f = hi * 19u;
  asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(t[0]) : "r"(f));
#pragma unroll
  for (int i = 1; i < 7; i++) asm volatile("addc.cc.u32 %0, %0, 0;" : "+r"(t[i]));
  asm volatile("addc.u32 %0, %0, 0;" : "+r"(t[7]));
  // now t < 2^255 + 19; subtract p if t >= p  <=>  t + 19 >= 2^255
  uint32_t u[8];
  asm volatile("add.cc.u32 %0, %1, 19;" : "=r"(u[0]) : "r"(t[0]));
#pragma unroll
  for (int i = 1; i < 7; i++) asm volatile("addc.cc.u32 %0, %1, 0;" : "=r"(u[i]) : "r"(t[i]));
  asm volatile("addc.u32 %0, %1, 0;" : "=r"(u[7]) : "r"(t[7]));
  bool ge = (u[7] >> 31) != 0;
  u[7] &= 0x7fffffffu;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = ge ? u[i] : t[i];
}

BPP_DEV bool fe_is_zero(const Fe25519& a) {
  Fe25519 c;
  fe_canon(c, a);
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) o |= c.v[i];
  return o == 0;
}
BPP_DEV bool fe_is_neg(const Fe25519& a) {  // RFC 9496 IS_NEGATIVE: low bit of the canonical value
  Fe25519 c;
  fe_canon(c, a);
  return c.v[0] & 1u;
}
BPP_DEV bool fe_eq(const Fe25519& a, const Fe25519& b) {
  Fe25519 d;
  fe_sub(d, a, b);
  return fe_is_zero(d);
}
BPP_DEV void fe_cmov(Fe25519& r, const Fe25519& a, bool c) {
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = c ? a.v[i] : r.v[i];
}
BPP_DEV void fe_cneg(Fe25519& r, bool c) {
  Fe25519 n;
  fe_neg(n, r);
  fe_cmov(r, n, c);
}
BPP_DEV void fe_nsqr(Fe25519& r, const Fe25519& a, int n) {
  r = a;
  for (int i = 0; i < n; i++) fe_sqr(r, r);
}

// z^(2^252 - 3) and, when INV, z^(p - 2); shared addition chain
template <bool INV>
__device__ __noinline__ void fe_pow_chain(Fe25519& out, const Fe25519& z) {
  Fe25519 z2, z9, z11, z_5_0, z_10_0, z_20_0, z_50_0, z_100_0, t;
  fe_sqr(z2, z);
  fe_nsqr(t, z2, 2);
  fe_mul(z9, t, z);
  fe_mul(z11, z9, z2);
  fe_sqr(t, z11);
  fe_mul(z_5_0, t, z9);
  fe_nsqr(t, z_5_0, 5);
  fe_mul(z_10_0, t, z_5_0);
  fe_nsqr(t, z_10_0, 10);
  fe_mul(z_20_0, t, z_10_0);
  fe_nsqr(t, z_20_0, 20);
  fe_mul(t, t, z_20_0);
  fe_nsqr(t, t, 10);
  fe_mul(z_50_0, t, z_10_0);
  fe_nsqr(t, z_50_0, 50);
  fe_mul(z_100_0, t, z_50_0);
  fe_nsqr(t, z_100_0, 100);
  fe_mul(t, t, z_100_0);
  fe_nsqr(t, t, 50);
  fe_mul(t, t, z_50_0);  // z^(2^250 - 1)
  if (INV) {
    fe_nsqr(t, t, 5);
    fe_mul(out, t, z11);  // 2^255 - 21
  } else {
    fe_nsqr(t, t, 2);
    fe_mul(out, t, z);  // 2^252 - 3
  }
}
BPP_DEV void fe_invert(Fe25519& r, const Fe25519& a) { fe_pow_chain<true>(r, a); }
BPP_DEV void fe_pow22523(Fe25519& r, const Fe25519& a) { fe_pow_chain<false>(r, a); }

// 32 little-endian bytes -> element; returns false if the encoding is not canonical (>= p)
BPP_DEV bool fe_from_bytes(Fe25519& r, const uint8_t* b) {
#pragma unroll
  for (int i = 0; i < 8; i++)
    r.v[i] = (uint32_t)b[4 * i] | ((uint32_t)b[4 * i + 1] << 8) | ((uint32_t)b[4 * i + 2] << 16) | ((uint32_t)b[4 * i + 3] << 24);
  // canonical <=> r < p = 2^255 - 19
  if (r.v[7] >> 31) return false;
  bool all_ones = r.v[7] == 0x7fffffffu;
#pragma unroll
  for (int i = 1; i < 7; i++) all_ones = all_ones && (r.v[i] == 0xffffffffu);
  if (all_ones && r.v[0] >= 0xffffffedu) return false;
  return true;
}
BPP_DEV void fe_to_bytes(uint8_t* b, const Fe25519& a) {
  Fe25519 c;
  fe_canon(c, a);
#pragma unroll
  for (int i = 0; i < 8; i++) {
    b[4 * i] = (uint8_t)c.v[i];
    b[4 * i + 1] = (uint8_t)(c.v[i] >> 8);
    b[4 * i + 2] = (uint8_t)(c.v[i] >> 16);
    b[4 * i + 3] = (uint8_t)(c.v[i] >> 24);
  }
}


}  // namespace bpp

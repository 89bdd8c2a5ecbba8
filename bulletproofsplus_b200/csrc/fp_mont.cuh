// Generic prime field for sm_100a: NL x 32-bit limbs, Montgomery form, always fully reduced. Used for the base fields
// of the short-Weierstrass curves (secp256k1: NL = 8, BLS12-381: NL = 12) — the B200 stand-in for mcl's Fp behind
// the reference's `Point(G1)` (src/bls12_381/building_block/point/point.rs) and for the field of the reference's
// own secp256k1 module (src/secp256k1/building_block/field/prime_field_elem.rs).
// The NL x NL product uses the same even/odd-column mad.lo.cc / madc.hi.cc chains as fe25519.cuh (one IMAD.WIDE.U32
// per 32x32 product); the Montgomery reduction is row-pipelined 64-bit C. No tensor cores: carry chains.
// A parameter set with SPECIAL_K1 = true (secp256k1: p = 2^256 - 2^32 - 977) keeps plain residues instead (r1 = r2 = 1): its
// reduction folds the high half down with 2^256 = 2^32 + 977 (mod p) — 8 wide products instead of the 64 of a Montgomery
// reduction, 73 per field multiplication instead of 128.
#pragma once
#include <cstdint>
#include "common.cuh"

namespace bpp {

// P supplies: NL, mod(i), r1(i) (= 2^(32 NL) mod p), r2(i) (= 2^(64 NL) mod p), NINV (= -p^-1 mod 2^32), SPECIAL_K1
template <class P>
struct Fp {
  static constexpr int NL = P::NL;
  static constexpr int BYTES = 4 * NL;
  uint32_t v[NL];

  BPP_DEV static Fp zero() { Fp r; for (int i = 0; i < NL; i++) r.v[i] = 0; return r; }
  BPP_DEV static Fp one() { Fp r; for (int i = 0; i < NL; i++) r.v[i] = P::r1(i); return r; }
  BPP_DEV static bool geq_mod(const uint32_t* a) {
#pragma unroll
    for (int i = NL - 1; i >= 0; i--) {
      uint32_t m = P::mod(i);
      if (a[i] != m) return a[i] > m;
    }
    return true;
  }
  BPP_DEV static void sub_mod(uint32_t* a) {
    uint64_t bw = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) {
      uint64_t d = (uint64_t)a[i] - P::mod(i) - bw;
      a[i] = (uint32_t)d;
      bw = (d >> 32) & 1;
    }
  }
  BPP_DEV static Fp add(const Fp& a, const Fp& b) {
    Fp r;
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) { c += (uint64_t)a.v[i] + b.v[i]; r.v[i] = (uint32_t)c; c >>= 32; }
    if (c || geq_mod(r.v)) sub_mod(r.v);
    return r;
  }
  BPP_DEV static Fp sub(const Fp& a, const Fp& b) {
    Fp r;
    uint64_t bw = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) {
      uint64_t d = (uint64_t)a.v[i] - b.v[i] - bw;
      r.v[i] = (uint32_t)d;
      bw = (d >> 32) & 1;
    }
    if (bw) {
      uint64_t c = 0;
#pragma unroll
      for (int i = 0; i < NL; i++) { c += (uint64_t)r.v[i] + P::mod(i); r.v[i] = (uint32_t)c; c >>= 32; }
    }
    return r;
  }
  BPP_DEV static Fp neg(const Fp& a) { return sub(zero(), a); }
  BPP_DEV static Fp dbl(const Fp& a) { return add(a, a); }
  // t[0..2NL) < p * 2^(32 NL)  ->  t * 2^(-32 NL) mod p
  BPP_DEV static Fp mont_reduce(uint32_t* t) {
    uint32_t extra = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) {
      const uint32_t m = t[i] * P::NINV;
      uint64_t c = 0;
#pragma unroll
      for (int j = 0; j < NL; j++) {
        c += (uint64_t)m * P::mod(j) + t[i + j];
        t[i + j] = (uint32_t)c;
        c >>= 32;
      }
      c += (uint64_t)t[i + NL] + extra;
      t[i + NL] = (uint32_t)c;
      extra = (uint32_t)(c >> 32);
    }
    if (extra || geq_mod(t + NL)) sub_mod(t + NL);
    Fp r;
#pragma unroll
    for (int i = 0; i < NL; i++) r.v[i] = t[NL + i];
    return r;
  }
  // secp256k1 only: t[0..16) (any 512-bit value) -> t mod p, fully reduced. First fold: lo + hi * 977 + (hi << 32), whose
  // overflow above 2^256 is below 2^34; second fold adds overflow * (2^32 + 977); a last carry (possible only when the sum
  // wrapped to a tiny value) adds 2^32 + 977 once more. Checked limb for limb against Python integers on 5 * 10^5 inputs
  // incl. all-ones and near-wrap cases.
  BPP_DEV static Fp k1_reduce(const uint32_t* t) {
    static_assert(!P::SPECIAL_K1 || NL == 8, "special reduction is written for 8 limbs");
    Fp r;
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < NL; i++) {
      c += (uint64_t)t[i] + (uint64_t)t[NL + i] * 977u + (i > 0 ? t[NL + i - 1] : 0u);
      r.v[i] = (uint32_t)c;
      c >>= 32;
    }
    c += t[2 * NL - 1];
    uint64_t d = (uint64_t)r.v[0] + c * 977u;
    r.v[0] = (uint32_t)d;
    d >>= 32;
    d += (uint64_t)r.v[1] + c;
    r.v[1] = (uint32_t)d;
    d >>= 32;
#pragma unroll
    for (int i = 2; i < NL; i++) { d += r.v[i]; r.v[i] = (uint32_t)d; d >>= 32; }
    if (d) {
      uint64_t e = (uint64_t)r.v[0] + 977u;
      r.v[0] = (uint32_t)e;
      e >>= 32;
      e += (uint64_t)r.v[1] + 1u;
      r.v[1] = (uint32_t)e;
      e >>= 32;
#pragma unroll
      for (int i = 2; i < NL; i++) { e += r.v[i]; r.v[i] = (uint32_t)e; e >>= 32; }
    }
    if (geq_mod(r.v)) sub_mod(r.v);
    return r;
  }
  BPP_DEV static Fp reduce_wide(uint32_t* t) {
    if constexpr (P::SPECIAL_K1) return k1_reduce(t);
    else return mont_reduce(t);
  }
  BPP_DEV static Fp mul_inl(const Fp& a, const Fp& b) {
    uint32_t t[2 * NL];
    mul_wide_n<NL>(t, a.v, b.v);
    return reduce_wide(t);
  }
  BPP_DEV static Fp sqr_inl(const Fp& a) {  // NL (NL + 1) / 2 wide products instead of NL^2
    uint32_t t[2 * NL];
    sqr_wide_n<NL>(t, a.v);
    return reduce_wide(t);
  }
  // P::MUL_IS_FUNCTION: the product and the square are real functions (operands and result in registers). A mixed addition
  // on a short-Weierstrass curve is 14 field multiplications; inlined, the accumulation loop was 50 KB (secp256k1) to 200 KB
  // (BLS12-381) of instructions and stalled on instruction fetch (`no_instruction` 1.2-1.9 warp-cycles per issued
  // instruction in the r02af captures).
  __device__ __noinline__ static Fp mul_fn(Fp a, Fp b) { return mul_inl(a, b); }
  __device__ __noinline__ static Fp sqr_fn(Fp a) { return sqr_inl(a); }
  BPP_DEV static Fp mul(const Fp& a, const Fp& b) {
    if constexpr (P::MUL_IS_FUNCTION) return mul_fn(a, b);
    else return mul_inl(a, b);
  }
  BPP_DEV static Fp sqr(const Fp& a) {
    if constexpr (P::MUL_IS_FUNCTION) return sqr_fn(a);
    else return sqr_inl(a);
  }
  BPP_DEV bool is_zero() const { uint32_t o = 0; for (int i = 0; i < NL; i++) o |= v[i]; return o == 0; }
  BPP_DEV static bool eq(const Fp& a, const Fp& b) { uint32_t o = 0; for (int i = 0; i < NL; i++) o |= a.v[i] ^ b.v[i]; return o == 0; }
  BPP_DEV static Fp from_canonical(const uint32_t* limbs) {
    Fp a;
#pragma unroll
    for (int i = 0; i < NL; i++) a.v[i] = limbs[i];
    if constexpr (P::SPECIAL_K1) {
      return a;  // plain residues
    } else {
      Fp r2;
#pragma unroll
      for (int i = 0; i < NL; i++) r2.v[i] = P::r2(i);
      return mul(a, r2);
    }
  }
  BPP_DEV void to_canonical(uint32_t* limbs) const {
    if constexpr (P::SPECIAL_K1) {
#pragma unroll
      for (int i = 0; i < NL; i++) limbs[i] = v[i];
    } else {
      uint32_t t[2 * NL];
#pragma unroll
      for (int i = 0; i < NL; i++) { t[i] = v[i]; t[NL + i] = 0; }
      Fp r = mont_reduce(t);
#pragma unroll
      for (int i = 0; i < NL; i++) limbs[i] = r.v[i];
    }
  }
  BPP_DEV static Fp from_u32(uint32_t x) {
    uint32_t l[NL];
    for (int i = 0; i < NL; i++) l[i] = 0;
    l[0] = x;
    return from_canonical(l);
  }
  // little-endian canonical bytes; false if >= p
  BPP_DEV static bool from_bytes(Fp& r, const uint8_t* b) {
    uint32_t l[NL];
#pragma unroll
    for (int i = 0; i < NL; i++)
      l[i] = (uint32_t)b[4 * i] | ((uint32_t)b[4 * i + 1] << 8) | ((uint32_t)b[4 * i + 2] << 16) | ((uint32_t)b[4 * i + 3] << 24);
    if (geq_mod(l)) { r = zero(); return false; }
    r = from_canonical(l);
    return true;
  }
  BPP_DEV void to_bytes(uint8_t* b) const {
    uint32_t l[NL];
    to_canonical(l);
#pragma unroll
    for (int i = 0; i < NL; i++) {
      b[4 * i] = (uint8_t)l[i];
      b[4 * i + 1] = (uint8_t)(l[i] >> 8);
      b[4 * i + 2] = (uint8_t)(l[i] >> 16);
      b[4 * i + 3] = (uint8_t)(l[i] >> 24);
    }
  }
  // Fermat inverse a^(p-2); inv(0) = 0
  __device__ __noinline__ static Fp inv(const Fp& a) {
    uint32_t e[NL];
    uint64_t bw = 2;
#pragma unroll
    for (int i = 0; i < NL; i++) {
      uint64_t d = (uint64_t)P::mod(i) - bw;
      e[i] = (uint32_t)d;
      bw = (d >> 32) & 1;
    }
    Fp r = one();
    for (int i = NL - 1; i >= 0; i--)
      for (int b = 31; b >= 0; b--) {
        r = sqr(r);
        if ((e[i] >> b) & 1) r = mul(r, a);
      }
    return r;
  }
};

}  // namespace bpp

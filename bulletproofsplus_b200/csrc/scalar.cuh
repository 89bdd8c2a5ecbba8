// Scalar fields (group orders) for sm_100a: 8 x 32-bit limbs, Montgomery form, one template for
// ell (Ristretto255), n (secp256k1) and r (BLS12-381). Device stand-in for the reference's
// PrimeFieldElem (src/bls12_381/building_block/scalar/prime_field_elem.rs:12-248: new(i32) :191-195,
// Add/Sub/Mul :51-120, Neg :158-182, inv :229-236). Scalar work is < 1 % of the path,
// but its latency sits on the critical path of the challenge-only table kernels: the product reuses the base field's
// IMAD.WIDE carry-chain schoolbook, the Montgomery reduction is 64-bit C that nvcc lowers to IMAD.WIDE.U32.
#pragma once
#include <cstdint>
#include "common.cuh"

namespace bpp {

struct ModEll {
  static constexpr int BITS = 253;
  BPP_DEV static uint32_t mod(int i) { const uint32_t c[8] = {0x5cf5d3edu, 0x5812631au, 0xa2f79cd6u, 0x14def9deu, 0x00000000u, 0x00000000u, 0x00000000u, 0x10000000u}; return c[i]; }
  BPP_DEV static uint32_t r1(int i) { const uint32_t c[8] = {0x8d98951du, 0xd6ec3174u, 0x737dcf70u, 0xc6ef5bf4u, 0xfffffffeu, 0xffffffffu, 0xffffffffu, 0x0fffffffu}; return c[i]; }
  BPP_DEV static uint32_t r2(int i) { const uint32_t c[8] = {0x449c0f01u, 0xa40611e3u, 0x68859347u, 0xd00e1ba7u, 0x17f5be65u, 0xceec73d2u, 0x7c309a3du, 0x0399411bu}; return c[i]; }
  static constexpr uint32_t NINV = 0x12547e1bu;
};
struct ModSecpN {
  static constexpr int BITS = 256;
  BPP_DEV static uint32_t mod(int i) { const uint32_t c[8] = {0xd0364141u, 0xbfd25e8cu, 0xaf48a03bu, 0xbaaedce6u, 0xfffffffeu, 0xffffffffu, 0xffffffffu, 0xffffffffu}; return c[i]; }
  BPP_DEV static uint32_t r1(int i) { const uint32_t c[8] = {0x2fc9bebfu, 0x402da173u, 0x50b75fc4u, 0x45512319u, 0x00000001u, 0x00000000u, 0x00000000u, 0x00000000u}; return c[i]; }
  BPP_DEV static uint32_t r2(int i) { const uint32_t c[8] = {0x67d7d140u, 0x896cf214u, 0x0e7cf878u, 0x741496c2u, 0x5bcd07c6u, 0xe697f5e4u, 0x81c69bc5u, 0x9d671cd5u}; return c[i]; }
  static constexpr uint32_t NINV = 0x5588b13fu;
};
struct ModBlsR {
  static constexpr int BITS = 255;
  BPP_DEV static uint32_t mod(int i) { const uint32_t c[8] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u, 0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u}; return c[i]; }
  BPP_DEV static uint32_t r1(int i) { const uint32_t c[8] = {0xfffffffeu, 0x00000001u, 0x00034802u, 0x5884b7fau, 0xecbc4ff5u, 0x998c4fefu, 0xacc5056fu, 0x1824b159u}; return c[i]; }
  BPP_DEV static uint32_t r2(int i) { const uint32_t c[8] = {0xf3f29c6du, 0xc999e990u, 0x87925c23u, 0x2b6cedcbu, 0x7254398fu, 0x05d31496u, 0x9f59ff11u, 0x0748d9d9u}; return c[i]; }
  static constexpr uint32_t NINV = 0xffffffffu;
};

// Sc<M>: an element of Z/M in Montgomery form, always fully reduced.
template <class M>
struct Sc {
  uint32_t v[8];

  BPP_DEV static Sc zero() { Sc r; for (int i = 0; i < 8; i++) r.v[i] = 0; return r; }
  BPP_DEV static Sc one() { Sc r; for (int i = 0; i < 8; i++) r.v[i] = M::r1(i); return r; }
  BPP_DEV static bool geq_mod(const uint32_t* a) {
#pragma unroll
    for (int i = 7; i >= 0; i--) {
      uint32_t m = M::mod(i);
      if (a[i] != m) return a[i] > m;
    }
    return true;
  }
  BPP_DEV static void sub_mod(uint32_t* a) {
    uint64_t bw = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      uint64_t d = (uint64_t)a[i] - M::mod(i) - bw;
      a[i] = (uint32_t)d;
      bw = (d >> 32) & 1;
    }
  }
  BPP_DEV friend Sc operator+(const Sc& a, const Sc& b) {
    Sc r;
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { c += (uint64_t)a.v[i] + b.v[i]; r.v[i] = (uint32_t)c; c >>= 32; }
    if (c || geq_mod(r.v)) sub_mod(r.v);
    return r;
  }
  BPP_DEV friend Sc operator-(const Sc& a, const Sc& b) {
    Sc r;
    uint64_t bw = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      uint64_t d = (uint64_t)a.v[i] - b.v[i] - bw;
      r.v[i] = (uint32_t)d;
      bw = (d >> 32) & 1;
    }
    if (bw) {
      uint64_t c = 0;
#pragma unroll
      for (int i = 0; i < 8; i++) { c += (uint64_t)r.v[i] + M::mod(i); r.v[i] = (uint32_t)c; c >>= 32; }
    }
    return r;
  }
  BPP_DEV Sc operator-() const { return zero() - *this; }
  // Montgomery reduction of a 16-limb value t < M * 2^256: returns t * 2^-256 mod M, fully reduced. Rows are written so that
  // row i+1 only waits for limb i+1 of row i (the rows pipeline), and limbs of M that are zero constants fold away.
  BPP_DEV static Sc mont_reduce(uint32_t* t) {
    uint32_t extra = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t m = t[i] * M::NINV;
      uint64_t c = 0;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        c += (uint64_t)m * M::mod(j) + t[i + j];
        t[i + j] = (uint32_t)c;
        c >>= 32;
      }
      c += (uint64_t)t[i + 8] + extra;
      t[i + 8] = (uint32_t)c;
      extra = (uint32_t)(c >> 32);
    }
    if (extra || geq_mod(t + 8)) sub_mod(t + 8);
    Sc r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = t[8 + i];
    return r;
  }
  // Montgomery product: 8x8 schoolbook (IMAD.WIDE carry chains, shared with the base field) + reduction.
  // The product and the square are real FUNCTIONS (operands and result in registers by the ABI), not inlined: the kernels that
  // use scalars are long straight-line protocol code with 50-150 multiplications each, and inlined they were 0.3-0.5 MB of
  // instructions that every warp executed about once - bound by instruction fetch, not by any pipe.
#ifdef BPP_SC_INLINE_MUL
  BPP_DEV static Sc mul_core(const Sc& a, const Sc& b) {
#else
  __device__ __noinline__ static Sc mul_core(Sc a, Sc b) {
#endif
    uint32_t t[16];
    mul_wide_n<8>(t, a.v, b.v);
    return mont_reduce(t);
  }
#ifdef BPP_SC_INLINE_MUL
  BPP_DEV static Sc sqr_core(const Sc& a) {
#else
  __device__ __noinline__ static Sc sqr_core(Sc a) {
#endif
    uint32_t t[16];
    sqr_wide_n<8>(t, a.v);
    return mont_reduce(t);
  }
  BPP_DEV friend Sc operator*(const Sc& a, const Sc& b) { return mul_core(a, b); }
  BPP_DEV Sc sqr() const { return sqr_core(*this); }  // 36 wide products instead of 64 (the inversion chains are ~250 squarings)
  BPP_DEV bool is_zero() const { uint32_t o = 0; for (int i = 0; i < 8; i++) o |= v[i]; return o == 0; }

  // canonical little-endian limbs <-> Montgomery
  BPP_DEV static Sc from_canonical(const uint32_t* limbs) {
    Sc a, r2;
#pragma unroll
    for (int i = 0; i < 8; i++) { a.v[i] = limbs[i]; r2.v[i] = M::r2(i); }
    return a * r2;
  }
  BPP_DEV void to_canonical(uint32_t* limbs) const {
    uint32_t t[16];
#pragma unroll
    for (int i = 0; i < 8; i++) { t[i] = v[i]; t[8 + i] = 0; }
    Sc r = mont_reduce(t);
#pragma unroll
    for (int i = 0; i < 8; i++) limbs[i] = r.v[i];
  }
  // the limbs as they are (no conversion): for values kept canonical, where a * b with a in Montgomery form and b
  // canonical is already the canonical product
  BPP_DEV static Sc raw(const uint32_t* limbs) {
    Sc a;
#pragma unroll
    for (int i = 0; i < 8; i++) a.v[i] = limbs[i];
    return a;
  }
  BPP_DEV static bool canonical_ok(const uint32_t* limbs) { return !geq_mod(limbs); }
  BPP_DEV static Sc from_u64(uint64_t x) {
    uint32_t l[8] = {(uint32_t)x, (uint32_t)(x >> 32), 0, 0, 0, 0, 0, 0};
    return from_canonical(l);
  }
  // PrimeFieldElem::new(i32): negative n -> modulus - |n| (prime_field_elem.rs:191-195)
  BPP_DEV static Sc from_i32(int32_t n) {
    if (n >= 0) return from_u64((uint64_t)n);
    return -from_u64((uint64_t)(-(int64_t)n));
  }
  __device__ __noinline__ Sc pow_u64(uint64_t e) const {
    Sc r = one(), b = *this;
    while (e) {
      if (e & 1) r = r * b;
      b = b.sqr();
      e >>= 1;
    }
    return r;
  }
  // Fermat inverse x^(M-2); inv(0) = 0
  __device__ __noinline__ Sc inv() const {
    uint32_t e[8];
    uint64_t bw = 2;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      uint64_t d = (uint64_t)M::mod(i) - bw;
      e[i] = (uint32_t)d;
      bw = (d >> 32) & 1;
    }
    Sc r = one();
    for (int i = 7; i >= 0; i--)
      for (int b = 31; b >= 0; b--) {
        r = r.sqr();
        if ((e[i] >> b) & 1) r = r * *this;
      }
    return r;
  }
};

// Montgomery's trick: x[0..n) (stride 8 words, Montgomery form, all non-zero) -> their inverses, in place; one
// Fermat inversion. scratch: n * 8 words.
template <class M>
__device__ __noinline__ void sc_batch_invert(uint32_t* x, uint32_t* scratch, uint32_t n) {
  Sc<M> acc = Sc<M>::one();
  for (uint32_t i = 0; i < n; i++) {
    for (int k = 0; k < 8; k++) scratch[8 * i + k] = acc.v[k];
    acc = acc * Sc<M>::raw(x + 8 * i);
  }
  acc = acc.inv();
  for (int i = (int)n - 1; i >= 0; i--) {
    Sc<M> xi = Sc<M>::raw(x + 8 * i);
    Sc<M> inv_i = acc * Sc<M>::raw(scratch + 8 * i);
    acc = acc * xi;
    for (int k = 0; k < 8; k++) x[8 * i + k] = inv_i.v[k];
  }
}

template <class M>
BPP_DEV Sc<M> sc_load(const uint32_t* canonical) { return Sc<M>::from_canonical(canonical); }
template <class M>
BPP_DEV void sc_store(uint32_t* canonical, const Sc<M>& s) { s.to_canonical(canonical); }

}  // namespace bpp

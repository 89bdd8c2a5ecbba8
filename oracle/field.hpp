// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked into, imported by, or called from the
// product library (bulletproofsplus_b200/). Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may use anything under oracle/.
//
// Generic prime-field arithmetic (64-bit limbs, Montgomery form) standing in for the
// reference's L0 layer: mcl's Fr / Fp reached through mcl_rust (un-vendored, unpinned:
// Cargo.toml:23, .gitignore:5 of the reference). Semantics followed:
//   PrimeFieldElem::new(i32)  src/bls12_381/building_block/scalar/prime_field_elem.rs:191-195
//   Add/Sub/Mul/Div/Neg       prime_field_elem.rs:51-182
//   inv / sq / pow            prime_field_elem.rs:202-236
// The same template is instantiated for every modulus on the path (base and scalar
// fields of Ristretto255, secp256k1, BLS12-381 G1).
#pragma once
#include <cstdint>
#include <cstring>
#include <cstddef>

namespace orc {

typedef unsigned __int128 u128;

// P must provide: static constexpr int N; static const uint64_t MOD[N] (little-endian limbs).
template <class P>
struct Fe {
  static constexpr int N = P::N;
  uint64_t v[N];  // Montgomery form, always fully reduced (< MOD)

  // ---- lazily computed constants -------------------------------------------------
  struct Consts {
    uint64_t n0inv;     // -MOD^{-1} mod 2^64
    uint64_t r1[N];     // R mod p     (Montgomery one)
    uint64_t r2[N];     // R^2 mod p
    Consts() {
      uint64_t m0 = P::MOD[0], x = 1;
      for (int i = 0; i < 6; i++) x *= 2 - m0 * x;  // Newton: x = m0^{-1} mod 2^64
      n0inv = (uint64_t)0 - x;
      // r = 2^(64N) mod p by doubling 1 64N times; continue another 64N times for R^2.
      uint64_t t[N];
      for (int i = 0; i < N; i++) t[i] = 0;
      t[0] = 1;
      for (int k = 0; k < 128 * N; k++) {
        if (k == 64 * N) for (int i = 0; i < N; i++) r1[i] = t[i];
        dbl_mod(t);
      }
      for (int i = 0; i < N; i++) r2[i] = t[i];
    }
    static void dbl_mod(uint64_t* t) {
      uint64_t carry = 0;
      for (int i = 0; i < N; i++) {
        uint64_t nv = (t[i] << 1) | carry;
        carry = t[i] >> 63;
        t[i] = nv;
      }
      if (carry || geq(t, P::MOD)) sub_n(t, P::MOD);
    }
  };
  static const Consts& C() { static Consts c; return c; }

  // ---- raw helpers ---------------------------------------------------------------
  static bool geq(const uint64_t* a, const uint64_t* b) {
    for (int i = N - 1; i >= 0; i--) {
      if (a[i] != b[i]) return a[i] > b[i];
    }
    return true;
  }
  static uint64_t sub_n(uint64_t* a, const uint64_t* b) {  // a -= b, returns borrow
    uint64_t borrow = 0;
    for (int i = 0; i < N; i++) {
      u128 d = (u128)a[i] - b[i] - borrow;
      a[i] = (uint64_t)d;
      borrow = (uint64_t)(d >> 64) & 1;
    }
    return borrow;
  }
  static uint64_t add_n(uint64_t* a, const uint64_t* b) {  // a += b, returns carry
    uint64_t carry = 0;
    for (int i = 0; i < N; i++) {
      u128 s = (u128)a[i] + b[i] + carry;
      a[i] = (uint64_t)s;
      carry = (uint64_t)(s >> 64);
    }
    return carry;
  }
  // Montgomery product (CIOS): out = a*b/R mod p
  static void mont_mul(uint64_t* out, const uint64_t* a, const uint64_t* b) {
    const uint64_t n0 = C().n0inv;
    uint64_t t[N + 2];
    for (int i = 0; i < N + 2; i++) t[i] = 0;
    for (int i = 0; i < N; i++) {
      u128 c = 0;
      for (int j = 0; j < N; j++) {
        c += (u128)a[j] * b[i] + t[j];
        t[j] = (uint64_t)c;
        c >>= 64;
      }
      c += t[N];
      t[N] = (uint64_t)c;
      t[N + 1] = (uint64_t)(c >> 64);
      uint64_t m = t[0] * n0;
      c = (u128)m * P::MOD[0] + t[0];
      c >>= 64;
      for (int j = 1; j < N; j++) {
        c += (u128)m * P::MOD[j] + t[j];
        t[j - 1] = (uint64_t)c;
        c >>= 64;
      }
      c += t[N];
      t[N - 1] = (uint64_t)c;
      t[N] = t[N + 1] + (uint64_t)(c >> 64);
    }
    if (t[N] || geq(t, P::MOD)) sub_n(t, P::MOD);
    for (int i = 0; i < N; i++) out[i] = t[i];
  }

  // ---- construction --------------------------------------------------------------
  static Fe zero() { Fe r; for (int i = 0; i < N; i++) r.v[i] = 0; return r; }
  static Fe one() { Fe r; for (int i = 0; i < N; i++) r.v[i] = C().r1[i]; return r; }
  static Fe from_raw(const uint64_t* limbs) {  // limbs < MOD, plain (non-Montgomery)
    Fe r;
    mont_mul(r.v, limbs, C().r2);
    return r;
  }
  static Fe from_u64(uint64_t x) {
    uint64_t t[N];
    for (int i = 0; i < N; i++) t[i] = 0;
    t[0] = x;
    if (N == 1 || geq(t, P::MOD)) { /* only relevant for tiny moduli; not used */ }
    return from_raw(t);
  }
  // PrimeFieldElem::new(i32): negative n -> modulus - |n|  (prime_field_elem.rs:191-195, Fr::set_int)
  static Fe from_i32(int32_t n) {
    if (n >= 0) return from_u64((uint64_t)n);
    return from_u64((uint64_t)(-(int64_t)n)).neg();
  }
  void to_raw(uint64_t* limbs) const {
    uint64_t one_[N];
    for (int i = 0; i < N; i++) one_[i] = 0;
    one_[0] = 1;
    mont_mul(limbs, v, one_);
  }
  // little-endian bytes, NB = 8*N bytes (or fewer: nbytes <= 8N, upper bytes must be 0)
  bool from_bytes_le(const uint8_t* b, size_t nbytes) {
    uint64_t t[N];
    for (int i = 0; i < N; i++) t[i] = 0;
    for (size_t i = 0; i < nbytes; i++) t[i / 8] |= (uint64_t)b[i] << (8 * (i % 8));
    if (geq(t, P::MOD)) return false;  // non-canonical
    *this = from_raw(t);
    return true;
  }
  // reduce an arbitrary little-endian byte string mod p (used for wide PRNG output)
  static Fe from_bytes_reduce(const uint8_t* b, size_t nbytes) {
    Fe acc = zero();
    Fe k256 = from_u64(256);
    for (size_t i = nbytes; i-- > 0;) acc = acc * k256 + from_u64(b[i]);
    return acc;
  }
  void to_bytes_le(uint8_t* b, size_t nbytes) const {
    uint64_t t[N];
    to_raw(t);
    for (size_t i = 0; i < nbytes; i++) b[i] = (uint8_t)(t[i / 8] >> (8 * (i % 8)));
  }

  // ---- arithmetic ----------------------------------------------------------------
  Fe operator+(const Fe& o) const {
    Fe r = *this;
    uint64_t c = add_n(r.v, o.v);
    if (c || geq(r.v, P::MOD)) sub_n(r.v, P::MOD);
    return r;
  }
  Fe operator-(const Fe& o) const {
    Fe r = *this;
    if (sub_n(r.v, o.v)) add_n(r.v, P::MOD);
    return r;
  }
  Fe neg() const { return zero() - *this; }
  Fe operator-() const { return neg(); }
  Fe operator*(const Fe& o) const { Fe r; mont_mul(r.v, v, o.v); return r; }
  Fe sqr() const { return *this * *this; }
  bool is_zero() const { uint64_t a = 0; for (int i = 0; i < N; i++) a |= v[i]; return a == 0; }
  bool operator==(const Fe& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
  bool operator!=(const Fe& o) const { return !(*this == o); }
  // x^e with e given as plain limbs (ne of them)
  Fe pow_limbs(const uint64_t* e, int ne) const {
    Fe r = one();
    for (int i = ne - 1; i >= 0; i--)
      for (int b = 63; b >= 0; b--) {
        r = r.sqr();
        if ((e[i] >> b) & 1) r = r * *this;
      }
    return r;
  }
  // Fermat inverse x^(p-2); inv(0) = 0
  Fe inv() const {
    uint64_t e[N];
    for (int i = 0; i < N; i++) e[i] = P::MOD[i];
    uint64_t two[N];
    for (int i = 0; i < N; i++) two[i] = 0;
    two[0] = 2;
    sub_n(e, two);
    return pow_limbs(e, N);
  }
  bool is_odd() const {  // parity of the canonical representative
    uint64_t t[N];
    to_raw(t);
    return t[0] & 1;
  }
};

// ---- moduli ----------------------------------------------------------------------
struct P25519 { static constexpr int N = 4; static constexpr uint64_t MOD[4] = {
    0xffffffffffffffedULL, 0xffffffffffffffffULL, 0xffffffffffffffffULL, 0x7fffffffffffffffULL}; };
struct PEll { static constexpr int N = 4; static constexpr uint64_t MOD[4] = {  // 2^252 + 27742317777372353535851937790883648493
    0x5812631a5cf5d3edULL, 0x14def9dea2f79cd6ULL, 0x0000000000000000ULL, 0x1000000000000000ULL}; };
// secp256k1 constants: reference src/secp256k1/building_block/secp256k1/secp256k1.rs:20-27
struct PSecP { static constexpr int N = 4; static constexpr uint64_t MOD[4] = {
    0xfffffffefffffc2fULL, 0xffffffffffffffffULL, 0xffffffffffffffffULL, 0xffffffffffffffffULL}; };
struct PSecN { static constexpr int N = 4; static constexpr uint64_t MOD[4] = {
    0xbfd25e8cd0364141ULL, 0xbaaedce6af48a03bULL, 0xfffffffffffffffeULL, 0xffffffffffffffffULL}; };
struct PBlsP { static constexpr int N = 6; static constexpr uint64_t MOD[6] = {
    0xb9feffffffffaaabULL, 0x1eabfffeb153ffffULL, 0x6730d2a0f6b0f624ULL,
    0x64774b84f38512bfULL, 0x4b1ba7b6434bacd7ULL, 0x1a0111ea397fe69aULL}; };
struct PBlsR { static constexpr int N = 4; static constexpr uint64_t MOD[4] = {
    0xffffffff00000001ULL, 0x53bda402fffe5bfeULL, 0x3339d80809a1d805ULL, 0x73eda753299d7d48ULL}; };

}  // namespace orc

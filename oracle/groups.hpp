// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// The three prime-order groups the protocol layer is instantiated over. The reference
// reaches its group through `Point(G1)` (src/bls12_381/building_block/point/point.rs:11-117:
// base_point :15-18, Add :34-49, Sub :51-67, Mul<PrimeFieldElem> :69-85, Neg :87-101,
// zero/is_zero :103-111, eq :113-117), i.e. mcl's BLS12-381 G1. north_star makes
// Ristretto255 the primary group and secp256k1 the third; secp256k1 constants follow
// src/secp256k1/building_block/secp256k1/secp256k1.rs:20-53.
//
// Encodings at the boundary (the reference defines none):
//   Ristretto255 : 32-byte canonical encoding, RFC 9496 §4.3.1/§4.3.2
//   secp256k1    : x||y, 32+32 bytes little-endian; point at infinity = 64 zero bytes
//   BLS12-381 G1 : x||y, 48+48 bytes little-endian; point at infinity = 96 zero bytes
// Scalars: 32 bytes little-endian, canonical (< group order).
#pragma once
#include "field.hpp"

namespace orc {

// ===================================================================================
// Ristretto255 over edwards25519 (a = -1), extended coordinates.
// ===================================================================================
struct Ristretto {
  typedef Fe<P25519> Fq;
  typedef Fe<PEll> Sc;
  static constexpr size_t PB = 32;
  static constexpr int CURVE_ID = 0;
  static constexpr int ORDER_BITS = 253;
  struct Point { Fq X, Y, Z, T; };

  struct K {
    Fq d, d2, sqrt_m1, invsqrt_a_minus_d;
    Point base;
    K() {
      d = Fq::from_u64(121665).neg() * Fq::from_u64(121666).inv();
      d2 = d + d;
      // 2^((p-1)/4): (p-1)/4 = 2^253 - 5
      uint64_t e[4] = {0xfffffffffffffffbULL, 0xffffffffffffffffULL, 0xffffffffffffffffULL, 0x1fffffffffffffffULL};
      sqrt_m1 = Fq::from_u64(2).pow_limbs(e, 4);
      // base point: y = 4/5, x = the even root of (y^2-1)/(d y^2+1)
      Fq y = Fq::from_u64(4) * Fq::from_u64(5).inv();
      Fq u = y.sqr() - Fq::one(), v = d * y.sqr() + Fq::one();
      Fq x;
      sqrt_ratio_impl(x, u, v, sqrt_m1);
      base.X = x; base.Y = y; base.Z = Fq::one(); base.T = x * y;
      Fq amd = Fq::one().neg() - d;  // a - d
      sqrt_ratio_impl(invsqrt_a_minus_d, Fq::one(), amd, sqrt_m1);
    }
  };
  static const K& k() { static K kk; return kk; }

  // RFC 9496 §4.2 SQRT_RATIO_M1
  static bool sqrt_ratio_impl(Fq& r, const Fq& u, const Fq& v, const Fq& sqrt_m1) {
    Fq v3 = v.sqr() * v, v7 = v3.sqr() * v;
    // (p-5)/8 = 2^252 - 3
    uint64_t e[4] = {0xfffffffffffffffdULL, 0xffffffffffffffffULL, 0xffffffffffffffffULL, 0x0fffffffffffffffULL};
    r = (u * v3) * (u * v7).pow_limbs(e, 4);
    Fq check = v * r.sqr();
    bool correct = (check == u);
    bool flipped = (check == u.neg());
    bool flipped_i = (check == u.neg() * sqrt_m1);
    if (flipped || flipped_i) r = r * sqrt_m1;
    if (r.is_odd()) r = r.neg();
    return correct || flipped;
  }
  static bool sqrt_ratio(Fq& r, const Fq& u, const Fq& v) { return sqrt_ratio_impl(r, u, v, k().sqrt_m1); }

  static Point identity() { Point p; p.X = Fq::zero(); p.Y = Fq::one(); p.Z = Fq::one(); p.T = Fq::zero(); return p; }
  static Point base() { return k().base; }
  // unified, complete addition (add-2008-hwcd-3 with k = 2d)
  static Point add(const Point& p, const Point& q) {
    Fq A = (p.Y - p.X) * (q.Y - q.X);
    Fq B = (p.Y + p.X) * (q.Y + q.X);
    Fq C = p.T * k().d2 * q.T;
    Fq D = p.Z * q.Z; D = D + D;
    Fq E = B - A, F = D - C, G = D + C, H = B + A;
    Point r; r.X = E * F; r.Y = G * H; r.T = E * H; r.Z = F * G;
    return r;
  }
  static Point dbl(const Point& p) { return add(p, p); }
  static Point neg(const Point& p) { Point r = p; r.X = p.X.neg(); r.T = p.T.neg(); return r; }
  // Ristretto equality (RFC 9496 §4.3.3)
  static bool eq(const Point& p, const Point& q) {
    return (p.X * q.Y == p.Y * q.X) || (p.Y * q.Y == p.X * q.X);
  }
  static bool is_identity(const Point& p) { return eq(p, identity()); }

  static bool decode(const uint8_t* b, Point& out) {
    Fq s;
    if (!s.from_bytes_le(b, 32)) return false;
    if (s.is_odd()) return false;
    Fq one = Fq::one();
    Fq ss = s.sqr(), u1 = one - ss, u2 = one + ss, u2s = u2.sqr();
    Fq v = (k().d * u1.sqr()).neg() - u2s;
    Fq invsqrt;
    bool was_square = sqrt_ratio(invsqrt, one, v * u2s);
    Fq den_x = invsqrt * u2, den_y = invsqrt * den_x * v;
    Fq x = (s + s) * den_x;
    if (x.is_odd()) x = x.neg();
    Fq y = u1 * den_y, t = x * y;
    if (!was_square || t.is_odd() || y.is_zero()) return false;
    out.X = x; out.Y = y; out.Z = one; out.T = t;
    return true;
  }
  static void encode(const Point& p, uint8_t* b) {
    Fq u1 = (p.Z + p.Y) * (p.Z - p.Y), u2 = p.X * p.Y;
    Fq invsqrt;
    sqrt_ratio(invsqrt, Fq::one(), u1 * u2.sqr());
    Fq den1 = invsqrt * u1, den2 = invsqrt * u2, z_inv = den1 * den2 * p.T;
    Fq ix0 = p.X * k().sqrt_m1, iy0 = p.Y * k().sqrt_m1;
    Fq ench = den1 * k().invsqrt_a_minus_d;
    bool rotate = (p.T * z_inv).is_odd();
    Fq x = rotate ? iy0 : p.X, y = rotate ? ix0 : p.Y, den_inv = rotate ? ench : den2;
    if ((x * z_inv).is_odd()) y = y.neg();
    Fq s = den_inv * (p.Z - y);
    if (s.is_odd()) s = s.neg();
    s.to_bytes_le(b, 32);
  }
};

// ===================================================================================
// Short Weierstrass y^2 = x^3 + b (a = 0), Jacobian coordinates, all exceptional cases
// handled (the reference's affine secp256k1 law does the same case split:
// src/secp256k1/building_block/macros.rs:34-152).
// ===================================================================================
template <class Cfg>
struct Weier {
  typedef typename Cfg::Fq Fq;
  typedef typename Cfg::Sc Sc;
  static constexpr size_t FB = Cfg::FB;
  static constexpr size_t PB = 2 * Cfg::FB;
  static constexpr int CURVE_ID = Cfg::CURVE_ID;
  static constexpr int ORDER_BITS = Cfg::ORDER_BITS;
  struct Point { Fq X, Y, Z; };  // Z == 0 <=> infinity

  static Point identity() { Point p; p.X = Fq::one(); p.Y = Fq::one(); p.Z = Fq::zero(); return p; }
  static Point base() {
    Point p;
    p.X.from_bytes_le(Cfg::GX, FB); p.Y.from_bytes_le(Cfg::GY, FB); p.Z = Fq::one();
    return p;
  }
  static bool is_identity(const Point& p) { return p.Z.is_zero(); }
  static Point dbl(const Point& p) {  // dbl-2009-l
    if (p.Z.is_zero()) return p;
    Fq A = p.X.sqr(), B = p.Y.sqr(), C = B.sqr();
    Fq D = (p.X + B).sqr() - A - C; D = D + D;
    Fq E = A + A + A, F = E.sqr();
    Point r;
    r.X = F - (D + D);
    Fq C8 = C + C; C8 = C8 + C8; C8 = C8 + C8;
    r.Y = E * (D - r.X) - C8;
    r.Z = p.Y * p.Z; r.Z = r.Z + r.Z;
    return r;
  }
  static Point add(const Point& p, const Point& q) {  // add-2007-bl with case split
    if (p.Z.is_zero()) return q;
    if (q.Z.is_zero()) return p;
    Fq Z1Z1 = p.Z.sqr(), Z2Z2 = q.Z.sqr();
    Fq U1 = p.X * Z2Z2, U2 = q.X * Z1Z1;
    Fq S1 = p.Y * q.Z * Z2Z2, S2 = q.Y * p.Z * Z1Z1;
    if (U1 == U2) {
      if (S1 == S2) return dbl(p);
      return identity();
    }
    Fq H = U2 - U1, I = (H + H).sqr(), J = H * I;
    Fq r2 = S2 - S1; r2 = r2 + r2;
    Fq V = U1 * I;
    Point r;
    r.X = r2.sqr() - J - (V + V);
    Fq S1J = S1 * J;
    r.Y = r2 * (V - r.X) - (S1J + S1J);
    r.Z = ((p.Z + q.Z).sqr() - Z1Z1 - Z2Z2) * H;
    return r;
  }
  static Point neg(const Point& p) { Point r = p; r.Y = p.Y.neg(); return r; }
  static bool eq(const Point& p, const Point& q) {
    bool pi = p.Z.is_zero(), qi = q.Z.is_zero();
    if (pi || qi) return pi && qi;
    Fq Z1Z1 = p.Z.sqr(), Z2Z2 = q.Z.sqr();
    return (p.X * Z2Z2 == q.X * Z1Z1) && (p.Y * q.Z * Z2Z2 == q.Y * p.Z * Z1Z1);
  }
  static void to_affine(const Point& p, Fq& x, Fq& y) {
    Fq zi = p.Z.inv(), zi2 = zi.sqr();
    x = p.X * zi2; y = p.Y * zi2 * zi;
  }
  static void encode(const Point& p, uint8_t* b) {
    if (p.Z.is_zero()) { memset(b, 0, PB); return; }
    Fq x, y; to_affine(p, x, y);
    x.to_bytes_le(b, FB); y.to_bytes_le(b + FB, FB);
  }
  static bool decode(const uint8_t* b, Point& out) {
    bool allz = true;
    for (size_t i = 0; i < PB; i++) if (b[i]) { allz = false; break; }
    if (allz) { out = identity(); return true; }
    Fq x, y;
    if (!x.from_bytes_le(b, FB) || !y.from_bytes_le(b + FB, FB)) return false;
    if (y.sqr() != x.sqr() * x + Fq::from_u64(Cfg::B)) return false;  // curve membership
    out.X = x; out.Y = y; out.Z = Fq::one();
    return true;
  }
};

struct SecpCfg {
  typedef Fe<PSecP> Fq;
  typedef Fe<PSecN> Sc;
  static constexpr size_t FB = 32;
  static constexpr int CURVE_ID = 1;
  static constexpr int ORDER_BITS = 256;
  static constexpr uint64_t B = 7;
  // generator, little-endian bytes of secp256k1.rs:45-46
  static constexpr uint8_t GX[32] = {0x98,0x17,0xf8,0x16,0x5b,0x81,0xf2,0x59,0xd9,0x28,0xce,0x2d,0xdb,0xfc,0x9b,0x02,
                                     0x07,0x0b,0x87,0xce,0x95,0x62,0xa0,0x55,0xac,0xbb,0xdc,0xf9,0x7e,0x66,0xbe,0x79};
  static constexpr uint8_t GY[32] = {0xb8,0xd4,0x10,0xfb,0x8f,0xd0,0x47,0x9c,0x19,0x54,0x85,0xa6,0x48,0xb4,0x17,0xfd,
                                     0xa8,0x08,0x11,0x0e,0xfc,0xfb,0xa4,0x5d,0x65,0xc4,0xa3,0x26,0x77,0xda,0x3a,0x48};
};
struct BlsCfg {
  typedef Fe<PBlsP> Fq;
  typedef Fe<PBlsR> Sc;
  static constexpr size_t FB = 48;
  static constexpr int CURVE_ID = 2;
  static constexpr int ORDER_BITS = 255;
  static constexpr uint64_t B = 4;
  // generator of point.rs:16 (decimal there), little-endian bytes
  static constexpr uint8_t GX[48] = {0xbb,0xc6,0x22,0xdb,0x0a,0xf0,0x3a,0xfb,0xef,0x1a,0x7a,0xf9,0x3f,0xe8,0x55,0x6c,
                                     0x58,0xac,0x1b,0x17,0x3f,0x3a,0x4e,0xa1,0x05,0xb9,0x74,0x97,0x4f,0x8c,0x68,0xc3,
                                     0x0f,0xac,0xa9,0x4f,0x8c,0x63,0x95,0x26,0x94,0xd7,0x97,0x31,0xa7,0xd3,0xf1,0x17};
  static constexpr uint8_t GY[48] = {0xe1,0xe7,0xc5,0x46,0x29,0x23,0xaa,0x0c,0xe4,0x8a,0x88,0xa2,0x44,0xc7,0x3c,0xd0,
                                     0xed,0xb3,0x04,0x2c,0xcb,0x18,0xdb,0x00,0xf6,0x0a,0xd0,0xd5,0x95,0xe0,0xf5,0xfc,
                                     0xe4,0x8a,0x1d,0x74,0xed,0x30,0x9e,0xa0,0xf1,0xa0,0xaa,0xe3,0x81,0xf4,0xb3,0x08};
};
typedef Weier<SecpCfg> Secp256k1;
typedef Weier<BlsCfg> Bls12381G1;

// ===================================================================================
// Scalar multiplication, the way `Point * PrimeFieldElem` (point.rs:69-85 -> G1::mul) is
// observable: the group element s*P. Plain MSB-first double-and-add.
// ===================================================================================
template <class G>
typename G::Point scalar_mul(const typename G::Point& p, const typename G::Sc& s) {
  uint64_t e[G::Sc::N];
  s.to_raw(e);
  typename G::Point acc = G::identity();
  bool started = false;
  for (int i = G::Sc::N - 1; i >= 0; i--)
    for (int b = 63; b >= 0; b--) {
      if (started) acc = G::dbl(acc);
      if ((e[i] >> b) & 1) { acc = started ? G::add(acc, p) : p; started = true; }
    }
  return acc;
}

}  // namespace orc

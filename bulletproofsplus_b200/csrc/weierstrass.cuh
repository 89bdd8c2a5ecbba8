// Short-Weierstrass curves y^2 = x^3 + b (a = 0) for sm_100a: secp256k1 (C-ABI curve 1) and BLS12-381 G1 (curve 2,
// the group the reference snapshot actually runs on: `pub struct Point(G1)`, src/bls12_381/building_block/point/point.rs:12).
// Same curve-policy interface as ed25519.cuh, so msm.cuh / proto_*.cuh instantiate unchanged.
//   Ext : XYZZ coordinates (X, Y, ZZ, ZZZ), x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2; identity <=> ZZ = 0
//   Aff : (x, y, -y) in Montgomery form (+ padding to a multiple of 16 B); affine infinity = all zero (not on the curve,
//         b != 0). Subtracting a table point is address selection, as on the Edwards side.
// Mixed addition 8M + 2S (madd-2008-s), addition 12M + 2S (add-2008-s), doubling 6M + 3S (dbl-2008-s-1, a = 0). These
// formulas are NOT complete: P + P, P + (-P), P + O, O + P are branched out explicitly — they are certain to occur with
// the reference's generators (G_4 = H_2 = 15 g), like the case split of the reference's own affine law
// (src/secp256k1/building_block/macros.rs:34-152).
// Boundary encoding (the reference defines none): x || y little-endian canonical, infinity = all zero bytes.
#pragma once
#include "fp_mont.cuh"

namespace bpp {

template <class Cfg>
struct Weierstrass {
  typedef Fp<typename Cfg::P> F;
  static constexpr int NL = F::NL;
  struct Ext { F X, Y, ZZ, ZZZ; };
  struct Aff { F x, y, ny; uint32_t pad[Cfg::AFF_PAD]; };
  static constexpr int CURVE_ID = Cfg::CURVE_ID;
  static constexpr int FIELD_BYTES = 4 * NL;
  static constexpr int POINT_BYTES = 8 * NL;
  static constexpr int SCALAR_BITS = Cfg::SCALAR_BITS;
  __host__ __device__ static constexpr uint32_t order_word(int i) { return Cfg::order_word(i); }  // prime group order
  // no FP64 accumulator path for the Montgomery-form base fields (Ed25519 has one; the names exist so that the hybrid
  // kernel template compiles, it is never launched with HYB = true for these curves)
  static constexpr bool HAS_F64 = false;
  static constexpr bool HAS_COOP4 = false;
  typedef Ext AccF;
  BPP_DEV static void store_table_point_f(uint32_t* dst, const Aff& q) { store_words<AFF_WORDS>(dst, q); }
  BPP_DEV static void accf_init(AccF& p, const uint32_t* q, uint32_t neg) { init_from_ptr(p, q, neg); }
  BPP_DEV static void accf_madd(AccF& p, const uint32_t* q, uint32_t neg) { madd_ptr(p, q, neg); }
  BPP_DEV static void accf_to_ext(Ext& o, const AccF& p) { o = p; }
  static constexpr int AFF_WORDS = 3 * NL + Cfg::AFF_PAD;
  static constexpr int EXT_WORDS = 4 * NL;

  BPP_DEV static void load_f(F& r, const uint32_t* src) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll
    for (int i = 0; i < NL / 4; i++) {
      uint4 q = __ldg(s4 + i);
      r.v[4 * i] = q.x; r.v[4 * i + 1] = q.y; r.v[4 * i + 2] = q.z; r.v[4 * i + 3] = q.w;
    }
  }
  BPP_DEV static F curve_b() { return F::from_u32(Cfg::B); }
  BPP_DEV static void base(Ext& p) {
#pragma unroll
    for (int i = 0; i < NL; i++) { p.X.v[i] = Cfg::gx(i); p.Y.v[i] = Cfg::gy(i); }
    p.ZZ = F::one(); p.ZZZ = F::one();
  }
  BPP_DEV static void set_identity(Ext& p) { p.X = F::one(); p.Y = F::one(); p.ZZ = F::zero(); p.ZZZ = F::zero(); }
  BPP_DEV static void set_identity(Aff& p) {
    p.x = F::zero(); p.y = F::zero(); p.ny = F::zero();
#pragma unroll
    for (int i = 0; i < Cfg::AFF_PAD; i++) p.pad[i] = 0;
  }
  BPP_DEV static bool is_identity(const Ext& p) { return p.ZZ.is_zero(); }
  BPP_DEV static void neg(Ext& p) { p.Y = F::neg(p.Y); }
  BPP_DEV static void dbl_n(Ext& p, uint32_t k) { for (uint32_t i = 0; i < k; i++) dbl(p); }

  // 2 * (x, y) for an affine point
  BPP_DEV static void mdbl(Ext& p, const F& x, const F& y) {
    F U = F::dbl(y), V = F::sqr(U), W = F::mul(U, V), S = F::mul(x, V);
    F X2 = F::sqr(x), M = F::add(F::dbl(X2), X2);
    p.X = F::sub(F::sqr(M), F::dbl(S));
    p.Y = F::sub(F::mul(M, F::sub(S, p.X)), F::mul(W, y));
    p.ZZ = V; p.ZZZ = W;
  }
  BPP_DEV static void dbl(Ext& p) {
    if (p.ZZ.is_zero()) return;
    F U = F::dbl(p.Y), V = F::sqr(U), W = F::mul(U, V), S = F::mul(p.X, V);
    F X2 = F::sqr(p.X), M = F::add(F::dbl(X2), X2);
    F X3 = F::sub(F::sqr(M), F::dbl(S));
    p.Y = F::sub(F::mul(M, F::sub(S, X3)), F::mul(W, p.Y));
    p.X = X3;
    p.ZZ = F::mul(V, p.ZZ); p.ZZZ = F::mul(W, p.ZZZ);
  }
  // p += (x, y), an affine point that is not the point at infinity
  BPP_DEV static void madd_xy(Ext& p, const F& x, const F& y) {
    if (p.ZZ.is_zero()) { p.X = x; p.Y = y; p.ZZ = F::one(); p.ZZZ = F::one(); return; }
    F U2 = F::mul(x, p.ZZ), S2 = F::mul(y, p.ZZZ);
    F P = F::sub(U2, p.X), R = F::sub(S2, p.Y);
    if (P.is_zero()) {
      if (R.is_zero()) mdbl(p, x, y); else set_identity(p);
      return;
    }
    F PP = F::sqr(P), PPP = F::mul(P, PP), Q = F::mul(p.X, PP);
    F X3 = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
    p.Y = F::sub(F::mul(R, F::sub(Q, X3)), F::mul(p.Y, PPP));
    p.X = X3;
    p.ZZ = F::mul(p.ZZ, PP); p.ZZZ = F::mul(p.ZZZ, PPP);
  }
  // HOT: p +/-= the table point at q (AFF_WORDS words, 16-byte aligned)
  BPP_DEV static void madd_ptr(Ext& p, const uint32_t* q, uint32_t neg) {
    F x, y;
    load_f(x, q);
    load_f(y, q + NL + NL * neg);
    if (x.is_zero() && y.is_zero()) return;
    madd_xy(p, x, y);
  }
  BPP_DEV static void init_from_ptr(Ext& p, const uint32_t* q, uint32_t neg) {
    F x, y;
    load_f(x, q);
    load_f(y, q + NL + NL * neg);
    if (x.is_zero() && y.is_zero()) { set_identity(p); return; }
    p.X = x; p.Y = y; p.ZZ = F::one(); p.ZZZ = F::one();
  }
  BPP_DEV static void madd(Ext& p, const Aff& q, bool neg) {
    if (q.x.is_zero() && q.y.is_zero()) return;
    madd_xy(p, q.x, neg ? q.ny : q.y);
  }
  BPP_DEV static void aff_to_ext(Ext& p, const Aff& q) {
    if (q.x.is_zero() && q.y.is_zero()) { set_identity(p); return; }
    p.X = q.x; p.Y = q.y; p.ZZ = F::one(); p.ZZZ = F::one();
  }
  BPP_DEV static void add(Ext& p, const Ext& q) {
    if (q.ZZ.is_zero()) return;
    if (p.ZZ.is_zero()) { p = q; return; }
    F U1 = F::mul(p.X, q.ZZ), U2 = F::mul(q.X, p.ZZ), S1 = F::mul(p.Y, q.ZZZ), S2 = F::mul(q.Y, p.ZZZ);
    F P = F::sub(U2, U1), R = F::sub(S2, S1);
    if (P.is_zero()) {
      if (R.is_zero()) dbl(p); else set_identity(p);
      return;
    }
    F PP = F::sqr(P), PPP = F::mul(P, PP), Q = F::mul(U1, PP);
    F X3 = F::sub(F::sub(F::sqr(R), PPP), F::dbl(Q));
    p.Y = F::sub(F::mul(R, F::sub(Q, X3)), F::mul(S1, PPP));
    p.X = X3;
    p.ZZ = F::mul(F::mul(p.ZZ, q.ZZ), PP);
    p.ZZZ = F::mul(F::mul(p.ZZZ, q.ZZZ), PPP);
  }
  // batched normalisation (wtab_build_kernel): the element whose inverse to_aff needs (1 for the point at infinity), the
  // conversion given that inverse, and the field product
  BPP_DEV static F aff_den(const Ext& p) { return p.ZZ.is_zero() ? F::one() : p.ZZZ; }
  BPP_DEV static F fmul(const F& a, const F& b) { return F::mul(a, b); }
  BPP_DEV static F fone() { return F::one(); }
  BPP_DEV static F finv(const F& a) { return F::inv(a); }
  BPP_DEV static void to_aff_inv(Aff& o, const Ext& p, const F& zzz_inv) {
    set_identity(o);
    if (p.ZZ.is_zero()) return;
    o.y = F::mul(p.Y, zzz_inv);
    F z_inv = F::mul(p.ZZ, zzz_inv);
    o.x = F::mul(p.X, F::sqr(z_inv));
    o.ny = F::neg(o.y);
  }
  // Ext -> Aff (one inversion)
  __device__ static void to_aff(Aff& o, const Ext& p) {
    set_identity(o);
    if (p.ZZ.is_zero()) return;
    F zzz_inv = F::inv(p.ZZZ);
    o.y = F::mul(p.Y, zzz_inv);
    F z_inv = F::mul(p.ZZ, zzz_inv);  // ZZ / ZZZ = 1 / Z
    o.x = F::mul(p.X, F::sqr(z_inv));
    o.ny = F::neg(o.y);
  }
  __device__ static void encode(uint8_t* b, const Ext& p) {
    Aff a;
    to_aff(a, p);
    if (p.ZZ.is_zero()) { for (int i = 0; i < POINT_BYTES; i++) b[i] = 0; return; }
    a.x.to_bytes(b);
    a.y.to_bytes(b + FIELD_BYTES);
  }
  // x || y -> Aff; false for non-canonical coordinates or a point off the curve. (No subgroup check:
  // mcl's G1 deserialisation is not on the reference's path either — it never decodes.)
  __device__ static bool decode(Aff& o, const uint8_t* b) {
    set_identity(o);
    uint32_t any = 0;
    for (int i = 0; i < POINT_BYTES; i++) any |= b[i];
    if (!any) return true;
    F x, y;
    bool ok = F::from_bytes(x, b);
    ok = F::from_bytes(y, b + FIELD_BYTES) && ok;
    if (!ok) return false;
    F rhs = F::add(F::mul(F::sqr(x), x), curve_b());
    if (!F::eq(F::sqr(y), rhs)) return false;
    o.x = x; o.y = y; o.ny = F::neg(y);
    return true;
  }
  // test surface (bpp_debug_op ops 0..4): canonical little-endian field elements in and out
  __device__ static void dbg_field(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
    F x, y, r;
    F::from_bytes(x, a);
    F::from_bytes(y, b);
    switch (op) {
      case 0: r = F::mul(x, y); break;
      case 1: r = F::add(x, y); break;
      case 2: r = F::sub(x, y); break;
      case 3: r = F::inv(x); break;
      default: r = F::sqr(x); break;
    }
    r.to_bytes(out);
  }
};

}  // namespace bpp

// Ristretto255 over edwards25519 (a = -1) for sm_100a: the group behind the C-ABI's curve 0.
// Replaces the reference's `Point(G1)` wrapper (src/bls12_381/building_block/point/point.rs:11-117)
// for the north_star's primary group. Device representations:
//   Ext  : extended coordinates (X:Y:Z:T), 128 B  — accumulators, bucket sums, results
//   Aff  : affine "Niels" form (y+x, y-x, 2d*x*y, -2d*x*y), 128 B = one cache line — every stored/streamed
//          point (generator and window tables, decoded proof points); identity is (1, 1, 0, 0).
//          -Q = (y-x, y+x, -2dxy, 2dxy): subtracting a table point is pure ADDRESS selection in
//          madd_ptr (no SEL / negation instructions in the hot loop).
// Mixed addition Ext += Aff costs 7 field multiplies, Ext += Ext 9, doubling 4S + 4M.
// Encoding/decoding: RFC 9496 §4.3.1 / §4.3.2.
#pragma once
#include "fe25519.cuh"
#include "fe25519_f64.cuh"

namespace bpp {

struct Ed25519 {
  typedef Fe25519 F;
  struct Ext { F X, Y, Z, T; };
  struct Aff { F ypx, ymx, t2d, nt2d; };
  static constexpr int CURVE_ID = 0;
  static constexpr int FIELD_BYTES = 32;
  static constexpr int POINT_BYTES = 32;   // boundary encoding
  static constexpr int SCALAR_BITS = 253;  // group order ell < 2^253
  // ell, little-endian 32-bit words: the table points (multiples of the Ristretto base point) have exactly this order
  __host__ __device__ static constexpr uint32_t order_word(int i) {
    constexpr uint32_t c[8] = {0x5cf5d3edu, 0x5812631au, 0xa2f79cd6u, 0x14def9deu, 0x00000000u, 0x00000000u, 0x00000000u, 0x10000000u};
    return c[i];
  }
  static constexpr int AFF_WORDS = 32;
  static constexpr int EXT_WORDS = 32;
  static constexpr const char* NAME = "ristretto255";

  BPP_DEV static F constant(const uint32_t (&c)[8]) {
    F r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = c[i];
    return r;
  }
  BPP_DEV static F k_d() {
    const uint32_t c[8] = {0x135978a3u, 0x75eb4dcau, 0x4141d8abu, 0x00700a4du, 0x7779e898u, 0x8cc74079u, 0x2b6ffe73u, 0x52036ceeu};
    return constant(c);
  }
  BPP_DEV static F k_d2() {
    const uint32_t c[8] = {0x26b2f159u, 0xebd69b94u, 0x8283b156u, 0x00e0149au, 0xeef3d130u, 0x198e80f2u, 0x56dffce7u, 0x2406d9dcu};
    return constant(c);
  }
  BPP_DEV static F k_sqrt_m1() {
    const uint32_t c[8] = {0x4a0ea0b0u, 0xc4ee1b27u, 0xad2fe478u, 0x2f431806u, 0x3dfbd7a7u, 0x2b4d0099u, 0x4fc1df0bu, 0x2b832480u};
    return constant(c);
  }
  BPP_DEV static F k_invsqrt_a_minus_d() {
    const uint32_t c[8] = {0x805d40eau, 0x99c8fdaau, 0x5a4172beu, 0x9d2f1617u, 0xfe01d840u, 0x16c27b91u, 0xcfaffca2u, 0x786c8905u};
    return constant(c);
  }
  // base point (reference: Point::base_point(), point.rs:15-18, for its own group)
  BPP_DEV static void base(Ext& p) {
    const uint32_t bx[8] = {0x8f25d51au, 0xc9562d60u, 0x9525a7b2u, 0x692cc760u, 0xfdd6dc5cu, 0xc0a4e231u, 0xcd6e53feu, 0x216936d3u};
    const uint32_t by[8] = {0x66666658u, 0x66666666u, 0x66666666u, 0x66666666u, 0x66666666u, 0x66666666u, 0x66666666u, 0x66666666u};
    const uint32_t bt[8] = {0xa5b7dda3u, 0x6dde8ab3u, 0x775152f5u, 0x20f09f80u, 0x64abe37du, 0x66ea4e8eu, 0xd78b7665u, 0x67875f0fu};
    p.X = constant(bx); p.Y = constant(by); p.T = constant(bt); fe_one(p.Z);
  }

  BPP_DEV static void set_identity(Ext& p) { fe_zero(p.X); fe_one(p.Y); fe_one(p.Z); fe_zero(p.T); }
  BPP_DEV static void set_identity(Aff& p) { fe_one(p.ypx); fe_one(p.ymx); fe_zero(p.t2d); fe_zero(p.nt2d); }

  // p += q (q affine Niels); neg => p -= q.  madd-2008-hwcd-3, 7M
  BPP_DEV static void madd(Ext& p, const Aff& q, bool neg) {
    F a, b, c, d, e, f, g, h, t0, m_minus = q.ymx, m_plus = q.ypx;
    fe_cmov(m_minus, q.ypx, neg);  // -q = (ymx, ypx, -t2d)
    fe_cmov(m_plus, q.ymx, neg);
    fe_sub(t0, p.Y, p.X);
    fe_mul(a, t0, m_minus);
    fe_add(t0, p.Y, p.X);
    fe_mul(b, t0, m_plus);
    fe_mul(c, p.T, q.t2d);
    fe_dbl(d, p.Z);
    fe_sub(e, b, a);
    fe_add(h, b, a);
    if (neg) { fe_add(f, d, c); fe_sub(g, d, c); }
    else     { fe_sub(f, d, c); fe_add(g, d, c); }
    fe_mul(p.X, e, f);
    fe_mul(p.Y, g, h);
    fe_mul(p.T, e, h);
    fe_mul(p.Z, f, g);
  }
  BPP_DEV static void load_fe(F& r, const uint32_t* src) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4 a = __ldg(s4), b = __ldg(s4 + 1);
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  }
  // HOT: p +/-= the table point at q (AFF_WORDS words, 128-byte aligned). neg in {0,1} only moves the three loads.
  BPP_DEV static void madd_ptr(Ext& p, const uint32_t* q, uint32_t neg) {
    F a, b, c, d, e, f, g, h, t0, m_minus, m_plus, td;
    load_fe(m_plus, q + 8 * neg);         // y+x  (or y-x)
    load_fe(m_minus, q + 8 * (1 - neg));  // y-x  (or y+x)
    load_fe(td, q + 16 + 8 * neg);        // 2dxy (or -2dxy)
    fe_sub(t0, p.Y, p.X);
    fe_mul(a, t0, m_minus);
    fe_add(t0, p.Y, p.X);
    fe_mul(b, t0, m_plus);
    fe_mul(c, p.T, td);
    fe_dbl(d, p.Z);
    fe_sub(e, b, a);
    fe_add(h, b, a);
    fe_sub(f, d, c);
    fe_add(g, d, c);
    fe_mul(p.X, e, f);
    fe_mul(p.Y, g, h);
    fe_mul(p.T, e, h);
    fe_mul(p.Z, f, g);
  }
  // ---- bucket accumulators on the FP64 pipe (fe25519_f64.cuh) ---------------------------------------------------------
  // Alone, the FP64 multiply is no faster than the integer one (both ~320 issue cycles per warp-multiply, measured), but
  // the two load DIFFERENT pipes: in the hybrid accumulation kernel (msm_accumulate_kernel<C, true>) some blocks keep
  // their buckets as 4 x 5 doubles and read table points in the 128-byte slot format written by store_table_point_f: ((y+x)/2, (y-x)/2, d*x*y) as 3 x 5 balanced doubles (+ 8 bytes of padding).
  // Halving the Niels form lets D = Z1 instead of 2*Z1, which keeps every operand of the second-level multiplies within
  // the lazy bound; all four output coordinates come out divided by 4, i.e. the same projective point.
  static constexpr bool HAS_F64 = true;
  static constexpr bool HAS_COOP4 = true;  // four-role point operations: msm_combine4_kernel
  struct AccF { F5 X, Y, Z, T; };
  BPP_DEV static void store_table_point_f(uint32_t* dst, const Aff& q) {
    const uint32_t hc[8] = {0xfffffff7u, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0x3fffffffu};  // (p+1)/2
    const F half = constant(hc);
    double* d = reinterpret_cast<double*>(dst);
    F t;
    F5 f;
    fe_mul(t, q.ypx, half); f5_from_fe(f, t);
#pragma unroll
    for (int i = 0; i < 5; i++) d[i] = f.v[i];
    fe_mul(t, q.ymx, half); f5_from_fe(f, t);
#pragma unroll
    for (int i = 0; i < 5; i++) d[5 + i] = f.v[i];
    fe_mul(t, q.t2d, half); f5_from_fe(f, t);
#pragma unroll
    for (int i = 0; i < 5; i++) d[10 + i] = f.v[i];
    d[15] = 0.0;
  }
  BPP_DEV static void accf_init(AccF& p, const uint32_t* q, uint32_t neg) {
    const double* d = reinterpret_cast<const double*>(q);
    F5 q0, q1;
    f5_load(q0, d + 5 * neg);
    f5_load(q1, d + 5 * (1 - neg));
    f5_sub(p.X, q0, q1);  // +/- x
    f5_add(p.Y, q0, q1);  // y
    f5_mul(p.T, p.X, p.Y);
#pragma unroll
    for (int i = 0; i < 5; i++) p.Z.v[i] = i ? 0.0 : 1.0;
  }
  // HOT: p +/-= the table point at q; 7 multiplies on the FP64 pipe
  BPP_DEV static void accf_madd(AccF& p, const uint32_t* q, uint32_t neg) {
    const double* d = reinterpret_cast<const double*>(q);
    F5 q0, q1, q2, t0, a, b, c, e, f, g, h;
    f5_load(q0, d + 5 * neg);
    f5_load(q1, d + 5 * (1 - neg));
    f5_load(q2, d + 10);
    f5_sub(t0, p.Y, p.X);
    f5_mul(a, t0, q1);
    f5_add(t0, p.Y, p.X);
    f5_mul(b, t0, q0);
    f5_mul(c, p.T, q2);
    f5_cneg(c, c, neg);
    f5_sub(e, b, a);
    f5_add(h, b, a);
    f5_sub(f, p.Z, c);
    f5_add(g, p.Z, c);
    f5_mul(p.X, e, f);
    f5_mul(p.Y, g, h);
    f5_mul(p.T, e, h);
    f5_mul(p.Z, f, g);
  }
  BPP_DEV static void accf_to_ext(Ext& o, const AccF& p) {
    f5_to_fe(o.X, p.X); f5_to_fe(o.Y, p.Y); f5_to_fe(o.Z, p.Z); f5_to_fe(o.T, p.T);
  }

  // p = +/- the table point at q, as (4x : 4y : 4 : 4xy) — one multiply instead of a madd onto the identity
  BPP_DEV static void init_from_ptr(Ext& p, const uint32_t* q, uint32_t neg) {
    F m_minus, m_plus, dx, sy;
    load_fe(m_plus, q + 8 * neg);
    load_fe(m_minus, q + 8 * (1 - neg));
    fe_sub(dx, m_plus, m_minus);  // 2x
    fe_add(sy, m_plus, m_minus);  // 2y
    fe_mul(p.T, dx, sy);          // 4xy
    fe_dbl(p.X, dx);
    fe_dbl(p.Y, sy);
    fe_zero(p.Z);
    p.Z.v[0] = 4;
  }
  // p += q, both extended; complete (add-2008-hwcd-3 with k = 2d), 9M
  BPP_DEV static void add(Ext& p, const Ext& q) {
    F a, b, c, d, e, f, g, h, t0, t1;
    fe_sub(t0, p.Y, p.X); fe_sub(t1, q.Y, q.X); fe_mul(a, t0, t1);
    fe_add(t0, p.Y, p.X); fe_add(t1, q.Y, q.X); fe_mul(b, t0, t1);
    fe_mul(t0, p.T, q.T); fe_mul(c, t0, k_d2());
    fe_mul(t0, p.Z, q.Z); fe_dbl(d, t0);
    fe_sub(e, b, a); fe_sub(f, d, c); fe_add(g, d, c); fe_add(h, b, a);
    fe_mul(p.X, e, f); fe_mul(p.Y, g, h); fe_mul(p.T, e, h); fe_mul(p.Z, f, g);
  }
  // dbl-2008-hwcd, a = -1
  BPP_DEV static void dbl(Ext& p) {
    F a, b, c, e, g, f, h, t0;
    fe_sqr(a, p.X); fe_sqr(b, p.Y);
    fe_sqr(t0, p.Z); fe_dbl(c, t0);
    fe_add(t0, p.X, p.Y); fe_sqr(e, t0); fe_sub(e, e, a); fe_sub(e, e, b);  // E = 2XY
    fe_sub(g, b, a);       // G = D + B = B - A
    fe_sub(f, g, c);       // F = G - C
    fe_add(h, a, b); fe_neg(h, h);  // H = D - B = -A - B
    fe_mul(p.X, e, f); fe_mul(p.Y, g, h); fe_mul(p.T, e, h); fe_mul(p.Z, f, g);
  }
  // k doublings in a row: the formulas never read T, so only the last one computes it (3M + 4S for the others)
  BPP_DEV static void dbl_n(Ext& p, uint32_t k) {
    for (uint32_t i = 0; i + 1 < k; i++) {
      F a, b, c, e, g, f, h, t0;
      fe_sqr(a, p.X); fe_sqr(b, p.Y);
      fe_sqr(t0, p.Z); fe_dbl(c, t0);
      fe_add(t0, p.X, p.Y); fe_sqr(e, t0); fe_sub(e, e, a); fe_sub(e, e, b);
      fe_sub(g, b, a);
      fe_sub(f, g, c);
      fe_add(h, a, b); fe_neg(h, h);
      fe_mul(p.X, e, f); fe_mul(p.Y, g, h); fe_mul(p.Z, f, g);
    }
    if (k) dbl(p);
  }
  BPP_DEV static void neg(Ext& p) { fe_neg(p.X, p.X); fe_neg(p.T, p.T); }

  // Ext -> Aff (one inversion)
  __device__ static void to_aff(Aff& o, const Ext& p) {
    F zi, x, y, t;
    fe_invert(zi, p.Z);
    fe_mul(x, p.X, zi); fe_mul(y, p.Y, zi);
    fe_add(o.ypx, y, x); fe_sub(o.ymx, y, x);
    fe_mul(t, x, y); fe_mul(o.t2d, t, k_d2()); fe_neg(o.nt2d, o.t2d);
  }
  // batched normalisation (wtab_build_kernel): the element whose inverse to_aff needs (Z is never zero on this curve), the
  // conversion given that inverse, and the field product
  BPP_DEV static F aff_den(const Ext& p) { return p.Z; }
  BPP_DEV static F fmul(const F& a, const F& b) { F r; fe_mul(r, a, b); return r; }
  BPP_DEV static F fone() { F r; fe_one(r); return r; }
  BPP_DEV static F finv(const F& a) { F r; fe_invert(r, a); return r; }
  BPP_DEV static void to_aff_inv(Aff& o, const Ext& p, const F& zi) {
    F x, y, t;
    fe_mul(x, p.X, zi); fe_mul(y, p.Y, zi);
    fe_add(o.ypx, y, x); fe_sub(o.ymx, y, x);
    fe_mul(t, x, y); fe_mul(o.t2d, t, k_d2()); fe_neg(o.nt2d, o.t2d);
  }
  BPP_DEV static void aff_to_ext(Ext& p, const Aff& q) {  // via identity + q (7M; only used off the hot path)
    set_identity(p);
    madd(p, q, false);
  }

  // the group identity in Ristretto: the coset {X == 0 or Y == 0} (RFC 9496 §4.3.3 applied to (0,1))
  BPP_DEV static bool is_identity(const Ext& p) { return fe_is_zero(p.X) || fe_is_zero(p.Y); }

  // RFC 9496 §4.2 SQRT_RATIO_M1
  __device__ static bool sqrt_ratio_m1(F& r, const F& u, const F& v) {
    F v3, v7, t, check, nu, nui;
    fe_sqr(t, v); fe_mul(v3, t, v);
    fe_sqr(t, v3); fe_mul(v7, t, v);
    fe_mul(t, u, v7); fe_pow22523(t, t);
    fe_mul(r, u, v3); fe_mul(r, r, t);
    fe_sqr(t, r); fe_mul(check, v, t);
    fe_neg(nu, u); fe_mul(nui, nu, k_sqrt_m1());
    bool correct = fe_eq(check, u), flipped = fe_eq(check, nu), flipped_i = fe_eq(check, nui);
    F rp;
    fe_mul(rp, r, k_sqrt_m1());
    fe_cmov(r, rp, flipped || flipped_i);
    fe_cneg(r, fe_is_neg(r));
    return correct || flipped;
  }

  // 32 bytes -> Aff; false on any of the RFC 9496 §4.3.1 failure conditions
  __device__ static bool decode(Aff& o, const uint8_t* b) {
    F s, ss, u1, u2, u2s, v, t, one, invsqrt, den_x, den_y, x, y;
    fe_one(one);
    bool ok = fe_from_bytes(s, b);
    ok = ok && !(s.v[0] & 1u);
    fe_sqr(ss, s);
    fe_sub(u1, one, ss); fe_add(u2, one, ss); fe_sqr(u2s, u2);
    fe_sqr(t, u1); fe_mul(t, t, k_d()); fe_neg(t, t); fe_sub(v, t, u2s);
    fe_mul(t, v, u2s);
    bool was_square = sqrt_ratio_m1(invsqrt, one, t);
    fe_mul(den_x, invsqrt, u2);
    fe_mul(den_y, invsqrt, den_x); fe_mul(den_y, den_y, v);
    fe_add(t, s, s); fe_mul(x, t, den_x);
    fe_cneg(x, fe_is_neg(x));
    fe_mul(y, u1, den_y);
    fe_mul(t, x, y);
    ok = ok && was_square && !fe_is_neg(t) && !fe_is_zero(y);
    fe_add(o.ypx, y, x); fe_sub(o.ymx, y, x); fe_mul(o.t2d, t, k_d2()); fe_neg(o.nt2d, o.t2d);
    if (!ok) set_identity(o);
    return ok;
  }

  // test surface (bpp_debug_op ops 0..4): a, b are raw 256-bit little-endian values (any value < 2^256, the loose
  // representation); out = canonical result
  __device__ static void dbg_field(int op, const uint8_t* a, const uint8_t* b, uint8_t* out) {
    F x, y, r;
    for (int i = 0; i < 8; i++) {
      x.v[i] = (uint32_t)a[4 * i] | ((uint32_t)a[4 * i + 1] << 8) | ((uint32_t)a[4 * i + 2] << 16) | ((uint32_t)a[4 * i + 3] << 24);
      y.v[i] = (uint32_t)b[4 * i] | ((uint32_t)b[4 * i + 1] << 8) | ((uint32_t)b[4 * i + 2] << 16) | ((uint32_t)b[4 * i + 3] << 24);
    }
    switch (op) {
      case 0: fe_mul(r, x, y); break;
      case 1: fe_add(r, x, y); break;
      case 2: fe_sub(r, x, y); break;
      case 3: fe_invert(r, x); break;
      default: fe_sqr(r, x); break;
    }
    fe_to_bytes(out, r);
  }

  // Ext -> 32 bytes (RFC 9496 §4.3.2)
  __device__ static void encode(uint8_t* b, const Ext& p) {
    F u1, u2, t, invsqrt, den1, den2, z_inv, ix0, iy0, ench, x, y, den_inv, s, one;
    fe_one(one);
    fe_add(t, p.Z, p.Y); fe_sub(u1, p.Z, p.Y); fe_mul(u1, t, u1);
    fe_mul(u2, p.X, p.Y);
    fe_sqr(t, u2); fe_mul(t, t, u1);
    sqrt_ratio_m1(invsqrt, one, t);
    fe_mul(den1, invsqrt, u1); fe_mul(den2, invsqrt, u2);
    fe_mul(z_inv, den1, den2); fe_mul(z_inv, z_inv, p.T);
    fe_mul(ix0, p.X, k_sqrt_m1()); fe_mul(iy0, p.Y, k_sqrt_m1());
    fe_mul(ench, den1, k_invsqrt_a_minus_d());
    fe_mul(t, p.T, z_inv);
    bool rotate = fe_is_neg(t);
    x = p.X; y = p.Y; den_inv = den2;
    fe_cmov(x, iy0, rotate); fe_cmov(y, ix0, rotate); fe_cmov(den_inv, ench, rotate);
    fe_mul(t, x, z_inv);
    fe_cneg(y, fe_is_neg(t));
    fe_sub(t, p.Z, y); fe_mul(s, den_inv, t);
    fe_cneg(s, fe_is_neg(s));
    fe_to_bytes(b, s);
  }
};

}  // namespace bpp

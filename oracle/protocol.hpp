// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Literal CPU restatement of the reference's protocol layer, generic over the group G.
// Each function cites the reference lines it follows (paths relative to /root/reference).
// Challenges and prover nonces come from `Params`; `Params::reference(m)` reproduces the
// constants hard-coded in the snapshot and is the parity mode.
#pragma once
#include <vector>
#include <stdexcept>
#include "groups.hpp"

namespace orc {

template <class G>
struct Params {
  typedef typename G::Sc Sc;
  Sc y, z, e_final, alpha, r, s, delta, eta;
  std::vector<Sc> e_round, d_L, d_R;
  // src/range/mod.rs:94,109-110 (single: alpha=7, y=z=7), :256,278-279 (aggregated: alpha=33, y=12, z=23);
  // src/weighted_inner_product_proof.rs:94-95 (d_L=4,d_R=5), :131 (round e=7), :175-178 (r,s,delta,eta), :211 (e=99)
  static Params reference(size_t m, size_t rounds) {
    Params p;
    bool single = (m == 1);
    p.y = Sc::from_i32(single ? 7 : 12);
    p.z = Sc::from_i32(single ? 7 : 23);
    p.alpha = Sc::from_i32(single ? 7 : 33);
    p.e_final = Sc::from_i32(99);
    p.r = Sc::from_i32(33); p.s = Sc::from_i32(44); p.delta = Sc::from_i32(88); p.eta = Sc::from_i32(123);
    p.e_round.assign(rounds, Sc::from_i32(7));
    p.d_L.assign(rounds, Sc::from_i32(4));
    p.d_R.assign(rounds, Sc::from_i32(5));
    return p;
  }
};

// Baseline switch (NOT part of the restatement): when a group installs a hook and the process-wide backend is 1
// (orc_set_msm_backend), MulVec::calculate hands MSMs of >= 16 terms to it — B2, the bucket MSM of baseline_msm.cpp — instead
// of running the reference's per-term loop. Default 0: every parity test runs the literal loop below.
inline int& msm_backend() { static int b = 0; return b; }
template <class G> struct MsmHook {
  typedef typename G::Point (*Fn)(const std::vector<typename G::Sc>&, const std::vector<typename G::Point>&);
  static Fn& fn() { static Fn f = nullptr; return f; }
};

// ---- src/bls12_381/building_block/mulvec.rs:13-53 ------------------------------------
template <class G>
struct MulVec {
  typedef typename G::Sc Sc;
  typedef typename G::Point Point;
  std::vector<Sc> scalars;
  std::vector<Point> points;
  void add_scalar(const Sc& s) { scalars.push_back(s); }
  void add_scalars(const std::vector<Sc>& v) { for (auto& s : v) scalars.push_back(s); }
  void add_point(const Point& p) { points.push_back(p); }
  void add_points(const std::vector<Point>& v) { for (auto& p : v) points.push_back(p); }
  // mulvec.rs:20-33: one full scalar multiplication and one addition per term
  Point calculate() const {
    if (scalars.size() != points.size())
      throw std::runtime_error("mulvec: lengths of scalars and points must match");
    if (msm_backend() == 1 && MsmHook<G>::fn() && scalars.size() >= 16) return MsmHook<G>::fn()(scalars, points);
    Point sum = G::identity();
    for (size_t i = 0; i < scalars.size(); i++) {
      Point p = scalar_mul<G>(points[i], scalars[i]);
      sum = G::add(sum, p);
    }
    return sum;
  }
};

// ---- src/util.rs ------------------------------------------------------------------------
template <class Sc> std::vector<Sc> exp_iter_type1(const Sc& x, size_t n) {  // util.rs:29-32: 1, x, x^2, ...
  std::vector<Sc> out; out.reserve(n);
  Sc cur = Sc::from_i32(1);
  for (size_t i = 0; i < n; i++) { out.push_back(cur); cur = cur * x; }
  return out;
}
template <class Sc> std::vector<Sc> exp_iter_type2(const Sc& x, size_t n) {  // util.rs:34-37: x, x^2, ...
  std::vector<Sc> out; out.reserve(n);
  Sc cur = x;
  for (size_t i = 0; i < n; i++) { out.push_back(cur); cur = cur * x; }
  return out;
}
template <class Sc> Sc scalar_exp_vartime(const Sc& x, uint64_t n) {  // util.rs:39-52
  Sc result = Sc::from_i32(1), aux = x;
  while (n > 0) {
    if (n & 1) result = result * aux;
    n >>= 1;
    aux = aux * aux;
  }
  return result;
}
template <class Sc> Sc sum_of_powers_type1(const Sc& x, size_t n) {  // util.rs:54-79
  if (n == 0 || (n & (n - 1))) { Sc s = Sc::zero(); for (auto& t : exp_iter_type1(x, n)) s = s + t; return s; }
  if (n == 1) return Sc::from_i32(1);
  size_t m = n;
  Sc result = Sc::from_i32(1) + x, factor = x;
  while (m > 2) { factor = factor * factor; result = result + factor * result; m /= 2; }
  return result;
}
template <class Sc> Sc sum_of_powers_type2(const Sc& x, size_t n) {  // util.rs:81-106
  if (n == 0 || (n & (n - 1))) { Sc s = Sc::zero(); for (auto& t : exp_iter_type2(x, n)) s = s + t; return s; }
  if (n == 1) return Sc::from_i32(1);  // (sic) util.rs:85-87 returns new(n), not x
  size_t m = n;
  Sc result = x + x * x, factor = x;
  while (m > 2) { factor = factor * factor; result = result + factor * result; m /= 2; }
  return result;
}
template <class Sc> Sc weighted_inner_product(const Sc* a, const Sc* b, const Sc* c, size_t n) {  // util.rs:117-127
  Sc out = Sc::zero();
  for (size_t i = 0; i < n; i++) out = out + (a[i] * b[i] * c[i]);
  return out;
}

// ---- src/publickey.rs:13-52 ---------------------------------------------------------------
template <class G>
struct PublicKey {
  typedef typename G::Sc Sc;
  typedef typename G::Point Point;
  Point g, h;
  std::vector<Point> G_vec, H_vec;
  static PublicKey create(size_t length) {  // publickey.rs:21-48
    PublicKey pk;
    pk.g = G::base();
    pk.h = scalar_mul<G>(pk.g, Sc::from_i32(2));
    for (size_t i = 0; i < length; i++) pk.G_vec.push_back(scalar_mul<G>(pk.g, Sc::from_i32(((int32_t)i + 1) * 3)));
    for (size_t i = 0; i < length; i++) pk.H_vec.push_back(scalar_mul<G>(pk.g, Sc::from_i32(((int32_t)i + 1) * 5)));
    return pk;
  }
  Point commitment(const Sc& v, const Sc& gamma) const {  // publickey.rs:50-52
    return G::add(scalar_mul<G>(g, v), scalar_mul<G>(h, gamma));
  }
  // src/range/prover.rs:28-42 — note the `v as i32` cast at :37
  Point commit_u64(uint64_t v, const Sc& gamma) const { return commitment(Sc::from_i32((int32_t)(uint32_t)v), gamma); }
};

// ---- src/weighted_inner_product_proof.rs ---------------------------------------------------
template <class G>
struct WipProof {
  typedef typename G::Sc Sc;
  typedef typename G::Point Point;
  std::vector<Point> L_vec, R_vec;
  Point A, B;
  Sc r_prime, s_prime, d_prime;

  // weighted_inner_product_proof.rs:36-226. `commitment` is only used by the dead update of P
  // (:137-142, never read afterwards) and is therefore not a parameter here.
  static WipProof prove(const PublicKey<G>& pk, const std::vector<Sc>& a_vec, const std::vector<Sc>& b_vec,
                        const std::vector<Sc>& power_of_y_vec, Sc gamma, const Params<G>& prm) {
    std::vector<Point> Gv = pk.G_vec, Hv = pk.H_vec;  // :45-49
    std::vector<Sc> a = a_vec, b = b_vec, py = power_of_y_vec;
    Sc alpha = gamma;
    size_t n = Gv.size();
    if (Hv.size() != n || a.size() != n || b.size() != n || py.size() != n) throw std::runtime_error("wip: length mismatch");  // :61-64
    if (n == 0 || (n & (n - 1))) throw std::runtime_error("wip: n must be a power of two");                            // :67
    WipProof out;
    size_t round = 0;
    while (n != 1) {  // :79-172
      n /= 2;
      Sc *a1 = &a[0], *a2 = &a[n], *b1 = &b[0], *b2 = &b[n], *y1 = &py[0], *y2 = &py[n];
      Point *G1 = &Gv[0], *G2 = &Gv[n], *H1 = &Hv[0], *H2 = &Hv[n];
      Sc c_L = weighted_inner_product(a1, b2, y1, n);  // :90
      Sc c_R = weighted_inner_product(a2, b1, y2, n);  // :91
      if (round >= prm.e_round.size()) throw std::runtime_error("wip: not enough round parameters");
      Sc d_L = prm.d_L[round], d_R = prm.d_R[round];   // :94-95
      Sc y_nhat = y1[n - 1], y_nhat_inv = y_nhat.inv();  // :98-99
      MulVec<G> mvL;  // :103-113
      for (size_t i = 0; i < n; i++) mvL.add_scalar(y_nhat_inv * a1[i]);
      for (size_t i = 0; i < n; i++) mvL.add_scalar(b2[i]);
      mvL.add_scalar(c_L); mvL.add_scalar(d_L);
      for (size_t i = 0; i < n; i++) mvL.add_point(G2[i]);
      for (size_t i = 0; i < n; i++) mvL.add_point(H1[i]);
      mvL.add_point(pk.g); mvL.add_point(pk.h);
      Point L = mvL.calculate();
      MulVec<G> mvR;  // :115-125
      for (size_t i = 0; i < n; i++) mvR.add_scalar(y_nhat * a2[i]);
      for (size_t i = 0; i < n; i++) mvR.add_scalar(b1[i]);
      mvR.add_scalar(c_R); mvR.add_scalar(d_R);
      for (size_t i = 0; i < n; i++) mvR.add_point(G1[i]);
      for (size_t i = 0; i < n; i++) mvR.add_point(H2[i]);
      mvR.add_point(pk.g); mvR.add_point(pk.h);
      Point R = mvR.calculate();
      out.L_vec.push_back(L); out.R_vec.push_back(R);  // :127-128
      Sc e = prm.e_round[round];                       // :131
      Sc e_inv = e.inv(), e_sqr = e * e, e_sqr_inv = e_inv * e_inv;
      Sc y_nhat_e_inv = y_nhat * e_inv, y_nhat_inv_e = y_nhat_inv * e;  // :144-145
      for (size_t i = 0; i < n; i++) {  // :147-164
        a1[i] = a1[i] * e + a2[i] * y_nhat_e_inv;
        b1[i] = b1[i] * e_inv + b2[i] * e;
        MulVec<G> mg; mg.add_scalar(e_inv); mg.add_scalar(y_nhat_inv_e); mg.add_point(G1[i]); mg.add_point(G2[i]);
        G1[i] = mg.calculate();
        MulVec<G> mh; mh.add_scalar(e); mh.add_scalar(e_inv); mh.add_point(H1[i]); mh.add_point(H2[i]);
        H1[i] = mh.calculate();
      }
      alpha = alpha + (e_sqr * d_L + e_sqr_inv * d_R);  // :171
      round++;
    }
    Sc r = prm.r, s = prm.s, delta = prm.delta, eta = prm.eta;  // :175-178
    Sc rcbsca = r * py[0] * b[0] + s * py[0] * a[0];            // :181
    Sc rcs = r * py[0] * s;                                     // :182
    MulVec<G> mvA;  // :189-199
    mvA.add_scalar(r); mvA.add_scalar(s); mvA.add_scalar(rcbsca); mvA.add_scalar(delta);
    mvA.add_point(Gv[0]); mvA.add_point(Hv[0]); mvA.add_point(pk.g); mvA.add_point(pk.h);
    out.A = mvA.calculate();
    MulVec<G> mvB;  // :202-208
    mvB.add_scalar(rcs); mvB.add_scalar(eta); mvB.add_point(pk.g); mvB.add_point(pk.h);
    out.B = mvB.calculate();
    Sc e = prm.e_final;  // :211
    out.r_prime = r + a[0] * e;                          // :214
    out.s_prime = s + b[0] * e;                          // :215
    out.d_prime = eta + delta * e + alpha * e * e;       // :216
    return out;
  }

  // weighted_inner_product_proof.rs:330-382. Returns false for the Err(VerificationError) at :334-337.
  bool verification_scalars(size_t n, const Params<G>& prm, std::vector<Sc>& ch_sqr, std::vector<Sc>& ch_inv_sqr,
                            std::vector<Sc>& s_vec, Sc& e) const {
    size_t logn = L_vec.size();
    if (logn >= 8 * sizeof(size_t) || n != ((size_t)1 << logn)) return false;
    if (prm.e_round.size() < logn) throw std::runtime_error("wip: not enough round parameters");
    std::vector<Sc> ch(prm.e_round.begin(), prm.e_round.begin() + logn);  // :351-354
    // batch_invert (prime_field_elem.rs:239-248): allinv = prod of inverses, inverses in input order
    Sc allinv = Sc::from_i32(1);
    std::vector<Sc> ch_inv;
    for (auto& x : ch) { ch_inv.push_back(x.inv()); allinv = allinv * x.inv(); }
    for (size_t i = 0; i < logn; i++) { ch[i] = ch[i] * ch[i]; ch_inv[i] = ch_inv[i] * ch_inv[i]; }  // :361-364
    ch_sqr = ch; ch_inv_sqr = ch_inv;
    e = prm.e_final;  // :369
    s_vec.clear(); s_vec.reserve(n);
    s_vec.push_back(allinv);  // :372-373
    for (size_t i = 1; i < n; i++) {  // :375-380
      size_t log_i = 31 - __builtin_clz((uint32_t)i);
      size_t k = (size_t)1 << log_i;
      s_vec.push_back(s_vec[i - k] * ch_sqr[(logn - 1) - log_i]);
    }
    return true;
  }

  // weighted_inner_product_proof.rs:238-328. true = Ok(()), false = Err(VerificationError)
  bool verify(const PublicKey<G>& pk, const std::vector<Sc>& power_of_y_vec, const std::vector<Sc>& G_exp_c,
              const std::vector<Sc>& H_exp_c, const Sc& g_exp_c, const std::vector<Sc>& V_exp_c, const Point& A_prime,
              const std::vector<Point>& V, const Params<G>& prm) const {
    size_t logn = L_vec.size();
    size_t n = (size_t)1 << logn;
    const Sc& y = power_of_y_vec[0];
    std::vector<Sc> ch_sqr, ch_inv_sqr, s_vec; Sc e;
    if (!verification_scalars(n, prm, ch_sqr, ch_inv_sqr, s_vec, e)) return false;
    Sc e_sqr = e * e, r_prime_e_y = r_prime * e * y, s_prime_e = s_prime * e;  // :260-262
    std::vector<Sc> Ls_exp, Rs_exp, G_exp, H_exp, V_exp;
    for (auto& c : ch_sqr) Ls_exp.push_back(c * e_sqr);      // :265-268
    for (auto& c : ch_inv_sqr) Rs_exp.push_back(c * e_sqr);  // :270-273
    size_t ng = s_vec.size() < G_exp_c.size() ? s_vec.size() : G_exp_c.size();  // zip truncation
    std::vector<Sc> yinv = exp_iter_type2(y.inv(), ng);
    for (size_t i = 0; i < ng; i++) G_exp.push_back((-s_vec[i]) * yinv[i] * r_prime_e_y + G_exp_c[i] * e_sqr);  // :275-282
    size_t nh = s_vec.size() < H_exp_c.size() ? s_vec.size() : H_exp_c.size();
    for (size_t i = 0; i < nh; i++) H_exp.push_back((-s_vec[s_vec.size() - 1 - i]) * s_prime_e + H_exp_c[i] * e_sqr);  // :284-288
    Sc g_exp = (-r_prime) * y * s_prime + g_exp_c * e_sqr;  // :290
    Sc h_exp = -d_prime;                                    // :292
    for (auto& c : V_exp_c) V_exp.push_back(c * e_sqr);     // :294-297
    MulVec<G> mv;  // :299-320
    mv.add_scalar(Sc::from_i32(1)); mv.add_scalar(e); mv.add_scalar(e_sqr); mv.add_scalar(g_exp); mv.add_scalar(h_exp);
    mv.add_scalars(Ls_exp); mv.add_scalars(Rs_exp); mv.add_scalars(G_exp); mv.add_scalars(H_exp); mv.add_scalars(V_exp);
    mv.add_point(B); mv.add_point(A); mv.add_point(A_prime); mv.add_point(pk.g); mv.add_point(pk.h);
    mv.add_points(L_vec); mv.add_points(R_vec); mv.add_points(pk.G_vec); mv.add_points(pk.H_vec); mv.add_points(V);
    return G::is_identity(mv.calculate());  // :322-327
  }
};

}  // namespace orc

// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Literal restatement of src/range/mod.rs (RangeProof::{prove,verify} and the four private
// functions they dispatch to) and src/range/prover.rs, generic over the group.
#pragma once
#include "protocol.hpp"

namespace orc {

template <class G>
struct RangeProver {  // src/range/prover.rs:13-42
  typedef typename G::Sc Sc;
  typedef typename G::Point Point;
  std::vector<uint64_t> v_vec;
  std::vector<Sc> gamma_vec;
  std::vector<Point> commitment_vec;
  void commit(const PublicKey<G>& pk, uint64_t v, const Sc& gamma) {
    v_vec.push_back(v);
    gamma_vec.push_back(gamma);
    commitment_vec.push_back(pk.commit_u64(v, gamma));
  }
};

template <class G>
struct RangeProof {  // src/range/mod.rs:24-28
  typedef typename G::Sc Sc;
  typedef typename G::Point Point;
  Point A;
  WipProof<G> proof;

  // src/range/mod.rs:31-55
  static RangeProof prove(const PublicKey<G>& pk, size_t n, const RangeProver<G>& prover, const Params<G>& prm) {
    size_t m = prover.v_vec.size();
    if (m == 1) return prove_single(pk, n, prover.v_vec[0], prover.gamma_vec[0], prover.commitment_vec[0], prm);
    return prove_multiple(pk, n, m, prover.v_vec, prover.gamma_vec, prover.commitment_vec, prm);
  }
  // src/range/mod.rs:57-78. true = Ok(()), false = Err(ProofError::VerificationError)
  bool verify(const PublicKey<G>& pk, size_t n, const std::vector<Point>& commitment_vec, const Params<G>& prm) const {
    size_t m = commitment_vec.size();
    if (m == 1) return verify_single(pk, n, commitment_vec[0], prm);
    return verify_multiple(pk, n, m, commitment_vec, prm);
  }

  // src/range/mod.rs:80-187
  static RangeProof prove_single(const PublicKey<G>& pk, size_t n, uint64_t v, const Sc& gamma, const Point& commitment,
                                 const Params<G>& prm) {
    if (pk.G_vec.size() != n || pk.H_vec.size() != n) throw std::runtime_error("range: generator count");  // :90-91
    Sc alpha = prm.alpha;  // :94
    std::vector<uint8_t> v_bits;
    Point A = scalar_mul<G>(pk.h, alpha);  // :98
    for (size_t i = 0; i < n; i++) {       // :100-106
      v_bits.push_back(i < 64 ? (uint8_t)((v >> i) & 1) : 0);
      Point point = v_bits[i] ? pk.G_vec[i] : G::neg(pk.H_vec[i]);
      A = G::add(A, point);
    }
    Sc y = prm.y, z = prm.z;  // :109-110
    Sc one = Sc::from_i32(1), two = Sc::from_i32(2);
    std::vector<Sc> power_of_two = exp_iter_type1(two, n);  // :115
    std::vector<Sc> power_of_y = exp_iter_type2(y, n);      // :116
    Point G_vec_sum = G::identity();                        // :119-122
    for (auto& p : pk.G_vec) G_vec_sum = G::add(G_vec_sum, p);
    Sc G_vec_sum_exp = -z;
    std::vector<Sc> H_exp;  // :125-129
    for (size_t i = 0; i < n; i++) H_exp.push_back(power_of_two[i] * power_of_y[n - 1 - i] + z);
    Sc V_exp = scalar_exp_vartime(y, (uint64_t)(n + 1));  // :131
    Sc g_exp = Sc::from_i32(0);                           // :133-138
    for (auto& x : power_of_y) g_exp = g_exp + x;
    g_exp = g_exp * (z - (z * z));
    g_exp = g_exp - ((scalar_exp_vartime(two, (uint64_t)n) - one) * V_exp * z);
    MulVec<G> mv;  // :140-153
    mv.add_scalar(Sc::from_i32(1)); mv.add_scalar(G_vec_sum_exp); mv.add_scalars(H_exp); mv.add_scalar(g_exp); mv.add_scalar(V_exp);
    mv.add_point(A); mv.add_point(G_vec_sum); mv.add_points(pk.H_vec); mv.add_point(pk.g); mv.add_point(commitment);
    Point A_hat = mv.calculate();
    (void)A_hat;  // only feeds the dead P update inside WIP::prove
    Sc nz = -z, one_minus_z = one - z;  // :156-157
    std::vector<Sc> a_vec, b_vec;
    for (size_t i = 0; i < n; i++) a_vec.push_back(v_bits[i] == 0 ? nz : one_minus_z);           // :159-162
    for (size_t i = 0; i < n; i++) b_vec.push_back(v_bits[i] == 0 ? (H_exp[i] - one) : H_exp[i]);  // :164-170
    Sc alpha_hat = alpha + gamma * V_exp;  // :172
    RangeProof out;
    out.A = A;
    out.proof = WipProof<G>::prove(pk, a_vec, b_vec, power_of_y, alpha_hat, prm);  // :175-182
    return out;
  }

  // src/range/mod.rs:189-238
  bool verify_single(const PublicKey<G>& pk, size_t n, const Point& commitment, const Params<G>& prm) const {
    Sc y = prm.y, z = prm.z;  // :198-199
    Sc one = Sc::from_i32(1), two = Sc::from_i32(2);
    std::vector<Sc> power_of_two = exp_iter_type1(two, n);
    std::vector<Sc> power_of_y = exp_iter_type2(y, n);
    std::vector<Sc> G_exp(n, -z);  // :213
    std::vector<Sc> H_exp;         // :214-218
    for (size_t i = 0; i < n; i++) H_exp.push_back(power_of_two[i] * power_of_y[n - 1 - i] + z);
    Sc V_exp = scalar_exp_vartime(y, (uint64_t)(n + 1));  // :219
    Sc g_exp = Sc::from_i32(0);                           // :221-226
    for (auto& x : power_of_y) g_exp = g_exp + x;
    g_exp = g_exp * (z - (z * z));
    g_exp = g_exp - ((scalar_exp_vartime(two, (uint64_t)n) - one) * V_exp * z);
    return proof.verify(pk, power_of_y, G_exp, H_exp, g_exp, std::vector<Sc>{V_exp}, A, std::vector<Point>{commitment}, prm);  // :228-237
  }

  // src/range/mod.rs:240-403
  static RangeProof prove_multiple(const PublicKey<G>& pk, size_t n, size_t m, const std::vector<uint64_t>& v,
                                   const std::vector<Sc>& gamma_vec, const std::vector<Point>& commitment_vec,
                                   const Params<G>& prm) {
    size_t mn = n * m;
    if (pk.G_vec.size() != mn || pk.H_vec.size() != mn) throw std::runtime_error("range: generator count");  // :252-253
    Sc alpha = prm.alpha;  // :256
    std::vector<uint8_t> v_bits;
    Point A = scalar_mul<G>(pk.h, alpha);  // :260
    for (size_t i = 0; i < mn; i++) {      // :263-277
      size_t index1 = i % n, index2 = i / n;
      v_bits.push_back(index1 < 64 ? (uint8_t)((v[index2] >> index1) & 1) : 0);
      Point point = v_bits[i] ? pk.G_vec[i] : G::neg(pk.H_vec[i]);
      A = G::add(A, point);
    }
    Sc y = prm.y, z = prm.z;  // :278-279
    std::vector<Sc> power_of_two = exp_iter_type1(Sc::from_i32(2), n);  // :283
    std::vector<Sc> power_of_y = exp_iter_type2(y, mn);                 // :284
    Sc z_sqr = z * z;
    std::vector<Sc> power_of_z = exp_iter_type2(z_sqr, m);  // :287
    std::vector<Sc> d;                                      // :288-291
    for (auto& ez : power_of_z) for (auto& e2 : power_of_two) d.push_back(e2 * ez);
    Sc G_vec_sum_exp = -z;  // :296
    std::vector<Sc> H_exp;  // :298-302
    for (size_t i = 0; i < mn; i++) H_exp.push_back(d[i] * power_of_y[mn - 1 - i] + z);
    Sc y_mn1 = scalar_exp_vartime(y, (uint64_t)(mn + 1));  // :304
    std::vector<Sc> V_exp;                                 // :306-309
    for (auto& pz : power_of_z) V_exp.push_back(pz * y_mn1);
    Sc g_exp = Sc::from_i32(0);  // :311-321
    for (auto& x : power_of_y) g_exp = g_exp + x;
    g_exp = g_exp * (z - z_sqr);
    Sc d_sum = Sc::from_i32(0);
    for (auto& x : d) d_sum = d_sum + x;
    g_exp = g_exp - (d_sum * y_mn1 * z);
    Point G_vec_sum = G::identity();  // :324-327
    for (auto& x : pk.G_vec) G_vec_sum = G::add(G_vec_sum, x);
    MulVec<G> mv;  // :330-343
    mv.add_scalar(Sc::from_i32(1)); mv.add_scalar(G_vec_sum_exp); mv.add_scalars(H_exp); mv.add_scalar(g_exp); mv.add_scalars(V_exp);
    mv.add_point(A); mv.add_point(G_vec_sum); mv.add_points(pk.H_vec); mv.add_point(pk.g); mv.add_points(commitment_vec);
    Point A_hat = mv.calculate();
    (void)A_hat;
    Sc one = Sc::from_i32(1), nz = -z, one_minus_z = one - z;  // :347-349
    std::vector<Sc> a_vec, b_vec;
    for (size_t i = 0; i < mn; i++) a_vec.push_back(v_bits[i] != 0 ? one_minus_z : nz);            // :351-355
    for (size_t i = 0; i < mn; i++) b_vec.push_back(v_bits[i] != 0 ? H_exp[i] : (H_exp[i] - one));  // :357-364
    Sc pzg = Sc::from_i32(0);  // :366-374
    for (size_t j = 0; j < m; j++) pzg = pzg + power_of_z[j] * gamma_vec[j];
    Sc alpha_hat = alpha + pzg * y_mn1;  // :376
    RangeProof out;
    out.A = A;
    out.proof = WipProof<G>::prove(pk, a_vec, b_vec, power_of_y, alpha_hat, prm);  // :382-389
    return out;
  }

  // src/range/mod.rs:405-510
  bool verify_multiple(const PublicKey<G>& pk, size_t n, size_t m, const std::vector<Point>& commitment_vec,
                       const Params<G>& prm) const {
    size_t mn = n * m;
    Sc y = prm.y, z = prm.z;  // :417-418
    Sc minus_z = -z, z_sqr = z * z;
    std::vector<Sc> power_of_two = exp_iter_type1(Sc::from_i32(2), n);  // :423
    std::vector<Sc> power_of_y = exp_iter_type2(y, mn + 1);             // :424-428
    Sc y_mn1 = power_of_y.back();
    power_of_y.pop_back();
    std::vector<Sc> power_of_z = exp_iter_type2(z_sqr, m);  // :432
    std::vector<Sc> concat_z_and_2;                         // :435-438
    for (auto& ez : power_of_z) for (auto& e2 : power_of_two) concat_z_and_2.push_back(e2 * ez);
    std::vector<Sc> ch_sqr, ch_inv_sqr, s_vec; Sc e;
    if (!proof.verification_scalars(mn, prm, ch_sqr, ch_inv_sqr, s_vec, e)) return false;  // :441-442
    Sc e_inv = e.inv(), e_sqr = e * e, e_sqr_inv = e_sqr.inv();                            // :445-447
    Sc r_prime_e_inv_y = proof.r_prime * e_inv * y, s_prime_e_inv = proof.s_prime * e_inv;  // :448-449
    const Sc &r_prime = proof.r_prime, &s_prime = proof.s_prime, &d_prime = proof.d_prime;
    std::vector<Sc> yinv = exp_iter_type2(y.inv(), mn);
    std::vector<Sc> G_exp, H_exp, V_exp;
    for (size_t i = 0; i < mn; i++) G_exp.push_back(minus_z - s_vec[i] * yinv[i] * r_prime_e_inv_y);  // :456-459
    for (size_t i = 0; i < mn; i++)                                                                  // :461-465
      H_exp.push_back((-s_prime_e_inv) * s_vec[mn - 1 - i] + (concat_z_and_2[i] * power_of_y[mn - 1 - i] + z));
    Sc sum_y = sum_of_powers_type2(y, mn);                   // :467
    Sc sum_2 = sum_of_powers_type1(Sc::from_i32(2), n);      // :468
    Sc sum_z = sum_of_powers_type2(z_sqr, m);                // :469
    Sc g_exp = (-r_prime) * s_prime * y * e_sqr_inv + (sum_y * (z - z_sqr) - y_mn1 * z * sum_2 * sum_z);  // :471
    Sc h_exp = (-d_prime) * e_sqr_inv;                                                                   // :472
    for (auto& pz : power_of_z) V_exp.push_back(pz * y_mn1);  // :475-477
    MulVec<G> mv;  // :480-503
    mv.add_scalar(Sc::from_i32(1)); mv.add_scalar(e_inv); mv.add_scalar(e_sqr_inv); mv.add_scalar(g_exp); mv.add_scalar(h_exp);
    mv.add_scalars(ch_sqr); mv.add_scalars(ch_inv_sqr); mv.add_scalars(G_exp); mv.add_scalars(H_exp); mv.add_scalars(V_exp);
    mv.add_point(A); mv.add_point(proof.A); mv.add_point(proof.B); mv.add_point(pk.g); mv.add_point(pk.h);
    mv.add_points(proof.L_vec); mv.add_points(proof.R_vec); mv.add_points(pk.G_vec); mv.add_points(pk.H_vec);
    mv.add_points(commitment_vec);
    return G::is_identity(mv.calculate());  // :505-509
  }
};

}  // namespace orc

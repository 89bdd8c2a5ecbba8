// C++ host mirror of the reference crate's public API over the C ABI of bpp.h (header-only, C++17).
//
// The reference is a Rust crate; its toolchain is not in the build image, so this is the compiled-language host side: same
// type and method names, argument meaning and error behaviour as the crate, everything computed by libbpp_b200.so on the GPU.
// Paths below are relative to the reference's src/.
//   Arith::init                                   bls12_381/building_block/arith.rs:9-17
//   PrimeFieldElem::new(i32)                      bls12_381/building_block/scalar/prime_field_elem.rs:191-195
//   PublicKey::new / commitment                   publickey.rs:13-52
//   RangeProver::{new, commit}                    range/prover.rs:13-42
//   RangeVerifier::allocate                       named by the task statement, absent from the snapshot
//   RangeProof::{prove, verify}                   range/mod.rs:24-78
//   WeightedInnerProductProof::{prove, verify}    weighted_inner_product_proof.rs:25-43, :238-248
//   MulVec::{new, add_*, calculate}               bls12_381/building_block/mulvec.rs:13-53
//   ProofError                                    errors.rs:14-50
// Error behaviour: what the reference returns as Err(ProofError::VerificationError) comes back as Result{error = ...};
// what the reference panics on (length mismatch, not a power of two, wrong generator count, failed init) throws Panic.
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "bpp.h"

namespace bulletproofsplus {

struct Panic : std::logic_error {
  using std::logic_error::logic_error;
};

enum class ProofError { VerificationError, FormatError };  // errors.rs:17-20
inline const char* to_string(ProofError e) { return e == ProofError::VerificationError ? "Proof verification failed." : "Proof data could not be parsed."; }

// Result<(), ProofError>
struct Result {
  std::optional<ProofError> error;
  bool is_ok() const { return !error.has_value(); }
  bool operator==(const Result& o) const { return error == o.error; }
  static Result Ok() { return Result{}; }
  static Result Err(ProofError e) { return Result{e}; }
};

enum class Curve : int { Ristretto255 = BPP_RISTRETTO255, Secp256k1 = BPP_SECP256K1, Bls12_381_G1 = BPP_BLS12_381_G1 };

namespace detail {
// group orders, little-endian
inline const std::array<uint8_t, 32>& order_le(Curve c) {
  static const std::array<uint8_t, 32> ell = {0xed, 0xd3, 0xf5, 0x5c, 0x1a, 0x63, 0x12, 0x58, 0xd6, 0x9c, 0xf7, 0xa2, 0xde, 0xf9, 0xde, 0x14,
                                              0,    0,    0,    0,    0,    0,    0,    0,    0,    0,    0,    0,    0,    0,    0,    0x10};
  static const std::array<uint8_t, 32> secp = {0x41, 0x41, 0x36, 0xd0, 0x8c, 0x5e, 0xd2, 0xbf, 0x3b, 0xa0, 0x48, 0xaf, 0xe6, 0xdc, 0xae, 0xba,
                                               0xfe, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff};
  static const std::array<uint8_t, 32> bls = {0x01, 0x00, 0x00, 0x00, 0xff, 0xff, 0xff, 0xff, 0xfe, 0x5b, 0xfe, 0xff, 0x02, 0xa4, 0xbd, 0x53,
                                              0x05, 0xd8, 0xa1, 0x09, 0x08, 0xd8, 0x39, 0x33, 0x48, 0x7d, 0x9d, 0x29, 0x53, 0xa7, 0xed, 0x73};
  return c == Curve::Ristretto255 ? ell : c == Curve::Secp256k1 ? secp : bls;
}
}  // namespace detail

// A scalar of the group order's field as its 32-byte little-endian canonical value. Host-side field arithmetic is not on the
// path (the reference's Add/Mul/inv are glue around mcl's Fr); this type only carries values across the boundary.
struct PrimeFieldElem {
  std::array<uint8_t, 32> e{};
  PrimeFieldElem() = default;
  // prime_field_elem.rs:191-195: negative n -> modulus - |n|
  static PrimeFieldElem new_(int32_t n, Curve c = Curve::Ristretto255) {
    PrimeFieldElem r;
    uint64_t mag = n < 0 ? (uint64_t)(-(int64_t)n) : (uint64_t)n;
    for (int i = 0; i < 8; i++) r.e[i] = (uint8_t)(mag >> (8 * i));
    if (n < 0) {
      const auto& q = detail::order_le(c);
      int borrow = 0;
      for (int i = 0; i < 32; i++) {
        int d = (int)q[i] - (int)r.e[i] - borrow;
        borrow = d < 0;
        r.e[i] = (uint8_t)(d + (borrow << 8));
      }
    }
    return r;
  }
  static PrimeFieldElem from_bytes(const uint8_t* le32) { PrimeFieldElem r; std::memcpy(r.e.data(), le32, 32); return r; }
  bool operator==(const PrimeFieldElem& o) const { return e == o.e; }
};

// A group element as its boundary encoding (Ristretto255: 32 B RFC 9496; secp256k1 / BLS12-381 G1: x || y little-endian)
struct Point {
  std::vector<uint8_t> bytes;
  bool operator==(const Point& o) const { return bytes == o.bytes; }
  bool operator!=(const Point& o) const { return !(*this == o); }
};

// One context per process and device: Arith::init is a process-global Once in the reference (arith.rs:6-17)
class Arith {
 public:
  static bpp_ctx* init(int device = 0) {
    static std::vector<std::pair<int, bpp_ctx*>> made;
    for (auto& p : made)
      if (p.first == device) return p.second;
    bpp_ctx* ctx = nullptr;
    if (bpp_init(device, nullptr, &ctx) != BPP_OK) throw Panic("Arith::init failed: no usable CUDA device (there is no CPU fallback)");  // arith.rs:12-14
    made.emplace_back(device, ctx);
    return ctx;
  }
};

namespace detail {
inline void check(bpp_ctx* ctx, int rc, const char* what) {
  if (rc == BPP_OK) return;
  std::string msg = std::string(what) + ": error " + std::to_string(rc) + " " + bpp_last_error(ctx);
  throw Panic(msg);
}
inline Result verdict(bpp_ctx* ctx, int rc, const char* what) {
  if (rc == BPP_OK) return Result::Ok();
  if (rc == BPP_ERR_VERIFICATION) return Result::Err(ProofError::VerificationError);
  if (rc == BPP_ERR_FORMAT) return Result::Err(ProofError::FormatError);
  check(ctx, rc, what);
  return Result::Ok();
}
inline std::vector<uint8_t> pack(const std::vector<PrimeFieldElem>& v) {
  std::vector<uint8_t> o(v.size() * 32);
  for (size_t i = 0; i < v.size(); i++) std::memcpy(&o[32 * i], v[i].e.data(), 32);
  return o;
}
inline std::vector<uint8_t> pack(const std::vector<Point>& v, size_t pb) {
  std::vector<uint8_t> o(v.size() * pb);
  for (size_t i = 0; i < v.size(); i++) {
    if (v[i].bytes.size() != pb) throw Panic("point encoding has the wrong length");
    std::memcpy(&o[pb * i], v[i].bytes.data(), pb);
  }
  return o;
}
}  // namespace detail

// publickey.rs:13-18. The generators live on the device; g, h, G_vec, H_vec are exported on first use.
class PublicKey {
 public:
  static PublicKey new_(size_t len, Curve curve = Curve::Ristretto255, bpp_ctx* ctx = nullptr) { return PublicKey(len, curve, ctx ? ctx : Arith::init()); }
  // hash-derived generators (bpp_pk_new_hashed): no known discrete-log relations — the constructor to use outside parity tests
  static PublicKey new_hashed(size_t len, const std::string& domain_sep, bpp_ctx* ctx = nullptr) {
    return PublicKey(len, Curve::Ristretto255, ctx ? ctx : Arith::init(), &domain_sep);
  }
  const Point& g() const { export_(); return g_; }
  const Point& h() const { export_(); return h_; }
  const std::vector<Point>& G_vec() const { export_(); return G_; }
  const std::vector<Point>& H_vec() const { export_(); return H_; }
  // publickey.rs:50-52; the value goes through `as i32` like range/prover.rs:37 when it comes from RangeProver::commit
  Point commitment(uint64_t v, const PrimeFieldElem& gamma) const {
    Point p;
    p.bytes.resize(pb_);
    detail::check(ctx_, bpp_commit_batch(ctx_, pk_.get(), &v, gamma.e.data(), 1, p.bytes.data()), "commitment");
    return p;
  }
  // all 64 bits of v (bpp_commit_batch_u64): the commitment an n = 64 proof with real Fiat-Shamir is about
  Point commitment_u64(uint64_t v, const PrimeFieldElem& gamma) const {
    Point p;
    p.bytes.resize(pb_);
    detail::check(ctx_, bpp_commit_batch_u64(ctx_, pk_.get(), &v, gamma.e.data(), 1, p.bytes.data()), "commitment_u64");
    return p;
  }
  // optional, once per key (bpp_pk_precompute): direct fixed-base tables within max_bytes of HBM; every later prove / verify /
  // commitment MSM against the key takes them. Returns the window width. No counterpart in the reference.
  int precompute(size_t max_bytes) {
    int c = 0;
    detail::check(ctx_, bpp_pk_precompute(ctx_, pk_.get(), max_bytes, &c), "PublicKey::precompute");
    return c;
  }
  size_t len() const { return len_; }
  Curve curve() const { return curve_; }
  size_t point_bytes() const { return pb_; }
  bpp_ctx* ctx() const { return ctx_; }
  const bpp_pk* handle() const { return pk_.get(); }

 private:
  PublicKey(size_t len, Curve curve, bpp_ctx* ctx, const std::string* domain_sep = nullptr)
      : ctx_(ctx), curve_(curve), len_(len), pb_((size_t)bpp_point_bytes((int)curve)) {
    bpp_pk* pk = nullptr;
    if (domain_sep)
      detail::check(ctx, bpp_pk_new_hashed(ctx, (int)curve, len, reinterpret_cast<const uint8_t*>(domain_sep->data()), domain_sep->size(), &pk),
                    "PublicKey::new_hashed");
    else
      detail::check(ctx, bpp_pk_new(ctx, (int)curve, len, &pk), "PublicKey::new");
    pk_ = std::shared_ptr<bpp_pk>(pk, bpp_pk_free);
  }
  void export_() const {
    if (!g_.bytes.empty()) return;
    std::vector<uint8_t> g(pb_), h(pb_), G(pb_ * len_ + 1), H(pb_ * len_ + 1);
    detail::check(ctx_, bpp_pk_export(ctx_, pk_.get(), g.data(), h.data(), G.data(), H.data()), "pk_export");
    g_.bytes = g; h_.bytes = h;
    G_.resize(len_); H_.resize(len_);
    for (size_t i = 0; i < len_; i++) {
      G_[i].bytes.assign(G.begin() + pb_ * i, G.begin() + pb_ * (i + 1));
      H_[i].bytes.assign(H.begin() + pb_ * i, H.begin() + pb_ * (i + 1));
    }
  }
  bpp_ctx* ctx_;
  Curve curve_;
  size_t len_, pb_;
  std::shared_ptr<bpp_pk> pk_;
  mutable Point g_, h_;
  mutable std::vector<Point> G_, H_;
};

// range/prover.rs:13-42
class RangeProver {
 public:
  std::vector<uint64_t> v_vec;
  std::vector<PrimeFieldElem> gamma_vec;
  std::vector<Point> commitment_vec;
  static RangeProver new_() { return RangeProver(); }
  void commit(const PublicKey& pk, uint64_t v, const PrimeFieldElem& gamma) {
    v_vec.push_back(v);
    gamma_vec.push_back(gamma);
    commitment_vec.push_back(pk.commitment(v, gamma));
  }
  // without the reference's `v as i32` (prover.rs:37): use with prove_with_transcript when values may reach 2^31
  void commit_u64(const PublicKey& pk, uint64_t v, const PrimeFieldElem& gamma) {
    v_vec.push_back(v);
    gamma_vec.push_back(gamma);
    commitment_vec.push_back(pk.commitment_u64(v, gamma));
  }
};

// The verifier's side of the statement (task statement: RangeVerifier::allocate; the snapshot passes a bare slice)
class RangeVerifier {
 public:
  std::vector<Point> commitment_vec;
  static RangeVerifier new_() { return RangeVerifier(); }
  void allocate(const std::vector<Point>& commitments) { commitment_vec.insert(commitment_vec.end(), commitments.begin(), commitments.end()); }
};

// weighted_inner_product_proof.rs:25-33
class WeightedInnerProductProof {
 public:
  std::vector<Point> L_vec, R_vec;
  Point A, B;
  PrimeFieldElem r_prime, s_prime, d_prime;

  // wire layout L[0..L) | R[0..L) | A | B | r' | s' | d'
  std::vector<uint8_t> to_bytes() const {
    std::vector<uint8_t> o;
    auto put = [&](const std::vector<uint8_t>& b) { o.insert(o.end(), b.begin(), b.end()); };
    for (auto& p : L_vec) put(p.bytes);
    for (auto& p : R_vec) put(p.bytes);
    put(A.bytes); put(B.bytes);
    for (auto* s : {&r_prime, &s_prime, &d_prime}) o.insert(o.end(), s->e.begin(), s->e.end());
    return o;
  }
  static WeightedInnerProductProof from_bytes(const uint8_t* b, size_t len, size_t pb) {
    if (len < 96 + 2 * pb || (len - 96) % pb || ((len - 96) / pb) % 2) throw Panic("wip proof: bad length");
    const size_t rounds = ((len - 96) / pb - 2) / 2;
    WeightedInnerProductProof w;
    auto pt = [&](size_t i) { Point p; p.bytes.assign(b + i * pb, b + (i + 1) * pb); return p; };
    for (size_t i = 0; i < rounds; i++) w.L_vec.push_back(pt(i));
    for (size_t i = 0; i < rounds; i++) w.R_vec.push_back(pt(rounds + i));
    w.A = pt(2 * rounds); w.B = pt(2 * rounds + 1);
    const uint8_t* s = b + (2 * rounds + 2) * pb;
    w.r_prime = PrimeFieldElem::from_bytes(s); w.s_prime = PrimeFieldElem::from_bytes(s + 32); w.d_prime = PrimeFieldElem::from_bytes(s + 64);
    return w;
  }
  // weighted_inner_product_proof.rs:36-43. `commitment` is accepted and unused, as in the reference (dead update :137-142).
  static WeightedInnerProductProof prove(const PublicKey& pk, const std::vector<PrimeFieldElem>& a_vec, const std::vector<PrimeFieldElem>& b_vec,
                                         const std::vector<PrimeFieldElem>& power_of_y_vec, const PrimeFieldElem& gamma,
                                         const Point* commitment = nullptr, const bpp_params* params = nullptr) {
    (void)commitment;
    const size_t n = pk.len();
    if (a_vec.size() != n || b_vec.size() != n || power_of_y_vec.size() != n) throw Panic("wip: all input vectors must have the same length");  // :61-64
    if (n & (n - 1)) throw Panic("wip: length must be a power of two");                                                                        // :67
    size_t plen = bpp_wip_proof_len((int)pk.curve(), n);
    std::vector<uint8_t> out(plen);
    detail::check(pk.ctx(), bpp_wip_prove(pk.ctx(), pk.handle(), detail::pack(a_vec).data(), detail::pack(b_vec).data(),
                                          detail::pack(power_of_y_vec).data(), gamma.e.data(), params, out.data(), &plen), "wip_prove");
    return from_bytes(out.data(), plen, pk.point_bytes());
  }
  // weighted_inner_product_proof.rs:238-248
  Result verify(const PublicKey& pk, const std::vector<PrimeFieldElem>& power_of_y_vec, const std::vector<PrimeFieldElem>& G_exp_of_commitment,
                const std::vector<PrimeFieldElem>& H_exp_of_commitment, const PrimeFieldElem& g_exp_of_commitment,
                const std::vector<PrimeFieldElem>& V_exp_of_commitment, const Point& A_prime, const std::vector<Point>& V,
                const bpp_params* params = nullptr) const {
    const size_t pb = pk.point_bytes();
    if (V_exp_of_commitment.size() != V.size()) throw Panic("wip verify: V and its exponents differ in length");
    auto proof = to_bytes();
    auto Vb = detail::pack(V, pb), Ve = detail::pack(V_exp_of_commitment);
    int rc = bpp_wip_verify(pk.ctx(), pk.handle(), detail::pack(power_of_y_vec).data(), detail::pack(G_exp_of_commitment).data(),
                            detail::pack(H_exp_of_commitment).data(), g_exp_of_commitment.e.data(), V.size(), V.empty() ? nullptr : Ve.data(),
                            A_prime.bytes.data(), V.empty() ? nullptr : Vb.data(), proof.data(), proof.size(), params);
    return detail::verdict(pk.ctx(), rc, "wip_verify");
  }
};

// range/mod.rs:24-28
class RangeProof {
 public:
  Point A;
  WeightedInnerProductProof proof;

  std::vector<uint8_t> to_bytes() const {
    std::vector<uint8_t> o = A.bytes, w = proof.to_bytes();
    o.insert(o.end(), w.begin(), w.end());
    return o;
  }
  static RangeProof from_bytes(const uint8_t* b, size_t len, size_t pb) {
    if (len < pb) throw Panic("range proof: bad length");
    RangeProof r;
    r.A.bytes.assign(b, b + pb);
    r.proof = WeightedInnerProductProof::from_bytes(b + pb, len - pb, pb);
    return r;
  }
  // range/mod.rs:31-55 (m == 1 -> prove_single, else prove_multiple: decided inside the library)
  static RangeProof prove(const PublicKey& pk, size_t n, const RangeProver& prover, const bpp_params* params = nullptr) {
    const size_t m = prover.v_vec.size();
    if (pk.len() != n * m) throw Panic("pk.G_vec.len() != n*m");  // range/mod.rs:90-91, :252-253
    size_t plen = bpp_range_proof_len((int)pk.curve(), n, m);
    if (!plen) throw Panic("n*m must be a power of two");
    std::vector<uint8_t> out(plen);
    detail::check(pk.ctx(), bpp_range_prove(pk.ctx(), pk.handle(), n, m, prover.v_vec.data(), detail::pack(prover.gamma_vec).data(), params,
                                            out.data(), &plen), "range_prove");
    return from_bytes(out.data(), plen, pk.point_bytes());
  }
  // range/mod.rs:57-78
  Result verify(const PublicKey& pk, size_t n, const std::vector<Point>& commitment_vec, const bpp_params* params = nullptr) const {
    const size_t m = commitment_vec.size();
    if (pk.len() != n * m) throw Panic("pk.G_vec.len() != n*m");
    auto p = to_bytes();
    auto V = detail::pack(commitment_vec, pk.point_bytes());
    return detail::verdict(pk.ctx(), bpp_range_verify(pk.ctx(), pk.handle(), n, m, V.data(), p.data(), p.size(), params), "range_verify");
  }
  Result verify(const PublicKey& pk, size_t n, const RangeVerifier& verifier, const bpp_params* params = nullptr) const {
    return verify(pk, n, verifier.commitment_vec, params);
  }
};

// bls12_381/building_block/mulvec.rs:13-53
class MulVec {
 public:
  explicit MulVec(Curve curve = Curve::Ristretto255, bpp_ctx* ctx = nullptr) : curve_(curve), ctx_(ctx ? ctx : Arith::init()) {}
  static MulVec new_(Curve curve = Curve::Ristretto255) { return MulVec(curve); }
  void add_scalar(const PrimeFieldElem& s) { scalars_.push_back(s); }
  void add_scalars(const std::vector<PrimeFieldElem>& ss) { scalars_.insert(scalars_.end(), ss.begin(), ss.end()); }
  void add_point(const Point& p) { points_.push_back(p); }
  void add_points(const std::vector<Point>& ps) { points_.insert(points_.end(), ps.begin(), ps.end()); }
  Point calculate() const {
    if (scalars_.size() != points_.size()) throw Panic("mulvec: lengths of scalars and points must match");  // mulvec.rs:23-25
    const size_t pb = (size_t)bpp_point_bytes((int)curve_);
    Point out;
    out.bytes.resize(pb);
    auto s = detail::pack(scalars_);
    auto p = detail::pack(points_, pb);
    int rc = bpp_msm(ctx_, (int)curve_, s.empty() ? nullptr : s.data(), p.empty() ? nullptr : p.data(), scalars_.size(), out.bytes.data());
    if (rc == BPP_ERR_FORMAT) throw Panic("mulvec: a point or scalar is not a canonical encoding");
    detail::check(ctx_, rc, "msm");
    return out;
  }

 private:
  Curve curve_;
  bpp_ctx* ctx_;
  std::vector<PrimeFieldElem> scalars_;
  std::vector<Point> points_;
};

}  // namespace bulletproofsplus

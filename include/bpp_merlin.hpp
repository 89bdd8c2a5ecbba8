// Host-side Fiat-Shamir for the C++ host: Merlin transcripts (STROBE-128 over Keccak-f[1600]) and the range-proof transcript
// protocol, the C++ twin of bulletproofsplus_b200/merlin.py (label set and protocol are documented there). The library never
// sees a transcript: the host feeds the challenges to bpp_range_prove_step / bpp_range_verify_batch_fs.
// Pinned to the merlin crate's `equivalence_simple` test value by tests/cpp/host_mirror_test.cpp.
#pragma once
#include <sys/random.h>
#include <array>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "bpp.hpp"

namespace bulletproofsplus {
namespace merlin {

inline void keccak_f1600(uint8_t* st) {
  static const uint64_t RC[24] = {0x0000000000000001ull, 0x0000000000008082ull, 0x800000000000808Aull, 0x8000000080008000ull, 0x000000000000808Bull,
                                  0x0000000080000001ull, 0x8000000080008081ull, 0x8000000000008009ull, 0x000000000000008Aull, 0x0000000000000088ull,
                                  0x0000000080008009ull, 0x000000008000000Aull, 0x000000008000808Bull, 0x800000000000008Bull, 0x8000000000008089ull,
                                  0x8000000000008003ull, 0x8000000000008002ull, 0x8000000000000080ull, 0x000000000000800Aull, 0x800000008000000Aull,
                                  0x8000000080008081ull, 0x8000000000008080ull, 0x0000000080000001ull, 0x8000000080008008ull};
  static const int ROT[5][5] = {{0, 36, 3, 41, 18}, {1, 44, 10, 45, 2}, {62, 6, 43, 15, 61}, {28, 55, 25, 21, 56}, {27, 20, 39, 8, 14}};
  uint64_t a[5][5], b[5][5], c[5], d[5];
  for (int x = 0; x < 5; x++)
    for (int y = 0; y < 5; y++) {
      uint64_t v = 0;
      for (int i = 7; i >= 0; i--) v = (v << 8) | st[8 * (x + 5 * y) + i];
      a[x][y] = v;
    }
  auto rol = [](uint64_t v, int n) { n &= 63; return n ? (v << n) | (v >> (64 - n)) : v; };
  for (int r = 0; r < 24; r++) {
    for (int x = 0; x < 5; x++) c[x] = a[x][0] ^ a[x][1] ^ a[x][2] ^ a[x][3] ^ a[x][4];
    for (int x = 0; x < 5; x++) d[x] = c[(x + 4) % 5] ^ rol(c[(x + 1) % 5], 1);
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) a[x][y] ^= d[x];
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) b[y][(2 * x + 3 * y) % 5] = rol(a[x][y], ROT[x][y]);
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) a[x][y] = b[x][y] ^ (~b[(x + 1) % 5][y] & b[(x + 2) % 5][y]);
    a[0][0] ^= RC[r];
  }
  for (int x = 0; x < 5; x++)
    for (int y = 0; y < 5; y++)
      for (int i = 0; i < 8; i++) st[8 * (x + 5 * y) + i] = (uint8_t)(a[x][y] >> (8 * i));
}

class Strobe128 {
 public:
  static constexpr int R = 166;
  enum : uint8_t { FLAG_I = 1, FLAG_A = 2, FLAG_C = 4, FLAG_T = 8, FLAG_M = 16, FLAG_K = 32 };
  explicit Strobe128(const std::string& protocol_label) {
    std::memset(st_, 0, sizeof st_);
    const uint8_t head[6] = {1, R + 2, 1, 0, 1, 96};
    std::memcpy(st_, head, 6);
    std::memcpy(st_ + 6, "STROBEv1.0.2", 12);
    keccak_f1600(st_);
    meta_ad(reinterpret_cast<const uint8_t*>(protocol_label.data()), protocol_label.size(), false);
  }
  void meta_ad(const uint8_t* d, size_t n, bool more) { begin_op(FLAG_M | FLAG_A, more); absorb(d, n); }
  void ad(const uint8_t* d, size_t n, bool more) { begin_op(FLAG_A, more); absorb(d, n); }
  void prf(uint8_t* out, size_t n, bool more) { begin_op(FLAG_I | FLAG_A | FLAG_C, more); squeeze(out, n); }

 private:
  void run_f() {
    st_[pos_] ^= pos_begin_;
    st_[pos_ + 1] ^= 0x04;
    st_[R + 1] ^= 0x80;
    keccak_f1600(st_);
    pos_ = 0; pos_begin_ = 0;
  }
  void absorb(const uint8_t* d, size_t n) {
    for (size_t i = 0; i < n; i++) { st_[pos_++] ^= d[i]; if (pos_ == R) run_f(); }
  }
  void squeeze(uint8_t* o, size_t n) {
    for (size_t i = 0; i < n; i++) { o[i] = st_[pos_]; st_[pos_++] = 0; if (pos_ == R) run_f(); }
  }
  void begin_op(uint8_t flags, bool more) {
    if (more) { if (flags != cur_flags_) throw Panic("strobe: flag mismatch in a continued operation"); return; }
    const uint8_t old_begin = pos_begin_;
    pos_begin_ = (uint8_t)(pos_ + 1);
    cur_flags_ = flags;
    const uint8_t hdr[2] = {old_begin, flags};
    absorb(hdr, 2);
    if ((flags & (FLAG_C | FLAG_K)) && pos_ != 0) run_f();
  }
  uint8_t st_[200];
  uint8_t pos_ = 0, pos_begin_ = 0, cur_flags_ = 0;
};

// merlin::Transcript
class Transcript {
 public:
  explicit Transcript(const std::string& label) : strobe_("Merlin v1.0") { append_message("dom-sep", label); }
  void append_message(const std::string& label, const uint8_t* msg, size_t n) {
    uint8_t len[4] = {(uint8_t)n, (uint8_t)(n >> 8), (uint8_t)(n >> 16), (uint8_t)(n >> 24)};
    strobe_.meta_ad(reinterpret_cast<const uint8_t*>(label.data()), label.size(), false);
    strobe_.meta_ad(len, 4, true);
    strobe_.ad(msg, n, false);
  }
  void append_message(const std::string& label, const std::string& msg) { append_message(label, reinterpret_cast<const uint8_t*>(msg.data()), msg.size()); }
  void append_message(const std::string& label, const std::vector<uint8_t>& msg) { append_message(label, msg.data(), msg.size()); }
  void append_u64(const std::string& label, uint64_t x) {
    uint8_t b[8];
    for (int i = 0; i < 8; i++) b[i] = (uint8_t)(x >> (8 * i));
    append_message(label, b, 8);
  }
  void challenge_bytes(const std::string& label, uint8_t* out, size_t n) {
    uint8_t len[4] = {(uint8_t)n, (uint8_t)(n >> 8), (uint8_t)(n >> 16), (uint8_t)(n >> 24)};
    strobe_.meta_ad(reinterpret_cast<const uint8_t*>(label.data()), label.size(), false);
    strobe_.meta_ad(len, 4, true);
    strobe_.prf(out, n, false);
  }
  // 64 transcript bytes reduced mod the group order; a zero result is drawn again
  PrimeFieldElem challenge_scalar(const std::string& label, Curve curve) {
    for (;;) {
      uint8_t wide[64];
      challenge_bytes(label, wide, 64);
      PrimeFieldElem e;
      if (reduce_wide(wide, curve, e)) return e;
    }
  }
  // 512-bit little-endian value mod the group order; false when the result is zero
  static bool reduce_wide(const uint8_t* wide, Curve curve, PrimeFieldElem& e) {
    const auto& q = detail::order_le(curve);
    uint64_t r[5] = {0, 0, 0, 0, 0}, ql[4];
    for (int i = 0; i < 4; i++) {
      ql[i] = 0;
      for (int k = 7; k >= 0; k--) ql[i] = (ql[i] << 8) | q[8 * i + k];
    }
    for (int bit = 511; bit >= 0; bit--) {  // r = 2r + bit, minus q when it does not fit
      for (int i = 4; i > 0; i--) r[i] = (r[i] << 1) | (r[i - 1] >> 63);
      r[0] = (r[0] << 1) | ((wide[bit >> 3] >> (bit & 7)) & 1);
      bool ge = r[4] != 0;
      if (!ge) {
        ge = true;
        for (int i = 3; i >= 0; i--)
          if (r[i] != ql[i]) { ge = r[i] > ql[i]; break; }
      }
      if (ge) {
        unsigned __int128 borrow = 0;
        for (int i = 0; i < 4; i++) {
          unsigned __int128 d = (unsigned __int128)r[i] - ql[i] - borrow;
          r[i] = (uint64_t)d;
          borrow = (d >> 64) & 1;
        }
        r[4] -= (uint64_t)borrow;
      }
    }
    bool nz = false;
    for (int i = 0; i < 32; i++) { e.e[i] = (uint8_t)(r[i >> 3] >> (8 * (i & 7))); nz = nz || e.e[i]; }
    return nz;
  }

 private:
  Strobe128 strobe_;
};

// protocol "rangeproof v1" of merlin.py, one instance per proof
class RangeProofTranscript {
 public:
  RangeProofTranscript(Transcript t, Curve curve, size_t n, size_t m, const std::vector<Point>& commitments) : t_(std::move(t)), curve_(curve), N_(n * m) {
    t_.append_message("dom-sep", std::string("rangeproof v1"));
    t_.append_u64("n", n);
    t_.append_u64("m", m);
    for (auto& V : commitments) t_.append_message("V", V.bytes);
  }
  void after_A(const Point& A, PrimeFieldElem& y, PrimeFieldElem& z) {
    t_.append_message("A", A.bytes);
    y = t_.challenge_scalar("y", curve_);
    z = t_.challenge_scalar("z", curve_);
    t_.append_message("dom-sep", std::string("wipp v1"));
    t_.append_u64("n", N_);
  }
  PrimeFieldElem after_round(const Point& L, const Point& R) {
    t_.append_message("L", L.bytes);
    t_.append_message("R", R.bytes);
    return t_.challenge_scalar("e", curve_);
  }
  PrimeFieldElem after_final(const Point& A, const Point& B) {
    t_.append_message("A", A.bytes);
    t_.append_message("B", B.bytes);
    return t_.challenge_scalar("e", curve_);
  }

 private:
  Transcript t_;
  Curve curve_;
  size_t N_;
};

}  // namespace merlin

// The prover's blinding scalars of one proof: alpha, r, s, delta, eta, d_L[L], d_R[L] (L = log2(n*m)). A proof is zero-knowledge
// only if these are fresh, secret and uniform: random() draws them from the operating system's CSPRNG. reference() is the
// reference snapshot's PUBLIC constants (range/mod.rs:94,256; wip.rs:94-95,175-178): with them anyone can solve the proof
// scalars for gamma and the bits of v — parity tests only.
struct ProverNonces {
  std::vector<uint8_t> bytes;  // (5 + 2L) x 32 canonical scalars; empty = the reference's constants
  static ProverNonces reference() { return ProverNonces{}; }
  static ProverNonces random(Curve curve, size_t L) {
    ProverNonces nn;
    nn.bytes.resize((5 + 2 * L) * 32);
    for (size_t j = 0; j < 5 + 2 * L; j++) {
      PrimeFieldElem e;
      for (;;) {
        uint8_t wide[64];
        size_t got = 0;
        while (got < sizeof wide) {
          ssize_t k = getrandom(wide + got, sizeof wide - got, 0);
          if (k < 0) throw Panic("getrandom failed");
          got += (size_t)k;
        }
        if (merlin::Transcript::reduce_wide(wide, curve, e)) break;
      }
      std::memcpy(&nn.bytes[32 * j], e.e.data(), 32);
    }
    return nn;
  }
};

// README-style calls with a transcript (README.md:39-52 of the reference): one proof, host-owned Fiat-Shamir. `nonces` NULL =
// fresh random blinding scalars (the only safe choice outside parity tests).
inline RangeProof prove_with_transcript(const PublicKey& pk, size_t n, const RangeProver& prover, merlin::Transcript transcript,
                                        const ProverNonces* nonces = nullptr) {
  const size_t m = prover.v_vec.size(), pb = pk.point_bytes();
  if (pk.len() != n * m) throw Panic("pk.G_vec.len() != n*m");
  bpp_prove_session* s = nullptr;
  Point A;
  A.bytes.resize(pb);
  size_t L0 = 0;
  while (((size_t)1 << L0) < n * m) L0++;
  const ProverNonces fresh = nonces ? ProverNonces{} : ProverNonces::random(pk.curve(), L0);
  const ProverNonces& use = nonces ? *nonces : fresh;
  detail::check(pk.ctx(), bpp_range_prove_begin(pk.ctx(), pk.handle(), n, m, 1, prover.v_vec.data(), detail::pack(prover.gamma_vec).data(),
                                                use.bytes.empty() ? nullptr : use.bytes.data(), &s, A.bytes.data()), "range_prove_begin");
  std::shared_ptr<bpp_prove_session> guard(s, bpp_prove_session_free);
  merlin::RangeProofTranscript rt(std::move(transcript), pk.curve(), n, m, prover.commitment_vec);
  PrimeFieldElem y, z;
  rt.after_A(A, y, z);
  std::vector<uint8_t> ch(64), out(2 * pb);
  std::memcpy(ch.data(), y.e.data(), 32);
  std::memcpy(ch.data() + 32, z.e.data(), 32);
  const int L = bpp_prove_session_rounds(s);
  for (int k = 0; k <= L; k++) {
    detail::check(pk.ctx(), bpp_range_prove_step(pk.ctx(), s, ch.data(), out.data()), "range_prove_step");
    Point P0, P1;
    P0.bytes.assign(out.begin(), out.begin() + pb);
    P1.bytes.assign(out.begin() + pb, out.end());
    PrimeFieldElem e = k < L ? rt.after_round(P0, P1) : rt.after_final(P0, P1);
    ch.assign(e.e.begin(), e.e.end());
  }
  size_t plen = bpp_range_proof_len((int)pk.curve(), n, m);
  std::vector<uint8_t> proof(plen);
  detail::check(pk.ctx(), bpp_range_prove_finish(pk.ctx(), s, ch.data(), proof.data()), "range_prove_finish");
  return RangeProof::from_bytes(proof.data(), plen, pb);
}

inline Result verify_with_transcript(const RangeProof& proof, const PublicKey& pk, size_t n, const std::vector<Point>& commitment_vec,
                                     merlin::Transcript transcript) {
  const size_t m = commitment_vec.size(), pb = pk.point_bytes();
  if (pk.len() != n * m) throw Panic("pk.G_vec.len() != n*m");
  merlin::RangeProofTranscript rt(std::move(transcript), pk.curve(), n, m, commitment_vec);
  const size_t L = proof.proof.L_vec.size();
  std::vector<uint8_t> ch((3 + L) * 32);
  PrimeFieldElem y, z;
  rt.after_A(proof.A, y, z);
  std::memcpy(&ch[0], y.e.data(), 32);
  std::memcpy(&ch[32], z.e.data(), 32);
  for (size_t k = 0; k < L; k++) {
    PrimeFieldElem e = rt.after_round(proof.proof.L_vec[k], proof.proof.R_vec[k]);
    std::memcpy(&ch[(3 + k) * 32], e.e.data(), 32);
  }
  PrimeFieldElem e = rt.after_final(proof.proof.A, proof.proof.B);
  std::memcpy(&ch[64], e.e.data(), 32);
  auto bytes = proof.to_bytes();
  auto V = detail::pack(commitment_vec, pb);
  uint8_t verdict = BPP_ERR_CUDA;
  detail::check(pk.ctx(), bpp_range_verify_batch_fs(pk.ctx(), pk.handle(), n, m, 1, V.data(), bytes.data(), ch.data(), &verdict), "range_verify_batch_fs");
  return detail::verdict(pk.ctx(), verdict, "verify");
}

}  // namespace bulletproofsplus

// Exercises the C++ host mirror (include/bpp.hpp) the way the reference's own tests exercise the crate: small
// known-answer group facts (point.rs:124-185), the README flow, tampering, and the panics. Prints "PASS name" / "FAIL name";
// tests/test_gpu_cpp_host.py runs it on the GPU box and compares the printed encodings with the CPU oracle.
#include <cstdio>
#include <cstdlib>
#include <functional>

#include "bpp.hpp"
#include "bpp_merlin.hpp"

using namespace bulletproofsplus;

static int failures = 0;
static void report(const char* name, bool ok) {
  printf("%s %s\n", ok ? "PASS" : "FAIL", name);
  if (!ok) failures++;
}
static bool panics(const std::function<void()>& f) {
  try { f(); } catch (const Panic&) { return true; }
  return false;
}
static void hex(const char* label, const std::vector<uint8_t>& b) {
  printf("%s=", label);
  for (uint8_t x : b) printf("%02x", x);
  printf("\n");
}

int main(int argc, char** argv) {
  const Curve curve = (Curve)(argc > 1 ? atoi(argv[1]) : 0);
  Arith::init();
  auto fe = [&](int32_t v) { return PrimeFieldElem::new_(v, curve); };

  // generators: g, h = 2g, G_i = 3(i+1)g, H_i = 5(i+1)g (publickey.rs:21-48) => G_4 == H_2 and g + g == h
  PublicKey pk8 = PublicKey::new_(8, curve);
  report("G_4 == H_2 (both 15 g)", pk8.G_vec()[4] == pk8.H_vec()[2]);
  {
    MulVec mv(curve);
    mv.add_scalar(fe(1)); mv.add_point(pk8.g());
    mv.add_scalar(fe(1)); mv.add_point(pk8.g());
    report("g + g == h (point.rs:131-139)", mv.calculate() == pk8.h());
    MulVec three(curve);
    three.add_scalars({fe(3)}); three.add_points({pk8.g()});
    report("3 g == G_0", three.calculate() == pk8.G_vec()[0]);
    MulVec cancel(curve);
    cancel.add_scalars({fe(5), fe(-5)}); cancel.add_points({pk8.g(), pk8.g()});
    MulVec empty(curve);
    report("5 g + (-5) g == Point::zero()", cancel.calculate() == empty.calculate());
    MulVec bad(curve);
    bad.add_scalar(fe(1));
    report("MulVec length mismatch panics (mulvec.rs:23-25)", panics([&] { bad.calculate(); }));
  }
  // commitment = v g + gamma h (publickey.rs:50-52): 2 g + 3 h = 8 g
  {
    MulVec mv(curve);
    mv.add_scalar(fe(8)); mv.add_point(pk8.g());
    report("commitment(2, 3) == 8 g", pk8.commitment(2, fe(3)) == mv.calculate());
  }
  // README / main.rs flow and its failure modes
  for (auto nm : {std::pair<size_t, size_t>{64, 2}, {32, 1}, {8, 4}}) {
    const size_t n = nm.first, m = nm.second;
    PublicKey pk = PublicKey::new_(n * m, curve);
    RangeProver prover = RangeProver::new_();
    for (size_t j = 0; j < m; j++) prover.commit(pk, 2 + 3 * j, fe(3 + 4 * (int)j));
    RangeProof proof = RangeProof::prove(pk, n, prover);
    char label[64];
    snprintf(label, sizeof label, "proof_n%zu_m%zu", n, m);
    hex(label, proof.to_bytes());
    snprintf(label, sizeof label, "verify n=%zu m=%zu == Ok(())", n, m);
    report(label, proof.verify(pk, n, prover.commitment_vec) == Result::Ok());
    RangeProof t = proof;
    t.proof.r_prime.e[0] ^= 1;
    snprintf(label, sizeof label, "tampered r' n=%zu m=%zu == Err(VerificationError)", n, m);
    report(label, t.verify(pk, n, prover.commitment_vec) == Result::Err(ProofError::VerificationError));
    t = proof;
    t.proof.L_vec[0] = pk.g();
    snprintf(label, sizeof label, "swapped L_0 n=%zu m=%zu == Err(VerificationError)", n, m);
    report(label, t.verify(pk, n, prover.commitment_vec) == Result::Err(ProofError::VerificationError));
    std::vector<Point> other = prover.commitment_vec;
    other[0] = pk.commitment(3, fe(3));
    snprintf(label, sizeof label, "other commitment n=%zu m=%zu == Err(VerificationError)", n, m);
    report(label, proof.verify(pk, n, other) == Result::Err(ProofError::VerificationError));
    // round trip through the wire format
    auto bytes = proof.to_bytes();
    RangeProof back = RangeProof::from_bytes(bytes.data(), bytes.size(), pk.point_bytes());
    report("wire round trip", back.to_bytes() == bytes && back.verify(pk, n, prover.commitment_vec).is_ok());
  }
  {
    PublicKey pk = PublicKey::new_(24, curve);  // not a power of two
    RangeProver prover = RangeProver::new_();
    for (int j = 0; j < 3; j++) prover.commit(pk, 1, fe(1));
    report("n*m not a power of two panics (wip.rs:67)", panics([&] { RangeProof::prove(pk, 8, prover); }));
    PublicKey pk16 = PublicKey::new_(16, curve);
    report("generator count != n*m panics (range/mod.rs:252-253)", panics([&] { RangeProof::prove(pk16, 8, prover); }));
  }
  // standalone weighted inner product argument (wip.rs:36-43, :238-248; relation in the comment at :229-237)
  {
    const size_t N = 16;
    PublicKey pk = PublicKey::new_(N, curve);
    std::vector<PrimeFieldElem> a, b, ypow, zeros(N, fe(0));
    int64_t y = 1, dl = 0;  // exponents stay small: P = dl * g is computed with one MulVec term
    for (size_t i = 0; i < N; i++) {
      y *= 2;  // y = 2: y^(i+1) = 2^(i+1)
      a.push_back(fe((int)i + 1)); b.push_back(fe(2 * (int)i + 3)); ypow.push_back(fe((int32_t)y));
      dl += (int64_t)(i + 1) * 3 * (i + 1) + (int64_t)(2 * i + 3) * 5 * (i + 1) + (int64_t)(i + 1) * (2 * i + 3) * y;
    }
    const int gamma = 9;
    dl += 2 * gamma;
    auto w = WeightedInnerProductProof::prove(pk, a, b, ypow, fe(gamma));
    hex("wip_N16", w.to_bytes());
    MulVec mv(curve);
    // dl < 2^31 here
    mv.add_scalar(fe((int32_t)dl)); mv.add_point(pk.g());
    Point P = mv.calculate();
    report("wip verify == Ok(())", w.verify(pk, ypow, zeros, zeros, fe(0), {}, P, {}) == Result::Ok());
    report("wip verify, wrong P == Err", w.verify(pk, ypow, zeros, zeros, fe(0), {}, pk.g(), {}) == Result::Err(ProofError::VerificationError));
    a.pop_back();
    report("wip length mismatch panics (wip.rs:61-64)", panics([&] { WeightedInnerProductProof::prove(pk, a, b, ypow, fe(gamma)); }));
  }
  // host-owned Fiat-Shamir: Merlin's published value, then the README flow with transcripts (README.md:39-52)
  {
    merlin::Transcript t("test protocol");
    t.append_message("some label", std::string("some data"));
    uint8_t c[32];
    t.challenge_bytes("challenge", c, 32);
    static const char* want = "d5a21972d0d5fe320c0d263fac7fffb8145aa640af6e9bca177c03c7efcf0615";
    char got[65];
    for (int i = 0; i < 32; i++) snprintf(got + 2 * i, 3, "%02x", c[i]);
    report("merlin equivalence_simple value", std::string(got) == want);
    const size_t n = 32, m = 4;
    PublicKey pk = PublicKey::new_(n * m, curve);
    RangeProver prover = RangeProver::new_();
    for (size_t j = 0; j < m; j++) prover.commit(pk, 31 + j, fe(11 + (int)j));
    const ProverNonces parity = ProverNonces::reference();  // the public constants: byte-comparable with the Python host and the oracle
    RangeProof proof = prove_with_transcript(pk, n, prover, merlin::Transcript("RangeProof Test"), &parity);
    hex("transcript_proof_n32_m4", proof.to_bytes());
    // default = fresh random blinding scalars: two proofs of one witness differ, and both verify
    RangeProof zk1 = prove_with_transcript(pk, n, prover, merlin::Transcript("RangeProof Test"));
    RangeProof zk2 = prove_with_transcript(pk, n, prover, merlin::Transcript("RangeProof Test"));
    report("default nonces are fresh: two proofs of one witness differ", zk1.to_bytes() != zk2.to_bytes() && zk1.to_bytes() != proof.to_bytes());
    report("random-nonce transcript proofs verify", verify_with_transcript(zk1, pk, n, prover.commitment_vec, merlin::Transcript("RangeProof Test")) == Result::Ok() &&
                                                    verify_with_transcript(zk2, pk, n, prover.commitment_vec, merlin::Transcript("RangeProof Test")) == Result::Ok());
    report("transcript verify == Ok(())", verify_with_transcript(proof, pk, n, prover.commitment_vec, merlin::Transcript("RangeProof Test")) == Result::Ok());
    report("other transcript label == Err(VerificationError)",
           verify_with_transcript(proof, pk, n, prover.commitment_vec, merlin::Transcript("RangeProof Test 2")) == Result::Err(ProofError::VerificationError));
    report("transcript proof under the fixed constants == Err", proof.verify(pk, n, prover.commitment_vec) == Result::Err(ProofError::VerificationError));
    RangeProof t2 = proof;
    t2.proof.R_vec[1] = pk.h();
    report("tampered transcript proof == Err(VerificationError)",
           verify_with_transcript(t2, pk, n, prover.commitment_vec, merlin::Transcript("RangeProof Test")) == Result::Err(ProofError::VerificationError));
  }
  // hash-derived generators (bpp_pk_new_hashed; Ristretto255): the README flow over a key nobody knows a relation in
  if (curve == Curve::Ristretto255) {
    const size_t n = 32, m = 2;
    PublicKey hk = PublicKey::new_hashed(n * m, "host_mirror_test");
    report("hashed key: G_4 != H_2, g + g != h", !(hk.G_vec()[4] == hk.H_vec()[2]) && !(hk.g() == hk.h()));
    RangeProver prover = RangeProver::new_();
    prover.commit_u64(hk, (1ull << 33) + 5, fe(3));
    prover.commit_u64(hk, 7, fe(9));
    RangeProof proof = prove_with_transcript(hk, n, prover, merlin::Transcript("hashed"));
    report("hashed key: transcript proof of v = 2^33 + 5, out of range for n = 32 == Err", !(verify_with_transcript(proof, hk, n, prover.commitment_vec, merlin::Transcript("hashed")) == Result::Ok()));
    PublicKey hk64 = PublicKey::new_hashed(64 * m, "host_mirror_test");
    RangeProver p64 = RangeProver::new_();
    p64.commit_u64(hk64, (1ull << 31) + 5, fe(3));
    p64.commit_u64(hk64, 7, fe(9));
    RangeProof proof64 = prove_with_transcript(hk64, 64, p64, merlin::Transcript("hashed"));
    report("hashed key: n = 64 transcript proof of v >= 2^31 == Ok(())", verify_with_transcript(proof64, hk64, 64, p64.commitment_vec, merlin::Transcript("hashed")) == Result::Ok());
    report("hashed key: fixed-constant prove/verify == Ok(())", RangeProof::prove(hk64, 64, p64).verify(hk64, 64, p64.commitment_vec) == Result::Ok());
    // the same key with direct tables (bpp_pk_precompute): a single proof / verification takes them too; bytes unchanged
    RangeProof before = RangeProof::prove(hk64, 64, p64);
    const int c = hk64.precompute((size_t)1 << 30);
    report("hashed key: direct tables built (8 <= c <= 16)", c >= 8 && c <= 16);
    RangeProof after = RangeProof::prove(hk64, 64, p64);
    report("hashed key: proof bytes unchanged by the tables", after.to_bytes() == before.to_bytes());
    report("hashed key: verify with tables == Ok(())", after.verify(hk64, 64, p64.commitment_vec) == Result::Ok());
    report("hashed key: transcript proof verifies with tables", verify_with_transcript(proof64, hk64, 64, p64.commitment_vec, merlin::Transcript("hashed")) == Result::Ok());
  }
  printf("%s\n", failures ? "SOME FAILED" : "ALL PASSED");
  return failures ? 1 : 0;
}

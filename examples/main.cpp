// The reference's src/main.rs:6-57 against libbpp_b200.so through the C++ host mirror (include/bpp.hpp):
// n = 64, m = 2, v = {2, 5}, gamma = {3, 7}; prove, verify, assert Ok(()).
// usage: main [curve: 0 Ristretto255 (default) | 1 secp256k1 | 2 BLS12-381 G1]   — prints the proof bytes in hex.
#include <cstdio>
#include <cstdlib>

#include "bpp.hpp"

using namespace bulletproofsplus;

int main(int argc, char** argv) {
  const Curve curve = (Curve)(argc > 1 ? atoi(argv[1]) : 0);
  printf("started\n");
  Arith::init();

  const size_t n = 64, m = 2;

  // setup generators
  PublicKey pk = PublicKey::new_(n * m, curve);
  printf("created public key\n");

  // create m commitments and add to prover
  RangeProver prover = RangeProver::new_();
  prover.commit(pk, 2, PrimeFieldElem::new_(3, curve));
  prover.commit(pk, 5, PrimeFieldElem::new_(7, curve));

  // build proof
  printf("started proving...\n");
  RangeProof proof = RangeProof::prove(pk, n, prover);
  printf("finished proving.\n");

  printf("proof=");
  for (uint8_t b : proof.to_bytes()) printf("%02x", b);
  printf("\n");

  RangeVerifier verifier = RangeVerifier::new_();
  verifier.allocate(prover.commitment_vec);

  // verify proof
  printf("started verifying...\n");
  Result result = proof.verify(pk, n, verifier);
  printf("finished verifying.\n");

  if (!(result == Result::Ok())) {
    printf("assertion failed: result == Ok(()): %s\n", to_string(*result.error));
    return 1;
  }
  printf("result=Ok(())\n");
  return 0;
}

"""bulletproofsplus_b200 — B200-native (sm_100a) Bulletproofs+ hot path behind the reference crate's API.

The product is libbpp_b200.so (CUDA, C ABI in include/bpp.h); this package is the Python-side mirror of the
reference's public interface. Importing `api` loads the shared library and fails loudly if it is missing —
there is no CPU fallback.
"""
from ._lib import RISTRETTO255, SECP256K1, BLS12_381_G1, build  # noqa: F401


def __getattr__(name):  # lazy: `import bulletproofsplus_b200` works before the library is built (build() needs it)
    if name in ("api", "Arith", "PublicKey", "RangeProver", "RangeProof", "WeightedInnerProductProof", "MulVec", "Points",
                "VerificationError", "FormatError", "ProofError", "make_params", "prove_batch", "verify_batch"):
        import importlib
        api = importlib.import_module(".api", __name__)
        return api if name == "api" else getattr(api, name)
    raise AttributeError(name)

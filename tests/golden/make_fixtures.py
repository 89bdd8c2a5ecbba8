"""Generates tests/golden/ristretto_proofs.json from the CPU oracle (oracle/liboracle.so, the C++ restatement of
the reference; the reference itself cannot be built or imported here — no Rust toolchain, mcl_rust un-vendored).
Run from the repo root: python tests/golden/make_fixtures.py"""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import exponent_checker as ec  # noqa: E402
from oracle import oracle_lib  # noqa: E402

R = 0
ELL = ec.ORDERS[0]


def main():
    o = oracle_lib.get()
    out = {}
    proof = o.range_prove(R, 64, 2, [2, 5], [3, 7])
    out["main_rs"] = dict(n=64, m=2, v=[2, 5], gamma=[3, 7], proof=proof.hex(), V=o.commit(R, [2, 5], [3, 7]).hex(),
                          dlog_digest=ec.main_rs_kat(R).kat_digest())
    rng = random.Random(0xB200)
    for name, (n, m) in (("n64_m16", (64, 16)), ("n32_m16", (32, 16)), ("n64_m1", (64, 1))):
        vs = [rng.randrange(0, 2**31) for _ in range(m)]
        gs = [rng.randrange(ELL) for _ in range(m)]
        proof = o.range_prove(R, n, m, vs, gs)
        V = o.commit(R, vs, gs)
        assert o.range_verify(R, n, m, V, proof) == 0
        out[name] = dict(n=n, m=m, v=vs, gamma=[hex(g) for g in gs], proof=proof.hex(), V=V.hex())
    g, h, G, H = o.pk_new(R, 8)
    out["pk8"] = dict(g=g.hex(), h=h.hex(), G=G.hex(), H=H.hex())
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ristretto_proofs.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(out), "fixtures")


if __name__ == "__main__":
    main()

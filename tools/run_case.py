"""One workload of the hot path, timed with the wall clock around a stream sync (development aid; bench.py is the
measurement of record).  usage: run_case.py verify|prove|msm|commit SIZE [reps] [curve]   env: BPP_TABLE_GB (direct tables), BPP_NM="n,m" (default 64,16)"""
import ctypes as C, os, sys, time, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bulletproofsplus_b200 import api, _lib

case, size = sys.argv[1], int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
curve = int(sys.argv[4]) if len(sys.argv) > 4 else 0  # msm only
ctx = api.Context(0)
lib = ctx.lib
ELL = api.ORDERS[0]
rng = random.Random(7)
n, m = (int(x) for x in os.environ.get("BPP_NM", "64,16").split(","))  # aggregation shape of the range-proof cases

def timeit(f, label):
    f(); ctx.sync()
    f(); ctx.sync()
    ctx.hot_kernel_stats(reset=True)
    t0 = time.perf_counter()
    for _ in range(reps):
        f()
    ctx.sync()
    dt = (time.perf_counter() - t0) / reps
    hot_ms, hot_n = ctx.hot_kernel_stats(reset=True)
    print(f"{label}: {dt*1e3:.3f} ms  ({size/dt:.1f} units/s)  accumulate {hot_ms/max(1,hot_n):.3f} ms x{hot_n//reps}", flush=True)

if case in ("verify", "prove", "verify_fs", "prove_fs", "verify_rlc"):
    pk = api.PublicKey.new(n * m, 0, ctx)
    if float(os.environ.get('BPP_TABLE_GB', '0')) > 0:  # direct fixed-base tables (bpp_pk_precompute)
        print('direct tables: c =', pk.precompute(int(float(os.environ['BPP_TABLE_GB']) * 2**30)), flush=True)
    vs = [rng.randrange(2**31) for _ in range(size * m)]
    gs = [rng.randrange(ELL) for _ in range(size * m)]
    if case == "prove":
        timeit(lambda: api.prove_batch(pk, n, m, vs, gs), f"prove_batch host-ABI B={size}")
    elif case == "prove_fs":  # interactive prover, challenges = the constants (no transcript cost): the GPU side of real Fiat-Shamir
        def run():
            ses = api.ProveSession(pk, n, m, vs, gs, api.REFERENCE_NONCES)
            ch = [(12, 23)] * size
            for _ in range(ses.rounds + 1):
                ses.step(ch)
                ch = [7] * size
            return ses.finish([99] * size)
        assert run() == api.prove_batch(pk, n, m, vs, gs)
        timeit(run, f"prove session (12 host round trips) B={size}")
    elif case == "verify_rlc":
        proofs = api.prove_batch(pk, n, m, vs, gs)
        Vs = pk.commit_batch(vs, gs)
        Vl = [Vs[i * m:(i + 1) * m] for i in range(size)]
        ws = [rng.getrandbits(128) | 1 for _ in range(size)]
        assert api.verify_batch_rlc(pk, n, m, proofs, Vl, weights=ws) == ([0] * size, False)
        timeit(lambda: api.verify_batch_rlc(pk, n, m, proofs, Vl, weights=ws), f"verify_batch_rlc host-ABI B={size}")
    elif case == "verify_fs":
        proofs = api.prove_batch(pk, n, m, vs, gs)
        Vs = pk.commit_batch(vs, gs)
        Vl = [Vs[i * m:(i + 1) * m] for i in range(size)]
        chal = [[12, 23, 99] + [7] * 10] * size
        assert api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=chal) == [0] * size
        timeit(lambda: api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=chal), f"verify_batch_fs host-ABI B={size}")
    else:
        proofs = api.prove_batch(pk, n, m, vs, gs)
        Vs = pk.commit_batch(vs, gs)
        Vl = [Vs[i * m:(i + 1) * m] for i in range(size)]
        out = api.verify_batch(pk, n, m, proofs, Vl)
        assert out == [0] * size, out[:8]
        timeit(lambda: api.verify_batch(pk, n, m, proofs, Vl), f"verify_batch host-ABI B={size}")
elif case == "msm":
    N = 1 << size
    order = api.ORDERS[curve]
    mult = b"".join(rng.randrange(order).to_bytes(32, "little") for _ in range(N))
    pts = api.Points.from_multipliers(mult, curve, ctx)
    if os.environ.get('BPP_TABLES'):
        pts.precompute()
    sc = b"".join(rng.randrange(order).to_bytes(32, "little") for _ in range(N))
    size = N
    timeit(lambda: pts.msm(sc), f"msm N=2^{sys.argv[2]}")
elif case == "commit":
    pk = api.PublicKey.new(64, 0, ctx)
    vs = [rng.randrange(2**31) for _ in range(size)]
    gs = [rng.randrange(ELL) for _ in range(size)]
    timeit(lambda: pk.commit_batch(vs, gs), f"commit_batch {size}")
print("launches", ctx.kernel_launches())

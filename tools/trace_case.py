"""Timeline of one call over the library's two streams (bpp_trace / bpp_trace_dump): development aid.
usage: trace_case.py verify|verify_fs|prove B   |   trace_case.py msm LOG2N [tables|generic] [curve]"""
import os, sys, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bulletproofsplus_b200 import api

case, size = sys.argv[1], int(sys.argv[2])
ctx = api.Context(0)
rng = random.Random(7)
n, m = (int(x) for x in os.environ.get("BPP_NM", "64,16").split(","))  # aggregation shape; default the headline config
ELL = api.ORDERS[0]
if case == "msm":
    curve = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    N = 1 << size
    order = api.ORDERS[curve]
    mult = b"".join(rng.randrange(order).to_bytes(32, "little") for _ in range(N))
    pts = api.Points.from_multipliers(mult, curve, ctx)
    if len(sys.argv) > 3 and sys.argv[3] == "tables":
        pts.precompute()
    sc = b"".join(rng.randrange(order).to_bytes(32, "little") for _ in range(N))
    run = lambda: pts.msm(sc)
else:
    pk = api.PublicKey.new(n * m, 0, ctx)
    if float(os.environ.get('BPP_TABLE_GB', '0')) > 0:  # direct fixed-base tables (bpp_pk_precompute)
        print('# direct tables: c =', pk.precompute(int(float(os.environ['BPP_TABLE_GB']) * 2**30)))
    vs = [rng.randrange(2**31) for _ in range(size * m)]
    gs = [rng.randrange(ELL) for _ in range(size * m)]
    if case == "prove":
        run = lambda: api.prove_batch(pk, n, m, vs, gs)
    else:
        proofs = []
        for c0 in range(0, size, 256):
            c1 = min(size, c0 + 256)
            proofs += api.prove_batch(pk, n, m, vs[c0 * m:c1 * m], gs[c0 * m:c1 * m])
        Vs = pk.commit_batch(vs, gs)
        Vl = [Vs[i * m:(i + 1) * m] for i in range(size)]
        if case == "verify_fs":
            chal = [[12, 23, 99] + [7] * 10] * size
            run = lambda: api.verify_batch_fs(pk, n, m, proofs, Vl, challenges=chal)
        else:
            run = lambda: api.verify_batch(pk, n, m, proofs, Vl)
for _ in range(3):
    run()
ctx.sync()
ctx.trace(True)
run()
ctx.trace(False)
rows = ctx.trace_dump()
t_end = max(r[2] for r in rows)
print(f"# {' '.join(sys.argv[1:])}: {len(rows)} launches, span {t_end:.1f} us (stream 0 = caller's, 1 = auxiliary)")
for s_, a, b, name in rows:
    print(f"{s_} {a:9.1f} {b:9.1f} {b - a:8.1f}  {name}")

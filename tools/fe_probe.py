"""Field-multiply throughput against the IMAD.WIDE peak, for several ILP / occupancy points (development aid)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bulletproofsplus_b200 import api
ctx = api.Context(0)
pr = ctx.int_peak_probe()
peak = pr["imad32_per_s"]
print(f"IMAD peak {peak/1e12:.2f} T/s, carry-chained wide {pr['wide_chain_per_s']/1e12:.2f} T/s, fused wide {pr['wide_fused_per_s']/1e12:.2f} T/s")
for ilp in (1, 2, 4):
    for tpb, bps in ((128, 4), (128, 8), (256, 4), (128, 3), (128, 2)):
        r = C.c_double()
        ctx.check(ctx.lib.bpp_fe_mul_probe(ctx.h, ilp, tpb, bps, C.byref(r)), "probe")
        print(f"ilp={ilp} threads/SM={tpb*bps:5d}: {r.value/1e9:8.1f} G mul/s = {72*r.value/peak*100:5.1f}% of IMAD peak")

"""bpp_points_sum_raw_noalloc_dev (no counterpart in the reference): the
allocation-free variant must give the bytes of bpp_points_sum_raw_dev — bit-exact, this is integer work — on two calls in a
row (the static device scratch keeps the previous call's contents), on all three curves, and must reject sizes beyond its
stated maximum. Collected last on purpose (pytest -x). Passes alone in a fresh process and after the other tests."""
import ctypes as C
import random

import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from bulletproofsplus_b200 import api
    return torch, api, api.Context(0)


def _raw_partials(torch, api, ctx, curve, count, seed):
    """count raw partial points on the device: one 8-term MSM each (bpp_msm_prepared_raw_dev), as a point-range split produces."""
    lib = ctx.lib
    rng = random.Random(seed)
    q = api.ORDERS[curve]
    k = 8
    mult = b"".join(rng.randrange(1, q).to_bytes(32, "little") for _ in range(k))
    pts = api.Points.from_multipliers(mult, curve, ctx)
    raw_bytes = lib.bpp_raw_point_bytes(curve)
    d_all = torch.empty(count * raw_bytes, dtype=torch.uint8, device="cuda:0")
    for j in range(count):
        sc = b"".join(rng.randrange(q).to_bytes(32, "little") for _ in range(k))
        d_sc = torch.frombuffer(bytearray(sc), dtype=torch.uint8).to("cuda:0")
        torch.cuda.synchronize()  # torch's stream is not the context's
        ctx.check(lib.bpp_msm_prepared_raw_dev(ctx.h, pts.h, d_sc.data_ptr(), k, d_all.data_ptr() + j * raw_bytes), "msm_raw")
        ctx.sync()
    return d_all, pts


@pytest.mark.parametrize("curve", [0, 1, 2])
def test_noalloc_matches_existing_entry_twice(env, curve):
    torch, api, ctx = env
    lib = ctx.lib
    pb = lib.bpp_point_bytes(curve)
    raw_bytes = lib.bpp_raw_point_bytes(curve)
    # the second input is larger than one phase-1 block (1024 points) and differs from the first: stale scratch would show
    d_a, keep_a = _raw_partials(torch, api, ctx, curve, 5, 100 + curve)
    d_b, keep_b = _raw_partials(torch, api, ctx, curve, 37, 200 + curve)
    reps = 60  # 37 * 60 = 2220 points: three phase-1 blocks
    d_b = d_b.repeat(reps)
    for d_in, count in ((d_a, 5), (d_b, 37 * reps), (d_a, 5)):
        want = torch.zeros(pb, dtype=torch.uint8, device="cuda:0")
        got = torch.full((pb,), 0xEE, dtype=torch.uint8, device="cuda:0")
        torch.cuda.synchronize()
        ctx.check(lib.bpp_points_sum_raw_dev(ctx.h, curve, d_in.data_ptr(), count, want.data_ptr()), "sum_raw_dev")
        ctx.check(lib.bpp_points_sum_raw_noalloc_dev(ctx.h, curve, d_in.data_ptr(), count, got.data_ptr()), "sum_raw_noalloc_dev")
        ctx.sync()
        assert bytes(got.cpu().numpy().tobytes()) == bytes(want.cpu().numpy().tobytes()), (curve, count)
    assert d_in.numel() == 5 * raw_bytes


def test_noalloc_rejects_sizes_beyond_the_maximum(env):
    torch, api, ctx = env
    lib = ctx.lib
    d = torch.zeros(256, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    rc = lib.bpp_points_sum_raw_noalloc_dev(ctx.h, 0, d.data_ptr(), 65537, d.data_ptr())
    assert rc == 3  # BPP_ERR_LENGTH
    assert b"BPP_NOALLOC_MAX_POINTS" in lib.bpp_last_error(ctx.h)
    assert lib.bpp_points_sum_raw_noalloc_dev(ctx.h, 0, d.data_ptr(), 0, d.data_ptr()) == 3
    assert lib.bpp_points_sum_raw_noalloc_dev(ctx.h, 9, d.data_ptr(), 1, d.data_ptr()) == 8  # BPP_ERR_UNSUPPORTED

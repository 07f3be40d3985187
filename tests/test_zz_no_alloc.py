"""dk_bh_rank_slice_noalloc — the allocation-free variant of dk_bh_rank_slice (fixed scratch in a static device array;
no counterpart in the reference). Collected last on purpose (
a failure here must not hide the parity tests). Checked through the C ABI against the existing
entry point, two calls in a row (the scratch keeps the previous call's contents), alone or after the other tests."""
import ctypes as C

import numpy as np
import pytest

import dekupl_run_b200 as dk
import oracle_lib as orc

pytestmark = pytest.mark.gpu
NOALLOC_MAX = 1 << 18        # DK_BH_NOALLOC_MAX in include/dekupl_b200.h


def _rank(ctx, fn, tp, n_below, R):
    import torch
    sp, F = torch.empty_like(tp), torch.empty_like(tp)
    ctx._chk(fn(ctx.h, tp.data_ptr(), tp.numel(), int(n_below), int(R), sp.data_ptr(), F.data_ptr()))
    ctx.sync()
    return sp.cpu().numpy(), F.cpu().numpy()


def test_bh_rank_slice_noalloc_equals_the_allocating_entry_point():
    import torch
    ctx = dk.Context(0)                                  # a context of its own: passes alone and after the other tests
    try:
        dev = torch.device("cuda", 0)
        rng = np.random.default_rng(2026)
        # call 1: a large slice with heavy ties; call 2: a shorter one with other values (stale scratch behind it)
        p1 = np.round(rng.random(200_000) ** 2 * 0.05, 5)
        p1[:300] = rng.random(300) * 1e-9
        p2 = rng.random(4097) * 0.01
        for p, n_below, R in ((p1, 0, 5_000_000), (p2, 12_345, 900_000), (p1[:70_001], 7, 5_000_000)):
            tp = torch.from_numpy(p).to(dev)
            sp_a, F_a = _rank(ctx, ctx.L.dk_bh_rank_slice, tp, n_below, R)
            sp_n, F_n = _rank(ctx, ctx.L.dk_bh_rank_slice_noalloc, tp, n_below, R)
            np.testing.assert_array_equal(sp_n, sp_a)                      # the same sorted bit patterns
            np.testing.assert_allclose(F_n, F_a, rtol=1e-15, atol=0)       # the tolerance of the BH tests
            assert np.array_equal(sp_n, np.sort(p))
        # against the oracle as well: a whole list ranked in one slice is the reference's adjustment (TtestFilter.c:459-476)
        tp = torch.from_numpy(p2).to(dev)
        sp_n, F_n = _rank(ctx, ctx.L.dk_bh_rank_slice_noalloc, tp, 0, len(p2))
        got = np.minimum(1.0, F_n[np.searchsorted(sp_n, p2, side="left")])
        np.testing.assert_allclose(got, orc.bh(p2), rtol=1e-12)
        # n == 0 is a no-op; a slice beyond the fixed scratch is refused loudly, nothing is launched
        assert ctx.L.dk_bh_rank_slice_noalloc(ctx.h, None, 0, 0, 1, None, None) == 0
        big = torch.zeros(NOALLOC_MAX + 1, dtype=torch.float64, device=dev)
        out = torch.empty_like(big)
        l0 = ctx.launches
        rc = ctx.L.dk_bh_rank_slice_noalloc(ctx.h, big.data_ptr(), big.numel(), 0, big.numel(), out.data_ptr(), out.data_ptr())
        assert rc == -2 and b"DK_BH_NOALLOC_MAX" in ctx.L.dk_last_error(ctx.h) and ctx.launches == l0
        # exactly the maximum is accepted
        full = torch.from_numpy(rng.random(NOALLOC_MAX) * 0.05).to(dev)
        sp_a, F_a = _rank(ctx, ctx.L.dk_bh_rank_slice, full, 0, 10 * NOALLOC_MAX)
        sp_n, F_n = _rank(ctx, ctx.L.dk_bh_rank_slice_noalloc, full, 0, 10 * NOALLOC_MAX)
        np.testing.assert_array_equal(sp_n, sp_a)
        np.testing.assert_allclose(F_n, F_a, rtol=1e-15, atol=0)
    finally:
        ctx.close()

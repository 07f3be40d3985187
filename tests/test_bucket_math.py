"""Host-side check of the arithmetic merge v4 relies on (dekupl-run_b200/csrc/dk_merge4.cuh, step B1): the bucket of a
k-mer inside a round, bucket(key) = ((key - base) >> wsh) * scale >> 16, must be (1) inside [0, M) for every key of the
round and (2) non-decreasing in the key, for ANY span — bucket order is what makes the kernel's rows come out in k-mer
order without a sort. Restated here with numpy integers exactly as the kernel computes it (32-bit products)."""
import numpy as np
import pytest

M = 2048          # M4_M


def bucket_params(span):
    bitlen = int(span).bit_length()
    wsh = bitlen - 16 if bitlen > 16 else 0
    fmax = (span - 1) >> wsh
    scale = 65536 if fmax + 1 <= M else (M << 16) // (fmax + 1)
    return wsh, fmax, scale


def buckets(off, wsh, scale):
    f = (off >> np.uint64(wsh)).astype(np.uint64)
    prod = f * np.uint64(scale)
    assert (prod < (1 << 32)).all(), "the kernel multiplies in 32 bits"
    return (prod >> np.uint64(16)).astype(np.int64)


@pytest.mark.parametrize("seed", range(6))
def test_bucket_is_in_range_and_monotone(seed):
    rng = np.random.default_rng(seed)
    spans = [1, 2, 3, M - 1, M, M + 1, 65535, 65536, 65537, (1 << 32) - 1, 1 << 32, (1 << 48) + 12345, (1 << 62) - 1, (1 << 64) - 2]
    spans += [int(rng.integers(1, 1 << int(b))) for b in rng.integers(1, 63, 40)]
    for span in spans:
        wsh, fmax, scale = bucket_params(span)
        assert fmax < 65536
        n = 4000
        off = np.unique(np.concatenate([rng.integers(0, span, n, dtype=np.uint64), np.array([0, span - 1], dtype=np.uint64),
                                        np.arange(min(span, 300), dtype=np.uint64)]))
        b = buckets(off, wsh, scale)
        assert b.min() >= 0 and b.max() < M, (span, int(b.max()))
        assert (np.diff(b) >= 0).all(), span                      # sorted keys -> non-decreasing buckets
        if fmax + 1 <= M and wsh == 0:                            # narrow rounds: one k-mer per bucket, never a collision
            assert len(np.unique(b)) == len(off)


def test_row_limits_of_the_kernel_geometry():
    """row capacity of a round (recurrence byte matrix aliased on the owner tables) for the sample counts v4 accepts"""
    J, alias = 4, 3 * M * 2
    for S in range(1, 65):
        abw = ((S + 3) // 4) | 1
        rmax = alias // (4 * abw)
        assert abw % 2 == 1 and 4 * abw >= S and rmax >= 64, S   # odd pitch, one byte per sample, useful capacity
        RS = S | 1
        BR = min(512, (alias - 32) // (4 * RS + 8))
        assert BR >= 4 and ((BR * RS * 4 + 15) // 16) * 16 + BR * 8 <= alias

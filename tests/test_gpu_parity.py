"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against the CPU
oracle on the same seeded inputs and against the committed golden vectors. Integer work is bit-exact;
fp64 results are within 1e-9 relative (BASELINE.json north_star)."""
import os

import numpy as np
import pytest

import dekupl_run_b200 as dk
import oracle_lib as orc
from common import GOLD, format_join, load_json, parse_dump_text, random_dumps

pytestmark = pytest.mark.gpu
RTOL = 1e-9   # north_star: NF, p, adjusted p within 1e-9 relative in fp64


def _join_both(ctx, keys, cnts, r, a, stale=False):
    m = ctx.join_counts(keys, cnts, r, a, emit_stale_tail=stale)
    gk, gc = m.export()
    U, R = m.union, m.rows
    m.free()
    ok, oc, oU = orc.join(keys, cnts, r, a, stale_tail=stale)
    assert U == oU and R == len(ok)
    np.testing.assert_array_equal(gk, ok)
    np.testing.assert_array_equal(gc, oc)
    return gk, gc


# ---------------------------------------------------------------- joinCounts ---
def test_join_reference_binary_goldens(ctx):
    for case in load_json("join_cases.json"):
        r, a = 1, 1
        args = case["args"]
        for i in range(0, len(args), 2):
            r = int(args[i + 1]) if args[i] == "-r" else r
            a = int(args[i + 1]) if args[i] == "-a" else a
        dumps = [parse_dump_text(t) for t in case["inputs"]]
        m = ctx.join_counts([d[0] for d in dumps], [d[1] for d in dumps], r, a, emit_stale_tail=True)
        keys, cnts = m.export()
        m.free()
        assert format_join(keys, cnts, dumps[0][2]) == case["stdout"], f"args={args}"


@pytest.mark.parametrize("S,n_union,pp,core,r,a", [
    (2, 1000, 0.5, 0.0, 1, 1), (4, 20000, 0.4, 0.2, 2, 5), (16, 200000, 0.075, 0.2, 3, 5),
    (17, 30000, 0.3, 0.1, 2, 3), (32, 50000, 0.2, 0.2, 3, 2), (33, 40000, 0.2, 0.2, 4, 2),
    (48, 60000, 0.1, 0.2, 4, 5), (64, 30000, 0.1, 0.1, 5, 3), (100, 20000, 0.1, 0.2, 8, 4),
    (200, 12000, 0.08, 0.2, 12, 5), (256, 6000, 0.1, 0.1, 10, 3), (1, 5000, 1.0, 0.0, 1, 3),
    # many surviving rows per round: output crosses matrix chunks inside one round (S=4 / S=1 once hit a
    # reciprocal-division overflow there, found by the config-1 real-data fixture)
    (4, 300000, 0.6, 0.3, 1, 1), (1, 200000, 1.0, 0.0, 1, 1), (2, 250000, 0.7, 0.2, 1, 2), (3, 150000, 0.6, 0.2, 1, 1),
    (8, 200000, 0.5, 0.3, 1, 1), (12, 150000, 0.3, 0.3, 2, 2),
])
def test_join_random_vs_oracle(ctx, S, n_union, pp, core, r, a):
    rng = np.random.default_rng(S * 1000 + n_union)
    keys, cnts = random_dumps(rng, S, n_union, 62, pp, 12, p_core=core)
    _join_both(ctx, keys, cnts, r, a)


def test_join_edge_cases(ctx):
    rng = np.random.default_rng(5)
    keys, cnts = random_dumps(rng, 6, 3000, 62, 0.5, 9)
    # empty samples (the reference segfaults on an empty file; packed API defines them as absent)
    keys[2], cnts[2] = keys[2][:0], cnts[2][:0]
    keys[5], cnts[5] = keys[5][:0], cnts[5][:0]
    _join_both(ctx, keys, cnts, 2, 3)
    # all samples empty
    m = ctx.join_counts([k[:0] for k in keys], [c[:0] for c in cnts], 1, 1)
    assert m.rows == 0 and m.union == 0
    m.free()
    # single record
    _join_both(ctx, [np.array([7], np.uint64), np.array([], np.uint64)], [np.array([3], np.int32), np.array([], np.int32)], 1, 1)
    # r <= 0 reproduces the stale trailing row; a <= 0 never counts absent samples
    _join_both(ctx, keys, cnts, 0, 1, stale=True)
    _join_both(ctx, keys, cnts, 1, 0)
    _join_both(ctx, keys, cnts, 7, 1)          # nothing can pass
    # duplicates inside one dump are emitted as separate rows (reference behaviour)
    _join_both(ctx, [np.array([4, 4, 8, 8, 8], np.uint64), np.array([4, 8], np.uint64)],
               [np.array([1, 2, 3, 4, 5], np.int32), np.array([9, 7], np.int32)], 1, 1)
    # identical dumps, maximal 31-mer keys, clustered keys (one hot bucket)
    k = np.sort(rng.choice(1 << 20, 5000, replace=False).astype(np.uint64)) + np.uint64((1 << 62) - (1 << 20))
    _join_both(ctx, [k, k, k], [np.full(5000, 5, np.int32)] * 3, 3, 5)
    hot = np.unique(np.concatenate([rng.integers(0, 1 << 62, 100, dtype=np.uint64),
                                    np.uint64(123456789) + rng.integers(0, 40000, 30000).astype(np.uint64)]))
    _join_both(ctx, [hot[::2], hot[::3], hot], [np.ones(len(hot[::2]), np.int32) * 6, np.ones(len(hot[::3]), np.int32),
                                                np.ones(len(hot), np.int32) * 9], 2, 5)
    # more duplicates of one key inside a dump than one staging window holds (merge v2 fallback path)
    dup = np.sort(np.concatenate([np.full(700, 5000, np.uint64), rng.integers(0, 10000, 300).astype(np.uint64)]))
    oth = np.unique(rng.integers(0, 10000, 2000).astype(np.uint64))
    _join_both(ctx, [dup, oth, oth[::2]], [np.arange(len(dup), dtype=np.int32) % 9, np.full(len(oth), 6, np.int32),
                                           np.full(len(oth[::2]), 2, np.int32)], 1, 3)
    # large counts up to INT_MAX
    _join_both(ctx, [np.array([1, 2], np.uint64)] * 2, [np.array([2 ** 31 - 1, 5], np.int32), np.array([7, 2 ** 31 - 1], np.int32)], 2, 6)


def _clustered_keys(rng, n, levels):
    """k-mers in nested clusters: every level packs a share of the keys into a 2^-12 narrower window, down to
    runs of consecutive integers (what low-complexity sequence does to a sorted dump)"""
    parts, width, lo = [], 1 << 62, 0
    for _ in range(levels):
        parts.append(lo + rng.integers(0, width, n // levels, dtype=np.uint64))
        lo = lo + int(rng.integers(0, max(1, width - (width >> 12))))
        width = max(width >> 12, 64)
    parts.append(np.uint64(lo) + np.arange(min(n, 3000), dtype=np.uint64))     # a run of consecutive k-mers
    return np.unique(np.concatenate([p.astype(np.uint64) for p in parts]))


@pytest.mark.parametrize("S,n,levels,pp,r,a", [(16, 120000, 5, 0.6, 3, 5), (3, 60000, 4, 0.9, 1, 1), (48, 40000, 3, 0.3, 4, 3),
                                              (64, 20000, 5, 0.5, 1, 1), (5, 200000, 2, 0.8, 2, 2), (1, 50000, 5, 1.0, 1, 1)])
def test_join_bucket_kernel_on_clustered_kmers(ctx, S, n, levels, pp, r, a):
    """merge v4 (monotone buckets + owner election): clusters force its round-shrinking paths (more distinct k-mers
    in a bucket than owner levels, more rows in a round than the recurrence matrix holds)"""
    rng = np.random.default_rng(S * 7 + levels)
    allk = _clustered_keys(rng, n, levels)
    keys, cnts = [], []
    for _ in range(S):
        m = rng.random(len(allk)) < pp
        keys.append(allk[m])
        cnts.append(rng.integers(1, 12, size=int(m.sum())).astype(np.int32))
    _join_both(ctx, keys, cnts, r, a)


def test_join_probe_hands_sparse_data_to_the_merge_tree(ctx):
    """enough partitions for the kernel-choice probe (dk_join_impl): ~2 records per k-mer, so the first partitions are
    merged by v4 and the rest by v3 into the same matrix; the result must not show the seam"""
    rng = np.random.default_rng(77)
    keys, cnts = random_dumps(rng, 16, 6_000_000, 62, 0.12, 12)
    _join_both(ctx, keys, cnts, 2, 3)
    _join_both(ctx, keys, cnts, 1, 1)


@pytest.mark.parametrize("which", ["3", "1"])
def test_join_older_merge_kernels_stay_exact(ctx, which, monkeypatch):
    """v3 is the path for S > 64 and for dumps holding copies of a k-mer, v1 the one behind it: keep both pinned"""
    monkeypatch.setenv("DK_MERGE", which)
    rng = np.random.default_rng(31)
    keys, cnts = random_dumps(rng, 16, 150000, 62, 0.075, 12, p_core=0.2)
    _join_both(ctx, keys, cnts, 3, 5)
    keys, cnts = random_dumps(rng, 5, 80000, 62, 0.6, 12, p_core=0.3)
    _join_both(ctx, keys, cnts, 1, 1)


def test_join_large_checksum_properties(ctx):
    """size-independent properties at a size the oracle would take minutes for"""
    rng = np.random.default_rng(99)
    keys, cnts = random_dumps(rng, 16, 3_000_000, 62, 0.075, 40, p_core=0.2)
    m_all = ctx.join_counts(keys, cnts, 1, 1)
    gk, gc = m_all.export()
    assert (np.diff(gk.astype(np.int64)) > 0).all()                       # sorted, distinct
    assert m_all.union == len(np.unique(np.concatenate(keys))) == m_all.rows
    for s in range(16):                                                   # every record lands in its column
        assert int(gc[:, s].astype(np.int64).sum()) == int(cnts[s].astype(np.int64).sum())
        assert int((gc[:, s] != 0).sum()) == len(keys[s])
    x, sm = m_all.checksum()
    assert x == int(np.bitwise_xor.reduce(gk)) and sm == int(gc.astype(np.int64).sum())
    m_f = ctx.join_counts(keys, cnts, 3, 5)
    fk, fc = m_f.export()
    keep = (gc >= 5).sum(axis=1) >= 3                                     # filter == row-wise predicate on the full matrix
    np.testing.assert_array_equal(fk, gk[keep])
    np.testing.assert_array_equal(fc, gc[keep])
    m_all.free(); m_f.free()


# ----------------------------------------------------------------- computeNF ---
def _nf_text(names, nf):
    return "sample\tnormalization_factor\n" + "".join("%s\t%f\n" % (n, v) for n, v in zip(names, nf))


def test_nf_reference_fixture_goldens(ctx):
    g = load_json("nf_cases.json")
    z = np.load(os.path.join(GOLD, "nf_fixture.npz"))
    m = ctx.matrix_from_host(z["counts"])
    nf, info = ctx.compute_nf(m, dk.NF_REFERENCE)
    m.free()
    assert _nf_text(z["names"], nf) == g["computeNF_fixture"]          # bytes of the real computeNF binary
    onf, _, _ = orc.nf_reference(z["counts"])
    np.testing.assert_allclose(nf, onf, rtol=RTOL)
    for nm in ("bigger", "small"):
        mm = g[f"ttest_{nm}_matrix"]
        m = ctx.matrix_from_host(np.array(mm["counts"], np.int32))
        nf, _ = ctx.compute_nf(m, dk.NF_REFERENCE)
        m.free()
        assert _nf_text(mm["names"], nf) == g[f"ttest_{nm}_fixture"]


@pytest.mark.parametrize("idx", [0, 1, 2, 3])
def test_nf_reference_step_logic_matches_binary(ctx, idx):
    from golden.make_golden import synth_matrix
    case = load_json("nf_cases.json")["nf_step_cases"][idx]
    cnts = synth_matrix(case["seed"], case["R"], case["S"])
    m = ctx.matrix_from_host(cnts)
    nf, info = ctx.compute_nf(m, dk.NF_REFERENCE)
    m.free()
    assert not info["undefined_zone"]
    assert _nf_text([f"s{j}" for j in range(case["S"])], nf) == case["stdout"]


@pytest.mark.parametrize("R,S,step", [(5000, 4, 1000), (20000, 16, 3000), (2500, 48, 1000), (1500, 7, 1000), (999, 3, 1000),
                                      (3000, 200, 1000)])
def test_nf_reference_small_steps_vs_oracle(ctx, R, S, step):
    rng = np.random.default_rng(R + S)
    cnts = rng.poisson(rng.uniform(2, 40, size=(R, 1)) * rng.uniform(0.5, 2, size=(1, S))).astype(np.int32)
    onf, oused, ouz = orc.nf_reference(cnts, step)
    m = ctx.matrix_from_host(cnts)
    nf, info = ctx.compute_nf(m, dk.NF_REFERENCE, step)
    m.free()
    assert info["undefined_zone"] == ouz
    np.testing.assert_allclose(nf, onf, rtol=RTOL)
    if not ouz and oused < R:
        assert info["consumed"] == oused


def test_nf_full_vs_oracle_and_after_join(ctx):
    rng = np.random.default_rng(12)
    keys, cnts = random_dumps(rng, 8, 60000, 62, 0.3, 50, p_core=0.5)
    m = ctx.join_counts(keys, cnts, 2, 2)        # chunked (non-identity) matrix layout
    _, gc = m.export()
    nf, info = ctx.compute_nf(m, dk.NF_FULL)
    nf_r, _ = ctx.compute_nf(m, dk.NF_REFERENCE, 5000)
    m.free()
    np.testing.assert_allclose(nf, orc.nf_full(gc), rtol=RTOL)
    np.testing.assert_allclose(nf_r, orc.nf_reference(gc, 5000)[0], rtol=RTOL)
    assert info["n_pos"] == int((gc > 0).all(axis=1).sum())


def test_nf_no_positive_row_fails_loudly(ctx):
    m = ctx.matrix_from_host(np.array([[0, 1], [2, 0]], np.int32))
    with pytest.raises(dk.DkError):
        ctx.compute_nf(m, dk.NF_REFERENCE)
    m.free()


# ---------------------------------------------------------------- TtestFilter ---
def _ttest_both(ctx, cnts, nf, idxA, idxB, pmax, fmin, p_cut=None):
    m = ctx.matrix_from_host(cnts)
    t = ctx.ttest_rows(m, nf, idxA, idxB)
    t.adjust(pmax if p_cut is None else p_cut)
    D = t.filter(pmax, fmin)
    p, fc, adj = t.export_rows()
    surv = t.export_survivors()
    ties = t.ties
    t.free(); m.free()
    op, omA, omB, ofc = orc.ttest_rows(cnts, nf, idxA, idxB)
    oadj = orc.bh(op)
    np.testing.assert_allclose(p, op, rtol=RTOL, atol=1e-300)
    np.testing.assert_allclose(fc, ofc, rtol=RTOL, atol=1e-12)
    keep = (oadj <= pmax) & (np.abs(ofc) >= fmin)
    # rows within 1e-9 of a threshold are "ties" (listed, not asserted)
    tie = (np.abs(oadj - pmax) <= 1e-9 * pmax) | ((fmin != 0) & (np.abs(np.abs(ofc) - fmin) <= 1e-9 * abs(fmin)))
    rows = surv["row"].astype(np.int64)
    want = np.nonzero(keep & ~tie)[0]
    assert set(want.tolist()) <= set(rows.tolist()) <= set(np.nonzero(keep | tie)[0].tolist())
    assert (np.diff(rows) > 0).all()                                     # reference (input) order
    np.testing.assert_allclose(surv["p_adj"], oadj[rows], rtol=RTOL, atol=1e-300)
    np.testing.assert_allclose(surv["meanA"], omA[rows], rtol=RTOL)
    np.testing.assert_allclose(surv["meanB"], omB[rows], rtol=RTOL)
    np.testing.assert_allclose(surv["norm"], cnts[rows] / nf[None, :], rtol=1e-15)
    computed = ~np.isnan(adj)
    exact = computed & (oadj <= (pmax if p_cut is None else p_cut))
    np.testing.assert_allclose(adj[exact], oadj[exact], rtol=RTOL, atol=1e-300)
    assert (adj[computed] >= oadj[computed] * (1 - RTOL)).all()          # never under-estimates
    return D, ties


@pytest.mark.parametrize("nm", ["bigger", "small"])
def test_ttest_reference_fixtures(ctx, nm):
    from test_oracle import SURVEY_ROWS
    g = load_json("nf_cases.json")
    mm = g[f"ttest_{nm}_matrix"]
    cnts = np.array(mm["counts"], np.int32)
    nf = np.array([float(l.split()[2]) for l in g["ttest_conditions"].strip().split("\n")[1:]])
    D, _ = _ttest_both(ctx, cnts, nf, [0, 1, 2], [3, 4, 5], 0.2, 0.0, p_cut=2.0)
    kept = sum(1 for row in cnts if SURVEY_ROWS[nm][tuple(row.tolist())][1] <= 0.2)
    assert D == kept


@pytest.mark.parametrize("R,nA,nB,effect,pmax,fmin", [
    (20000, 2, 2, 3.0, 0.2, 0.0), (30000, 8, 8, 4.0, 0.05, 0.0), (5000, 3, 5, 2.0, 0.05, 1.0),
    (8000, 24, 24, 1.5, 0.01, 0.5), (3000, 100, 100, 1.3, 0.05, 0.0), (64, 2, 3, 5.0, 0.5, 0.0), (1, 2, 2, 1.0, 1.0, 0.0),
])
def test_ttest_random_vs_oracle(ctx, R, nA, nB, effect, pmax, fmin):
    rng = np.random.default_rng(R + nA)
    S = nA + nB + 1                                    # one sample in neither group (NF 0 in the reference; unused)
    depth = rng.uniform(0.5, 2.0, S)
    lam = rng.gamma(0.7, 20, size=(R, 1)) * depth[None, :]
    de = rng.random(R) < 0.05
    perm = rng.permutation(S)
    idxA, idxB = perm[:nA].tolist(), perm[nA:nA + nB].tolist()
    lam[np.ix_(de, idxB)] *= effect
    cnts = rng.poisson(lam).astype(np.int32)
    cnts[rng.random(R) < 0.1] = 0                      # all-zero rows: sd1 == sd2 == 0 => p = 0.5
    cnts[rng.random(R) < 0.02] = 7                     # constant rows
    nf = np.round(depth / depth.mean(), 6)             # as read back from "%f" text
    _ttest_both(ctx, cnts, nf, idxA, idxB, pmax, fmin)
    _ttest_both(ctx, cnts, nf, idxA, idxB, pmax, fmin, p_cut=2.0)   # full BH


def test_ttest_group_of_one_is_refused(ctx):
    m = ctx.matrix_from_host(np.ones((4, 3), np.int32))
    with pytest.raises(dk.DkError):
        ctx.ttest_rows(m, np.ones(3), [0], [1, 2])
    m.free()


def test_bh_heavy_ties_and_large(ctx):
    """BH on 2M rows with massive ties (the fixtures' 93-fold tie, scaled up)"""
    rng = np.random.default_rng(21)
    R = 2_000_000
    cnts = rng.poisson(3.0, size=(R, 4)).astype(np.int32)   # tiny counts => few distinct p-values
    nf = np.array([1.0, 1.0, 1.0, 1.0])
    m = ctx.matrix_from_host(cnts)
    t = ctx.ttest_rows(m, nf, [0, 1], [2, 3])
    t.adjust(2.0)
    p, fc, adj = t.export_rows()
    t.free(); m.free()
    oadj = orc.bh(p)                                          # BH of the GPU's own raw p: isolates K5
    np.testing.assert_allclose(adj, oadj, rtol=1e-15)
    assert len(np.unique(p)) < R // 100


@pytest.mark.gpu
@pytest.mark.parametrize("S,slices", [(16, 4), (5, 7), (48, 3)])
def test_join_host_pipelined_slices_equal_single_pass(ctx, S, slices, monkeypatch):
    """dk_join_counts on host buffers cut into key-range slices (copy/merge overlap) == oracle, including
    duplicate k-mers sitting exactly on a slice edge and NF / checksum on the spliced matrix"""
    rng = np.random.default_rng(S * 31 + slices)
    keys, cnts = random_dumps(rng, S, 40_000, 62, 0.2, 40, p_core=0.3)
    big = int(np.argmax([len(k) for k in keys]))
    edge = keys[big][len(keys[big]) // slices]              # the first slice edge: make it a duplicated k-mer elsewhere
    j = int(np.searchsorted(keys[1], edge))
    keys[1] = np.insert(keys[1], j, [edge, edge]); cnts[1] = np.insert(cnts[1], j, [7, 9]).astype(np.int32)
    monkeypatch.setenv("DK_JOIN_SLICES", str(slices))
    m = ctx.join_counts(keys, cnts, 2, 3)
    monkeypatch.setenv("DK_JOIN_SLICES", "1")
    m1 = ctx.join_counts(keys, cnts, 2, 3)
    gk, gc = m.export()
    ok, oc, oU = orc.join(keys, cnts, 2, 3)
    assert m.union == oU and np.array_equal(gk, ok) and np.array_equal(gc, oc)
    assert m.checksum() == m1.checksum()
    m.free(); m1.free()


@pytest.mark.gpu
def test_bh_rank_slices_stitch_to_the_reference_adjustment(ctx):
    """dk_bh_rank_slice on three p-value ranges + cross-range running minimum == orc.bh on the whole list (heavy ties)"""
    import torch
    rng = np.random.default_rng(11)
    p = np.round(rng.random(50_000) ** 2, 3)                       # many exact ties, some zeros
    p[:200] = rng.random(200) * 1e-6
    want = orc.bh(p)
    dev = torch.device("cuda", 0)
    tp = torch.from_numpy(p).to(dev)
    edges = [-1.0, 0.1, 0.5, float("inf")]
    parts = []
    for lo, hi in zip(edges[:-1], edges[1:]):
        mine = tp[(tp > lo) & (tp <= hi)].contiguous()
        sp, F = torch.empty_like(mine), torch.empty_like(mine)
        ctx._chk(ctx.L.dk_bh_rank_slice(ctx.h, mine.data_ptr(), mine.numel(), int((tp <= lo).sum().item()), len(p),
                                        sp.data_ptr(), F.data_ptr()))
        ctx.sync()
        parts.append((sp.cpu().numpy(), F.cpu().numpy()))
    for g in range(len(parts) - 1):                                # lower F by the minimum of all higher ranges
        above = min(parts[h][1][0] for h in range(g + 1, len(parts)))
        parts[g] = (parts[g][0], np.minimum(parts[g][1], above))
    sp = np.concatenate([x[0] for x in parts]); F = np.concatenate([x[1] for x in parts])
    assert np.all(np.diff(sp) >= 0)
    got = np.minimum(1.0, F[np.searchsorted(sp, p, side="left")])
    np.testing.assert_allclose(got, want, rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("R,nA,nB", [(20000, 3, 3), (5000, 8, 8), (3000, 2, 5)])
def test_poisson_rows_vs_oracle(ctx, R, nA, nB):
    """TtestFilter -l (TtestFilter.c:152-185): raw p, log2FC, group totals, BH and survivors against the restatement,
    including rows whose density underflows (the reference's 0.001 constant -> negative statistic -> p = 1)"""
    rng = np.random.default_rng(R + nA)
    S = nA + nB
    base = rng.gamma(2.0, 15.0, size=(R, 1))
    cnts = rng.poisson(base * rng.uniform(0.5, 2.0, size=(1, S))).astype(np.int32)
    cnts[: R // 50, nA:] *= 6                                   # real effects
    cnts[R // 50: R // 50 + 20, nA:] += 20000                   # extreme rows: density underflow path
    cnts[-50:] = 0
    idxA, idxB = list(range(nA)), list(range(nA, S))
    m = ctx.matrix_from_host(cnts)
    t, info = ctx.poisson_rows(m, idxA, idxB)
    t.adjust(0.05)
    t.filter(0.05, 0.0)
    p, fc, adj = t.export_rows()
    surv = t.export_survivors()
    t.free(); m.free()
    op, omA, omB, ofc, (N1, N2) = orc.poisson_rows(cnts, idxA, idxB)
    assert (info[0], info[1]) == (N1, N2)
    assert info[2] <= int((op == 1.0).sum())                    # negative statistics are reported as p = 1
    np.testing.assert_allclose(p, op, rtol=1e-9, atol=1e-14)    # 1 - erf(): absolute accuracy near p = 0
    np.testing.assert_allclose(fc, ofc, rtol=1e-9, atol=1e-12)
    oadj = orc.bh(op)
    near = np.abs(oadj - 0.05) <= 1e-9
    keep = np.nonzero((oadj <= 0.05) & ~near)[0]
    rows = surv["row"].astype(np.int64)
    assert set(keep.tolist()) <= set(rows.tolist()) <= set(np.nonzero((oadj <= 0.05) | near)[0].tolist())
    sel = np.isin(rows, keep)
    np.testing.assert_allclose(surv["p_adj"][sel], oadj[rows[sel]], rtol=1e-9, atol=1e-13)
    np.testing.assert_allclose(surv["norm"][sel], cnts[rows[sel]].astype(np.float64))


# ------------------------------------------------- reference-transcript mask (diffFilter.pl) ---
@pytest.mark.parametrize("S,n_union,frac_mask,extra", [(4, 60000, 0.3, 50000), (16, 200000, 0.1, 400000), (1, 5000, 0.5, 10),
                                                      (33, 40000, 0.9, 0), (3, 3000, 0.0, 100000)])
def test_matrix_mask_on_join_output_vs_oracle(ctx, S, n_union, frac_mask, extra):
    """the join leaves a chunked matrix (physical chunks in arbitrary order); the mask reads it in logical order"""
    rng = np.random.default_rng(S * 77 + n_union)
    keys, cnts = random_dumps(rng, S, n_union, 62, 0.4, 12, p_core=0.2)
    m = ctx.join_counts(keys, cnts, 1, 1)
    gk, gc = m.export()
    hit = gk[rng.random(len(gk)) < frac_mask]
    mask = np.sort(np.concatenate([hit, hit[: len(hit) // 7], rng.integers(0, 1 << 62, extra, dtype=np.uint64)]))  # with duplicates
    keep = orc.antijoin(gk, mask)
    mm, removed = ctx.matrix_mask(m, mask)
    k2, c2 = mm.export()
    assert removed == int((~keep).sum()) and mm.rows == int(keep.sum())
    np.testing.assert_array_equal(k2, gk[keep])
    np.testing.assert_array_equal(c2, gc[keep])
    np.testing.assert_array_equal(ctx.keys_antijoin(gk, mask), keep)
    # a masked matrix is an ordinary matrix: NF and the t-test run on it
    if S >= 4 and mm.rows:
        nf, _ = ctx.compute_nf(mm, dk.NF_FULL)
        np.testing.assert_allclose(nf, orc.nf_full(c2), rtol=RTOL)
    mm.free(); m.free()


def test_matrix_mask_edges(ctx):
    cn = np.arange(12, dtype=np.int32).reshape(6, 2)
    ks = np.array([1, 5, 5, 9, 2 ** 62 - 1, 2 ** 64 - 1], np.uint64)         # duplicates in the table, an all-ones key
    m = ctx.matrix_from_host(cn, ks)
    for mask in ([], [0], [5], [1, 5, 9], [2 ** 64 - 1], [0, 2, 3, 2 ** 63], list(ks)):
        mask = np.array(mask, np.uint64)
        keep = orc.antijoin(ks, mask)
        mm, removed = ctx.matrix_mask(m, mask)
        k2, c2 = mm.export() if mm.rows else (np.zeros(0, np.uint64), np.zeros((0, 2), np.int32))
        assert removed == int((~keep).sum())
        np.testing.assert_array_equal(k2, ks[keep])
        np.testing.assert_array_equal(c2.reshape(-1, 2), cn[keep])
        mm.free()
    with pytest.raises(dk.DkError, match="not sorted"):
        ctx.matrix_mask(m, np.array([9, 5, 1], np.uint64))
    m.free()
    e = ctx.matrix_from_host(np.zeros((0, 2), np.int32), np.zeros(0, np.uint64))
    mm, removed = ctx.matrix_mask(e, np.array([1, 2], np.uint64))
    assert mm.rows == 0 and removed == 0
    mm.free(); e.free()
    u = ctx.matrix_from_host(cn[:3], np.array([7, 3, 5], np.uint64))
    with pytest.raises(dk.DkError, match="not sorted"):
        ctx.matrix_mask(u, np.array([1, 2], np.uint64))
    u.free()


# ------------------------------------------------- device text formatter (joinCounts.c:105-109) ---
@pytest.mark.parametrize("S,R,k", [(4, 5000, 31), (16, 40000, 31), (1, 300, 9), (100, 3000, 25), (256, 700, 31), (7, 129, 1)])
def test_format_rows_equals_printf(ctx, S, R, k):
    rng = np.random.default_rng(S + R)
    keys = np.sort(rng.integers(0, 1 << (2 * k), R, dtype=np.uint64))
    cnts = rng.integers(0, 50, (R, S)).astype(np.int32)
    big = rng.random((R, S)) < 0.05                                     # every digit count, both signs, the int limits
    cnts[big] = rng.integers(-2 ** 31, 2 ** 31, int(big.sum()), dtype=np.int64).astype(np.int32)
    cnts[0, 0], cnts[-1, -1] = -2 ** 31, 2 ** 31 - 1
    for d in range(10):
        cnts[min(R - 1, 1 + d), 0] = 10 ** d
        cnts[min(R - 1, 1 + d), S - 1] = 10 ** d - 1
    m = ctx.matrix_from_host(cnts, keys)
    want = format_join(keys, cnts, k).encode()
    assert m.format_rows(k) == want
    lines = want.split(b"\n")
    assert m.format_rows(k, 3, 100) == b"\n".join(lines[3:103]) + b"\n"  # a window of rows
    assert m.format_rows(k, R, 0) == b""
    m.free()


def test_format_rows_on_chunked_join_output(ctx):
    rng = np.random.default_rng(99)
    keys, cnts = random_dumps(rng, 8, 120000, 62, 0.4, 5000, p_core=0.3)
    m = ctx.join_counts(keys, cnts, 2, 3)
    gk, gc = m.export()
    assert m.format_rows(31) == format_join(gk, gc, 31).encode()
    with pytest.raises(dk.DkError):
        ctx.L.dk_matrix_format_rows.restype                             # keep the symbol referenced
        buf = np.empty(8, np.uint8)
        import ctypes as C
        ln = C.c_uint64(0)
        ctx._chk(ctx.L.dk_matrix_format_rows(ctx.h, m.h, 0, m.rows, 31, buf.ctypes.data, 8, C.byref(ln)))
    assert ln.value == len(format_join(gk, gc, 31))                     # too-small buffer: size needed is reported
    m.free()


@pytest.mark.gpu
@pytest.mark.parametrize("G,flow", [(3, "routed"), (2, "gathered")])
def test_sharded_c_abi_path_on_one_device(ctx, G, flow):
    """The multi-GPU path, shard by shard on ONE device: a seeded input with a planted effect is cut into G key ranges,
    every shard goes through the C-ABI building blocks the ranks call (dk_join_counts_dev, dk_nf_begin/_set_rank/
    _hist_dev/_advance_dev/_result, dk_ttest_rows, dk_ttest_candidates, then either the routed BH — dk_bh_rank_slice per
    p-value range + dk_ttest_apply_adjusted — or dk_ttest_adjust_global on the gathered list, dk_ttest_filter), the
    'collectives' are done by hand on the host, and U, R, NFs, adjusted p and the survivor set must equal the oracle on
    the UNSHARDED input (TtestFilter.c:459-497, computeNF.c:38-47)."""
    import torch
    from common import plant_effect
    from dekupl_run_b200.distributed import GpuBackend
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(100 + G)
    S, nA, r, a, pmax = 8, 4, 2, 5, 0.05
    idxA, idxB = list(range(nA)), list(range(nA, S))
    keys, cnts = random_dumps(rng, S, 150_000, 62, 0.15, 60, p_core=0.3)
    cnts = plant_effect(rng, keys, cnts, set(idxB))
    ok, oc, oU = orc.join(keys, cnts, r, a)
    onf = orc.nf_full(oc)
    op, omA, omB, ofc = orc.ttest_rows(oc, onf, idxA, idxB)
    oadj = orc.bh(op)
    okeep = np.nonzero(oadj <= pmax)[0]
    assert len(okeep) > 50
    # key-range shards
    edges = [0] + [int(ok[len(ok) * g // G]) for g in range(1, G)] + [1 << 63]
    bes, bufs, rows0 = [], [], []
    R_glob = U_glob = 0
    for g in range(G):
        tk, tc = [], []
        for k, c in zip(keys, cnts):
            lo, hi = np.searchsorted(k, np.uint64(edges[g])), np.searchsorted(k, np.uint64(edges[g + 1]))
            tk.append(torch.from_numpy(k[lo:hi].view(np.int64).copy()).to(dev))
            tc.append(torch.from_numpy(c[lo:hi].copy()).to(dev))
        bufs.append((tk, tc))
        be = GpuBackend(ctx, dev)
        Rl, Ul = be.join([t.data_ptr() for t in tk], [t.data_ptr() for t in tc], [t.numel() for t in tk], r, a)
        rows0.append(R_glob)
        R_glob += Rl; U_glob += Ul
        bes.append(be)
    assert (U_glob, R_glob) == (oU, len(ok))
    # NF, full mode: histograms summed over the shards by hand (the all-reduce)
    n_pos = sum(be.nf_begin() for be in bes)
    for be in bes:
        be.nf_set_rank(n_pos // 2)
    for _ in range(8):
        hs = [be.nf_hist().clone() for be in bes]
        tot = torch.stack(hs).sum(0)
        for be in bes:
            be.nf_advance(tot.clone())
    nfs = [be.nf_result() for be in bes]
    for v in nfs:
        np.testing.assert_allclose(v, onf, rtol=RTOL)
    for be in bes:
        be.ttest_rows(onf, idxA, idxB)
    cands = [be.candidates(pmax).clone() for be in bes]
    if flow == "gathered":
        allp = torch.cat(cands)
        for be in bes:
            be.adjust_global(allp, R_glob)
    else:
        pe = [pmax * g / G for g in range(1, G)]
        sends = [torch.sort(c)[0] for c in cands]
        cuts = [[0] + [int(torch.searchsorted(s, torch.tensor([e], dtype=torch.float64, device=dev), right=True).item()) for e in pe]
                + [s.numel()] for s in sends]
        mines = [torch.cat([sends[src][cuts[src][g]:cuts[src][g + 1]] for src in range(G)]) for g in range(G)]
        n_below = [sum(cuts[src][g] for src in range(G)) for g in range(G)]
        ranked = [bes[g].bh_rank_slice(mines[g].contiguous(), n_below[g], R_glob) for g in range(G)]
        ctx.sync()
        firsts = [float(F[0].item()) if F.numel() else float("inf") for _, F in ranked]
        for g in range(G):
            above = min(firsts[g + 1:], default=None)
            sp, F = ranked[g]
            ranked[g] = (sp, torch.clamp(F, max=above) if (above is not None and F.numel()) else F)
        for src in range(G):
            parts = []
            for g in range(G):
                seg = sends[src][cuts[src][g]:cuts[src][g + 1]]
                sp, F = ranked[g]
                parts.append(F[torch.searchsorted(sp, seg)] if seg.numel() else seg)
            bes[src].apply_adjusted(sends[src], torch.cat(parts).contiguous())
    got_rows, got_adj = [], []
    for g, be in enumerate(bes):
        D, ties = be.filter(pmax, 0.0)
        sv = be.t.export_survivors()
        got_rows.append(sv["row"].astype(np.int64) + rows0[g])
        got_adj.append(sv["p_adj"])
        p, fc, adj = be.t.export_rows()
        sl = slice(rows0[g], rows0[g] + len(p))
        np.testing.assert_allclose(p, op[sl], rtol=RTOL, atol=1e-300)
    got_rows, got_adj = np.concatenate(got_rows), np.concatenate(got_adj)
    near = np.abs(oadj - pmax) <= 1e-9 * pmax
    assert set(np.nonzero((oadj <= pmax) & ~near)[0].tolist()) <= set(got_rows.tolist()) <= set(np.nonzero((oadj <= pmax) | near)[0].tolist())
    np.testing.assert_allclose(got_adj, oadj[got_rows], rtol=RTOL, atol=1e-300)
    for be in bes:
        be.free()


@pytest.mark.gpu
@pytest.mark.parametrize("G,on_device", [(2, True), (8, True), (5, False)])
def test_key_splitters_balance_records_on_clustered_kmers(ctx, G, on_device):
    """dk_key_splitters: shards of equal RECORD counts on k-mers that are far from uniform (nested clusters,
    dumps of very different sizes): every shard within 3 % of 1/G of all records, where an equal split of the key space
    is off by a large factor; joining the shards one by one and concatenating equals joining everything at once."""
    import torch
    rng = np.random.default_rng(77 + G)
    allk = _clustered_keys(rng, 400_000, 5)
    keys, cnts = [], []
    for s in range(6):
        m = rng.random(len(allk)) < (0.9 if s < 2 else 0.15)
        keys.append(allk[m]); cnts.append(rng.integers(1, 30, size=int(m.sum())).astype(np.int32))
    n = [len(k) for k in keys]
    if on_device:
        tk = [torch.from_numpy(k.view(np.int64).copy()).cuda() for k in keys]
        spl = ctx.key_splitters([t.data_ptr() for t in tk], n, G, device_inputs=True)
    else:
        spl = ctx.key_splitters([k.ctypes.data for k in keys], n, G, device_inputs=False)
    assert len(spl) == G - 1 and np.all(np.diff(spl.astype(np.float64)) >= 0)
    UMAX = np.array([np.iinfo(np.uint64).max], np.uint64)
    edges = np.concatenate([np.zeros(1, np.uint64), spl.astype(np.uint64), UMAX])
    per = np.array([sum(int(np.searchsorted(k, edges[g + 1]) - np.searchsorted(k, edges[g])) for k in keys) for g in range(G)], float)
    per[-1] += sum(int((k == np.iinfo(np.uint64).max).sum()) for k in keys)
    assert per.sum() == sum(n)
    assert np.max(np.abs(per / per.sum() * G - 1.0)) < 0.03, per / per.sum()
    naive = np.linspace(float(allk.min()), float(allk.max()), G + 1)[1:-1].astype(np.uint64)
    ne = np.concatenate([np.zeros(1, np.uint64), naive, UMAX])
    nper = np.array([sum(int(np.searchsorted(k, ne[g + 1]) - np.searchsorted(k, ne[g])) for k in keys) for g in range(G)], float)
    assert np.max(nper) / np.mean(nper) > 1.5                       # the key-space split really is unbalanced here
    parts_k, parts_c = [], []
    for g in range(G):
        sk, sc = [], []
        for k, c in zip(keys, cnts):
            lo, hi = np.searchsorted(k, edges[g]), (np.searchsorted(k, edges[g + 1]) if g + 1 < G else len(k))
            sk.append(k[lo:hi]); sc.append(c[lo:hi])
        m = ctx.join_counts(sk, sc, 2, 5)
        gk, gc = m.export(); m.free()
        parts_k.append(gk); parts_c.append(gc)
    ok, oc, _ = orc.join(keys, cnts, 2, 5)
    np.testing.assert_array_equal(np.concatenate(parts_k), ok)
    np.testing.assert_array_equal(np.concatenate(parts_c), oc)


@pytest.mark.gpu
@pytest.mark.parametrize("S,nA,chunk", [(8, 4, 60_000), (20, 10, 250_000), (12, 6, 10 ** 9)])
def test_chunk_streamed_path_equals_the_oracle(ctx, S, nA, chunk):
    """dk_stream_nf_reference + dk_stream_path (host dumps walked in key-range chunks, only candidate rows kept) + the
    ordinary BH / filter / export calls == oracle on the whole input: U, R, NFs, survivor k-mers, adjusted p, means,
    log2FC, normalised counts (joinCounts.c:72-115, computeNF.c:138-168, TtestFilter.c:373-497)."""
    from common import plant_effect
    rng = np.random.default_rng(S * 13 + nA)
    r, a, pmax, step = 2, 3, 0.05, 5000
    idxA, idxB = list(range(nA)), list(range(nA, S))
    keys, cnts = random_dumps(rng, S, 120_000, 62, 0.2, 50, p_core=0.35)
    cnts = plant_effect(rng, keys, cnts, set(idxB), frac=0.04, factor=25)
    kp, cp, n = [k.ctypes.data for k in keys], [c.ctypes.data for c in cnts], [len(k) for k in keys]
    ok, oc, oU = orc.join(keys, cnts, r, a)
    onf, used, uz = orc.nf_reference(oc, step)
    assert not uz
    # whether computeNF's early stop fires inside the first chunk, inside a widened head, or never (whole table): same factors
    nf, info = ctx.stream_nf_reference(kp, cp, n, r, a, chunk, step)
    np.testing.assert_allclose(nf, onf, rtol=RTOL)
    assert info["consumed"] == used or (used == len(oc) and info["consumed"] >= used)
    nf_used = np.array([float("%f" % v) for v in onf])
    st, m, t = ctx.stream_path(kp, cp, n, r, a, chunk, nf_used, idxA, idxB, pmax)
    total = sum(n)
    assert st["chunks"] == max(1, -(-total // max(chunk, 4096))) or st["chunks"] >= 1
    assert (st["U"], st["R"]) == (oU, len(ok)) and st["h2d_bytes"] == 12 * total
    op, omA, omB, ofc = orc.ttest_rows(oc, nf_used, idxA, idxB)
    cand = np.nonzero(op <= pmax)[0]
    near_p = np.abs(op - pmax) <= 1e-9 * pmax
    assert abs(st["n_cand"] - len(cand)) <= int(near_p.sum()) and st["n_cand"] > 50
    gk, gc = m.export()
    if not near_p.any():
        np.testing.assert_array_equal(gk, ok[cand])            # candidate rows in k-mer order, counts intact
        np.testing.assert_array_equal(gc, oc[cand])
    nc, ptr = t.candidates(pmax)
    t.adjust_global(ptr, nc, st["R"])                          # R_global = ALL filtered rows, not the candidate count
    D = t.filter(pmax, 0.0)
    sv = t.export_survivors()
    oadj = orc.bh(op)
    near = np.abs(oadj - pmax) <= 1e-9 * pmax
    keep = np.nonzero((oadj <= pmax) & ~near)[0]
    assert D > 20 and set(ok[keep].tolist()) <= set(sv["key"].tolist()) <= set(ok[np.nonzero((oadj <= pmax) | near)[0]].tolist())
    rows = np.searchsorted(ok, sv["key"])
    assert np.all(np.diff(rows) > 0)                           # survivors come out in the reference's order
    np.testing.assert_allclose(sv["p_adj"], oadj[rows], rtol=RTOL, atol=1e-300)
    np.testing.assert_allclose(sv["meanA"], omA[rows], rtol=RTOL)
    np.testing.assert_allclose(sv["meanB"], omB[rows], rtol=RTOL)
    np.testing.assert_allclose(sv["log2fc"], ofc[rows], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(sv["norm"], oc[rows] / nf_used, rtol=RTOL)
    t.free(); m.free()
    # a first chunk too small for computeNF's early stop: the head of the table is widened until the stop fires
    nf2, info2 = ctx.stream_nf_reference(kp, cp, n, r, a, 8000, step)
    np.testing.assert_allclose(nf2, onf, rtol=RTOL)
    # without widening (a rank of a sharded run) the call reports honestly whether the stop fired inside the head
    nf3, info3 = ctx.stream_nf_reference(kp, cp, n, r, a, 8000, step, widen=False)
    assert info3["stop_fired"] == (used < len(oc) and used < info3["consumed"] + 1 and info3["consumed"] == used) or not info3["stop_fired"]
    # the two-sweep form over three emulated ranks: all-positive rows of every rank's key range with global row numbers,
    # computeNF's step logic replayed on them == computeNF on the whole table, whether or not its stop fires
    edges = [0] + [int(ok[len(ok) * g // 3]) for g in (1, 2)] + [1 << 63]
    rows_all, cnts_all, R_tot = [], [], 0
    for g in range(3):
        sk, sc = [], []
        for k, c in zip(keys, cnts):
            lo, hi = np.searchsorted(k, np.uint64(edges[g])), np.searchsorted(k, np.uint64(edges[g + 1]))
            sk.append(np.ascontiguousarray(k[lo:hi])); sc.append(np.ascontiguousarray(c[lo:hi]))
        st1, prow, pcnt = ctx.stream_positive_rows([x.ctypes.data for x in sk], [x.ctypes.data for x in sc], [len(x) for x in sk], r, a, 50_000)
        rows_all.append(prow + np.uint64(R_tot)); cnts_all.append(pcnt)
        R_tot += st1["R"]
    assert R_tot == len(ok)
    rows_all, cnts_all = np.concatenate(rows_all), np.concatenate(cnts_all)
    want_pos = np.nonzero((oc > 0).all(axis=1))[0]
    np.testing.assert_array_equal(rows_all, want_pos.astype(np.uint64))
    np.testing.assert_array_equal(cnts_all, oc[want_pos])
    for stp in (step, 3000, 10 ** 9):
        want, wused, wuz = orc.nf_reference(oc, stp)
        if wuz:
            continue
        got, ginfo = ctx.nf_reference_sparse(rows_all, cnts_all, R_tot, stp)
        np.testing.assert_allclose(got, want, rtol=RTOL)
        assert ginfo["consumed"] == wused


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["c1", "r88", "r35"])
def test_ttest_kernels_vs_scipy_only_goldens(ctx, tag):
    """the row kernels (fast 8-per-group kernel for r88 / c1, general kernel for r35) and the BH kernels against vectors
    computed by scipy alone (tests/golden/make_ttest_independent.py): p, log2FC, adjusted p of every row"""
    G = np.load(os.path.join(GOLD, "ttest_independent.npz"))
    c, nf, iA, iB = G[f"{tag}_counts"], G[f"{tag}_nf"], G[f"{tag}_idxA"].tolist(), G[f"{tag}_idxB"].tolist()
    m = ctx.matrix_from_host(c)
    t = ctx.ttest_rows(m, nf, iA, iB)
    t.adjust(2.0)                                              # every row is a candidate: the full adjusted vector
    p, fc, adj = t.export_rows()
    t.free(); m.free()
    np.testing.assert_allclose(p, G[f"{tag}_p"], rtol=RTOL, atol=1e-300)
    np.testing.assert_allclose(fc, G[f"{tag}_fc"], rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(adj, G[f"{tag}_adj"], rtol=RTOL, atol=1e-300)
    assert int((p == 0.5).sum()) == int((G[f"{tag}_p"] == 0.5).sum())


@pytest.mark.gpu
@pytest.mark.parametrize("slices,device_inputs", [(6, False), (1, False), (1, True)])
def test_run_path_matrix_out_returns_the_matrix_in_reference_order(ctx, slices, device_inputs, monkeypatch):
    """dk_run_path_matrix_out: the filtered matrix written to host buffers (slice by slice beside the H2D copies when the
    dumps are host arrays) == joinCounts' rows in order (oracle), same statistics as dk_run_path; too small a buffer is
    refused with the need reported"""
    import ctypes as C
    import torch
    from common import plant_effect
    rng = np.random.default_rng(9 + slices)
    S, nA = 8, 4
    keys, cnts = random_dumps(rng, S, 90_000, 62, 0.2, 40, p_core=0.3)
    cnts = plant_effect(rng, keys, cnts, set(range(nA, S)))
    ok, oc, oU = orc.join(keys, cnts, 2, 5)
    monkeypatch.setenv("DK_JOIN_SLICES", str(slices))
    if device_inputs:
        tk = [torch.from_numpy(k.view(np.int64).copy()).cuda() for k in keys]
        tc = [torch.from_numpy(c.copy()).cuda() for c in cnts]
        kp, cp = [t.data_ptr() for t in tk], [t.data_ptr() for t in tc]
    else:
        kp, cp = [k.ctypes.data for k in keys], [c.ctypes.data for c in cnts]
    n = [len(k) for k in keys]
    cap = len(ok) + 10
    hk, hc = np.zeros(cap, np.uint64), np.zeros((cap, S), np.int32)
    st, m, t, rows = ctx.run_path_matrix_out(kp, cp, n, device_inputs, 2, 5, dk.NF_REFERENCE, list(range(nA)), list(range(nA, S)), 0.05, 0.0,
                                             hk.ctypes.data, hc.ctypes.data, cap)
    assert rows == len(ok) == st["R"] and st["U"] == oU and st["D"] > 0
    np.testing.assert_array_equal(hk[:rows], ok)
    np.testing.assert_array_equal(hc[:rows], oc)
    st2 = ctx.run_path(kp, cp, n, device_inputs, 2, 5, dk.NF_REFERENCE, list(range(nA)), list(range(nA, S)), 0.05, 0.0)
    assert (st2["U"], st2["R"], st2["D"]) == (st["U"], st["R"], st["D"])
    t.free(); m.free()
    with pytest.raises(dk.DkError, match="host buffers hold"):
        ctx.run_path_matrix_out(kp, cp, n, device_inputs, 2, 5, dk.NF_REFERENCE, list(range(nA)), list(range(nA, S)), 0.05, 0.0,
                                hk.ctypes.data, hc.ctypes.data, 100)


@pytest.mark.gpu
@pytest.mark.parametrize("nf_mode", ["reference", "full"])
def test_run_path_sharded_under_the_c_abi_one_rank_communicator(nf_mode):
    """dk_comm_* + dk_run_path_sharded with a ONE-rank NCCL communicator: every collective of the C driver runs (all-reduce,
    broadcast, the radix-select histogram all-reduce) and the result must equal dk_run_path and the oracle. The N > 1 case
    is checked against the oracle inside every multi-GPU bench run (parity_check)."""
    from common import plant_effect
    c = dk.Context(0)
    try:
        c.comm_init(c.comm_unique_id(), 0, 1)
        assert c.L.dk_comm_world(c.h) == 1 and c.L.dk_comm_rank(c.h) == 0
        rng = np.random.default_rng(31)
        S, nA = 8, 4
        keys, cnts = random_dumps(rng, S, 150_000, 62, 0.15, 60, p_core=0.3)
        cnts = plant_effect(rng, keys, cnts, set(range(nA, S)))
        kp, cp, n = [k.ctypes.data for k in keys], [x.ctypes.data for x in cnts], [len(k) for k in keys]
        mode = dk.NF_REFERENCE if nf_mode == "reference" else dk.NF_FULL
        st, m, t = c.run_path_sharded(kp, cp, n, False, 2, 5, mode, list(range(nA)), list(range(nA, S)), 0.05, 0.0, keep=True)
        ok, oc, oU = orc.join(keys, cnts, 2, 5)
        onf = orc.nf_reference(oc)[0] if nf_mode == "reference" else orc.nf_full(oc)
        assert (st["U"], st["R"], st["U_local"], st["R_local"]) == (oU, len(ok), oU, len(ok))
        np.testing.assert_allclose(st["nf"], onf, rtol=RTOL)
        op, _, _, _ = orc.ttest_rows(oc, np.array([float("%f" % v) for v in onf]), list(range(nA)), list(range(nA, S)))
        oadj = orc.bh(op)
        keep = np.nonzero(oadj <= 0.05)[0]
        sv = t.export_survivors()
        assert st["D"] == len(keep) > 50 and np.array_equal(sv["row"].astype(np.int64), keep)
        np.testing.assert_allclose(sv["p_adj"], oadj[keep], rtol=RTOL, atol=1e-300)
        st1 = c.run_path(kp, cp, n, False, 2, 5, mode, list(range(nA)), list(range(nA, S)), 0.05, 0.0)
        assert (st1["U"], st1["R"], st1["D"]) == (st["U"], st["R"], st["D"])
        t.free(); m.free()
        c.comm_free()
    finally:
        c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("R,nA,nB", [(20000, 24, 24), (6000, 100, 100), (9000, 9, 31), (4000, 33, 12)])
def test_ttest_large_groups_fast_kernel_vs_oracle(ctx, R, nA, nB):
    """groups of more than 8 samples (cfg 3: 24 v 24, cfg 4: 100 v 100, uneven groups): the two-pass variant of the fast row
    kernel against the oracle — p, log2FC, survivors, adjusted p — and bit-identical to the general kernel's exact tests
    (p == 0.5 rows)"""
    rng = np.random.default_rng(R + nA)
    S = nA + nB
    base = np.floor(rng.pareto(1.2, size=(R, 1)) * 4 + 1)
    cnts = rng.poisson(base * rng.uniform(0.5, 2.0, size=(1, S))).astype(np.int32)
    cnts[rng.random((R, S)) < 0.1] = 0
    cnts[: R // 25, nA:] *= 3
    cnts[-20:] = 0                                             # both sds exactly 0
    cnts[-40:-20, :nA] = 4; cnts[-40:-20, nA:] = 11            # constant groups with different means
    cnts[R // 2, :] = 2 ** 31 - 1                              # far beyond both tables
    nf = np.array([float("%f" % v) for v in rng.uniform(0.6, 1.8, S)])
    D, ties = _ttest_both(ctx, cnts, nf, list(range(nA)), list(range(nA, S)), 0.05, 0.0)
    assert D > 0

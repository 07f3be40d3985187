"""k-mer counting of a sample from its reads: jellyfish count -L 2 -m 31 on the two mates (mate 1 through
revCompFastq.pl for the bundled "rf" library), dump, sort — Snakefile:523-582. jellyfish is absent from the image; the
fixtures' dumps were produced by an independent restatement (tests/golden/make_config1.py, a Python dict over strings)
and then consumed by the UNMODIFIED reference joinCounts/computeNF, so the C oracle and the CUDA counter are checked
against those dumps and, through them, against the reference's raw-counts table."""
import gzip
import os

import numpy as np
import pytest

from common import GOLD, parse_dump_text

C1 = os.path.join(GOLD, "config1")
NAMES = ["sample1", "sample2", "sample3", "sample4"]


def _reads(nm):
    seqs, rc = [], []
    for line in gzip.open(os.path.join(C1, f"reads_{nm}.txt.gz"), "rt"):
        mate, sq = line.rstrip("\n").split("\t")
        seqs.append(sq)
        rc.append(1 if mate == "1" else 0)               # lib_type rf: mate 1 is reverse-complemented (Snakefile:540-542)
    return seqs, np.array(rc, np.uint8)


def _dump(nm):
    return parse_dump_text(gzip.open(os.path.join(C1, nm + ".txt.gz"), "rt").read())


def _random_reads(rng, n, canonical_friendly=False):
    reads = []
    for _ in range(n):
        L = int(rng.integers(0, 90))
        s = "".join(rng.choice(list("ACGT"), L))
        if L and rng.random() < 0.3:
            p = int(rng.integers(0, L))
            s = s[:p] + rng.choice(list("Nn.R")) + s[p + 1:]
        if rng.random() < 0.2:
            s = s.lower()
        reads.append(s)
    reads += [reads[i] for i in rng.integers(0, n, n // 2)]            # repeats, so that counts exceed 1
    return reads


def test_oracle_counter_reproduces_the_fixture_dumps():
    import oracle_lib as orc
    for nm in NAMES[:2]:
        seqs, rc = _reads(nm)
        k, c = orc.count_kmers(seqs, 31, rc, canonical=False, min_count=2)
        wk, wc, _ = _dump(nm)
        assert np.array_equal(k, wk) and np.array_equal(c, wc), nm


def test_oracle_counter_small_cases():
    import oracle_lib as orc
    from common import enc
    k, c = orc.count_kmers(["ACGTA", "ACGTA", "acgtn", "TACGT"], 4, None, False, 1)
    want = {"ACGT": 4, "CGTA": 2, "TACG": 1}
    assert {int(a): int(b) for a, b in zip(k, c)} == {enc(s): n for s, n in want.items()}
    k, c = orc.count_kmers(["ACGTA"], 4, [1], False, 1)                  # reverse complement of the read: TACGT
    assert {int(a): int(b) for a, b in zip(k, c)} == {enc("TACG"): 1, enc("ACGT"): 1}
    k, c = orc.count_kmers(["AAAC", "GTTT"], 4, None, True, 1)            # -C: both are the canonical AAAC
    assert list(k) == [enc("AAAC")] and list(c) == [2]
    k, c = orc.count_kmers(["ACG", ""], 4, None, False, 1)                # shorter than k
    assert len(k) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("k,canonical,min_count", [(31, False, 2), (31, True, 2), (5, False, 1), (12, True, 3), (1, False, 1)])
def test_count_kmers_vs_oracle_random(ctx, k, canonical, min_count):
    import oracle_lib as orc
    rng = np.random.default_rng(k * 7 + min_count)
    reads = _random_reads(rng, 3000)
    rc = (rng.random(len(reads)) < 0.5).astype(np.uint8)
    d = ctx.count_kmers(reads, k, rc, canonical, min_count)
    gk, gc = d.export()
    ok, oc = orc.count_kmers(reads, k, rc, canonical, min_count)
    st = d.stats()
    d.free()
    np.testing.assert_array_equal(gk, ok)
    np.testing.assert_array_equal(gc, oc)
    assert st["windows"] == sum(max(0, len(r) - k + 1) for r in reads) and st["valid"] <= st["windows"]


@pytest.mark.gpu
def test_count_kmers_edges(ctx):
    for reads in ([], [""], ["ACG"], ["NNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNNN"]):
        d = ctx.count_kmers(reads, 31, None, False, 1)
        assert d.size == 0
        d.free()
    d = ctx.count_kmers(["A" * 40] * 3, 31, None, False, 2)
    k, c = d.export()
    assert list(k) == [0] and list(c) == [30]
    d.free()


@pytest.mark.gpu
def test_reads_to_diff_counts_without_text_dumps(ctx):
    """config 1 from the reads: device k-mer counter -> dk_join_counts_dev -> NF -> t-test; equals the reference's files"""
    import dekupl_run_b200 as dk
    import oracle_lib as orc
    dumps = []
    for nm in NAMES:
        seqs, rc = _reads(nm)
        d = ctx.count_kmers(seqs, 31, rc, False, 2)
        wk, wc, _ = _dump(nm)
        gk, gc = d.export()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc), nm
        dumps.append(d)
    stats, m, t = ctx.run_path([d.keys_ptr for d in dumps], [d.cnts_ptr for d in dumps], [d.size for d in dumps], True, 2, 5,
                               dk.NF_REFERENCE, [0, 1], [2, 3], 0.2, 0.0, nf_text_roundtrip=True, keep=True)
    txt = m.format_rows(31).decode()
    assert txt == gzip.open(os.path.join(C1, "raw-counts.tsv.gz"), "rt").read().split("\n", 1)[1]     # reference joinCounts bytes
    want_nf = [float(l.split("\t")[1]) for l in open(os.path.join(C1, "normalization_factors.tsv")).read().strip().split("\n")[1:]]
    assert ["%f" % v for v in stats["nf"]] == ["%f" % v for v in want_nf]                               # reference computeNF
    surv = t.export_survivors()
    want = [l.split("\t")[0] for l in gzip.open(os.path.join(C1, "diff-counts.tsv.gz"), "rt").read().strip().split("\n")[1:]]
    from common import dec
    assert [dec(int(x), 31) for x in surv["key"]] == want
    t.free(); m.free()
    for d in dumps:
        d.free()
    _ = orc


@pytest.mark.gpu
@pytest.mark.parametrize("nbatch,canonical", [(3, False), (7, True), (1, False)])
def test_batched_counting_adds_up(ctx, nbatch, canonical):
    """a sample counted in read batches (min_count 1) and summed equals the single-call table (-L on the sums)"""
    rng = np.random.default_rng(nbatch)
    reads = _random_reads(rng, 4000)
    rc = (rng.random(len(reads)) < 0.5).astype(np.uint8)
    whole = ctx.count_kmers(reads, 21, rc, canonical, 3)
    wk, wc = whole.export()
    cuts = np.linspace(0, len(reads), nbatch + 1).astype(int)
    parts = [ctx.count_kmers(reads[a:b], 21, rc[a:b], canonical, 1) for a, b in zip(cuts[:-1], cuts[1:])]
    merged = ctx.merge_dumps(parts, 3)
    mk, mc = merged.export()
    np.testing.assert_array_equal(mk, wk)
    np.testing.assert_array_equal(mc, wc)
    assert merged.stats()["windows"] == whole.stats()["windows"] and merged.stats()["valid"] == whole.stats()["valid"]
    for d in parts + [merged, whole]:
        d.free()

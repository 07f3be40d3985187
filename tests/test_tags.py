"""mergeTags (share/mergeTags/mergeTags.c): contigs of the differentially expressed k-mers, stranded mode.
The checker is the UNMODIFIED reference binary compiled into oracle/_ref/mergeTags (it travels to the GPU box)."""
import gzip
import os
import random
import subprocess

import pytest

from common import GOLD

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "dekupl-run_b200", "host", "bin")
REF = os.path.join(ROOT, "oracle", "_ref", "mergeTags")
C1 = os.path.join(GOLD, "config1")


def _random_table(rnd, path):
    G = rnd.choice([300, 2000, 8000])
    g = "".join(rnd.choice("ACGT") for _ in range(G))
    if rnd.random() < 0.5:                                   # repeats: shared ends, occasionally rings
        p = rnd.randrange(0, G - 200)
        g = g + g[p:p + rnd.choice([20, 40, 120])] + "".join(rnd.choice("ACGT") for _ in range(100))
    variants = [g]
    for _ in range(rnd.choice([0, 2, 5])):                   # SNP variants: branches
        p = rnd.randrange(40, len(g) - 40)
        variants.append(g[p - 35:p] + rnd.choice("ACGT") + g[p + 1:p + 35])
    kms = set()
    for s in variants:
        keep = rnd.choice([0.3, 0.7, 0.95])
        for i in range(len(s) - 30):
            if rnd.random() < keep:
                kms.add(s[i:i + 31])
    kms = sorted(kms)
    if rnd.random() < 0.5:
        rnd.shuffle(kms)
    lines = ["tag\tpvalue\tmeanA\tmeanB\tlog2FC\ts1\ts2"]
    for km in kms:
        p = rnd.choice([0.0, 0.01, 0.02, round(rnd.random() * 0.05, 6)])          # many ties
        fc = rnd.choice([1.5, -1.5, 0.0, 2.25]) if rnd.random() < 0.3 else 1.0
        lines.append("%s\t%f\t%f\t%f\t%f\t%.2f\t%.2f" % (km, p, 1.0, 2.0, fc, 1.0, 2.0))
    with gzip.open(path, "wt") as f:
        f.write("\n".join(lines) + "\n")
    return len(kms)


def _ref(path, *opts):
    r = subprocess.run([REF, *opts, path], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-300:]
    return r.stdout


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/mergeTags not built")
def test_model_equals_reference_binary(tmp_path):
    from tags_model import model
    for nm in ("diff-counts.tsv.gz", "diff-counts-masked.tsv.gz"):
        got, rings = model(os.path.join(C1, nm))
        assert got == _ref(os.path.join(C1, nm), "-k", "31") and rings == 0
    rnd = random.Random(4)
    for t in range(25):
        p = str(tmp_path / f"t{t}.tsv.gz")
        _random_table(rnd, p)
        got, rings = model(p)
        if rings == 0:
            assert got == _ref(p, "-k", "31"), t


def test_mergeTags_cli_argv_surface():
    r = subprocess.run([os.path.join(BIN, "mergeTags")], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage:   mergeTags [options] <counts.tsv>" in r.stderr and "min assembly overlap" in r.stderr
    r = subprocess.run([os.path.join(BIN, "mergeTags"), "-n", "x.tsv"], capture_output=True, text=True)
    assert r.returncode != 0 and "unstranded" in r.stderr
    r = subprocess.run([os.path.join(BIN, "mergeTags"), "/nonexistent/x.tsv"], capture_output=True, text=True)
    assert r.returncode != 0 and "Failed to open /nonexistent/x.tsv" in r.stderr


@pytest.mark.gpu
def test_mergeTags_cli_equals_reference_binary_on_config1():
    for nm in ("diff-counts.tsv.gz", "diff-counts-masked.tsv.gz"):
        r = subprocess.run([os.path.join(BIN, "mergeTags"), "-k", "31", os.path.join(C1, nm)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert r.stdout == _ref(os.path.join(C1, nm), "-k", "31"), nm


@pytest.mark.gpu
def test_mergeTags_cli_equals_reference_binary_random(tmp_path):
    rnd = random.Random(9)
    checked = 0
    for t in range(30):
        p = str(tmp_path / f"t{t}.tsv.gz")
        _random_table(rnd, p)
        opts = ["-k", "31"] + (["-m", str(rnd.choice([15, 20, 29]))] if t % 3 == 0 else [])
        r = subprocess.run([os.path.join(BIN, "mergeTags"), *opts, p], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        if "rings" in r.stderr:                               # the reference's answer there depends on its hash order
            continue
        assert r.stdout == _ref(p, *opts), (t, opts)
        checked += 1
    assert checked >= 20


@pytest.mark.gpu
def test_merge_tags_api_small(ctx):
    # ACGTACGTAC... three 8-mers in a chain with 7-base overlaps, and one unrelated k-mer
    s = "ACGGTCATTGC"
    kms = [s[i:i + 8] for i in range(4)] + ["TTTTTTTT"]
    got, rings = ctx.merge_tags(kms, [0.5, 0.1, 0.1, 0.3, 0.2], k=8, min_k=5)
    assert rings == 0
    assert sorted(got) == [(2, 4, s), (4, 1, "TTTTTTTT")]     # representative: lowest p, rightmost on ties (id 2)
    got, _ = ctx.merge_tags([], [], k=8)
    assert got == []

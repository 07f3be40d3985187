#!/usr/bin/env python
"""Generates the committed golden vectors in tests/golden/ from
  * the UNMODIFIED reference tools built by oracle/Makefile into oracle/_ref/
    (joinCounts, computeNF) run on the reference's own fixtures and on seeded inputs;
  * scipy (Student-t tail) for the one piece the reference delegates to Boost.
Run in the build container (needs /root/reference and oracle/_ref):
    make -C oracle && python tests/golden/make_golden.py
The GPU box never runs this; tests only read the JSON/NPZ files written here."""
import gzip
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/share"
JOIN = os.path.join(ROOT, "oracle/_ref/joinCounts")
NF = os.path.join(ROOT, "oracle/_ref/computeNF")
BASES = "ACGT"


def run(cmd):
    return subprocess.run(cmd, check=True, capture_output=True, text=True).stdout


def kmer_str(v, k):
    return "".join(BASES[(int(v) >> (2 * (k - 1 - i))) & 3] for i in range(k))


def random_dumps(rng, S, n_union, k, p_present, max_count):
    keys = np.unique(rng.integers(0, 1 << (2 * k), size=n_union, dtype=np.uint64))
    dumps = []
    for _ in range(S):
        m = rng.random(len(keys)) < p_present
        c = rng.integers(1, max_count, size=int(m.sum()))
        dumps.append((keys[m], c.astype(np.int32)))
    return dumps


def golden_join(out):
    t1 = open(f"{REF}/joinCounts/test/test1.tsv").read()
    t2 = open(f"{REF}/joinCounts/test/test2.tsv").read()
    cases = []
    for args in ([], ["-r", "2"], ["-r", "2", "-a", "5"], ["-r", "0"], ["-a", "4"]):
        cases.append(dict(args=args, inputs=[t1, t2],
                          stdout=run([JOIN] + args + [f"{REF}/joinCounts/test/test1.tsv", f"{REF}/joinCounts/test/test2.tsv"])))
    rng = np.random.default_rng(20261104)
    with tempfile.TemporaryDirectory() as d:
        for S, nu, k, pp, mc, args in ((5, 400, 31, 0.5, 12, ["-r", "2", "-a", "3"]), (16, 3000, 31, 0.3, 9, ["-r", "3", "-a", "5"]),
                                       (3, 50, 5, 0.9, 100, []), (40, 800, 21, 0.2, 7, ["-r", "4", "-a", "2"])):
            dumps = random_dumps(rng, S, nu, k, pp, mc)
            files, texts = [], []
            for s, (kk, cc) in enumerate(dumps):
                txt = "".join(f"{kmer_str(a, k)}\t{b}\n" for a, b in zip(kk, cc))
                if len(kk) == 0:      # the reference segfaults on an empty file (joinCounts.c:59-61); keep one record
                    txt = f"{kmer_str(0, k)}\t1\n"
                p = os.path.join(d, f"s{s}.txt")
                open(p, "w").write(txt)
                files.append(p)
                texts.append(txt)
            cases.append(dict(args=args, inputs=texts, stdout=run([JOIN] + args + files)))
    json.dump(cases, open(os.path.join(out, "join_cases.json"), "w"))
    print("join cases:", len(cases))


def golden_nf(out):
    res = {}
    # (1) the reference's own fixtures
    res["computeNF_fixture"] = run([NF, f"{REF}/computeNF/test/raw-counts.tsv"])
    res["ttest_bigger_fixture"] = run([NF, f"{REF}/TtestFilter/test/bigger-counts.tsv.gz"])
    res["ttest_small_fixture"] = run([NF, f"{REF}/TtestFilter/test/small-counts.tsv.gz"])
    names, rows = None, []
    with gzip.open(f"{REF}/computeNF/test/raw-counts.tsv", "rt") as f:
        names = f.readline().split()[1:]
        for line in f:
            rows.append([int(x) for x in line.split()[1:]])
    np.savez_compressed(os.path.join(out, "nf_fixture.npz"), counts=np.array(rows, dtype=np.int32), names=np.array(names))
    for nm in ("bigger", "small"):
        with gzip.open(f"{REF}/TtestFilter/test/{nm}-counts.tsv.gz", "rt") as f:
            hdr = f.readline().split()[1:]
            kmers, rows = [], []
            for line in f:
                t = line.split()
                kmers.append(t[0])
                rows.append([int(x) for x in t[1:]])
        res[f"ttest_{nm}_matrix"] = dict(names=hdr, kmers=kmers, counts=rows)
    res["ttest_conditions"] = open(f"{REF}/TtestFilter/test/sample_conditions_full.tsv").read()
    # (2) the step / early-stop logic (computeNF.c:138-168) needs >= 2M rows: seeded synthetic matrices
    steps = []
    with tempfile.TemporaryDirectory() as d:
        for seed, R, S in ((1, 3_300_000, 4), (2, 2_500_000, 3), (3, 999_999, 5), (4, 4_100_000, 6)):
            cnts = synth_matrix(seed, R, S)
            p = os.path.join(d, "m.tsv")
            with open(p, "w") as f:
                f.write("tag\t" + "\t".join(f"s{j}" for j in range(S)) + "\n")
                np.savetxt(f, np.column_stack([np.arange(R), cnts]), fmt="%d", delimiter="\t")
            r = subprocess.run([NF, p], check=True, capture_output=True, text=True)
            steps.append(dict(seed=seed, R=R, S=S, stdout=r.stdout, stderr=r.stderr))
            print("nf steps", seed, r.stdout.split("\n")[1:3], r.stderr.count("step"))
    res["nf_step_cases"] = steps
    json.dump(res, open(os.path.join(out, "nf_cases.json"), "w"))


def synth_matrix(seed, R, S):
    """seeded count matrix whose median ratio drifts along the rows, so the early stop is exercised
    (also imported by the tests to regenerate the same matrix)"""
    rng = np.random.default_rng(seed)
    depth = rng.uniform(0.5, 2.0, size=S)
    base = rng.geometric(0.05, size=R).astype(np.float64)
    drift = 1.0 + 0.3 * np.sin(np.arange(R) / R * 3.0)[:, None] * np.linspace(-1, 1, S)[None, :]
    lam = base[:, None] * depth[None, :] * drift
    c = rng.poisson(lam).astype(np.int32)
    c[rng.random((R, S)) < 0.02] = 0
    return c


def golden_ttest(out):
    from scipy import special
    rng = np.random.default_rng(7)

    def two_tail(t, df):
        # Boost's branch structure (students_t.hpp cdf): small t -> complement of I_z(1/2, v/2), else I_x(v/2, 1/2).
        # (scipy.stats.t.sf goes through cephes stdtr, which is off by 3e-9 at t=1e-8, v=1, so it is not used.)
        t2 = t * t
        if df > 2 * t2:
            return float(special.betaincc(0.5, df / 2, t2 / (df + t2)))
        return float(special.betainc(df / 2, 0.5, df / (df + t2)))

    grid = []
    for df in (1.0, 1.5, 2.0, 2.7, 3.3, 4.0, 7.9, 14.0, 30.5, 100.0, 1e3, 1e5):
        for t in (0.0, 1e-8, 1e-3, 0.1, 0.5, 1.0, 1.9, 2.5, 4.0, 8.0, 20.0, 60.0, 300.0):
            grid.append((t, df, two_tail(t, df)))
    for _ in range(300):
        t, df = float(np.exp(rng.uniform(-6, 5))), float(1 + np.exp(rng.uniform(-3, 6)))
        grid.append((t, df, two_tail(t, df)))
    json.dump(dict(student_t_2tail=grid, scipy=__import__("scipy").__version__), open(os.path.join(out, "ttest_scipy.json"), "w"))
    print("ttest grid:", len(grid))


if __name__ == "__main__":
    which = sys.argv[1:] or ["join", "nf", "ttest"]
    if "join" in which:
        golden_join(HERE)
    if "nf" in which:
        golden_nf(HERE)
    if "ttest" in which:
        golden_ttest(HERE)

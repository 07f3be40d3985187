#!/usr/bin/env python
"""Golden vectors for the TtestFilter rows (T3-T8) computed WITHOUT the oracle: numpy + scipy only.

TtestFilter cannot be built here (Boost.Math absent), so nothing of the reference pins rows T3-T9. These vectors pin
the oracle AND the CUDA kernels against an implementation neither of them shares code with:
    x = count / NF,  a = x + 1,  l = log(a)                       (TtestFilter.c:379-395)
    meanA / meanB = mean of a per group,  log2FC = log2(meanB / meanA)   (:72-75, 433)
    p = scipy.stats.ttest_ind(l[A], l[B], equal_var=False)        (Welch, :120-150); both group sds 0 -> 0.5 (:128)
    adjusted p = scipy.stats.false_discovery_control(p, method="bh")     (:100-118, 456-476)
Inputs: the bundled 4-sample table of config 1 (tests/golden/config1, made by make_config1.py) with its computeNF
factors as text, and 6000 random rows for 8 vs 8 and 3 vs 5 samples. Run: python tests/golden/make_ttest_independent.py
This file must never import oracle_lib or anything under oracle/."""
import gzip
import os

import numpy as np
from scipy import stats

HERE = os.path.dirname(os.path.abspath(__file__))


def independent(counts, nf, idxA, idxB):
    x = counts.astype(np.float64) / nf[None, :]
    a = x + 1.0
    la = np.log(a)
    A, B = la[:, idxA], la[:, idxB]
    with np.errstate(all="ignore"):
        p = stats.ttest_ind(A, B, axis=1, equal_var=False).pvalue
    # exact-zero test on the sd as the reference computes it: sequential mean, sum of squared deviations
    mA = A.sum(axis=1) / A.shape[1]
    mB = B.sum(axis=1) / B.shape[1]
    flat = (((A - mA[:, None]) ** 2).sum(axis=1) == 0) & (((B - mB[:, None]) ** 2).sum(axis=1) == 0)
    p = np.where(flat, 0.5, p)
    meanA, meanB = a[:, idxA].mean(axis=1), a[:, idxB].mean(axis=1)
    fc = np.log2(meanB / meanA)
    adj = stats.false_discovery_control(p, method="bh")
    return p, meanA, meanB, fc, adj


def config1():
    d = os.path.join(HERE, "config1")
    lines = gzip.open(os.path.join(d, "raw-counts.tsv.gz"), "rt").read().splitlines()
    names = lines[0].split("\t")[1:]
    counts = np.array([[int(v) for v in ln.split("\t")[1:]] for ln in lines[1:]], dtype=np.int32)
    cond = [ln.split("\t") for ln in open(os.path.join(d, "sample_conditions_full.tsv")).read().splitlines()[1:]]
    nf = np.zeros(len(names))
    groups = {}
    for sample, condition, f in cond:                                  # TtestFilter.c:251-284: conditions-file order
        j = names.index(sample)
        nf[j] = float(f)
        groups.setdefault(condition, []).append(j)
    (cA, iA), (cB, iB) = sorted(groups.items())[:2]
    return counts, nf, np.array(iA), np.array(iB), f"{cA} vs {cB}"


def random_case(rng, R, nA, nB):
    S = nA + nB
    base = np.floor(rng.pareto(1.0, size=(R, 1)) * 3 + 1)
    c = rng.poisson(base * rng.uniform(0.5, 2.0, size=(1, S))).astype(np.int64)
    c[rng.random((R, S)) < 0.15] = 0
    c[: R // 20, nA:] *= 7                                             # real effects
    c[R // 20: R // 20 + 50] = c[R // 20: R // 20 + 50, :1]            # rows with identical counts in every sample
    c[-30:] = 0                                                        # all-zero rows: both sds 0 -> p = 0.5
    c[-60:-30, :nA] = 5; c[-60:-30, nA:] = 9                           # both groups constant: sds 0, means differ
    nf = np.array([float("%f" % v) for v in rng.uniform(0.5, 1.9, S)])  # as TtestFilter reads them: 6 decimals
    return np.clip(c, 0, 2 ** 31 - 1).astype(np.int32), nf, np.arange(nA), np.arange(nA, S)


def main():
    out = {}
    c, nf, iA, iB, label = config1()
    p, mA, mB, fc, adj = independent(c, nf, iA, iB)
    out.update(c1_counts=c, c1_nf=nf, c1_idxA=iA, c1_idxB=iB, c1_p=p, c1_meanA=mA, c1_meanB=mB, c1_fc=fc, c1_adj=adj)
    print("config 1:", c.shape, label, "p<=0.05:", int((p <= 0.05).sum()), "adj<=0.05:", int((adj <= 0.05).sum()))
    rng = np.random.default_rng(20261105)
    for tag, nA, nB in (("r88", 8, 8), ("r35", 3, 5)):
        c, nf, iA, iB = random_case(rng, 6000, nA, nB)
        p, mA, mB, fc, adj = independent(c, nf, iA, iB)
        out.update({f"{tag}_counts": c, f"{tag}_nf": nf, f"{tag}_idxA": iA, f"{tag}_idxB": iB, f"{tag}_p": p, f"{tag}_meanA": mA,
                    f"{tag}_meanB": mB, f"{tag}_fc": fc, f"{tag}_adj": adj})
        print(tag, c.shape, "p<=0.05:", int((p <= 0.05).sum()), "adj<=0.05:", int((adj <= 0.05).sum()), "p==0.5:", int((p == 0.5).sum()))
    np.savez_compressed(os.path.join(HERE, "ttest_independent.npz"), **out)


if __name__ == "__main__":
    main()

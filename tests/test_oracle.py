"""CPU tests: pin the oracle (oracle/*.c) against golden vectors captured from the REAL reference
binaries (tests/golden/make_golden.py) and against scipy for the Student-t tail."""
import json
import os

import numpy as np
import pytest

import oracle_lib as orc
from common import GOLD, parse_dump_text, format_join, load_json


def test_join_matches_reference_binary_goldens():
    for case in load_json("join_cases.json"):
        r, a = 1, 1
        args = case["args"]
        for i in range(0, len(args), 2):
            if args[i] == "-r":
                r = int(args[i + 1])
            if args[i] == "-a":
                a = int(args[i + 1])
        dumps = [parse_dump_text(t) for t in case["inputs"]]
        k = dumps[0][2]
        keys, cnts, U = orc.join([d[0] for d in dumps], [d[1] for d in dumps], r, a, stale_tail=True)
        assert format_join(keys, cnts, k) == case["stdout"], f"args={args}"


def test_join_stale_tail_only_when_r_nonpositive():
    k = [np.array([1, 5, 9], np.uint64), np.array([5, 7], np.uint64)]
    c = [np.array([3, 4, 5], np.int32), np.array([6, 7], np.int32)]
    keys, cnts, U = orc.join(k, c, 0, 1, stale_tail=True)
    assert U == 4 and len(keys) == 5 and keys[-1] == 9 and (cnts[-1] == 0).all()
    keys, cnts, U = orc.join(k, c, 1, 1, stale_tail=True)
    assert len(keys) == 4


def test_join_duplicate_keys_emit_two_rows_like_reference():
    # duplicate k-mer inside one file: the reference prints it twice (SURVEY §8a J-edge)
    keys, cnts, U = orc.join([np.array([4, 4, 8], np.uint64), np.array([4], np.uint64)],
                             [np.array([1, 2, 3], np.int32), np.array([9], np.int32)])
    assert keys.tolist() == [4, 4, 8] and cnts.tolist() == [[1, 9], [2, 0], [3, 0]]


def _nf_text(names, nf):
    return "sample\tnormalization_factor\n" + "".join("%s\t%f\n" % (n, v) for n, v in zip(names, nf))


def test_nf_reference_fixture_goldens():
    g = load_json("nf_cases.json")
    z = np.load(os.path.join(GOLD, "nf_fixture.npz"))
    nf, used, uz = orc.nf_reference(z["counts"])
    assert _nf_text(z["names"], nf) == g["computeNF_fixture"]
    assert used == len(z["counts"]) and uz == 0
    for nm in ("bigger", "small"):
        m = g[f"ttest_{nm}_matrix"]
        nf, _, _ = orc.nf_reference(np.array(m["counts"], np.int32))
        assert _nf_text(m["names"], nf) == g[f"ttest_{nm}_fixture"]


@pytest.mark.parametrize("idx", [0, 1, 2, 3])
def test_nf_reference_step_logic_matches_binary(idx):
    from golden.make_golden import synth_matrix
    case = load_json("nf_cases.json")["nf_step_cases"][idx]
    cnts = synth_matrix(case["seed"], case["R"], case["S"])
    nf, used, uz = orc.nf_reference(cnts)
    if uz:
        pytest.skip("reference prints uninitialised memory here")
    assert _nf_text([f"s{j}" for j in range(case["S"])], nf) == case["stdout"]


def test_nf_small_step_equals_full_prefix():
    rng = np.random.default_rng(3)
    cnts = rng.poisson(20, size=(5000, 4)).astype(np.int32)
    nf_all, used, _ = orc.nf_reference(cnts, step=10 ** 6)
    assert used == 5000
    np.testing.assert_allclose(nf_all, orc.nf_full(cnts), rtol=0, atol=0)
    # with a tiny step the early stop consumes a prefix only
    nf_s, used_s, _ = orc.nf_reference(cnts, step=1000)
    assert used_s % 1000 == 0 and used_s <= 5000
    np.testing.assert_allclose(nf_s, orc.nf_full(cnts[:used_s]), rtol=0, atol=0)


def test_student_t_tail_vs_scipy():
    g = load_json("ttest_scipy.json")["student_t_2tail"]
    for t, df, want in g:
        got = orc.student_t_2tail(t, df)
        if want < 1e-290:
            assert got < 1e-280
        else:
            assert abs(got - want) <= 1e-11 * want + 1e-300, (t, df, got, want)


def test_student_t_tail_vs_boost_math_inside_scipy():
    """scipy's special functions are compiled against Boost.Math (its _ufuncs_cxx module carries boost::math::ibeta and
    boost::math::students_t_distribution), the library TtestFilter.c:146-147 calls. The reference computes
    2*cdf(complement(students_t(df), |t|)); for Boost that is ibeta(df/2, 1/2, df/(df+t^2)) when df > 2 t^2 is false, and its
    complement form otherwise: both equal the regularised incomplete beta checked here."""
    sp = pytest.importorskip("scipy.special")
    st = pytest.importorskip("scipy.stats")
    rng = np.random.default_rng(11)
    df = np.concatenate([rng.uniform(1.0, 40.0, 400), rng.uniform(40, 2000, 100), [1.0, 2.0, 14.0]])
    t = np.concatenate([np.abs(rng.standard_normal(400)) * 3, rng.uniform(0, 60, 100), [0.0, 1e-8, 37.5]])
    for ti, di in zip(t, df):
        got = orc.student_t_2tail(float(ti), float(di))
        y = ti * ti / (di + ti * ti)                      # the small argument is formed without cancellation
        w1 = float(sp.betainc(di / 2, 0.5, di / (di + ti * ti))) if y > 0.01 else float(1 - sp.betainc(0.5, di / 2, y))
        w2 = float(2 * st.t.sf(ti, di))
        assert abs(got - w1) <= 1e-11 * w1 + 1e-300, (ti, di, got, w1)
        assert abs(got - w2) <= 1e-10 * w2 + 1e-300, (ti, di, got, w2)


# SURVEY §4 table (scipy-derived pins for the reference's TtestFilter fixtures, -p 0.2)
SURVEY_ROWS = {
    "bigger": {(16, 16, 16, 18, 18, 18): (0.3767727982505043, 0.3767727982505043, 19.501368, 15.843223, -0.299709),
               (0, 0, 0, 16, 16, 16): (0.00073631188614349617, 0.069949629183632137, 1.000000, 14.193976, 3.827207),
               (16, 16, 16, 2, 2, 2): (0.00392433950660651, 0.18640612656380923, 19.501368, 2.649247, -2.879921)},
    "small": {(6, 6, 6, 8, 8, 8): (0.88865603566518181, 0.88865603566518181, 7.938013, 7.596988, -0.063350),
              (0, 0, 0, 6, 6, 6): (0.0012933699965006257, 0.12287014966755944, 1.000000, 5.947741, 2.572342),
              (6, 6, 6, 2, 2, 2): (0.012419775562328221, 0.5899393392105905, 7.938013, 2.649247, -1.583196)},
}


@pytest.mark.parametrize("nm", ["bigger", "small"])
def test_ttest_fixture_rows_match_survey_table(nm):
    g = load_json("nf_cases.json")
    m = g[f"ttest_{nm}_matrix"]
    cnts = np.array(m["counts"], np.int32)
    nf = np.array([float(l.split()[2]) for l in g["ttest_conditions"].strip().split("\n")[1:]])
    p, mA, mB, fc = orc.ttest_rows(cnts, nf, [0, 1, 2], [3, 4, 5])
    adj = orc.bh(p)
    seen = set()
    for i, row in enumerate(cnts):
        want = SURVEY_ROWS[nm][tuple(row.tolist())]
        seen.add(tuple(row.tolist()))
        assert abs(p[i] - want[0]) <= 1e-11 * want[0]
        assert abs(adj[i] - want[1]) <= 1e-11 * want[1]
        assert "%f" % mA[i] == "%f" % want[2] and "%f" % mB[i] == "%f" % want[3] and "%f" % fc[i] == "%f" % want[4]
    assert len(seen) == 3


def test_bh_definition_and_ties():
    rng = np.random.default_rng(11)
    p = np.round(rng.random(500), 2)          # many ties
    adj = orc.bh(p)
    n = len(p)
    order = np.argsort(-p, kind="stable")
    want = np.empty(n)
    mn = 1.0
    for i, j in enumerate(order):
        mn = min(mn, n / (n - i) * p[j])
        want[j] = min(1.0, mn)
    np.testing.assert_allclose(adj, want, rtol=1e-15)
    for v in np.unique(p):
        assert len(set(adj[p == v].tolist())) == 1


def test_poisson_lr_vs_scipy():
    """-l mode (TtestFilter.c:152-185): log-density at non-integer k, LR statistic, chi-squared(1) upper tail"""
    from scipy import special, stats
    rng = np.random.default_rng(3)
    for _ in range(300):
        K1, K2 = 1 + rng.gamma(2.0, 20.0), 1 + rng.gamma(2.0, 20.0)
        N1, N2 = int(rng.integers(10**6, 10**9)), int(rng.integers(10**6, 10**9))
        lp = lambda lam, k: -lam + k * np.log(lam) - special.gammaln(k + 1)
        th = (K1 + K2) / (N1 + N2)
        stat = 2 * ((lp(K1, K1) + lp(K2, K2)) - (lp(th * N1, K1) + lp(th * N2, K2)))
        got = orc.poisson_lr(K1, K2, N1, N2)
        want = stats.chi2.sf(max(stat, 0.0), 1)
        assert abs(got - want) <= 1e-13 + 1e-9 * want * (want > 1e-6), (K1, K2, N1, N2, got, want)
    # the reference forms 1 - cdf: tiny p-values bottom out at the double spacing near 1
    assert orc.poisson_lr(5.0, 300.0, 10**8, 10**8) in (0.0, 1.1102230246251565e-16, 2.220446049250313e-16)
    # density underflow: logpoisson returns the reference's constant 0.001 (TtestFilter.c:157-161), the statistic turns
    # negative, Boost's chi-squared cdf throws and the reference aborts; the restatement (and the product) give p = 1
    assert orc.poisson_lr(5.0, 5000.0, 10**8, 10**8) == 1.0


# ---- TtestFilter rows against vectors that share no code with the oracle (tests/golden/make_ttest_independent.py) ----
INDEP = np.load(os.path.join(GOLD, "ttest_independent.npz"))


@pytest.mark.parametrize("tag", ["c1", "r88", "r35"])
def test_ttest_rows_and_bh_vs_scipy_only_goldens(tag):
    """p (Welch on log(x/NF + 1), scipy.stats.ttest_ind), means, log2FC and BH (scipy false_discovery_control) computed
    without the oracle: every config-1 row and 2 x 6000 random rows, including both-sds-zero rows (p = 0.5)"""
    c, nf, iA, iB = INDEP[f"{tag}_counts"], INDEP[f"{tag}_nf"], INDEP[f"{tag}_idxA"].tolist(), INDEP[f"{tag}_idxB"].tolist()
    p, mA, mB, fc = orc.ttest_rows(c, nf, iA, iB)
    np.testing.assert_allclose(p, INDEP[f"{tag}_p"], rtol=1e-9, atol=1e-300)
    np.testing.assert_allclose(mA, INDEP[f"{tag}_meanA"], rtol=1e-13)
    np.testing.assert_allclose(mB, INDEP[f"{tag}_meanB"], rtol=1e-13)
    np.testing.assert_allclose(fc, INDEP[f"{tag}_fc"], rtol=1e-12, atol=1e-14)
    assert int((INDEP[f"{tag}_p"] == 0.5).sum()) == int((p == 0.5).sum())
    np.testing.assert_allclose(orc.bh(p), INDEP[f"{tag}_adj"], rtol=1e-9, atol=1e-300)
    np.testing.assert_allclose(orc.bh(INDEP[f"{tag}_p"]), INDEP[f"{tag}_adj"], rtol=1e-14)      # BH alone, same input


def test_student_t_tail_two_evaluation_paths_agree():
    """orc_student_t_2tail (always I_{v/(v+t^2)}(v/2, 1/2), x and 1-x both exact) against the branch structure of
    Boost.Math's students_t complement (v > 2t^2: ibetac(1/2, v/2, t^2/(v+t^2)), else ibeta(v/2, 1/2, v/(v+t^2)), each
    routine given its z alone): <= 1e-12 relative for every df a Welch test can produce here (df <= nA + nB - 2 <= 254),
    and against the closed form for df = 1 (Cauchy) and df = 2"""
    worst = 0.0
    for df in np.concatenate([np.linspace(1, 30, 60), np.logspace(1.5, np.log10(2000), 40), [1.0001, 2, 2.5, 198, 254]]):
        for t in np.concatenate([np.logspace(-6, 4, 60), [0.0, 0.5, 1, 2, 3.3]]):
            a, b = orc.student_t_2tail(t, df), orc.student_t_2tail_boost_branches(t, df)
            if a > 1e-300:
                worst = max(worst, abs(a - b) / a)
    assert worst <= 1e-12, worst
    for t in (1e-6, 0.3, 1.0, 7.5, 1e3):
        cauchy = 2 / np.pi * np.arctan(1 / t)                       # = 1 - (2/pi) atan(t), without the cancellation
        np.testing.assert_allclose(orc.student_t_2tail(t, 1.0), cauchy, rtol=1e-13)
        np.testing.assert_allclose(orc.student_t_2tail_boost_branches(t, 1.0), cauchy, rtol=1e-13)
        np.testing.assert_allclose(orc.student_t_2tail(t, 2.0), 2 / (np.sqrt(2 + t * t) * (np.sqrt(2 + t * t) + t)), rtol=1e-13)


def test_gamma_ratio_series_of_the_row_kernel():
    """k_ttest_rows8 (dk_ttest.cu) takes Gamma(v+1/2)/Gamma(v) from an upward shift to w >= 8 and a 6-term asymptotic
    series; this is the same arithmetic in numpy against 40-digit mpmath over the v = df/2 range of groups of 2..8 samples
    and beyond (exp(gammaln - gammaln) cancels to 3e-13 at v ~ 200 and cannot referee). The kernel's p-values are held
    to 1e-9 relative: the ratio must be far inside that."""
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 40
    rng = np.random.default_rng(5)
    v = np.concatenate([rng.uniform(0.5, 8.0, 3000), rng.uniform(8.0, 200.0, 1000)])
    num, dg, w = np.ones_like(v), np.ones_like(v), v.copy()
    for _ in range(8):
        m = w < 8.0
        num[m] *= w[m]
        dg[m] *= w[m] + 0.5
        w[m] += 1.0
    assert (w >= 8.0).all()
    r = 1.0 / w
    rr = r * r
    ser = r * (-1 / 8 + rr * (1 / 192 + rr * (-1 / 640 + rr * (17 / 14336 + rr * (-31 / 18432 + rr * (691 / 180224))))))
    got = np.sqrt(w) * np.exp(ser) * num / dg
    want = np.array([float(mp.gamma(mp.mpf(float(x)) + 0.5) / mp.gamma(mp.mpf(float(x)))) for x in v])
    assert np.max(np.abs(got / want - 1)) < 5e-14

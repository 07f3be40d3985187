"""The three drop-in CLIs (dekupl-run_b200/host/bin/{joinCounts,computeNF,TtestFilter}): same argv,
stdout bytes and exit codes as the reference tools."""
import gzip
import os
import subprocess

import numpy as np
import pytest

from common import GOLD, load_json

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "dekupl-run_b200", "host", "bin")


def run(tool, *args, cwd=None, env=None):
    return subprocess.run([os.path.join(BIN, tool), *args], capture_output=True, text=True, cwd=cwd,
                          env=None if env is None else {**os.environ, **env})


# ------------------------------------------------------------------ CPU: argv surface ---
def test_usage_and_exit_codes_without_arguments():
    r = run("joinCounts")
    assert r.returncode == 1 and "Usage:   joinCounts [options] <counts.tsv> [<counts.tsv> ...]" in r.stderr
    assert "Options: -r INT    min recurrence [1]" in r.stderr and r.stdout == ""
    r = run("joinCounts", "-r", "3", "-a", "7")
    assert r.returncode == 1 and "min recurrence [3]" in r.stderr and "abundance [7]" in r.stderr
    r = run("computeNF")
    assert r.returncode == 1 and "Usage:   computeNF [options] <counts.tsv>" in r.stderr
    r = run("TtestFilter", "a", "b", "c")
    assert r.returncode == 1 and "Usage:   ttestFilter [options] <counts.tsv> <conditions.tsv> <conditionA> <conditionB>" in r.stderr
    assert "max pvalue [0.05]" in r.stderr


def test_missing_input_file_fails_like_the_reference():
    for tool, args in (("joinCounts", ["/nonexistent/x.txt"]), ("computeNF", ["/nonexistent/x.tsv"]),
                       ("TtestFilter", ["/nonexistent/x.tsv", "/nonexistent/c.tsv", "A", "B"])):
        r = run(tool, *args)
        assert r.returncode != 0 and "Failed to open /nonexistent/x" in r.stderr


def test_diffFilter_argv_surface(tmp_path):
    r = run("diffFilter")
    assert r.returncode == 1 and "Usage:   diffFilter" in r.stderr and r.stdout == ""
    r = run("diffFilter", "/nonexistent/ref.txt.gz", "/nonexistent/x.tsv")
    assert r.returncode != 0 and "Cannot open /nonexistent/ref.txt.gz" in r.stderr       # diffFilter.pl:35,37
    ref, tab = tmp_path / "ref.txt", tmp_path / "tab.tsv"
    ref.write_text("ACGT 1\n")
    tab.write_text("tag\ts1\nACGTN\t3\n")
    r = run("diffFilter", str(ref), str(tab))
    assert r.returncode != 0 and "at most 31 ACGT bases" in r.stderr and r.stdout == ""


def test_table_parser_threads_report_the_same_ragged_line(tmp_path):
    """the parallel parser falls back to the sequential reader whenever a part does not hold one row per line"""
    rows = ["K%d\t%d\t%d" % (i, i, 2 * i) for i in range(400)]
    rows[257] = "K257\t5"                                                 # ragged: swallows the next line's tag
    bad = tmp_path / "bad.tsv"
    bad.write_text("tag\ta\tb\n" + "\n".join(rows) + "\n")
    msgs = set()
    for nt in ("1", "2", "5", "32"):
        r = run("computeNF", str(bad), env={"DK_PARSE_THREADS": nt})
        assert r.returncode != 0
        msgs.add(r.stderr.strip())
    assert msgs == {"inconsistent number of column (line 258)"}, msgs


def test_poisson_flag_is_accepted_in_usage():
    r = run("TtestFilter", "-l")
    assert r.returncode == 1 and "-l        disregard the program name" in r.stderr


def test_block_streamed_tokeniser_matches_the_in_memory_one(tmp_path):
    """joinCounts reads each dump through DkStreamCursor (dk_text.h), a few MB at a time; test_text replays the parser's
    loop over the same file with the whole-file cursor and with buffers down to 16 bytes and compares the record streams."""
    cases = {
        "plain.txt": b"ACGT 5 extra tokens here\nCCGT\t7\nGGGT 9\r\nTTTT   \n",
        "no_newline_at_end.txt": b"AAAA 3\nCCCC 4",
        "empty.txt": b"",
        "long_first_line.txt": b"ACGT 1 " + b"x" * 5000 + b"\nCCCC 2\n",
        "blank_lines.txt": b"ACGT 1\n\n\nCCCC 2\n",
    }
    for nm, data in cases.items():
        (tmp_path / nm).write_bytes(data)
    (tmp_path / "two_members.gz").write_bytes(gzip.compress(cases["plain.txt"]) + gzip.compress(cases["no_newline_at_end.txt"]))
    files = [str(tmp_path / nm) for nm in list(cases) + ["two_members.gz"]] + [os.path.join(GOLD, "config1", "sample1.txt.gz")]
    for f in files:
        outs = set()
        for b in (16, 17, 61, 4096, 8 << 20):
            r = subprocess.run([os.path.join(BIN, "test_text"), f, str(b)], capture_output=True)
            assert r.returncode == 0, (f, b)
            outs.add(r.stdout)
        assert len(outs) == 1, f
    with gzip.open(files[-1], "rt") as fh:
        n = sum(1 for _ in fh)
    assert r.stdout.count(b"\n") == n


@pytest.mark.gpu
def test_joinCounts_cli_matches_reference_binary_bytes(tmp_path):
    for ci, case in enumerate(load_json("join_cases.json")):
        files = []
        for i, txt in enumerate(case["inputs"]):
            p = tmp_path / f"c{ci}_s{i}.txt"
            if i % 2:                       # mix plain and gz inputs (gzopen is transparent)
                p = tmp_path / f"c{ci}_s{i}.txt.gz"
                with gzip.open(p, "wt") as f:
                    f.write(txt)
            else:
                p.write_text(txt)
            files.append(str(p))
        r = run("joinCounts", *case["args"], *files)
        assert r.returncode == 0, r.stderr
        assert r.stdout == case["stdout"], case["args"]


@pytest.mark.gpu
def test_computeNF_cli_matches_reference_binary_bytes(tmp_path):
    g = load_json("nf_cases.json")
    z = np.load(os.path.join(GOLD, "nf_fixture.npz"))
    p = tmp_path / "raw-counts.tsv.gz"
    body = "".join("K%d\t" % i + "\t".join(map(str, row)) + "\n" for i, row in enumerate(z["counts"].tolist()))
    with open(p, "wb") as f:                 # two gzip members: header, then body (Snakefile:594-596)
        f.write(gzip.compress(("tag\t" + "\t".join(z["names"].tolist()) + "\n").encode()))
        f.write(gzip.compress(body.encode()))
    r = run("computeNF", str(p))
    assert r.returncode == 0, r.stderr
    assert r.stdout == g["computeNF_fixture"]
    # ragged row => the reference's message and a failure exit
    bad = tmp_path / "bad.tsv"
    bad.write_text("tag\ta\tb\nK1\t1\t2\nK2\t3\nK3\t4\t5\n")
    r = run("computeNF", str(bad))
    assert r.returncode != 0 and "inconsistent number of column (line 2)" in r.stderr
    # the multi-threaded table parser (forced on this small file) gives the same bytes as the sequential one
    for nt in ("1", "3", "16"):
        r = run("computeNF", str(p), env={"DK_PARSE_THREADS": nt})
        assert r.returncode == 0 and r.stdout == g["computeNF_fixture"], nt


@pytest.mark.gpu
@pytest.mark.parametrize("nm", ["bigger", "small"])
def test_TtestFilter_cli_fixture(tmp_path, nm):
    import oracle_lib as orc
    g = load_json("nf_cases.json")
    mm = g[f"ttest_{nm}_matrix"]
    cnts = np.array(mm["counts"], np.int32)
    counts = tmp_path / "counts.tsv.gz"
    with gzip.open(counts, "wt") as f:
        f.write("tag\t" + "\t".join(mm["names"]) + "\n")
        for k, row in zip(mm["kmers"], mm["counts"]):
            f.write(k + "\t" + "\t".join(map(str, row)) + "\n")
    conds = tmp_path / "conds.tsv"
    conds.write_text(g["ttest_conditions"])
    raw = tmp_path / "raw_pvals.txt.gz"
    r = run("TtestFilter", "-p", "0.2", "-r", str(raw), str(counts), str(conds), "A", "B", cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert "Group A: sampleA1, sampleA2, sampleA3" in r.stderr and "Adjusting pvalues with Benjamini" in r.stderr
    assert not os.path.exists(tmp_path / ".TtestFilter.tmp")
    nf = np.array([float(l.split()[2]) for l in g["ttest_conditions"].strip().split("\n")[1:]])
    p, mA, mB, fc = orc.ttest_rows(cnts, nf, [0, 1, 2], [3, 4, 5])
    adj = orc.bh(p)
    want = "tag\tpvalue\tmeanA\tmeanB\tlog2FC\t" + "\t".join(mm["names"]) + "\n"
    for i in range(len(cnts)):
        if adj[i] <= 0.2 and abs(fc[i]) >= 0:
            want += "%s\t%f\t%f\t%f\t%f" % (mm["kmers"][i], adj[i], mA[i], mB[i], fc[i])
            want += "".join("\t%.2f" % (cnts[i, j] / nf[j]) for j in range(6)) + "\n"
    assert r.stdout == want
    assert gzip.open(raw, "rt").read() == "".join("%s\t%f\n" % (k, v) for k, v in zip(mm["kmers"], p))


@pytest.mark.gpu
def test_TtestFilter_cli_poisson_mode(tmp_path):
    """-l: Poisson likelihood ratio (TtestFilter.c:152-185); survivors print the RAW counts"""
    import oracle_lib as orc
    g = load_json("nf_cases.json")
    mm = g["ttest_bigger_matrix"]
    cnts = np.array(mm["counts"], np.int32)
    counts = tmp_path / "counts.tsv.gz"
    with gzip.open(counts, "wt") as f:
        f.write("tag\t" + "\t".join(mm["names"]) + "\n")
        for k, row in zip(mm["kmers"], mm["counts"]):
            f.write(k + "\t" + "\t".join(map(str, row)) + "\n")
    conds = tmp_path / "conds.tsv"
    conds.write_text(g["ttest_conditions"])
    raw = tmp_path / "raw_pvals.txt.gz"
    r = run("TtestFilter", "-l", "-p", "0.5", "-r", str(raw), str(counts), str(conds), "A", "B", cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    p, mA, mB, fc, _ = orc.poisson_rows(cnts, [0, 1, 2], [3, 4, 5])
    adj = orc.bh(p)
    want = "tag\tpvalue\tmeanA\tmeanB\tlog2FC\t" + "\t".join(mm["names"]) + "\n"
    for i in range(len(cnts)):
        if adj[i] <= 0.5:
            want += "%s\t%f\t%f\t%f\t%f" % (mm["kmers"][i], adj[i], mA[i], mB[i], fc[i])
            want += "".join("\t%.2f" % float(cnts[i, j]) for j in range(6)) + "\n"
    assert want.count("\n") > 1, "fixture must have survivors"
    assert r.stdout == want
    assert gzip.open(raw, "rt").read() == "".join("%s\t%f\n" % (k, v) for k, v in zip(mm["kmers"], p))

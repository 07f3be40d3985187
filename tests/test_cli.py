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


def test_block_streamed_table_reader_matches_the_whole_file_one(tmp_path):
    """computeNF and TtestFilter read the count table through DkMatrixText::load (dk_matrix_text.h), a block at a time;
    test_matrix_text reads the same file with the whole-file reader too and compares names, rows, counts, tags and the
    reported inconsistent line — for blocks down to 64 bytes, one to seven parser threads, and the irregular inputs that
    must fall back to the token-by-token reader (a short row in the middle, a row split over two lines, CR LF)."""
    rng = np.random.default_rng(3)

    def table(n, S=6, trail=True):
        lines = ["tag\t" + "\t".join("s%d" % i for i in range(S))]
        for _ in range(n):
            lines.append("".join(rng.choice(list("ACGT"), 31)) + "\t" + "\t".join(str(int(v)) for v in rng.integers(0, 5000, S)))
        return ("\n".join(lines) + ("\n" if trail else "")).encode()

    cases = {
        "regular.tsv": (table(3000), 3000, 0),
        "no_trailing_newline.tsv": (table(777, trail=False), 777, 0),
        "header_only.tsv": (b"tag\ta\tb\n", 0, 0),
        "header_no_newline.tsv": (b"tag\ta\tb", 0, 0),
        "empty.tsv": (b"", 0, 0),
        "one_row.tsv": (b"tag\ta\tb\nACGT\t1\t2\n", 1, 0),
        "short_last_row.tsv": (b"tag\ta\tb\tc\nACGT\t1\t2\t3\nCCGT\t4", 2, 0),          # padded with zeros at EOF (computeNF.c:124-131)
        "short_last_row_nl.tsv": (b"tag\ta\tb\tc\nACGT\t1\t2\t3\nCCGT\t4\n", 2, 0),
        "blank_line_end.tsv": (table(100) + b"\n", 100, 0),
        "space_separated.tsv": (table(300).replace(b"\t", b" "), 300, 0),
        "long_lines.tsv": (table(40, S=200), 40, 0),
    }
    t = table(2000).split(b"\n")
    t[1200] = t[1200].rsplit(b"\t", 1)[0]
    cases["short_row_in_the_middle.tsv"] = (b"\n".join(t), None, 1200)
    t = table(2000).split(b"\n")
    t[37] += b"\t9"
    cases["long_row.tsv"] = (b"\n".join(t), None, 37)
    t = table(500).split(b"\n")
    a = t[250].split(b"\t")
    t[250] = b"\t".join(a[:4]) + b"\n" + b"\t".join(a[4:])        # the reference reads S tokens whatever the line ends: one row
    cases["row_over_two_lines.tsv"] = (b"\n".join(t), 500, 0)
    for nm, (data, _, _) in cases.items():
        (tmp_path / nm).write_bytes(data)
    reg = cases["regular.tsv"][0]
    (tmp_path / "two_members.gz").write_bytes(gzip.compress(reg[:100000]) + gzip.compress(reg[100000:]))
    cases["two_members.gz"] = (None, 3000, 0)
    for nm, (_, rows, bad) in cases.items():
        outs = set()
        for block in (64, 100, 333, 4096, 64 << 20):
            for threads in ("1", "4", "7"):
                r = subprocess.run([os.path.join(BIN, "test_matrix_text"), str(tmp_path / nm), str(block)], capture_output=True,
                                   text=True, env={**os.environ, "DK_PARSE_THREADS": threads})
                assert r.returncode == 0, (nm, block, threads, r.stdout)
                outs.add(r.stdout)
        assert len(outs) == 1, (nm, outs)
        S, R, b, _ = outs.pop().split()
        assert int(b) == bad, nm
        if rows is not None:
            assert int(R) == rows, nm


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

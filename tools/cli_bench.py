#!/usr/bin/env python
"""Process-level timing of the drop-in CLIs against the unmodified reference binaries (oracle/_ref) on the same
gz text dumps: what a Snakefile rule sees (inflate + parse + compute + format + write). Not the headline metric
(bench.py); it documents where the time goes once the compute is on the GPU.
usage: python tools/cli_bench.py [S] [records_per_sample]"""
import gzip
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
BIN = os.path.join(ROOT, "dekupl-run_b200", "host", "bin")
REF = os.path.join(ROOT, "oracle", "_ref")


def timed(cmd, out):
    t0 = time.time()
    with open(out, "wb") as f:
        r = subprocess.run(cmd, stdout=f, stderr=subprocess.PIPE)
    dt = time.time() - t0
    assert r.returncode == 0, r.stderr.decode()[-400:]
    return dt


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
    rng = np.random.default_rng(3)
    union = np.unique(rng.integers(0, 1 << 62, int(n * 2.2), dtype=np.uint64))
    lut = np.array(list(b"ACGT"), np.uint8)
    shifts = np.arange(30, -1, -1, dtype=np.uint64) * np.uint64(2)
    d = tempfile.mkdtemp(prefix="dkcli_")
    files = []
    t0 = time.time()
    for s in range(S):
        keys = union[rng.random(len(union)) < n / len(union)]
        cnt = rng.zipf(2.0, len(keys)).clip(1, 10 ** 6) + 1
        txt = lut[((keys[:, None] >> shifts[None, :]) & np.uint64(3)).astype(np.intp)]          # (n, 31) bytes
        lines = np.concatenate([txt, np.full((len(keys), 1), ord(" "), np.uint8)], axis=1)
        body = b"".join(l.tobytes() + str(int(c)).encode() + b"\n" for l, c in zip(lines, cnt))
        p = os.path.join(d, f"s{s}.txt.gz")
        with gzip.open(p, "wb", compresslevel=1) as f:
            f.write(body)
        files.append(p)
    print(f"# {S} dumps x ~{n} records generated in {time.time() - t0:.1f}s under {d}")
    res = {}
    res["joinCounts ours"] = timed([os.path.join(BIN, "joinCounts"), "-r", "2", "-a", "3"] + files, os.path.join(d, "ours.tsv"))
    res["joinCounts ours (warm)"] = timed([os.path.join(BIN, "joinCounts"), "-r", "2", "-a", "3"] + files, os.path.join(d, "ours.tsv"))
    res["joinCounts reference"] = timed([os.path.join(REF, "joinCounts"), "-r", "2", "-a", "3"] + files, os.path.join(d, "ref.tsv"))
    same = open(os.path.join(d, "ours.tsv"), "rb").read() == open(os.path.join(d, "ref.tsv"), "rb").read()
    rows = sum(1 for _ in open(os.path.join(d, "ref.tsv"), "rb"))
    raw = os.path.join(d, "raw-counts.tsv.gz")
    with open(raw, "wb") as f:
        f.write(gzip.compress(("tag\t" + "\t".join(f"s{i}" for i in range(S)) + "\n").encode()))
        f.write(gzip.compress(open(os.path.join(d, "ref.tsv"), "rb").read(), compresslevel=1))
    tiny = os.path.join(d, "tiny.tsv")
    open(tiny, "w").write("tag\ta\tb\nACGT\t1\t2\nACGG\t3\t4\n")
    res["computeNF ours, 2-row table (start-up floor)"] = timed([os.path.join(BIN, "computeNF"), tiny], os.path.join(d, "nf_tiny.tsv"))
    res["computeNF ours"] = timed([os.path.join(BIN, "computeNF"), raw], os.path.join(d, "nf_ours.tsv"))
    res["computeNF ours (1 parse thread)"] = None
    os.environ["DK_PARSE_THREADS"] = "1"
    res["computeNF ours (1 parse thread)"] = timed([os.path.join(BIN, "computeNF"), raw], os.path.join(d, "nf_ours1.tsv"))
    del os.environ["DK_PARSE_THREADS"]
    res["computeNF reference"] = timed([os.path.join(REF, "computeNF"), raw], os.path.join(d, "nf_ref.tsv"))
    same_nf = open(os.path.join(d, "nf_ours.tsv")).read() == open(os.path.join(d, "nf_ref.tsv")).read()
    # mergeTags: the k-mers of a random 400 kb sequence (70% kept), random p-values with ties
    import random
    rnd = random.Random(1)
    g = "".join(rnd.choice("ACGT") for _ in range(400_000))
    kms = [g[i:i + 31] for i in range(len(g) - 30) if rnd.random() < 0.7]
    tab = os.path.join(d, "diff-counts.tsv.gz")
    with gzip.open(tab, "wt", compresslevel=1) as f:
        f.write("tag\tpvalue\tmeanA\tmeanB\tlog2FC\ts1\ts2\n")
        for km in kms:
            f.write("%s\t%f\t1.000000\t2.000000\t%f\t1.00\t2.00\n" % (km, rnd.choice([0.0, 0.01, round(rnd.random() * 0.05, 6)]),
                                                                   rnd.choice([1.0, 1.0, -1.0])))
    res[f"mergeTags ours ({len(kms)} k-mers)"] = timed([os.path.join(BIN, "mergeTags"), "-k", "31", tab], os.path.join(d, "mt_ours.tsv"))
    res["mergeTags reference"] = timed([os.path.join(REF, "mergeTags"), "-k", "31", tab], os.path.join(d, "mt_ref.tsv"))
    same_mt = open(os.path.join(d, "mt_ours.tsv")).read() == open(os.path.join(d, "mt_ref.tsv")).read()
    print(f"# rows {rows}; joinCounts bytes identical: {same}; computeNF bytes identical: {same_nf}; mergeTags bytes identical: {same_mt}")
    for k, v in res.items():
        print(f"{k:48s} {v:8.2f} s")


if __name__ == "__main__":
    main()

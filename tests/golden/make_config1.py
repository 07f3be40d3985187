#!/usr/bin/env python
"""Golden fixtures for BASELINE.json configs[0] (tests/config_ttest.json on the bundled data/sample1-4:
2 vs 2, k=31, lib_type rf, min_recurrence 2, min_recurrence_abundance 5, pvalue_threshold 0.2).

jellyfish, snakemake and Rscript are absent from the build image, so the three upstream steps are restated
here exactly (integer work, test harness only):
  * revCompFastq.pl (bin/revCompFastq.pl:33-44) on mate 1 (lib_type rf, Snakefile:540-542);
  * `jellyfish count -L 2 -m 31` (Snakefile:523-552): every 31-mer of both mates, no canonicalisation,
    windows containing a non-ACGT base skipped; `jellyfish dump -c` keeps count >= 2 (-L 2);
  * `sort -k1` (Snakefile:572-582).
Then the UNMODIFIED reference joinCounts and computeNF binaries (oracle/_ref) produce raw-counts and the
normalisation factors, and the oracle restatement of TtestFilter (Boost is absent) produces diff-counts.
To keep the committed fixture small only the first NPAIRS read pairs of every sample are counted; the k-mer
distribution is the real one (clustered keys, shared prefixes), which is what this fixture is for.
The mask (default RNA-Seq mode, Snakefile:603-644) is the sorted `jellyfish count -m 31` dump (no -L, no -C for
lib_type rf, Snakefile:608-615) of tests/gencode.v24.transcripts.head1000.fa, restated the same way; the UNMODIFIED
bin/diffFilter.pl (perl is present) then produces masked-counts, and the TtestFilter restatement run on the masked
matrix with the raw matrix's NFs (Snakefile:366,654) gives diff-counts-masked.
Run in the build container:  make -C oracle && python tests/golden/make_config1.py"""
import gzip
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
DATA = "/root/reference/data"
OUT = os.path.join(HERE, "config1")
NPAIRS = int(os.environ.get("NPAIRS", "4000"))
K = 31
COMP = str.maketrans("ACGTacgt", "TGCAtgca")


def reads(path, n, revcomp):
    with gzip.open(path, "rt") as f:
        for i, line in enumerate(f):
            if i // 4 >= n:
                break
            if i % 4 == 1:
                s = line.rstrip("\n")
                yield s.translate(COMP)[::-1] if revcomp else s


def count_sample(name):
    cnt = {}
    for path, rc in ((f"{DATA}/{name}_1.fastq.gz", True), (f"{DATA}/{name}_2.fastq.gz", False)):
        for s in reads(path, NPAIRS, rc):
            s = s.upper()
            for i in range(len(s) - K + 1):
                w = s[i:i + K]
                if w.strip("ACGT"):
                    continue
                cnt[w] = cnt.get(w, 0) + 1
    return sorted((k, c) for k, c in cnt.items() if c >= 2)


def count_transcripts(path):
    seqs, cur = [], []
    for l in open(path):
        if l.startswith(">"):
            if cur:
                seqs.append("".join(cur))
                cur = []
        else:
            cur.append(l.strip().upper())
    if cur:
        seqs.append("".join(cur))
    cnt = {}
    for s in seqs:
        for i in range(len(s) - K + 1):
            w = s[i:i + K]
            if w.strip("ACGT"):
                continue
            cnt[w] = cnt.get(w, 0) + 1
    return sorted(cnt.items())


def diff_counts_text(names, kmers, cnts, nfa, p_adj, mA, mB, fc, pmax=0.2):
    want = "tag\tpvalue\tmeanA\tmeanB\tlog2FC\t" + "\t".join(names) + "\n"
    for i in range(len(cnts)):
        if p_adj[i] <= pmax and abs(fc[i]) >= 0:
            want += "%s\t%f\t%f\t%f\t%f" % (kmers[i], p_adj[i], mA[i], mB[i], fc[i])
            want += "".join("\t%.2f" % (cnts[i, j] / nfa[j]) for j in range(len(names))) + "\n"
    return want


def main():
    import oracle_lib as orc
    from common import enc
    os.makedirs(OUT, exist_ok=True)
    names = ["sample1", "sample2", "sample3", "sample4"]
    files = []
    for nm in names:
        # the reads themselves (sequence lines only, "mate<TAB>sequence"), for the device k-mer counter's parity test
        with gzip.GzipFile(os.path.join(OUT, f"reads_{nm}.txt.gz"), "wb", mtime=0) as f:
            for mate in (1, 2):
                for sq in reads(f"{DATA}/{nm}_{mate}.fastq.gz", NPAIRS, False):
                    f.write(f"{mate}\t{sq}\n".encode())
        recs = count_sample(nm)
        p = os.path.join(OUT, nm + ".txt.gz")
        with gzip.GzipFile(p, "wb", mtime=0) as f:
            f.write("".join(f"{k} {c}\n" for k, c in recs).encode())
        files.append(p)
        print(nm, len(recs), "k-mers")
    body = subprocess.run([os.path.join(ROOT, "oracle/_ref/joinCounts"), "-r", "2", "-a", "5"] + files, check=True,
                          capture_output=True).stdout
    raw = os.path.join(OUT, "raw-counts.tsv.gz")
    with open(raw, "wb") as f:                      # two gzip members, like the rule (Snakefile:594-596)
        f.write(gzip.compress(("tag\t" + "\t".join(names) + "\n").encode(), mtime=0))
        f.write(gzip.compress(body, mtime=0))
    nf_txt = subprocess.run([os.path.join(ROOT, "oracle/_ref/computeNF"), raw], check=True, capture_output=True, text=True).stdout
    open(os.path.join(OUT, "normalization_factors.tsv"), "w").write(nf_txt)
    nf = [float(l.split("\t")[1]) for l in nf_txt.strip().split("\n")[1:]]
    conds = "sample\tcondition\tnormalization_factor\n" + "".join(
        f"{n}\t{c}\t{v:f}\n" for n, c, v in zip(names, "AABB", nf))
    open(os.path.join(OUT, "sample_conditions_full.tsv"), "w").write(conds)
    rows = [l.split("\t") for l in body.decode().strip().split("\n")]
    kmers = [r[0] for r in rows]
    cnts = np.array([[int(x) for x in r[1:]] for r in rows], np.int32)
    nfa = np.array([float("%f" % v) for v in nf])
    p, mA, mB, fc = orc.ttest_rows(cnts, nfa, [0, 1], [2, 3])
    adj = orc.bh(p)
    want = "tag\tpvalue\tmeanA\tmeanB\tlog2FC\t" + "\t".join(names) + "\n"
    for i in range(len(cnts)):
        if adj[i] <= 0.2 and abs(fc[i]) >= 0:
            want += "%s\t%f\t%f\t%f\t%f" % (kmers[i], adj[i], mA[i], mB[i], fc[i])
            want += "".join("\t%.2f" % (cnts[i, j] / nfa[j]) for j in range(4)) + "\n"
    with gzip.GzipFile(os.path.join(OUT, "diff-counts.tsv.gz"), "wb", mtime=0) as f:
        f.write(want.encode())
    with gzip.GzipFile(os.path.join(OUT, "raw_pvals.txt.gz"), "wb", mtime=0) as f:
        f.write("".join("%s\t%f\n" % (k, v) for k, v in zip(kmers, p)).encode())
    print("rows", len(cnts), "survivors", want.count("\n") - 1, "NF", nf)
    # ---- mask: transcript dump -> real diffFilter.pl -> t-test on the masked matrix, NFs of the raw one
    ref = os.path.join(OUT, "ref_transcripts.txt.gz")
    with gzip.GzipFile(ref, "wb", mtime=0) as f:
        f.write("".join(f"{k} {c}\n" for k, c in count_transcripts("/root/reference/tests/gencode.v24.transcripts.head1000.fa")).encode())
    masked = subprocess.run(["perl", "/root/reference/bin/diffFilter.pl", ref, raw], check=True, capture_output=True).stdout
    with gzip.GzipFile(os.path.join(OUT, "masked-counts.tsv.gz"), "wb", mtime=0) as f:
        f.write(masked)
    mrows = [l.split("\t") for l in masked.decode().strip().split("\n")[1:]]
    mk = [r[0] for r in mrows]
    mc = np.array([[int(x) for x in r[1:]] for r in mrows], np.int32)
    p, mA, mB, fc = orc.ttest_rows(mc, nfa, [0, 1], [2, 3])
    wantm = diff_counts_text(names, mk, mc, nfa, orc.bh(p), mA, mB, fc)
    with gzip.GzipFile(os.path.join(OUT, "diff-counts-masked.tsv.gz"), "wb", mtime=0) as f:
        f.write(wantm.encode())
    print("masked rows", len(mc), "of", len(cnts), "survivors", wantm.count("\n") - 1)
    _ = enc


if __name__ == "__main__":
    main()

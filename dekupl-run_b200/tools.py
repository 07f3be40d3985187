"""Python mirror of the three reference tools' text interfaces, used by the parity tests so
they read like the reference's own invocations:
  join_counts_text  ~ `joinCounts -r R -a A f1 f2 ...`   (share/joinCounts/joinCounts.c)
  compute_nf_text   ~ `computeNF counts.tsv`             (share/computeNF/computeNF.c)
  ttest_filter_text ~ `TtestFilter -p P -f F counts conds A B` (share/TtestFilter/TtestFilter.c)
All compute goes through the C ABI (GPU); only parsing/printf formatting happens here. The
production CLIs are the C++ programs in host/."""
import gzip
import io

import numpy as np

from .binding import NF_REFERENCE
from .kmer import decode_kmer, encode_kmer


def _open_text(path):
    with open(path, "rb") as f:
        magic = f.read(2)
    if magic == b"\x1f\x8b":
        return io.TextIOWrapper(gzip.open(path, "rb"))   # gzip module reads multi-member files like zlib's gzread
    return open(path, "r")


def _atoi(tok):
    """C atoi: optional sign, leading digits, stops at first non-digit (kseq tokens have no leading space)"""
    i, sign = 0, 1
    if tok[:1] in "+-":
        sign = -1 if tok[0] == "-" else 1
        i = 1
    j = i
    while j < len(tok) and tok[j].isdigit():
        j += 1
    return sign * int(tok[i:j]) if j > i else 0


def read_dump(path):
    """per-sample `KMER COUNT` dump -> (keys uint64, counts int32, k). Tokens split on any whitespace
    exactly like ks_getuntil(KS_SEP_SPACE) pairs (joinCounts.c:55-68,86-92)."""
    toks = _open_text(path).read().split()
    kmers, cnts = toks[0::2], toks[1::2]
    k = len(kmers[0]) if kmers else 0
    keys = np.array([encode_kmer(s) for s in kmers], dtype=np.uint64)
    return keys, np.array([_atoi(c) for c in cnts], dtype=np.int32), k


def join_counts_text(ctx, files, min_rec=1, min_abund=1):
    dumps = [read_dump(f) for f in files]
    k = max(d[2] for d in dumps)
    m = ctx.join_counts([d[0] for d in dumps], [d[1] for d in dumps], min_rec, min_abund, emit_stale_tail=True)
    keys, cnts = m.export()
    m.free()
    return "".join(decode_kmer(key, k) + "".join("\t%d" % c for c in row) + "\n" for key, row in zip(keys, cnts))


def read_matrix(path):
    """`tag s1 s2 ...` header + rows -> (sample names, kmer strings, int32 matrix)"""
    f = _open_text(path)
    names = f.readline().split()[1:]
    kmers, rows = [], []
    for line in f:
        t = line.split()
        if not t:
            continue
        if len(t) != len(names) + 1:
            raise ValueError("inconsistent number of column (line %d)" % (len(rows) + 1))
        kmers.append(t[0])
        rows.append([_atoi(x) for x in t[1:]])
    return names, kmers, np.array(rows, dtype=np.int32).reshape(len(rows), len(names))


def compute_nf_text(ctx, path, step_rows=0):
    names, _, cnts = read_matrix(path)
    m = ctx.matrix_from_host(cnts)
    nf, _ = ctx.compute_nf(m, NF_REFERENCE, step_rows)
    m.free()
    return "sample\tnormalization_factor\n" + "".join("%s\t%f\n" % (s, v) for s, v in zip(names, nf))


def read_conditions(path, names, condA, condB):
    """TtestFilter.c:251-284: (idxA, idxB, nf[S]) in conditions-file order"""
    idxA, idxB, nf = [], [], np.zeros(len(names))
    lines = _open_text(path).read().split("\n")[1:]
    for line in lines:
        t = line.split()
        if len(t) < 3 or t[0] not in names:
            continue
        j = names.index(t[0])
        if t[1] == condA:
            idxA.append(j)
        elif t[1] == condB:
            idxB.append(j)
        nf[j] = float(t[2])
    return idxA, idxB, nf


def ttest_filter_text(ctx, counts_path, cond_path, condA, condB, pmax=0.05, fmin=0.0):
    names, kmers, cnts = read_matrix(counts_path)
    idxA, idxB, nf = read_conditions(cond_path, names, condA, condB)
    m = ctx.matrix_from_host(cnts)
    t = ctx.ttest_rows(m, nf, idxA, idxB)
    t.adjust(pmax)
    t.filter(pmax, fmin)
    s = t.export_survivors()
    out = ["tag\tpvalue\tmeanA\tmeanB\tlog2FC" + "".join("\t" + n for n in names) + "\n"]
    for i in range(t.D):
        out.append("%s\t%f\t%f\t%f\t%f" % (kmers[int(s["row"][i])], s["p_adj"][i], s["meanA"][i], s["meanB"][i], s["log2fc"][i])
                   + "".join("\t%.2f" % v for v in s["norm"][i]) + "\n")
    t.free()
    m.free()
    return "".join(out)

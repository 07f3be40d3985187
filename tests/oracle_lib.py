"""ctypes wrapper of the CPU oracle (oracle/_ref/liboracle.so). TEST-ONLY: the product never imports this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "liboracle.so")
_L = None


def load():
    global _L
    if _L is not None:
        return _L
    if not os.path.exists(SO):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/liboracle.so"], check=True, capture_output=True)
    L = C.CDLL(SO)
    vp, u64, i32, dbl, sz = C.c_void_p, C.c_uint64, C.c_int, C.c_double, C.c_size_t
    L.orc_join.restype = sz
    L.orc_join.argtypes = [i32, C.POINTER(vp), C.POINTER(vp), C.POINTER(sz), i32, i32, i32, vp, vp, sz, C.POINTER(sz)]
    L.orc_nf_reference.restype = sz
    L.orc_nf_reference.argtypes = [vp, sz, i32, sz, vp, C.POINTER(i32)]
    L.orc_nf_full.restype = None
    L.orc_nf_full.argtypes = [vp, sz, i32, vp]
    L.orc_student_t_2tail.restype = dbl
    L.orc_student_t_2tail.argtypes = [dbl, dbl]
    L.orc_student_t_2tail_boost_branches.restype = dbl
    L.orc_student_t_2tail_boost_branches.argtypes = [dbl, dbl]
    L.orc_ttest_rows.restype = None
    L.orc_ttest_rows.argtypes = [vp, sz, i32, vp, vp, i32, vp, i32, vp, vp, vp, vp]
    L.orc_poisson_lr.restype = dbl
    L.orc_poisson_lr.argtypes = [dbl, dbl, C.c_int64, C.c_int64]
    L.orc_poisson_rows.restype = None
    L.orc_poisson_rows.argtypes = [vp, sz, i32, vp, i32, vp, i32, vp, vp, vp, vp, vp]
    L.orc_antijoin.restype = sz
    L.orc_antijoin.argtypes = [vp, sz, vp, sz, vp]
    L.orc_count_kmers.restype = C.c_longlong
    L.orc_count_kmers.argtypes = [vp, vp, sz, vp, i32, i32, C.c_uint32, vp, vp, sz]
    L.orc_write_dump_text.restype = C.c_long
    L.orc_write_dump_text.argtypes = [C.c_char_p, vp, vp, sz, i32]
    L.orc_parse_matrix_text.restype = C.c_long
    L.orc_parse_matrix_text.argtypes = [C.c_char_p, i32, vp, sz]
    L.orc_bh.restype = None
    L.orc_bh.argtypes = [vp, sz, vp]
    L.orc_filter.restype = sz
    L.orc_filter.argtypes = [vp, vp, sz, dbl, dbl, vp]
    _L = L
    return L


def join(keys, cnts, min_rec=1, min_abund=1, stale_tail=False):
    L = load()
    S = len(keys)
    keys = [np.ascontiguousarray(k, dtype=np.uint64) for k in keys]
    cnts = [np.ascontiguousarray(c, dtype=np.int32) for c in cnts]
    cap = sum(len(k) for k in keys) + 1
    ok, oc = np.empty(cap, np.uint64), np.empty((cap, S), np.int32)
    kp = (C.c_void_p * S)(*[k.ctypes.data for k in keys])
    cp = (C.c_void_p * S)(*[c.ctypes.data for c in cnts])
    nn = (C.c_size_t * S)(*[len(k) for k in keys])
    U = C.c_size_t(0)
    R = L.orc_join(S, kp, cp, nn, min_rec, min_abund, int(stale_tail), ok.ctypes.data, oc.ctypes.data, cap, C.byref(U))
    return ok[:R].copy(), oc[:R].copy(), int(U.value)


def nf_reference(cnts, step=1000000):
    L = load()
    cnts = np.ascontiguousarray(cnts, dtype=np.int32)
    R, S = cnts.shape
    nf = np.empty(S)
    uz = C.c_int(0)
    used = L.orc_nf_reference(cnts.ctypes.data, R, S, step, nf.ctypes.data, C.byref(uz))
    return nf, int(used), int(uz.value)


def nf_full(cnts):
    L = load()
    cnts = np.ascontiguousarray(cnts, dtype=np.int32)
    R, S = cnts.shape
    nf = np.empty(S)
    L.orc_nf_full(cnts.ctypes.data, R, S, nf.ctypes.data)
    return nf


def ttest_rows(cnts, nf, idxA, idxB):
    L = load()
    cnts = np.ascontiguousarray(cnts, dtype=np.int32)
    R, S = cnts.shape
    nf = np.ascontiguousarray(nf, dtype=np.float64)
    a, b = np.array(idxA, dtype=np.int32), np.array(idxB, dtype=np.int32)
    p, mA, mB, fc = np.empty(R), np.empty(R), np.empty(R), np.empty(R)
    L.orc_ttest_rows(cnts.ctypes.data, R, S, nf.ctypes.data, a.ctypes.data, len(a), b.ctypes.data, len(b),
                     p.ctypes.data, mA.ctypes.data, mB.ctypes.data, fc.ctypes.data)
    return p, mA, mB, fc


def poisson_rows(cnts, idxA, idxB):
    """-l mode: (p_raw, meanA, meanB, log2fc, (N1, N2))"""
    L = load()
    cnts = np.ascontiguousarray(cnts, dtype=np.int32)
    R, S = cnts.shape
    a, b = np.array(idxA, dtype=np.int32), np.array(idxB, dtype=np.int32)
    p, mA, mB, fc = np.empty(R), np.empty(R), np.empty(R), np.empty(R)
    N = np.zeros(2, np.int64)
    L.orc_poisson_rows(cnts.ctypes.data, R, S, a.ctypes.data, len(a), b.ctypes.data, len(b),
                       p.ctypes.data, mA.ctypes.data, mB.ctypes.data, fc.ctypes.data, N.ctypes.data)
    return p, mA, mB, fc, (int(N[0]), int(N[1]))


def poisson_lr(K1, K2, N1, N2):
    return load().orc_poisson_lr(float(K1), float(K2), int(N1), int(N2))


def antijoin(keys, mask):
    """keep flags of bin/diffFilter.pl: rows whose k-mer is absent from the sorted mask dump"""
    L = load()
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    mask = np.ascontiguousarray(mask, dtype=np.uint64)
    keep = np.zeros(len(keys), np.uint8)
    L.orc_antijoin(keys.ctypes.data, len(keys), mask.ctypes.data, len(mask), keep.ctypes.data)
    return keep.astype(bool)


def count_kmers(reads, k=31, rc_flags=None, canonical=False, min_count=2):
    """sorted (k-mer, count) table of one sample: jellyfish count -L -m [-C] | dump | sort on packed k-mers"""
    L = load()
    seqs = [r.encode() if isinstance(r, str) else r for r in reads]
    off = np.zeros(len(seqs) + 1, np.uint64)
    np.cumsum([len(x) for x in seqs], out=off[1:])
    bases = np.frombuffer(b"".join(seqs) + b"\0", dtype=np.uint8)
    rc = None if rc_flags is None else np.ascontiguousarray(rc_flags, dtype=np.uint8)
    cap = int(off[-1]) + 1
    keys, cnts = np.empty(cap, np.uint64), np.empty(cap, np.int32)
    n = L.orc_count_kmers(bases.ctypes.data, off.ctypes.data, len(seqs), rc.ctypes.data if rc is not None else None, k,
                          int(canonical), int(min_count), keys.ctypes.data, cnts.ctypes.data, cap)
    assert n >= 0
    return keys[:n].copy(), cnts[:n].copy()


def bh(p):
    L = load()
    p = np.ascontiguousarray(p, dtype=np.float64)
    out = np.empty_like(p)
    L.orc_bh(p.ctypes.data, len(p), out.ctypes.data)
    return out


def student_t_2tail(t, df):
    return load().orc_student_t_2tail(float(t), float(df))


def write_dump_text(path, keys, cnts, k=31):
    """one sample's dump as text ("KMER<tab>COUNT" lines, what joinCounts reads); returns the bytes written"""
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    cnts = np.ascontiguousarray(cnts, dtype=np.int32)
    n = load().orc_write_dump_text(path.encode(), keys.ctypes.data, cnts.ctypes.data, len(keys), k)
    assert n >= 0, path
    return int(n)


def parse_matrix_text(path, S, cap):
    """the rows joinCounts printed (optionally behind a header line) as an int32 matrix"""
    out = np.empty((cap, S), np.int32)
    R = load().orc_parse_matrix_text(path.encode(), S, out.ctypes.data, cap)
    assert R >= 0, path
    return out[:R]


def student_t_2tail_boost_branches(t, df):
    return load().orc_student_t_2tail_boost_branches(float(t), float(df))

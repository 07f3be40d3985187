"""ctypes binding of libdekupl_b200.so (include/dekupl_b200.h)."""
import ctypes as C
import os
import re

import numpy as np

NF_REFERENCE, NF_FULL = 0, 1
MAX_SAMPLES = 256
_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class DkError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"dekupl_b200 error {code}: {msg}")
        self.code = code


def lib_path():
    # DK_LIB: an alternative build of the SAME library (e.g. the -DDK_M3_PROF phase-timing variant)
    return os.environ.get("DK_LIB") or os.path.join(_HERE, "libdekupl_b200.so")


def declared_symbols():
    """every dk_* function include/dekupl_b200.h declares"""
    hdr = open(os.path.join(_HERE, "..", "include", "dekupl_b200.h")).read()
    return sorted(set(re.findall(r"\b(dk_[a-z0-9_]+)\s*\(", hdr)))


class PathStats(C.Structure):
    _fields_ = [("U", C.c_uint64), ("R", C.c_uint64), ("R_pos", C.c_uint64), ("D", C.c_uint64),
                ("ties", C.c_uint64 * 3), ("nf", C.c_double * MAX_SAMPLES),
                ("ms_join", C.c_double), ("ms_nf", C.c_double), ("ms_ttest", C.c_double), ("ms_bh", C.c_double)]


class StreamStats(C.Structure):
    _fields_ = [("U", C.c_uint64), ("R", C.c_uint64), ("n_cand", C.c_uint64), ("chunks", C.c_uint64), ("h2d_bytes", C.c_uint64),
                ("h2d_seconds", C.c_double), ("ms_join", C.c_double), ("ms_merge", C.c_double), ("ms_ttest", C.c_double),
                ("ms_total", C.c_double)]


def load_library():
    """Load the CUDA library. Fails loudly when it has not been built (no fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    p = lib_path()
    if not os.path.exists(p):
        raise DkError(-5, f"{p} not found: build it with dekupl-run_b200/csrc/build.sh (no CPU fallback exists)")
    L = C.CDLL(p)
    vp, u64, i32, dbl = C.c_void_p, C.c_uint64, C.c_int, C.c_double
    P = C.POINTER
    sig = {
        "dk_init": (i32, [P(vp), i32]), "dk_free": (None, [vp]), "dk_last_error": (C.c_char_p, [vp]),
        "dk_version": (C.c_char_p, []), "dk_set_stream": (i32, [vp, vp]), "dk_sync": (i32, [vp]),
        "dk_launch_count": (u64, [vp]), "dk_stage_ms": (dbl, [vp, i32]),
        "dk_join_counts": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, i32, P(vp)]),
        "dk_join_counts_dev": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, i32, P(vp)]),
        "dk_key_splitters": (i32, [vp, i32, P(vp), P(u64), i32, i32, vp]),
        "dk_matrix_from_host": (i32, [vp, i32, u64, vp, vp, P(vp)]),
        "dk_matrix_rows": (u64, [vp]), "dk_matrix_union": (u64, [vp]), "dk_matrix_samples": (i32, [vp]),
        "dk_matrix_export": (i32, [vp, vp, u64, u64, vp, vp]),
        "dk_matrix_export_dev": (i32, [vp, vp, u64, u64, vp, vp]),
        "dk_matrix_format_rows": (i32, [vp, vp, u64, u64, i32, vp, u64, P(u64)]),
        "dk_matrix_checksum": (i32, [vp, vp, P(u64), P(u64)]), "dk_matrix_free": (None, [vp, vp]),
        "dk_count_kmers": (i32, [vp, vp, vp, u64, vp, i32, i32, C.c_uint32, P(vp)]),
        "dk_dump_size": (u64, [vp]), "dk_dump_keys_dev": (vp, [vp]), "dk_dump_cnts_dev": (vp, [vp]),
        "dk_dump_stats": (None, [vp, P(u64), P(u64), P(u64)]), "dk_dump_export": (i32, [vp, vp, vp, vp]),
        "dk_dump_free": (None, [vp, vp]), "dk_dump_merge": (i32, [vp, P(vp), i32, C.c_uint32, P(vp)]),
        "dk_merge_tags": (i32, [vp, vp, u64, i32, vp, vp, i32, P(vp)]),
        "dk_contigs_count": (u64, [vp]), "dk_contigs_chars": (u64, [vp]), "dk_contigs_rings": (u64, [vp]),
        "dk_contigs_export": (i32, [vp, vp, vp, vp, vp]), "dk_contigs_free": (None, [vp]),
        "dk_matrix_mask": (i32, [vp, vp, vp, u64, i32, P(vp), P(u64)]),
        "dk_keys_antijoin": (i32, [vp, vp, u64, vp, u64, vp]),
        "dk_run_path_masked": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, i32, i32, P(i32), i32, P(i32), i32, dbl, dbl, i32,
                                     vp, u64, i32, P(PathStats), P(u64), P(vp), P(vp)]),
        "dk_compute_nf": (i32, [vp, vp, i32, u64, P(dbl), P(u64)]),
        "dk_nf_begin": (i32, [vp, vp, P(vp), P(u64)]), "dk_nf_set_rank": (i32, [vp, vp, u64]),
        "dk_nf_hist_dev": (i32, [vp, vp, vp]), "dk_nf_advance_dev": (i32, [vp, vp, vp]),
        "dk_nf_result": (i32, [vp, vp, P(dbl)]), "dk_nf_end": (None, [vp, vp]),
        "dk_ttest_rows": (i32, [vp, vp, P(dbl), P(i32), i32, P(i32), i32, P(vp)]),
        "dk_poisson_rows": (i32, [vp, vp, P(i32), i32, P(i32), i32, P(vp), P(u64)]),
        "dk_ttest_candidates": (i32, [vp, vp, dbl, P(u64), P(vp)]),
        "dk_ttest_adjust_global": (i32, [vp, vp, vp, u64, u64]),
        "dk_ttest_adjust": (i32, [vp, vp, dbl]),
        "dk_ttest_count_le": (i32, [vp, vp, dbl, P(u64)]),
        "dk_bh_next_cut": (dbl, [dbl, u64, u64]),
        "dk_ttest_candidates_narrowed": (i32, [vp, vp, dbl, dbl, P(u64), P(vp)]),
        "dk_bh_rank_slice": (i32, [vp, vp, u64, u64, u64, vp, vp]),
        "dk_bh_rank_slice_noalloc": (i32, [vp, vp, u64, u64, u64, vp, vp]),
        "dk_ttest_apply_adjusted": (i32, [vp, vp, vp, vp, u64]),
        "dk_ttest_filter": (i32, [vp, vp, dbl, dbl, P(u64), P(u64)]),
        "dk_ttest_export_survivors": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "dk_ttest_export_rows": (i32, [vp, vp, vp, vp, vp]),
        "dk_ttest_rows_count": (u64, [vp]), "dk_ttest_free": (None, [vp, vp]),
        "dk_stream_nf_reference": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, u64, u64, i32, P(dbl), P(u64), P(i32)]),
        "dk_stream_positive_rows": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, u64, P(StreamStats), P(vp)]),
        "dk_nf_reference_sparse": (i32, [vp, i32, u64, vp, vp, u64, u64, P(dbl), P(u64)]),
        "dk_stream_path": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, u64, P(dbl), P(i32), i32, P(i32), i32, dbl,
                                 P(StreamStats), P(vp), P(vp)]),
        "dk_run_path_matrix_out": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, i32, i32, P(i32), i32, P(i32), i32, dbl, dbl, i32,
                                         vp, vp, u64, P(u64), P(PathStats), P(vp), P(vp)]),
        "dk_comm_unique_id": (i32, [vp, vp]), "dk_comm_init": (i32, [vp, vp, i32, i32]), "dk_comm_free": (None, [vp]),
        "dk_comm_rank": (i32, [vp]), "dk_comm_world": (i32, [vp]),
        "dk_run_path_sharded": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, i32, i32, P(i32), i32, P(i32), i32, dbl, dbl, i32,
                                      P(PathStats), P(u64), P(vp), P(vp)]),
        "dk_run_path": (i32, [vp, i32, P(vp), P(vp), P(u64), i32, i32, i32, i32, P(i32), i32, P(i32), i32, dbl, dbl, i32,
                              P(PathStats), P(vp), P(vp)]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype, f.argtypes = res, args
    _LIB = L
    return L


def _np_ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


class Matrix:
    def __init__(self, ctx, handle):
        self.ctx, self.h = ctx, handle

    @property
    def rows(self):
        return int(self.ctx.L.dk_matrix_rows(self.h))

    @property
    def union(self):
        return int(self.ctx.L.dk_matrix_union(self.h))

    @property
    def samples(self):
        return int(self.ctx.L.dk_matrix_samples(self.h))

    def export(self, row0=0, nrows=None, keys=True):
        n = self.rows - row0 if nrows is None else nrows
        k = np.empty(n, dtype=np.uint64) if keys else None
        c = np.empty((n, self.samples), dtype=np.int32)
        self.ctx._chk(self.ctx.L.dk_matrix_export(self.ctx.h, self.h, row0, n, _np_ptr(k), _np_ptr(c)))
        return k, c

    def format_rows(self, k, row0=0, nrows=None):
        """the text joinCounts prints for these rows (joinCounts.c:105-109), formatted on the device"""
        n = self.rows - row0 if nrows is None else nrows
        ln = C.c_uint64(0)
        if n == 0:
            return b""
        self.ctx.L.dk_matrix_format_rows(self.ctx.h, self.h, row0, n, k, None, 0, C.byref(ln))      # sizes only: *len = bytes needed
        buf = np.empty(int(ln.value) + 16, dtype=np.uint8)
        self.ctx._chk(self.ctx.L.dk_matrix_format_rows(self.ctx.h, self.h, row0, n, k, _np_ptr(buf), len(buf), C.byref(ln)))
        return buf[: ln.value].tobytes()

    def checksum(self):
        x, s = C.c_uint64(0), C.c_uint64(0)
        self.ctx._chk(self.ctx.L.dk_matrix_checksum(self.ctx.h, self.h, C.byref(x), C.byref(s)))
        return int(x.value), int(s.value)

    def free(self):
        if self.h:
            self.ctx.L.dk_matrix_free(self.ctx.h, self.h)
            self.h = None


class Dump:
    """one sample's sorted (k-mer, count) table on the device (dk_count_kmers)"""

    def __init__(self, ctx, handle):
        self.ctx, self.h = ctx, handle

    @property
    def size(self):
        return int(self.ctx.L.dk_dump_size(self.h))

    @property
    def keys_ptr(self):
        return int(self.ctx.L.dk_dump_keys_dev(self.h) or 0)

    @property
    def cnts_ptr(self):
        return int(self.ctx.L.dk_dump_cnts_dev(self.h) or 0)

    def stats(self):
        w, v, d = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        self.ctx.L.dk_dump_stats(self.h, C.byref(w), C.byref(v), C.byref(d))
        return dict(windows=int(w.value), valid=int(v.value), distinct=int(d.value))

    def export(self):
        n = self.size
        k, c = np.empty(n, np.uint64), np.empty(n, np.int32)
        self.ctx._chk(self.ctx.L.dk_dump_export(self.ctx.h, self.h, _np_ptr(k), _np_ptr(c)))
        return k, c

    def free(self):
        if self.h:
            self.ctx.L.dk_dump_free(self.ctx.h, self.h)
            self.h = None


class TtestResult:
    def __init__(self, ctx, handle, matrix):
        self.ctx, self.h, self.matrix = ctx, handle, matrix
        self.D = 0
        self.ties = (0, 0, 0)

    @property
    def rows(self):
        return int(self.ctx.L.dk_ttest_rows_count(self.h))

    def adjust(self, p_cut):
        self.ctx._chk(self.ctx.L.dk_ttest_adjust(self.ctx.h, self.h, float(p_cut)))

    def candidates(self, p_cut):
        n, p = C.c_uint64(0), C.c_void_p(0)
        self.ctx._chk(self.ctx.L.dk_ttest_candidates(self.ctx.h, self.h, float(p_cut), C.byref(n), C.byref(p)))
        return int(n.value), p.value

    def count_le(self, cut):
        n = C.c_uint64(0)
        self.ctx._chk(self.ctx.L.dk_ttest_count_le(self.ctx.h, self.h, float(cut), C.byref(n)))
        return int(n.value)

    def candidates_narrowed(self, cut, p_cut):
        n, p = C.c_uint64(0), C.c_void_p(0)
        self.ctx._chk(self.ctx.L.dk_ttest_candidates_narrowed(self.ctx.h, self.h, float(cut), float(p_cut), C.byref(n), C.byref(p)))
        return int(n.value), p.value

    def adjust_global(self, all_p_dev_ptr, n_all, R_global):
        self.ctx._chk(self.ctx.L.dk_ttest_adjust_global(self.ctx.h, self.h, C.c_void_p(all_p_dev_ptr), n_all, R_global))

    def apply_adjusted(self, sorted_p_ptr, F_ptr, n_all):
        self.ctx._chk(self.ctx.L.dk_ttest_apply_adjusted(self.ctx.h, self.h, C.c_void_p(sorted_p_ptr), C.c_void_p(F_ptr), n_all))

    def filter(self, pmax, fmin):
        d, t = C.c_uint64(0), (C.c_uint64 * 3)()
        self.ctx._chk(self.ctx.L.dk_ttest_filter(self.ctx.h, self.h, float(pmax), float(fmin), C.byref(d), t))
        self.D, self.ties = int(d.value), tuple(int(x) for x in t)
        return self.D

    def export_rows(self, adj=True):
        R = self.rows
        p, f = np.empty(R), np.empty(R)
        a = np.empty(R) if adj else None
        self.ctx._chk(self.ctx.L.dk_ttest_export_rows(self.ctx.h, self.h, _np_ptr(p), _np_ptr(f), _np_ptr(a)))
        return p, f, a

    def export_survivors(self):
        D, S = self.D, self.matrix.samples
        out = dict(row=np.empty(D, np.uint64), key=np.empty(D, np.uint64), p_adj=np.empty(D), meanA=np.empty(D),
                   meanB=np.empty(D), log2fc=np.empty(D), norm=np.empty((D, S)))
        self.ctx._chk(self.ctx.L.dk_ttest_export_survivors(
            self.ctx.h, self.h, _np_ptr(out["row"]), _np_ptr(out["key"]), _np_ptr(out["p_adj"]), _np_ptr(out["meanA"]),
            _np_ptr(out["meanB"]), _np_ptr(out["log2fc"]), _np_ptr(out["norm"])))
        return out

    def free(self):
        if self.h:
            self.ctx.L.dk_ttest_free(self.ctx.h, self.h)
            self.h = None


class Context:
    """One GPU, one caller thread (mirrors the single-threaded reference tools)."""

    def __init__(self, device=0):
        self.L = load_library()
        h = C.c_void_p(0)
        rc = self.L.dk_init(C.byref(h), int(device))
        if rc != 0:
            raise DkError(rc, self.L.dk_last_error(None).decode())
        self.h = h

    def _chk(self, rc):
        if rc != 0:
            raise DkError(rc, self.L.dk_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.L.dk_free(self.h)
            self.h = None

    def set_stream(self, cuda_stream_ptr):
        self._chk(self.L.dk_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        self._chk(self.L.dk_sync(self.h))

    @property
    def launches(self):
        return int(self.L.dk_launch_count(self.h))

    def stage_ms(self, which):
        return float(self.L.dk_stage_ms(self.h, which))

    # ---- joinCounts -----------------------------------------------------------
    @staticmethod
    def _tables(keys, cnts, n):
        S = len(keys)
        kp, cp = (C.c_void_p * S)(*keys), (C.c_void_p * S)(*cnts)
        nn = (C.c_uint64 * S)(*n)
        return S, kp, cp, nn

    def join_counts(self, keys, cnts, min_rec=1, min_abund=1, emit_stale_tail=False):
        """keys/cnts: lists of numpy arrays (uint64 sorted / int32), one per sample (host)."""
        keys = [np.ascontiguousarray(k, dtype=np.uint64) for k in keys]
        cnts = [np.ascontiguousarray(c, dtype=np.int32) for c in cnts]
        S, kp, cp, nn = self._tables([k.ctypes.data for k in keys], [c.ctypes.data for c in cnts], [len(k) for k in keys])
        out = C.c_void_p(0)
        self._chk(self.L.dk_join_counts(self.h, S, kp, cp, nn, min_rec, min_abund, int(emit_stale_tail), C.byref(out)))
        return Matrix(self, out)

    def join_counts_dev(self, key_ptrs, cnt_ptrs, n, min_rec=1, min_abund=1, emit_stale_tail=False):
        """device pointers (ints), e.g. torch tensors' data_ptr()"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        out = C.c_void_p(0)
        self._chk(self.L.dk_join_counts_dev(self.h, S, kp, cp, nn, min_rec, min_abund, int(emit_stale_tail), C.byref(out)))
        return Matrix(self, out)

    def key_splitters(self, key_ptrs, n, G, device_inputs=True):
        """G-1 count-balanced shard boundaries (uint64 k-mers) for dumps given by pointers (device unless told otherwise)"""
        S = len(key_ptrs)
        kp, nn = (C.c_void_p * S)(*key_ptrs), (C.c_uint64 * S)(*n)
        out = np.zeros(max(G - 1, 1), np.uint64)
        self._chk(self.L.dk_key_splitters(self.h, S, kp, nn, int(device_inputs), G, _np_ptr(out)))
        return out[: G - 1]

    def matrix_from_host(self, cnts, keys=None):
        cnts = np.ascontiguousarray(cnts, dtype=np.int32)
        R, S = cnts.shape
        if keys is not None:
            keys = np.ascontiguousarray(keys, dtype=np.uint64)
        out = C.c_void_p(0)
        self._chk(self.L.dk_matrix_from_host(self.h, S, R, _np_ptr(keys), _np_ptr(cnts), C.byref(out)))
        return Matrix(self, out)

    # ---- k-mer counting (jellyfish count | dump | sort) -------------------------
    def count_kmers(self, reads, k=31, rc_flags=None, canonical=False, min_count=2):
        """reads: list of bytes/str sequence lines; rc_flags: per-read 0/1 (reads the pipeline reverse-complements)"""
        seqs = [r.encode() if isinstance(r, str) else r for r in reads]
        lens = np.fromiter((len(x) for x in seqs), dtype=np.uint64, count=len(seqs))
        off = np.zeros(len(seqs) + 1, np.uint64)
        np.cumsum(lens, out=off[1:])
        bases = np.frombuffer(b"".join(seqs), dtype=np.uint8) if seqs else np.zeros(0, np.uint8)
        rc = None if rc_flags is None else np.ascontiguousarray(rc_flags, dtype=np.uint8)
        out = C.c_void_p(0)
        self._chk(self.L.dk_count_kmers(self.h, _np_ptr(bases), _np_ptr(off), len(seqs), _np_ptr(rc), k, int(canonical),
                                        int(min_count), C.byref(out)))
        return Dump(self, out)

    def merge_dumps(self, dumps, min_count=2):
        """sum of tables counted with min_count=1 over read batches of one sample"""
        arr = (C.c_void_p * len(dumps))(*[d.h for d in dumps])
        out = C.c_void_p(0)
        self._chk(self.L.dk_dump_merge(self.h, arr, len(dumps), int(min_count), C.byref(out)))
        return Dump(self, out)

    # ---- mergeTags -----------------------------------------------------------------
    def merge_tags(self, kmers, pvalues, rep_ids=None, k=None, min_k=15):
        """contigs of one regulation direction; returns list of (rep_id, nb_merged, sequence) and the ring-member count"""
        kmers = [s.encode() if isinstance(s, str) else s for s in kmers]
        n = len(kmers)
        k = k if k is not None else (len(kmers[0]) if n else 31)
        seqs = np.frombuffer(b"".join(kmers), dtype=np.uint8) if n else np.zeros(0, np.uint8)
        p = np.ascontiguousarray(pvalues, dtype=np.float64)
        rid = np.arange(n, dtype=np.uint32) if rep_ids is None else np.ascontiguousarray(rep_ids, dtype=np.uint32)
        out = C.c_void_p(0)
        self._chk(self.L.dk_merge_tags(self.h, _np_ptr(seqs), n, k, _np_ptr(p), _np_ptr(rid), min_k, C.byref(out)))
        m, nch, rings = int(self.L.dk_contigs_count(out)), int(self.L.dk_contigs_chars(out)), int(self.L.dk_contigs_rings(out))
        rep, nb, off, ch = np.empty(m, np.uint32), np.empty(m, np.uint32), np.empty(m + 1, np.uint64), np.empty(nch, np.uint8)
        self.L.dk_contigs_export(out, _np_ptr(rep), _np_ptr(nb), _np_ptr(off), _np_ptr(ch))
        self.L.dk_contigs_free(out)
        txt = ch.tobytes()
        return [(int(rep[j]), int(nb[j]), txt[int(off[j]):int(off[j + 1])].decode()) for j in range(m)], rings

    # ---- reference-transcript mask (bin/diffFilter.pl) ---------------------------
    def matrix_mask(self, m, mask_keys, mask_on_device=False, n_mask=None):
        """rows of m whose k-mer is absent from the sorted mask; returns (Matrix, rows_removed).
        mask_keys: numpy uint64 array, or a device pointer (int) with n_mask when mask_on_device"""
        if mask_on_device:
            ptr, n = C.c_void_p(int(mask_keys)), int(n_mask)
        else:
            mask_keys = np.ascontiguousarray(mask_keys, dtype=np.uint64)
            ptr, n = _np_ptr(mask_keys), len(mask_keys)
        out, removed = C.c_void_p(0), C.c_uint64(0)
        self._chk(self.L.dk_matrix_mask(self.h, m.h, ptr, n, int(mask_on_device), C.byref(out), C.byref(removed)))
        return Matrix(self, out), int(removed.value)

    def keys_antijoin(self, keys, mask_keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        mask_keys = np.ascontiguousarray(mask_keys, dtype=np.uint64)
        keep = np.zeros(len(keys), np.uint8)
        self._chk(self.L.dk_keys_antijoin(self.h, _np_ptr(keys), len(keys), _np_ptr(mask_keys), len(mask_keys), _np_ptr(keep)))
        return keep.astype(bool)

    # ---- computeNF ------------------------------------------------------------
    def compute_nf(self, m, mode=NF_REFERENCE, step_rows=0):
        S = m.samples
        nf, info = (C.c_double * S)(), (C.c_uint64 * 4)()
        self._chk(self.L.dk_compute_nf(self.h, m.h, mode, step_rows, nf, info))
        return np.array(nf[:]), dict(consumed=int(info[0]), steps=int(info[1]), undefined_zone=int(info[2]), n_pos=int(info[3]))

    # ---- TtestFilter ----------------------------------------------------------
    def ttest_rows(self, m, nf, idxA, idxB):
        nf = np.ascontiguousarray(nf, dtype=np.float64)
        a, b = (C.c_int * len(idxA))(*idxA), (C.c_int * len(idxB))(*idxB)
        out = C.c_void_p(0)
        self._chk(self.L.dk_ttest_rows(self.h, m.h, nf.ctypes.data_as(C.POINTER(C.c_double)), a, len(idxA), b, len(idxB),
                                       C.byref(out)))
        return TtestResult(self, out, m)

    def poisson_rows(self, m, idxA, idxB):
        """TtestFilter -l; returns (TtestResult, (N1, N2, negative_statistics))"""
        a, b = (C.c_int * len(idxA))(*idxA), (C.c_int * len(idxB))(*idxB)
        out, info = C.c_void_p(0), (C.c_uint64 * 3)()
        self._chk(self.L.dk_poisson_rows(self.h, m.h, a, len(idxA), b, len(idxB), C.byref(out), info))
        return TtestResult(self, out, m), tuple(int(x) for x in info)

    # ---- several GPUs: an NCCL communicator owned by the context ----------------------
    def comm_unique_id(self):
        """rank 0: the bytes every rank passes to comm_init (hand them over by any means)"""
        buf = (C.c_ubyte * 128)()
        self._chk(self.L.dk_comm_unique_id(self.h, buf))
        return bytes(buf)

    def comm_init(self, uid, rank, world):
        buf = (C.c_ubyte * 128).from_buffer_copy(uid)
        self._chk(self.L.dk_comm_init(self.h, buf, int(rank), int(world)))

    def comm_free(self):
        self.L.dk_comm_free(self.h)

    def run_path_sharded(self, key_ptrs, cnt_ptrs, n, device_inputs, min_rec, min_abund, nf_mode, idxA, idxB, pmax, fmin,
                         nf_text_roundtrip=True, keep=False):
        """dk_run_path on this rank's shard, global pieces by NCCL inside the library; stats are GLOBAL (+ *_local)"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        a, b = (C.c_int * len(idxA))(*idxA), (C.c_int * len(idxB))(*idxB)
        st, loc, om, ot = PathStats(), (C.c_uint64 * 3)(), C.c_void_p(0), C.c_void_p(0)
        self._chk(self.L.dk_run_path_sharded(self.h, S, kp, cp, nn, int(device_inputs), min_rec, min_abund, nf_mode, a, len(idxA), b,
                                             len(idxB), float(pmax), float(fmin), int(nf_text_roundtrip), C.byref(st), loc,
                                             C.byref(om) if keep else None, C.byref(ot) if keep else None))
        stats = dict(U=int(st.U), R=int(st.R), R_pos=int(st.R_pos), D=int(st.D), ties=tuple(int(x) for x in st.ties),
                     nf=np.array(st.nf[:S]), ms_join=st.ms_join, ms_nf=st.ms_nf, ms_ttest=st.ms_ttest, ms_bh=st.ms_bh,
                     U_local=int(loc[0]), R_local=int(loc[1]), D_local=int(loc[2]))
        if keep:
            m = Matrix(self, om)
            t = TtestResult(self, ot, m) if ot.value else None
            if t is not None:
                t.D, t.ties = stats["D_local"], stats["ties"]
            return stats, m, t
        return stats

    # ---- chunk-streamed path (host dumps larger than device memory) ----------------
    def stream_nf_reference(self, key_ptrs, cnt_ptrs, n, min_rec, min_abund, chunk_records, step_rows=0, widen=True):
        """computeNF's factors from the head of the table; info["stop_fired"] says whether its early stop fired there"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        nf, info, fired = (C.c_double * S)(), (C.c_uint64 * 4)(), C.c_int(0)
        self._chk(self.L.dk_stream_nf_reference(self.h, S, kp, cp, nn, min_rec, min_abund, int(chunk_records), int(step_rows),
                                                int(widen), nf, info, C.byref(fired)))
        return np.array(nf[:]), dict(consumed=int(info[0]), steps=int(info[1]), undefined_zone=int(info[2]), n_pos=int(info[3]),
                                     stop_fired=bool(fired.value))

    def stream_positive_rows(self, key_ptrs, cnt_ptrs, n, min_rec, min_abund, chunk_records):
        """-> (stats, row numbers uint64[n_pos], counts int32[n_pos][S]) of the all-positive rows of this rank's table"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        st, om = StreamStats(), C.c_void_p(0)
        self._chk(self.L.dk_stream_positive_rows(self.h, S, kp, cp, nn, min_rec, min_abund, int(chunk_records), C.byref(st), C.byref(om)))
        stats = {k: getattr(st, k) for k, _ in StreamStats._fields_}
        m = Matrix(self, om)
        rows, cnts = m.export() if m.rows else (np.zeros(0, np.uint64), np.zeros((0, S), np.int32))
        m.free()
        return stats, rows, cnts

    def nf_reference_sparse(self, row_idx, cnts, R_total, step_rows=0):
        row_idx = np.ascontiguousarray(row_idx, dtype=np.uint64)
        cnts = np.ascontiguousarray(cnts, dtype=np.int32)
        S = cnts.shape[1]
        nf, info = (C.c_double * S)(), (C.c_uint64 * 4)()
        self._chk(self.L.dk_nf_reference_sparse(self.h, S, len(row_idx), _np_ptr(row_idx), _np_ptr(cnts), int(R_total), int(step_rows), nf, info))
        return np.array(nf[:]), dict(consumed=int(info[0]), steps=int(info[1]), undefined_zone=int(info[2]), n_pos=int(info[3]))

    def stream_path(self, key_ptrs, cnt_ptrs, n, min_rec, min_abund, chunk_records, nf, idxA, idxB, p_cut):
        """-> (stats dict, Matrix of the candidate rows or None, TtestResult over them or None)"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        nf = np.ascontiguousarray(nf, dtype=np.float64)
        a, b = (C.c_int * len(idxA))(*idxA), (C.c_int * len(idxB))(*idxB)
        st, om, ot = StreamStats(), C.c_void_p(0), C.c_void_p(0)
        self._chk(self.L.dk_stream_path(self.h, S, kp, cp, nn, min_rec, min_abund, int(chunk_records),
                                        nf.ctypes.data_as(C.POINTER(C.c_double)), a, len(idxA), b, len(idxB), float(p_cut),
                                        C.byref(st), C.byref(om), C.byref(ot)))
        stats = {k: getattr(st, k) for k, _ in StreamStats._fields_}
        if not om.value:
            return stats, None, None
        m = Matrix(self, om)
        return stats, m, TtestResult(self, ot, m)

    # ---- whole path -----------------------------------------------------------
    def run_path_matrix_out(self, key_ptrs, cnt_ptrs, n, device_inputs, min_rec, min_abund, nf_mode, idxA, idxB, pmax, fmin,
                            matrix_keys_ptr, matrix_cnts_ptr, cap_rows, nf_text_roundtrip=True):
        """run_path that also writes the filtered matrix (reference order) to host buffers given by pointer (pinned memory
        keeps the copy asynchronous); returns (stats, Matrix, TtestResult or None, rows written)"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        a, b = (C.c_int * len(idxA))(*idxA), (C.c_int * len(idxB))(*idxB)
        st, om, ot, rows = PathStats(), C.c_void_p(0), C.c_void_p(0), C.c_uint64(0)
        self._chk(self.L.dk_run_path_matrix_out(self.h, S, kp, cp, nn, int(device_inputs), min_rec, min_abund, nf_mode, a, len(idxA), b,
                                                len(idxB), float(pmax), float(fmin), int(nf_text_roundtrip), C.c_void_p(matrix_keys_ptr),
                                                C.c_void_p(matrix_cnts_ptr), int(cap_rows), C.byref(rows), C.byref(st), C.byref(om), C.byref(ot)))
        stats = dict(U=int(st.U), R=int(st.R), R_pos=int(st.R_pos), D=int(st.D), ties=tuple(int(x) for x in st.ties),
                     nf=np.array(st.nf[:S]), ms_join=st.ms_join, ms_nf=st.ms_nf, ms_ttest=st.ms_ttest, ms_bh=st.ms_bh)
        m = Matrix(self, om)
        t = TtestResult(self, ot, m) if ot.value else None
        if t is not None:
            t.D, t.ties = stats["D"], stats["ties"]
        return stats, m, t, int(rows.value)

    def run_path(self, key_ptrs, cnt_ptrs, n, device_inputs, min_rec, min_abund, nf_mode, idxA, idxB, pmax, fmin,
                 nf_text_roundtrip=True, keep=False, mask=None, mask_on_device=False, n_mask=None):
        """mask: sorted uint64 k-mers of the reference transcripts (numpy array, or a device pointer with n_mask when
        mask_on_device): RNA-Seq mode, NFs from the raw matrix and the test on the masked one (Snakefile:366,654)"""
        S, kp, cp, nn = self._tables(key_ptrs, cnt_ptrs, n)
        a, b = (C.c_int * len(idxA))(*idxA), (C.c_int * len(idxB))(*idxB)
        st = PathStats()
        om, ot = C.c_void_p(0), C.c_void_p(0)
        removed = C.c_uint64(0)
        if mask is None:
            self._chk(self.L.dk_run_path(self.h, S, kp, cp, nn, int(device_inputs), min_rec, min_abund, nf_mode, a, len(idxA), b,
                                         len(idxB), float(pmax), float(fmin), int(nf_text_roundtrip), C.byref(st),
                                         C.byref(om) if keep else None, C.byref(ot) if keep else None))
        else:
            if mask_on_device:
                mp, mn = C.c_void_p(int(mask)), int(n_mask)
            else:
                mask = np.ascontiguousarray(mask, dtype=np.uint64)
                mp, mn = _np_ptr(mask), len(mask)
            self._chk(self.L.dk_run_path_masked(self.h, S, kp, cp, nn, int(device_inputs), min_rec, min_abund, nf_mode, a,
                                                len(idxA), b, len(idxB), float(pmax), float(fmin), int(nf_text_roundtrip), mp, mn,
                                                int(mask_on_device), C.byref(st), C.byref(removed),
                                                C.byref(om) if keep else None, C.byref(ot) if keep else None))
        stats = dict(rows_masked=int(removed.value), U=int(st.U), R=int(st.R), R_pos=int(st.R_pos), D=int(st.D), ties=tuple(int(x) for x in st.ties),
                     nf=np.array(st.nf[:S]), ms_join=st.ms_join, ms_nf=st.ms_nf, ms_ttest=st.ms_ttest, ms_bh=st.ms_bh)
        if keep:
            m = Matrix(self, om)
            t = TtestResult(self, ot, m) if ot.value else None
            if t is not None:
                t.D, t.ties = stats["D"], stats["ties"]
            return stats, m, t
        return stats

"""helpers shared by the CPU and GPU tests"""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BASES = "ACGT"
CODE = {c: i for i, c in enumerate(BASES)}


def load_json(name):
    return json.load(open(os.path.join(GOLD, name)))


def enc(s):
    v = 0
    for ch in s:
        v = (v << 2) | CODE[ch]
    return v


def dec(v, k):
    v = int(v)
    return "".join(BASES[(v >> (2 * (k - 1 - i))) & 3] for i in range(k))


_LUT = np.full(256, 255, np.uint8)
for _i, _c in enumerate(b"ACGT"):
    _LUT[_c] = _i


def enc_many(kmers):
    """packed uint64 keys of a list of equal-length ACGT strings (vectorised enc)"""
    if not kmers:
        return np.zeros(0, np.uint64)
    k = len(kmers[0])
    a = np.frombuffer("".join(kmers).encode(), np.uint8)
    if len(a) != k * len(kmers):
        return np.array([enc(s) for s in kmers], dtype=np.uint64)            # ragged input: the slow exact way
    codes = _LUT[a].reshape(len(kmers), k)
    if (codes == 255).any():
        raise KeyError("non-ACGT base in a k-mer")
    v = np.zeros(len(kmers), np.uint64)
    for j in range(k):
        v = (v << np.uint64(2)) | codes[:, j].astype(np.uint64)
    return v


def parse_dump_text(txt):
    t = txt.split()
    kmers, cnts = t[0::2], t[1::2]
    k = len(kmers[0]) if kmers else 0
    return enc_many(kmers), np.array(cnts, dtype=np.int64).astype(np.int32) if cnts else np.zeros(0, np.int32), k


def format_join(keys, cnts, k):
    return "".join(dec(key, k) + "".join("\t%d" % c for c in row) + "\n" for key, row in zip(keys, cnts))


def random_dumps(rng, S, n_union, kbits, p_present, max_count, p_core=0.0):
    """sorted distinct keys per sample; a p_core fraction of rows is present in ~all samples"""
    keys = np.unique(rng.integers(0, 1 << kbits, size=n_union, dtype=np.uint64))
    core = rng.random(len(keys)) < p_core
    out_k, out_c = [], []
    for _ in range(S):
        m = np.where(core, rng.random(len(keys)) < 0.95, rng.random(len(keys)) < p_present)
        out_k.append(keys[m])
        out_c.append(rng.integers(1, max_count, size=int(m.sum())).astype(np.int32))
    return out_k, out_c


def plant_effect(rng, keys, cnts, group_b, frac=0.05, factor=40):
    """multiply the counts of the group-B samples on a random `frac` of the union k-mers: rows the t-test must find"""
    allk = np.unique(np.concatenate([k for k in keys if len(k)]))
    hit = allk[rng.random(len(allk)) < frac]
    out = []
    for s, (k, c) in enumerate(zip(keys, cnts)):
        c = c.copy()
        if s in group_b:
            c[np.isin(k, hit)] *= factor
        out.append(c)
    return out

"""2-bit k-mer packing: A=0 C=1 G=2 T=3, first base most significant (k <= 31), so unsigned
integer order equals the reference's strcmp order (share/mergeTags/dna.h:49-59)."""
import numpy as np

_CODE = {"A": 0, "C": 1, "G": 2, "T": 3}
_BASES = "ACGT"


def encode_kmer(s):
    v = 0
    for ch in s:
        v = (v << 2) | _CODE[ch]
    return v


def decode_kmer(v, k):
    v = int(v)
    return "".join(_BASES[(v >> (2 * (k - 1 - i))) & 3] for i in range(k))


def encode_kmers(strings):
    return np.array([encode_kmer(s) for s in strings], dtype=np.uint64)


def decode_kmers(arr, k):
    return [decode_kmer(v, k) for v in arr]

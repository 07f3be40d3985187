"""dekupl-run_b200 — B200-native count-matrix hot path of DE-kupl
(joinCounts -> computeNF -> TtestFilter) behind a C ABI (include/dekupl_b200.h).

This package is the Python host-side mirror used by tests/ and bench.py: a thin
ctypes binding over libdekupl_b200.so. The product itself is the shared library
plus the three drop-in CLIs in host/ (C++). There is NO CPU fallback: if the CUDA
library is missing or no B200 is present, calls raise.

The directory name has a hyphen (it is the name the build contract asks for), so
import it through `dekupl_run_b200.py` at the repo root.
"""
from .binding import (  # noqa: F401
    DkError, Context, Matrix, Dump, TtestResult, lib_path, load_library, declared_symbols,
    NF_REFERENCE, NF_FULL,
)
from .kmer import encode_kmer, decode_kmer, encode_kmers, decode_kmers  # noqa: F401
from . import tools  # noqa: F401

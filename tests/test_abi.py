"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol the header
declares, and fails loudly (no CPU fallback) when there is no GPU."""
import os
import subprocess

import pytest

import dekupl_run_b200 as dk

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    L = dk.load_library()
    decl = dk.declared_symbols()
    assert len(decl) >= 30
    missing = [s for s in decl if not hasattr(L, s)]
    assert missing == []


def test_exports_are_plain_c_symbols():
    out = subprocess.run(["nm", "-D", "--defined-only", dk.lib_path()], capture_output=True, text=True, check=True).stdout
    names = {l.split()[-1] for l in out.splitlines() if " T " in l}
    for s in dk.declared_symbols():
        assert s in names


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(dk.DkError) as e:
        dk.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_link_or_import_the_oracle():
    out = subprocess.run(["ldd", dk.lib_path()], capture_output=True, text=True).stdout
    assert "oracle" not in out
    pkg = os.path.join(ROOT, "dekupl-run_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in src and "liboracle" not in src and "orc_" not in src, f


def test_kmer_codec_roundtrip_and_order():
    s = ["AAAA", "AAAC", "ACGT", "CAAA", "TTTT"]
    v = [dk.encode_kmer(x) for x in s]
    assert v == sorted(v)
    assert [dk.decode_kmer(x, 4) for x in v] == s

"""CPU-side checks of the boundary: the in-tree library loads and exports every symbol include/geobpe_b200.h
declares; the product path refuses to run without CUDA (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "geobpe_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from geobpe_b200 import build, _lib
    lib = ctypes.CDLL(build.build())
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), s
    assert set(_lib.SIGNATURES) == set(syms)
    assert lib.gb_abi_version() == 1


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import numpy as np
    from geobpe_b200 import synthetic
    from geobpe_b200.bpe import BPE
    vals, offs = synthetic.generate(2, 8, seed=0)
    bpe = BPE(synthetic.to_structures(vals, offs), bins={1: 4}, res_init=True, rmsd_partition_min_size=float("inf"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        bpe.initialize()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "pt-bpe_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(d, f)).read()
                assert "geobpe_oracle" not in txt and "ref_shim" not in txt, f

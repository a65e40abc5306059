"""CPU: the C-ABI library builds, loads and exports every symbol include/mcb200.h declares; without a GPU it
fails loudly instead of falling back to anything."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build()
    from montecarlo_b200 import _lib
    return _lib.load()


def test_header_and_bindings_agree(lib):
    from montecarlo_b200 import _lib
    header = (ROOT / "include" / "mcb200.h").read_text()
    declared = set(re.findall(r"\b(mcb_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert getattr(lib, name) is not None


def test_struct_layouts_match_header():
    from montecarlo_b200._lib import McbChainStats, McbMcParams, McbParams
    assert C.sizeof(McbParams) == 8 * 4 + 3 * 8 + 2 * 8
    assert C.sizeof(McbMcParams) == 6 * 8 + 2 * 4 + 8
    assert C.sizeof(McbChainStats) == 16 * 8


def test_no_gpu_means_error_not_fallback(lib, shipped):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from montecarlo_b200.engine import LJEngine, McbError
    box, system = shipped
    with pytest.raises(McbError, match="no CUDA device"):
        LJEngine(system, box.n)


def test_product_does_not_touch_the_oracle():
    for p in (ROOT / "montecarlo_b200").rglob("*"):
        if p.suffix in (".py", ".cu", ".cuh", ".h"):
            text = p.read_text()
            assert "oracle" not in text.lower(), p
    assert "oracle" not in (ROOT / "include" / "mcb200.h").read_text().lower()

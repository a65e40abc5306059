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


def test_fortran_binding_covers_the_header():
    """bindings/mcb200.f90 (compile-unverified here: no Fortran compiler) binds every symbol the header declares,
    and INTEGRATION.md prints that same module."""
    header = (ROOT / "include" / "mcb200.h").read_text()
    f90 = (ROOT / "bindings" / "mcb200.f90").read_text()
    declared = set(re.findall(r"\b(mcb_[a-z0-9_]+)\s*\(", header))
    bound = set(re.findall(r'name="(mcb_[a-z0-9_]+)"', f90))
    assert declared == bound, declared ^ bound
    module = f90[f90.index("module mcb200"):]
    assert module in (ROOT / "INTEGRATION.md").read_text()


def test_struct_layouts_match_header():
    from montecarlo_b200._lib import McbChainStats, McbMcParams, McbParams
    assert C.sizeof(McbParams) == 8 * 4 + 3 * 8 + 2 * 8
    assert C.sizeof(McbMcParams) == 6 * 8 + 2 * 4 + 8
    assert C.sizeof(McbChainStats) == 19 * 8
    from montecarlo_b200._lib import McbGlobalStats, COMM_ID_BYTES
    assert C.sizeof(McbGlobalStats) == 12 * 8 and COMM_ID_BYTES == 128


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


def test_wrap_fma_matches_division():
    """k_run_spec wraps coordinates with wrap_fma (mcb_kernels.cuh): x - period*AINT(q) with q = q0 + r/half formed by
    two fmas instead of the reference's x/half (MonteCarlo.F90:705-712).  The host check of the identity AINT(q) ==
    AINT(x/half): random arguments and the adversarial ones, x within a few ulps of k*half, for the half-widths the
    kernels use (pi for the angles, BOXL for the box)."""
    libm = C.CDLL("libm.so.6")
    libm.fma.restype = C.c_double
    libm.fma.argtypes = [C.c_double] * 3
    rng = np.random.default_rng(7)
    for half in (np.pi, 7.7838685668, 15.5677371336, 23.3516057004, 9.3):
        inv = 1.0 / half
        xs = list((rng.random(4000) - 0.5) * 8.0 * half)
        for k in range(-4, 5):
            for ul in range(-4, 5):
                x = np.float64(k * half)
                for _ in range(abs(ul)):
                    x = np.nextafter(x, np.inf if ul > 0 else -np.inf)
                xs.append(float(x))
        for x in xs:
            q0 = x * inv
            r = libm.fma(-q0, half, x)
            q = libm.fma(r, inv, q0)
            assert np.trunc(q) == np.trunc(x / half), (x, half)
            if abs(x) > 1e-300:
                assert q == x / half, (x, half)

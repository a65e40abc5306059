"""Loader and ctypes signatures of libmcb200.so (include/mcb200.h).

The library is built in-tree by `__graft_entry__.build()` (nvcc, sm_100a).  If it is missing the import
fails loudly: there is no Python or CPU fallback for anything it computes.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
LIB_PATH = Path(os.environ.get("MCB200_LIB", PKG / "libmcb200.so"))  # override: debug builds only
SOURCES = [PKG / "csrc" / "mcb200.cu"]
HEADERS = [PKG / "csrc" / "mcb_device.cuh", PKG / "csrc" / "mcb_kernels.cuh", PKG / "csrc" / "mcb_verlet.cuh", PKG / "csrc" / "mcb_verlet2.cuh", PKG / "csrc" / "mcb_cells.cuh", ROOT / "include" / "mcb200.h"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-ldl"]

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_i64p = C.POINTER(C.c_int64)
_u64p = C.POINTER(C.c_uint64)
_u8p = C.POINTER(C.c_ubyte)


class McbParams(C.Structure):
    _fields_ = [("n_mol", C.c_int), ("n_sites", C.c_int), ("n_sol", C.c_int), ("n_chains", C.c_int),
                ("device", C.c_int), ("quirk_31", C.c_int), ("matrix_mode", C.c_int), ("lookahead", C.c_int),
                ("boxl", C.c_double), ("cutoff", C.c_double), ("conv", C.c_double),
                ("site_rel", _dp), ("site_tab", _dp)]


class McbMcParams(C.Structure):
    _fields_ = [("temperature", C.c_double), ("dposmax", C.c_double), ("dposmin", C.c_double),
                ("dhoekmax", C.c_double), ("dhoekmin", C.c_double), ("padj", C.c_double),
                ("nadj", C.c_int), ("reserved", C.c_int), ("start_prod", C.c_int64)]


class McbChainStats(C.Structure):
    _fields_ = [("toteng", C.c_double), ("dposmax", C.c_double), ("dhoekmax", C.c_double),
                ("production", C.c_double), ("sum_e", C.c_double), ("sum_e2", C.c_double),
                ("ratio", C.c_double), ("last_delta", C.c_double),
                ("step", C.c_int64), ("nsuc", C.c_int64), ("naccept", C.c_int64), ("prod_acc", C.c_int64),
                ("n_clamped", C.c_int64), ("sitepairs", C.c_int64), ("uniforms_used", C.c_int64),
                ("n_stat", C.c_int64), ("exch_attempts", C.c_int64), ("exch_accepted", C.c_int64),
                ("temperature", C.c_double)]


class McbGlobalStats(C.Structure):
    _fields_ = [("sum_e", C.c_double), ("sum_e2", C.c_double), ("production", C.c_double), ("sum_toteng", C.c_double),
                ("n_stat", C.c_int64), ("prod_acc", C.c_int64), ("nsuc", C.c_int64), ("step", C.c_int64),
                ("sitepairs", C.c_int64), ("n_clamped", C.c_int64), ("uniforms_used", C.c_int64), ("n_chains", C.c_int64)]


COMM_ID_BYTES = 128

# every symbol include/mcb200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "mcb_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(McbParams)]),
    "mcb_destroy": (C.c_int, [C.c_void_p]),
    "mcb_last_error": (C.c_char_p, [C.c_void_p]),
    "mcb_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mcb_launch_count": (C.c_int64, [C.c_void_p]),
    "mcb_set_solute": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp, C.c_int]),
    "mcb_set_chain_state": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp, C.c_int]),
    "mcb_get_chain_state": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp]),
    "mcb_total_energy": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp, _i64p]),
    "mcb_get_matrix": (C.c_int, [C.c_void_p, C.c_int, _dp, _dp]),
    "mcb_row_energy": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp]),
    "mcb_trial_delta": (C.c_int, [C.c_void_p, C.c_int, _ip, _dp, _dp, _dp, _dp]),
    "mcb_commit": (C.c_int, [C.c_void_p, C.c_int, _u8p]),
    "mcb_set_mc_params": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(McbMcParams), C.c_int]),
    "mcb_seed_chains": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _u64p]),
    "mcb_run_steps": (C.c_int, [C.c_void_p, C.c_int64, _dp, C.c_int64, _i64p, _i64p]),
    "mcb_get_stats": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(McbChainStats)]),
    "mcb_reset_stats": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "mcb_exchange": (C.c_int, [C.c_void_p, C.c_int, _ip, _ip, _dp, _u8p]),
    "mcb_measure_fp64_peak": (C.c_int, [C.c_int, C.c_double, _dp]),
    "mcb_comm_get_unique_id": (C.c_int, [C.c_char_p]),
    "mcb_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_int64]),
    "mcb_comm_attach": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64]),
    "mcb_comm_destroy": (C.c_int, [C.c_void_p]),
    "mcb_reduce_stats": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(McbGlobalStats), C.c_int]),
    "mcb_gather_chains": (C.c_int, [C.c_void_p, _dp, C.c_int64]),
    "mcb_box_energy_noalloc": (C.c_int, [C.POINTER(McbParams), C.c_void_p, C.c_void_p, _dp, _dp, C.c_void_p, C.c_void_p,
                                         C.c_char_p, C.c_size_t]),
}


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile libmcb200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = SOURCES + HEADERS
    if not force and LIB_PATH.exists() and all(LIB_PATH.stat().st_mtime >= s.stat().st_mtime for s in srcs):
        return LIB_PATH
    cmd = ["nvcc", *NVCC_FLAGS, "-o", str(LIB_PATH), *[str(s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(r.stderr)
    return LIB_PATH


_LIB = None


def load():
    """dlopen libmcb200.so and attach the signatures.  Raises if the library has not been built."""
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(the CUDA library is the product; there is no fallback)")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _LIB = lib
    return _LIB

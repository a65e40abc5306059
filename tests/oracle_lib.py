"""ctypes binding of oracle/libmc_oracle.so — the checker used by tests, smoke() and the CPU baseline.

TEST INFRASTRUCTURE ONLY: nothing under montecarlo_b200/ imports this.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
INPUTS = ROOT / "tests" / "golden" / "inputs"

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_u64p = C.POINTER(C.c_uint64)


def _ptr(a, t=_dp):
    return a.ctypes.data_as(t) if a is not None else None


def build_oracle() -> None:
    subprocess.run(["make", "-C", str(ORACLE_DIR), "libmc_oracle.so", "libmc_oracle_q.so"], check=True,
                   capture_output=True)


def _load(name: str):
    path = ORACLE_DIR / name
    if not path.exists():
        build_oracle()
    lib = C.CDLL(str(path))
    lib.mco_create.restype = C.c_void_p
    lib.mco_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, _dp, _ip, _dp, _dp, _dp,
                               C.c_double, C.c_double, C.c_int]
    lib.mco_destroy.argtypes = [C.c_void_p]
    lib.mco_set_state.argtypes = [C.c_void_p, _dp, _dp]
    lib.mco_get_state.argtypes = [C.c_void_p, _dp, _dp]
    lib.mco_set_mc_params.argtypes = [C.c_void_p] + [C.c_double] * 6 + [C.c_int, C.c_int64, C.c_int]
    lib.mco_calculate_lj.argtypes = [C.c_void_p, C.c_int]
    lib.mco_calc_energy.restype = C.c_double
    lib.mco_calc_energy.argtypes = [C.c_void_p]
    lib.mco_total_energy.restype = C.c_double
    lib.mco_total_energy.argtypes = [C.c_void_p]
    for f in (lib.mco_total_energy_matrixfree, lib.mco_total_energy_cells):
        f.restype = C.c_double
        f.argtypes = [C.c_void_p, _dp, _u64p]
    lib.mco_row_energy.restype = C.c_double
    lib.mco_row_energy.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
    lib.mco_get_matrix.argtypes = [C.c_void_p, _dp, _dp]
    lib.mco_mcinit.restype = C.c_int
    lib.mco_mcinit.argtypes = [C.c_void_p, _dp]
    lib.mco_step.restype = C.c_int
    lib.mco_step.argtypes = [C.c_void_p, _dp, _dp]
    lib.mco_step_given_move.restype = C.c_int
    lib.mco_step_given_move.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp, _dp]
    lib.mco_run_steps.restype = C.c_int64
    lib.mco_run_steps.argtypes = [C.c_void_p, C.c_int64, _dp, C.c_int64, C.POINTER(C.c_int64), _dp]
    lib.mco_run_steps_rng.restype = C.c_int64
    lib.mco_run_steps_rng.argtypes = [C.c_void_p, C.c_int64, _u64p, _dp]
    lib.mco_get_stats.argtypes = [C.c_void_p, _dp]
    lib.mco_rng_seed.argtypes = [_u64p, C.c_uint64]
    lib.mco_rng_uniform.restype = C.c_double
    lib.mco_rng_uniform.argtypes = [_u64p]
    lib.mco_exchange.restype = C.c_int
    lib.mco_exchange.argtypes = [C.c_void_p, C.c_void_p, C.c_double, _ip]
    lib.mco_rng_exchange_seed.argtypes = [_u64p, C.c_uint64]
    lib.mco_get_temperature.restype = C.c_double
    lib.mco_get_temperature.argtypes = [C.c_void_p]
    lib.mco_rotmatrix.argtypes = [_dp, _dp, _dp, C.c_int, _dp]
    lib.mco_lj_formula.argtypes = [C.c_double] * 6 + [_dp, C.c_double, C.c_double, _dp]
    for n in ("pi", "tau", "e", "hartree2kjmol", "conv_reference", "conv_intended"):
        getattr(lib, "mco_const_" + n).restype = C.c_double
    return lib


_LIBS: dict[str, object] = {}


def lib(quad: bool = False):
    name = "libmc_oracle_q.so" if quad else "libmc_oracle.so"
    if name not in _LIBS:
        _LIBS[name] = _load(name)
    return _LIBS[name]


TRACE_COLS = ("toteng_trial", "toteng_old", "kans", "rv", "rsolv", "accepted", "ratio", "dposmax", "dhoekmax",
              "sitepairs")
STAT_COLS = ("step", "nsuc", "naccept", "prod_acc", "nclamp", "sitepairs", "toteng", "toteng_old", "production",
             "dposmax", "dhoekmax", "ratio")


class OracleChain:
    """One copy of the reference program's LJ state, driven through the C restatement."""

    def __init__(self, system, com, hoek, *, quad: bool = False, keep_matrix: bool = True,
                 solute=None, sol_tab=None):
        self.L = lib(quad)
        self.n = int(com.shape[0])
        s = system
        self._keep = [np.ascontiguousarray(s.dmso_rel, dtype=np.float64),
                      np.ascontiguousarray(s.dmso_is_h, dtype=np.int32),
                      np.ascontiguousarray(s.dmso_tab, dtype=np.float64),
                      np.ascontiguousarray(s.solute if solute is None else solute, dtype=np.float64),
                      np.ascontiguousarray(s.sol_tab if sol_tab is None else sol_tab, dtype=np.float64)]
        k = self._keep
        self.h = C.c_void_p(self.L.mco_create(self.n, k[0].shape[0], k[3].shape[0], s.boxl, _ptr(k[0]),
                                              _ptr(k[1], _ip), _ptr(k[2]), _ptr(k[3]), _ptr(k[4]),
                                              s.conv, s.cutoff, 1 if keep_matrix else 0))
        self.set_state(com, hoek)

    def __del__(self):
        try:
            self.L.mco_destroy(self.h)
        except Exception:
            pass

    def set_state(self, com, hoek):
        com = np.ascontiguousarray(com, dtype=np.float64); hoek = np.ascontiguousarray(hoek, dtype=np.float64)
        self.L.mco_set_state(self.h, _ptr(com), _ptr(hoek))

    def get_state(self):
        com = np.zeros((self.n, 3)); hoek = np.zeros((self.n, 3))
        self.L.mco_get_state(self.h, _ptr(com), _ptr(hoek))
        return com, hoek

    def set_mc_params(self, temperature=300.0, dposmax=0.30, dposmin=1e-10, dhoekmax=0.16 * np.pi,
                      dhoekmin=1e-10 * np.pi, padj=0.05, nadj=100, start_prod=0, quirk31=1):
        self.L.mco_set_mc_params(self.h, temperature, dposmax, dposmin, dhoekmax, dhoekmin, padj, nadj,
                                 start_prod, quirk31)

    def total_energy(self) -> float:
        return self.L.mco_total_energy(self.h)

    def total_energy_matrixfree(self, cells: bool = False):
        es = C.c_double(0); npair = C.c_uint64(0)
        f = self.L.mco_total_energy_cells if cells else self.L.mco_total_energy_matrixfree
        e = f(self.h, C.byref(es), C.byref(npair))
        return e, es.value, npair.value

    def row_energy(self, i: int):
        row = np.zeros(self.n); ei = C.c_double(0)
        tot = self.L.mco_row_energy(self.h, i, _ptr(row), C.byref(ei))
        return tot, row, ei.value

    def matrix(self):
        ss = np.zeros((self.n, self.n)); en = np.zeros(self.n)
        self.L.mco_get_matrix(self.h, _ptr(ss), _ptr(en))
        return ss, en

    def step(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64); tr = np.zeros(10)
        used = self.L.mco_step(self.h, _ptr(u), _ptr(tr))
        return used, tr

    def step_given_move(self, i, com, hoek, u_accept):
        com = np.ascontiguousarray(com, dtype=np.float64); hoek = np.ascontiguousarray(hoek, dtype=np.float64)
        ua = np.array([u_accept], dtype=np.float64); tr = np.zeros(10)
        used = self.L.mco_step_given_move(self.h, int(i), _ptr(com), _ptr(hoek), _ptr(ua), _ptr(tr))
        return used, tr

    def run_steps(self, nsteps, uniforms, trace=True):
        u = np.ascontiguousarray(uniforms, dtype=np.float64)
        tr = np.zeros((nsteps, 10)) if trace else None
        consumed = C.c_int64(0)
        done = self.L.mco_run_steps(self.h, nsteps, _ptr(u), u.size, C.byref(consumed), _ptr(tr))
        return done, consumed.value, tr

    def run_steps_rng(self, nsteps, rng_state, trace=True):
        tr = np.zeros((nsteps, 10)) if trace else None
        self.L.mco_run_steps_rng(self.h, nsteps, _ptr(rng_state, _u64p), _ptr(tr))
        return tr

    def stats(self) -> dict:
        out = np.zeros(12)
        self.L.mco_get_stats(self.h, _ptr(out))
        return dict(zip(STAT_COLS, out))


def exchange(oa: "OracleChain", ob: "OracleChain", u: float):
    """The exchange rule on two oracle chains; returns (accepted, used_the_uniform)."""
    used = C.c_int(0)
    acc = oa.L.mco_exchange(oa.h, ob.h, float(u), C.byref(used))
    return bool(acc), bool(used.value)


def rng_exchange_seed(seed: int) -> np.ndarray:
    s = np.zeros(2, dtype=np.uint64)
    lib().mco_rng_exchange_seed(_ptr(s, _u64p), C.c_uint64(seed))
    return s


def rng_seed(seed: int) -> np.ndarray:
    s = np.zeros(2, dtype=np.uint64)
    lib().mco_rng_seed(_ptr(s, _u64p), C.c_uint64(seed))
    return s


def rng_uniforms(state: np.ndarray, count: int) -> np.ndarray:
    L = lib()
    p = _ptr(state, _u64p)
    return np.array([L.mco_rng_uniform(p) for _ in range(count)], dtype=np.float64)


def shipped_system():
    """The reference's shipped box + parameter files (copies under tests/golden/inputs)."""
    import sys
    sys.path.insert(0, str(ROOT))
    from montecarlo_b200 import boxio
    box = boxio.read_box(INPUTS / "box.txt")
    system = boxio.load_system(box, INPUTS / "DMSO.txt", INPUTS / "param.txt", INPUTS / "par_solute.txt")
    return box, system

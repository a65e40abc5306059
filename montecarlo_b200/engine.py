"""Host-side handle on the CUDA energy path, shaped after the reference's own procedures.

Reference procedure (MonteCarlo/MonteCarlo.F90)         here
  start-up `DO I: CALL calculateLJ(I)` + calcEnergy (:357-364)  LJEngine.total_energy()
  CALL calculateLJ(RSOLV) + calcEnergy for a trial (:411-414)   LJEngine.trial_delta()
  keep / roll back (:779-786)                                   LJEngine.commit()
  loop_LJ (:406-433) = MCINIT + calculateLJ + METROPOLIS        LJEngine.run_steps()
  `solventsolvent` dump (:372-376)                              LJEngine.matrix()

Everything numeric happens in libmcb200.so on the GPU; this file only marshals numpy arrays into the
C ABI (include/mcb200.h).  Molecule indices are 1-based, like RSOLV.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import McbChainStats, McbMcParams, McbParams

_dp = C.POINTER(C.c_double)


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


class McbError(RuntimeError):
    pass


class LJEngine:
    def __init__(self, system, n_mol: int, n_chains: int = 1, device: int = -1, quirk_31: bool = True,
                 matrix_mode: int = 0, with_solute: bool = True, lookahead: int = 0):
        self.lib = _lib.load()
        self.system = system
        self.n_mol, self.n_chains = int(n_mol), int(n_chains)
        self._site_rel = np.ascontiguousarray(system.site_rel, dtype=np.float64)
        self._site_tab = np.ascontiguousarray(system.site_tab, dtype=np.float64)
        self.n_sites = int(self._site_rel.shape[0])
        self.n_sol = int(system.solute.shape[0]) if with_solute else 0
        p = McbParams(n_mol=self.n_mol, n_sites=self.n_sites, n_sol=self.n_sol, n_chains=self.n_chains,
                      device=device, quirk_31=1 if quirk_31 else 0, matrix_mode=matrix_mode, lookahead=lookahead,
                      boxl=system.boxl, cutoff=system.cutoff, conv=system.conv,
                      site_rel=_d(self._site_rel), site_tab=_d(self._site_tab))
        self.h = C.c_void_p()
        rc = self.lib.mcb_create(C.byref(self.h), C.byref(p))
        if rc != 0:
            raise McbError(f"mcb_create failed ({rc}): {self.lib.mcb_last_error(None).decode()}")
        if self.n_sol:
            self.set_solute(system.solute, system.sol_tab)

    # -- plumbing
    def _ck(self, rc: int, what: str):
        if rc != 0:
            raise McbError(f"{what} failed ({rc}): {self.lib.mcb_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.lib.mcb_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream_ptr: int):
        self._ck(self.lib.mcb_set_stream(self.h, C.c_void_p(cuda_stream_ptr)), "mcb_set_stream")

    @property
    def launch_count(self) -> int:
        return int(self.lib.mcb_launch_count(self.h))

    def _range(self, chain0, nchains):
        return int(chain0), int(self.n_chains - chain0 if nchains is None else nchains)

    # -- state
    def set_solute(self, solute, sol_tab, chain0: int = 0, nchains: int | None = None):
        chain0, nchains = self._range(chain0, nchains)
        solute = np.ascontiguousarray(solute, dtype=np.float64)
        sol_tab = np.ascontiguousarray(sol_tab, dtype=np.float64)
        bc = 1 if solute.ndim == 2 else 0
        if not bc:
            assert solute.shape == (nchains, self.n_sol, 3) and sol_tab.shape == solute.shape
        self._ck(self.lib.mcb_set_solute(self.h, chain0, nchains, _d(solute), _d(sol_tab), bc), "mcb_set_solute")

    def set_state(self, com, hoek, chain0: int = 0, nchains: int | None = None):
        chain0, nchains = self._range(chain0, nchains)
        com = np.ascontiguousarray(com, dtype=np.float64)
        hoek = np.ascontiguousarray(hoek, dtype=np.float64)
        bc = 1 if com.ndim == 2 else 0
        assert com.shape[-2:] == (self.n_mol, 3) and hoek.shape == com.shape
        if not bc:
            assert com.shape[0] == nchains
        self._ck(self.lib.mcb_set_chain_state(self.h, chain0, nchains, _d(com), _d(hoek), bc), "mcb_set_chain_state")

    def get_state(self, chain0: int = 0, nchains: int | None = None, out=None):
        """CoM / HOEK of the chains as [chain][molecule][3].  `out=(com, hoek)` reuses the caller's arrays (a driver
        that keeps them page-locked - cudaHostRegister, or torch's pin_memory - gets the copy at full PCIe speed)."""
        chain0, nchains = self._range(chain0, nchains)
        if out is None:
            com = np.empty((nchains, self.n_mol, 3)); hoek = np.empty((nchains, self.n_mol, 3))
        else:
            com, hoek = out
            assert com.shape == hoek.shape == (nchains, self.n_mol, 3) and com.dtype == hoek.dtype == np.float64
            assert com.flags.c_contiguous and hoek.flags.c_contiguous
        self._ck(self.lib.mcb_get_chain_state(self.h, chain0, nchains, _d(com), _d(hoek)), "mcb_get_chain_state")
        return com, hoek

    # -- energies
    def total_energy(self, chain0: int = 0, nchains: int | None = None):
        """Start-up recompute (reference MonteCarlo.F90:357-364). Returns (TOTENG, sum ENERGY(i), pairs in cutoff)."""
        chain0, nchains = self._range(chain0, nchains)
        tot = np.empty(nchains); esol = np.empty(nchains); pairs = np.empty(nchains, dtype=np.int64)
        self._ck(self.lib.mcb_total_energy(self.h, chain0, nchains, _d(tot), _d(esol),
                                           pairs.ctypes.data_as(C.POINTER(C.c_int64))), "mcb_total_energy")
        return tot, esol, pairs

    def matrix(self, chain: int = 0):
        ss = np.empty((self.n_mol, self.n_mol)); en = np.empty(self.n_mol)
        self._ck(self.lib.mcb_get_matrix(self.h, chain, _d(ss), _d(en)), "mcb_get_matrix")
        return ss, en

    def row_energy(self, chain: int, i: int):
        """calculateLJ(I) without side effects (reference MonteCarlo.F90:792-828); i is 1-based."""
        row = np.empty(self.n_mol); ei = C.c_double(0.0)
        self._ck(self.lib.mcb_row_energy(self.h, chain, int(i), _d(row), C.byref(ei)), "mcb_row_energy")
        return row, ei.value

    # -- single-step mode (the driver keeps MCINIT / METROPOLIS)
    def trial_delta(self, rsolv, new_com, new_hoek):
        rsolv = np.ascontiguousarray(rsolv, dtype=np.int32)
        n = int(rsolv.shape[0])
        new_com = np.ascontiguousarray(new_com, dtype=np.float64).reshape(n, 3)
        new_hoek = np.ascontiguousarray(new_hoek, dtype=np.float64).reshape(n, 3)
        d_e = np.empty(n); e_new = np.empty(n)
        self._ck(self.lib.mcb_trial_delta(self.h, n, rsolv.ctypes.data_as(C.POINTER(C.c_int)), _d(new_com),
                                          _d(new_hoek), _d(d_e), _d(e_new)), "mcb_trial_delta")
        return d_e, e_new

    def commit(self, accept):
        accept = np.ascontiguousarray(accept, dtype=np.uint8)
        self._ck(self.lib.mcb_commit(self.h, int(accept.shape[0]), accept.ctypes.data_as(C.POINTER(C.c_ubyte))),
                 "mcb_commit")

    # -- batched mode
    def set_mc_params(self, temperature=300.0, dposmax=0.30, dposmin=1e-10, dhoekmax=0.16 * np.pi,
                      dhoekmin=1e-10 * np.pi, padj=0.05, nadj=100, start_prod=0, chain0: int = 0,
                      nchains: int | None = None):
        """Scalars are broadcast to the chain range; arrays of length nchains set per-chain values."""
        chain0, nchains = self._range(chain0, nchains)
        vals = [temperature, dposmax, dposmin, dhoekmax, dhoekmin, padj, nadj, start_prod]
        if all(np.ndim(v) == 0 for v in vals):
            p = McbMcParams(temperature, dposmax, dposmin, dhoekmax, dhoekmin, padj, int(nadj), 0, int(start_prod))
            self._ck(self.lib.mcb_set_mc_params(self.h, chain0, nchains, C.byref(p), 1), "mcb_set_mc_params")
            return
        cols = [np.broadcast_to(np.asarray(v), (nchains,)) for v in vals]
        arr = (McbMcParams * nchains)()
        for c in range(nchains):
            arr[c] = McbMcParams(*(float(col[c]) for col in cols[:6]), int(cols[6][c]), 0, int(cols[7][c]))
        self._ck(self.lib.mcb_set_mc_params(self.h, chain0, nchains, arr, 0), "mcb_set_mc_params")

    def seed(self, seeds, chain0: int = 0):
        seeds = np.ascontiguousarray(seeds, dtype=np.uint64)
        self._ck(self.lib.mcb_seed_chains(self.h, chain0, int(seeds.shape[0]),
                                          seeds.ctypes.data_as(C.POINTER(C.c_uint64))), "mcb_seed_chains")

    def run_steps(self, nsteps: int, uniforms=None, want_counts: bool = True):
        """loop_LJ on the GPU for every chain. uniforms: None (built-in stream) or [n_chains][cap] in [0,1)."""
        consumed = np.zeros(self.n_chains, dtype=np.int64) if want_counts else None
        steps = np.zeros(self.n_chains, dtype=np.int64) if want_counts else None
        i64 = C.POINTER(C.c_int64)
        cp = consumed.ctypes.data_as(i64) if want_counts else None
        sp = steps.ctypes.data_as(i64) if want_counts else None
        if uniforms is None:
            rc = self.lib.mcb_run_steps(self.h, int(nsteps), None, 0, cp, sp)
        else:
            u = np.ascontiguousarray(uniforms, dtype=np.float64)
            assert u.ndim == 2 and u.shape[0] == self.n_chains
            rc = self.lib.mcb_run_steps(self.h, int(nsteps), _d(u), int(u.shape[1]), cp, sp)
        self._ck(rc, "mcb_run_steps")
        return consumed, steps

    def reset_stats(self, chain0: int = 0, nchains: int | None = None):
        chain0, nchains = self._range(chain0, nchains)
        self._ck(self.lib.mcb_reset_stats(self.h, chain0, nchains), "mcb_reset_stats")

    def exchange(self, pair_a, pair_b, uniforms=None):
        """Replica exchange between launches: the reference's METROPOLIS rule (MonteCarlo.F90:728-744) on the
        parallel-tempering exponent; an accepted pair swaps temperatures and step sizes.  Returns accepted[p]."""
        pa = np.ascontiguousarray(pair_a, dtype=np.int32); pb = np.ascontiguousarray(pair_b, dtype=np.int32)
        n = int(pa.shape[0])
        assert pb.shape == pa.shape
        u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float64)
        assert u is None or u.shape == (n,)
        acc = np.zeros(n, dtype=np.uint8)
        ip = C.POINTER(C.c_int)
        self._ck(self.lib.mcb_exchange(self.h, n, pa.ctypes.data_as(ip), pb.ctypes.data_as(ip), _d(u),
                                       acc.ctypes.data_as(C.POINTER(C.c_ubyte))), "mcb_exchange")
        return acc.astype(bool)

    def stats(self, chain0: int = 0, nchains: int | None = None) -> dict:
        chain0, nchains = self._range(chain0, nchains)
        arr = (McbChainStats * nchains)()
        self._ck(self.lib.mcb_get_stats(self.h, chain0, nchains, arr), "mcb_get_stats")
        raw = np.frombuffer(arr, dtype=np.dtype([(n, "f8" if t is C.c_double else "i8") for n, t in McbChainStats._fields_]))
        return {n: raw[n].copy() for n in raw.dtype.names}


    # -- end of run: the job's one collective (NCCL inside libmcb200.so; PostProcess.F90:202-256 in the reference)
    def comm_init(self, nranks: int, rank: int, unique_id: bytes, n_total_chains: int):
        """Join the job's NCCL communicator; `unique_id` are the 128 bytes rank 0 got from comm_unique_id()."""
        assert len(unique_id) == _lib.COMM_ID_BYTES
        _preload_nccl()
        self._ck(self.lib.mcb_comm_init(self.h, int(nranks), int(rank), unique_id, int(n_total_chains)), "mcb_comm_init")
        self._n_total = int(n_total_chains)

    def comm_destroy(self):
        self._ck(self.lib.mcb_comm_destroy(self.h), "mcb_comm_destroy")
        self._n_total = None

    def reduce_stats(self, chains_per_group: int = 0) -> dict:
        """Sums over all chains of all ranks, per group of `chains_per_group` consecutive global chains (0: one group).
        Returns arrays of length n_groups plus the derived mean_e, var_e, acceptance."""
        n_total = getattr(self, "_n_total", None) or self.n_chains
        per = chains_per_group if chains_per_group > 0 else n_total
        ng = -(-n_total // per)
        arr = (_lib.McbGlobalStats * ng)()
        self._ck(self.lib.mcb_reduce_stats(self.h, int(chains_per_group), arr, ng), "mcb_reduce_stats")
        raw = np.frombuffer(arr, dtype=np.dtype([(n, "f8" if t is C.c_double else "i8") for n, t in _lib.McbGlobalStats._fields_]))
        out = {n: raw[n].copy() for n in raw.dtype.names}
        n = np.maximum(out["n_stat"], 1)
        out["mean_e"] = out["sum_e"] / n
        out["var_e"] = np.maximum(out["sum_e2"] / n - out["mean_e"] ** 2, 0.0)
        out["acceptance"] = out["nsuc"] / np.maximum(out["step"], 1)
        return out

    def gather_chains(self) -> dict:
        """Four numbers per chain of the whole job, in global chain order, on every rank."""
        n_total = getattr(self, "_n_total", None) or self.n_chains
        out = np.empty((n_total, 4))
        self._ck(self.lib.mcb_gather_chains(self.h, _d(out), n_total), "mcb_gather_chains")
        return {"toteng": out[:, 0].copy(), "mean_e": out[:, 1].copy(), "acceptance": out[:, 2].copy(),
                "production_mean": out[:, 3].copy()}


_NCCL_PRELOADED = False


def _preload_nccl():
    """Python hosts only.  libmcb200.so binds NCCL by soname at first use and reuses a copy the process has already
    loaded.  If PyTorch is imported LATER in the same process it would then be handed that copy too (same soname) and
    fail on symbols newer than the system NCCL has.  So before the library's first NCCL call, load the copy PyTorch
    ships (the `nvidia.nccl` wheel) when there is one; the library picks it up, and a later `import torch` is unaffected."""
    global _NCCL_PRELOADED
    if _NCCL_PRELOADED:
        return
    _NCCL_PRELOADED = True
    try:
        import importlib.util
        from pathlib import Path as _P
        spec = importlib.util.find_spec("nvidia.nccl")
        for loc in (spec.submodule_search_locations if spec else []):
            cand = _P(loc) / "lib" / "libnccl.so.2"
            if cand.exists():
                C.CDLL(str(cand), mode=C.RTLD_GLOBAL)
                return
    except (ImportError, OSError, ValueError):
        pass


def comm_unique_id() -> bytes:
    """The 128-byte NCCL id rank 0 creates and hands to the other ranks (by MPI, a file, torch.distributed ...)."""
    _preload_nccl()
    lib = _lib.load()
    buf = C.create_string_buffer(_lib.COMM_ID_BYTES)
    rc = lib.mcb_comm_get_unique_id(buf)
    if rc != 0:
        raise McbError(f"mcb_comm_get_unique_id failed ({rc}): is libnccl.so.2 installed?")
    return buf.raw


def measure_fp64_peak(device: int = -1, ms: float = 200.0) -> float:
    """Measured fp64 FMA peak in TFLOP/s (independent DFMA chains on every SM)."""
    lib = _lib.load()
    out = C.c_double(0.0)
    rc = lib.mcb_measure_fp64_peak(device, ms, C.byref(out))
    if rc != 0:
        raise McbError(f"mcb_measure_fp64_peak failed ({rc})")
    return out.value

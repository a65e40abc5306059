"""GPU: mcb_box_energy_noalloc — the entry point for hosts that are not allowed to allocate on the GPU.

Who needs it: a driver whose own memory pool has claimed the whole device at start-up, a process that runs under
a fixed memory cap, or a timing-critical loop that cannot take the device-wide stall a cudaMalloc causes.  They
cannot use mcb_create (it allocates the chain images).  This entry point takes device pointers the caller
already owns and keeps its only scratch in an array compiled into the library.

The test calls it through the C ABI exactly as such a host would (the tensors below stand for the caller's
pre-allocated memory) and compares with the regular entry point, mcb_total_energy, at the tolerance the other
GPU tests use (1e-10 relative).  It runs the call twice in a row, first on a large box and then on a small
one, so the second call finds the first call's leftovers in the shared scratch array.  It lives in a file that
sorts last so that a stop-at-first-failure run has already executed every other test.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from montecarlo_b200 import boxio

pytestmark = pytest.mark.gpu

RTOL = 1e-10


def _params(system, n_mol, n_sol, keep):
    from montecarlo_b200._lib import McbParams
    rel = np.ascontiguousarray(system.site_rel, dtype=np.float64)
    tab = np.ascontiguousarray(system.site_tab, dtype=np.float64)
    keep.extend([rel, tab])
    dp = C.POINTER(C.c_double)
    return McbParams(n_mol=n_mol, n_sites=rel.shape[0], n_sol=n_sol, n_chains=1, device=-1, quirk_31=1, matrix_mode=0,
                     lookahead=0, boxl=system.boxl, cutoff=system.cutoff, conv=system.conv,
                     site_rel=rel.ctypes.data_as(dp), site_tab=tab.ctypes.data_as(dp))


def _noalloc_energy(lib, system, box):
    import torch
    keep = []
    p = _params(system, box.n, box.nsol, keep)
    d_com = torch.from_numpy(np.ascontiguousarray(box.com)).cuda()
    d_hoek = torch.from_numpy(np.ascontiguousarray(box.hoek)).cuda()
    d_out = torch.full((3,), float("nan"), dtype=torch.float64, device="cuda")
    sol = np.ascontiguousarray(system.solute, dtype=np.float64)
    stab = np.ascontiguousarray(system.sol_tab, dtype=np.float64)
    dp = C.POINTER(C.c_double)
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    rc = lib.mcb_box_energy_noalloc(C.byref(p), C.c_void_p(d_com.data_ptr()), C.c_void_p(d_hoek.data_ptr()),
                                    sol.ctypes.data_as(dp), stab.ctypes.data_as(dp), C.c_void_p(d_out.data_ptr()),
                                    C.c_void_p(torch.cuda.current_stream().cuda_stream), err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    return d_out.cpu().numpy()


def _regular_energy(system, box):
    from montecarlo_b200.engine import LJEngine
    eng = LJEngine(system, box.n, n_chains=1)
    eng.set_state(box.com, box.hoek)
    tot, esol, pairs = eng.total_energy()
    return tot[0], esol[0], pairs[0]


def test_noalloc_matches_total_energy_twice_in_a_row(shipped):
    from montecarlo_b200 import _lib
    lib = _lib.load()
    box, system = shipped
    big = boxio.replicate(box, 5)                         # N = 3750: 30 tiles, 465 scratch slots
    sys_big = boxio.System(**{**system.__dict__, "boxl": big.boxl, "solute": big.solute})
    for b, s in ((big, sys_big), (box, system)):          # second call reuses the scratch the first one filled
        got = _noalloc_energy(lib, s, b)
        tot, esol, pairs = _regular_energy(s, b)
        assert abs(got[0] - tot) <= RTOL * abs(tot), (got[0], tot)
        assert abs(got[1] - esol) <= RTOL * abs(esol)
        assert got[2] == pairs
    assert abs(got[0] - 2.7968745663e3) <= 6e-11 * 2.8e3  # the reference's own printed start-up energy
    # and once more on the same box: identical bits (fixed-order sums over the shared scratch)
    assert np.array_equal(_noalloc_energy(lib, system, box), got)


def test_noalloc_rejects_boxes_beyond_its_fixed_scratch(shipped):
    from montecarlo_b200 import _lib
    lib = _lib.load()
    box, system = shipped
    keep = []
    err = C.create_string_buffer(256)
    p = _params(system, 32769, box.nsol, keep)
    dummy = C.c_void_p(16)   # never dereferenced: the size check comes first
    dp = C.POINTER(C.c_double)
    sol = np.ascontiguousarray(system.solute); stab = np.ascontiguousarray(system.sol_tab)
    rc = lib.mcb_box_energy_noalloc(C.byref(p), dummy, dummy, sol.ctypes.data_as(dp), stab.ctypes.data_as(dp), dummy, None, err, len(err))
    assert rc == 2 and b"MCB_NOALLOC_MAX_MOL" in err.value
    p = _params(system, 30, 25, keep)
    rc = lib.mcb_box_energy_noalloc(C.byref(p), dummy, dummy, sol.ctypes.data_as(dp), stab.ctypes.data_as(dp), dummy, None, err, len(err))
    assert rc == 2 and b"MCB_NOALLOC_MAX_SOL" in err.value

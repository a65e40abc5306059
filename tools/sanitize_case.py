"""Small end-to-end pass over every kernel, for `compute-sanitizer --tool memcheck python tools/sanitize_case.py`."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from montecarlo_b200 import boxio, _lib  # noqa: E402
from montecarlo_b200.engine import LJEngine  # noqa: E402

inp = ROOT / "tests" / "golden" / "inputs"
box = boxio.read_box(inp / "box.txt")
system = boxio.load_system(box, inp / "DMSO.txt", inp / "param.txt", inp / "par_solute.txt")
for la in (1, 2, 4, 8):
    eng = LJEngine(system, box.n, n_chains=5, lookahead=la)
    eng.set_state(box.com, box.hoek)
    print("lookahead", la, "E0", eng.total_energy()[0][0])
    eng.run_steps(60)
    eng.run_steps(20, np.random.default_rng(1).random((5, 200)))
    d, e = eng.trial_delta([3, 7, 30, 1, 12], np.tile(box.com[2], (5, 1)) + 0.01, np.tile(box.hoek[2], (5, 1)))
    eng.commit([1, 0, 1, 0, 1])
    eng.row_energy(2, 5); eng.matrix(1); eng.stats(); eng.get_state()
    eng.close()
b2 = boxio.replicate(box, 2)
s2 = boxio.System(**{**system.__dict__, "boxl": b2.boxl, "solute": b2.solute})
eng = LJEngine(s2, b2.n, n_chains=2)
eng.set_state(b2.com, b2.hoek); eng.run_steps(30); print("N=240", eng.stats()["toteng"]); eng.close()
# round-2 entry points: temperature ladder + exchange, end-of-run sums and gather (one rank), the one-warp list kernel
eng = LJEngine(system, box.n, n_chains=6)
eng.set_state(box.com, box.hoek); eng.set_mc_params(temperature=np.linspace(250.0, 500.0, 6)); eng.run_steps(40)
print("exchange", eng.exchange([0, 2, 4], [1, 3, 5]), eng.exchange([1, 3], [2, 4], np.array([0.3, 0.9])))
print("reduce", eng.reduce_stats()["nsuc"], eng.reduce_stats(4)["n_chains"], "gather", eng.gather_chains()["acceptance"][:2]); eng.reset_stats(); eng.close()
import os
os.environ["MCB200_NO_VL2"] = "1"
eng = LJEngine(s2, b2.n, n_chains=2)
eng.set_state(b2.com, b2.hoek); eng.run_steps(30); print("N=240, one-warp list kernel", eng.stats()["toteng"]); eng.close()
del os.environ["MCB200_NO_VL2"]
b5 = boxio.replicate(box, 5)
s5 = boxio.System(**{**system.__dict__, "boxl": b5.boxl, "solute": b5.solute})
eng = LJEngine(s5, b5.n, n_chains=1)
eng.set_state(b5.com, b5.hoek); print("N=3750 matrix-free", eng.total_energy()[0]); eng.close()
import torch
from montecarlo_b200._lib import McbParams
lib = _lib.load()
rel = np.ascontiguousarray(system.site_rel); tab = np.ascontiguousarray(system.site_tab)
dp = C.POINTER(C.c_double)
p = McbParams(n_mol=box.n, n_sites=4, n_sol=box.nsol, n_chains=1, device=-1, quirk_31=1, matrix_mode=0, lookahead=0,
              boxl=system.boxl, cutoff=7.0, conv=0.0, site_rel=rel.ctypes.data_as(dp), site_tab=tab.ctypes.data_as(dp))
dc = torch.from_numpy(box.com.copy()).cuda(); dh = torch.from_numpy(box.hoek.copy()).cuda(); out = torch.zeros(3, dtype=torch.float64, device="cuda")
sol = np.ascontiguousarray(system.solute); st = np.ascontiguousarray(system.sol_tab); err = C.create_string_buffer(256)
rc = lib.mcb_box_energy_noalloc(C.byref(p), C.c_void_p(dc.data_ptr()), C.c_void_p(dh.data_ptr()), sol.ctypes.data_as(dp), st.ctypes.data_as(dp),
                                C.c_void_p(out.data_ptr()), None, err, 256)
torch.cuda.synchronize()
print("noalloc", rc, out.cpu().numpy())
print("SANITIZE CASE DONE")

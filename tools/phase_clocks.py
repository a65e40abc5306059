"""Debug aid: per-phase cycle counts of k_run_steps (build with -DMCB_PHASE_CLOCKS). Not part of the product.

    nvcc ... -DMCB_PHASE_CLOCKS -o montecarlo_b200/libmcb200_prof.so montecarlo_b200/csrc/mcb200.cu
    MCB200_LIB=montecarlo_b200/libmcb200_prof.so python tools/phase_clocks.py
"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from montecarlo_b200 import boxio, _lib  # noqa: E402
from montecarlo_b200.engine import LJEngine  # noqa: E402

inputs = ROOT / "tests" / "golden" / "inputs"
box = boxio.read_box(inputs / "box.txt")
system = boxio.load_system(box, inputs / "DMSO.txt", inputs / "param.txt", inputs / "par_solute.txt")
nch, S = int(sys.argv[1]) if len(sys.argv) > 1 else 1024, 2000
la = int(sys.argv[2]) if len(sys.argv) > 2 else 0
k = int(sys.argv[3]) if len(sys.argv) > 3 else 1
if k > 1:
    box = boxio.replicate(box, k)
    system = boxio.System(**{**system.__dict__, "boxl": box.boxl, "solute": box.solute})
    S = 500
eng = LJEngine(system, box.n, n_chains=nch, lookahead=la)
eng.set_state(box.com, box.hoek)
eng.seed(np.arange(1, nch + 1, dtype=np.uint64))
eng.run_steps(S)
lib = _lib.load()
buf = (C.c_ulonglong * 8)()
lib.mcb_debug_phase_clocks(buf, 1)
eng.run_steps(S)
lib.mcb_debug_phase_clocks(buf, 0)
names = ["mcinit (rng, old row, move, sincos, rotate)", "centre-of-mass gate + list", "site-pair items", "reduce + metropolis", "bookkeeping", "exchange + resolution", "commit"]
tot = sum(buf[:7])
for n, v in zip(names, buf[:7]):
    print(f"{n:32s} {v / (nch * S):9.1f} cyc/step  {100.0 * v / tot:5.1f}%")
import os
print(f"{'total':32s} {tot / (nch * S):9.1f} cyc/step (per committed step, lookahead={la}, tune={os.environ.get('MCB200_TUNE')})")

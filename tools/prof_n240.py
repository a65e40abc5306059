"""ncu target: the N = 240 box, <chains> chains, two short launches of mcb_run_steps (tools only, not the product)."""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from montecarlo_b200 import boxio
from montecarlo_b200.engine import LJEngine
inputs = "/root/repo/tests/golden/inputs/"
box = boxio.read_box(inputs + "box.txt")
system = boxio.load_system(box, inputs + "DMSO.txt", inputs + "param.txt", inputs + "par_solute.txt")
k = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nch = int(sys.argv[2]) if len(sys.argv) > 2 else 512
S = int(sys.argv[3]) if len(sys.argv) > 3 else 500
b = boxio.replicate(box, k) if k > 1 else box
s = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute}) if k > 1 else system
eng = LJEngine(s, b.n, n_chains=nch)
eng.set_state(b.com, b.hoek); eng.seed(np.arange(1, nch + 1, dtype=np.uint64)); eng.total_energy()
for _ in range(3):
    eng.run_steps(S)
st = eng.stats()
print("steps", st["step"].sum(), "sitepairs", st["sitepairs"].sum())

"""Throughput of the batched mode on the 2x2x2 replica (N = 240; BASELINE configs[3] uses 512 chains)."""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
import torch
from montecarlo_b200 import boxio
from montecarlo_b200.engine import LJEngine
inputs = "/root/repo/tests/golden/inputs/"
box = boxio.read_box(inputs + "box.txt")
system = boxio.load_system(box, inputs + "DMSO.txt", inputs + "param.txt", inputs + "par_solute.txt")
k = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nch = int(sys.argv[2]) if len(sys.argv) > 2 else 512
S = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
b = boxio.replicate(box, k)
s = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute})
eng = LJEngine(s, b.n, n_chains=nch)
eng.set_stream(torch.cuda.current_stream().cuda_stream)
eng.set_state(b.com, b.hoek); eng.seed(np.arange(1, nch + 1, dtype=np.uint64)); eng.total_energy()
eng.run_steps(S, want_counts=False); torch.cuda.synchronize()
st0 = eng.stats()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): eng.run_steps(S, want_counts=False)
e1.record(); torch.cuda.synchronize()
st1 = eng.stats(); t = e0.elapsed_time(e1) * 1e-3
w = (st1["sitepairs"] - st0["sitepairs"]).sum(); m = (st1["step"] - st0["step"]).sum()
print(f"k={k} N={b.n} chains={nch} S={S}: {w/t:.4g} site-pairs/s {m/t:.4g} moves/s ({1e6*t/(3*S):.3f} us/step/chain-wave)")

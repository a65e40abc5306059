"""Where the end-to-end time of one bench step goes (host wall clock per call of the C ABI)."""
import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import torch
from montecarlo_b200 import boxio
from montecarlo_b200.engine import LJEngine
inputs = "/root/repo/tests/golden/inputs/"
box = boxio.read_box(inputs + "box.txt")
system = boxio.load_system(box, inputs + "DMSO.txt", inputs + "param.txt", inputs + "par_solute.txt")
k = int(sys.argv[1]) if len(sys.argv) > 1 else 1
nch = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
S = int(sys.argv[3]) if len(sys.argv) > 3 else 5000
if k > 1:
    box = boxio.replicate(box, k)
    system = boxio.System(**{**system.__dict__, "boxl": box.boxl, "solute": box.solute})
eng = LJEngine(system, box.n, n_chains=nch)
eng.set_stream(torch.cuda.current_stream().cuda_stream)
com0 = np.ascontiguousarray(np.broadcast_to(box.com, (nch, box.n, 3))); hoek0 = np.ascontiguousarray(np.broadcast_to(box.hoek, (nch, box.n, 3)))
eng.set_state(com0, hoek0); eng.seed(np.arange(1, nch + 1, dtype=np.uint64)); eng.total_energy()
eng.run_steps(S, want_counts=False); torch.cuda.synchronize()
com_h, hoek_h = eng.get_state()
T = {k: 0.0 for k in ("set_state", "prepare", "run", "stats", "get_state")}
R = 10
for _ in range(R):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    eng.set_state(com_h, hoek_h); torch.cuda.synchronize(); t1 = time.perf_counter()
    eng.total_energy(); torch.cuda.synchronize(); t1b = time.perf_counter()
    eng.run_steps(S, want_counts=False); torch.cuda.synchronize(); t2 = time.perf_counter()
    st = eng.stats(); t3 = time.perf_counter()
    com_h, hoek_h = eng.get_state(); t4 = time.perf_counter()
    T["set_state"] += t1 - t0; T["prepare"] += t1b - t1; T["run"] += t2 - t1b; T["stats"] += t3 - t2; T["get_state"] += t4 - t3
for k, v in T.items(): print(f"{k:10s} {1e3 * v / R:8.3f} ms")

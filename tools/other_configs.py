import sys, time, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from montecarlo_b200 import boxio
from montecarlo_b200.engine import LJEngine
inputs = "/root/repo/tests/golden/inputs/"
box = boxio.read_box(inputs + "box.txt")
system = boxio.load_system(box, inputs + "DMSO.txt", inputs + "param.txt", inputs + "par_solute.txt")
def run(k, nch, S, reps=3):
    b = boxio.replicate(box, k)
    s = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute})
    eng = LJEngine(s, b.n, n_chains=nch)
    eng.set_stream(torch.cuda.current_stream().cuda_stream)
    eng.set_state(b.com, b.hoek); eng.seed(np.arange(1, nch + 1, dtype=np.uint64)); eng.total_energy()
    eng.run_steps(S, want_counts=False); torch.cuda.synchronize()
    st0 = eng.stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): eng.run_steps(S, want_counts=False)
    e1.record(); torch.cuda.synchronize()
    st1 = eng.stats(); t = e0.elapsed_time(e1) * 1e-3
    w = (st1["sitepairs"] - st0["sitepairs"]).sum(); m = (st1["step"] - st0["step"]).sum()
    print(f"k={k} N={b.n} chains={nch}: {w/t:.4g} site-pairs/s {m/t:.4g} moves/s acc={(st1['nsuc']-st0['nsuc']).sum()/m:.3f} pairs/move={w/m:.1f}")
run(1, 1024, 2000)
run(2, 512, 1000)
run(2, 4096, 300)
# config 5 timing
b = boxio.replicate(box, 15)
s = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute})
eng = LJEngine(s, b.n, n_chains=1); eng.set_state(b.com, b.hoek)
tot, es, pr = eng.total_energy()
t0 = time.perf_counter(); tot, es, pr = eng.total_energy(); t1 = time.perf_counter()
w = 16 * pr[0] + 4 * 11 * b.n
print(f"config5 N={b.n}: total energy {tot[0]:.10e} in {1e3*(t1-t0):.2f} ms wall; {w} site pairs -> {w/(t1-t0):.4g} site-pairs/s; CoM tests {b.n*(b.n-1)//2:.3g} -> {b.n*(b.n-1)/2/(t1-t0):.4g} /s")

"""One-off soak: a few chains for 10^6 steps each on the GPU (several launches of uneven length) against the oracle
with the same seeds: counters equal, energies and coordinates within the parity tolerance.  Not part of the test
suite (the oracle needs ~10 s per chain); its output is kept under profiles/."""
import sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import oracle_lib as ol
from montecarlo_b200 import boxio
from montecarlo_b200.engine import LJEngine

inputs = ROOT / "tests" / "golden" / "inputs"
box = boxio.read_box(inputs / "box.txt")
system = boxio.load_system(box, inputs / "DMSO.txt", inputs / "param.txt", inputs / "par_solute.txt")
k = int(sys.argv[1]) if len(sys.argv) > 1 else 1
total = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
nch = 3
b = boxio.replicate(box, k) if k > 1 else box
s = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute}) if k > 1 else system
eng = LJEngine(s, b.n, n_chains=nch)
eng.set_state(b.com, b.hoek)
eng.set_mc_params(start_prod=total // 3)
seeds = np.array([1001, 1002, 1003], dtype=np.uint64)
eng.seed(seeds)
left, chunk = total, [1, 63, 64, 65, 4999, 100000, 333333]
t0 = time.perf_counter()
ci = 0
while left > 0:
    n = min(left, chunk[ci % len(chunk)]); ci += 1
    eng.run_steps(n, want_counts=False); left -= n
st = eng.stats(); com, hoek = eng.get_state()
t1 = time.perf_counter()
print(f"GPU: N={b.n}, {nch} chains x {total} steps in {ci} launches, {t1 - t0:.2f} s")
worst = 0.0
for c in range(nch):
    o = ol.OracleChain(s, b.com, b.hoek)
    o.total_energy()
    o.set_mc_params(start_prod=total // 3)
    o.run_steps_rng(total, ol.rng_seed(int(seeds[c])), trace=False)
    so = o.stats(); ocom, ohoek = o.get_state()
    assert st["step"][c] == so["step"] and st["nsuc"][c] == so["nsuc"] and st["sitepairs"][c] == so["sitepairs"], (c, st["nsuc"][c], so["nsuc"])
    assert st["prod_acc"][c] == so["prod_acc"] and st["naccept"][c] == so["naccept"]
    rel = abs(st["toteng"][c] - so["toteng"]) / abs(so["toteng"])
    dpos = max(np.abs(com[c] - ocom).max(), np.abs(hoek[c] - ohoek).max())
    worst = max(worst, rel)
    print(f"chain {c}: accepted {so['nsuc']} of {so['step']}, site pairs {so['sitepairs']}, TOTENG rel diff {rel:.2e}, "
          f"max coordinate diff {dpos:.2e}, production mean rel diff {abs(st['production'][c] - so['production']) / abs(so['production']):.2e}")
    assert rel < 1e-10 and dpos < 1e-10
print(f"oracle: {time.perf_counter() - t1:.1f} s; all counters equal, worst TOTENG rel diff {worst:.2e}")

"""Tuning aid: k_run_vl2 variants (image staged in shared memory / left in global memory, warps per chain) over box
sizes and chain counts.  One line per case; chain 0 is checked against the oracle (counters, energy) on the first launch.

    python tools/vl2_grid.py "k,chains,steps;..." "img:w;..."        e.g.  "2,512,1000;2,4096,300"  "smem:4;global:1;global:2"
"""
import os, sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from montecarlo_b200 import boxio
from montecarlo_b200.engine import LJEngine
import oracle_lib as ol

inputs = "/root/repo/tests/golden/inputs/"
box = boxio.read_box(inputs + "box.txt")
system = boxio.load_system(box, inputs + "DMSO.txt", inputs + "param.txt", inputs + "par_solute.txt")


def run(k, nch, S, img, w, reps=3):
    os.environ.pop("MCB200_VL2_IMG", None); os.environ.pop("MCB200_VL2_W", None)
    if img != "auto":
        os.environ["MCB200_VL2_IMG"] = img
        os.environ["MCB200_VL2_W"] = str(w)
    b = boxio.replicate(box, k)
    s = boxio.System(**{**system.__dict__, "boxl": b.boxl, "solute": b.solute})
    eng = LJEngine(s, b.n, n_chains=nch)
    eng.set_stream(torch.cuda.current_stream().cuda_stream)
    eng.set_state(b.com, b.hoek); eng.seed(np.arange(1, nch + 1, dtype=np.uint64)); eng.total_energy()
    eng.run_steps(S, want_counts=False); torch.cuda.synchronize()
    st0 = eng.stats()
    o = ol.OracleChain(s, b.com, b.hoek); o.total_energy(); o.run_steps_rng(S, ol.rng_seed(1), trace=False); so = o.stats()
    ok = (st0["nsuc"][0] == so["nsuc"] and st0["sitepairs"][0] == so["sitepairs"]
          and abs(st0["toteng"][0] - so["toteng"]) <= 1e-10 * abs(so["toteng"]))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): eng.run_steps(S, want_counts=False)
    e1.record(); torch.cuda.synchronize()
    st1 = eng.stats(); t = e0.elapsed_time(e1) * 1e-3
    wk = (st1["sitepairs"] - st0["sitepairs"]).sum(); m = (st1["step"] - st0["step"]).sum()
    print(f"k={k} N={b.n} chains={nch} steps={S} img={img} W={w}: {wk/t:.4g} site-pairs/s {m/t:.4g} moves/s oracle_ok={ok}", flush=True)
    eng.close()


cases = [tuple(int(x) for x in c.split(",")) for c in sys.argv[1].split(";")]
variants = [(v.split(":")[0], int(v.split(":")[1])) for v in sys.argv[2].split(";")]
for k, nch, S in cases:
    for img, w in variants:
        run(k, nch, S, img, w)

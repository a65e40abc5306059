"""GPU: the end-of-run collective through the C ABI (mcb_comm_init / mcb_reduce_stats / mcb_gather_chains).

One rank on one GPU always runs; the two-rank case (one process per GPU, the NCCL id handed over through a file, as a
Fortran driver without MPI would) runs where the box has two GPUs and is skipped otherwise.
"""
import os
import subprocess
import sys
import textwrap
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _ngpu():
    import torch
    return torch.cuda.device_count()


def test_reduce_and_gather_one_rank(shipped):
    from montecarlo_b200.engine import LJEngine, comm_unique_id
    box, system = shipped
    nch = 37
    eng = LJEngine(system, box.n, n_chains=nch)
    eng.set_state(box.com, box.hoek)
    eng.seed(np.arange(1, nch + 1, dtype=np.uint64))
    eng.set_mc_params(start_prod=500)
    eng.run_steps(1500)
    st = eng.stats()
    for with_comm in (False, True):
        if with_comm:
            eng.comm_init(1, 0, comm_unique_id(), nch)
        g = eng.reduce_stats()
        assert g["n_chains"][0] == nch and g["step"][0] == st["step"].sum() and g["nsuc"][0] == st["nsuc"].sum()
        assert g["sitepairs"][0] == st["sitepairs"].sum() and g["n_stat"][0] == st["n_stat"].sum()
        assert g["uniforms_used"][0] == st["uniforms_used"].sum() and g["prod_acc"][0] == st["prod_acc"].sum()
        for k, key in (("sum_e", "sum_e"), ("sum_e2", "sum_e2"), ("production", "production"), ("sum_toteng", "toteng")):
            assert abs(g[k][0] - st[key].sum()) <= 1e-12 * abs(st[key].sum())
        # groups of 8 chains: the last group is short
        gg = eng.reduce_stats(8)
        assert list(gg["n_chains"]) == [8, 8, 8, 8, 5]
        for q in range(5):
            sl = slice(8 * q, min(8 * q + 8, nch))
            assert gg["nsuc"][q] == st["nsuc"][sl].sum() and abs(gg["sum_e"][q] - st["sum_e"][sl].sum()) <= 1e-12 * abs(st["sum_e"][sl].sum())
        f = eng.gather_chains()
        assert np.array_equal(f["toteng"], st["toteng"])
        assert np.allclose(f["mean_e"], st["sum_e"] / st["n_stat"], rtol=1e-15)
        assert np.allclose(f["acceptance"], st["nsuc"] / st["step"], rtol=1e-15)
        assert np.allclose(f["production_mean"], st["production"] / np.maximum(st["prod_acc"], 1), rtol=1e-15)
    # the first chain is the oracle's chain
    o = ol.OracleChain(system, box.com, box.hoek)
    o.set_mc_params(start_prod=500)
    o.total_energy()
    o.run_steps_rng(1500, ol.rng_seed(1), trace=False)
    assert abs(f["toteng"][0] - o.stats()["toteng"]) <= 1e-10 * abs(f["toteng"][0])
    from montecarlo_b200.engine import McbError
    with pytest.raises(McbError, match="n_groups"):
        arr = (__import__("montecarlo_b200._lib", fromlist=["x"]).McbGlobalStats * 3)()
        eng._ck(eng.lib.mcb_reduce_stats(eng.h, 8, arr, 3), "mcb_reduce_stats")


_WORKER = textwrap.dedent("""
    import sys, time, json, os
    from pathlib import Path
    import numpy as np
    rank, world, n_total, idfile, out = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), Path(sys.argv[4]), Path(sys.argv[5])
    sys.path.insert(0, sys.argv[6]); sys.path.insert(0, sys.argv[6] + "/tests")
    import oracle_lib as ol
    from montecarlo_b200 import partition
    from montecarlo_b200.engine import LJEngine, comm_unique_id
    box, system = ol.shipped_system()
    start, count = partition.chain_block(rank, world, n_total)
    eng = LJEngine(system, box.n, n_chains=count, device=rank)
    if rank == 0:
        tmp = idfile.with_suffix(".tmp"); tmp.write_bytes(comm_unique_id()); tmp.rename(idfile)
    t0 = time.time()
    while not idfile.exists():
        assert time.time() - t0 < 120
        time.sleep(0.05)
    eng.comm_init(world, rank, idfile.read_bytes(), n_total)
    eng.set_state(box.com, box.hoek)
    eng.seed(partition.global_seeds(rank, world, n_total))
    eng.set_mc_params(start_prod=200)
    eng.run_steps(800)
    st = eng.stats()
    g = eng.reduce_stats()
    gg = eng.reduce_stats(4)
    f = eng.gather_chains()
    np.savez(out, nsuc=st["nsuc"], sum_e=st["sum_e"], toteng=st["toteng"], start=start,
             g_nsuc=g["nsuc"], g_sum_e=g["sum_e"], g_n=g["n_chains"], gg_nsuc=gg["nsuc"], gg_n=gg["n_chains"], f_toteng=f["toteng"])
    eng.comm_destroy()
""")


def test_two_ranks_over_nccl(tmp_path):
    if _ngpu() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    world, n_total = 2, 11                       # blocks of 6 and 5 chains
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    idfile = tmp_path / "nccl.id"
    procs = [subprocess.Popen([sys.executable, str(script), str(r), str(world), str(n_total), str(idfile),
                               str(tmp_path / f"r{r}.npz"), str(ROOT)], env={**os.environ, "NCCL_DEBUG": "WARN"})
             for r in range(world)]
    assert [p.wait(timeout=300) for p in procs] == [0, 0]
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    nsuc = np.concatenate([r0["nsuc"], r1["nsuc"]]); sum_e = np.concatenate([r0["sum_e"], r1["sum_e"]])
    tot = np.concatenate([r0["toteng"], r1["toteng"]])
    assert r0["start"] == 0 and r1["start"] == 6 and len(nsuc) == n_total
    for r in (r0, r1):                           # every rank receives everything
        assert r["g_n"][0] == n_total and r["g_nsuc"][0] == nsuc.sum()
        assert abs(r["g_sum_e"][0] - sum_e.sum()) <= 1e-12 * abs(sum_e.sum())
        assert list(r["gg_n"]) == [4, 4, 3] and list(r["gg_nsuc"]) == [nsuc[0:4].sum(), nsuc[4:8].sum(), nsuc[8:].sum()]
        assert np.array_equal(r["f_toteng"], tot)
    assert len(np.unique(tot)) == n_total        # seeds = global index + 1: all chains differ

"""GPU: the C++ stand-in for the Fortran driver (driver/mcdriver.cpp) — same config keys, same files.

Its OUT / DUMP / solventsolvent files are read with the parsers written for the reference binary's files
(oracle/run_ref.py), and every row is compared with the oracle stepping on the same uniform stream."""
import shutil
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "oracle"))


def _config(lj, seed, nprint=1, ndump=1, nadj=100):
    return "\n".join([f"LJ={lj}", "GA=0", "prod=200", f"seed={seed}", "temp=300.0", f"LJ_nadj={nadj}", f"LJ_nprint={nprint}",
                      f"LJ_dump={ndump}", "GA_nadj=10", "GA_nprint=1", "GA_dump=1", "dposmax=0.30", "dposmin=1.0E-10",
                      "dhoekmax=0.16", "dhoekmin=1.0E-10", "pAdj=0.05", "proc=1", "drotsolu=60.0", "# files",
                      "box=box.txt", "dmso=DMSO.txt", "param=param.txt", "parsol=par_solute.txt", "out=out.txt",
                      "err=err.txt", "dump=dump.txt", "solventsolvent=ss.txt", "result=result.txt", "solute=s", "solout=s", ""])


def _run(tmp_path, lj, seed, *args, **kw):
    exe = ROOT / "driver" / "mcdriver"
    if not exe.exists():   # normally built by __graft_entry__.build(); it only needs g++ and the built library
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", str(exe), str(ROOT / "driver" / "mcdriver.cpp"),
                        f"-L{ROOT / 'montecarlo_b200'}", "-lmcb200", "-Wl,-rpath,$ORIGIN/../montecarlo_b200"], check=True)
    for f in ol.INPUTS.iterdir():
        shutil.copy(f, tmp_path / f.name)
    (tmp_path / "config.ini").write_text(_config(lj, seed, **kw))
    r = subprocess.run([str(exe), "config.ini", *args], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


def test_single_step_driver_writes_reference_formats(tmp_path, shipped):
    import re
    box, system = shipped
    lj, seed = 600, 5
    out = _run(tmp_path, lj, seed)
    e0 = float(re.search(r"Initial energy \(LJ\):\s*(\S+)", out).group(1))
    assert abs(e0 - 2.7968745663e3) <= 6e-11 * abs(e0)
    # OUT rows: header + step 0 + one row per step, format 902
    lines = (tmp_path / "out.txt").read_text().splitlines()
    assert lines[0].split()[:3] == ["i", "TotEng", "TotEng_old"] and len(lines) == lj + 2
    assert re.fullmatch(r"\d{12} +-?\d\.\d{10}E[+-]\d\d +-?\d\.\d{10}E[+-]\d\d \d\.\d{4} \d\.\d{4} \d{3} \d\.\d{4} \d\.\d{4} \.\d{5} \.\d{5}", lines[5])
    rows = np.array([[float(x) for x in l.split()] for l in lines[1:]])
    # the oracle on the same stream
    o = ol.OracleChain(system, box.com, box.hoek)
    o.set_mc_params(start_prod=lj - 200 + 1)
    o.total_energy()
    tr = o.run_steps_rng(lj, ol.rng_seed(seed))
    assert np.all(np.abs(rows[1:, 1] - tr[:, 0]) <= 1.2e-10 * np.maximum(np.abs(tr[:, 0]), np.abs(tr[:, 1])))   # TotEng (11 digits)
    assert np.all(np.abs(rows[1:, 2] - tr[:, 1]) <= 1.2e-10 * np.abs(tr[:, 1]))                                   # TotEng_old
    assert np.all(np.abs(rows[1:, 3] - tr[:, 2]) <= 5.1e-5) and np.all(np.abs(rows[1:, 4] - tr[:, 3]) <= 5.1e-5)  # kans, rv
    assert np.array_equal(rows[1:, 5], tr[:, 4] + 1)                                                             # rSolv, 1-based
    assert np.all(np.abs(rows[1:, 8] - tr[:, 7]) <= 5.1e-6)                                                       # dposmax
    # DUMP: BOXL, frame 0, one TRIAL frame per step, final frame (parsed like the reference's dump)
    toks = (tmp_path / "dump.txt").read_text().replace("Timestep:", " ").split()
    assert float(toks[0]) == box.boxl
    nfr = (len(toks) - 1) // (2 + 6 * box.n)
    assert nfr == lj + 2
    fr = np.array(toks[1:], dtype=object).reshape(nfr, 2 + 6 * box.n)
    assert [int(x) for x in fr[:4, 1]] == [0, 1, 2, 3] and int(fr[-1, 1]) == lj + 2
    final = np.array(fr[-1, 2:], dtype=np.float64).reshape(box.n, 6)
    com, hoek = o.get_state()
    assert np.abs(final[:, :3] - com).max() < 1e-12 and np.abs(final[:, 3:] - hoek).max() < 1e-12
    # solventsolvent matrix and the result box (= next run's input) round-trip through our readers
    ss = np.array((tmp_path / "ss.txt").read_text().split(), dtype=np.float64).reshape(box.n, box.n)
    assert np.array_equal(ss, ss.T) and int((ss != 0).sum()) == 340
    from montecarlo_b200 import boxio
    rb = boxio.read_box(tmp_path / "result.txt")
    assert rb.n == box.n and np.abs(rb.com - com).max() < 1e-12 and rb.sol_sym == box.sol_sym and rb.dihedrals == box.dihedrals
    st = o.stats()
    assert f"Accepted {int(st['nsuc'])} of {lj} steps" in out
    assert f"Production: {int(st['prod_acc'])} accepted steps" in out


def test_batched_driver_matches_single_step_driver(tmp_path, shipped):
    """mode=batched runs loop_LJ on the GPU (mcb_run_steps); chain 0 must end where the host loop ends."""
    import re
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    out_s = _run(tmp_path / "a", 2000, 9, nprint=0, ndump=0)
    out_b = _run(tmp_path / "b", 2000, 9, "mode=batched", "chains=3", nprint=0, ndump=0)
    acc_s = int(re.search(r"Accepted (\d+) of", out_s).group(1))
    m = re.search(r"chain 0: steps 2000 accepted (\d+) energy\s+(\S+)", out_b)
    assert int(m.group(1)) == acc_s
    from montecarlo_b200 import boxio
    ra, rb = boxio.read_box(tmp_path / "a" / "result.txt"), boxio.read_box(tmp_path / "b" / "result.txt")
    assert np.abs(ra.com - rb.com).max() < 1e-12 and np.abs(ra.hoek - rb.hoek).max() < 1e-12
    assert abs(ra.energy - rb.energy) <= 1.2e-10 * abs(ra.energy)

"""Run the reference's own shipped binary (oracle/_ref/MonteCarloOpenMP) — TEST INFRASTRUCTURE ONLY.

This is the strongest oracle this repository has: the unmodified ifort build the
reference ships (reference MonteCarlo/MonteCarloOpenMP), started with a serial
stand-in for libiomp5 and a stub `g09` (oracle/ref_shim/).  It is used

  * by oracle/make_golden.py to produce the fixtures under tests/golden/,
  * by tests to re-validate the C restatement when oracle/_ref exists,
  * by bench.py's `--impl reference` / `cpu_baseline` legs as the CPU arm.

Nothing in montecarlo_b200/ imports this module.

What the binary gives us (reference MonteCarlo/MonteCarlo.F90):
  stdout  "Initial energy (LJ): ES20.10" (:367), "Final energy (LJ):" (:437), "Walltime:" (:677)
  OUT     one row per LJ_nprint steps, format 902 (:393, :762)
  DUMP    first line BOXL (:386), then per frame NCOM / "Timestep: i" / N x 6 doubles (:1009-1019);
          frames are the TRIAL configuration (dump precedes METROPOLIS, :416-431)
  SOLV    the full N x N pair matrix after the start-up recompute (:372-376)
"""
from __future__ import annotations

import os
import re
import shutil
import subprocess
import tempfile
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"
REF_BIN = REF_DIR / "MonteCarloOpenMP"
INPUTS = HERE.parent / "tests" / "golden" / "inputs"


def available() -> bool:
    return REF_BIN.exists() and (REF_DIR / "libiomp5.so").exists() and (REF_DIR / "bin" / "g09").exists()


@dataclass
class RefRun:
    initial_energy: float | None = None
    final_energy: float | None = None
    walltime: float | None = None
    stdout: str = ""
    out_rows: np.ndarray | None = None  # columns: i TotEng TotEng_old kans rv rSolv pSuc ratio dposmax dhoekmax
    dump_boxl: float | None = None
    dump_steps: list[int] = field(default_factory=list)
    dump_frames: np.ndarray | None = None  # [frames][N][6]
    matrix: np.ndarray | None = None  # [N][N]
    result_box: str | None = None


_FLT = r"[-+]?(?:\d+\.?\d*|\.\d+)(?:[EeDd][-+]?\d+)?"


def _config_text(lj, temp, nadj, nprint, ndump, dposmax, dposmin, dhoekmax, dhoekmin, padj, prod):
    # keys: reference MonteCarlo/readConfig.f90:89-161 (current names: drotsolu, not drotsolv)
    return "\n".join([
        f"LJ={lj}", "GA=0", f"prod={prod}", "seed=1", f"temp={temp:.4f}",
        f"LJ_nadj={nadj}", f"LJ_nprint={nprint}", f"LJ_dump={ndump}",
        "GA_nadj=10", "GA_nprint=1" if nprint else "GA_nprint=0", "GA_dump=1" if ndump else "GA_dump=0",
        f"dposmax={dposmax:.8f}", f"dposmin={dposmin:.3E}", f"dhoekmax={dhoekmax:.8f}", f"dhoekmin={dhoekmin:.3E}",
        f"pAdj={padj:.6f}", "proc=1", "drotsolu=60.0",
        "box=box.txt", "dmso=DMSO.txt", "param=param.txt", "solute=solute", "solout=solout",
        "out=out.txt", "err=err.txt", "dump=dump.txt", "solventsolvent=solventsolvent.txt",
        "result=result.txt", "parsol=par_solute.txt", "",
    ])


def run_reference(box_text: str | None = None, *, lj_steps: int = 1, temp: float = 300.0, nadj: int = 100,
                  nprint: int = 0, ndump: int = 0, dposmax: float = 0.30, dposmin: float = 1e-10,
                  dhoekmax: float = 0.16, dhoekmin: float = 1e-10, padj: float = 0.05, prod: int = 1,
                  charges: str | None = None, want_matrix: bool = True, keep_dir: str | None = None,
                  timeout: float = 3600.0) -> RefRun:
    """Run the shipped binary once in a scratch directory and parse everything it wrote."""
    if not available():
        raise RuntimeError("oracle/_ref is not built (run `make -C oracle ref` where /root/reference exists)")
    work = Path(keep_dir) if keep_dir else Path(tempfile.mkdtemp(prefix="mcref_"))
    work.mkdir(parents=True, exist_ok=True)
    try:
        for name in ("DMSO.txt", "param.txt", "par_solute.txt"):
            shutil.copy(INPUTS / name, work / name)
        if box_text is None:
            shutil.copy(INPUTS / "box.txt", work / "box.txt")
        else:
            (work / "box.txt").write_text(box_text)
        (work / "config.ini").write_text(
            _config_text(lj_steps, temp, nadj, nprint, ndump, dposmax, dposmin, dhoekmax, dhoekmin, padj, prod))
        env = dict(os.environ)
        env["LD_LIBRARY_PATH"] = f"{REF_DIR}:{env.get('LD_LIBRARY_PATH', '')}"
        env["PATH"] = f"{REF_DIR / 'bin'}:{env.get('PATH', '')}"
        env["MonteCarlo_WORKDIR"] = str(work / "gauss")
        env["OMP_NUM_THREADS"] = "1"
        if charges is not None:
            env["MCB_G09_CHARGES"] = charges
        proc = subprocess.run([str(REF_BIN), "config.ini"], cwd=work, env=env, capture_output=True,
                              text=True, timeout=timeout, errors="replace")
        res = RefRun(stdout=proc.stdout)
        m = re.search(r"Initial energy \(LJ\):\s*(" + _FLT + ")", proc.stdout)
        if m:
            res.initial_energy = float(m.group(1))
        m = re.search(r"Final energy \(LJ\):\s*(" + _FLT + ")", proc.stdout)
        if m:
            res.final_energy = float(m.group(1))
        m = re.search(r"Walltime:\s*(" + _FLT + ")", proc.stdout)
        if m:
            res.walltime = float(m.group(1))
        if res.initial_energy is None:
            raise RuntimeError(f"reference binary produced no LJ energy (rc={proc.returncode}):\n"
                               f"{proc.stdout[-2000:]}\n{proc.stderr[-2000:]}")
        out = work / "out.txt"
        if nprint and out.exists():
            rows = []
            for line in out.read_text().splitlines()[1:]:
                parts = line.split()
                if len(parts) == 10:
                    try:
                        rows.append([float(p) for p in parts])
                    except ValueError:  # '******' overflow fields
                        rows.append([float(p) if re.fullmatch(_FLT, p) else np.nan for p in parts])
            res.out_rows = np.array(rows, dtype=np.float64)
        dump = work / "dump.txt"
        if ndump and dump.exists():
            # ifort list-directed output wraps after 3 values, so parse by token, not by line
            toks = dump.read_text().replace("Timestep:", " ").split()
            res.dump_boxl = float(toks[0])
            frames, steps, i = [], [], 1
            while i < len(toks):
                n = int(toks[i])
                steps.append(int(toks[i + 1]))
                frames.append(np.array(toks[i + 2:i + 2 + 6 * n], dtype=np.float64).reshape(n, 6))
                i += 2 + 6 * n
            res.dump_steps = steps
            res.dump_frames = np.array(frames, dtype=np.float64)
        ss = work / "solventsolvent.txt"
        if want_matrix and ss.exists():
            vals = np.array(ss.read_text().split(), dtype=np.float64)
            n = int(round(np.sqrt(vals.size)))
            res.matrix = vals.reshape(n, n)
        rb = work / "result.txt"
        if rb.exists():
            res.result_box = rb.read_text()
        return res
    finally:
        if not keep_dir:
            shutil.rmtree(work, ignore_errors=True)

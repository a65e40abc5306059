"""Generate tests/golden/* from the reference's own shipped binary — TEST INFRASTRUCTURE ONLY.

Run in the build container (needs oracle/_ref, i.e. `make -C oracle ref`, which needs /root/reference):

    python oracle/make_golden.py

Everything written here is REFERENCE-PRODUCED (not restated): the fixtures are what pins the C
restatement in oracle/mc_oracle.c and, through it, the CUDA path.  The GPU box never runs this.

Fixtures
  ref_startup_k1.npz    start-up recompute on the shipped box: `Initial energy (LJ)` (11 digits) and the
                        30x30 pair matrix at 15 digits (reference MonteCarlo.F90:357-376)
  ref_startup_k2.npz    same for the periodic 2x2x2 replica (N=240)
  ref_startup_k5.json   scalars only for the 5x5x5 replica (N=3750): energy, sum of the upper triangle,
                        non-zero count
  ref_charges.json      start-up energy with non-zero CHELPG charges: identical (CONV == 0)
  ref_trace_10k.npz     a 10^4-step LJ run with print+dump every step, stored compactly: per step the
                        printed OUT row (TotEng, TotEng_old, kans, rv, rSolv, pSuc, ratio, dposmax, dhoekmax)
                        and the 6 coordinates of the moved molecule in the dumped TRIAL configuration
  ref_trace_adapt.npz   2 000 steps with LJ_nadj=20 and a large first step, to exercise the step-size
                        adaptation and its floor
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(ROOT))

import run_ref  # noqa: E402
from montecarlo_b200 import boxio  # noqa: E402

GOLD = ROOT / "tests" / "golden"


def compact_trace(r: run_ref.RefRun, box: boxio.Box):
    """Reduce OUT + DUMP of a run with nprint = ndump = 1 to per-step records, checking on the way that
    the dump is consistent with 'one molecule moves per step, rejected moves are rolled back'."""
    rows = r.out_rows
    assert rows[0, 0] == 0
    steps = rows[1:]
    nstep = steps.shape[0]
    frames = {s: f for s, f in zip(r.dump_steps, r.dump_frames)}
    cur = np.concatenate([box.com, box.hoek], axis=1)
    assert np.allclose(frames[0], cur, rtol=0, atol=1e-13 * 10)
    moved = np.zeros((nstep, 6))
    accepted = np.zeros(nstep, dtype=np.int8)
    for k in range(nstep):
        step = int(steps[k, 0])
        i = int(steps[k, 5]) - 1
        f = frames[step]
        others = np.delete(np.arange(f.shape[0]), i)
        assert np.abs(f[others] - cur[others]).max() < 1e-12, (step, np.abs(f[others] - cur[others]).max())
        moved[k] = f[i]
        # accept/reject is recoverable exactly: the next row's TotEng_old equals this row's TotEng
        if k + 1 < nstep:
            acc = steps[k + 1, 2] == steps[k, 1]
            # a rejected move whose trial energy prints identically to the old energy would be ambiguous
            if steps[k, 1] == steps[k, 2]:
                acc = True
        else:
            # the last frame of the file is the final ACCEPTED configuration (MonteCarlo.F90:581); the printed
            # `Final energy (LJ)` is TOTENG_OLD (:437), i.e. the energy before the last step, so it cannot tell
            acc = bool(np.array_equal(r.dump_frames[-1][i], f[i]))
        accepted[k] = acc
        if acc:
            cur[i] = f[i]
    return steps, moved, accepted


def main():
    assert run_ref.available(), "build oracle/_ref first: make -C oracle ref"
    GOLD.mkdir(parents=True, exist_ok=True)
    box = boxio.read_box(GOLD / "inputs" / "box.txt")

    r = run_ref.run_reference(lj_steps=1)
    np.savez_compressed(GOLD / "ref_startup_k1.npz", initial_energy=r.initial_energy, matrix=r.matrix)
    print("k1", r.initial_energy, (r.matrix != 0).sum())

    r2 = run_ref.run_reference(boxio.box_text(boxio.replicate(box, 2)), lj_steps=1)
    np.savez_compressed(GOLD / "ref_startup_k2.npz", initial_energy=r2.initial_energy, matrix=r2.matrix)
    print("k2", r2.initial_energy, (r2.matrix != 0).sum(), np.triu(r2.matrix, 1).sum())

    r5 = run_ref.run_reference(boxio.box_text(boxio.replicate(box, 5)), lj_steps=1)
    (GOLD / "ref_startup_k5.json").write_text(json.dumps({
        "initial_energy": r5.initial_energy, "sum_upper": float(np.triu(r5.matrix, 1).sum()),
        "nonzero": int((r5.matrix != 0).sum()), "n": int(r5.matrix.shape[0])}, indent=1))
    print("k5", r5.initial_energy, (r5.matrix != 0).sum())

    q = "0.35 0.10 -0.20 -0.30 0.08 0.09 0.10 0.70 -0.55 -0.60 0.43"
    rq = run_ref.run_reference(lj_steps=1, charges=q)
    (GOLD / "ref_charges.json").write_text(json.dumps({
        "charges": [float(x) for x in q.split()], "initial_energy": rq.initial_energy,
        "matrix_equal_to_uncharged": bool(np.array_equal(rq.matrix, r.matrix))}, indent=1))
    print("charges", rq.initial_energy, np.array_equal(rq.matrix, r.matrix))

    rt = run_ref.run_reference(lj_steps=10000, nprint=1, ndump=1)
    steps, moved, acc = compact_trace(rt, box)
    np.savez_compressed(GOLD / "ref_trace_10k.npz", initial_energy=rt.initial_energy,
                        final_energy=rt.final_energy, out_rows=steps, moved=moved, accepted=acc,
                        params=np.array([300.0, 0.30, 1e-10, 0.16, 1e-10, 0.05, 100.0]))
    print("trace", steps.shape, int(acc.sum()), "accepted; final", rt.final_energy)

    ra = run_ref.run_reference(lj_steps=2000, nprint=1, ndump=1, nadj=20, dposmax=0.9, dposmin=0.05,
                               dhoekmax=0.30, dhoekmin=0.02, padj=0.25)
    steps, moved, acc = compact_trace(ra, box)
    np.savez_compressed(GOLD / "ref_trace_adapt.npz", initial_energy=ra.initial_energy,
                        final_energy=ra.final_energy, out_rows=steps, moved=moved, accepted=acc,
                        params=np.array([300.0, 0.9, 0.05, 0.30, 0.02, 0.25, 20.0]))
    print("adapt", steps.shape, int(acc.sum()), "accepted; dposmax range", steps[:, 8].min(), steps[:, 8].max())


if __name__ == "__main__":
    main()

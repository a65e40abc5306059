"""What happens to the gathered per-window energies: the reference's window hopping and solute-level Metropolis.

Restates PostProcess/PostProcess.F90:260-304 (COMPARE) and MonteCarlo/solmod.F90:63-98 (SOLUTE_METROPOLIS) on the
host.  In the reference these consume energies scraped from run logs of separate processes; here they consume
the per-conformer statistics that the end-of-run NCCL gather collects (config 4).  Uniforms are an argument:
the reference draws them from the compiler's RANDOM_NUMBER.

Note the two routines differ on purpose, like their originals: COMPARE uses EXP() and a strict `RV < KANS`,
SOLUTE_METROPOLIS uses the float32-rounded E ** x and `RV <= KANS`.
"""
from __future__ import annotations

import math
from typing import Iterator

import numpy as np

E_F32 = float(np.float32(2.71828182845904523536))   # MC_lib/lib.f90:17


def _kans_rv(exponent: float, draw, use_exp: bool):
    if exponent < -75.0:
        return 0.0, 1.0            # no draw
    if exponent >= 0.0:
        return 1.0, 0.0            # no draw
    kans = math.exp(exponent) if use_exp else E_F32 ** exponent
    return min(kans, 1.0), float(draw())


def solute_metropolis(pre_en: float, post_en: float, temp: float, draw) -> bool:
    """SOLUTE_METROPOLIS (solmod.F90:63-98): accept the new solute conformer?  `draw()` yields a uniform."""
    exponent = -1.0 * (post_en - pre_en) * 1000.0 / (8.314 * temp)
    kans, rv = _kans_rv(exponent, draw, use_exp=False)
    return rv <= kans


def window_hopping(energy: np.ndarray, temperature: float, uniforms: Iterator[float], literal: bool = False):
    """COMPARE (PostProcess.F90:260-304).  energy[run, window]; starts in the window with the lowest energy of the
    first run; per run one uniform picks the neighbour window (cyclic), a second one (only when needed) decides.
    Returns (old, now, accepted) per run, windows 0-based.

    Indexing.  The reference allocates ENERGY(NUMLOG, RUNS) (:66) and fills it ENERGY(log, run) (:249), but COMPARE
    reads ENERGY(I, NEXT) and ENERGY(I, CURRENT) with I the RUN and NEXT / CURRENT the window (:277): the two
    subscripts are transposed.  With Fortran's column-major storage and no bounds checking that is the linear element
    window * NUMLOG + run (0-based), i.e. the energy of run (window * NUMLOG + run) // NUMLOG in window
    (window * NUMLOG + run) % NUMLOG - some other run's energy in some other window.  The first subscript exceeds its
    declared bound as soon as run >= NUMLOG; the linear element still lies inside the array unless there are fewer
    runs than windows (window * NUMLOG + run >= NUMLOG * RUNS).
      literal=False (default)  what the routine means: energy[run, window].
      literal=True             what the shipped source computes: the linear element above; IndexError where the
                               reference would read outside its array (undefined there).
    No output of the reference's PostProcess exists to pin either reading (it cannot be built here and ships no
    result file), so this function's parity is UNPINNED; both readings are restated from the source text."""
    energy = np.asarray(energy, dtype=np.float64)
    runs, nwin = energy.shape
    flat = energy.ravel()            # [run, window] row-major == ENERGY(window, run) column-major: the reference's storage order

    def en(i, w):
        if not literal:
            return energy[i, w]
        k = w * nwin + i
        if k >= flat.size:
            raise IndexError(f"COMPARE reads ENERGY({i + 1},{w + 1}) outside ENERGY({nwin},{runs}) (PostProcess.F90:277)")
        return flat[k]

    draw = lambda: next(uniforms)  # noqa: E731
    current = int(np.argmin(energy[0]))          # MINLOC(ENERGY(:,1)) (:264): this one is indexed as intended
    old = np.zeros(runs, dtype=np.int64); now = np.zeros(runs, dtype=np.int64); acc = np.zeros(runs, dtype=bool)
    for i in range(runs):
        nxt = current + 1 if draw() < 0.5 else current - 1
        if nxt > nwin - 1:
            nxt = 0
        if nxt < 0:
            nxt = nwin - 1
        exponent = -1.0 * (en(i, nxt) - en(i, current)) * 1000.0 / (8.314 * temperature)
        kans, rv = _kans_rv(exponent, draw, use_exp=True)
        old[i], now[i] = current, nxt
        if rv < kans:
            acc[i] = True
            current = nxt
    return old, now, acc

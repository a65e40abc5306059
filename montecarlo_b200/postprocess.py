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


def window_hopping(energy: np.ndarray, temperature: float, uniforms: Iterator[float]):
    """COMPARE (PostProcess.F90:260-304).  energy[run, window]; starts in the window with the lowest energy of the
    first run; per run one uniform picks the neighbour window (cyclic), a second one (only when needed) decides.
    Returns (old, now, accepted) per run, windows 0-based."""
    energy = np.asarray(energy, dtype=np.float64)
    runs, nwin = energy.shape
    draw = lambda: next(uniforms)  # noqa: E731
    current = int(np.argmin(energy[0]))
    old = np.zeros(runs, dtype=np.int64); now = np.zeros(runs, dtype=np.int64); acc = np.zeros(runs, dtype=bool)
    for i in range(runs):
        nxt = current + 1 if draw() < 0.5 else current - 1
        if nxt > nwin - 1:
            nxt = 0
        if nxt < 0:
            nxt = nwin - 1
        exponent = -1.0 * (energy[i, nxt] - energy[i, current]) * 1000.0 / (8.314 * temperature)
        kans, rv = _kans_rv(exponent, draw, use_exp=True)
        old[i], now[i] = current, nxt
        if rv < kans:
            acc[i] = True
            current = nxt
    return old, now, acc

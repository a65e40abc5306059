"""Helpers shared by the CPU (oracle vs reference fixtures) and GPU (CUDA vs oracle/fixtures) tests."""
from __future__ import annotations

from pathlib import Path

import numpy as np

GOLD = Path(__file__).resolve().parent / "golden"

# OUT row columns (reference MonteCarlo.F90:762, format 902)
C_I, C_TOT, C_OLD, C_KANS, C_RV, C_RSOLV, C_PSUC, C_RATIO, C_DPOS, C_DHOEK = range(10)


def load_trace(name: str):
    z = np.load(GOLD / name)
    return {k: z[k] for k in z.files}


def recover_uniforms(cur6, moved6, dposmax, dhoekmax, boxl, rsolv1, n, accepted, kans_drawn):
    """Invert MCINIT (reference MonteCarlo.F90:697-712) for one step of a reference trace: the 7 uniforms
    that turn the current molecule into the dumped trial molecule (to the 15 digits the dump prints), plus
    an 8th that forces the recorded decision when METROPOLIS draws."""
    u = np.zeros(8)
    u[0] = (rsolv1 - 0.5) / n
    sc = dposmax / np.sqrt(3.0)
    for k in range(3):
        d = moved6[k] - cur6[k]
        cands = [d, d + 2 * boxl, d - 2 * boxl]
        d = min(cands, key=abs)
        u[1 + k] = d / (2.0 * sc) + 0.5
    for k in range(3):
        d = moved6[3 + k] - cur6[3 + k]
        cands = [d, d + 2 * np.pi, d - 2 * np.pi]
        d = min(cands, key=abs)
        u[4 + k] = d / (2.0 * dhoekmax) + 0.5
    u[7] = 0.0 if accepted else 1.0
    return np.clip(u, 0.0, np.nextafter(1.0, 0.0))

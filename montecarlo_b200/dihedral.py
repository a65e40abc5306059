"""Solute dihedral angles on the host: input generation for BASELINE config 4 (one chain family per conformer).

Restates the reference's GETDIHEDRAL / SETDIHEDRAL (MC_lib/dihedral.f90:9-46, :51-187) and FIND
(MonteCarlo/solmod.F90:99-120).  Nothing here is on the GPU path; the conformers it makes are handed to
mcb_set_solute as per-chain solute coordinates.
"""
from __future__ import annotations

import numpy as np

BONDL = float(np.float32(2.05))   # default-REAL parameter, dihedral.f90:79
BONDL2 = BONDL * BONDL


def get_dihedral(mol: np.ndarray, i1: int, i2: int, i3: int, i4: int) -> float:
    """GETDIHEDRAL (dihedral.f90:9-46): signed angle in radians; atom indices are 1-based like the reference's."""
    v1, v2, v3, v4 = (np.asarray(mol[i - 1], dtype=np.float64) for i in (i1, i2, i3, i4))
    b1, b2, b3 = v2 - v1, v3 - v2, v4 - v3
    n1, n2 = np.cross(b1, b2), np.cross(b2, b3)
    t = np.dot(n1, n2) / (np.linalg.norm(n1) * np.linalg.norm(n2))
    h = float(np.arccos(np.clip(t, -1.0, 1.0)))
    return -h if np.dot(b1, n2) < 0.0 else h


def find_neighbour(a1: int, a2: int, sol: np.ndarray) -> int:
    """FIND (solmod.F90:99-120): first atom other than a1, a2 closer than 1.90 A to a1 (1-based)."""
    for i in range(1, sol.shape[0] + 1):
        if i != a1 and i != a2 and np.linalg.norm(sol[i - 1] - sol[a1 - 1]) < float(np.float32(1.90)):
            return i
    raise ValueError(f"no bonded neighbour found for atom {a1} (excluding {a2})")


def dihedral_atoms(sol: np.ndarray, bond: tuple[int, int]) -> tuple[int, int, int, int]:
    """A1-A2-A3-A4 for a rotatable bond (A2, A3) as SOLUTE_INIT builds them (solmod.F90:39-44)."""
    a2, a3 = bond
    return find_neighbour(a2, a3, sol), a2, a3, find_neighbour(a3, a2, sol)


def fragments(mol: np.ndarray, sym: list[str], i2: int, i3: int) -> np.ndarray:
    """Fragment assignment of SETDIHEDRAL (dihedral.f90:95-152): 0 = axis atoms, 1 = side of I2, 2 = side of I3."""
    n = mol.shape[0]
    frag = np.full(n, 99, dtype=np.int64)
    frag[i2 - 1] = 0
    frag[i3 - 1] = 0
    v2, v3 = mol[i2 - 1], mol[i3 - 1]
    for i in range(n):
        if frag[i] == 0:
            continue
        d1 = float(np.sum((mol[i] - v2) ** 2))
        d2 = float(np.sum((mol[i] - v3) ** 2))
        if sym[i] == "H":
            if d1 < 2.25 and d2 >= 2.25:
                frag[i] = 1
            elif d1 >= 2.25 and d2 < 2.25:
                frag[i] = 2
        else:
            if d1 < BONDL2 and d2 >= 4.21:
                frag[i] = 1
            elif d1 >= 4.21 and d2 < BONDL2:
                frag[i] = 2
    for _ in range(n + 1):   # flood fill along bonds (findAtoms loop)
        if not np.any(frag == 99):
            break
        for i in range(n):
            if frag[i] != 99:
                continue
            for j in range(n):
                if j != i and frag[j] in (1, 2) and float(np.sum((mol[i] - mol[j]) ** 2)) < BONDL2:
                    frag[i] = frag[j]
                    break
    if np.any(frag == 99):
        raise ValueError("SETDIHEDRAL: atoms not connected to either fragment")
    return frag


def set_dihedral(mol: np.ndarray, sym: list[str], i1: int, i2: int, i3: int, i4: int, hoek: float) -> np.ndarray:
    """SETDIHEDRAL (dihedral.f90:51-187): rotate fragment 2 about the axis I2->I3 by the INCREMENT `hoek` (radians),
    Rodrigues' formula around the point I3 (:177-185)."""
    mol = np.asarray(mol, dtype=np.float64)
    out = mol.copy()
    v2, v3 = mol[i2 - 1], mol[i3 - 1]
    k = (v3 - v2) / np.linalg.norm(v3 - v2)
    c, s = np.cos(hoek), np.sin(hoek)
    frag = fragments(mol, sym, i2, i3)
    for i in np.nonzero(frag == 2)[0]:
        v = mol[i] - v3
        out[i] = v * c + np.cross(k, v) * s + k * np.dot(k, v) * (1.0 - c) + v3
    return out


def conformer_sweep(sol: np.ndarray, sym: list[str], bond: tuple[int, int], targets_deg) -> tuple[np.ndarray, tuple]:
    """Conformers with the dihedral SET TO each target (degrees): rotate by target - current (changeDih.f90:42-66)."""
    atoms = dihedral_atoms(sol, bond)
    cur = get_dihedral(sol, *atoms)
    out = np.stack([set_dihedral(sol, sym, *atoms, np.deg2rad(t) - cur) for t in targets_deg])
    return out, atoms

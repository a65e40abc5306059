"""Input files of the Fortran driver, read and written on the host.

The driver keeps its own formats (reference MonteCarlo/MonteCarlo.F90:1053-1166 `read_norm`,
:587-611 result writer, :1009-1019 `dump`); this module reads and writes the same text so that
C++/Python harnesses can stand in for the driver where no Fortran compiler exists, and so the
inputs of configs 3-5 (replicated boxes) can be generated.  Nothing here computes an energy.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from pathlib import Path

import numpy as np

ENERGY_ABSENT = 123456.789  # sentinel, reference MonteCarlo.F90:1090


def _f(tok: str) -> float:
    return float(tok.replace("D", "E").replace("d", "e"))


@dataclass
class Box:
    """One box file (reference MonteCarlo.F90:1084-1126; layout in SURVEY.md appendix A)."""
    boxl: float                      # HALF the box side, Angstrom (reference BOXL)
    com: np.ndarray                  # [N][3] centres of mass
    hoek: np.ndarray                 # [N][3] Euler angles (radians)
    energy: float = ENERGY_ABSENT    # line 3, energy of the previous run
    box_id: str = "BOX"
    sol_sym: list[str] = field(default_factory=list)
    solute: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))
    sol_name: str = ""
    e_sol: float = 0.0               # Hartree
    dihedrals: list[tuple[int, int, float]] = field(default_factory=list)

    @property
    def n(self) -> int:
        return int(self.com.shape[0])

    @property
    def nsol(self) -> int:
        return int(self.solute.shape[0])


def _dihedral_line(line: str) -> tuple[int, int, float]:
    """One dihedral line as the reference reads it: fixed columns `(I3, 1X, I3, 1X, F10.6)` (MonteCarlo.F90:1123), so
    fields that touch (`12 13-90.000000`) still parse; blanks inside a numeric field are ignored, an F10.6 field
    without a decimal point has its last six digits taken as decimals, and a rotation of 0 means "use the global
    one", stored as -1 (:1124).  (The clip to DROTSOLU, :1125, needs the driver's config and stays with the driver.)"""
    line = line.rstrip("\n").ljust(18)
    a, b, r = line[0:3].replace(" ", ""), line[4:7].replace(" ", ""), line[8:18].replace(" ", "")
    ia, ib = (int(a) if a else 0), (int(b) if b else 0)
    if not r:
        rot = 0.0
    elif any(ch in r for ch in ".eEdD"):
        rot = _f(r)
    else:
        rot = int(r) / 1.0e6
    return ia, ib, (-1.0 if rot == 0.0 else rot)


def read_box(path: str | Path) -> Box:
    """Reads the current box format (reference MonteCarlo.F90:1078-1128)."""
    lines = Path(path).read_text().splitlines()
    boxl = _f(lines[0].split()[0])
    n = int(lines[1].split()[0])
    p = 2
    energy = ENERGY_ABSENT
    try:  # READ "(E20.10)": a line that does not parse as a number is the ID line (:1086-1092)
        energy = _f(lines[p].split()[0])
        p += 1
    except (ValueError, IndexError):
        pass
    box_id = lines[p].strip()
    p += 1
    vals = np.array([[_f(t) for t in lines[p + i].split()[:6]] for i in range(n)], dtype=np.float64)
    p += n
    box = Box(boxl=boxl, com=vals[:, 0:3].copy(), hoek=vals[:, 3:6].copy(), energy=energy, box_id=box_id)
    if p < len(lines) and lines[p].strip().upper().startswith("SOLUTE"):
        p += 1
        nsol = int(lines[p].split()[0]); p += 1
        box.sol_name = lines[p].rstrip("\n"); p += 1
        syms, xyz = [], []
        for i in range(nsol):
            t = lines[p + i].split()
            syms.append(t[0]); xyz.append([_f(t[1]), _f(t[2]), _f(t[3])])
        p += nsol
        box.sol_sym, box.solute = syms, np.array(xyz, dtype=np.float64).reshape(nsol, 3)
        box.e_sol = _f(lines[p].split()[0]); p += 1
        ndih = int(lines[p].split()[0]); p += 1
        for i in range(ndih):
            box.dihedrals.append(_dihedral_line(lines[p + i]))
    return box


def write_box(box: Box, path: str | Path) -> None:
    """Writes a box the reference's `read_norm` accepts (17 significant digits round-trip fp64)."""
    out = [f"  {box.boxl:.17g}", f"  {box.n:d}"]
    e = box.energy if box.energy != ENERGY_ABSENT else 0.0
    out.append(f"{e:20.10E}")
    out.append(f" {box.box_id}")
    for c, h in zip(box.com, box.hoek):
        out.append(" ".join(f"{v:.17g}" for v in (*c, *h)))
    out.append(" SOLUTE")
    out.append(f"  {box.nsol:d}")
    out.append(box.sol_name)
    for s, x in zip(box.sol_sym, box.solute):
        out.append(f" {s:<3s} " + " ".join(f"{v:.17g}" for v in x))
    out.append(f"  {box.e_sol:.17g}")
    out.append(f"  {len(box.dihedrals):d}")
    for a, b, r in box.dihedrals:
        out.append(f"{a:3d} {b:3d} {r:10.6f}")
    Path(path).write_text("\n".join(out) + "\n")


def _records(buf: bytes):
    """Fortran sequential unformatted records: 4-byte length, payload, the same length again."""
    p = 0
    while p < len(buf):
        n = int.from_bytes(buf[p:p + 4], "little")
        rec = buf[p + 4:p + 4 + n]
        if int.from_bytes(buf[p + 4 + n:p + 8 + n], "little") != n:
            raise ValueError("not a sequential unformatted file (record markers disagree)")
        yield rec
        p += n + 8


def read_box_bin(path: str | Path) -> Box:
    """Reads the unformatted box (reference MC_lib/lib_read.f90:117-175 READ_BIN; records in SURVEY.md appendix A):
    BOXL, nCoM, ENG, BOX_ID(255), N x 6 doubles, "SOLUTE", nSol, SOLNAME(255), nSol x (CHAR*4 + 3 doubles), E_sol,
    NDIHOEK, NDIHOEK x (2 integers + double; a rotation of 0 means "use the global one" and is stored as -1)."""
    rec = _records(Path(path).read_bytes())
    f8 = lambda r: np.frombuffer(r, dtype="<f8")
    i4 = lambda r: int(np.frombuffer(r[:4], dtype="<i4")[0])
    boxl = float(f8(next(rec))[0]); n = i4(next(rec)); energy = float(f8(next(rec))[0])
    box_id = next(rec).decode("latin-1").rstrip()
    vals = np.array([f8(next(rec))[:6] for _ in range(n)], dtype=np.float64).reshape(n, 6)
    box = Box(boxl=boxl, com=vals[:, 0:3].copy(), hoek=vals[:, 3:6].copy(), energy=energy, box_id=box_id)
    if next(rec).decode("latin-1").strip().upper() != "SOLUTE":
        raise ValueError("SOLUTE record expected")
    nsol = i4(next(rec))
    box.sol_name = next(rec).decode("latin-1").rstrip()
    syms, xyz = [], []
    for _ in range(nsol):
        r = next(rec)
        syms.append(r[:4].decode("latin-1").strip()); xyz.append(f8(r[4:28]))
    box.sol_sym, box.solute = syms, np.array(xyz, dtype=np.float64).reshape(nsol, 3)
    box.e_sol = float(f8(next(rec))[0])
    for _ in range(i4(next(rec))):
        r = next(rec)
        a, b = np.frombuffer(r[:8], dtype="<i4")
        rot = float(f8(r[8:16])[0])
        box.dihedrals.append((int(a), int(b), -1.0 if rot == 0.0 else rot))
    return box


def write_box_bin(box: Box, path: str | Path) -> None:
    """Writes the unformatted box READ_BIN accepts (reference MC_lib/lib_read.f90:177-225 WRITE_BIN)."""
    out = bytearray()

    def put(payload: bytes):
        n = len(payload).to_bytes(4, "little")
        out.extend(n); out.extend(payload); out.extend(n)

    f8 = lambda *v: np.asarray(v, dtype="<f8").tobytes()
    i4 = lambda *v: np.asarray(v, dtype="<i4").tobytes()
    put(f8(box.boxl)); put(i4(box.n)); put(f8(box.energy))
    put(box.box_id.encode("latin-1")[:255].ljust(255))
    for c, h in zip(box.com, box.hoek):
        put(f8(*c, *h))
    put(b"SOLUTE"); put(i4(box.nsol))
    put(box.sol_name.encode("latin-1")[:255].ljust(255))
    for sym, x in zip(box.sol_sym, box.solute):
        put(sym.encode("latin-1")[:4].ljust(4) + f8(*x))
    put(f8(box.e_sol)); put(i4(len(box.dihedrals)))
    for a, b, r in box.dihedrals:
        put(i4(a, b) + f8(r))
    Path(path).write_bytes(bytes(out))


def check_box(box: Box) -> dict:
    """The checks ChecknPrep runs on a box before a run (reference ChecknPrep/check.f90:130-229), on an already parsed
    box: returns {"errors": [...], "warnings": [...], "info": [...]}.  The reference counts a warning for BOXL, nCoM,
    the box energy, nSol or E_sol equal to 0 and for every molecule line with a coordinate or angle exactly 0, and an
    error for a dihedral bond that names atom 0 or an atom beyond nSol; "NDIHOEK is 0" is tested on nCoM there
    (check.f90:205-208), which is kept.  One note of our own: a centre of mass outside [-BOXL, BOXL] is legal for the
    reference but switches this library's neighbour and cell lists off (csrc/mcb_verlet.cuh, csrc/mcb_cells.cuh)."""
    err, war, info = [], [], []
    if box.boxl == 0.0:
        war.append("BOXL is 0!")
    if box.n == 0:
        war.append("nCoM is 0!")
    if box.energy == ENERGY_ABSENT:
        info.append("No energy detected in input box.")
    if box.energy == 0.0:
        war.append("ENG is 0!")
    for i, (c, h) in enumerate(zip(box.com, box.hoek)):
        if np.any(np.asarray(c) == 0.0) or np.any(np.asarray(h) == 0.0):
            war.append(f"coordinates are 0! Molecule: {i + 1}")
    if box.nsol == 0:
        war.append("nSOL is 0!")
    if box.e_sol == 0.0:
        war.append("E_SOL is 0!")
    if box.n == 0:
        war.append("NDHIHOEK is 0!")
    for k, (a, b, rot) in enumerate(box.dihedrals):
        if rot == 0.0 or rot == -1.0:
            info.append(f"MaxRot is not defined at dihedral {k + 1}")
        if a > box.nsol or b > box.nsol or a == 0 or b == 0:
            err.append(f"Values are out of bounds! Dihedral {k + 1}: {a:03d} {b:03d} {rot:8.2f}")
    if box.n and box.boxl > 0.0 and np.any(np.abs(box.com) > box.boxl):
        info.append("a centre of mass lies outside [-BOXL, BOXL]: neighbour and cell lists stay off for this box")
    return {"errors": err, "warnings": war, "info": info}


def box_text(box: Box) -> str:
    import io, tempfile, os
    fd, name = tempfile.mkstemp(suffix=".box")
    os.close(fd)
    try:
        write_box(box, name)
        return Path(name).read_text()
    finally:
        os.unlink(name)


def read_dmso(path: str | Path):
    """Solvent geometry file (reference MonteCarlo.F90:1066-1074): n, comment, n x `sym x y z`, E (Hartree)."""
    lines = Path(path).read_text().splitlines()
    n = int(lines[0].split()[0])
    syms, rel = [], []
    for i in range(n):
        t = lines[2 + i].split()
        syms.append(t[0]); rel.append([_f(t[1]), _f(t[2]), _f(t[3])])
    e = _f(lines[2 + n].split()[0])
    return syms, np.array(rel, dtype=np.float64), e


def read_param(path: str | Path):
    """param.txt (reference MonteCarlo.F90:1138-1146): n, comment, n x `sym q eps sigma [mass]`."""
    lines = Path(path).read_text().splitlines()
    n = int(lines[0].split()[0])
    syms, tab = [], []
    for i in range(n):
        t = lines[2 + i].split()
        syms.append(t[0]); tab.append([_f(t[1]), _f(t[2]), _f(t[3])])
    return syms, np.array(tab, dtype=np.float64)


def read_par_solute(path: str | Path):
    """par_solute.txt (reference MonteCarlo.F90:1155-1162): n, comment, n x `sym eps sigma`."""
    lines = Path(path).read_text().splitlines()
    n = int(lines[0].split()[0])
    syms, tab = [], []
    for i in range(n):
        t = lines[2 + i].split()
        syms.append(t[0]); tab.append([_f(t[1]), _f(t[2])])
    return syms, np.array(tab, dtype=np.float64)


def _find_sym(sym: str, table_syms: list[str]) -> int:
    """findSym (reference MC_lib/lib.f90:87-107): first match, a miss falls back to row 1."""
    for i, s in enumerate(table_syms):
        if s == sym:
            return i
    return 0


def associate_dmso(syms, par_syms, par_tab):
    """ASSOSIATE_DMSO (reference LennardJones.f90:192-219): (q, eps, sigma) per solvent atom.
    Hydrogen rows are never filled by the reference (and never read: LennardJones.f90:66,68);
    they are zero here."""
    tab = np.zeros((len(syms), 3), dtype=np.float64)
    is_h = np.array([1 if s == "H" else 0 for s in syms], dtype=np.int32)
    for i, s in enumerate(syms):
        if s != "H":
            tab[i] = par_tab[_find_sym(s, par_syms)]
    return tab, is_h


def associate_solute(syms, par_syms, par_tab, charges=None):
    """ASSOSIATE_SOLUTE (reference LennardJones.f90:224-248): q = 0 unless the driver supplies
    CHELPG charges afterwards (MonteCarlo.F90:353)."""
    tab = np.zeros((len(syms), 3), dtype=np.float64)
    for i, s in enumerate(syms):
        a = _find_sym(s, par_syms)
        tab[i, 1:3] = par_tab[a]
    if charges is not None:
        tab[:, 0] = np.asarray(charges, dtype=np.float64)
    return tab


def replicate(box: Box, k: int) -> Box:
    """Periodic k x k x k replication (new; the reference's BoxScale only rescales at constant N,
    BoxScale/BoxScale.f08:31-40).  Offsets (m - (k-1)/2) * 2*BOXL per axis, BOXL' = k*BOXL, angles copied;
    one solute, shifted by (+BOXL)^3 for even k so it stays inside a cavity."""
    if k == 1:
        return replace(box, com=box.com.copy(), hoek=box.hoek.copy(), solute=box.solute.copy())
    side = 2.0 * box.boxl
    coms, hoeks = [], []
    for mz in range(k):
        for my in range(k):
            for mx in range(k):
                off = (np.array([mx, my, mz], dtype=np.float64) - (k - 1) / 2.0) * side
                coms.append(box.com + off)
                hoeks.append(box.hoek.copy())
    solute = box.solute + (box.boxl if k % 2 == 0 else 0.0)
    return replace(box, boxl=k * box.boxl, com=np.concatenate(coms), hoek=np.concatenate(hoeks),
                   solute=solute, box_id=f"{box.box_id} x{k}^3")


def scale(box: Box, factor: float) -> Box:
    """BoxScale (reference BoxScale/BoxScale.f08:31-40): BOXL and every CoM times `factor`, N unchanged."""
    return replace(box, boxl=box.boxl * factor, com=box.com * factor, hoek=box.hoek.copy())


@dataclass
class System:
    """Everything the energy path needs besides the per-chain state: the reference's DMSO, DMSO_SYM,
    TABLE_DMSO, SOLUTE, TABLE_SOL, BOXL (MonteCarlo.F90:306-310), reduced to the sites that interact."""
    boxl: float
    dmso_sym: list[str]
    dmso_rel: np.ndarray     # [ndmso][3]
    dmso_tab: np.ndarray     # [ndmso][3]
    dmso_is_h: np.ndarray    # [ndmso] int32
    sol_sym: list[str]
    solute: np.ndarray       # [nsol][3]
    sol_tab: np.ndarray      # [nsol][3]
    conv: float = 0.0        # reference value (LennardJones.f90:57)
    cutoff: float = 7.0      # MonteCarlo.F90:812

    @property
    def site_rel(self) -> np.ndarray:
        return np.ascontiguousarray(self.dmso_rel[self.dmso_is_h == 0])

    @property
    def site_tab(self) -> np.ndarray:
        return np.ascontiguousarray(self.dmso_tab[self.dmso_is_h == 0])


def load_system(box: Box, dmso_path, param_path, parsol_path, charges=None, conv: float = 0.0) -> System:
    dsym, drel, _ = read_dmso(dmso_path)
    psym, ptab = read_param(param_path)
    ssym, stab = read_par_solute(parsol_path)
    dtab, is_h = associate_dmso(dsym, psym, ptab)
    soltab = associate_solute(box.sol_sym, ssym, stab, charges)
    return System(boxl=box.boxl, dmso_sym=dsym, dmso_rel=drel, dmso_tab=dtab, dmso_is_h=is_h,
                  sol_sym=list(box.sol_sym), solute=box.solute.copy(), sol_tab=soltab, conv=conv)

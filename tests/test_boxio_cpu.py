"""CPU: the driver's file formats and the box generators (host logic, no energies)."""
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as ol
from montecarlo_b200 import boxio


def test_read_shipped_box(shipped):
    box, system = shipped
    assert box.n == 30 and box.nsol == 11 and box.boxl == 7.7838685668
    assert box.energy == -860.60490768939985 and box.box_id == "BOX 3"
    assert box.dihedrals == [(8, 1, 90.0)] and box.sol_sym[2] == "Cl"
    assert system.site_rel.shape == (4, 3) and list(system.dmso_is_h) == [0, 0, 1, 1, 1, 0, 1, 1, 1, 0]
    # S, C, C, O rows of param.txt
    assert np.allclose(system.site_tab[:, 1], [1.2970, 0.9414, 0.9414, 1.7154])
    assert np.allclose(system.sol_tab[:, 0], 0.0) and system.sol_tab[2, 1] == 1.442  # Cl


def test_box_roundtrip(tmp_path, shipped):
    box, _ = shipped
    p = tmp_path / "b.txt"
    boxio.write_box(box, p)
    b2 = boxio.read_box(p)
    assert np.array_equal(b2.com, box.com) and np.array_equal(b2.hoek, box.hoek) and np.array_equal(b2.solute, box.solute)
    assert b2.boxl == box.boxl and b2.dihedrals == box.dihedrals and b2.sol_sym == box.sol_sym


def test_box_without_energy_line(tmp_path, shipped):
    box, _ = shipped
    lines = (ol.INPUTS / "box.txt").read_text().splitlines()
    del lines[2]
    p = tmp_path / "noenergy.txt"
    p.write_text("\n".join(lines) + "\n")
    b = boxio.read_box(p)
    assert b.energy == boxio.ENERGY_ABSENT and b.n == 30 and np.array_equal(b.com, box.com)


def test_replicate_and_scale(shipped):
    box, _ = shipped
    b2 = boxio.replicate(box, 2)
    assert b2.n == 240 and abs(b2.boxl - 15.5677371336) < 1e-12
    assert np.all(np.abs(b2.com) <= b2.boxl)
    # every replica is the original shifted by a multiple of the original side
    d = (b2.com.reshape(8, 30, 3) - box.com) / (2 * box.boxl)
    assert np.allclose(d, np.round(d * 2) / 2)
    assert np.allclose(b2.solute - box.solute, box.boxl)
    b3 = boxio.replicate(box, 3)
    assert b3.n == 810 and np.array_equal(b3.solute, box.solute)
    s = boxio.scale(box, 0.9)
    assert s.n == 30 and s.boxl == box.boxl * 0.9 and np.array_equal(s.com, box.com * 0.9)


def test_unformatted_box_is_the_text_box(tmp_path, shipped):
    """The reference ships the same box as text and as a Fortran sequential unformatted file (boxbin/box.txt,
    boxbin/box.bin, converted by boxbin/boxbin.f90 with MC_lib/lib_read.f90 READ_BIN / WRITE_BIN): every number must be
    the same through both readers, and writing the binary box again must give the reference's file byte for byte."""
    box, _ = shipped
    golden = Path(__file__).parent / "golden" / "inputs" / "box.bin"
    b = boxio.read_box_bin(golden)
    assert b.boxl == box.boxl and b.n == box.n and b.energy == box.energy
    assert np.array_equal(b.com, box.com) and np.array_equal(b.hoek, box.hoek)
    assert b.sol_sym == box.sol_sym and np.array_equal(b.solute, box.solute) and b.sol_name == box.sol_name
    assert b.e_sol == box.e_sol and b.dihedrals == box.dihedrals
    out = tmp_path / "out.bin"
    boxio.write_box_bin(b, out)
    assert out.read_bytes() == golden.read_bytes()
    # a rotation of 0 is stored as "use the global one" (lib_read.f90:168)
    b0 = boxio.Box(**{**b.__dict__, "dihedrals": [(8, 1, 0.0)]})
    boxio.write_box_bin(b0, out)
    assert boxio.read_box_bin(out).dihedrals == [(8, 1, -1.0)]
    with pytest.raises(ValueError):
        boxio.read_box_bin(Path(__file__).parent / "golden" / "inputs" / "box.txt")


def test_check_box_follows_checknprep(shipped):
    """ChecknPrep's box checks (check.f90:130-229): the shipped box is clean; zeros, an empty solute and a dihedral
    bond beyond nSol are reported the way the reference counts them."""
    box, _ = shipped
    r = boxio.check_box(box)
    assert r == {"errors": [], "warnings": [], "info": []}
    com = box.com.copy(); com[4, 1] = 0.0
    bad = boxio.Box(**{**box.__dict__, "com": com, "energy": 0.0, "e_sol": 0.0, "dihedrals": [(8, 1, 0.0), (12, 1, 30.0), (0, 3, 30.0)]})
    r = boxio.check_box(bad)
    assert r["warnings"] == ["ENG is 0!", "coordinates are 0! Molecule: 5", "E_SOL is 0!"]
    assert len(r["errors"]) == 2 and "Dihedral 2: 012 001" in r["errors"][0] and "Dihedral 3: 000 003" in r["errors"][1]
    assert r["info"] == ["MaxRot is not defined at dihedral 1"]
    far = box.com.copy(); far[0, 0] = 3.0 * box.boxl
    r = boxio.check_box(boxio.Box(**{**box.__dict__, "com": far, "energy": boxio.ENERGY_ABSENT}))
    assert r["info"][0] == "No energy detected in input box." and "outside" in r["info"][1]


def test_dihedral_lines_are_fixed_columns(tmp_path, shipped):
    """The reference reads dihedral lines with (I3,1X,I3,1X,F10.6) (MonteCarlo.F90:1123-1124): touching fields still
    parse, and a rotation of 0 becomes -1 ("use the global one") in the text reader as in the unformatted one."""
    assert boxio._dihedral_line("  8   1  90.000000") == (8, 1, 90.0)
    assert boxio._dihedral_line("100 -12 -12.500000") == (100, -12, -12.5)
    assert boxio._dihedral_line(" 12  13x-90.00000") == (12, 13, -90.0)      # column 8 is the 1X: whatever stands there is skipped
    assert boxio._dihedral_line("  8   1   0.000000") == (8, 1, -1.0)
    assert boxio._dihedral_line("  8   1") == (8, 1, -1.0)                  # short line: the field reads as blank = 0
    assert boxio._dihedral_line("  8   1        30") == (8, 1, 30e-6)       # no decimal point: F10.6 implies six decimals
    box, _ = shipped
    b0 = boxio.Box(**{**box.__dict__, "dihedrals": [(8, 1, 0.0)]})
    p, q = tmp_path / "z.txt", tmp_path / "z.bin"
    boxio.write_box(b0, p); boxio.write_box_bin(b0, q)
    assert boxio.read_box(p).dihedrals == boxio.read_box_bin(q).dihedrals == [(8, 1, -1.0)]

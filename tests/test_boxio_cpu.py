"""CPU: the driver's file formats and the box generators (host logic, no energies)."""
import numpy as np

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

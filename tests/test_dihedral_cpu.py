"""CPU: the host-side dihedral generator for config 4, pinned by the reference's own production log
(PostProcess/log030.txt:51-52, same start box: "Dihedral angle with 09 - 08 - 01 - 02",
"Dihedral change from  -71.24 with   -3.32 to  -74.55")."""
import numpy as np

from montecarlo_b200 import dihedral


def test_atoms_and_angle_match_reference_log(shipped):
    box, _ = shipped
    atoms = dihedral.dihedral_atoms(box.solute, box.dihedrals[0][:2])
    assert atoms == (9, 8, 1, 2)
    pre = np.rad2deg(dihedral.get_dihedral(box.solute, *atoms))
    assert abs(pre - (-71.24)) < 0.006 and abs(pre - (-71.2351)) < 1e-3
    rot = dihedral.set_dihedral(box.solute, box.sol_sym, *atoms, np.deg2rad(-3.32))
    post = np.rad2deg(dihedral.get_dihedral(rot, *atoms))
    assert abs(post - (-74.55)) < 0.011            # printed with two decimals; the increment too
    # the check the reference's own CLI makes (CalcDihedral/changeDih.f90:53-65): post - pre - requested < 0.01 degrees
    assert abs(post - pre - (-3.32)) < 1e-9


def test_fragments_and_rigidity(shipped):
    box, _ = shipped
    frag = dihedral.fragments(box.solute, box.sol_sym, 8, 1)
    assert [i + 1 for i in np.nonzero(frag == 2)[0]] == [2, 3, 4, 5, 6, 7]
    assert [i + 1 for i in np.nonzero(frag == 1)[0]] == [9, 10, 11]
    rot = dihedral.set_dihedral(box.solute, box.sol_sym, 9, 8, 1, 2, 1.0)
    d0 = np.linalg.norm(box.solute[:, None] - box.solute[None], axis=-1)
    d1 = np.linalg.norm(rot[:, None] - rot[None], axis=-1)
    moving = frag == 2
    assert np.allclose(d0[np.ix_(moving, moving)], d1[np.ix_(moving, moving)], atol=1e-12)   # fragment moves rigidly
    assert np.array_equal(rot[~moving], box.solute[~moving])


def test_conformer_sweep_hits_targets(shipped):
    box, _ = shipped
    targets = np.arange(0.0, 360.0, 10.0)
    conf, atoms = dihedral.conformer_sweep(box.solute, box.sol_sym, box.dihedrals[0][:2], targets)
    assert conf.shape == (36, 11, 3)
    got = np.array([np.rad2deg(dihedral.get_dihedral(c, *atoms)) for c in conf])
    err = (got - targets + 180.0) % 360.0 - 180.0
    assert np.abs(err).max() < 1e-9

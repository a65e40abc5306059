"""CPU: the C restatement (oracle/) against fixtures produced by the reference's own binary.

This is what makes the oracle "pinned": every quantity below was written by the unmodified
MonteCarlo/MonteCarloOpenMP (oracle/make_golden.py), not restated.
"""
import json

import numpy as np
import pytest

import oracle_lib as ol
from golden_util import (GOLD, C_TOT, C_OLD, C_KANS, C_RV, C_RSOLV, C_RATIO, C_DPOS, C_DHOEK, C_PSUC, load_trace,
                         recover_uniforms)
from montecarlo_b200 import boxio

PRINT11 = 6e-11  # `ES20.10` prints 11 significant digits: half a unit in the last place, relative


def test_constants():
    L = ol.lib()
    assert L.mco_const_e() == 2.7182817459106445           # float literal, reference MC_lib/lib.f90:17
    assert L.mco_const_hartree2kjmol() == 2625.499267578125  # float32 product, lib.f90:19
    assert L.mco_const_conv_reference() == 0.0             # LennardJones.f90:57
    # PostProcess/log030.txt:45-46 print these products of the reference's own constants
    assert abs(-0.16947515 * L.mco_const_hartree2kjmol() - (-444.956882197693)) < 1e-9
    assert abs(float(np.float32(-.0499674791107)) * L.mco_const_hartree2kjmol() - (-131.189579414181)) < 1e-9
    assert abs(0.053 * L.mco_const_pi() - 0.166504410640259) < 1e-15


def test_startup_shipped_box(shipped):
    box, system = shipped
    g = np.load(GOLD / "ref_startup_k1.npz")
    ch = ol.OracleChain(system, box.com, box.hoek)
    e = ch.total_energy()
    assert abs(e - float(g["initial_energy"])) <= PRINT11 * abs(e)
    assert abs(e - 2.7968745663e3) <= PRINT11 * abs(e)
    ss, en = ch.matrix()
    ref = g["matrix"]
    assert np.array_equal(ss != 0, ref != 0) and int((ref != 0).sum()) == 340
    nz = ref != 0
    assert np.max(np.abs(ss[nz] - ref[nz]) / np.abs(ref[nz])) < 1e-13  # 15 printed digits
    assert np.array_equal(ss, ss.T)
    assert abs(ss[0, 2] - (-1.93595228537649)) < 1e-13 and abs(ss[5, 26] - 1235.49700686554) < 1e-10
    # matrix-free variants hold the same numbers
    e_mf, es, npair = ch.total_energy_matrixfree()
    assert e_mf == e and npair == 170 and abs(es - 463.892021002) < 1e-7
    # the fp128 arbiter agrees far inside the tolerance the GPU is held to
    q = ol.OracleChain(system, box.com, box.hoek, quad=True).total_energy()
    assert abs(q - e) < 1e-12 * abs(e)


def test_charges_do_not_matter(shipped):
    """CONV == 0 in the reference: the start-up energy with CHELPG charges is the same number."""
    box, system = shipped
    g = json.loads((GOLD / "ref_charges.json").read_text())
    assert g["matrix_equal_to_uncharged"]
    tab = system.sol_tab.copy()
    tab[:, 0] = g["charges"]
    e = ol.OracleChain(system, box.com, box.hoek, sol_tab=tab).total_energy()
    assert abs(e - g["initial_energy"]) <= PRINT11 * abs(e)
    assert e == ol.OracleChain(system, box.com, box.hoek).total_energy()


def test_startup_replica_k2(shipped):
    box, system = shipped
    g = np.load(GOLD / "ref_startup_k2.npz")
    b2 = boxio.replicate(box, 2)
    s2 = boxio.System(**{**system.__dict__, "boxl": b2.boxl, "solute": b2.solute})
    ch = ol.OracleChain(s2, b2.com, b2.hoek)
    e = ch.total_energy()
    assert abs(e - float(g["initial_energy"])) <= PRINT11 * abs(e)
    ss, _ = ch.matrix()
    ref = g["matrix"]
    assert np.array_equal(ss != 0, ref != 0) and int((ref != 0).sum()) == 2720
    nz = ref != 0
    assert np.max(np.abs(ss[nz] - ref[nz]) / np.abs(ref[nz])) < 1e-12
    e_c, _, npair = ch.total_energy_matrixfree(cells=True)
    assert e_c == ch.total_energy_matrixfree()[0] == e and npair == 1360


def test_startup_replica_k5(shipped):
    box, system = shipped
    g = json.loads((GOLD / "ref_startup_k5.json").read_text())
    b5 = boxio.replicate(box, 5)
    s5 = boxio.System(**{**system.__dict__, "boxl": b5.boxl, "solute": b5.solute})
    ch = ol.OracleChain(s5, b5.com, b5.hoek, keep_matrix=False)
    e, es, npair = ch.total_energy_matrixfree(cells=True)
    assert b5.n == g["n"] == 3750 and 2 * npair == g["nonzero"]
    assert abs(e - g["initial_energy"]) <= PRINT11 * abs(e)
    assert abs((e - es) - g["sum_upper"]) < 1e-9 * abs(e)
    e_b, _, np_b = ch.total_energy_matrixfree(cells=False)
    assert e_b == e and np_b == npair


@pytest.mark.parametrize("name", ["ref_trace_10k.npz", "ref_trace_adapt.npz"])
def test_replay_reference_trace(shipped, name):
    """Drive the oracle's MCINIT + calculateLJ + calcEnergy + METROPOLIS with uniforms recovered from the
    reference's own DUMP/OUT files and compare every printed column of every step."""
    box, system = shipped
    t = load_trace(name)
    rows, moved, acc = t["out_rows"], t["moved"], t["accepted"]
    temp, dposmax, dposmin, dhoekmax, dhoekmin, padj, nadj = t["params"]
    ch = ol.OracleChain(system, box.com, box.hoek)
    ch.set_mc_params(temp, dposmax, dposmin, dhoekmax * np.pi, dhoekmin * np.pi, padj, int(nadj), 0, 1)
    e0 = ch.total_energy()
    assert abs(e0 - float(t["initial_energy"])) <= PRINT11 * abs(e0)
    cur = np.concatenate([box.com, box.hoek], axis=1)
    nsuc = 0
    unresolved = 0
    for k in range(rows.shape[0]):
        st = ch.stats()
        i1 = int(rows[k, C_RSOLV])
        u = recover_uniforms(cur[i1 - 1], moved[k], st["dposmax"], st["dhoekmax"], box.boxl, i1, box.n, acc[k], True)
        used, tr = ch.step(u)
        assert int(tr[4]) == i1 - 1
        com, hoek = ch.get_state()
        # trial energy and the energy it is compared with, 11 printed digits
        assert abs(tr[0] - rows[k, C_TOT]) <= 2 * PRINT11 * max(abs(tr[0]), abs(rows[k, C_OLD])), (k, tr[0], rows[k, C_TOT])
        assert abs(tr[1] - rows[k, C_OLD]) <= 2 * PRINT11 * abs(tr[1]), k
        # kans / rv columns (F6.4) and the draw-skipping rules of METROPOLIS (MonteCarlo.F90:730-741)
        assert abs(tr[2] - rows[k, C_KANS]) <= 5.1e-5, (k, tr[2], rows[k, C_KANS])
        if used == 7:
            assert rows[k, C_RV] in (0.0, 1.0) and tr[3] == rows[k, C_RV]
        else:
            # decision from the PRINTED rv against the oracle's kans, where 4 decimals can resolve it
            if abs(rows[k, C_RV] - tr[2]) > 1.01e-4:
                assert (rows[k, C_RV] <= tr[2]) == bool(acc[k]), (k, rows[k, C_RV], tr[2])
            else:
                unresolved += 1
        assert bool(tr[5]) == bool(acc[k])
        nsuc += int(acc[k])
        assert abs(tr[6] - rows[k, C_RATIO]) <= 5.1e-5
        assert abs(nsuc / (k + 1) - rows[k, C_PSUC]) <= 5.1e-5
        assert abs(tr[7] - rows[k, C_DPOS]) <= 5.1e-6 and abs(tr[8] - rows[k, C_DHOEK]) <= 5.1e-6
        if acc[k]:
            assert np.abs(np.concatenate([com[i1 - 1], hoek[i1 - 1]]) - moved[k]).max() < 5e-13
            cur[i1 - 1] = np.concatenate([com[i1 - 1], hoek[i1 - 1]])
    assert unresolved <= 10
    # `Final energy (LJ)` prints TOTENG_OLD (MonteCarlo.F90:437): the energy the last step started from
    assert abs(ch.stats()["toteng_old"] - float(t["final_energy"])) <= 2 * PRINT11 * abs(float(t["final_energy"]))


def test_rng_stream_is_documented():
    """key = mix(seed + G); u_n = (mix(key + (n+1) G) >> 11) * 2^-53 — restated here in python integers."""
    M = (1 << 64) - 1
    G = 0x9E3779B97F4A7C15

    def mix(x):
        x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M
        x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M
        return x ^ (x >> 31)

    for seed in (0, 1, 2, 1024, M):
        st = ol.rng_seed(seed)
        key = mix((seed + G) & M)
        assert int(st[0]) == key and int(st[1]) == 0
        got = ol.rng_uniforms(st, 9)
        for n, g in enumerate(got):
            assert g == (mix((key + (n + 1) * G) & M) >> 11) * 2.0 ** -53 and 0.0 <= g < 1.0
        assert int(st[1]) == 9
    a = ol.rng_uniforms(ol.rng_seed(1), 4096)
    b = ol.rng_uniforms(ol.rng_seed(2), 4096)
    assert abs(a.mean() - 0.5) < 0.02 and abs(np.corrcoef(a, b)[0, 1]) < 0.06


def test_exchange_rule_is_metropolis_rule(shipped):
    """mco_exchange (the checker for mcb_exchange; the reference has no exchange move) is METROPOLIS' rule
    (MonteCarlo.F90:728-744) on (beta_a - beta_b)(E_a - E_b): no draw below -75 and at or above 0, else RV <= E**x with
    the default-REAL E; an accepted swap moves temperature and step sizes, nothing else."""
    box, system = shipped
    com2 = box.com.copy(); com2[0] += 0.05
    a = ol.OracleChain(system, box.com, box.hoek); b = ol.OracleChain(system, com2, box.hoek)
    a.set_mc_params(temperature=300.0, dposmax=0.3); b.set_mc_params(temperature=330.0, dposmax=0.4)
    ea, eb = a.total_energy(), b.total_energy()
    x = (1000.0 / (8.314 * 300.0) - 1000.0 / (8.314 * 330.0)) * (ea - eb)
    assert -75.0 < x and x != 0.0
    if x >= 0.0:
        assert ol.exchange(a, b, 0.999999) == (True, False)
    else:
        kans = ol.lib().mco_const_e() ** x
        assert ol.exchange(a, b, float(np.nextafter(kans, 1.0))) == (False, True)      # RV just above KANS: rejected
        assert a.L.mco_get_temperature(a.h) == 300.0
        assert ol.exchange(a, b, kans) == (True, True)                                  # RV == KANS: accepted (.LE.)
    assert a.L.mco_get_temperature(a.h) == 330.0 and b.L.mco_get_temperature(b.h) == 300.0
    assert a.stats()["dposmax"] == 0.4 and b.stats()["dposmax"] == 0.3
    assert a.stats()["toteng"] == ea and b.stats()["toteng"] == eb
    # a hopeless swap (exponent < -75) never draws
    c = ol.OracleChain(system, box.com, box.hoek); d = ol.OracleChain(system, com2, box.hoek)
    ec, ed = c.total_energy(), d.total_energy()
    cold, hot = (c, d) if ec < ed else (d, c)                     # the cold chain already holds the lower energy
    cold.set_mc_params(temperature=1.0); hot.set_mc_params(temperature=1000.0)
    cold.total_energy(); hot.total_energy()
    assert ol.exchange(cold, hot, 0.0) == (False, False) and ol.exchange(hot, cold, 0.0) == (False, False)

"""CPU: window hopping and solute-level Metropolis (PostProcess.F90 COMPARE, solmod.F90 SOLUTE_METROPOLIS)."""
import numpy as np

from montecarlo_b200 import postprocess as pp


def test_solute_metropolis_rules():
    none = lambda: (_ for _ in ()).throw(AssertionError("no draw expected"))  # noqa: E731
    assert pp.solute_metropolis(-100.0, -101.0, 300.0, none) is True           # downhill: no draw
    assert pp.solute_metropolis(-100.0, 900.0, 300.0, none) is False           # exponent < -75: no draw
    kans = pp.E_F32 ** (-1.0 * 1.0 * 1000.0 / (8.314 * 300.0))                  # +1 kJ/mol at 300 K
    assert abs(kans - 0.6697) < 1e-3
    assert pp.solute_metropolis(0.0, 1.0, 300.0, lambda: kans) is True         # RV <= KANS
    assert pp.solute_metropolis(0.0, 1.0, 300.0, lambda: np.nextafter(kans, 1.0)) is False


def test_window_hopping_sequence():
    # three windows (as in PostProcess/log030|150|270), four runs, scripted uniforms
    e = np.array([[5.0, 1.0, 3.0], [5.0, 1.0, 0.5], [0.0, 2.0, 0.5], [0.0, 2.0, 9.0]])
    u = iter([0.2,            # run 0: from window 1 go +1 -> window 2, uphill 2 kJ/mol: needs a draw
              0.99,           #        rejected (kans = exp(-0.80) = 0.448)
              0.2,            # run 1: +1 -> window 2 again, downhill: accepted, no draw
              0.7,            # run 2: -1 -> window 1, uphill 1.5: draw
              0.1,            #        accepted (kans = 0.548)
              0.2])           # run 3: +1 -> window 2, +7 kJ/mol: draw
    u = iter(list(u) + [0.5])
    old, now, acc = pp.window_hopping(e, 300.0, u)
    assert list(old) == [1, 1, 2, 1] and list(now) == [2, 2, 1, 2] and list(acc) == [False, True, True, False]
    # cyclic neighbours
    old, now, acc = pp.window_hopping(np.array([[0.0, 1.0, 1.0]]), 300.0, iter([0.9, 0.0]))
    assert old[0] == 0 and now[0] == 2


def test_window_hopping_literal_indexing():
    """literal=True reproduces COMPARE's transposed subscripts (PostProcess.F90:66, :249, :277): ENERGY(I, NEXT) on an
    array allocated ENERGY(NUMLOG, RUNS) is the column-major linear element NEXT*NUMLOG + I.  Checked against an
    independent emulation with a Fortran-ordered numpy array."""
    rng = np.random.default_rng(4)
    nwin, runs = 3, 7
    e = rng.normal(0.0, 2.0, size=(runs, nwin))               # [run, window]
    F = np.asfortranarray(e.T)                                # ENERGY(NUMLOG, RUNS), column-major like the reference
    lin = F.ravel(order="F")

    def emulate(uniforms):
        it = iter(uniforms)
        cur = int(np.argmin(F[:, 0])); out = []
        for i in range(runs):
            nxt = cur + 1 if next(it) < 0.5 else cur - 1
            nxt = 0 if nxt > nwin - 1 else (nwin - 1 if nxt < 0 else nxt)
            k_n, k_c = nxt * nwin + i, cur * nwin + i          # element (I, NEXT) of an array with leading dimension NUMLOG
            if max(k_n, k_c) >= lin.size:
                return out, i
            x = -1.0 * (lin[k_n] - lin[k_c]) * 1000.0 / (8.314 * 300.0)
            if x < -75.0: kans, rv = 0.0, 1.0
            elif x >= 0.0: kans, rv = 1.0, 0.0
            else: kans, rv = min(np.exp(x), 1.0), next(it)
            a = rv < kans
            out.append((cur, nxt, a))
            if a: cur = nxt
        return out, runs

    u = list(rng.random(4 * runs))
    want, n_ok = emulate(u)
    old, now, acc = pp.window_hopping(e, 300.0, iter(u), literal=True)
    assert [(int(o), int(n), bool(a)) for o, n, a in zip(old, now, acc)] == want
    o2, n2, a2 = pp.window_hopping(e, 300.0, iter(u), literal=False)
    assert not (np.array_equal(acc, a2) and np.array_equal(now, n2))          # (on this data the intended reading hops differently)
    assert n_ok == runs
    # the two readings differ: run 3's literal element (window 1) is linear 1*3 + 3 = 6 = ENERGY(1, 3) = run 2, window 0
    assert lin[1 * nwin + 3] == e[2, 0]
    # fewer runs than windows: the literal read leaves the array (undefined in the reference)
    import pytest
    with pytest.raises(IndexError, match="outside ENERGY"):
        pp.window_hopping(np.zeros((2, 3)), 300.0, iter([0.9] + [0.2] * 8), literal=True)   # first hop: window 0 -> 2

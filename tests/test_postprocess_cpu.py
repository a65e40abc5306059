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

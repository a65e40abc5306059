"""GPU: the CUDA path (through the C ABI) against the oracle and the reference-produced fixtures.

Tolerances: BASELINE.json asks for 1e-10 relative on per-move dE and full-box energy; energies are compared
at RTOL = 1e-10 relative to the energies involved, coordinates at 1e-12 absolute, counters exactly.
"""
import json

import numpy as np
import pytest

import oracle_lib as ol
from golden_util import GOLD, C_TOT, C_OLD, C_RSOLV, load_trace
from montecarlo_b200 import boxio

pytestmark = pytest.mark.gpu

RTOL = 1e-10
PRINT11 = 6e-11


def _engine(system, n, **kw):
    from montecarlo_b200.engine import LJEngine
    return LJEngine(system, n, **kw)


def _sys_for(system, box):
    return boxio.System(**{**system.__dict__, "boxl": box.boxl, "solute": box.solute})


def _close(a, b, scale=None, rtol=RTOL):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    s = np.maximum(np.abs(a), np.abs(b)) if scale is None else scale
    return np.all(np.abs(a - b) <= rtol * s + 1e-300)


def test_startup_energy_and_matrix(shipped):
    box, system = shipped
    g = np.load(GOLD / "ref_startup_k1.npz")
    eng = _engine(system, box.n, n_chains=3)
    eng.set_state(box.com, box.hoek)
    tot, esol, pairs = eng.total_energy()
    o = ol.OracleChain(system, box.com, box.hoek)
    e = o.total_energy()
    assert _close(tot, e) and np.all(pairs == 170)
    assert np.all(np.abs(tot - float(g["initial_energy"])) <= PRINT11 * abs(e))   # the reference's printed value
    ss, en = eng.matrix(1)
    oss, oen = o.matrix()
    assert np.array_equal(ss != 0, oss != 0) and np.array_equal(ss, ss.T)
    assert _close(ss, oss, scale=np.abs(oss) + 1e-3) and _close(en, oen)
    ref = g["matrix"]
    nz = ref != 0
    assert np.max(np.abs(ss[nz] - ref[nz]) / (np.abs(ref[nz]) + 1e-3)) < 1e-10
    assert abs(esol[0] - oen.sum()) <= RTOL * abs(oen.sum())
    # quad arbiter: the GPU is as close to the exact value as the fp64 restatement is
    q = ol.OracleChain(system, box.com, box.hoek, quad=True).total_energy()
    assert abs(tot[0] - q) <= 1e-12 * abs(q)


@pytest.mark.parametrize("matrix_mode", [0, 1])
def test_startup_replica_k2(shipped, matrix_mode):
    box, system = shipped
    g = np.load(GOLD / "ref_startup_k2.npz")
    b2 = boxio.replicate(box, 2)
    s2 = _sys_for(system, b2)
    eng = _engine(s2, b2.n, n_chains=2, matrix_mode=matrix_mode)
    eng.set_state(b2.com, b2.hoek)
    tot, esol, pairs = eng.total_energy()
    assert np.all(np.abs(tot - float(g["initial_energy"])) <= PRINT11 * abs(tot[0])) and np.all(pairs == 1360)
    ss, _ = eng.matrix(0)
    ref = g["matrix"]
    assert np.array_equal(ss != 0, ref != 0)
    nz = ref != 0
    assert np.max(np.abs(ss[nz] - ref[nz]) / (np.abs(ref[nz]) + 1e-3)) < 1e-10


def test_startup_replica_k5(shipped):
    box, system = shipped
    g = json.loads((GOLD / "ref_startup_k5.json").read_text())
    b5 = boxio.replicate(box, 5)
    eng = _engine(_sys_for(system, b5), b5.n, n_chains=1)
    eng.set_state(b5.com, b5.hoek)
    tot, esol, pairs = eng.total_energy()
    assert abs(tot[0] - g["initial_energy"]) <= PRINT11 * abs(tot[0])
    assert 2 * pairs[0] == g["nonzero"]
    assert abs((tot[0] - esol[0]) - g["sum_upper"]) <= 1e-9 * abs(tot[0])


def test_config5_hundred_thousand_molecules(shipped):
    """BASELINE config 5: 15x15x15 replica, N = 101 250, full-box energy only (the reference cannot run it: its two
    N x N matrices would need 2 x 82 GB).  Matrix-free kernels against the cell-list oracle, which itself is
    anchored by the reference binary at k = 1, 2, 5 (test_oracle_golden.py)."""
    box, system = shipped
    b = boxio.replicate(box, 15)
    assert b.n == 101250 and abs(b.boxl - 116.758028502) < 1e-9
    s15 = _sys_for(system, b)
    eng = _engine(s15, b.n, n_chains=1)
    eng.set_state(b.com, b.hoek)
    tot, esol, pairs = eng.total_energy()
    o = ol.OracleChain(s15, b.com, b.hoek, keep_matrix=False)
    e, es, npair = o.total_energy_matrixfree(cells=True)
    assert pairs[0] == npair and abs(tot[0] - e) <= RTOL * abs(e) and abs(esol[0] - es) <= RTOL * abs(es)
    # a second call gives the same bits (fixed-order reductions, no atomics)
    assert eng.total_energy()[0][0] == tot[0]
    from montecarlo_b200.engine import McbError
    with pytest.raises(McbError, match="energy-only"):
        eng.run_steps(10)
    with pytest.raises(McbError, match="energy-only"):
        eng.matrix(0)


def test_cell_list_against_all_pairs(shipped, monkeypatch):
    """The matrix-free energy finds its pairs through a cell list (mcb_cells.cuh) when every centre of mass is inside
    the box, and through the all-pairs kernel otherwise.  Same pairs and the same energy to rounding on a shuffled
    5x5x5 replica; with one molecule pushed outside the box the cell path must hand over to the all-pairs kernel
    (bit-identical result), because the reference's minimum image is no longer a metric there."""
    box, system = shipped
    b = boxio.replicate(box, 5)
    s5 = _sys_for(system, b)
    rng = np.random.default_rng(3)
    perm = rng.permutation(b.n)                      # cell order != index order
    com, hoek = b.com[perm].copy(), b.hoek[perm].copy()

    def energy(c, h, cells):
        if cells:
            monkeypatch.delenv("MCB200_NO_CELLS", raising=False)
        else:
            monkeypatch.setenv("MCB200_NO_CELLS", "1")
        eng = _engine(s5, b.n, n_chains=1)
        eng.set_state(c, h)
        r = eng.total_energy()
        r2 = eng.total_energy()
        assert r[0][0] == r2[0][0] and r[2][0] == r2[2][0]      # same bits on a second call
        return r[0][0], r[1][0], int(r[2][0])

    t_c, es_c, n_c = energy(com, hoek, True)
    t_a, es_a, n_a = energy(com, hoek, False)
    assert n_c == n_a and es_c == es_a
    assert abs(t_c - t_a) <= 1e-12 * abs(t_a)
    o = ol.OracleChain(s5, com, hoek, keep_matrix=False)
    e, es, npair = o.total_energy_matrixfree(cells=False)
    assert n_c == npair and abs(t_c - e) <= RTOL * abs(e)
    com_out = com.copy()
    com_out[17, 0] += 2.5 * b.boxl                   # far outside [-BOXL, BOXL]
    t_c2, _, n_c2 = energy(com_out, hoek, True)
    t_a2, _, n_a2 = energy(com_out, hoek, False)
    assert n_c2 == n_a2 and t_c2 == t_a2


def test_coulomb_path_when_conv_is_nonzero(shipped):
    """The reference's CONV is 0 (LennardJones.f90:57); the kernel still carries q1 q2 / r * conv."""
    box, system = shipped
    charges = json.loads((GOLD / "ref_charges.json").read_text())["charges"]
    tab = system.sol_tab.copy(); tab[:, 0] = charges
    conv = ol.lib().mco_const_conv_intended()
    s = boxio.System(**{**system.__dict__, "sol_tab": tab, "conv": conv})
    eng = _engine(s, box.n, n_chains=1)
    eng.set_state(box.com, box.hoek)
    tot, _, _ = eng.total_energy()
    o = ol.OracleChain(s, box.com, box.hoek)
    e = o.total_energy()
    assert abs(e - 2796.8745663) > 1.0            # the Coulomb term really contributes here
    assert _close(tot[0], e)
    ss, en = eng.matrix(0)
    oss, oen = o.matrix()
    assert _close(ss, oss, scale=np.abs(oss) + 1e-3) and _close(en, oen)
    # and with conv = 0 the same charges change nothing (what the reference binary does, ref_charges.json)
    s0 = boxio.System(**{**system.__dict__, "sol_tab": tab, "conv": 0.0})
    eng0 = _engine(s0, box.n, n_chains=1)
    eng0.set_state(box.com, box.hoek)
    assert abs(eng0.total_energy()[0][0] - 2796.8745663) <= PRINT11 * 2796.87


def test_row_energy_is_calculateLJ(shipped):
    box, system = shipped
    eng = _engine(system, box.n, n_chains=2)
    eng.set_state(box.com, box.hoek)
    o = ol.OracleChain(system, box.com, box.hoek)
    for i in (1, 6, 17, 30):
        row, ei = eng.row_energy(1, i)
        _, orow, oei = o.row_energy(i - 1)
        assert np.array_equal(row != 0, orow != 0)
        assert _close(row, orow, scale=np.abs(orow) + 1e-3) and _close(ei, oei)


def test_single_step_mode_replays_reference_trace(shipped):
    """mcb_trial_delta / mcb_commit fed the trial moves the reference itself made (DUMP) and its decisions:
    TOTENG of every trial against the reference's printed OUT rows and against the oracle."""
    box, system = shipped
    t = load_trace("ref_trace_10k.npz")
    rows, moved, acc = t["out_rows"], t["moved"], t["accepted"]
    nstep = len(acc)          # all 10^4 steps of the fixture (BASELINE.json: >= 10^4)
    assert nstep >= 10000
    eng = _engine(system, box.n, n_chains=1)
    eng.set_state(box.com, box.hoek)
    e0 = eng.total_energy()[0][0]
    o = ol.OracleChain(system, box.com, box.hoek)
    o.total_energy()
    cur = e0
    for k in range(nstep):
        i1 = int(rows[k, C_RSOLV])
        d_e, e_new = eng.trial_delta([i1], moved[k, 0:3], moved[k, 3:6])
        _, tr = o.step_given_move(i1 - 1, moved[k, 0:3], moved[k, 3:6], 0.0 if acc[k] else 1.0)
        scale = max(abs(tr[0]), abs(tr[1]))
        assert abs(e_new[0] - tr[0]) <= RTOL * scale, (k, e_new[0], tr[0])
        assert abs(d_e[0] - (tr[0] - tr[1])) <= RTOL * scale, (k, d_e[0], tr[0] - tr[1])
        assert abs(e_new[0] - rows[k, C_TOT]) <= 2 * PRINT11 * max(abs(rows[k, C_TOT]), abs(rows[k, C_OLD])), k
        eng.commit([1 if acc[k] else 0])
        if acc[k]:
            cur = e_new[0]
    st = eng.stats()
    assert abs(st["toteng"][0] - cur) <= RTOL * abs(cur)
    # the tracked energy equals a fresh recompute of the final configuration
    assert abs(eng.total_energy()[0][0] - cur) <= RTOL * abs(cur)
    com, hoek = eng.get_state()
    ocom, ohoek = o.get_state()
    assert np.abs(com[0] - ocom).max() == 0.0 and np.abs(hoek[0] - ohoek).max() == 0.0


def _compare_chain(st, c, so, com, hoek, o):
    assert st["step"][c] == so["step"]
    assert st["nsuc"][c] == so["nsuc"], (c, st["nsuc"][c], so["nsuc"])
    assert st["naccept"][c] == so["naccept"] and st["prod_acc"][c] == so["prod_acc"]
    assert st["sitepairs"][c] == so["sitepairs"]
    assert abs(st["toteng"][c] - so["toteng"]) <= RTOL * abs(so["toteng"])
    assert abs(st["production"][c] - so["production"]) <= RTOL * abs(so["production"]) + 1e-300
    assert abs(st["dposmax"][c] - so["dposmax"]) <= 1e-15 and abs(st["dhoekmax"][c] - so["dhoekmax"]) <= 1e-15
    ocom, ohoek = o.get_state()
    assert np.abs(com[c] - ocom).max() <= 1e-12 and np.abs(hoek[c] - ohoek).max() <= 1e-12


def test_batched_mode_with_driver_uniforms(shipped):
    """mcb_run_steps fed a uniform buffer: same accept/reject sequence, state and counters as the oracle fed
    the same buffer, for 10^4 steps per chain."""
    box, system = shipped
    nch, nsteps = 6, 10000
    rng = np.random.default_rng(20261106)
    u = rng.random((nch, 8 * nsteps))
    eng = _engine(system, box.n, n_chains=nch)
    eng.set_state(box.com, box.hoek)
    eng.set_mc_params(start_prod=nsteps // 2)
    consumed, steps = eng.run_steps(nsteps, u)
    st = eng.stats()
    com, hoek = eng.get_state()
    assert np.all(steps == nsteps)
    for c in range(nch):
        o = ol.OracleChain(system, box.com, box.hoek)
        o.set_mc_params(start_prod=nsteps // 2)
        o.total_energy()
        done, used, _ = o.run_steps(nsteps, u[c], trace=False)
        assert done == nsteps and used == consumed[c] == st["uniforms_used"][c]
        _compare_chain(st, c, o.stats(), com, hoek, o)


@pytest.mark.parametrize("lookahead", [1, 2, 4, 8])
def test_batched_mode_builtin_stream_and_resume(shipped, lookahead):
    """Built-in counter stream, run in three launches (resume across launches must be seamless), for every
    speculative look-ahead depth: the chain must be the sequential one, decision for decision."""
    box, system = shipped
    nch = 12
    seeds = np.arange(1, nch + 1, dtype=np.uint64)
    eng = _engine(system, box.n, n_chains=nch, lookahead=lookahead)
    eng.set_state(box.com, box.hoek)
    eng.seed(seeds)
    eng.set_mc_params(nadj=50, padj=0.1, temperature=np.linspace(250.0, 400.0, nch))
    for n in (1, 4999, 5000):
        eng.run_steps(n)
    st = eng.stats()
    com, hoek = eng.get_state()
    temps = np.linspace(250.0, 400.0, nch)
    for c in range(nch):
        o = ol.OracleChain(system, box.com, box.hoek)
        o.set_mc_params(temperature=temps[c], nadj=50, padj=0.1)
        o.total_energy()
        o.run_steps_rng(10000, ol.rng_seed(int(seeds[c])), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    # size-independent property: the running energy equals a fresh full-box recompute
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))


def test_uniform_buffer_runs_dry(shipped):
    box, system = shipped
    eng = _engine(system, box.n, n_chains=2)
    eng.set_state(box.com, box.hoek)
    u = np.random.default_rng(3).random((2, 100))
    consumed, steps = eng.run_steps(50, u)
    o = ol.OracleChain(system, box.com, box.hoek)
    o.total_energy()
    done, used, _ = o.run_steps(50, u[0], trace=False)
    assert steps[0] == done and consumed[0] == used and done < 50 and np.all(consumed <= 100)


@pytest.mark.parametrize("matrix_mode,lookahead", [(0, 4), (0, 1), (1, 4), (1, 2)])
def test_batched_mode_replica_box(shipped, matrix_mode, lookahead):
    """N = 240 (pair matrix in global memory), and the shipped box forced onto the global-matrix path."""
    box, system = shipped
    b = boxio.replicate(box, 2) if matrix_mode == 0 else box
    s = _sys_for(system, b)
    nch, nsteps = 3, 3000
    eng = _engine(s, b.n, n_chains=nch, matrix_mode=matrix_mode, lookahead=lookahead)
    eng.set_state(b.com, b.hoek)
    eng.seed(np.array([11, 12, 13], dtype=np.uint64))
    eng.run_steps(nsteps)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        o.run_steps_rng(nsteps, ol.rng_seed(11 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)


@pytest.mark.parametrize("skin", ["0.25", "1.0", None])
def test_large_box_neighbour_lists(shipped, skin, monkeypatch):
    """N = 240 runs through k_run_vl (Verlet lists in global memory, the default for boxes of this size).  A tiny skin forces the fall-back to the full
    scan and list rebuilds every few accepted moves; the chain must stay the oracle's either way, and equal to the
    list-free kernel's."""
    box, system = shipped
    b = boxio.replicate(box, 2)
    s = _sys_for(system, b)
    if skin is None:
        monkeypatch.setenv("MCB200_NO_VERLET", "1")
    else:
        monkeypatch.setenv("MCB200_SKIN", skin)
    nch, nsteps = 2, 6000
    eng = _engine(s, b.n, n_chains=nch)
    eng.set_state(b.com, b.hoek)
    eng.seed(np.array([21, 22], dtype=np.uint64))
    eng.run_steps(nsteps // 2)
    eng.run_steps(nsteps // 2)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        o.run_steps_rng(nsteps, ol.rng_seed(21 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))


def test_neighbour_lists_follow_state_changes(shipped, monkeypatch):
    """Lists are kept from one launch to the next only while k_run_vl alone has moved molecules: after new coordinates
    from the host (here: two distant molecules swapped, so their lists are wrong in every entry) the next launch must
    rebuild them.  The chain after the swap is compared with an oracle started from the swapped state."""
    box, system = shipped
    b = boxio.replicate(box, 2)
    s = _sys_for(system, b)
    eng = _engine(s, b.n, n_chains=2)
    eng.set_state(b.com, b.hoek)
    eng.seed(np.array([31, 32], dtype=np.uint64))
    eng.run_steps(1500)
    eng.run_steps(500)                       # second launch: lists trusted
    com, hoek = eng.get_state()
    d = np.linalg.norm(com[0] - com[0][0], axis=1)
    far = int(np.argmax(d))
    com2, hoek2 = com.copy(), hoek.copy()
    com2[:, [0, far]] = com2[:, [far, 0]]
    eng.set_state(com2, hoek2)
    eng.reset_stats()
    eng.set_mc_params()                      # step sizes back to the defaults the oracle starts from
    eng.seed(np.array([41, 42], dtype=np.uint64))
    eng.run_steps(2500)
    st = eng.stats()
    com3, hoek3 = eng.get_state()
    for c in range(2):
        o = ol.OracleChain(s, com2[c], hoek2[c])
        o.total_energy()
        o.run_steps_rng(2500, ol.rng_seed(41 + c), trace=False)
        _compare_chain(st, c, o.stats(), com3, hoek3, o)


@pytest.mark.parametrize("width", ["2", "3", "4"])
def test_steps_in_flight_edges(shipped, width, monkeypatch):
    """k_run_vl2 with 2, 3 and 4 warps per chain on the 2x2x2 replica: launches shorter than a batch, an adaptation
    window (nadj = 3) that is not a multiple of the batch, a production start in the middle of a batch, and a skin so
    small that most steps fall back to the full scan on warp 0 - still the oracle's chain."""
    box, system = shipped
    b = boxio.replicate(box, 2)
    s = _sys_for(system, b)
    monkeypatch.setenv("MCB200_VL2_W", width)
    for skin in ("2.0", "0.2"):
        monkeypatch.setenv("MCB200_SKIN", skin)
        eng = _engine(s, b.n, n_chains=2)
        eng.set_state(b.com, b.hoek)
        eng.set_mc_params(nadj=3, start_prod=50)
        eng.seed(np.array([81, 82], dtype=np.uint64))
        total = 0
        for n in (1, 2, 3, 5, 7, 1, 64, 65, 300, 0, 11):
            eng.run_steps(n)
            total += n
        st = eng.stats()
        com, hoek = eng.get_state()
        for c in range(2):
            o = ol.OracleChain(s, b.com, b.hoek)
            o.set_mc_params(nadj=3, start_prod=50)
            o.total_energy()
            o.run_steps_rng(total, ol.rng_seed(81 + c), trace=False)
            _compare_chain(st, c, o.stats(), com, hoek, o)


def test_neighbour_lists_default_size(shipped):
    """N = 810 (3x3x3 replica): k_run_vl by default; 1500 steps of two chains against the oracle."""
    box, system = shipped
    b = boxio.replicate(box, 3)
    s = _sys_for(system, b)
    eng = _engine(s, b.n, n_chains=2)
    eng.set_state(b.com, b.hoek)
    eng.seed(np.array([5, 6], dtype=np.uint64))
    eng.run_steps(1500)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(2):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        o.run_steps_rng(1500, ol.rng_seed(5 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)


@pytest.mark.parametrize("k", [1, 2])
def test_large_solute(shipped, k):
    """A solute of 37 atoms: 10 pseudo-molecules of four sites (the last one padded), i.e. more solute slots than one
    pass over the partner slots holds, in the small box (k_run_spec) and in the 2x2x2 replica (k_run_vl); start-up
    energy, single-step mode and 1500 batched steps against the oracle."""
    box, system = shipped
    b = boxio.replicate(box, k) if k > 1 else box
    base = _sys_for(system, b) if k > 1 else system
    rng = np.random.default_rng(37)
    reps = [base.solute] + [base.solute + rng.uniform(-0.8, 0.8, size=(1, 3)) + 0.3 * rng.standard_normal(base.solute.shape) for _ in range(3)]
    sol = np.concatenate(reps)[:37]
    tab = np.concatenate([base.sol_tab] * 4)[:37]
    syms = (base.sol_sym * 4)[:37]
    s = boxio.System(**{**base.__dict__, "solute": sol, "sol_tab": tab, "sol_sym": syms})
    nch = 2
    eng = _engine(s, b.n, n_chains=nch)
    eng.set_state(b.com, b.hoek)
    tot, esol, _ = eng.total_energy()
    o0 = ol.OracleChain(s, b.com, b.hoek)
    e0 = o0.total_energy()
    assert abs(tot[0] - e0) <= RTOL * abs(e0)
    eng.seed(np.array([71, 72], dtype=np.uint64))
    eng.run_steps(1500)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        o.run_steps_rng(1500, ol.rng_seed(71 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))


def test_per_chain_solute_conformers(shipped):
    """Config 4 shape: every chain sees its own solute coordinates."""
    box, system = shipped
    nch = 4
    rng = np.random.default_rng(5)
    sol = np.stack([system.solute + 0.05 * rng.standard_normal(system.solute.shape) for _ in range(nch)])
    tabs = np.stack([system.sol_tab] * nch)
    eng = _engine(system, box.n, n_chains=nch)
    eng.set_solute(sol, tabs)
    eng.set_state(box.com, box.hoek)
    tot = eng.total_energy()[0]
    eng.seed(np.arange(100, 100 + nch, dtype=np.uint64))
    eng.run_steps(2000)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(system, box.com, box.hoek, solute=sol[c])
        assert abs(o.total_energy() - tot[c]) <= RTOL * abs(tot[c])
        o.run_steps_rng(2000, ol.rng_seed(100 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    assert len(set(np.round(tot, 6))) == nch


def test_config4_dihedral_sweep(shipped):
    """BASELINE config 4 in miniature: 36 solute conformers (dihedral 9-8-1-2 set to 0..350 degrees by the host-side
    generator) x 3 chains each; every chain must be the oracle's chain for its own conformer, and the per-angle
    statistics are what the end-of-run gather collects."""
    from montecarlo_b200 import dihedral
    box, system = shipped
    conf, atoms = dihedral.conformer_sweep(box.solute, box.sol_sym, box.dihedrals[0][:2], np.arange(0.0, 360.0, 10.0))
    per = 3
    nch = conf.shape[0] * per
    sol = np.repeat(conf, per, axis=0)
    eng = _engine(system, box.n, n_chains=nch)
    eng.set_solute(sol, np.broadcast_to(system.sol_tab, (nch,) + system.sol_tab.shape).copy())
    eng.set_state(box.com, box.hoek)
    tot0 = eng.total_energy()[0]
    seeds = np.arange(1, nch + 1, dtype=np.uint64)
    eng.seed(seeds)
    eng.run_steps(1000)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in (0, 1, 50, nch - 1):
        o = ol.OracleChain(system, box.com, box.hoek, solute=sol[c])
        assert abs(o.total_energy() - tot0[c]) <= RTOL * abs(tot0[c])
        o.run_steps_rng(1000, ol.rng_seed(int(seeds[c])), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    e0 = tot0.reshape(-1, per)
    assert np.all(e0 == e0[:, :1]) and len(np.unique(np.round(e0[:, 0], 6))) == 36   # same conformer, same start energy
    per_angle_mean = (st["sum_e"] / st["n_stat"]).reshape(-1, per).mean(axis=1)
    assert per_angle_mean.shape == (36,) and np.all(np.isfinite(per_angle_mean))
    # the end-of-run collective per angle (one group per conformer) and what consumes it: the reference's window hopping
    # (PostProcess.F90:260-304) over the gathered per-chain mean energies, one "run" per chain index inside an angle
    g = eng.reduce_stats(per)
    assert list(g["n_chains"]) == [per] * 36 and np.allclose(g["mean_e"], (st["sum_e"].reshape(-1, per).sum(axis=1) / st["n_stat"].reshape(-1, per).sum(axis=1)), rtol=1e-13)
    from montecarlo_b200 import postprocess
    e_run_win = eng.gather_chains()["mean_e"].reshape(36, per).T.copy()          # [run, window]
    rng = np.random.default_rng(11)
    old, now, acc = postprocess.window_hopping(e_run_win, 300.0, iter(rng.random(4 * per)))
    assert old[0] == int(np.argmin(e_run_win[0])) and np.all((now - old) % 36 % 34 == 1) and len(acc) == per


def test_quirk_31_and_odd_sizes(shipped):
    """N = 33 random box: molecule 31 is never picked, 30 twice as often (MonteCarlo.F90:698); N not a multiple of 4."""
    box, system = shipped
    rng = np.random.default_rng(9)
    n = 33
    boxl = 9.0
    com = rng.uniform(-boxl, boxl, (n, 3)); hoek = rng.uniform(-np.pi, np.pi, (n, 3))
    s = boxio.System(**{**system.__dict__, "boxl": boxl})
    for quirk in (True, False):
        eng = _engine(s, n, n_chains=2, quirk_31=quirk)
        eng.set_state(com, hoek)
        eng.seed(np.array([7, 8], dtype=np.uint64))
        eng.run_steps(1500)
        st = eng.stats()
        c_, h_ = eng.get_state()
        for c in range(2):
            o = ol.OracleChain(s, com, hoek)
            o.set_mc_params(quirk31=1 if quirk else 0)
            o.total_energy()
            o.run_steps_rng(1500, ol.rng_seed(7 + c), trace=False)
            _compare_chain(st, c, o.stats(), c_, h_, o)
        moved31 = np.abs(c_[:, 30] - com[30]).max() > 0
        assert moved31 == (not quirk)


def test_edge_cases(shipped):
    box, system = shipped
    # no solute at all
    eng = _engine(system, box.n, n_chains=1, with_solute=False)
    eng.set_state(box.com, box.hoek)
    tot, esol, _ = eng.total_energy()
    o = ol.OracleChain(system, box.com, box.hoek)
    o.total_energy()
    ss, en = o.matrix()
    assert esol[0] == 0.0 and abs(tot[0] - np.triu(ss, 1).sum()) <= RTOL * abs(tot[0])
    # a single molecule: only the solute term
    eng1 = _engine(system, 1, n_chains=1)
    eng1.set_state(box.com[:1], box.hoek[:1])
    t1 = eng1.total_energy()[0][0]
    assert abs(t1 - en[0]) <= RTOL * abs(en[0])
    eng1.run_steps(10)
    assert eng1.stats()["step"][0] == 10
    # argument errors come back as status codes with a message, not as crashes
    from montecarlo_b200.engine import McbError
    with pytest.raises(McbError):
        eng.trial_delta([0], box.com[0], box.hoek[0])
    with pytest.raises(McbError):
        eng.trial_delta([31], box.com[0], box.hoek[0])
    with pytest.raises(McbError):
        eng.run_steps(5, np.zeros((1, 4)))


def test_overlap_gives_nan_as_a_value_and_rollback_is_clean(shipped):
    """A trial move that lands exactly on another molecule: r = 0 for the coinciding sites, the reference's
    LJ_FORMULA gives Inf - Inf = NaN (LennardJones.f90:185) and its driver clamps and rejects (MonteCarlo.F90:421-428).
    The single-step entry points return the NaN as a value (no clamp in the library) and a discarded trial leaves
    the chain exactly as it was."""
    box, system = shipped
    eng = _engine(system, box.n, n_chains=2)
    eng.set_state(box.com, box.hoek)
    e0 = eng.total_energy()[0].copy()
    d_e, e_new = eng.trial_delta([5, 5], np.stack([box.com[9], box.com[4]]), np.stack([box.hoek[9], box.hoek[4]]))
    assert not np.isfinite(d_e[0]) and not np.isfinite(e_new[0])       # molecule 5 dropped onto molecule 10
    assert abs(d_e[1]) <= RTOL * abs(e0[1]) and abs(e_new[1] - e0[1]) <= RTOL * abs(e0[1])   # molecule 5 "moved" onto itself
    eng.commit([0, 1])
    st = eng.stats()
    assert st["toteng"][0] == e0[0] and abs(st["toteng"][1] - e0[1]) <= RTOL * abs(e0[1])
    com, hoek = eng.get_state()
    assert np.array_equal(com[0], box.com) and np.array_equal(com[1], box.com)
    assert np.all(np.abs(eng.total_energy()[0] - e0) <= RTOL * np.abs(e0))
    # the oracle sees the same thing: its trial energy is clamped to HUGE and the step is rejected
    o = ol.OracleChain(system, box.com, box.hoek)
    o.total_energy()
    _, tr = o.step_given_move(4, box.com[9], box.hoek[9], 0.0)
    assert tr[0] > 1e307 and tr[5] == 0.0 and o.stats()["nclamp"] == 1


def test_full_size_properties(shipped):
    """BASELINE config 2 size (1024 chains of the shipped box): determinism across identical seeds, independence
    across different seeds, and running energy == fresh recompute for every chain."""
    box, system = shipped
    nch = 1024
    eng = _engine(system, box.n, n_chains=nch)
    eng.set_state(box.com, box.hoek)
    seeds = np.arange(1, nch + 1, dtype=np.uint64)
    seeds[512:] = seeds[:512]          # second half repeats the first
    eng.seed(seeds)
    eng.run_steps(2000)
    st = eng.stats()
    com, _ = eng.get_state()
    assert np.array_equal(st["toteng"][:512], st["toteng"][512:]) and np.array_equal(com[:512], com[512:])
    assert len(np.unique(st["toteng"][:512])) == 512
    assert np.all(st["step"] == 2000) and np.all(st["n_clamped"] == 0)
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))
    o = ol.OracleChain(system, box.com, box.hoek)
    o.total_energy()
    o.run_steps_rng(2000, ol.rng_seed(301), trace=False)
    assert st["nsuc"][300] == st["nsuc"][812] == o.stats()["nsuc"]


def test_fp64_peak_microbenchmark_is_sane():
    from montecarlo_b200.engine import measure_fp64_peak
    tf = measure_fp64_peak(0, 50.0)
    assert 10.0 < tf < 60.0, tf


# ------------------------------------------------------------------------------------------------------------------
# Kernel paths that are shipped but were not covered before: k_run_vl (one warp per chain with lists), k_run_steps with
# the pair matrix in global memory, every COUL = true instantiation of the stepping kernels, single-step mode on a
# global-matrix box, and the stream position across mcb_reset_stats.

def _coulomb_system(system, box=None):
    """conv = the physically meant factor (LennardJones.f90:52-56 before the `* 10**(-8)`), solute charges from the
    reference's CHELPG stub run: the Coulomb sum then really moves energies and decisions."""
    charges = json.loads((GOLD / "ref_charges.json").read_text())["charges"]
    tab = system.sol_tab.copy(); tab[:, 0] = charges
    kw = {"sol_tab": tab, "conv": ol.lib().mco_const_conv_intended()}
    if box is not None:
        kw.update(boxl=box.boxl, solute=box.solute)
    return boxio.System(**{**system.__dict__, **kw})


@pytest.mark.parametrize("case", ["vl_default_skin", "vl_small_skin"])
def test_one_warp_list_kernel(shipped, case, monkeypatch):
    """MCB200_NO_VL2=1: the N = 240 box steps with k_run_vl (one warp per chain, neighbour lists), the kernel that is
    taken whenever k_run_vl2 does not fit.  Two launches, then new coordinates from the host and a third launch."""
    box, system = shipped
    b = boxio.replicate(box, 2)
    s = _sys_for(system, b)
    monkeypatch.setenv("MCB200_NO_VL2", "1")
    if case == "vl_small_skin":
        monkeypatch.setenv("MCB200_SKIN", "0.25")
    nch, nsteps = 2, 4000
    eng = _engine(s, b.n, n_chains=nch)
    eng.set_state(b.com, b.hoek)
    eng.seed(np.array([21, 22], dtype=np.uint64))
    eng.run_steps(nsteps // 2)
    eng.run_steps(nsteps // 2)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        o.run_steps_rng(nsteps, ol.rng_seed(21 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))
    # lists must be rebuilt after coordinates arrive from the host (two distant molecules swapped)
    d = np.linalg.norm(com[0] - com[0][0], axis=1)
    far = int(np.argmax(d))
    com2, hoek2 = com.copy(), hoek.copy()
    com2[:, [0, far]] = com2[:, [far, 0]]
    eng.set_state(com2, hoek2)
    eng.reset_stats()
    eng.set_mc_params()
    eng.seed(np.array([41, 42], dtype=np.uint64))
    eng.run_steps(1500)
    st = eng.stats()
    com3, hoek3 = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(s, com2[c], hoek2[c])
        o.total_energy()
        o.run_steps_rng(1500, ol.rng_seed(41 + c), trace=False)
        _compare_chain(st, c, o.stats(), com3, hoek3, o)


def test_one_warp_list_kernel_large_solute(shipped, monkeypatch):
    """k_run_vl with more solute slots than one pass over the partner slots holds (37 solute atoms, N = 240)."""
    monkeypatch.setenv("MCB200_NO_VL2", "1")
    test_large_solute(shipped, 2)


@pytest.mark.parametrize("which", ["replica", "forced_global"])
def test_driver_uniforms_with_global_matrix(shipped, which):
    """k_run_steps<*, MAT_SMEM = false>: the driver's own uniforms on a box whose pair matrix lives in global memory -
    the 2x2x2 replica, and the shipped box with matrix_mode = 1."""
    box, system = shipped
    b = boxio.replicate(box, 2) if which == "replica" else box
    s = _sys_for(system, b)
    nch, nsteps = 3, 3000
    u = np.random.default_rng(77).random((nch, 8 * nsteps))
    eng = _engine(s, b.n, n_chains=nch, matrix_mode=1)
    eng.set_state(b.com, b.hoek)
    eng.set_mc_params(start_prod=1000, nadj=64)
    consumed, steps = eng.run_steps(nsteps, u)
    st = eng.stats()
    com, hoek = eng.get_state()
    assert np.all(steps == nsteps)
    for c in range(nch):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.set_mc_params(start_prod=1000, nadj=64)
        o.total_energy()
        done, used, _ = o.run_steps(nsteps, u[c], trace=False)
        assert done == nsteps and used == consumed[c] == st["uniforms_used"][c]
        _compare_chain(st, c, o.stats(), com, hoek, o)
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))


@pytest.mark.parametrize("path", ["spec_n30", "spec_n30_lookahead4", "ubuf_n30", "ubuf_n240", "vl2_n240", "vl_n240", "spec_n240"])
def test_coulomb_stepping_kernels(shipped, path, monkeypatch):
    """conv != 0 through every stepping kernel (the reference's CONV is 0, LennardJones.f90:57, but the sum q1 q2 / r
    must be carried, :52-56, :80-82): k_run_spec<true>, k_run_steps<true, *>, k_run_vl2<true>, k_run_vl<true>."""
    box, system = shipped
    big = path.endswith("n240")
    b = boxio.replicate(box, 2) if big else box
    s = _coulomb_system(system, b if big else None)
    if path == "vl_n240":
        monkeypatch.setenv("MCB200_NO_VL2", "1")
    if path == "spec_n240":
        monkeypatch.setenv("MCB200_NO_VERLET", "1")
    nch, nsteps = 3, 2500 if big else 6000
    eng = _engine(s, b.n, n_chains=nch, lookahead=4 if path.endswith("lookahead4") else 0)
    eng.set_state(b.com, b.hoek)
    tot0 = eng.total_energy()[0]
    o0 = ol.OracleChain(s, b.com, b.hoek)
    e0 = o0.total_energy()
    lj_only = ol.OracleChain(_sys_for(system, b) if big else system, b.com, b.hoek).total_energy()
    assert abs(e0 - lj_only) > 1.0 and np.all(np.abs(tot0 - e0) <= RTOL * abs(e0))
    seeds = np.array([61, 62, 63], dtype=np.uint64)
    if path.startswith("ubuf"):
        u = np.random.default_rng(99).random((nch, 8 * nsteps))
        consumed, steps = eng.run_steps(nsteps, u)
        assert np.all(steps == nsteps)
    else:
        eng.seed(seeds)
        eng.run_steps(nsteps // 2)
        eng.run_steps(nsteps - nsteps // 2)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(nch):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        if path.startswith("ubuf"):
            done, used, _ = o.run_steps(nsteps, u[c], trace=False)
            assert used == consumed[c]
        else:
            o.run_steps_rng(nsteps, ol.rng_seed(int(seeds[c])), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))


@pytest.mark.parametrize("coulomb,kbox,nmoves", [(False, 2, 1200), (True, 2, 1200), (False, 4, 150)])
def test_single_step_mode_on_global_matrix_box(shipped, coulomb, kbox, nmoves):
    """mcb_trial_delta / mcb_commit on the N = 240 replica (k_trial / k_commit with the matrix in global memory; with
    conv != 0 also k_trial<true>) and on the N = 1920 replica, whose image does not fit in shared memory: trial moves
    drawn the way MCINIT draws them, the oracle deciding."""
    box, system = shipped
    b = boxio.replicate(box, kbox)
    s = _coulomb_system(system, b) if coulomb else _sys_for(system, b)
    eng = _engine(s, b.n, n_chains=2)
    eng.set_state(b.com, b.hoek)
    e0 = eng.total_energy()[0]
    o = ol.OracleChain(s, b.com, b.hoek)
    assert abs(o.total_energy() - e0[0]) <= RTOL * abs(e0[0])
    rng = np.random.default_rng(123)
    nacc = 0
    for k in range(nmoves):
        u = rng.random(8)
        used, tr = o.step(u)                      # MCINIT + calculateLJ + METROPOLIS in the oracle
        i = int(tr[4]); accepted = tr[5] != 0.0
        ocom, ohoek = o.get_state()
        if accepted:
            nc, nh = ocom[i], ohoek[i]
        else:                                     # rejected: replay MCINIT on a copy to get the trial position
            o2 = ol.OracleChain(s, ocom, ohoek)
            o2.set_mc_params(dposmax=tr[7], dhoekmax=tr[8])
            j = o2.L.mco_mcinit(o2.h, ol._ptr(np.ascontiguousarray(u[:7])))
            assert j == i
            c2, h2 = o2.get_state()
            nc, nh = c2[i], h2[i]
        # chain 0 follows the oracle's decisions; chain 1 rejects everything and must stay where it started
        d_e, e_new = eng.trial_delta([i + 1, i + 1], np.stack([nc, nc]), np.stack([nh, nh]))
        scale = max(abs(tr[0]), abs(tr[1]))
        assert abs(e_new[0] - tr[0]) <= RTOL * scale, (k, e_new[0], tr[0])
        assert abs(d_e[0] - (tr[0] - tr[1])) <= RTOL * scale, (k, d_e[0], tr[0] - tr[1])
        eng.commit([1 if accepted else 0, 0])
        nacc += accepted
    assert nacc > nmoves // 12
    st = eng.stats()
    so = o.stats()
    assert abs(st["toteng"][0] - so["toteng"]) <= RTOL * abs(so["toteng"]) and st["toteng"][1] == e0[1]
    com, hoek = eng.get_state()
    ocom, ohoek = o.get_state()
    assert np.abs(com[0] - ocom).max() == 0.0 and np.abs(hoek[0] - ohoek).max() == 0.0
    assert np.array_equal(com[1], b.com) and np.array_equal(hoek[1], b.hoek)
    fresh = eng.total_energy()[0]
    assert abs(fresh[0] - so["toteng"]) <= RTOL * abs(fresh[0]) and abs(fresh[1] - e0[1]) <= RTOL * abs(e0[1])


@pytest.mark.parametrize("k", [1, 2])
def test_reset_stats_keeps_the_stream_position(shipped, k):
    """mcb_reset_stats zeroes the counters but not the position in the built-in uniform stream: equilibrate, reset,
    produce must continue the stream (an oracle that simply keeps going), not replay the equilibration's numbers."""
    box, system = shipped
    b = boxio.replicate(box, 2) if k == 2 else box
    s = _sys_for(system, b) if k == 2 else system
    n_eq, n_prod = 1500, 2500             # the reset falls on an adaptation boundary (nadj = 100), where NACCEPT is 0 anyway
    seeds = np.array([91, 92], dtype=np.uint64)
    eng = _engine(s, b.n, n_chains=2)
    eng.set_state(b.com, b.hoek)
    eng.seed(seeds)
    eng.run_steps(n_eq)
    eq = eng.stats()
    eng.reset_stats()
    assert np.all(eng.stats()["step"] == 0) and np.all(eng.stats()["dposmax"] == eq["dposmax"])
    eng.run_steps(n_prod)
    st = eng.stats()
    com, hoek = eng.get_state()
    for c in range(2):
        o = ol.OracleChain(s, b.com, b.hoek)
        o.total_energy()
        rng = ol.rng_seed(int(seeds[c]))
        o.run_steps_rng(n_eq, rng, trace=False)
        so_eq = o.stats()
        o.run_steps_rng(n_prod, rng, trace=False)
        so = o.stats()
        assert st["step"][c] == n_prod and st["nsuc"][c] == so["nsuc"] - so_eq["nsuc"]
        assert st["sitepairs"][c] == so["sitepairs"] - so_eq["sitepairs"]
        assert abs(st["toteng"][c] - so["toteng"]) <= RTOL * abs(so["toteng"])
        ocom, ohoek = o.get_state()
        assert np.abs(com[c] - ocom).max() <= 1e-12 and np.abs(hoek[c] - ohoek).max() <= 1e-12
    # a replay of the equilibration's numbers would have reproduced the first n_eq decisions: it must not
    eng2 = _engine(s, b.n, n_chains=2)
    eng2.set_state(b.com, b.hoek)
    eng2.seed(seeds)
    eng2.run_steps(n_eq)
    c_eq, _ = eng2.get_state()
    eng2.reset_stats()
    eng2.run_steps(1)
    # re-seeding is what rewinds the stream
    eng2.seed(seeds)
    eng2.set_state(b.com, b.hoek)
    eng2.set_mc_params()
    eng2.reset_stats()
    eng2.run_steps(n_eq)
    c_again, _ = eng2.get_state()
    assert np.array_equal(c_eq, c_again)


def test_commit_needs_a_pending_trial(shipped):
    """mcb_commit without a trial, or after new coordinates have replaced the state the trial was evaluated against, is an
    argument error - not a silent commit of a stale move."""
    box, system = shipped
    from montecarlo_b200.engine import McbError
    eng = _engine(system, box.n, n_chains=1)
    eng.set_state(box.com, box.hoek)
    e0 = eng.total_energy()[0][0]
    with pytest.raises(McbError, match="no trial is pending"):
        eng.commit([1])
    eng.trial_delta([3], box.com[2] + 0.2, box.hoek[2])
    eng.set_state(box.com, box.hoek)                 # discards the parked trial
    with pytest.raises(McbError, match="no trial is pending"):
        eng.commit([1])
    assert eng.total_energy()[0][0] == e0 and np.array_equal(eng.get_state()[0][0], box.com)


def test_row_energy_refuses_while_a_trial_is_pending(shipped):
    box, system = shipped
    from montecarlo_b200.engine import McbError
    eng = _engine(system, box.n, n_chains=1)
    eng.set_state(box.com, box.hoek)
    eng.total_energy()
    eng.trial_delta([3], box.com[2] + 0.1, box.hoek[2])
    with pytest.raises(McbError, match="pending"):
        eng.row_energy(0, 3)
    eng.commit([1])
    row, ei = eng.row_energy(0, 3)
    com, hoek = eng.get_state()
    o = ol.OracleChain(system, com[0], hoek[0])
    _, orow, oei = o.row_energy(2)
    assert _close(row, orow, scale=np.abs(orow) + 1e-3) and _close(ei, oei)


@pytest.mark.parametrize("stream", ["caller", "builtin"])
def test_replica_exchange_ladder(shipped, stream):
    """Temperature ladder with exchange moves between launches (mcb_exchange): eight chains at 250..600 K, rounds of
    300 steps, even pairs then odd pairs; every decision, every temperature and the chains afterwards must be those of
    the oracle applying the same rule (the reference's METROPOLIS, MonteCarlo.F90:728-744) with the same uniforms."""
    box, system = shipped
    nch, rounds, nsteps = 8, 10, 300
    temps = np.linspace(250.0, 600.0, nch)
    seeds = np.arange(501, 501 + nch, dtype=np.uint64)
    eng = _engine(system, box.n, n_chains=nch)
    eng.set_state(box.com, box.hoek)
    eng.seed(seeds)
    eng.set_mc_params(temperature=temps)
    eng.total_energy()
    orc = []
    for c in range(nch):
        o = ol.OracleChain(system, box.com, box.hoek)
        o.set_mc_params(temperature=temps[c])
        o.total_energy()
        orc.append((o, ol.rng_seed(int(seeds[c])), ol.rng_exchange_seed(int(seeds[c]))))
    rng = np.random.default_rng(2026)
    n_acc = n_draw = 0
    for rd in range(rounds):
        eng.run_steps(nsteps)
        for o, st, _ in orc:
            # the oracle keeps Metropolis parameters per chain; only temperature and step sizes move in an exchange
            o.run_steps_rng(nsteps, st, trace=False)
        pa = np.arange(rd % 2, nch - 1, 2); pb = pa + 1
        u = rng.random(len(pa)) if stream == "caller" else None
        acc = eng.exchange(pa, pb, u)
        for p, (a, b) in enumerate(zip(pa, pb)):
            uu = u[p] if u is not None else float(ol.rng_uniforms(orc[a][2].copy(), 1)[0])
            oacc, used = ol.exchange(orc[a][0], orc[b][0], uu)
            if used and u is None:
                ol.rng_uniforms(orc[a][2], 1)          # the draw advances chain a's exchange stream
            assert oacc == acc[p], (rd, p)
            n_acc += oacc; n_draw += used
    st = eng.stats()
    com, hoek = eng.get_state()
    assert n_acc > 5 and n_draw > 5 and st["exch_accepted"].sum() == 2 * n_acc
    assert st["exch_attempts"].sum() == 2 * sum(len(np.arange(rd % 2, nch - 1, 2)) for rd in range(rounds))
    assert sorted(st["temperature"]) == sorted(temps)                       # a permutation of the ladder
    for c in range(nch):
        o = orc[c][0]
        assert st["temperature"][c] == o.L.mco_get_temperature(o.h)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    from montecarlo_b200.engine import McbError
    with pytest.raises(McbError, match="at most one pair"):
        eng.exchange([0, 1], [1, 2])


def _lattice_box(n, boxl, rng, jitter=0.35):
    """n molecules on a jittered cubic lattice inside (-boxl, boxl)^3 with random orientations: liquid-like distances
    (no overlaps that would make every energy 1e10) at any n and box size."""
    m = int(np.ceil(n ** (1.0 / 3.0)))
    cells = np.stack(np.meshgrid(*[np.arange(m)] * 3, indexing="ij"), axis=-1).reshape(-1, 3)
    pick = rng.permutation(len(cells))[:n]
    com = (cells[pick] + 0.5) * (2.0 * boxl / m) - boxl + rng.uniform(-jitter, jitter, (n, 3))
    return np.clip(com, -0.999 * boxl, 0.999 * boxl), rng.uniform(-np.pi, np.pi, (n, 3))


SWEEP = [
    # n, boxl, solute atoms, conv != 0, steps      what the case is for
    (2, 6.0, 11, False, 600),      # two molecules: a single partner slot
    (3, 8.0, 0, False, 600),       # no solute: no solute slots in the passes, ENERGY(:) = 0
    (5, 7.0, 1, False, 600),       # one solute atom: one pseudo-molecule with three padding atoms
    (17, 7.5, 11, False, 500),
    (30, 5.2, 11, False, 500),     # box smaller than the cutoff sphere: nearly every molecule is a partner, every pass full
    (31, 8.0, 5, True, 500),       # solute atoms not a multiple of four, Coulomb on
    (32, 8.0, 11, False, 500),     # the last size whose matrix is staged with one 32-molecule chunk
    (48, 9.0, 0, False, 400),      # global matrix, neighbour lists, no solute
    (64, 10.0, 7, True, 400),
    (100, 12.0, 11, False, 400),
    (100, 9.5, 11, False, 400),    # ~45 listed candidates per molecule: the second half of the 64-entry lists is in use
    (150, 9.0, 11, False, 300),    # dense: neighbour lists overflow their 64 entries -> the full scan takes over
]


@pytest.mark.parametrize("case", SWEEP, ids=[f"n{c[0]}_boxl{c[1]}_sol{c[2]}{'_coul' if c[3] else ''}" for c in SWEEP])
def test_random_boxes_against_oracle(shipped, case):
    """Sizes, densities and solutes the fixtures do not hold: start-up energy, then the default stepping path for a few
    hundred steps with the built-in stream, decision for decision against the oracle, on two chains."""
    box, system = shipped
    n, boxl, nsol, coul, steps = case
    rng = np.random.default_rng(1000 + n)
    com, hoek = _lattice_box(n, boxl, rng)
    s = boxio.System(**{**system.__dict__, "boxl": boxl, "solute": system.solute[:nsol].copy(), "sol_tab": system.sol_tab[:nsol].copy(),
                        "conv": ol.lib().mco_const_conv_intended() if coul else system.conv, "sol_sym": list(system.sol_sym[:nsol])})
    eng = _engine(s, n, n_chains=2, with_solute=nsol > 0)
    eng.set_state(com, hoek)
    eng.set_mc_params(nadj=37, start_prod=steps // 3)
    seeds = np.array([11, 12], dtype=np.uint64)
    eng.seed(seeds)
    tot0 = eng.total_energy()[0]
    eng.run_steps(steps // 2)
    eng.run_steps(steps - steps // 2)
    st = eng.stats()
    c_, h_ = eng.get_state()
    fresh = eng.total_energy()[0]
    for c in range(2):
        o = ol.OracleChain(s, com, hoek)
        o.set_mc_params(nadj=37, start_prod=steps // 3)
        e0 = o.total_energy()
        assert abs(tot0[c] - e0) <= RTOL * abs(e0) + 1e-300
        o.run_steps_rng(steps, ol.rng_seed(int(seeds[c])), trace=False)
        _compare_chain(st, c, o.stats(), c_, h_, o)
        assert abs(fresh[c] - st["toteng"][c]) <= RTOL * abs(fresh[c]) + 1e-300


@pytest.mark.parametrize("width", ["1", "2", "3", "4"])
def test_global_image_list_kernel(shipped, width, monkeypatch):
    """k_run_vl2 with the chain's image left in global memory (MCB200_VL2_IMG=global: the variant for more chains than the
    SMs can hold with staged images, and for boxes too large to stage) with 1-4 warps per chain: the scenario of
    test_steps_in_flight_edges - short launches, nadj = 3, a production start inside a batch, a skin that forces full
    scans - plus the Coulomb sum, get/set of the state between launches and a fresh recompute at the end."""
    box, system = shipped
    b = boxio.replicate(box, 2)
    s = _sys_for(system, b)
    monkeypatch.setenv("MCB200_VL2_IMG", "global")
    monkeypatch.setenv("MCB200_VL2_W", width)
    for skin, conv in (("2.0", 0.0), ("0.2", 0.0), ("1.0", ol.lib().mco_const_conv_intended())):
        monkeypatch.setenv("MCB200_SKIN", skin)
        sc = boxio.System(**{**s.__dict__, "conv": conv})
        eng = _engine(sc, b.n, n_chains=3)
        eng.set_state(b.com, b.hoek)
        eng.set_mc_params(nadj=3, start_prod=50)
        eng.seed(np.array([81, 82, 83], dtype=np.uint64))
        total = 0
        for n in (1, 2, 3, 5, 7, 1, 64, 65, 300, 0, 11):
            eng.run_steps(n)
            total += n
            if n == 65:   # the coordinates leave the device and come back: lists are rebuilt, the energies recomputed
                c0, h0 = eng.get_state()
                eng.set_state(c0, h0)
        st = eng.stats()
        com, hoek = eng.get_state()
        fresh = eng.total_energy()[0]
        assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))
        for c in range(3):
            o = ol.OracleChain(sc, b.com, b.hoek)
            o.set_mc_params(nadj=3, start_prod=50)
            o.total_energy()
            o.run_steps_rng(total, ol.rng_seed(81 + c), trace=False)
            so = o.stats()
            # (set_state in the middle re-anchors TOTENG on a fresh recompute: compare it at the tolerance, the rest exactly)
            _compare_chain(st, c, so, com, hoek, o)


def test_box_too_large_for_shared_memory_steps_from_global_memory(shipped):
    """N = 1920 (4x4x4 replica): a chain's image (292 KB) cannot be staged in shared memory, so round 1 could only
    compute its energy.  It now steps with the built-in stream on the global-image kernel; driver-supplied uniforms
    are refused with a message.  600 steps of two chains against the oracle."""
    from montecarlo_b200.engine import McbError
    box, system = shipped
    b = boxio.replicate(box, 4)
    s = _sys_for(system, b)
    eng = _engine(s, b.n, n_chains=2)
    eng.set_state(b.com, b.hoek)
    eng.seed(np.array([21, 22], dtype=np.uint64))
    tot0 = eng.total_energy()[0]
    eng.run_steps(250); eng.run_steps(350)
    st = eng.stats()
    com, hoek = eng.get_state()
    fresh = eng.total_energy()[0]
    assert np.all(np.abs(fresh - st["toteng"]) <= RTOL * np.abs(fresh))
    for c in range(2):
        o = ol.OracleChain(s, b.com, b.hoek)
        e0 = o.total_energy()
        assert abs(tot0[c] - e0) <= RTOL * abs(e0)
        o.run_steps_rng(600, ol.rng_seed(21 + c), trace=False)
        _compare_chain(st, c, o.stats(), com, hoek, o)
    with pytest.raises(McbError):
        eng.run_steps(5, np.zeros((2, 40)))

/* mc_oracle.h — CPU restatement of the reference's classical LJ(+Coulomb) Monte-Carlo path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
 * legs may load this library; the product (montecarlo_b200/, libmcb200.so) never does.
 *
 * Parity status: PINNED.  Every function here is checked (tests/test_oracle_golden.py) against
 * fixtures produced by the reference's own shipped binary MonteCarlo/MonteCarloOpenMP
 * (oracle/run_ref.py, oracle/make_golden.py): start-up energy, the full pair matrix, per-step
 * trial energies, accept/reject decisions and step-size adaptation.
 *
 * All citations are into /root/reference.  Indices are 0-based here; the reference's are 1-based.
 * Arrays of the reference's TYPE(vector) (MC_lib/vector_class.f90:18-20) are double[n][3].
 * Parameter tables are [n][3] = (q, epsilon, sigma) per atom (LennardJones.f90:211-216, :243-245).
 */
#ifndef MC_ORACLE_H
#define MC_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mco_chain mco_chain;

/* ---- constants (MC_lib/lib.f90:15-19; MonteCarlo/LennardJones.f90:53-57) ---- */
double mco_const_pi(void);
double mco_const_tau(void);
double mco_const_e(void);              /* (double)(float)2.71828...: the literal has no D0 suffix */
double mco_const_hartree2kjmol(void);  /* float32 product 4.359744f*6.02214085f*100f */
double mco_const_conv_reference(void); /* == 0.0: 10**(-8) is integer exponentiation */
double mco_const_conv_intended(void);  /* what CONV holds before the *0, times 1.0e4 (physical kJ A/mol/e2) */

/* ---- building blocks ---- */
/* RotMatrix, MC_lib/lib.f90:27-83 */
void mco_rotmatrix(const double com[3], const double hoek[3], const double *rel, int natom, double *abs_out);
/* LJ_FORMULA, LennardJones.f90:153-187; out[0] = q1 q2 / r, out[1] = 4 e (s^12/r^12 - s^6/r^6) */
void mco_lj_formula(double e1, double e2, double s1, double s2, double q1, double q2,
                    const double rin[3], double boxl, double boxl2, double out[2]);
/* calcLJ, LennardJones.f90:24-84 (H skipped on both sides) */
double mco_calc_lj(const double *mol1, const double *mol2, const int *is_h, const double *table,
                   int natom, double boxl, double boxl2, double conv);
/* CalcLJ_SV, LennardJones.f90:89-148 (H skipped on the solvent side only) */
double mco_calc_lj_sv(const double *mol1, const double *solute, const int *is_h, const double *table_dmso,
                      const double *table_sol, int natom, int nsol, double boxl, double boxl2, double conv);

/* ---- a chain = one copy of the reference program's state ---- */
mco_chain *mco_create(int n, int ndmso, int nsol, double boxl,
                      const double *dmso_rel, const int *dmso_is_h, const double *dmso_tab,
                      const double *solute, const double *sol_tab,
                      double conv, double cutoff, int keep_matrix);
void mco_destroy(mco_chain *c);
void mco_set_state(mco_chain *c, const double *com, const double *hoek);
void mco_get_state(const mco_chain *c, double *com, double *hoek);
void mco_set_mc_params(mco_chain *c, double temperature, double dposmax, double dposmin,
                       double dhoekmax_rad, double dhoekmin_rad, double padj, int nadj,
                       int64_t start_prod, int quirk31);

/* calculateLJ(I), MonteCarlo.F90:792-828: refresh row/column i of SS and ENERGY(i) */
void mco_calculate_lj(mco_chain *c, int i);
/* calcEnergy, MonteCarlo.F90:988-1004 */
double mco_calc_energy(const mco_chain *c);
/* start-up recompute, MonteCarlo.F90:357-364: calculateLJ for every I, then calcEnergy; sets TOTENG */
double mco_total_energy(mco_chain *c);
/* matrix-free variants (same pair orientation and order of evaluation, no N x N storage):
 * brute force over all pairs, and a cell list that is used ONLY to find CoM pairs within the cutoff */
double mco_total_energy_matrixfree(const mco_chain *c, double *e_solute_sum, uint64_t *n_pairs_in_cutoff);
double mco_total_energy_cells(const mco_chain *c, double *e_solute_sum, uint64_t *n_pairs_in_cutoff);
/* row i without touching state: ss_row[n] (may be NULL) and ENERGY(i) */
double mco_row_energy(const mco_chain *c, int i, double *ss_row, double *e_i);
void mco_get_matrix(const mco_chain *c, double *ss, double *energy);

/* MCINIT (MonteCarlo.F90:685-715) with the 7 uniforms it would draw; returns the chosen molecule */
int mco_mcinit(mco_chain *c, const double u[7]);
/* one whole step of loop_LJ (MonteCarlo.F90:406-433): MCINIT, calculateLJ, calcEnergy, clamp,
 * METROPOLIS (:720-788).  Consumes 7 or 8 uniforms from u; returns how many.  trace (may be NULL)
 * receives {toteng_trial, toteng_old, kans, rv, rsolv(0-based), accepted, ratio, dposmax, dhoekmax, n_sitepairs}. */
int mco_step(mco_chain *c, const double *u, double trace[10]);
/* trial move supplied by the caller instead of drawn (replay of a reference DUMP frame) */
int mco_step_given_move(mco_chain *c, int i, const double new_com[3], const double new_hoek[3],
                        const double *u_accept, double trace[10]);
/* nsteps steps fed from a uniform buffer; stops early (returns steps done) if the buffer runs out */
int64_t mco_run_steps(mco_chain *c, int64_t nsteps, const double *uniforms, int64_t cap, int64_t *consumed,
                      double *trace /* [nsteps][10] or NULL */);
/* nsteps steps fed from the built-in counter stream: step s (0-based, chain life) owns u(8s)..u(8s+7) */
int64_t mco_run_steps_rng(mco_chain *c, int64_t nsteps, uint64_t rng[2], double *trace);

/* ---- replica exchange: no counterpart in the reference; its METROPOLIS rule (:728-744) on (beta_a - beta_b)(E_a - E_b) ---- */
int mco_exchange(mco_chain *a, mco_chain *b, double u, int *used);
void mco_rng_exchange_seed(uint64_t state[2], uint64_t seed);
double mco_get_temperature(const mco_chain *c);

/* counters: {step, nsuc, naccept_window, prod_acc, n_clamped, sitepairs(lo), toteng, toteng_old, production,
 *            dposmax, dhoekmax} */
void mco_get_stats(const mco_chain *c, double out[12]);

/* ---- the documented uniform stream used when the driver does not supply uniforms ----
 * G = 0x9E3779B97F4A7C15, mix = splitmix64's finaliser, key = mix(seed + G),
 * u_n = (mix(key + (n+1)*G) >> 11) * 2^-53 in [0,1); state = {key, n}.  Not from the reference: its RNG is the
 * compiler's RANDOM_NUMBER seeded from /dev/urandom (MC_lib/randgen.f90:15-40), so no stream exists to match. */
void mco_rng_seed(uint64_t state[2], uint64_t seed);
double mco_rng_uniform(uint64_t state[2]);

#ifdef __cplusplus
}
#endif
#endif

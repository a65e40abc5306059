/* mcb200.h — C ABI of libmcb200.so: the classical LJ(+Coulomb) trial-move energy path of
 * MatthijsLasure/MonteCarlo, rebuilt as fp64 CUDA for B200 (sm_100a).
 *
 * The reference has no FFI: its hot path is two internal procedures of the main program
 * (calculateLJ, calcEnergy) plus module LennardJones, sharing state by host association.  This header is
 * the boundary a Fortran driver binds with ISO_C_BINDING instead; every entry point
 * names the reference code it stands in for (paths relative to the reference root).
 *
 * Conventions
 *   - all functions return 0 on success, > 0 for an argument error, < 0 for a CUDA error;
 *     mcb_last_error() gives the text.  The library never calls exit() and never clamps energies in the
 *     single-step entry points (Inf/NaN are returned as values, as the reference's functions do).
 *   - arrays of the reference's TYPE(vector) (MC_lib/vector_class.f90:18-20) are double[n][3], i.e. what
 *     c_loc(arr) points at; molecule indices are 1-based like RSOLV.
 *   - BOXL is HALF the box side (MonteCarlo.F90:296 BOXL2 = 2*BOXL).
 *   - a handle is bound to one GPU and is not thread-safe; host buffers are the caller's.
 *   - there is no CPU fallback: without a CUDA device mcb_create fails.
 */
#ifndef MCB200_H
#define MCB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCB_MAX_SITES 4 /* interacting sites per solvent molecule (DMSO: S, C, C, O; H is skipped) */
#define MCB_NOALLOC_MAX_MOL 32768 /* largest box mcb_box_energy_noalloc accepts (size of its fixed scratch array) */
#define MCB_NOALLOC_MAX_SOL 24    /* most solute atoms mcb_box_energy_noalloc accepts (they travel as a kernel argument) */

typedef struct mcb_handle mcb_t;

/* What calculateLJ reads besides the per-chain state (MonteCarlo.F90:306-310, :792-828). */
typedef struct {
    int n_mol;              /* NCOM */
    int n_sites;            /* non-H atoms of the solvent molecule, 1..MCB_MAX_SITES (LennardJones.f90:66,68) */
    int n_sol;              /* NSOL, may be 0 */
    int n_chains;           /* independent copies of the box held on this GPU */
    int device;             /* CUDA ordinal; -1 = the current device */
    int quirk_31;           /* 1 = keep `IF (RSOLV .EQ. 31) RSOLV = 30` (MonteCarlo.F90:698) in mcb_run_steps */
    int matrix_mode;        /* 0 auto, 1 force the pair matrix into global memory (never shared memory) */
    int lookahead;          /* steps of a chain evaluated speculatively at once by mcb_run_steps' built-in-stream
                               path: 1, 2, 4 or 8; 0 = default (1 = none).  Does not change any decision of the chain. */
    double boxl;            /* half side, Angstrom */
    double cutoff;          /* centre-of-mass cutoff, 7.0 in the reference (MonteCarlo.F90:812) */
    double conv;            /* factor on the Coulomb sum; the reference's value is exactly 0.0 (LennardJones.f90:57) */
    const double *site_rel; /* [n_sites][3] molecule-frame coordinates (rows of DMSO.txt with symbol != H) */
    const double *site_tab; /* [n_sites][3] = (q, epsilon, sigma), ASSOSIATE_DMSO (LennardJones.f90:192-219) */
} mcb_params;

/* Per-chain Metropolis parameters: the driver's TEMPERATURE, DPOSMAX, ... (MonteCarlo.F90:33-47, :219-222). */
typedef struct {
    double temperature;
    double dposmax, dposmin;
    double dhoekmax, dhoekmin;  /* radians (the driver multiplies the config values by PI, :219-220) */
    double padj;
    int nadj;                   /* LJ_NADJ */
    int reserved;
    int64_t start_prod;         /* first step that counts for the production average (:222, :749) */
} mcb_mc_params;

/* Counters of one chain: NSUC, NACCEPT, PRODUCTION, PROD_ACC, TOTENG, ... (MonteCarlo.F90:36-44, :98-101). */
typedef struct {
    double toteng;          /* energy of the current (accepted) configuration */
    double dposmax, dhoekmax;
    double production;      /* sum of TOTENG over accepted steps >= start_prod (:749-752) */
    double sum_e, sum_e2;   /* sum of the current energy and its square over every step >= start_prod */
    double ratio;           /* REAL(NACCEPT)/REAL(NADJ) as of the last step (:758) */
    double last_delta;      /* TOTENG - TOTENG_OLD of the last trial */
    int64_t step;           /* steps done (UNICORN) */
    int64_t nsuc;           /* accepted steps */
    int64_t naccept;        /* accepted steps in the current adaptation window */
    int64_t prod_acc;
    int64_t n_clamped;      /* trial energies that were Inf/NaN and were clamped (:421-428) */
    int64_t sitepairs;      /* LJ_FORMULA evaluations of trial states (the unit of the throughput metric) */
    int64_t uniforms_used;
    int64_t n_stat;         /* number of terms in sum_e / sum_e2 */
    int64_t exch_attempts;  /* replica-exchange attempts this chain took part in (mcb_exchange) */
    int64_t exch_accepted;  /* ... and how many of them swapped its temperature */
    double temperature;     /* the chain's current temperature (mcb_exchange moves it between chains) */
} mcb_chain_stats;

/* ---- life cycle --------------------------------------------------------------------------------- */
/* replaces: the ALLOCATEs and table set-up at MonteCarlo.F90:306-322 */
int mcb_create(mcb_t **out, const mcb_params *p);
int mcb_destroy(mcb_t *h);
const char *mcb_last_error(const mcb_t *h); /* h may be NULL: error of the last failed mcb_create */
/* launch everything on this cudaStream_t (default: the legacy default stream) */
int mcb_set_stream(mcb_t *h, void *cuda_stream);
/* kernels launched by this handle so far */
int64_t mcb_launch_count(const mcb_t *h);

/* ---- state --------------------------------------------------------------------------------------- */
/* replaces: SOLUTE and TABLE_SOL as read/associated at MonteCarlo.F90:1103-1114, LennardJones.f90:224-248
 * (+ CHELPG charges, MonteCarlo.F90:353).  solute [c][n_sol][3], sol_tab [c][n_sol][3] = (q, eps, sigma);
 * with broadcast != 0 one copy ([n_sol][3]) is given to all nchains chains. */
int mcb_set_solute(mcb_t *h, int chain0, int nchains, const double *solute, const double *sol_tab, int broadcast);
/* replaces: CoM / HOEK as read by read_norm (MonteCarlo.F90:1097-1099).  com, hoek [c][n_mol][3].
 * Invalidates the cached energies of those chains until mcb_total_energy is called. */
int mcb_set_chain_state(mcb_t *h, int chain0, int nchains, const double *com, const double *hoek, int broadcast);
/* for the driver's dump / backup / result writers (MonteCarlo.F90:1009-1046, :587-611) */
int mcb_get_chain_state(mcb_t *h, int chain0, int nchains, double *com, double *hoek);

/* ---- energies ------------------------------------------------------------------------------------ */
/* replaces: the start-up recompute `DO I=1,NCOM: CALL calculateLJ(I)` + calcEnergy (MonteCarlo.F90:357-364).
 * Rebuilds the rotated-site cache, the pair matrix and ENERGY(:) of each chain.  Outputs (any may be NULL):
 * toteng[c], e_solute_sum[c] = sum_i ENERGY(i), pairs_in_cutoff[c] = unordered CoM pairs inside the cutoff. */
int mcb_total_energy(mcb_t *h, int chain0, int nchains, double *toteng, double *e_solute_sum,
                     int64_t *pairs_in_cutoff);
/* replaces: the `solventsolvent` dump (MonteCarlo.F90:372-376): ss [n_mol][n_mol], energy [n_mol] */
int mcb_get_matrix(mcb_t *h, int chain, double *ss, double *energy);
/* replaces: CALL calculateLJ(I) read-only (MonteCarlo.F90:792-828): row I of the pair matrix recomputed from
 * the current coordinates with I as MOL1, and ENERGY(I).  i is 1-based.  Argument error while a trial from
 * mcb_trial_delta waits for its mcb_commit (both use the chain's work area). */
int mcb_row_energy(mcb_t *h, int chain, int i, double *ss_row, double *e_i);

/* ---- single-step mode: the driver keeps MCINIT / METROPOLIS (RNG, accept/reject) ------------------- */
/* replaces: CALL calculateLJ(RSOLV) + calcEnergy for the trial state (MonteCarlo.F90:411-414), for the
 * first nchains chains at once.  rsolv[c] 1-based; new_com, new_hoek [c][3] = the molecule after MCINIT's
 * move and wrap (:699-712).  d_e[c] = TOTENG_trial - TOTENG_current, e_new[c] = TOTENG_trial. */
int mcb_trial_delta(mcb_t *h, int nchains, const int *rsolv, const double *new_com, const double *new_hoek,
                    double *d_e, double *e_new);
/* replaces: the keep-or-roll-back at MonteCarlo.F90:779-786: accept[c] != 0 makes the pending trial of
 * chain c current; 0 discards it.  Argument error when no trial is pending: mcb_set_chain_state and mcb_set_solute
 * discard a parked trial (it was evaluated against a state that no longer exists). */
int mcb_commit(mcb_t *h, int nchains, const unsigned char *accept);

/* ---- batched mode: MCINIT + calculateLJ + calcEnergy + METROPOLIS looped on the GPU ----------------- */
int mcb_set_mc_params(mcb_t *h, int chain0, int nchains, const mcb_mc_params *p, int broadcast);
/* Uniform stream used when mcb_run_steps gets no buffer: counter-based and random-access,
 *   G = 0x9E3779B97F4A7C15, mix = splitmix64's finaliser, key = mix(seed + G),
 *   u(n) = (mix(key + (n+1)*G) >> 11) * 2^-53 in [0,1);
 * Metropolis step s reads u(8s) .. u(8s+6) for MCINIT and u(8s+7) if METROPOLIS draws, where s counts every step
 * mcb_run_steps has made for the chain since it was seeded (0-based).  mcb_reset_stats does NOT rewind it: an
 * equilibrate / reset / production sequence goes on in the stream, it never replays the equilibration's numbers.
 * Only mcb_seed_chains puts s back to 0.  (The reference's own RNG is the compiler's RANDOM_NUMBER seeded from
 * /dev/urandom, MC_lib/randgen.f90:15-40: there is no reference stream to reproduce; a driver that wants its
 * own stream passes a buffer.) */
int mcb_seed_chains(mcb_t *h, int chain0, int nchains, const uint64_t *seeds);
/* replaces: loop_LJ (MonteCarlo.F90:406-433) for nsteps steps of every chain.  uniforms: NULL (built-in
 * stream) or host [n_chains][cap] doubles in [0,1): each step consumes 7 (MCINIT :697-700, lib.f90:115-130)
 * plus 1 only when METROPOLIS draws (:739).  A chain stops early when fewer than 8 remain; consumed[c]
 * (may be NULL) receives how many it used, steps_done[c] (may be NULL) how many steps it made.  With
 * uniforms, consumed and steps_done all NULL the call only enqueues the kernel on the handle's stream and
 * returns; any later call that reads results (mcb_get_stats, mcb_get_chain_state, ...) waits for it.
 * DELTA: with a uniform buffer METROPOLIS tests DELTA = (TOTENG + dE) - TOTENG, the reference's subtraction of two
 * totals (MonteCarlo.F90:415-428, :726); with the built-in stream it tests dE as summed.  The two differ by the
 * rounding of TOTENG + dE (at most an ulp of TOTENG), which can change a decision only where the exponent is within
 * that distance of its threshold - the margin BASELINE's parity rule leaves open.  Both clamp +Inf and NaN and let
 * -Inf pass, as the reference does.
 * Box sizes: up to ~1400 molecules a chain's coordinates and site cache are staged in the SM's shared memory and both
 * kinds of stream work; a larger box (up to 65535 molecules, as long as the pair matrices fit in the GPU's memory) steps
 * with the built-in stream only - a uniform buffer is then an argument error - and beyond that the handle is
 * energy-only (mcb_total_energy). */
int mcb_run_steps(mcb_t *h, int64_t nsteps, const double *uniforms, int64_t cap, int64_t *consumed,
                  int64_t *steps_done);
int mcb_get_stats(mcb_t *h, int chain0, int nchains, mcb_chain_stats *out);
/* zero the counters (step, nsuc, production, ...) but keep coordinates, energies, step sizes and the position in
 * the built-in uniform stream */
int mcb_reset_stats(mcb_t *h, int chain0, int nchains);

/* ---- temperature ladder: replica exchange between launches ------------------------------------------ */
/* The reference runs one temperature per process and has no exchange move; its acceptance rule is reused here
 * (METROPOLIS, MonteCarlo.F90:728-744) for the standard parallel-tempering swap of two chains a, b:
 *     EXPONENT = (1000/(8.314 T_a) - 1000/(8.314 T_b)) * (E_a - E_b)
 *     EXPONENT < -75: rejected, no draw;  EXPONENT >= 0: accepted, no draw;  else accepted iff RV <= E**EXPONENT
 * with E the reference's default-REAL 2.7182817 (lib.f90:17) and E_x the chains' current TOTENG.  An accepted swap
 * exchanges the two chains' TEMPERATURES together with their step sizes DPOSMAX / DHOEKMAX (tuned per temperature);
 * coordinates, energies and counters stay where they are.  pair_a[p], pair_b[p] are 0-based chain indices of this
 * handle; a chain may appear in at most one pair per call.  uniforms[p] in [0,1) is pair p's RV (read only when the
 * rule draws); with uniforms == NULL pair p draws from chain pair_a[p]'s exchange stream
 *     key_x = mix(key + 0xD1B54A32D192ED03),  u_x(n) = (mix(key_x + (n+1)*G) >> 11) * 2^-53,
 * n = draws that chain has made from it so far (key, mix, G as for mcb_seed_chains, which also rewinds n).
 * accepted[p] (may be NULL) receives 1 or 0.  Both chains count the attempt in mcb_chain_stats.exch_attempts. */
int mcb_exchange(mcb_t *h, int npairs, const int *pair_a, const int *pair_b, const double *uniforms,
                 unsigned char *accepted);

/* ---- end of run: one collective over the GPUs of the job ---------------------------------------------- */
/* replaces: the way the reference combines its independent runs - separate processes whose logs PostProcess scrapes
 * afterwards (PostProcess/PostProcess.F90:202-256 RECON / READLOG).  The chains of a job are spread over the GPUs in
 * contiguous blocks - rank r of nranks holds n_total/nranks chains, the first n_total % nranks ranks one more - one
 * handle (and one host process or thread) per GPU, nothing is exchanged while sampling, and the run ends with ONE
 * NCCL all-reduce (mcb_reduce_stats) and ONE all-gather (mcb_gather_chains) over NVLink.  NCCL is bound at run time
 * (libnccl.so.2); without it these calls fail with a message and everything else still works. */
#define MCB_COMM_ID_BYTES 128
/* rank 0 calls this and hands the 128 bytes to the other ranks by its own means (MPI_Bcast, a file, a socket) */
int mcb_comm_get_unique_id(char id[MCB_COMM_ID_BYTES]);
/* collective over all ranks: joins the communicator on the handle's GPU; h must hold exactly rank's block */
int mcb_comm_init(mcb_t *h, int nranks, int rank, const char id[MCB_COMM_ID_BYTES], int64_t n_total_chains);
/* the same with a communicator the host already has (an ncclComm_t, e.g. from its MPI+NCCL set-up); not destroyed here */
int mcb_comm_attach(mcb_t *h, void *nccl_comm, int nranks, int rank, int64_t n_total_chains);
int mcb_comm_destroy(mcb_t *h);

/* Sums over all chains of a group, over all ranks: what the production average (MonteCarlo.F90:555-560) and the
 * acceptance printed at the end of a run need, per run in the reference, per group of chains here. */
typedef struct {
    double sum_e, sum_e2;   /* sums of mcb_chain_stats.sum_e / sum_e2 */
    double production;      /* sum of PRODUCTION */
    double sum_toteng;      /* sum of the final TOTENG */
    int64_t n_stat, prod_acc, nsuc, step, sitepairs, n_clamped, uniforms_used;
    int64_t n_chains;       /* chains in the group */
} mcb_global_stats;
/* Collective.  Group g = global chains [g*chains_per_group, (g+1)*chains_per_group) (BASELINE config 4: one group per
 * dihedral angle); chains_per_group <= 0: one group with everything.  n_groups = ceil(n_total / chains_per_group) is
 * checked.  Every rank receives every group.  Without mcb_comm_init the handle's own chains are summed (one rank). */
int mcb_reduce_stats(mcb_t *h, int64_t chains_per_group, mcb_global_stats *out, int n_groups);
/* Collective.  out[n_total_chains][4] in global chain order: final TOTENG, mean energy sum_e/n_stat, acceptance
 * nsuc/step, production mean PRODUCTION/PROD_ACC (0 where the divisor is 0) - what PostProcess reads from each run's
 * log (' Energy:' lines, PostProcess.F90:229-256). */
int mcb_gather_chains(mcb_t *h, double *out, int64_t n_total_chains);

/* ---- for hosts that must not allocate ---------------------------------------------------------------- */
/* No counterpart in the reference (which never touches a GPU).  Same result as mcb_total_energy, for callers
 * that cannot let a library allocate device memory: a driver whose own pool owns the whole GPU, a process under
 * a fixed memory cap, a timing-critical loop that cannot take the device-wide stall of cudaMalloc.  It needs no
 * handle and calls no cudaMalloc / cudaFree or any allocating routine; its only scratch (per-block partial sums)
 * is a ~0.5 MB array built into the library, which every GPU that loads the library holds until exit.
 *   p           n_mol <= MCB_NOALLOC_MAX_MOL, n_sol <= MCB_NOALLOC_MAX_SOL, else rc 2 and a message; n_chains,
 *               quirk_31, matrix_mode, lookahead are ignored; device -1 = current device
 *   d_com, d_hoek  DEVICE pointers, double[n_mol][3] (the driver's TYPE(vector) arrays, already on the GPU)
 *   solute, sol_tab  HOST pointers, double[n_sol][3] (coordinates; q, eps, sigma), may be NULL when n_sol == 0
 *   d_out       DEVICE pointer, double[3]: TOTENG, sum_i ENERGY(i), unordered CoM pairs inside the cutoff
 *   cuda_stream the kernels are only enqueued there; the result is ready when the stream reaches that point
 *   errbuf      optional buffer for the error text (there is no handle to ask)
 * The first call on each device also looks up the scratch array's device address; make that call before any
 * stream capture begins.  All calls share the scratch array: calls whose kernels could overlap on the GPU
 * (different streams or host threads) race and must be serialised by the caller. */
int mcb_box_energy_noalloc(const mcb_params *p, const double *d_com, const double *d_hoek, const double *solute,
                           const double *sol_tab, double *d_out, void *cuda_stream, char *errbuf, size_t errbuf_len);

/* ---- measurement ---------------------------------------------------------------------------------- */
/* Independent DFMA chains on every SM for about `ms` milliseconds: the fp64 peak that the roofline
 * fractions in bench.py are quoted against (MEASURED_PEAKS.json has no fp64 entry). */
int mcb_measure_fp64_peak(int device, double ms, double *tflops);

#ifdef __cplusplus
}
#endif
#endif

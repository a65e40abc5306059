/* mc_oracle.c — CPU restatement of the reference's classical LJ(+Coulomb) Monte-Carlo path.
 *
 * TEST INFRASTRUCTURE ONLY — see mc_oracle.h.  Parity status: PINNED against fixtures produced by the
 * reference's own shipped binary (tests/golden/, oracle/make_golden.py).
 *
 * Built twice from this one source (oracle/Makefile):
 *   libmc_oracle.so    REAL = double       — the reference's arithmetic, operation for operation
 *   libmc_oracle_q.so  REAL = __float128   — arbiter: when GPU and fp64 oracle differ in the last bits,
 *                                            the quad result says which one is closer
 * Compiled with -ffp-contract=off -fno-fast-math so the fp64 build rounds after every operation, as the
 * reference's x86-64 SSE2 code does.
 */
#include "mc_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#ifdef MCO_QUAD
#include <quadmath.h>
typedef __float128 REAL;
#define R_SQRT sqrtq
#define R_SIN sinq
#define R_COS cosq
#define R_TRUNC truncq
#define R_POW powq
#define R_ATAN atanq
#else
typedef double REAL;
#define R_SQRT sqrt
#define R_SIN sin
#define R_COS cos
#define R_TRUNC trunc
#define R_POW pow
#define R_ATAN atan
#endif

/* ------------------------------------------------------------------ constants */
/* MC_lib/lib.f90:15-16: PI = 4.D0*DATAN(1.D0), TAU = 2.D0*PI (fp64 in both builds: they are stored as
 * DOUBLE PRECISION parameters in the reference) */
static double c_pi(void) { return 4.0 * atan(1.0); }
double mco_const_pi(void) { return c_pi(); }
double mco_const_tau(void) { return 2.0 * c_pi(); }
/* lib.f90:17: the literal has no kind suffix, so it is a default REAL (float32) widened to fp64 */
double mco_const_e(void) { return (double)2.71828182845904523536028747135266249775724709369995f; }
/* lib.f90:19: product of three default-REAL literals, evaluated in float32 */
double mco_const_hartree2kjmol(void)
{
    volatile float a = 4.359744f, b = 6.02214085f, c = 100.0f;
    volatile float ab = a * b;
    volatile float abc = ab * c;
    return (double)abc;
}
/* LennardJones.f90:53-57.  10**(-8) is INTEGER**INTEGER = 0, so CONV ends as 0.0 */
double mco_const_conv_reference(void)
{
    volatile float a = 6.023f, b = 1.60217646f;
    volatile float b2 = b * b;
    volatile float ab = a * b2; /* REAL*REAL, float32 */
    double conv = (double)ab;
    conv = conv / 4;
    conv = conv / c_pi();
    conv = conv / (double)8.854187817620f;
    conv = conv * 0; /* 10**(-8) */
    return conv;
}
double mco_const_conv_intended(void)
{
    volatile float a = 6.023f, b = 1.60217646f;
    volatile float b2 = b * b;
    volatile float ab = a * b2;
    double conv = (double)ab;
    conv = conv / 4;
    conv = conv / c_pi();
    conv = conv / (double)8.854187817620f;
    return conv * 1.0e4;
}

/* ------------------------------------------------------------------ RotMatrix */
/* MC_lib/lib.f90:27-83.  HOEK = (phi, psi, theta) = (x, y, z); entries exactly as written at :62-72;
 * ABSPOS(i) = COM + matmul(RM, RELPOS(i)) (:76-81). */
static void rotmatrix_r(const REAL com[3], const REAL hoek[3], const double *rel, int natom, REAL *out)
{
    REAL c1 = R_COS(hoek[0]), c2 = R_COS(hoek[1]), c3 = R_COS(hoek[2]);
    REAL s1 = R_SIN(hoek[0]), s2 = R_SIN(hoek[1]), s3 = R_SIN(hoek[2]);
    REAL rm[3][3];
    rm[0][0] = c1 * c2 - s1 * s2 * c3;
    rm[0][1] = s1 * c2 + c1 * s2 * c3;
    rm[0][2] = s2 * s3;
    rm[1][0] = (REAL)-1.0 * c1 * s2 - s1 * c2 * c3;
    rm[1][1] = (REAL)-1.0 * s1 * s2 + c1 * c2 * c3;
    rm[1][2] = c2 * s3;
    rm[2][0] = s1 * s3;
    rm[2][1] = (REAL)-1.0 * c1 * s3;
    rm[2][2] = c3;
    for (int i = 0; i < natom; ++i) {
        REAL t0 = rel[3 * i + 0], t1 = rel[3 * i + 1], t2 = rel[3 * i + 2];
        for (int k = 0; k < 3; ++k) {
            REAL v = rm[k][0] * t0 + rm[k][1] * t1 + rm[k][2] * t2;
            out[3 * i + k] = com[k] + v;
        }
    }
}

void mco_rotmatrix(const double com[3], const double hoek[3], const double *rel, int natom, double *abs_out)
{
    REAL c[3] = {com[0], com[1], com[2]}, h[3] = {hoek[0], hoek[1], hoek[2]};
    REAL *tmp = (REAL *)malloc(sizeof(REAL) * 3 * (size_t)natom);
    rotmatrix_r(c, h, rel, natom, tmp);
    for (int i = 0; i < 3 * natom; ++i) abs_out[i] = (double)tmp[i];
    free(tmp);
}

/* ------------------------------------------------------------------ LJ_FORMULA */
/* LennardJones.f90:153-187.  Integer powers by repeated squaring (x2, x4, x6 = x4*x2, x12 = x6*x6). */
static inline REAL pow6_r(REAL x) { REAL x2 = x * x; REAL x4 = x2 * x2; return x4 * x2; }

static inline void lj_formula_r(REAL e1, REAL e2, REAL s1, REAL s2, REAL q1, REAL q2,
                                REAL rx, REAL ry, REAL rz, REAL boxl, REAL boxl2, REAL out[2])
{
    REAL e = R_SQRT(e1 * e2);        /* :173 */
    REAL s = (s1 + s2) / 2;          /* :174 */
    rx = rx - boxl2 * R_TRUNC(rx / boxl); /* :178-180, AINT on the raw difference, BOXL = half side */
    ry = ry - boxl2 * R_TRUNC(ry / boxl);
    rz = rz - boxl2 * R_TRUNC(rz / boxl);
    REAL rv = R_SQRT(rx * rx + ry * ry + rz * rz); /* length(), vector_class.f90:104-114 */
    REAL s6 = pow6_r(s), r6 = pow6_r(rv);
    REAL s12 = s6 * s6, r12 = r6 * r6;
    out[0] = q1 * q2 / rv;                         /* :184 */
    out[1] = 4 * e * (s12 / r12 - s6 / r6);        /* :185 */
}

void mco_lj_formula(double e1, double e2, double s1, double s2, double q1, double q2,
                    const double rin[3], double boxl, double boxl2, double out[2])
{
    REAL o[2];
    lj_formula_r(e1, e2, s1, s2, q1, q2, rin[0], rin[1], rin[2], boxl, boxl2, o);
    out[0] = (double)o[0];
    out[1] = (double)o[1];
}

/* calcLJ, LennardJones.f90:24-84: I outer over MOL1, J inner over MOL2, both filtered with sym1 (:66,:68);
 * R = mol2(J) - mol1(I) (:69); EN = SUML*CONV + SUMR (:80-82) */
static REAL calc_lj_r(const REAL *mol1, const REAL *mol2, const int *is_h, const double *table, int natom,
                      REAL boxl, REAL boxl2, REAL conv, uint64_t *npairs)
{
    REAL suml = 0, sumr = 0, t[2];
    for (int i = 0; i < natom; ++i) {
        if (is_h[i]) continue;
        for (int j = 0; j < natom; ++j) {
            if (is_h[j]) continue;
            lj_formula_r(table[3 * i + 1], table[3 * j + 1], table[3 * i + 2], table[3 * j + 2],
                         table[3 * i + 0], table[3 * j + 0],
                         mol2[3 * j + 0] - mol1[3 * i + 0], mol2[3 * j + 1] - mol1[3 * i + 1],
                         mol2[3 * j + 2] - mol1[3 * i + 2], boxl, boxl2, t);
            suml = suml + t[0];
            sumr = sumr + t[1];
            if (npairs) ++*npairs;
        }
    }
    suml = suml * conv;
    return suml + sumr;
}

/* CalcLJ_SV, LennardJones.f90:89-148: solvent H skipped (:132), every solute atom counted (:133) */
static REAL calc_lj_sv_r(const REAL *mol1, const double *solute, const int *is_h, const double *tdmso,
                         const double *tsol, int natom, int nsol, REAL boxl, REAL boxl2, REAL conv,
                         uint64_t *npairs)
{
    REAL suml = 0, sumr = 0, t[2];
    for (int i = 0; i < natom; ++i) {
        if (is_h[i]) continue;
        for (int j = 0; j < nsol; ++j) {
            lj_formula_r(tdmso[3 * i + 1], tsol[3 * j + 1], tdmso[3 * i + 2], tsol[3 * j + 2],
                         tdmso[3 * i + 0], tsol[3 * j + 0],
                         (REAL)solute[3 * j + 0] - mol1[3 * i + 0], (REAL)solute[3 * j + 1] - mol1[3 * i + 1],
                         (REAL)solute[3 * j + 2] - mol1[3 * i + 2], boxl, boxl2, t);
            suml = suml + t[0];
            sumr = sumr + t[1];
            if (npairs) ++*npairs;
        }
    }
    suml = suml * conv;
    return suml + sumr;
}

double mco_calc_lj(const double *mol1, const double *mol2, const int *is_h, const double *table,
                   int natom, double boxl, double boxl2, double conv)
{
    REAL *a = (REAL *)malloc(sizeof(REAL) * 6 * (size_t)natom), *b = a + 3 * natom;
    for (int i = 0; i < 3 * natom; ++i) { a[i] = mol1[i]; b[i] = mol2[i]; }
    double r = (double)calc_lj_r(a, b, is_h, table, natom, boxl, boxl2, conv, NULL);
    free(a);
    return r;
}

double mco_calc_lj_sv(const double *mol1, const double *solute, const int *is_h, const double *table_dmso,
                      const double *table_sol, int natom, int nsol, double boxl, double boxl2, double conv)
{
    REAL *a = (REAL *)malloc(sizeof(REAL) * 3 * (size_t)natom);
    for (int i = 0; i < 3 * natom; ++i) a[i] = mol1[i];
    double r = (double)calc_lj_sv_r(a, solute, is_h, table_dmso, table_sol, natom, nsol, boxl, boxl2, conv, NULL);
    free(a);
    return r;
}

/* ------------------------------------------------------------------ chain state */
struct mco_chain {
    int n, ndmso, nsol, keep_matrix;
    REAL boxl, boxl2, conv, cutoff;
    double *dmso_rel; int *is_h; double *dmso_tab; double *solute; double *sol_tab;
    REAL *com, *hoek, *com_old, *hoek_old;   /* [n][3] */
    REAL *ss, *ssold;                        /* [n][n], only if keep_matrix */
    REAL *energy, *eold;                     /* [n] */
    REAL *mol1, *mol2;                       /* scratch [ndmso][3] */
    REAL toteng, toteng_old;
    /* MC parameters and counters (MonteCarlo.F90:33-47, :100-101) */
    REAL temperature, dposmax, dposmin, dhoekmax, dhoekmin, padj;
    int nadj, quirk31, rsolv;
    int64_t start_prod, step, nsuc, prod_acc, nclamp;
    int naccept;
    REAL production;
    uint64_t sitepairs, last_sitepairs;
    double ratio;
};

static void *xdup(const void *src, size_t bytes)
{
    void *p = malloc(bytes ? bytes : 1);
    if (src) memcpy(p, src, bytes);
    return p;
}

mco_chain *mco_create(int n, int ndmso, int nsol, double boxl,
                      const double *dmso_rel, const int *dmso_is_h, const double *dmso_tab,
                      const double *solute, const double *sol_tab,
                      double conv, double cutoff, int keep_matrix)
{
    mco_chain *c = (mco_chain *)calloc(1, sizeof(*c));
    c->n = n; c->ndmso = ndmso; c->nsol = nsol; c->keep_matrix = keep_matrix;
    c->boxl = boxl; c->boxl2 = (REAL)boxl * (REAL)2.0; /* MonteCarlo.F90:296 */
    c->conv = conv; c->cutoff = cutoff;
    c->dmso_rel = (double *)xdup(dmso_rel, sizeof(double) * 3 * ndmso);
    c->is_h = (int *)xdup(dmso_is_h, sizeof(int) * ndmso);
    c->dmso_tab = (double *)xdup(dmso_tab, sizeof(double) * 3 * ndmso);
    c->solute = (double *)xdup(solute, sizeof(double) * 3 * nsol);
    c->sol_tab = (double *)xdup(sol_tab, sizeof(double) * 3 * nsol);
    size_t v = sizeof(REAL) * 3 * (size_t)n;
    c->com = (REAL *)calloc(1, v); c->hoek = (REAL *)calloc(1, v);
    c->com_old = (REAL *)calloc(1, v); c->hoek_old = (REAL *)calloc(1, v);
    if (keep_matrix) {
        c->ss = (REAL *)calloc((size_t)n * n, sizeof(REAL));
        c->ssold = (REAL *)calloc((size_t)n * n, sizeof(REAL));
    }
    c->energy = (REAL *)calloc(n, sizeof(REAL)); c->eold = (REAL *)calloc(n, sizeof(REAL));
    c->mol1 = (REAL *)calloc(3 * ndmso, sizeof(REAL)); c->mol2 = (REAL *)calloc(3 * ndmso, sizeof(REAL));
    /* defaults = Tools/files/config0.ini:4-15 */
    c->temperature = 300.0; c->dposmax = 0.30; c->dposmin = 1e-10;
    c->dhoekmax = 0.16 * c_pi(); c->dhoekmin = 1e-10 * c_pi(); c->padj = 0.05; c->nadj = 100;
    c->start_prod = 0; c->quirk31 = 1;
    return c;
}

void mco_destroy(mco_chain *c)
{
    if (!c) return;
    free(c->dmso_rel); free(c->is_h); free(c->dmso_tab); free(c->solute); free(c->sol_tab);
    free(c->com); free(c->hoek); free(c->com_old); free(c->hoek_old);
    free(c->ss); free(c->ssold); free(c->energy); free(c->eold); free(c->mol1); free(c->mol2);
    free(c);
}

void mco_set_state(mco_chain *c, const double *com, const double *hoek)
{
    for (int i = 0; i < 3 * c->n; ++i) { c->com[i] = com[i]; c->hoek[i] = hoek[i]; }
}

void mco_get_state(const mco_chain *c, double *com, double *hoek)
{
    for (int i = 0; i < 3 * c->n; ++i) { com[i] = (double)c->com[i]; hoek[i] = (double)c->hoek[i]; }
}

void mco_set_mc_params(mco_chain *c, double temperature, double dposmax, double dposmin,
                       double dhoekmax_rad, double dhoekmin_rad, double padj, int nadj,
                       int64_t start_prod, int quirk31)
{
    c->temperature = temperature; c->dposmax = dposmax; c->dposmin = dposmin;
    c->dhoekmax = dhoekmax_rad; c->dhoekmin = dhoekmin_rad; c->padj = padj; c->nadj = nadj;
    c->start_prod = start_prod; c->quirk31 = quirk31;
}

/* ------------------------------------------------------------------ calculateLJ / calcEnergy */
/* CoM minimum image + cutoff test, MonteCarlo.F90:807-812 */
static inline int com_in_cutoff(const mco_chain *c, const REAL *ci, const REAL *cj)
{
    REAL rx = cj[0] - ci[0], ry = cj[1] - ci[1], rz = cj[2] - ci[2];
    rx = rx - c->boxl2 * R_TRUNC(rx / c->boxl);
    ry = ry - c->boxl2 * R_TRUNC(ry / c->boxl);
    rz = rz - c->boxl2 * R_TRUNC(rz / c->boxl);
    REAL len = R_SQRT(rx * rx + ry * ry + rz * rz);
    return !(len > c->cutoff); /* IF (length(R) .GT. 7.D0) EN = 0 */
}

/* pair energy with molecule a as MOL1 (already rotated into mol1) and b as MOL2 */
static inline REAL pair_energy(const mco_chain *c, const REAL *mol1, int b, REAL *mol2, uint64_t *np)
{
    rotmatrix_r(&c->com[3 * b], &c->hoek[3 * b], c->dmso_rel, c->ndmso, mol2); /* :816 */
    return calc_lj_r(mol1, mol2, c->is_h, c->dmso_tab, c->ndmso, c->boxl, c->boxl2, c->conv, np); /* :817 */
}

/* calculateLJ(I), MonteCarlo.F90:792-828 */
void mco_calculate_lj(mco_chain *c, int i)
{
    int n = c->n;
    uint64_t np = 0;
    if (c->ss) c->ss[(size_t)i * n + i] = 0;
    rotmatrix_r(&c->com[3 * i], &c->hoek[3 * i], c->dmso_rel, c->ndmso, c->mol1); /* :802 */
    for (int j = 0; j < n; ++j) {
        if (j == i) continue;
        REAL en = 0;
        if (com_in_cutoff(c, &c->com[3 * i], &c->com[3 * j])) en = pair_energy(c, c->mol1, j, c->mol2, &np);
        if (c->ss) { c->ss[(size_t)i * n + j] = en; c->ss[(size_t)j * n + i] = en; } /* :819-820 */
    }
    c->energy[i] = calc_lj_sv_r(c->mol1, c->solute, c->is_h, c->dmso_tab, c->sol_tab, c->ndmso, c->nsol,
                                c->boxl, c->boxl2, c->conv, &np); /* :825-826 */
    c->last_sitepairs = np;
}

/* calcEnergy, MonteCarlo.F90:988-1004: strictly in this order */
double mco_calc_energy(const mco_chain *c)
{
    int n = c->n;
    REAL out = 0;
    for (int i = 0; i < n; ++i) {
        out = out + c->energy[i];
        for (int j = i + 1; j < n; ++j) out = out + c->ss[(size_t)i * n + j];
    }
    return (double)out;
}

static REAL calc_energy_r(const mco_chain *c)
{
    int n = c->n;
    REAL out = 0;
    for (int i = 0; i < n; ++i) {
        out = out + c->energy[i];
        for (int j = i + 1; j < n; ++j) out = out + c->ss[(size_t)i * n + j];
    }
    return out;
}

/* start-up recompute, MonteCarlo.F90:357-364.  Every unordered pair is evaluated twice; the later call
 * (the larger index as MOL1) is the one that stays in SS(i,j) and SS(j,i). */
double mco_total_energy(mco_chain *c)
{
    uint64_t np = 0;
    for (int i = 0; i < c->n; ++i) { mco_calculate_lj(c, i); np += c->last_sitepairs; }
    c->last_sitepairs = np;
    c->toteng = calc_energy_r(c);
    return (double)c->toteng;
}

/* Matrix-free: same values as the SS matrix would hold after the start-up loop (pair (i<j) evaluated with
 * j as MOL1, i as MOL2), summed in calcEnergy's order.  For boxes whose N x N matrix cannot be stored. */
double mco_total_energy_matrixfree(const mco_chain *c, double *e_solute_sum, uint64_t *n_pairs_in_cutoff)
{
    int n = c->n, na = c->ndmso;
    REAL *sites = (REAL *)malloc(sizeof(REAL) * 3 * (size_t)na * n);
    for (int i = 0; i < n; ++i) rotmatrix_r(&c->com[3 * i], &c->hoek[3 * i], c->dmso_rel, na, &sites[(size_t)3 * na * i]);
    REAL out = 0, esum = 0;
    uint64_t npair = 0;
    for (int i = 0; i < n; ++i) {
        REAL ei = calc_lj_sv_r(&sites[(size_t)3 * na * i], c->solute, c->is_h, c->dmso_tab, c->sol_tab, na, c->nsol,
                               c->boxl, c->boxl2, c->conv, NULL);
        out = out + ei; esum = esum + ei;
        for (int j = i + 1; j < n; ++j) {
            if (!com_in_cutoff(c, &c->com[3 * j], &c->com[3 * i])) continue;
            ++npair;
            out = out + calc_lj_r(&sites[(size_t)3 * na * j], &sites[(size_t)3 * na * i], c->is_h, c->dmso_tab, na,
                                  c->boxl, c->boxl2, c->conv, NULL);
        }
    }
    free(sites);
    if (e_solute_sum) *e_solute_sum = (double)esum;
    if (n_pairs_in_cutoff) *n_pairs_in_cutoff = npair;
    return (double)out;
}

/* Cell list, used ONLY to find the pairs whose CoM minimum-image distance passes the cutoff test; the pair
 * arithmetic and the summation order are those of mco_total_energy_matrixfree.  Valid when every CoM lies
 * inside [-BOXL, BOXL] (true after the wrap at MonteCarlo.F90:710-712) and the box holds >= 3 cells per
 * side; otherwise falls back to the brute-force loop. */
static int cmp_int(const void *a, const void *b) { return (*(const int *)a > *(const int *)b) - (*(const int *)a < *(const int *)b); }

double mco_total_energy_cells(const mco_chain *c, double *e_solute_sum, uint64_t *n_pairs_in_cutoff)
{
    int n = c->n, na = c->ndmso;
    double boxl = (double)c->boxl, side = 2.0 * boxl, cut = (double)c->cutoff;
    int nc = (int)floor(side / cut);
    int inside = 1;
    for (int i = 0; i < 3 * n; ++i) if (fabs((double)c->com[i]) > boxl) { inside = 0; break; }
    if (nc < 3 || !inside) return mco_total_energy_matrixfree(c, e_solute_sum, n_pairs_in_cutoff);
    if (nc > 256) nc = 256;
    double cs = side / nc;
    size_t ncell = (size_t)nc * nc * nc;
    int *head = (int *)malloc(sizeof(int) * (ncell + 1)), *cell_of = (int *)malloc(sizeof(int) * n);
    int *order = (int *)malloc(sizeof(int) * n);
    memset(head, 0, sizeof(int) * (ncell + 1));
    for (int i = 0; i < n; ++i) {
        int ix[3];
        for (int k = 0; k < 3; ++k) {
            int q = (int)floor(((double)c->com[3 * i + k] + boxl) / cs);
            ix[k] = q < 0 ? 0 : (q >= nc ? nc - 1 : q);
        }
        cell_of[i] = (ix[2] * nc + ix[1]) * nc + ix[0];
        head[cell_of[i] + 1]++;
    }
    for (size_t k = 0; k < ncell; ++k) head[k + 1] += head[k];
    int *fill = (int *)malloc(sizeof(int) * ncell);
    for (size_t k = 0; k < ncell; ++k) fill[k] = head[k];
    for (int i = 0; i < n; ++i) order[fill[cell_of[i]]++] = i;
    free(fill);

    REAL *sites = (REAL *)malloc(sizeof(REAL) * 3 * (size_t)na * n);
    for (int i = 0; i < n; ++i) rotmatrix_r(&c->com[3 * i], &c->hoek[3 * i], c->dmso_rel, na, &sites[(size_t)3 * na * i]);
    REAL out = 0, esum = 0;
    uint64_t npair = 0;
    int cap = 256, *nb = (int *)malloc(sizeof(int) * cap);
    for (int i = 0; i < n; ++i) {
        REAL ei = calc_lj_sv_r(&sites[(size_t)3 * na * i], c->solute, c->is_h, c->dmso_tab, c->sol_tab, na, c->nsol,
                               c->boxl, c->boxl2, c->conv, NULL);
        out = out + ei; esum = esum + ei;
        int ci = cell_of[i], cx = ci % nc, cy = (ci / nc) % nc, cz = ci / (nc * nc), cnt = 0;
        for (int dz = -1; dz <= 1; ++dz) for (int dy = -1; dy <= 1; ++dy) for (int dx = -1; dx <= 1; ++dx) {
            int cc = (((cz + dz + nc) % nc) * nc + (cy + dy + nc) % nc) * nc + (cx + dx + nc) % nc;
            for (int p = head[cc]; p < head[cc + 1]; ++p) {
                int j = order[p];
                if (j <= i) continue;
                if (!com_in_cutoff(c, &c->com[3 * j], &c->com[3 * i])) continue;
                if (cnt == cap) { cap *= 2; nb = (int *)realloc(nb, sizeof(int) * cap); }
                nb[cnt++] = j;
            }
        }
        qsort(nb, cnt, sizeof(int), cmp_int);
        for (int k = 0; k < cnt; ++k) {
            int j = nb[k];
            out = out + calc_lj_r(&sites[(size_t)3 * na * j], &sites[(size_t)3 * na * i], c->is_h, c->dmso_tab, na,
                                  c->boxl, c->boxl2, c->conv, NULL);
        }
        npair += (uint64_t)cnt;
    }
    free(nb); free(sites); free(head); free(cell_of); free(order);
    if (e_solute_sum) *e_solute_sum = (double)esum;
    if (n_pairs_in_cutoff) *n_pairs_in_cutoff = npair;
    return (double)out;
}

/* Row i as calculateLJ(i) would write it, without touching the chain (MonteCarlo.F90:802-826). */
double mco_row_energy(const mco_chain *c, int i, double *ss_row, double *e_i)
{
    int n = c->n, na = c->ndmso;
    REAL *m1 = (REAL *)malloc(sizeof(REAL) * 6 * (size_t)na), *m2 = m1 + 3 * na;
    rotmatrix_r(&c->com[3 * i], &c->hoek[3 * i], c->dmso_rel, na, m1);
    REAL sum = 0;
    for (int j = 0; j < n; ++j) {
        REAL en = 0;
        if (j != i && com_in_cutoff(c, &c->com[3 * i], &c->com[3 * j])) en = pair_energy(c, m1, j, m2, NULL);
        if (ss_row) ss_row[j] = (double)en;
        sum = sum + en;
    }
    REAL ei = calc_lj_sv_r(m1, c->solute, c->is_h, c->dmso_tab, c->sol_tab, na, c->nsol, c->boxl, c->boxl2, c->conv, NULL);
    if (e_i) *e_i = (double)ei;
    free(m1);
    return (double)(sum + ei);
}

void mco_get_matrix(const mco_chain *c, double *ss, double *energy)
{
    int n = c->n;
    if (ss && c->ss) for (size_t k = 0; k < (size_t)n * n; ++k) ss[k] = (double)c->ss[k];
    if (energy) for (int i = 0; i < n; ++i) energy[i] = (double)c->energy[i];
}

/* ------------------------------------------------------------------ MCINIT / METROPOLIS */
/* back up exactly what one step can change (the reference copies whole arrays, MonteCarlo.F90:690-694;
 * only molecule rsolv's entries can differ afterwards) */
static void backup_row(mco_chain *c, int i)
{
    int n = c->n;
    memcpy(&c->com_old[3 * i], &c->com[3 * i], 3 * sizeof(REAL));
    memcpy(&c->hoek_old[3 * i], &c->hoek[3 * i], 3 * sizeof(REAL));
    c->eold[i] = c->energy[i];
    if (c->ss) memcpy(&c->ssold[(size_t)i * n], &c->ss[(size_t)i * n], sizeof(REAL) * n);
    c->toteng_old = c->toteng;
}

static void restore_row(mco_chain *c, int i)
{
    int n = c->n;
    memcpy(&c->com[3 * i], &c->com_old[3 * i], 3 * sizeof(REAL));
    memcpy(&c->hoek[3 * i], &c->hoek_old[3 * i], 3 * sizeof(REAL));
    c->energy[i] = c->eold[i];
    if (c->ss) {
        memcpy(&c->ss[(size_t)i * n], &c->ssold[(size_t)i * n], sizeof(REAL) * n);
        for (int j = 0; j < n; ++j) c->ss[(size_t)j * n + i] = c->ssold[(size_t)i * n + j];
    }
    c->toteng = c->toteng_old;
}

/* MCINIT, MonteCarlo.F90:685-715; randVec lib.f90:111-120; randVecHoek lib.f90:124-131 */
int mco_mcinit(mco_chain *c, const double u[7])
{
    int n = c->n;
    int rsolv = (int)((REAL)u[0] * n) + 1;           /* :697 INT(RAND()*NCOM)+1, 1-based */
    if (c->quirk31 && rsolv == 31) rsolv = 30;       /* :698 */
    if (rsolv > n) rsolv = n;                        /* unreachable for u < 1; guards the array */
    int i = rsolv - 1;
    backup_row(c, i);
    REAL inv_sqrt3 = (REAL)1.0 / R_SQRT((REAL)3.0);  /* vecdiv multiplies by 1/M, vector_class.f90:73-82 */
    REAL pi = c_pi(), tau = 2.0 * c_pi();
    for (int k = 0; k < 3; ++k) {
        REAL rv = ((REAL)u[1 + k] - (REAL)0.5) * 2;  /* lib.f90:115-117 */
        rv = inv_sqrt3 * rv;                         /* :118 */
        rv = c->dposmax * rv;                        /* :119 */
        c->com[3 * i + k] = c->com[3 * i + k] + rv;  /* MonteCarlo.F90:699 */
    }
    for (int k = 0; k < 3; ++k) {
        REAL rv = ((REAL)u[4 + k] - (REAL)0.5) * 2 * c->dhoekmax; /* lib.f90:128-130 */
        c->hoek[3 * i + k] = c->hoek[3 * i + k] + rv;             /* :700 */
    }
    for (int k = 0; k < 3; ++k) {                    /* :705-707 */
        REAL h = c->hoek[3 * i + k];
        c->hoek[3 * i + k] = h - tau * R_TRUNC(h / pi);
    }
    for (int k = 0; k < 3; ++k) {                    /* :710-712 */
        REAL x = c->com[3 * i + k];
        c->com[3 * i + k] = x - c->boxl2 * R_TRUNC(x / c->boxl);
    }
    c->rsolv = i;
    return i;
}

/* everything of one loop_LJ iteration after the trial move: MonteCarlo.F90:411-431 and METROPOLIS :720-788 */
static int finish_step(mco_chain *c, int i, const double *u_accept, double trace[10])
{
    int used = 0;
    c->step += 1;
    mco_calculate_lj(c, i);                           /* :411 */
    c->sitepairs += c->last_sitepairs;
    c->toteng = calc_energy_r(c);                     /* :413-414 */
    if (c->toteng > (REAL)DBL_MAX) { c->toteng = (REAL)DBL_MAX; c->nclamp++; }   /* :421-424, HUGE-1 == HUGE in fp64 */
    if (c->toteng != c->toteng) { c->toteng = (REAL)DBL_MAX; c->nclamp++; }      /* :425-428 */
    REAL toteng_trial = c->toteng;
    REAL delta = c->toteng - c->toteng_old;           /* :728 */
    REAL exponent = (REAL)-1.0 * delta * (REAL)1000.0 / ((REAL)8.314 * c->temperature); /* :729 */
    REAL kans, rv;
    if (exponent < (REAL)-75.0) { kans = 0; rv = 1; }            /* :730-733, no draw */
    else if (exponent >= 0) { kans = 1; rv = 0; }                /* :734-736, no draw */
    else {
        kans = R_POW((REAL)mco_const_e(), exponent);             /* :738 */
        rv = u_accept[0]; used = 1;                              /* :739 */
        if (kans > 1) kans = 1;
    }
    int accepted = rv <= kans;                                   /* :744 */
    if (accepted) {
        c->nsuc++; c->naccept++;
        if (c->step >= c->start_prod) { c->production = c->production + c->toteng; c->prod_acc++; } /* :749-752 */
    }
    c->ratio = (double)((float)c->naccept / (float)c->nadj);     /* :758, default REAL quotient */
    double dpos_print = (double)c->dposmax, dhoek_print = (double)c->dhoekmax;
    if (c->step % c->nadj == 0) {                                /* :766-777 */
        if (c->ratio > 0.5) { c->dposmax = c->dposmax * ((REAL)1.0 + c->padj); c->dhoekmax = c->dhoekmax * ((REAL)1.0 + c->padj); }
        else if (c->ratio < 0.5) { c->dposmax = c->dposmax * ((REAL)1.0 - c->padj); c->dhoekmax = c->dhoekmax * ((REAL)1.0 - c->padj); }
        if (c->dposmax < c->dposmin) c->dposmax = c->dposmin;
        if (c->dhoekmax < c->dhoekmin) c->dhoekmax = c->dhoekmin;
        c->naccept = 0;
    }
    if (!accepted) restore_row(c, i);                            /* :779-786 */
    if (trace) {
        trace[0] = (double)toteng_trial; trace[1] = (double)c->toteng_old; trace[2] = (double)kans;
        trace[3] = (double)rv; trace[4] = i; trace[5] = accepted; trace[6] = c->ratio;
        trace[7] = dpos_print; trace[8] = dhoek_print; trace[9] = (double)c->last_sitepairs;
    }
    return used;
}

int mco_step(mco_chain *c, const double *u, double trace[10])
{
    int i = mco_mcinit(c, u);
    return 7 + finish_step(c, i, u + 7, trace);
}

int mco_step_given_move(mco_chain *c, int i, const double new_com[3], const double new_hoek[3],
                        const double *u_accept, double trace[10])
{
    backup_row(c, i);
    for (int k = 0; k < 3; ++k) { c->com[3 * i + k] = new_com[k]; c->hoek[3 * i + k] = new_hoek[k]; }
    c->rsolv = i;
    return finish_step(c, i, u_accept, trace);
}

int64_t mco_run_steps(mco_chain *c, int64_t nsteps, const double *uniforms, int64_t cap, int64_t *consumed,
                      double *trace)
{
    int64_t pos = 0, s;
    for (s = 0; s < nsteps; ++s) {
        if (pos + 8 > cap) break;
        pos += mco_step(c, uniforms + pos, trace ? trace + 10 * s : NULL);
    }
    if (consumed) *consumed = pos;
    return s;
}

/* ------------------------------------------------------------------ uniform stream */
/* Counter-based stream (shared definition with include/mcb200.h, mcb_seed_chains):
 *   G = 0x9E3779B97F4A7C15, mix = splitmix64 finaliser, key = mix(seed + G),
 *   u_n = (mix(key + (n+1)*G) >> 11) * 2^-53 for n = 0, 1, 2, ...   state = {key, n}. */
static inline uint64_t mix64(uint64_t x)
{
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

void mco_rng_seed(uint64_t s[2], uint64_t seed)
{
    s[0] = mix64(seed + 0x9E3779B97F4A7C15ULL);
    s[1] = 0;
}

double mco_rng_uniform(uint64_t s[2])
{
    s[1] += 1;
    uint64_t x = mix64(s[0] + s[1] * 0x9E3779B97F4A7C15ULL);
    return (double)(x >> 11) * 0x1.0p-53;
}

/* u(n) of the stream, random access */
static double rng_at(uint64_t key, uint64_t n)
{
    uint64_t x = mix64(key + (n + 1) * 0x9E3779B97F4A7C15ULL);
    return (double)(x >> 11) * 0x1.0p-53;
}

/* Step s (0-based over the chain's life) owns u(8s)..u(8s+7): seven for MCINIT, the eighth only if METROPOLIS
 * draws (MonteCarlo.F90:739).  rng[0] = key; rng[1] is set to the number of uniforms consumed so far. */
int64_t mco_run_steps_rng(mco_chain *c, int64_t nsteps, uint64_t rng[2], double *trace)
{
    double u[8];
    for (int64_t s = 0; s < nsteps; ++s) {
        uint64_t base = 8ULL * (uint64_t)c->step;
        for (int k = 0; k < 8; ++k) u[k] = rng_at(rng[0], base + k);
        int i = mco_mcinit(c, u);
        rng[1] += 7 + finish_step(c, i, u + 7, trace ? trace + 10 * s : NULL);
    }
    return nsteps;
}

/* ------------------------------------------------------------------ replica exchange (no counterpart in the reference)
 * The reference has one temperature per process and no exchange move.  The product's mcb_exchange applies the
 * reference's own acceptance rule, METROPOLIS (MonteCarlo.F90:728-744), to the parallel-tempering exponent
 * (beta_a - beta_b)(E_a - E_b), beta = 1000/(8.314 T); this is that rule, literally, for two chains.  An accepted swap
 * exchanges the temperatures and the step sizes.  *used = 1 if u was consumed.  Returns 1 if accepted. */
int mco_exchange(mco_chain *a, mco_chain *b, double u, int *used)
{
    REAL ba = (REAL)1000.0 / ((REAL)8.314 * a->temperature), bb = (REAL)1000.0 / ((REAL)8.314 * b->temperature);
    REAL exponent = (ba - bb) * (a->toteng - b->toteng);
    REAL kans, rv;
    int drew = 0;
    if (exponent < (REAL)-75.0) { kans = 0; rv = 1; }            /* :730-733 */
    else if (exponent >= 0) { kans = 1; rv = 0; }                /* :734-736 */
    else {
        kans = R_POW((REAL)mco_const_e(), exponent);             /* :738 */
        rv = u; drew = 1;                                        /* :739 */
        if (kans > 1) kans = 1;
    }
    if (used) *used = drew;
    if (!(rv <= kans)) return 0;                                 /* :744 */
    REAL t;
    t = a->temperature; a->temperature = b->temperature; b->temperature = t;
    t = a->dposmax; a->dposmax = b->dposmax; b->dposmax = t;
    t = a->dhoekmax; a->dhoekmax = b->dhoekmax; b->dhoekmax = t;
    return 1;
}

/* the exchange stream of a chain seeded with `seed` (include/mcb200.h, mcb_exchange): key_x = mix(key + C), state = {key_x, n} */
void mco_rng_exchange_seed(uint64_t s[2], uint64_t seed)
{
    s[0] = mix64(mix64(seed + 0x9E3779B97F4A7C15ULL) + 0xD1B54A32D192ED03ULL);
    s[1] = 0;
}

double mco_get_temperature(const mco_chain *c) { return (double)c->temperature; }

void mco_get_stats(const mco_chain *c, double out[12])
{
    out[0] = (double)c->step; out[1] = (double)c->nsuc; out[2] = (double)c->naccept; out[3] = (double)c->prod_acc;
    out[4] = (double)c->nclamp; out[5] = (double)c->sitepairs; out[6] = (double)c->toteng; out[7] = (double)c->toteng_old;
    out[8] = (double)c->production; out[9] = (double)c->dposmax; out[10] = (double)c->dhoekmax; out[11] = c->ratio;
}

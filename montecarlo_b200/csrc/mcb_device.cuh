// mcb_device.cuh — device-side building blocks of the LJ trial-move energy path (sm_100a, fp64).
//
// What the reference does per Monte-Carlo step (MonteCarlo/MonteCarlo.F90:406-433) is restated here for one
// warp per chain.  The arithmetic that decides parity is kept literal where it matters (minimum image on RAW
// site differences with AINT, LennardJones.f90:178-180; H sites skipped, :66,68; 7 A centre-of-mass cutoff,
// MonteCarlo.F90:807-813) and reorganised where it does not (per-type-pair constants 4*eps*sigma^12 and
// 4*eps*sigma^6 are tabulated once; one reciprocal of r^2 replaces six divisions and two square roots).
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace mcb {

constexpr unsigned FULL = 0xffffffffu;
constexpr int NSITE = 4;          // interacting sites per solvent molecule (lane & 3 owns one)
constexpr int SITE_DOUBLES = 12;  // one molecule's rotated sites: [site][xyz]
// One (type, own site a) row of tabAB: four (A, B) pairs + one pair of padding.  With 64-byte rows every row starts on
// bank 0 or 16 and the 16 rows a quarter-warp reads at once collide 8 ways (ncu: 10 shared-memory wavefronts per load,
// the largest single user of the shared-memory pipe); with 80-byte rows they spread over all 32 banks.
constexpr int TAB_ROW = 10;

// Per-chain "image": every array a chain's warp touches, contiguous in HBM so that staging it into shared
// memory is one linear, coalesced copy.  Offsets are in doubles.
//   com    [3][Np]      SoA planes x|y|z of the centres of mass          (reference CoM(:)); Np >= N even, not a multiple of 16
//   hoek   [3][Np]      SoA planes of the Euler angles                   (reference HOEK(:))
//   energy [Np]         solvent-solute energy per molecule               (reference ENERGY(:))
//   sites  [N+nsolp][4][3]  rotated absolute site coordinates (cache of RotMatrix output, lib.f90:27-83);
//                       entries N.. are the solute atoms grouped in fours ("pseudo-molecules"), padded with
//                       far-away atoms whose parameters are zero
//   tabAB  [1+nsolp][4][TAB_ROW]  (A,B) = (4 eps s^12, 4 eps s^6) for (type, own site a, other site b), rows padded;
//                       type 0 = solvent-solvent, type 1+p = solute pseudo-molecule p
//   tabQ   [1+nsolp][4][4]     q_a q_b conv (only read when conv != 0)
//   mat    [N][MS]      pair matrix SOLVENTSOLVENT (MS odd: conflict-free column access in shared memory);
//                       part of the image only when it is staged into shared memory
struct Geo {
    int N, Np, nsol, nsolp, MS;
    int o_com, o_hoek, o_energy, o_sites, o_tabAB, o_tabQ, o_mat;
    int image_doubles;      // per-chain stride of the image array in HBM (includes mat iff mat_in_image)
    int stage_doubles;      // how much of the image is copied to shared memory
    int o_newrow, o_list, o_epart, o_pend, o_touch;  // work arrays (shared memory in k_run_steps, scratch in the single-step kernels)
    int work_doubles;       // size of one work area, in doubles
    int spec_group_doubles; // stride between the work areas of the look-ahead groups of k_run_spec (= work_doubles)
    int mat_in_image;
    int quirk31;
    double boxl, boxl2, inv_boxl, cut2, conv, cutoff;
};

// Scalars of one chain (the driver's TOTENG, DPOSMAX, NSUC, ... MonteCarlo.F90:33-47,98-101).
struct ChainCtl {
    double toteng, dposmax, dhoekmax, production, sum_e, sum_e2, ratio, last_delta;
    long long step, nsuc, naccept, prod_acc, nclamp, sitepairs, uused, nstat;
    unsigned long long rng_key, rng_n;
    double temperature, dposmin, dhoekmin, padj;
    long long start_prod;
    int nadj, energies_valid;
    long long xatt, xacc;          // replica-exchange attempts / acceptances this chain took part in (mcb_exchange)
    unsigned long long xn;         // position in the chain's exchange stream
};

struct ChainPtrs {
    double *com, *hoek, *energy, *sites;
    const double *tabAB, *tabQ;
    double *mat;     // row stride g.MS
    double *newrow;  // [Np]
    int *list;       // [nsolp + Np]
};

__device__ __forceinline__ ChainPtrs chain_ptrs(const Geo &g, double *img, double *mat, double *work)
{
    ChainPtrs p;
    p.com = img + g.o_com; p.hoek = img + g.o_hoek; p.energy = img + g.o_energy; p.sites = img + g.o_sites;
    p.tabAB = img + g.o_tabAB; p.tabQ = img + g.o_tabQ;
    p.mat = mat; p.newrow = work + g.o_newrow; p.list = reinterpret_cast<int *>(work + g.o_list);
    return p;
}

__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(FULL, v, m); }
__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(FULL, v, src); }

__device__ __forceinline__ double warp_sum(double v)
{
    v += shfl_xor_d(v, 16); v += shfl_xor_d(v, 8); v += shfl_xor_d(v, 4); v += shfl_xor_d(v, 2); v += shfl_xor_d(v, 1);
    return v;
}

// The same sum over the 32 lanes through the fp64 tensor-core product (mma.sync m8n8k4, DMMA): with an all-ones B,
// D[i][j] = sum_k A[i][k] leaves in every lane the sum over its group of four lanes; lane (i, k) then fetches the
// sums of groups k and k + 4 and a second product adds those four numbers.  Two DMMAs, two shuffled doubles and one
// add instead of five shuffle-add levels; fixed (hardware) summation order, so deterministic like warp_sum.
__device__ __forceinline__ double warp_sum_mma(double v, int lane)
{
    double g0, g1 = 0.0, t0, t1 = 0.0;
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
        : "=d"(g0), "=d"(g1) : "d"(v), "d"(1.0), "d"(0.0), "d"(0.0));
    const int k = lane & 3;
    const double u = shfl_d(g0, 4 * k) + shfl_d(g0, 4 * k + 16);
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
        : "=d"(t0), "=d"(t1) : "d"(u), "d"(1.0), "d"(0.0), "d"(0.0));
    (void)g1; (void)t1;   // (every column of D holds the same sum)
    return t0;
}

// Minimum image as the reference writes it, R - BOXL2*AINT(R/BOXL) with BOXL = half side
// (LennardJones.f90:178-180) — not a nearest image once |R| >= 2*BOXL, and that is part of the contract.
// Only the SQUARE of the result is ever used, so the kernel works on |R|:
//     |R - BOXL2*AINT(R/BOXL)| = | |R| - BOXL2*floor(|R|/BOXL) |
// and floor comes from one round-down FMA onto 2^52 (whose ulp is 1) instead of an fp64 FRND, which on B200
// issues once per ~13 cycles per warp (tools/latency.cu) and would cost as much as all the DFMAs of a site pair.
// The quotient uses the reciprocal of BOXL, so floor can differ from the reference's AINT(R/BOXL) only when
// |R|/BOXL lies within an ulp of an integer - a set of measure ~1e-16 per component, not a harmless one: at an odd
// quotient the two candidates are the two equally distant images (r^2 unchanged to 1e-15), but at an EVEN quotient
// (|R| = 2*BOXL, reachable: see above) they are |R| ~ 0 and |R| ~ 2*BOXL, a different energy.  The argument for
// accepting it is the probability, not equidistance; SURVEY.md section 7 takes the same view.
__device__ __forceinline__ double mic_abs(double d, const Geo &g)
{
    const double a = fabs(d);
    const double t = __fma_rd(a, g.inv_boxl, 4503599627370496.0) - 4503599627370496.0;
    return fma(-g.boxl2, t, a);
}
// With the true division, for the coordinate wraps of MCINIT (MonteCarlo.F90:705-712).
__device__ __forceinline__ double wrap_exact(double x, double period, double half) { return x - period * trunc(x / half); }

// 1/x for squared site distances: MUFU.RCP64H seed (~2^-20) and one cubically convergent step
// y = y0*(1 + e + e^2), e = 1 - x*y0 (no denormal/overflow fix-up: r^2 = 0 gives NaN, which METROPOLIS clamps
// and rejects like the reference's Inf - Inf).
__device__ __forceinline__ double rcp_nr(double x)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x, y, 1.0);
    return fma(y, fma(e, e, e), y);
}

// Squared minimum-image distance of one site pair (LJ_FORMULA, LennardJones.f90:178-182).
__device__ __forceinline__ double pair_r2(double dx, double dy, double dz, const Geo &g)
{
    dx = mic_abs(dx, g); dy = mic_abs(dy, g); dz = mic_abs(dz, g);
    return fma(dz, dz, fma(dy, dy, dx * dx));
}

// acc + one site pair's energy.  LJ_FORMULA (LennardJones.f90:153-187) with tabulated A = 4 eps s^12,
// B = 4 eps s^6:  4 eps (s^12/r^12 - s^6/r^6) = s6*(A*s6 - B), s6 = (1/r^2)^3;  Coulomb q1 q2 conv / r.
template <bool COUL>
__device__ __forceinline__ double site_pair_acc(double acc, double dx, double dy, double dz, double A, double B, double QQ,
                                                const Geo &g)
{
    const double r2 = pair_r2(dx, dy, dz, g);
    if (COUL) {
        const double rinv = rsqrt(r2);
        const double inv = rinv * rinv;
        const double s6 = inv * inv * inv;
        return fma(s6, fma(A, s6, -B), fma(QQ, rinv, acc));
    } else {
        const double inv = rcp_nr(r2);
        const double s6 = inv * inv * inv;
        return fma(s6, fma(A, s6, -B), acc);
    }
}

// RotMatrix for ONE site (MC_lib/lib.f90:52-81): c/s = cos/sin of (HOEK%x, HOEK%y, HOEK%z).
__device__ __forceinline__ void rotate_site(double c1, double s1, double c2, double s2, double c3, double s3,
                                            double rx, double ry, double rz, double cx, double cy, double cz,
                                            double &ox, double &oy, double &oz)
{
    double m11 = c1 * c2 - s1 * s2 * c3, m12 = s1 * c2 + c1 * s2 * c3, m13 = s2 * s3;
    double m21 = -1.0 * c1 * s2 - s1 * c2 * c3, m22 = -1.0 * s1 * s2 + c1 * c2 * c3, m23 = c2 * s3;
    double m31 = s1 * s3, m32 = -1.0 * c1 * s3, m33 = c3;
    ox = cx + (m11 * rx + m12 * ry + m13 * rz);
    oy = cy + (m21 * rx + m22 * ry + m23 * rz);
    oz = cz + (m31 * rx + m32 * ry + m33 * rz);
}

// sin and cos of an angle in [-2 pi, 2 pi] (the wrapped Euler angles, MonteCarlo.F90:705-707), for RotMatrix
// (lib.f90:52-57).  Cody-Waite reduction to |r| <= pi/4 with pi/2 in two parts under FMA, then the fdlibm
// kernel polynomials (error < 1 ulp), coefficients in constant memory so that the DFMAs read them as c[][]
// operands.  The library sincos() costs ~50 warp instructions here, 16 of them UMOVs that only materialise
// coefficients; this one ~30.
__constant__ double SINCOS_C[16] = {
    -1.66666666666666324348e-01, 8.33333333332248946124e-03, -1.98412698298579493134e-04,     // S1..S6
    2.75573137070700676789e-06, -2.50507602534068634195e-08, 1.58969099521155010221e-10,
    4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05,      // C1..C6
    -2.75573143513906633035e-07, 2.08757232129817482790e-09, -1.13596475577881948265e-11,
    6.36619772367581382433e-01,   // 2/pi
    1.57079632679489655800e+00,   // pi/2 high
    6.12323399573676603587e-17,   // pi/2 low
    6755399441055744.0};          // 1.5 * 2^52
__device__ __forceinline__ void sincos_small(double x, double &sn, double &cs)
{
    const double t = fma(x, SINCOS_C[12], SINCOS_C[15]);
    const int k = __double2loint(t);
    const double kd = t - SINCOS_C[15];
    double r = fma(-kd, SINCOS_C[13], x);
    r = fma(-kd, SINCOS_C[14], r);
    const double z = r * r;
    double ps = fma(z, SINCOS_C[5], SINCOS_C[4]);
    double pc = fma(z, SINCOS_C[11], SINCOS_C[10]);
    ps = fma(z, ps, SINCOS_C[3]); pc = fma(z, pc, SINCOS_C[9]);
    ps = fma(z, ps, SINCOS_C[2]); pc = fma(z, pc, SINCOS_C[8]);
    ps = fma(z, ps, SINCOS_C[1]); pc = fma(z, pc, SINCOS_C[7]);
    ps = fma(z, ps, SINCOS_C[0]); pc = fma(z, pc, SINCOS_C[6]);
    const double s0 = fma(z * r, ps, r);                    // r + r^3 * S(z)
    const double c0 = fma(z * z, pc, fma(z, -0.5, 1.0));    // 1 - z/2 + z^2 * C(z)
    const double ss = (k & 1) ? c0 : s0, cc = (k & 1) ? s0 : c0;
    sn = (k & 2) ? -ss : ss;
    cs = ((k + 1) & 2) ? -cc : cc;
}

// The same for any argument: first reduced to [-pi, pi] (the error of the reduction, |x| 2^-53, is that of forming x).
__device__ __forceinline__ void sincos_any(double x, double &sn, double &cs)
{
    sincos_small(fma(-6.283185307179586, rint(x * 0.15915494309189535), x), sn, cs);
}

// Uniform stream (include/mcb200.h, mcb_seed_chains): u_n = (mix(key + (n+1) G) >> 11) * 2^-53, random access
// in n; Metropolis step s (0-based) of a chain owns n = 8 s .. 8 s + 7.
__device__ __forceinline__ unsigned long long mix64(unsigned long long x)
{
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
__device__ __forceinline__ double stream_uniform(unsigned long long key, unsigned long long n)
{
    unsigned long long x = mix64(key + (n + 1ULL) * 0x9E3779B97F4A7C15ULL);
    return (double)(x >> 11) * 0x1.0p-53;
}

// Row I of the pair matrix for a (trial) position of molecule i — calculateLJ(I), MonteCarlo.F90:792-828.
//   (ncx,ncy,ncz)  centre of mass of molecule i (trial)
//   (sx,sy,sz)     rotated absolute coordinates of site a = lane&3 of molecule i (trial)
//   jlimit         partners j < jlimit are considered (N for a trial move; i for the start-up triangle)
// Results: newrow[0..N) = the new row (0 outside the cutoff), list[0..nn) the in-cutoff partners'
// indices, e_sol = ENERGY(I) (CalcLJ_SV, LennardJones.f90:89-148).  Returns nn.
// Work split: centre-of-mass gate, one partner per lane; then (partner, own site a) items, four lanes per
// partner, each lane sweeping the partner's four sites b; solute atoms ride along as pseudo-molecules.
template <bool COUL>
__device__ __forceinline__ int row_eval(const Geo &g, const ChainPtrs &p, int i, int jlimit, double ncx, double ncy,
                                        double ncz, double sx, double sy, double sz, int lane, double &e_sol)
{
    const unsigned lt = (1u << lane) - 1u;
    int nn = 0;
    for (int j = lane; j < g.N; j += 32) p.newrow[j] = 0.0;   // EN = 0.D0 outside the cutoff, :812-813
    for (int j0 = 0; j0 < jlimit; j0 += 32) {
        int j = j0 + lane;
        bool in = false;
        if (j < jlimit && j != i) {
            double r2 = pair_r2(p.com[j] - ncx, p.com[g.Np + j] - ncy, p.com[2 * g.Np + j] - ncz, g);   // R = CoM(J) - CoM(I), :806-810
            in = !(r2 > g.cut2);   // cut2 = largest x with sqrt(x) <= cutoff: same decision as length(R) .GT. 7.D0
        }
        unsigned m = __ballot_sync(FULL, in);
        if (in) p.list[nn + __popc(m & lt)] = j;
        nn += __popc(m);
    }
    __syncwarp();
    const int a = lane & 3, q = lane >> 2;
    const int nslots = nn + g.nsolp;
    double acc_sol = 0.0;
    for (int s0 = 0; s0 < nslots; s0 += 8) {
        int s = s0 + q;
        double e = 0.0;
        int j = 0;
        if (s < nslots) {
            int vm, t;
            if (s < nn) { j = p.list[s]; vm = j; t = 0; }
            else { vm = g.N + (s - nn); t = 1 + (s - nn); }
            const double2 *sp = reinterpret_cast<const double2 *>(p.sites + SITE_DOUBLES * vm);
            const double2 *ab = reinterpret_cast<const double2 *>(p.tabAB) + (t * 4 + a) * (TAB_ROW / 2);
            double2 v0 = sp[0], v1 = sp[1], v2 = sp[2], v3 = sp[3], v4 = sp[4], v5 = sp[5];
            double2 p0 = ab[0], p1 = ab[1], p2 = ab[2], p3 = ab[3];
            double q0 = 0, q1 = 0, q2 = 0, q3 = 0;
            if (COUL) { const double *tq = p.tabQ + (t * 4 + a) * 4; q0 = tq[0]; q1 = tq[1]; q2 = tq[2]; q3 = tq[3]; }
            // R = mol2(J) - mol1(I) (LennardJones.f90:69); b = 0..3 inner like the reference's J loop
            e = site_pair_acc<COUL>(0.0, v0.x - sx, v0.y - sy, v1.x - sz, p0.x, p0.y, q0, g);
            e = site_pair_acc<COUL>(e, v1.y - sx, v2.x - sy, v2.y - sz, p1.x, p1.y, q1, g);
            e = site_pair_acc<COUL>(e, v3.x - sx, v3.y - sy, v4.x - sz, p2.x, p2.y, q2, g);
            e = site_pair_acc<COUL>(e, v4.y - sx, v5.x - sy, v5.y - sz, p3.x, p3.y, q3, g);
        }
        e += shfl_xor_d(e, 1);
        e += shfl_xor_d(e, 2);
        if (s < nn) { if (a == 0) p.newrow[j] = e; }
        else if (s < nslots && a == 0) acc_sol += e;
    }
    e_sol = warp_sum(acc_sol);
    __syncwarp();
    return nn;
}

}  // namespace mcb

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

// Per-chain "image": every array a chain's warp touches, contiguous in HBM so that staging it into shared
// memory is one linear, coalesced copy.  Offsets are in doubles.
//   com    [3][Np]      SoA planes x|y|z of the centres of mass          (reference CoM(:))
//   hoek   [3][Np]      SoA planes of the Euler angles                   (reference HOEK(:))
//   energy [Np]         solvent-solute energy per molecule               (reference ENERGY(:))
//   sites  [N+nsolp][4][3]  rotated absolute site coordinates (cache of RotMatrix output, lib.f90:27-83);
//                       entries N.. are the solute atoms grouped in fours ("pseudo-molecules"), padded with
//                       far-away atoms whose parameters are zero
//   tabAB  [1+nsolp][4][4][2]  (A,B) = (4 eps s^12, 4 eps s^6) for (type, own site a, other site b);
//                       type 0 = solvent-solvent, type 1+p = solute pseudo-molecule p
//   tabQ   [1+nsolp][4][4]     q_a q_b conv (only read when conv != 0)
//   mat    [N][MS]      pair matrix SOLVENTSOLVENT (MS odd: conflict-free column access in shared memory);
//                       part of the image only when it is staged into shared memory
struct Geo {
    int N, Np, nsol, nsolp, MS;
    int o_com, o_hoek, o_energy, o_sites, o_tabAB, o_tabQ, o_mat;
    int image_doubles;      // per-chain stride of the image array in HBM (includes mat iff mat_in_image)
    int stage_doubles;      // how much of the image is copied to shared memory
    int o_newrow, o_list, o_pend;  // work arrays (shared memory in k_run_steps, scratch in the single-step kernels)
    int work_doubles;       // size of the work area (newrow + list), in doubles
    int mat_in_image;
    int quirk31;
    double boxl, boxl2, inv_boxl, cut2, conv;
};

// Scalars of one chain (the driver's TOTENG, DPOSMAX, NSUC, ... MonteCarlo.F90:33-47,98-101).
struct ChainCtl {
    double toteng, dposmax, dhoekmax, production, sum_e, sum_e2, ratio, last_delta;
    long long step, nsuc, naccept, prod_acc, nclamp, sitepairs, uused, nstat;
    unsigned long long rng_key, rng_n;
    double temperature, dposmin, dhoekmin, padj;
    long long start_prod;
    int nadj, energies_valid;
};

struct ChainPtrs {
    double *com, *hoek, *energy, *sites;
    const double *tabAB, *tabQ;
    double *mat;     // row stride g.MS
    double *newrow;  // [Np]
    int *list;       // [Np]
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

// Minimum image exactly as the reference writes it: R - BOXL2*AINT(R/BOXL), BOXL = half side
// (LennardJones.f90:178-180).  Not a nearest image once |R| >= 2*BOXL; that is part of the contract.
// The quotient is formed with the reciprocal: it can differ from R/BOXL only within one ulp of an integer.
__device__ __forceinline__ double mic_fast(double d, const Geo &g) { return fma(-g.boxl2, trunc(d * g.inv_boxl), d); }
// With the true division, for the centre-of-mass gate and the wraps (MonteCarlo.F90:705-712, :808-810).
__device__ __forceinline__ double wrap_exact(double x, double period, double half) { return x - period * trunc(x / half); }

#ifndef MCB_RCP_NEWTON
#define MCB_RCP_NEWTON 2  // seed is good to ~2^-20; two steps reach fp64 (checked on the GPU in tests/test_gpu_parity.py)
#endif
// 1/x for x in the range of squared site distances: MUFU.RCP64H seed + two Newton steps (no denormal /
// overflow fix-up path: r^2 = 0 gives NaN, which METROPOLIS clamps and rejects like the reference's Inf-Inf).
__device__ __forceinline__ double rcp_nr(double x)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    y = fma(y, e, y);
#if MCB_RCP_NEWTON > 2
    e = fma(-x, y, 1.0);
    y = fma(y, e, y);
#endif
    return y;
}

// One site pair, LJ_FORMULA (LennardJones.f90:153-187) with tabulated A = 4 eps s^12, B = 4 eps s^6:
// 4 eps (s^12/r^12 - s^6/r^6) = s6*(A*s6 - B), s6 = (1/r^2)^3.
template <bool COUL>
__device__ __forceinline__ double site_pair(double dx, double dy, double dz, double A, double B, double QQ, const Geo &g)
{
    dx = mic_fast(dx, g); dy = mic_fast(dy, g); dz = mic_fast(dz, g);
    double r2 = fma(dz, dz, fma(dy, dy, dx * dx));
    if (COUL) {
        double rinv = rsqrt(r2);          // q1 q2 / r * conv  (LennardJones.f90:184, :80)
        double inv = rinv * rinv;
        double s6 = inv * inv * inv;
        return fma(s6, fma(A, s6, -B), QQ * rinv);
    } else {
        double inv = rcp_nr(r2);
        double s6 = inv * inv * inv;
        return s6 * fma(A, s6, -B);
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

// Uniform stream (include/mcb200.h, mcb_seed_chains): u_n = (mix(key + (n+1) G) >> 11) * 2^-53.
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
            double dx = wrap_exact(p.com[j] - ncx, g.boxl2, g.boxl);              // R = CoM(J) - CoM(I), :806-810
            double dy = wrap_exact(p.com[g.Np + j] - ncy, g.boxl2, g.boxl);
            double dz = wrap_exact(p.com[2 * g.Np + j] - ncz, g.boxl2, g.boxl);
            double r2 = dx * dx + dy * dy + dz * dz;                              // length_sq, vector_class.f90:107
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
            const double2 *ab = reinterpret_cast<const double2 *>(p.tabAB) + (t * 4 + a) * 4;
            double2 v0 = sp[0], v1 = sp[1], v2 = sp[2], v3 = sp[3], v4 = sp[4], v5 = sp[5];
            double2 p0 = ab[0], p1 = ab[1], p2 = ab[2], p3 = ab[3];
            double q0 = 0, q1 = 0, q2 = 0, q3 = 0;
            if (COUL) { const double *tq = p.tabQ + (t * 4 + a) * 4; q0 = tq[0]; q1 = tq[1]; q2 = tq[2]; q3 = tq[3]; }
            // R = mol2(J) - mol1(I) (LennardJones.f90:69); b = 0..3 inner like the reference's J loop
            double e0 = site_pair<COUL>(v0.x - sx, v0.y - sy, v1.x - sz, p0.x, p0.y, q0, g);
            double e1 = site_pair<COUL>(v1.y - sx, v2.x - sy, v2.y - sz, p1.x, p1.y, q1, g);
            double e2 = site_pair<COUL>(v3.x - sx, v3.y - sy, v4.x - sz, p2.x, p2.y, q2, g);
            double e3 = site_pair<COUL>(v4.y - sx, v5.x - sy, v5.y - sz, p3.x, p3.y, q3, g);
            e = ((e0 + e1) + e2) + e3;
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

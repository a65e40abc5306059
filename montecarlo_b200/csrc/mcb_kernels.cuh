// mcb_kernels.cuh — the kernels behind libmcb200.so.  One warp owns one chain (one copy of the box).
//
//   k_prepare     start-up recompute: site cache, pair matrix, ENERGY(:), TOTENG   (MonteCarlo.F90:357-364)
//   k_trial       energy change of one caller-supplied trial move                  (MonteCarlo.F90:411-414)
//   k_commit      keep or roll back that move                                      (MonteCarlo.F90:779-786)
//   k_run_steps   MCINIT + calculateLJ + calcEnergy + METROPOLIS looped on device  (MonteCarlo.F90:406-433)
#pragma once

#include "mcb_device.cuh"

namespace mcb {

struct SiteRel { double r[NSITE * 3]; int n_sites; int pad; };

constexpr double DBL_HUGE = 1.7976931348623157e308;            // HUGE(TOTENG); HUGE-1 == HUGE in fp64
constexpr double E_F32 = 2.7182817459106445;                   // lib.f90:17, a default-REAL literal
constexpr double LN_E_F32 = 0x1.fffffefb245eap-1;               // log(2.7182817459106445) = 0.99999996963214...
constexpr double PI_D = 3.141592653589793;                     // 4*atan(1), lib.f90:15
constexpr double TAU_D = 6.283185307179586;                    // 2*PI, lib.f90:16
constexpr double INV_SQRT3 = 0x1.279a74590331dp-1;              // 1.D0/sqrt(3.D0), vector_class.f90:77 via lib.f90:118

// ---------------------------------------------------------------------------------------------------------
// Start-up recompute.  Each unordered pair is evaluated once, with the LARGER index as MOL1 — the value the
// reference's double evaluation leaves in SS(i,j) and SS(j,i) (last writer wins, MonteCarlo.F90:358-360,819-820).
constexpr int PREP_W = 8;   // warps per chain in k_prepare: rows of the matrix are dealt out to them

__host__ __device__ inline int prepare_warp_doubles(const Geo &g) { return g.Np + ((g.Np + g.nsolp + 1) / 2 + 1 & ~1); }
// shared memory of a k_prepare block: the warps' row buffers, their partial sums and - when it all fits (N <~ 1000) - a
// copy of the centres of mass, the site cache and the parameter rows, which every row's gate and site pairs read (from global memory the
// kernel spent 10 of its 17 cycles per instruction waiting for those loads: profiles/README.md, round 2)
__host__ __device__ inline int prepare_stage_doubles(const Geo &g) { return 3 * g.Np + SITE_DOUBLES * (g.N + g.nsolp) + (1 + g.nsolp) * 4 * (TAB_ROW + 4); }
__host__ __device__ inline int prepare_smem_doubles(const Geo &g, bool staged)
{
    return PREP_W * prepare_warp_doubles(g) + 3 * PREP_W + 1 + (staged ? prepare_stage_doubles(g) : 0);
}
__host__ inline bool prepare_can_stage(const Geo &g) { return (size_t)prepare_smem_doubles(g, true) * 8 <= 100 * 1024; }

template <bool COUL, bool STAGED>
__global__ void __launch_bounds__(32 * PREP_W) k_prepare(Geo g, SiteRel rel, double *images, double *matg, ChainCtl *ctl, int chain0,
                                                         double *out_tot, double *out_esol, long long *out_pairs)
{
    extern __shared__ __align__(16) double smem[];   // per warp: newrow[Np] + list[Np + nsolp]; then PREP_W x 3 partial sums; then the staged copy
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = chain0 + blockIdx.x;
    double *img = images + (size_t)c * g.image_doubles;
    double *mat = g.mat_in_image ? img + g.o_mat : matg + (size_t)c * g.N * g.MS;
    const int N = g.N, Np = g.Np;
    double *const mine = smem + warp * prepare_warp_doubles(g);
    double *const red = smem + PREP_W * prepare_warp_doubles(g);
    ChainPtrs p;
    p.com = img + g.o_com; p.hoek = img + g.o_hoek; p.energy = img + g.o_energy; p.sites = img + g.o_sites;
    p.tabAB = img + g.o_tabAB; p.tabQ = img + g.o_tabQ; p.mat = mat;
    p.newrow = mine; p.list = reinterpret_cast<int *>(mine + Np);

    for (int j = tid; j < N; j += 32 * PREP_W) {   // RotMatrix for every molecule, lib.f90:27-83
        double s1, c1, s2, c2, s3, c3;
        sincos(p.hoek[j], &s1, &c1); sincos(p.hoek[Np + j], &s2, &c2); sincos(p.hoek[2 * Np + j], &s3, &c3);
        double cx = p.com[j], cy = p.com[Np + j], cz = p.com[2 * Np + j];
        for (int a = 0; a < NSITE; ++a) {
            double ox, oy, oz;
            rotate_site(c1, s1, c2, s2, c3, s3, rel.r[3 * a], rel.r[3 * a + 1], rel.r[3 * a + 2], cx, cy, cz, ox, oy, oz);
            p.sites[SITE_DOUBLES * j + 3 * a] = ox; p.sites[SITE_DOUBLES * j + 3 * a + 1] = oy;
            p.sites[SITE_DOUBLES * j + 3 * a + 2] = oz;
        }
    }
    __syncthreads();
    if (STAGED) {   // the rows below read centres of mass and sites from shared memory
        double *const scom = smem + ((PREP_W * prepare_warp_doubles(g) + 3 * PREP_W + 1) & ~1), *const ssites = scom + 3 * Np;
        for (int q = tid; q < 3 * Np; q += 32 * PREP_W) scom[q] = p.com[q];
        const double2 *src = reinterpret_cast<const double2 *>(p.sites);
        double2 *dst = reinterpret_cast<double2 *>(ssites);
        for (int q = tid; q < SITE_DOUBLES / 2 * (N + g.nsolp); q += 32 * PREP_W) dst[q] = src[q];
        double *const stab = ssites + SITE_DOUBLES * (N + g.nsolp), *const stq = stab + (1 + g.nsolp) * 4 * TAB_ROW;
        for (int q = tid; q < (1 + g.nsolp) * 4 * TAB_ROW; q += 32 * PREP_W) stab[q] = p.tabAB[q];
        for (int q = tid; q < (1 + g.nsolp) * 16; q += 32 * PREP_W) stq[q] = p.tabQ[q];
        p.com = scom; p.sites = ssites; p.tabAB = stab; p.tabQ = stq;
        __syncthreads();
    }

    // row i = molecule i as MOL1 against every j < i: each unordered pair once, the larger index as MOL1
    const int a = lane & 3;
    long long pairs = 0;
    double esum = 0.0, racc = 0.0;
    for (int i = warp; i < N; i += PREP_W) {
        double sx = p.sites[SITE_DOUBLES * i + 3 * a], sy = p.sites[SITE_DOUBLES * i + 3 * a + 1],
               sz = p.sites[SITE_DOUBLES * i + 3 * a + 2];
        double e_sol;
        int nn = row_eval<COUL>(g, p, i, i, p.com[i], p.com[Np + i], p.com[2 * Np + i], sx, sy, sz, lane, e_sol);
        pairs += nn;
        esum += e_sol;
        if (lane == 0) { p.energy[i] = e_sol; mat[(size_t)i * g.MS + i] = 0.0; }
        for (int j = lane; j < i; j += 32) {
            double v = p.newrow[j];
            racc += v;
            mat[(size_t)i * g.MS + j] = v;
            mat[(size_t)j * g.MS + i] = v;
        }
        __syncwarp();
    }
    // calcEnergy (MonteCarlo.F90:988-1004): every pair energy once plus every ENERGY(i), summed as the rows are formed
    // (per lane, then per warp, then over the warps in a fixed order => deterministic; no second pass over the matrix)
    const double acc = warp_sum(racc) + esum;
    if (lane == 0) { red[3 * warp] = acc; red[3 * warp + 1] = esum; red[3 * warp + 2] = (double)pairs; }
    __syncthreads();
    if (tid == 0) {
        double tot = 0.0, es = 0.0, np = 0.0;
        for (int w = 0; w < PREP_W; ++w) { tot += red[3 * w]; es += red[3 * w + 1]; np += red[3 * w + 2]; }
        ctl[c].toteng = tot;
        ctl[c].energies_valid = 1;
        if (out_tot) out_tot[c - chain0] = tot;
        if (out_esol) out_esol[c - chain0] = es;
        if (out_pairs) out_pairs[c - chain0] = (long long)np;
    }
}

// AoS <-> SoA: the driver's TYPE(vector) arrays double[chain][molecule][3] and the images' planes x|y|z (a11)
__global__ void __launch_bounds__(128) k_state_pack(Geo g, double *images, int chain0, const double *com, const double *hoek, int broadcast)
{
    const int c = chain0 + blockIdx.x;
    double *img = images + (size_t)c * g.image_doubles;
    const size_t src = broadcast ? 0 : (size_t)blockIdx.x * g.N * 3;
    for (int q = threadIdx.x; q < 3 * g.Np; q += blockDim.x) {
        const int k = q / g.Np, j = q - k * g.Np;
        img[g.o_com + q] = j < g.N ? com[src + 3 * (size_t)j + k] : 0.0;
        img[g.o_hoek + q] = j < g.N ? hoek[src + 3 * (size_t)j + k] : 0.0;
    }
}

__global__ void __launch_bounds__(128) k_state_unpack(Geo g, const double *images, int chain0, double *com, double *hoek)
{
    const int c = chain0 + blockIdx.x;
    const double *img = images + (size_t)c * g.image_doubles;
    const size_t dst = (size_t)blockIdx.x * g.N * 3;
    for (int q = threadIdx.x; q < 3 * g.N; q += blockDim.x) {
        const int j = q / 3, k = q - 3 * j;
        if (com) com[dst + q] = img[g.o_com + k * g.Np + j];
        if (hoek) hoek[dst + q] = img[g.o_hoek + k * g.Np + j];
    }
}

// ---------------------------------------------------------------------------------------------------------
// Single-step mode.  The trial lives in the chain's scratch area until k_commit.
struct Pending { double nc[3], nh[3], sites[SITE_DOUBLES], e_sol, dE; int i, valid; };

__device__ __forceinline__ Pending *pending_of(const Geo &g, double *work) { return reinterpret_cast<Pending *>(work + g.o_pend); }

template <bool COUL>
__global__ void __launch_bounds__(32) k_trial(Geo g, SiteRel rel, double *images, double *matg, double *work,
                                              ChainCtl *ctl, int chain0, const int *rsolv, const double *new_com,
                                              const double *new_hoek, double *out_dE, double *out_enew, int read_only)
{
    const int lane = threadIdx.x;
    const int c = chain0 + blockIdx.x;
    rsolv -= chain0; new_com -= 3 * chain0; new_hoek -= 3 * chain0; out_dE -= chain0; out_enew -= chain0;
    double *img = images + (size_t)c * g.image_doubles;
    double *mat = g.mat_in_image ? img + g.o_mat : matg + (size_t)c * g.N * g.MS;
    double *wk = work + (size_t)c * g.work_doubles;
    ChainPtrs p = chain_ptrs(g, img, mat, wk);
    Pending *pd = pending_of(g, wk);
    const int i = rsolv[c] - 1;   // RSOLV is 1-based
    if (i < 0 || i >= g.N) { if (lane == 0) { pd->valid = 0; out_dE[c] = nan(""); out_enew[c] = nan(""); } return; }
    double ncx = new_com[3 * c], ncy = new_com[3 * c + 1], ncz = new_com[3 * c + 2];
    double h1 = new_hoek[3 * c], h2 = new_hoek[3 * c + 1], h3 = new_hoek[3 * c + 2];
    double s1, c1, s2, c2, s3, c3;
    sincos(h1, &s1, &c1); sincos(h2, &s2, &c2); sincos(h3, &s3, &c3);
    const int a = lane & 3;
    double sx, sy, sz;
    rotate_site(c1, s1, c2, s2, c3, s3, rel.r[3 * a], rel.r[3 * a + 1], rel.r[3 * a + 2], ncx, ncy, ncz, sx, sy, sz);
    double e_sol;
    int nn = row_eval<COUL>(g, p, i, g.N, ncx, ncy, ncz, sx, sy, sz, lane, e_sol);
    double d = 0.0;
    for (int j = lane; j < g.N; j += 32) d += p.newrow[j] - mat[(size_t)i * g.MS + j];
    double dE = warp_sum(d) + (e_sol - p.energy[i]);
    if (lane < 4) { pd->sites[3 * lane] = sx; pd->sites[3 * lane + 1] = sy; pd->sites[3 * lane + 2] = sz; }
    if (lane == 0) {
        pd->nc[0] = ncx; pd->nc[1] = ncy; pd->nc[2] = ncz; pd->nh[0] = h1; pd->nh[1] = h2; pd->nh[2] = h3;
        pd->e_sol = e_sol; pd->dE = dE; pd->i = i; pd->valid = read_only ? 0 : 1;
        out_dE[c] = dE; out_enew[c] = ctl[c].toteng + dE;
        if (!read_only) ctl[c].sitepairs += (long long)rel.n_sites * (rel.n_sites * nn + g.nsol);
    }
}

__global__ void __launch_bounds__(32) k_commit(Geo g, double *images, double *matg, double *work, ChainCtl *ctl,
                                               const unsigned char *accept)
{
    const int lane = threadIdx.x;
    const int c = blockIdx.x;
    double *img = images + (size_t)c * g.image_doubles;
    double *mat = g.mat_in_image ? img + g.o_mat : matg + (size_t)c * g.N * g.MS;
    double *wk = work + (size_t)c * g.work_doubles;
    ChainPtrs p = chain_ptrs(g, img, mat, wk);
    Pending *pd = pending_of(g, wk);
    if (!pd->valid) return;
    const int i = pd->i;
    if (accept[c]) {
        for (int j = lane; j < g.N; j += 32) {
            double v = p.newrow[j];
            mat[(size_t)i * g.MS + j] = v;
            mat[(size_t)j * g.MS + i] = v;
        }
        if (lane < 3) { p.com[lane * g.Np + i] = pd->nc[lane]; p.hoek[lane * g.Np + i] = pd->nh[lane]; }
        if (lane < SITE_DOUBLES) p.sites[SITE_DOUBLES * i + lane] = pd->sites[lane];
        if (lane == 0) { p.energy[i] = pd->e_sol; ctl[c].toteng += pd->dE; }
    }
    __syncwarp();
    if (lane == 0) pd->valid = 0;
}

#ifdef MCB_PHASE_CLOCKS
// Debug build only (-DMCB_PHASE_CLOCKS): cycles per phase of k_run_steps, summed over warps.
__device__ unsigned long long g_phase[8];
#define PH_T(k) do { long long now_ = clock64(); ph[k] += now_ - tph; tph = now_; } while (0)
#else
#define PH_T(k) do { } while (0)
#endif

// x - period*AINT(x/half) (MonteCarlo.F90:705-712) without a division and without a branch.  q0 = x*(1/half) is
// within an ulp of the quotient; the residual x - q0*half is exact in one fma, and q0 + r*(1/half) then rounds to the
// correctly rounded x/half (Markstein's division step; 1/half is the correctly rounded reciprocal, formed by a true
// division on the host or at compile time), so AINT sees exactly the reference's quotient.  A guarded true division
// (a function call) in the middle of the step's straight-line code cost 1.4 % of the step.
// tests/test_abi_cpu.py::test_wrap_fma_matches_division checks the identity on the host, adversarial cases included.
__device__ __forceinline__ double wrap_fma(double x, double period, double half, double inv_half)
{
    const double q0 = x * inv_half;
    const double r = fma(-q0, half, x);
    const double q = fma(r, inv_half, q0);
    return fma(-period, trunc(q), x);
}

// One (partner, own site a) item: own site (sx,sy,sz) against the four sites of virtual molecule vm with the
// (A,B[,QQ]) row of type t.  R = mol2(J) - mol1(I) (LennardJones.f90:69), b inner like the reference's J loop.
template <bool COUL>
__device__ __forceinline__ double item_energy(const Geo &g, const double *sites, const double *tabAB, const double *tabQ,
                                              int vm, int t, int a, double sx, double sy, double sz)
{
    const double2 *sp = reinterpret_cast<const double2 *>(sites + SITE_DOUBLES * vm);
    const double2 *ab = reinterpret_cast<const double2 *>(tabAB) + (t * 4 + a) * (TAB_ROW / 2);
    double2 v0 = sp[0], v1 = sp[1], v2 = sp[2], v3 = sp[3], v4 = sp[4], v5 = sp[5];
    double2 p0 = ab[0], p1 = ab[1], p2 = ab[2], p3 = ab[3];
    double q0 = 0, q1 = 0, q2 = 0, q3 = 0;
    if (COUL) { const double *tq = tabQ + (t * 4 + a) * 4; q0 = tq[0]; q1 = tq[1]; q2 = tq[2]; q3 = tq[3]; }
    double ea = site_pair_acc<COUL>(0.0, v0.x - sx, v0.y - sy, v1.x - sz, p0.x, p0.y, q0, g);
    double eb = site_pair_acc<COUL>(0.0, v1.y - sx, v2.x - sy, v2.y - sz, p1.x, p1.y, q1, g);
    ea = site_pair_acc<COUL>(ea, v3.x - sx, v3.y - sy, v4.x - sz, p2.x, p2.y, q2, g);
    eb = site_pair_acc<COUL>(eb, v4.y - sx, v5.x - sy, v5.y - sz, p3.x, p3.y, q3, g);
    return ea + eb;
}

// NI (partner, own site) items = 4*NI site pairs evaluated STAGE BY STAGE (all differences, all minimum images,
// all r^2, all reciprocals, ...) so that one warp keeps 4*NI independent fp64 chains in flight: a lone warp gets
// 126 cycles per site pair with the pairs written one after the other, 87 with 8 and 63 with 16 staged
// (profiles/pairbench_r01b.txt), and with ~1.75 warps per SM sub-partition that is what decides the step time.
// Items past the end are computed on slot 0 and dropped.
template <bool COUL, int NI>
__device__ __forceinline__ double items_staged(const Geo &g, const double *sites, const double *tabAB, const double *tabQ,
                                               const int *list, int nsolp, int nslots, int s0, int stride, int a,
                                               double sx, double sy, double sz, double *epart)
{
    constexpr int NP = 4 * NI;
    constexpr double M = 4503599627370496.0;   // 2^52
    double ax[NP], ay[NP], az[NP];
    int slot[NI];
    bool valid[NI];
#pragma unroll
    for (int n = 0; n < NI; ++n) {
        const int sl = s0 + n * stride;
        valid[n] = sl < nslots;
        slot[n] = valid[n] ? sl : 0;   // slot 0 exists whenever a pass runs
        const double2 *sp = reinterpret_cast<const double2 *>(sites + SITE_DOUBLES * list[slot[n]]);
        const double2 v0 = sp[0], v1 = sp[1], v2 = sp[2], v3 = sp[3], v4 = sp[4], v5 = sp[5];
        // R = mol2(J) - mol1(I) (LennardJones.f90:69); only |R| is needed
        ax[4 * n + 0] = fabs(v0.x - sx); ay[4 * n + 0] = fabs(v0.y - sy); az[4 * n + 0] = fabs(v1.x - sz);
        ax[4 * n + 1] = fabs(v1.y - sx); ay[4 * n + 1] = fabs(v2.x - sy); az[4 * n + 1] = fabs(v2.y - sz);
        ax[4 * n + 2] = fabs(v3.x - sx); ay[4 * n + 2] = fabs(v3.y - sy); az[4 * n + 2] = fabs(v4.x - sz);
        ax[4 * n + 3] = fabs(v4.y - sx); ay[4 * n + 3] = fabs(v5.x - sy); az[4 * n + 3] = fabs(v5.y - sz);
    }
    double tx[NP], ty[NP], tz[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) { tx[q] = __fma_rd(ax[q], g.inv_boxl, M); ty[q] = __fma_rd(ay[q], g.inv_boxl, M); tz[q] = __fma_rd(az[q], g.inv_boxl, M); }
#pragma unroll
    for (int q = 0; q < NP; ++q) { tx[q] -= M; ty[q] -= M; tz[q] -= M; }
#pragma unroll
    for (int q = 0; q < NP; ++q) { ax[q] = fma(-g.boxl2, tx[q], ax[q]); ay[q] = fma(-g.boxl2, ty[q], ay[q]); az[q] = fma(-g.boxl2, tz[q], az[q]); }
    double r2[NP], y[NP], e[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) r2[q] = ax[q] * ax[q];
#pragma unroll
    for (int q = 0; q < NP; ++q) r2[q] = fma(ay[q], ay[q], r2[q]);
#pragma unroll
    for (int q = 0; q < NP; ++q) r2[q] = fma(az[q], az[q], r2[q]);
    if (COUL) {
#pragma unroll
        for (int q = 0; q < NP; ++q) { y[q] = rsqrt(r2[q]); }
    } else {
#pragma unroll
        for (int q = 0; q < NP; ++q) asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y[q]) : "d"(r2[q]));
#pragma unroll
        for (int q = 0; q < NP; ++q) e[q] = fma(-r2[q], y[q], 1.0);
#pragma unroll
        for (int q = 0; q < NP; ++q) e[q] = fma(e[q], e[q], e[q]);
#pragma unroll
        for (int q = 0; q < NP; ++q) y[q] = fma(y[q], e[q], y[q]);
    }
    double tot = 0.0;
#pragma unroll
    for (int n = 0; n < NI; ++n) {
        const int t = slot[n] < nsolp ? 1 + slot[n] : 0;
        const double2 *ab = reinterpret_cast<const double2 *>(tabAB) + (t * 4 + a) * (TAB_ROW / 2);
        double en = 0.0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int q = 4 * n + b;
            const double2 pb = ab[b];
            if (COUL) {
                const double inv = y[q] * y[q], s6 = inv * inv * inv;
                en = fma(s6, fma(pb.x, s6, -pb.y), fma(tabQ[(t * 4 + a) * 4 + b], y[q], en));
            } else {
                const double s6 = (y[q] * y[q]) * y[q];
                en = fma(s6, fma(pb.x, s6, -pb.y), en);
            }
        }
        if (valid[n]) { epart[4 * slot[n] + a] = en; tot += en; }
    }
    return tot;
}

// ---------------------------------------------------------------------------------------------------------
// Batched mode with DRIVER-SUPPLIED uniforms: nsteps Metropolis steps of one chain per warp, the chain's image
// resident in shared memory, the uniforms read from the caller's buffer [chain][cap] in exactly the order the
// reference draws them (7 per step, an 8th only when METROPOLIS needs it), so a buffer of the driver's RAND()
// values reproduces the driver's accept/reject sequence.  (k_run_spec below is the fast path for the built-in
// stream; this kernel keeps the straightforward one-step-at-a-time structure.)
//   MAT_SMEM  the pair matrix is part of the staged image (small N); otherwise it stays in global memory
template <bool COUL, bool MAT_SMEM>
__global__ void __launch_bounds__(32) k_run_steps(Geo g, SiteRel rel, double *images, double *matg, ChainCtl *ctl,
                                                  long long nsteps, const double *ubuf, long long cap,
                                                  long long *out_consumed, long long *out_steps)
{
    extern __shared__ __align__(16) double smem[];
    const int lane = threadIdx.x;
    const int c = blockIdx.x;
    const int N = g.N, Np = g.Np, MS = g.MS, nsolp = g.nsolp;
    double *gimg = images + (size_t)c * g.image_doubles;
    double *img = smem;
    {   // stage the image: linear, coalesced 16-byte loads
        const double2 *src = reinterpret_cast<const double2 *>(gimg);
        double2 *dst = reinterpret_cast<double2 *>(img);
        for (int k = lane; k < g.stage_doubles / 2; k += 32) dst[k] = src[k];
    }
    double *mat = MAT_SMEM ? img + g.o_mat : matg + (size_t)c * N * MS;
    double *const com = img + g.o_com, *const hoek = img + g.o_hoek, *const energy = img + g.o_energy;
    double *const sites = img + g.o_sites;
    const double *const tabAB = img + g.o_tabAB, *const tabQ = img + g.o_tabQ;
    double *const wk = smem + g.stage_doubles;
    int *const list = reinterpret_cast<int *>(wk + g.o_list);      // [nsolp + N]: virtual molecule of each slot
    double *const epart = wk + g.o_epart;                          // [nsolp + N][4]: per (slot, own site) energy
    ChainCtl s = ctl[c];
    for (int p = lane; p < nsolp; p += 32) list[p] = N + p;        // the solute pseudo-molecules are slots 0..nsolp-1
    __syncwarp();

    // re-anchor TOTENG on the stored matrix so that rounding of the running sum cannot accumulate across launches
    {
        double acc = 0.0;
        for (int i = 0; i < N; ++i)
            for (int j = i + 1 + lane; j < N; j += 32) acc += mat[(size_t)i * MS + j];
        for (int i = lane; i < N; i += 32) acc += energy[i];
        s.toteng = warp_sum(acc);
    }

    const int a = lane & 3, q = lane >> 2;
    const unsigned lt = (1u << lane) - 1u;
    const double relx = rel.r[3 * a], rely = rel.r[3 * a + 1], relz = rel.r[3 * a + 2];
    const int wpair = rel.n_sites * rel.n_sites, wsol = rel.n_sites * g.nsol;
    const double neg_inv_rt = -1000.0 / (8.314 * s.temperature);   // EXPONENT = -DELTA*1000/(8.314*T), MonteCarlo.F90:729
    const int r = min(max(lane - 1, 0), 5);
    const bool isang = r >= 3;
    const int k = isang ? r - 3 : r;
    double *const base = (isang ? hoek : com) + k * Np;
    const double period = isang ? TAU_D : g.boxl2, half = isang ? PI_D : g.boxl;
    const double inv_half = isang ? 1.0 / PI_D : g.inv_boxl;
    const double f1 = isang ? 1.0 : INV_SQRT3;
    int adj_count = (int)(s.step % s.nadj);
    long long cursor = 0;
    int done = 0;
    unsigned nsuc = 0, nacc = (unsigned)s.naccept, racc = 0, nclamp = 0, uused = 0, prod_acc = 0, nstat = 0;
    unsigned long long sitepairs = 0;
    const unsigned long long step0 = (unsigned long long)s.step;
    const double *myubuf = ubuf + (size_t)c * cap;
    const int nsteps_i = (int)min(nsteps, (long long)0x7fffffff);
#ifdef MCB_PHASE_CLOCKS
    long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tph = clock64();
#endif

    for (; done < nsteps_i; ++done) {
        if (cursor + 8 > cap) break;
        const double myu = lane < 8 ? myubuf[cursor + lane] : 0.0;
        // ---- MCINIT (MonteCarlo.F90:685-715): lane 0 picks the molecule, lanes 1-3 move the centre of mass,
        //      lanes 4-6 the angles (randVec / randVecHoek, lib.f90:111-131), then the wraps (:705-712)
        const double u0 = shfl_d(myu, 0);
        int i = (int)(u0 * (double)N) + 1;
        if (g.quirk31 && i == 31) i = 30;                         // :698
        i = min(i, N) - 1;
        // old row of the pair matrix + ENERGY(i): needed only for dE, fetched now to hide the latency
        double old_part = lane == 0 ? energy[i] : 0.0;
        for (int j = lane; j < N; j += 32) old_part += mat[(size_t)i * MS + j];
        double rv = (myu - 0.5) * 2.0;
        rv = f1 * rv;
        rv = (isang ? s.dhoekmax : s.dposmax) * rv;
        const double nv = wrap_fma(base[i] + rv, period, half, inv_half);
        double sn, cs;
        sincos(nv, &sn, &cs);
        const double ncx = shfl_d(nv, 1), ncy = shfl_d(nv, 2), ncz = shfl_d(nv, 3);
        const double c1 = shfl_d(cs, 4), s1 = shfl_d(sn, 4), c2 = shfl_d(cs, 5), s2 = shfl_d(sn, 5),
                     c3 = shfl_d(cs, 6), s3 = shfl_d(sn, 6);
        double sx, sy, sz;
        rotate_site(c1, s1, c2, s2, c3, s3, relx, rely, relz, ncx, ncy, ncz, sx, sy, sz);
        PH_T(0);

        // ---- calculateLJ(RSOLV) (:792-828): centre-of-mass gate, one partner per lane (:804-813)
        int nn = 0;
        for (int j0 = 0; j0 < N; j0 += 32) {
            const int j = j0 + lane;
            bool in = false;
            if (j < N && j != i) {
                in = !(pair_r2(com[j] - ncx, com[Np + j] - ncy, com[2 * Np + j] - ncz, g) > g.cut2);
            }
            const unsigned m = __ballot_sync(FULL, in);
            if (in) list[nsolp + nn + __popc(m & lt)] = j;
            nn += __popc(m);
        }
        __syncwarp();
        PH_T(1);
        // ---- site pairs: 16 slots per pass, two (slot, own site) items per lane (calcLJ / CalcLJ_SV)
        const int nslots = nsolp + nn;
        double lane_tot = 0.0;
        for (int s0 = 0; s0 < nslots; s0 += 16) {
            const int sA = s0 + q, sB = s0 + 8 + q;
            double eA = 0.0, eB = 0.0;
            if (sA < nslots) eA = item_energy<COUL>(g, sites, tabAB, tabQ, list[sA], sA < nsolp ? 1 + sA : 0, a, sx, sy, sz);
            if (sB < nslots) eB = item_energy<COUL>(g, sites, tabAB, tabQ, list[sB], sB < nsolp ? 1 + sB : 0, a, sx, sy, sz);
            if (sA < nslots) epart[4 * sA + a] = eA;
            if (sB < nslots) epart[4 * sB + a] = eB;
            lane_tot += eA + eB;
        }
        PH_T(2);
        // ---- calcEnergy as a difference (:988-1004): TOTENG_new - TOTENG_old = sum(new row) - sum(old row)
        const double dE = warp_sum(lane_tot - old_part);
        double tnew = s.toteng + dE;
        if (!(tnew <= DBL_HUGE)) { tnew = DBL_HUGE; nclamp++; }    // Inf or NaN, :421-428

        // ---- METROPOLIS (:720-788)
        const double delta = tnew - s.toteng;
        const double exponent = delta * neg_inv_rt;
        bool accept;
        int used = 7;
        if (exponent < -75.0) accept = false;                       // KANS = 0, RV = 1, no draw
        else if (exponent >= 0.0) accept = true;                    // KANS = 1, RV = 0, no draw
        else { accept = shfl_d(myu, 7) <= fmin(exp(exponent * LN_E_F32), 1.0); used = 8; }
        PH_T(3);
        uused += used; s.last_delta = delta;
        sitepairs += (unsigned)(wpair * nn + wsol);
        cursor += used;
        const bool in_prod = (long long)(step0 + (unsigned)done + 1ULL) >= s.start_prod;
        if (accept) {
            nsuc++; nacc++;
            if (in_prod) { s.production += tnew; prod_acc++; }
            for (int j = lane; j < N; j += 32)
                if (MAT_SMEM || mat[(size_t)i * MS + j] != 0.0) { mat[(size_t)i * MS + j] = 0.0; mat[(size_t)j * MS + i] = 0.0; }
            if (lane >= 1 && lane <= 6) base[i] = nv;
            if (lane < 4) {
                sites[SITE_DOUBLES * i + 3 * a] = sx; sites[SITE_DOUBLES * i + 3 * a + 1] = sy;
                sites[SITE_DOUBLES * i + 3 * a + 2] = sz;
            }
            if (lane == 0) {   // ENERGY(i) = CalcLJ_SV: the solute slots
                double es = 0.0;
                for (int p = 0; p < nsolp; ++p) es += (epart[4 * p] + epart[4 * p + 1]) + (epart[4 * p + 2] + epart[4 * p + 3]);
                energy[i] = es;
            }
            __syncwarp();
            for (int sl = nsolp + lane; sl < nslots; sl += 32) {   // SS(i,j) = SS(j,i) = EN, :819-820
                const int j = list[sl];
                const double e = (epart[4 * sl] + epart[4 * sl + 1]) + (epart[4 * sl + 2] + epart[4 * sl + 3]);
                mat[(size_t)i * MS + j] = e; mat[(size_t)j * MS + i] = e;
            }
            s.toteng = tnew;
        }
        if (in_prod) { s.sum_e += s.toteng; s.sum_e2 = fma(s.toteng, s.toteng, s.sum_e2); nstat++; }
        racc = nacc;
        if (++adj_count == s.nadj) {                                // mod(I, NADJ) == 0, :766-777
            adj_count = 0;
            const double ratio = (double)((float)nacc / (float)s.nadj);   // default-REAL quotient, :758
            if (ratio > 0.5) { s.dposmax *= (1.0 + s.padj); s.dhoekmax *= (1.0 + s.padj); }
            else if (ratio < 0.5) { s.dposmax *= (1.0 - s.padj); s.dhoekmax *= (1.0 - s.padj); }
            s.dposmax = fmax(s.dposmax, s.dposmin); s.dhoekmax = fmax(s.dhoekmax, s.dhoekmin);
            nacc = 0;
        }
        __syncwarp();
        PH_T(4);
    }
#ifdef MCB_PHASE_CLOCKS
    if (lane == 0) for (int qq = 0; qq < 8; ++qq) atomicAdd(&g_phase[qq], (unsigned long long)ph[qq]);
#endif

    // write back what can have changed: com, hoek, energy, solvent sites (contiguous) and the matrix
    {
        const int n_mut = g.o_sites + SITE_DOUBLES * N;
        for (int qq = lane; qq < n_mut; qq += 32) gimg[qq] = img[qq];
        if (MAT_SMEM) for (int qq = lane; qq < N * MS; qq += 32) gimg[g.o_mat + qq] = img[g.o_mat + qq];
    }
    if (lane == 0) {
        s.step += done; s.rng_n += (unsigned long long)done; s.nsuc += nsuc; s.naccept = nacc; s.nclamp += nclamp; s.uused += uused; s.prod_acc += prod_acc;
        s.nstat += nstat; s.sitepairs += (long long)sitepairs;
        // REAL(NACCEPT)/REAL(NADJ) as METROPOLIS leaves it after its last step (:758): before the window reset
        if (done > 0) s.ratio = (double)((float)racc / (float)s.nadj);
        ctl[c] = s;
        if (out_consumed) out_consumed[c] = cursor;
        if (out_steps) out_steps[c] = done;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Batched mode with speculative look-ahead (built-in uniform stream only).
//
// One step of one chain is a long chain of dependent operations (~3500 cycles) around only ~230 site pairs, and
// BASELINE's configuration fixes the number of chains, so a GPU full of one-warp-per-chain steps is latency-
// bound at ~12 % of the fp64 pipe.  Here a BLOCK of W warps works on T = W*32/GL CONSECUTIVE steps of its chain
// at once, GL lanes per step, all evaluated against the same committed state.  The stream is random-access
// (step s owns u(8s..8s+7)), so the trial move of step k+g does not depend on the outcome of the steps before
// it.  After the evaluation the steps are resolved in order: step k+g stands if every ACCEPTED earlier step of
// the batch moved a molecule that step k+g never looked at — not the same molecule, not inside the cutoff of the
// trial position (before or after its own move) and not a non-zero entry of the old matrix row (before or
// after).  The first step that fails the test, and everything behind it, is thrown away and re-evaluated in the
// next batch with the same uniforms; a batch also ends at a step-size adaptation.  The resulting Markov chain
// is the sequential one, decision for decision.  Only committed steps count as work.
// The ring of pre-drawn trial moves (k_run_spec): entry e = step & 63 holds the molecule pick, the six unit displacements
// of randVec / randVecHoek before the step size (lib.f90:111-131), METROPOLIS' draw as log(u7)/log(E) - RV <= E**x with
// x < 0 is log(RV)/log(E) <= x, and the left side needs only the uniform - and (sin, cos) of the three angle steps
// dhoekmax * rv.  Both routines are deliberately NOT inlined: they run once per 32 steps / once per adaptation window,
// and keeping their ~400 instructions out of the step loop is worth 3-5 % of the step (profiles/README.md, round 2).
struct RingPtrs { double *thr, *rv; int *pick; double2 *sc; };

__device__ __noinline__ void ring_draw(RingPtrs rg, unsigned long long key, unsigned long long sidx, int N, int quirk31, double dhoekmax)
{
    const int e = (int)(sidx & 63ULL);
    const double u0 = stream_uniform(key, 8ULL * sidx);
    int ii = (int)(u0 * (double)N) + 1;                     // MCINIT, MonteCarlo.F90:697-698
    if (quirk31 && ii == 31) ii = 30;
    rg.pick[e] = min(ii, N) - 1;
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        const double rvu = (stream_uniform(key, 8ULL * sidx + 1 + q) - 0.5) * 2.0;
        rg.rv[6 * e + q] = (q < 3 ? INV_SQRT3 : 1.0) * rvu;
    }
    rg.thr[e] = log(stream_uniform(key, 8ULL * sidx + 7)) * (1.0 / LN_E_F32);
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        double sd, cd;
        sincos_any(dhoekmax * rg.rv[6 * e + 3 + q], sd, cd);
        rg.sc[3 * e + q] = make_double2(sd, cd);
    }
}

// after a step-size adaptation: the pre-drawn steps [first, last) of the launch get the sin/cos of their angle steps at the new size
__device__ __noinline__ void ring_rescale(RingPtrs rg, int ring0, int first, int last, int lane, double dhoekmax)
{
    for (int st = first + lane; st < last; st += 32) {
        const int e = (ring0 + st) & 63;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            double sd, cd;
            sincos_any(dhoekmax * rg.rv[6 * e + 3 + q], sd, cd);
            rg.sc[3 * e + q] = make_double2(sd, cd);
        }
    }
}

struct SpecX { double dE, ncx, ncy, ncz; int i, nn, flags, pad; };   // what a step hands to the resolution
constexpr int SPECX_DOUBLES = sizeof(SpecX) / 8;

// With W > 1 the block must still fit 7 times on an SM (1024 chains / 148 SMs): without the minimum-blocks hint the
// kernel takes ~155 registers, 64-thread blocks fit only 6 times and the 1024 chains need a second wave.
template <bool COUL, bool MAT_SMEM, int GL, int W, bool SMALL>
__global__ void __launch_bounds__(32 * W, W == 1 ? 1 : 7) k_run_spec(Geo g, SiteRel rel, double *images, double *matg, ChainCtl *ctl,
                                                     long long nsteps, long long *out_steps)
{
    constexpr int G = 32 / GL;          // steps per warp
    constexpr int T = W * G;            // steps in flight per chain
    constexpr int NSUB = GL / 4;        // lanes of a group that share one own site
    constexpr int MCH = 32 / GL;        // molecules per lane per 32-molecule chunk
    constexpr unsigned GMASK = GL == 32 ? 0xffffffffu : ((1u << GL) - 1u);
    static_assert(GL == 8 || GL == 16 || GL == 32, "a step needs 7 role lanes");
    static_assert(T <= 8, "the resolution handles at most 8 steps (28 ordered pairs, one per lane)");
    // SMALL: N <= 32, i.e. one 32-molecule chunk: the chunk loops below unroll into straight-line code
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x;
    const int N = g.N, Np = g.Np, MS = g.MS, nsolp = g.nsolp;
    const int nchunk = SMALL ? 1 : (N + 31) >> 5;
    double *gimg = images + (size_t)c * g.image_doubles;
    double *img = smem;
    {
        const double2 *src = reinterpret_cast<const double2 *>(gimg);
        double2 *dst = reinterpret_cast<double2 *>(img);
        for (int k = tid; k < g.stage_doubles / 2; k += 32 * W) dst[k] = src[k];
    }
    double *mat = MAT_SMEM ? img + g.o_mat : matg + (size_t)c * N * MS;
    double *const com = img + g.o_com, *const hoek = img + g.o_hoek, *const energy = img + g.o_energy;
    double *const sites = img + g.o_sites;
    const double *const tabAB = img + g.o_tabAB, *const tabQ = img + g.o_tabQ;
    const int grp = lane / GL, r = lane % GL, leader = grp * GL;
    const int gidx = warp * G + grp;    // which step of the batch this group evaluates
    const int a = r & 3, sub = r >> 2;
    double *const wk0 = smem + g.stage_doubles;
    double *const wk = wk0 + (size_t)gidx * g.spec_group_doubles;
    int *const list = reinterpret_cast<int *>(wk + g.o_list);
    double *const epart = wk + g.o_epart;
    unsigned *const touch = reinterpret_cast<unsigned *>(wk + g.o_touch);
    SpecX *const xch = reinterpret_cast<SpecX *>(wk0 + (size_t)T * g.spec_group_doubles);
    // ring of pre-drawn trial moves: 64 entries indexed by (step & 63), refilled 32 steps at a time with one
    // step per lane, so the uniforms, the molecule pick and log(u7) cost ~8 instructions per step instead of ~60
    double *const ring_thr = reinterpret_cast<double *>(xch) + (size_t)T * SPECX_DOUBLES;
    double *const ring_rv = ring_thr + 64;
    int *const ring_i = reinterpret_cast<int *>(ring_rv + 6 * 64);
    // tot[j] = sum_i SS(i,j) + ENERGY(j): all that dE needs about the old state of molecule j.  Kept current on every
    // accepted move, so the evaluation of a step never reads a matrix row (for large boxes that row lives in global
    // memory) and never has to split the trial energy into its solvent and solute parts: ENERGY(:) itself is
    // recomputed once, at the end of the launch, for the molecules that moved.
    double *const tot = ring_rv + 6 * 64 + 32;
    // sc[j] = (sin, cos) of molecule j's three Euler angles, ring_sc[e] = (sin, cos) of step e's three angle steps: the
    // trial RotMatrix (lib.f90:52-57) then needs no sincos and no shuffle, only the angle-addition formulas
    double2 *const sc = reinterpret_cast<double2 *>(tot + Np);                 // [N][3]
    double2 *const ring_sc = sc + 3 * Np;                                        // [64][3]
    unsigned *const movedw = reinterpret_cast<unsigned *>(ring_sc + 3 * 64);    // [nchunk]: molecules whose ENERGY is stale
    const RingPtrs rg = {ring_thr, ring_rv, ring_i, ring_sc};
    ChainCtl s = ctl[c];
    for (int p = r; p < nsolp; p += GL) list[p] = N + p;
    __syncthreads();
    // column sums (= row sums, the matrix is symmetric; consecutive lanes read consecutive addresses), then TOTENG
    // re-anchored on them so that rounding of the running sums cannot accumulate across launches
    for (int j = tid; j < N; j += 32 * W) {
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int i = 0;
        for (; i + 3 < N; i += 4) {
            a0 += mat[(size_t)i * MS + j]; a1 += mat[(size_t)(i + 1) * MS + j];
            a2 += mat[(size_t)(i + 2) * MS + j]; a3 += mat[(size_t)(i + 3) * MS + j];
        }
        for (; i < N; ++i) a0 += mat[(size_t)i * MS + j];
        tot[j] = (a0 + a1) + (a2 + a3);
        double sj, cj;
#pragma unroll
        for (int q = 0; q < 3; ++q) { sincos_any(hoek[q * Np + j], sj, cj); sc[3 * j + q] = make_double2(sj, cj); }
    }
    for (int q = tid; q < nchunk; q += 32 * W) movedw[q] = 0u;
    __syncthreads();
    {
        double acc = 0.0;
        for (int i = lane; i < N; i += 32) acc += fma(0.5, tot[i], energy[i]);
        s.toteng = warp_sum(acc);
    }
    __syncthreads();
    for (int j = tid; j < N; j += 32 * W) tot[j] += energy[j];
    __syncthreads();
    const double relx = rel.r[3 * a], rely = rel.r[3 * a + 1], relz = rel.r[3 * a + 2];
    const int wpair = rel.n_sites * rel.n_sites, wsol = rel.n_sites * g.nsol;
    const double neg_inv_rt = -1000.0 / (8.314 * s.temperature);
    const int rr = min(max(r - 1, 0), 5);
    const bool isang = rr >= 3;
    const int k = isang ? rr - 3 : rr;
    double *const base = (isang ? hoek : com) + k * Np;
    const double period = isang ? TAU_D : g.boxl2, half = isang ? PI_D : g.boxl;
    const double inv_half = isang ? 1.0 / PI_D : g.inv_boxl;
    // which ordered pair (ga < gb) of steps this lane checks in the resolution
    int gb = 1;
    while (gb * (gb + 1) / 2 <= lane) ++gb;
    const int ga = lane - gb * (gb - 1) / 2;
    int adj_count = (int)(s.step % s.nadj);
    int done = 0;
    unsigned nsuc = 0, nacc = (unsigned)s.naccept, racc = 0, nclamp = 0, n_used8 = 0, prod_acc = 0;
    unsigned long long sum_nn = 0;
    const unsigned long long step0 = (unsigned long long)s.step;
    // position in the uniform stream: steps since mcb_seed_chains (NOT the statistics counter `step`, which
    // mcb_reset_stats zeroes - the stream must go on where it was)
    const unsigned long long pos0 = s.rng_n;
    const int nsteps_i = (int)min(nsteps, (long long)0x3fffffff);
    int ring_hi = 0;                                                // the ring holds the launch-local steps [ring_hi - 64, ring_hi)
    const int ring0 = (int)(pos0 & 63ULL);
    unsigned moved0 = 0u;                                           // SMALL: the molecules whose ENERGY is stale (a register; movedw otherwise)
    // launch-local index of the first production step: step0 + k + 1 >= start_prod (MonteCarlo.F90:760)
    const long long pf = s.start_prod - (long long)step0 - 1;
    const int prod_from = pf <= 0 ? 0 : (pf > 0x3fffffff ? 0x7fffffff : (int)pf);
#ifdef MCB_PHASE_CLOCKS
    long long ph[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tph = clock64();
#endif

    while (done < nsteps_i) {
        // ================= pre-draw the next 32 steps when the ring runs low (one step per lane) =================
        while (done + T > ring_hi) {                                // (block-uniform condition; with W > 1 warp 0 fills for everybody)
            if (W == 1 || warp == 0) ring_draw(rg, s.rng_key, pos0 + (unsigned)(ring_hi + lane), N, g.quirk31, s.dhoekmax);
            ring_hi += 32;
            if (W > 1) __syncthreads(); else __syncwarp();
        }
        // ================= evaluation of step (done + gidx) by this group =================
        const int ring_e = (ring0 + done + gidx) & 63;
        const int i = ring_i[ring_e];
        const double thr = ring_thr[ring_e];
        // sum of the old matrix row + ENERGY(i), straight from the running row sums
        const double old_tot = tot[i];
        const double old_part = r == 0 ? old_tot : 0.0;
        // the old matrix row itself is only needed if the move is accepted (row-sum bookkeeping); for up to 8 chunks
        // it is fetched now so that a global-memory row has arrived by then
        constexpr int PF = SMALL ? 1 : 8;
        double oldrow[PF * MCH];
        const bool prefetched = nchunk <= PF;
        if (prefetched) {
#pragma unroll
            for (int q = 0; q < PF; ++q)
#pragma unroll
                for (int m = 0; m < MCH; ++m) {
                    const int j = 32 * q + r + GL * m;
                    oldrow[q * MCH + m] = (q < nchunk && j < N) ? mat[(size_t)i * MS + j] : 0.0;
                }
        }
        const double rv = (isang ? s.dhoekmax : s.dposmax) * ring_rv[6 * ring_e + rr];
        const double nv = wrap_fma(base[i] + rv, period, half, inv_half);       // :699-712
        const double ncx = shfl_d(nv, leader + 1), ncy = shfl_d(nv, leader + 2), ncz = shfl_d(nv, leader + 3);
        // sin and cos of the trial angles h + dh from those of h (cached per molecule) and of dh (pre-drawn): every lane
        // forms all six itself.  (h + dh and its wrap, :700-707, are still formed exactly, above, for the stored angle.)
        const double2 a1 = sc[3 * i], a2 = sc[3 * i + 1], a3 = sc[3 * i + 2];
        const double2 d1 = ring_sc[3 * ring_e], d2 = ring_sc[3 * ring_e + 1], d3 = ring_sc[3 * ring_e + 2];
        const double s1 = fma(a1.x, d1.y, a1.y * d1.x), c1 = fma(a1.y, d1.y, -(a1.x * d1.x));
        const double s2 = fma(a2.x, d2.y, a2.y * d2.x), c2 = fma(a2.y, d2.y, -(a2.x * d2.x));
        const double s3 = fma(a3.x, d3.y, a3.y * d3.x), c3 = fma(a3.y, d3.y, -(a3.x * d3.x));
        double sx, sy, sz;
        rotate_site(c1, s1, c2, s2, c3, s3, relx, rely, relz, ncx, ncy, ncz, sx, sy, sz);
        PH_T(0);
        // centre-of-mass gate (:804-813): MCH partners per lane and chunk, GB chunks at a time so that their distance
        // tests and ballots overlap instead of queueing behind one another
        int nn = 0;
        unsigned neww0 = 0;
        constexpr int GB = SMALL ? 1 : 4;
        for (int wb = 0; wb < nchunk; wb += GB) {
            bool in[GB * MCH];
            unsigned word[GB];
#pragma unroll
            for (int q = 0; q < GB; ++q)
#pragma unroll
                for (int m = 0; m < MCH; ++m) {
                    const int j = 32 * (wb + q) + r + GL * m;
                    // every lane computes a distance (lanes past N on molecule 0) and the guard selects: no branch in the gate
                    const int jj = j < N ? j : 0;
                    const double r2 = pair_r2(com[jj] - ncx, com[Np + jj] - ncy, com[2 * Np + jj] - ncz, g);
                    in[q * MCH + m] = !((j < N && j != i ? r2 : DBL_HUGE) > g.cut2);
                }
#pragma unroll
            for (int q = 0; q < GB; ++q) {
                word[q] = 0;
#pragma unroll
                for (int m = 0; m < MCH; ++m) {
                    const unsigned bb = __ballot_sync(FULL, in[q * MCH + m]);
                    word[q] |= ((bb >> leader) & GMASK) << (GL * m);
                }
            }
#pragma unroll
            for (int q = 0; q < GB; ++q) {
                if (wb + q < nchunk) {
#pragma unroll
                    for (int m = 0; m < MCH; ++m)
                        if (in[q * MCH + m]) list[nsolp + nn + __popc(word[q] & ((1u << (r + GL * m)) - 1u))] = 32 * (wb + q) + r + GL * m;
                    if (wb + q == 0) neww0 = word[q];
                    if ((T > 1 || !SMALL) && r == 0) touch[wb + q] = word[q];     // the partners this step looked at (resolution, commit)
                    nn += __popc(word[q]);
                }
            }
        }
        __syncwarp();
        PH_T(1);
        // site pairs: lanes with the same own site split the slots; NI items staged per pass
        const int nslots = nsolp + nn;
        double lane_tot = 0.0;
        constexpr int NI = GL == 32 ? 2 : 4;
        {   // passes of NI items per lane (the same number on every lane: lanes past the last slot evaluate slot 0 and
            // drop it), except that a remainder of at most NSUB slots takes a pass of ONE item per lane: with the usual
            // 17-24 slots that is 12 site pairs per lane instead of 16 (+2.4 % on configs[1], round 2)
            int sb = 0;
            for (; nslots - sb > NSUB; sb += NI * NSUB)
                lane_tot += items_staged<COUL, NI>(g, sites, tabAB, tabQ, list, nsolp, nslots, sb + sub, NSUB, a, sx, sy, sz, epart);
            if (sb < nslots)
                lane_tot += items_staged<COUL, 1>(g, sites, tabAB, tabQ, list, nsolp, nslots, sb + sub, NSUB, a, sx, sy, sz, epart);
        }
        PH_T(2);
        // dE = sum(new row) - sum(old row) over the group; ENERGY(i) = the solute slots' share of it
        double dE = lane_tot - old_part;
        __syncwarp();
        // One step per warp in a SMALL box: what an accepted move would write (new matrix entries and row sums of
        // the partners inside the cutoff, the row sums of those that left) does not depend on dE, so it is formed
        // here, in the shadow of the reduction's shuffles; the commit below is then only stores.
        constexpr bool HOIST = SMALL && GL == 32 && T == 1;
        const int c_sl = nsolp + r;
        const bool c_has = HOIST && c_sl < nslots, c_left = HOIST && oldrow[0] != 0.0 && !((neww0 >> r) & 1u);
        int c_j = 0;
        double c_e = 0.0, c_rs_in = 0.0, c_rs_left = 0.0;
        if (HOIST) {
            c_j = c_has ? list[c_sl] : 0;
            const int esl = c_has ? c_sl : 0;
            const double2 e01 = *reinterpret_cast<const double2 *>(epart + 4 * esl), e23 = *reinterpret_cast<const double2 *>(epart + 4 * esl + 2);
            c_e = (e01.x + e01.y) + (e23.x + e23.y);
            c_rs_in = tot[c_j] + (c_e - mat[(size_t)i * MS + c_j]);
            c_rs_left = tot[r < N ? r : 0] - oldrow[0];
        }
        if (GL == 32) dE = warp_sum_mma(dE, lane);   // (two DMMAs + two shuffled doubles instead of five shuffle-add levels: +3 %)
        else {
#pragma unroll
            for (int o = GL / 2; o >= 1; o >>= 1) dE += shfl_xor_d(dE, o);
        }
        // METROPOLIS (:720-788) on DELTA = dE
        const double exponent = dE * neg_inv_rt;
        const bool clamped = !(dE <= DBL_HUGE);                      // +Inf/NaN trial energy: clamped, rejected (:421-428); -Inf passes, as there
        const bool used8 = !clamped && exponent >= -75.0 && exponent < 0.0;
        const bool accept = !clamped && (exponent >= 0.0 || (used8 && thr <= exponent));
        if (T > 1 && r == 0) {
            SpecX x;
            x.dE = dE; x.ncx = ncx; x.ncy = ncy; x.ncz = ncz; x.i = i; x.nn = nn;
            x.flags = (accept ? 1 : 0) | (used8 ? 2 : 0) | (clamped ? 4 : 0); x.pad = 0;
            xch[gidx] = x;
        }
        PH_T(3);
        if (W > 1) __syncthreads(); else if (T > 1) __syncwarp();
        // ================= resolution (every warp, redundantly): which steps of the batch stand =================
        const int limit = min(min(T, nsteps_i - done), s.nadj - adj_count);   // a batch never runs past an adaptation
        int progress = 1;
        if (T > 1) {
            bool dep = false;
            if (gb < T) {   // would the move of step ga, if accepted, change what step gb saw?
                const SpecX xa = xch[ga], xb = xch[gb];
                const unsigned *tb = reinterpret_cast<const unsigned *>(wk0 + (size_t)gb * g.spec_group_doubles + g.o_touch);
                dep = xa.i == xb.i || ((tb[xa.i >> 5] >> (xa.i & 31)) & 1u);
                dep = dep || !(pair_r2(xa.ncx - xb.ncx, xa.ncy - xb.ncy, xa.ncz - xb.ncz, g) > g.cut2);
                dep = dep || !(pair_r2(com[xb.i] - xa.ncx, com[Np + xb.i] - xa.ncy, com[2 * Np + xb.i] - xa.ncz, g) > g.cut2);
                // a non-zero entry of step gb's old matrix row = the two molecules are inside the cutoff as committed
                dep = dep || !(pair_r2(com[xb.i] - com[xa.i], com[Np + xb.i] - com[Np + xa.i], com[2 * Np + xb.i] - com[2 * Np + xa.i], g) > g.cut2);
            }
            const unsigned depbits = __ballot_sync(FULL, dep);
            unsigned acc_valid = xch[0].flags & 1u;
#pragma unroll
            for (int gg = 1; gg < T; ++gg) {
                const unsigned dm = (depbits >> (gg * (gg - 1) / 2)) & ((1u << gg) - 1u);
                if (progress == gg && gg < limit && (dm & acc_valid) == 0) {
                    progress = gg + 1;
                    acc_valid |= (unsigned)(xch[gg].flags & 1) << gg;
                }
            }
        }
        PH_T(5);
        // ================= commit the accepted steps that stand =================
        // They are mutually independent, but two of them can share a neighbour whose row sum both adjust, so with
        // several steps in flight the commits run one after the other.
#pragma unroll
        for (int cg = 0; cg < T; ++cg) {
            if (HOIST) {
                if (accept) {
                    if (c_left) { tot[r] = c_rs_left; mat[(size_t)i * MS + r] = 0.0; mat[(size_t)r * MS + i] = 0.0; }
                    if (c_has) { tot[c_j] = c_rs_in; mat[(size_t)i * MS + c_j] = c_e; mat[(size_t)c_j * MS + i] = c_e; }
                    if (r >= 1 && r <= 6) base[i] = nv;
                    if (r < 4) {
                        sites[SITE_DOUBLES * i + 3 * a] = sx; sites[SITE_DOUBLES * i + 3 * a + 1] = sy;
                        sites[SITE_DOUBLES * i + 3 * a + 2] = sz;
                    }
                    if (r == 0) tot[i] = dE + old_tot;
                    // (three predicated stores from the values the rotation used: forming them again in the shadow of the
                    // reduction, or under the accept branch, measured 4-8 % slower)
                    if (r == 0) sc[3 * i] = make_double2(s1, c1);
                    if (r == 1) sc[3 * i + 1] = make_double2(s2, c2);
                    if (r == 2) sc[3 * i + 2] = make_double2(s3, c3);
                }
            } else if (gidx == cg && gidx < progress && accept) {
                // partners that left the cutoff: entry to zero, their row sums lose the old pair energy
                if (prefetched) {
#pragma unroll
                    for (int q = 0; q < PF; ++q) {
                        if (q >= nchunk) break;
                        const unsigned neww = SMALL ? neww0 : touch[q];
#pragma unroll
                        for (int m = 0; m < MCH; ++m) {
                            const int j = 32 * q + r + GL * m;
                            const double old = oldrow[q * MCH + m];
                            if (old != 0.0 && !((neww >> (r + GL * m)) & 1u)) {
                                tot[j] -= old;
                                mat[(size_t)i * MS + j] = 0.0; mat[(size_t)j * MS + i] = 0.0;
                            }
                        }
                    }
                } else {
                    for (int w0 = 0; w0 < nchunk; ++w0) {
                        const unsigned neww = touch[w0];
#pragma unroll
                        for (int m = 0; m < MCH; ++m) {
                            const int j = 32 * w0 + r + GL * m;
                            const double old = j < N ? mat[(size_t)i * MS + j] : 0.0;
                            if (old != 0.0 && !((neww >> (r + GL * m)) & 1u)) {
                                tot[j] -= old;
                                mat[(size_t)i * MS + j] = 0.0; mat[(size_t)j * MS + i] = 0.0;
                            }
                        }
                    }
                }
                // partners inside the cutoff: SS(i,j) = SS(j,i) = EN (:819-820), row sums follow
                // (at most 31 partners in a SMALL box: one partner per lane, no loop)
                for (int sl = nsolp + r; sl < nslots; sl += GL) {
                    const int j = list[sl];
                    const double2 e01 = *reinterpret_cast<const double2 *>(epart + 4 * sl), e23 = *reinterpret_cast<const double2 *>(epart + 4 * sl + 2);
                    const double e = (e01.x + e01.y) + (e23.x + e23.y);
                    tot[j] += e - mat[(size_t)i * MS + j];
                    mat[(size_t)i * MS + j] = e; mat[(size_t)j * MS + i] = e;
                    if (SMALL && GL == 32) break;
                }
                if (r >= 1 && r <= 6) base[i] = nv;
                if (r < 4) {
                    sites[SITE_DOUBLES * i + 3 * a] = sx; sites[SITE_DOUBLES * i + 3 * a + 1] = sy;
                    sites[SITE_DOUBLES * i + 3 * a + 2] = sz;
                }
                if (r == 0) { tot[i] = dE + old_tot; if (!SMALL) movedw[i >> 5] |= 1u << (i & 31); }
                if (r == 0) sc[3 * i] = make_double2(s1, c1);
                if (r == 1) sc[3 * i + 1] = make_double2(s2, c2);
                if (r == 2) sc[3 * i + 2] = make_double2(s3, c3);
                if (SMALL) moved0 |= 1u << (i & 31);
            }
            if (T > 1) { if (W > 1) __syncthreads(); else __syncwarp(); }
        }
        if (HOIST) moved0 |= accept ? 1u << (i & 31) : 0u;
        PH_T(6);
        // ================= bookkeeping in step order (every warp keeps the same copy) =================
#pragma unroll
        for (int gg = 0; gg < T; ++gg) {
            if (gg < progress) {
                SpecX x;
                if (T > 1) x = xch[gg];
                else { x.dE = dE; x.nn = nn; x.flags = (accept ? 1 : 0) | (used8 ? 2 : 0) | (clamped ? 4 : 0); }   // one step in flight: no exchange
                // (uniforms used, site pairs and production samples follow from these sums and the step count at the end)
                n_used8 += (x.flags >> 1) & 1;
                nclamp += (x.flags >> 2) & 1;
                sum_nn += (unsigned)x.nn;
                s.last_delta = x.dE;
                // selects instead of branches: adding 0.0 leaves the sums as they are
                const bool in_prod = done + gg >= prod_from;
                const bool acc = (x.flags & 1) != 0;
                nsuc += acc; nacc += acc;
                s.toteng += acc ? x.dE : 0.0;
                s.production += acc && in_prod ? s.toteng : 0.0; prod_acc += acc && in_prod;
                const double te = in_prod ? s.toteng : 0.0;
                s.sum_e += te; s.sum_e2 = fma(te, te, s.sum_e2);
            }
        }
        racc = nacc;
        adj_count += progress;
        if (adj_count == s.nadj) {                                  // mod(I, NADJ) == 0, :766-777
            adj_count = 0;
            const double ratio = (double)((float)nacc / (float)s.nadj);
            if (ratio > 0.5) { s.dposmax *= (1.0 + s.padj); s.dhoekmax *= (1.0 + s.padj); }
            else if (ratio < 0.5) { s.dposmax *= (1.0 - s.padj); s.dhoekmax *= (1.0 - s.padj); }
            s.dposmax = fmax(s.dposmax, s.dposmin); s.dhoekmax = fmax(s.dhoekmax, s.dhoekmin);
            nacc = 0;
            if (W == 1 || warp == 0) ring_rescale(rg, ring0, done + progress, ring_hi, lane, s.dhoekmax);
        }
        done += progress;
        if (W > 1) __syncthreads(); else __syncwarp();
        PH_T(4);
    }
#ifdef MCB_PHASE_CLOCKS
    if (tid == 0) for (int qq = 0; qq < 8; ++qq) atomicAdd(&g_phase[qq], (unsigned long long)ph[qq]);
#endif
    // ENERGY(i) (CalcLJ_SV, LennardJones.f90:89-148) of every molecule that moved, from its final cached sites: one
    // (solute pseudo-molecule, own site) item per lane and pass, fixed summation order
    if (SMALL) {   // (with several groups or warps per chain each commits its own step: merge the masks)
        for (int o = 16; o >= 1; o >>= 1) moved0 |= __shfl_xor_sync(FULL, moved0, o);
        if (lane == 0 && moved0) atomicOr(&movedw[0], moved0);
    }
    __syncthreads();
    for (int i = warp; i < N; i += W) {
        if (!((movedw[i >> 5] >> (i & 31)) & 1u)) continue;
        double acc = 0.0;
        for (int t = lane; t < 4 * nsolp; t += 32) {
            const int p = t >> 2, as = t & 3;
            acc += item_energy<COUL>(g, sites, tabAB, tabQ, N + p, 1 + p, as, sites[SITE_DOUBLES * i + 3 * as], sites[SITE_DOUBLES * i + 3 * as + 1],
                                     sites[SITE_DOUBLES * i + 3 * as + 2]);
        }
        acc = warp_sum(acc);
        if (lane == 0) energy[i] = acc;
    }
    __syncthreads();
    {
        const int n_mut = g.o_sites + SITE_DOUBLES * N;
        for (int qq = tid; qq < n_mut; qq += 32 * W) gimg[qq] = img[qq];
        if (MAT_SMEM) for (int qq = tid; qq < N * MS; qq += 32 * W) gimg[g.o_mat + qq] = img[g.o_mat + qq];
    }
    if (tid == 0) {
        s.step += done; s.rng_n += (unsigned long long)done; s.nsuc += nsuc; s.naccept = nacc; s.nclamp += nclamp; s.prod_acc += prod_acc;
        s.uused += 7LL * done + n_used8;                                      // 7 uniforms per step, an 8th when METROPOLIS draws
        s.nstat += done > prod_from ? done - prod_from : 0;                   // steps at or after start_prod
        s.sitepairs += (long long)wpair * (long long)sum_nn + (long long)wsol * done;
        if (done > 0) s.ratio = (double)((float)racc / (float)s.nadj);
        ctl[c] = s;
        if (out_steps) out_steps[c] = done;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Matrix-free full-box energy for boxes whose N x N matrix cannot exist (BASELINE config 5: ~10^5 molecules; the
// reference itself needs 2 x 82 GB there, MonteCarlo.F90:320-321).  Same numbers as the start-up recompute
// (:357-364): every unordered pair inside the centre-of-mass cutoff once, the larger index as MOL1.
//   k_big_sites   RotMatrix for every molecule -> site cache                      (1 thread / molecule)
//   k_big_pairs   128 x 128 tiles of (j, i<j): gate on the centres of mass staged in shared memory, the rare hits
//                 evaluated in place; each block leaves one partial energy + pair count in SCRATCH
//   k_big_solute  ENERGY(i) for every molecule, block partials into SCRATCH
//   k_big_reduce  fixed-order sum of the scratch partials (deterministic; no fp64 atomics)
constexpr int BIG_TILE = 128;

__global__ void __launch_bounds__(128) k_big_sites(Geo g, SiteRel rel, double *img)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N) return;
    const double *com = img + g.o_com, *hoek = img + g.o_hoek;
    double *sites = img + g.o_sites;
    double s1, c1, s2, c2, s3, c3;
    sincos(hoek[j], &s1, &c1); sincos(hoek[g.Np + j], &s2, &c2); sincos(hoek[2 * g.Np + j], &s3, &c3);
    const double cx = com[j], cy = com[g.Np + j], cz = com[2 * g.Np + j];
    for (int a = 0; a < NSITE; ++a) {
        double ox, oy, oz;
        rotate_site(c1, s1, c2, s2, c3, s3, rel.r[3 * a], rel.r[3 * a + 1], rel.r[3 * a + 2], cx, cy, cz, ox, oy, oz);
        sites[SITE_DOUBLES * j + 3 * a] = ox; sites[SITE_DOUBLES * j + 3 * a + 1] = oy; sites[SITE_DOUBLES * j + 3 * a + 2] = oz;
    }
}

// fixed-order block sum (tree over shared memory): same result on every run
__device__ __forceinline__ double block_sum_128(double v, double *sh)
{
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 64; o >= 1; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    const double r = sh[0];
    __syncthreads();
    return r;
}

template <bool COUL>
__global__ void __launch_bounds__(128) k_big_pairs(Geo g, const double *img, int ntiles, double *scratch_e, long long *scratch_n)
{
    __shared__ double sc[3][BIG_TILE];
    __shared__ double red[BIG_TILE];
    // linear block index -> tile pair (tj >= ti) of the lower triangle
    const long long b = blockIdx.x;
    int tj = (int)((sqrt(8.0 * (double)b + 1.0) - 1.0) * 0.5);
    while ((long long)tj * (tj + 1) / 2 > b) --tj;
    while ((long long)(tj + 1) * (tj + 2) / 2 <= b) ++tj;
    const int ti = (int)(b - (long long)tj * (tj + 1) / 2);
    (void)ntiles;
    const double *com = img + g.o_com, *sites = img + g.o_sites, *tabAB = img + g.o_tabAB, *tabQ = img + g.o_tabQ;
    const int i0 = ti * BIG_TILE, j = tj * BIG_TILE + threadIdx.x;
    {
        const int i = i0 + threadIdx.x;
        const bool ok = i < g.N;
        sc[0][threadIdx.x] = ok ? com[i] : 0.0; sc[1][threadIdx.x] = ok ? com[g.Np + i] : 0.0; sc[2][threadIdx.x] = ok ? com[2 * g.Np + i] : 0.0;
    }
    __syncthreads();
    double e = 0.0;
    long long np = 0;
    if (j < g.N) {
        const double cx = com[j], cy = com[g.Np + j], cz = com[2 * g.Np + j];
        const int ni = min(BIG_TILE, min(g.N, j) - i0);      // partners i < j only
        for (int k = 0; k < ni; ++k) {
            // R = CoM(J) - CoM(I) with I = j (MOL1), J = i (MonteCarlo.F90:806-812)
            if (pair_r2(sc[0][k] - cx, sc[1][k] - cy, sc[2][k] - cz, g) > g.cut2) continue;
            ++np;
            const double *s1 = sites + SITE_DOUBLES * (size_t)j, *s2 = sites + SITE_DOUBLES * (size_t)(i0 + k);
            double en = 0.0;
            for (int a = 0; a < NSITE; ++a)          // calcLJ: a outer over MOL1, bb inner over MOL2 (LennardJones.f90:65-79)
                for (int bb = 0; bb < NSITE; ++bb)
                    en = site_pair_acc<COUL>(en, s2[3 * bb] - s1[3 * a], s2[3 * bb + 1] - s1[3 * a + 1], s2[3 * bb + 2] - s1[3 * a + 2],
                                             tabAB[a * TAB_ROW + 2 * bb], tabAB[a * TAB_ROW + 2 * bb + 1], COUL ? tabQ[a * 4 + bb] : 0.0, g);
            e += en;
        }
    }
    const double be = block_sum_128(e, red);
    const double bn = block_sum_128((double)np, red);
    if (threadIdx.x == 0) { scratch_e[blockIdx.x] = be; scratch_n[blockIdx.x] = (long long)bn; }
}

template <bool COUL>
__global__ void __launch_bounds__(128) k_big_solute(Geo g, double *img, double *scratch_e)
{
    __shared__ double red[BIG_TILE];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const double *sites = img + g.o_sites, *tabAB = img + g.o_tabAB, *tabQ = img + g.o_tabQ;
    double e = 0.0;
    if (i < g.N) {
        const double *s1 = sites + SITE_DOUBLES * (size_t)i;
        for (int a = 0; a < NSITE; ++a)              // CalcLJ_SV (LennardJones.f90:131-143)
            for (int p = 0; p < g.nsolp; ++p) {
                const double *s2 = sites + SITE_DOUBLES * (size_t)(g.N + p);
                for (int bb = 0; bb < 4; ++bb) {
                    const int t = ((1 + p) * 4 + a) * 4 + bb;
                    e = site_pair_acc<COUL>(e, s2[3 * bb] - s1[3 * a], s2[3 * bb + 1] - s1[3 * a + 1], s2[3 * bb + 2] - s1[3 * a + 2],
                                            tabAB[((1 + p) * 4 + a) * TAB_ROW + 2 * bb], tabAB[((1 + p) * 4 + a) * TAB_ROW + 2 * bb + 1], COUL ? tabQ[t] : 0.0, g);
                }
            }
        img[g.o_energy + i] = e;
    }
    const double be = block_sum_128(e, red);
    if (threadIdx.x == 0) scratch_e[blockIdx.x] = be;
}

// one block: out[0] = sum of pair partials, out[1] = sum of solute partials, n_out = pairs in cutoff
__global__ void __launch_bounds__(128) k_big_reduce(const double *pair_e, const long long *pair_n, long long npair_blocks,
                                                    const double *sol_e, int nsol_blocks, double *out, long long *n_out)
{
    __shared__ double red[BIG_TILE];
    double e = 0.0, n = 0.0, es = 0.0;
    for (long long k = threadIdx.x; k < npair_blocks; k += BIG_TILE) { e += pair_e[k]; n += (double)pair_n[k]; }
    for (int k = threadIdx.x; k < nsol_blocks; k += BIG_TILE) es += sol_e[k];
    const double te = block_sum_128(e, red), tn = block_sum_128(n, red), ts = block_sum_128(es, red);
    if (threadIdx.x == 0) { out[0] = te; out[1] = ts; n_out[0] = (long long)tn; }
}

// ---------------------------------------------------------------------------------------------------------
// fp64 peak: 8 independent DFMA chains per thread, all SMs.
__global__ void k_fp64_peak(double *out, int iters)
{
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 0.999999, b = 1e-6;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = fma(a0, m, b); a1 = fma(a1, m, b); a2 = fma(a2, m, b); a3 = fma(a3, m, b);
            a4 = fma(a4, m, b); a5 = fma(a5, m, b); a6 = fma(a6, m, b); a7 = fma(a7, m, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
}

}  // namespace mcb

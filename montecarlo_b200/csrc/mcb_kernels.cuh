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
template <bool COUL>
__global__ void __launch_bounds__(32) k_prepare(Geo g, SiteRel rel, double *images, double *matg, double *work,
                                                ChainCtl *ctl, int chain0, double *out_tot, double *out_esol,
                                                long long *out_pairs)
{
    const int lane = threadIdx.x;
    const int c = chain0 + blockIdx.x;
    double *img = images + (size_t)c * g.image_doubles;
    double *mat = g.mat_in_image ? img + g.o_mat : matg + (size_t)c * g.N * g.MS;
    ChainPtrs p = chain_ptrs(g, img, mat, work + (size_t)c * g.work_doubles);
    const int N = g.N, Np = g.Np;

    for (int j = lane; j < N; j += 32) {   // RotMatrix for every molecule, lib.f90:27-83
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
    __syncwarp();

    const int a = lane & 3;
    long long pairs = 0;
    double esum = 0.0;
    for (int i = 0; i < N; ++i) {
        double sx = p.sites[SITE_DOUBLES * i + 3 * a], sy = p.sites[SITE_DOUBLES * i + 3 * a + 1],
               sz = p.sites[SITE_DOUBLES * i + 3 * a + 2];
        double e_sol;
        int nn = row_eval<COUL>(g, p, i, i, p.com[i], p.com[Np + i], p.com[2 * Np + i], sx, sy, sz, lane, e_sol);
        pairs += nn;
        esum += e_sol;
        if (lane == 0) { p.energy[i] = e_sol; mat[(size_t)i * g.MS + i] = 0.0; }
        for (int j = lane; j < i; j += 32) {
            double v = p.newrow[j];
            mat[(size_t)i * g.MS + j] = v;
            mat[(size_t)j * g.MS + i] = v;
        }
        __syncwarp();
    }
    // calcEnergy (MonteCarlo.F90:988-1004); lane-parallel partial sums, fixed order => deterministic
    double acc = 0.0;
    for (int i = 0; i < N; ++i)
        for (int j = i + 1 + lane; j < N; j += 32) acc += mat[(size_t)i * g.MS + j];
    for (int i = lane; i < N; i += 32) acc += p.energy[i];
    double tot = warp_sum(acc);
    if (lane == 0) {
        ctl[c].toteng = tot;
        ctl[c].energies_valid = 1;
        if (out_tot) out_tot[c - chain0] = tot;
        if (out_esol) out_esol[c - chain0] = esum;
        if (out_pairs) out_pairs[c - chain0] = pairs;
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

// ---------------------------------------------------------------------------------------------------------
// Batched mode: nsteps Metropolis steps of one chain per warp, the chain's image resident in shared memory.
//   UBUF      uniforms come from a caller buffer [chain][cap] instead of the built-in counter stream
//   MAT_SMEM  the pair matrix is part of the staged image (small N); otherwise it stays in global memory
template <bool COUL, bool UBUF, bool MAT_SMEM>
__global__ void __launch_bounds__(32) k_run_steps(Geo g, SiteRel rel, double *images, double *matg, ChainCtl *ctl,
                                                  long long nsteps, const double *ubuf, long long cap,
                                                  long long *out_consumed, long long *out_steps)
{
    extern __shared__ __align__(16) double smem[];
    const int lane = threadIdx.x;
    const int c = blockIdx.x;
    const int N = g.N, Np = g.Np, MS = g.MS;
    double *gimg = images + (size_t)c * g.image_doubles;
    double *img = smem;
    {   // stage the image: linear, coalesced 16-byte loads
        const double2 *src = reinterpret_cast<const double2 *>(gimg);
        double2 *dst = reinterpret_cast<double2 *>(img);
        for (int k = lane; k < g.stage_doubles / 2; k += 32) dst[k] = src[k];
    }
    double *mat = MAT_SMEM ? img + g.o_mat : matg + (size_t)c * N * MS;
    ChainPtrs p = chain_ptrs(g, img, mat, smem + g.stage_doubles);
    ChainCtl s = ctl[c];
    __syncwarp();

    // re-anchor TOTENG on the stored matrix so that rounding of the running sum cannot accumulate across launches
    {
        double acc = 0.0;
        for (int i = 0; i < N; ++i)
            for (int j = i + 1 + lane; j < N; j += 32) acc += mat[(size_t)i * MS + j];
        for (int i = lane; i < N; i += 32) acc += p.energy[i];
        s.toteng = warp_sum(acc);
    }

    const int a = lane & 3;
    const double relx = rel.r[3 * a], rely = rel.r[3 * a + 1], relz = rel.r[3 * a + 2];
    const int wpair = rel.n_sites * rel.n_sites, wsol = rel.n_sites * g.nsol;
    const double rt = 8.314 * s.temperature;                       // MonteCarlo.F90:729
    const int r = min(max(lane - 1, 0), 5);
    const bool isang = r >= 3;
    const int k = isang ? r - 3 : r;
    double *const base = (isang ? p.hoek : p.com) + k * Np;
    const double period = isang ? TAU_D : g.boxl2, half = isang ? PI_D : g.boxl;
    int adj_count = (int)(s.step % s.nadj);
    long long cursor = 0, done = 0;
    const double *myubuf = UBUF ? ubuf + (size_t)c * cap : nullptr;

    for (; done < nsteps; ++done) {
        double myu;
        if (UBUF) {
            if (cursor + 8 > cap) break;
            myu = lane < 8 ? myubuf[cursor + lane] : 0.0;
        } else {
            myu = stream_uniform(s.rng_key, s.rng_n + (unsigned long long)(lane & 7));
        }
        // ---- MCINIT (MonteCarlo.F90:685-715): lane 0 picks the molecule, lanes 1-3 move the centre of mass,
        //      lanes 4-6 the angles (randVec / randVecHoek, lib.f90:111-131), then the wraps (:705-712)
        const double u0 = shfl_d(myu, 0);
        int i = (int)(u0 * (double)N) + 1;
        if (g.quirk31 && i == 31) i = 30;                         // :698
        i = min(i, N) - 1;
        double rv = (myu - 0.5) * 2.0;
        rv = (isang ? 1.0 : INV_SQRT3) * rv;
        rv = (isang ? s.dhoekmax : s.dposmax) * rv;
        double nv = wrap_exact(base[i] + rv, period, half);
        double sn, cs;
        sincos(nv, &sn, &cs);
        const double ncx = shfl_d(nv, 1), ncy = shfl_d(nv, 2), ncz = shfl_d(nv, 3);
        const double c1 = shfl_d(cs, 4), s1 = shfl_d(sn, 4), c2 = shfl_d(cs, 5), s2 = shfl_d(sn, 5),
                     c3 = shfl_d(cs, 6), s3 = shfl_d(sn, 6);
        double sx, sy, sz;
        rotate_site(c1, s1, c2, s2, c3, s3, relx, rely, relz, ncx, ncy, ncz, sx, sy, sz);

        // ---- calculateLJ(RSOLV) (:792-828) as a difference against the stored row, then calcEnergy (:988-1004)
        double e_sol;
        const int nn = row_eval<COUL>(g, p, i, N, ncx, ncy, ncz, sx, sy, sz, lane, e_sol);
        double d = 0.0;
        for (int j = lane; j < N; j += 32) d += p.newrow[j] - mat[(size_t)i * MS + j];
        const double dE = warp_sum(d) + (e_sol - p.energy[i]);
        double tnew = s.toteng + dE;
        if (!(tnew <= DBL_HUGE)) { tnew = DBL_HUGE; s.nclamp++; }  // Inf or NaN, :421-428

        // ---- METROPOLIS (:720-788)
        const double delta = tnew - s.toteng;
        const double exponent = -1.0 * delta * 1000.0 / rt;
        double kans, rvv;
        int used = 7;
        if (exponent < -75.0) { kans = 0.0; rvv = 1.0; }
        else if (exponent >= 0.0) { kans = 1.0; rvv = 0.0; }
        else { kans = fmin(exp(exponent * LN_E_F32), 1.0); rvv = shfl_d(myu, 7); used = 8; }
        const bool accept = rvv <= kans;
        s.step++; s.uused += used; s.last_delta = delta;
        s.sitepairs += (long long)(wpair * nn + wsol);
        if (UBUF) cursor += used; else s.rng_n += used;
        if (accept) {
            s.nsuc++; s.naccept++;
            if (s.step >= s.start_prod) { s.production += tnew; s.prod_acc++; }
            for (int j = lane; j < N; j += 32) {
                double v = p.newrow[j];
                if (MAT_SMEM || v != mat[(size_t)i * MS + j]) { mat[(size_t)i * MS + j] = v; mat[(size_t)j * MS + i] = v; }
            }
            if (lane >= 1 && lane <= 6) base[i] = nv;
            if (lane < 4) {
                p.sites[SITE_DOUBLES * i + 3 * a] = sx; p.sites[SITE_DOUBLES * i + 3 * a + 1] = sy;
                p.sites[SITE_DOUBLES * i + 3 * a + 2] = sz;
            }
            if (lane == 0) p.energy[i] = e_sol;
            s.toteng = tnew;
        }
        if (s.step >= s.start_prod) { s.sum_e += s.toteng; s.sum_e2 = fma(s.toteng, s.toteng, s.sum_e2); s.nstat++; }
        s.ratio = (double)((float)s.naccept / (float)s.nadj);       // default-REAL quotient, :758
        if (++adj_count == s.nadj) {                                // mod(I, NADJ) == 0, :766-777
            adj_count = 0;
            if (s.ratio > 0.5) { s.dposmax *= (1.0 + s.padj); s.dhoekmax *= (1.0 + s.padj); }
            else if (s.ratio < 0.5) { s.dposmax *= (1.0 - s.padj); s.dhoekmax *= (1.0 - s.padj); }
            s.dposmax = fmax(s.dposmax, s.dposmin); s.dhoekmax = fmax(s.dhoekmax, s.dhoekmin);
            s.naccept = 0;
        }
        __syncwarp();
    }

    // write back what can have changed: com, hoek, energy, solvent sites (contiguous) and the matrix
    {
        const int n_mut = g.o_sites + SITE_DOUBLES * N;
        for (int q = lane; q < n_mut; q += 32) gimg[q] = img[q];
        if (MAT_SMEM) for (int q = lane; q < N * MS; q += 32) gimg[g.o_mat + q] = img[g.o_mat + q];
    }
    if (lane == 0) {
        ctl[c] = s;
        if (out_consumed) out_consumed[c] = UBUF ? cursor : 0;
        if (out_steps) out_steps[c] = done;
    }
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

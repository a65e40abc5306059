// mcb_verlet.cuh — batched stepping for LARGE boxes (N > 32, pair matrix in global memory): k_run_vl.
//
// Same Markov chain as k_run_spec with one step in flight (same ring of pre-drawn moves, same arithmetic, same
// commit), but nothing in a step is O(N) any more.  Every molecule keeps a list of the molecules whose centre of
// mass was within cutoff + skin when the lists were built (a Verlet list, in global memory: the moved molecule of
// the next step is known from the ring of pre-drawn moves, so its list is fetched one step ahead).  While no molecule has
// strayed more than skin/2 from where it was at build time, every partner inside the cutoff of a trial position is
// on the moved molecule's list, so the centre-of-mass gate (MonteCarlo.F90:804-813), the old matrix row needed for
// the row-sum bookkeeping and the "left the cutoff" test only look at the ~25 listed candidates instead of all N
// molecules.  The list only decides WHICH pairs get the reference's exact gate test; a trial position further than
// skin/2 from the build position (or a molecule whose list overflowed) falls back to the full scan for that step,
// and an accepted move beyond skin/2 rebuilds all lists.  Requires every centre of mass inside [-BOXL, BOXL]
// (true after the wrap at :710-712), otherwise the reference's minimum image is not a metric and lists stay off.
#pragma once

#include "mcb_kernels.cuh"

namespace mcb {

constexpr int VL_K = 64;   // list capacity per molecule (two 32-lane chunks); cutoff 7 + skin 2 holds ~25 in liquid DMSO

// per chain: idx[N][VL_K], cnt[N + 2] (slot N: rebuilds in the last launch; slot N+1: lists match the stored coordinates), ref[3][Np]
// trust: nothing but k_run_vl has changed the coordinates since the last launch, so valid lists are kept across launches
struct VerletBufs { unsigned short *idx; int *cnt; double *ref; double skin; int trust; };

// shared memory of k_run_vl: image, one work area, exchange record, ring (threshold, moves, picks), row sums
__host__ __device__ inline int vl_smem_doubles(const Geo &g) { return g.stage_doubles + g.work_doubles + SPECX_DOUBLES + 64 + 6 * 64 + 32 + g.Np + 2; }

template <bool COUL>
__global__ void __launch_bounds__(32) k_run_vl(Geo g, SiteRel rel, double *images, double *matg, ChainCtl *ctl,
                                               long long nsteps, long long *out_steps, VerletBufs vb)
{
    extern __shared__ __align__(16) double smem[];
    const int lane = threadIdx.x;
    const int c = blockIdx.x;
    const int N = g.N, Np = g.Np, MS = g.MS, nsolp = g.nsolp;
    const unsigned lt = (1u << lane) - 1u;
    double *gimg = images + (size_t)c * g.image_doubles;
    double *img = smem;
    {
        const double2 *src = reinterpret_cast<const double2 *>(gimg);
        double2 *dst = reinterpret_cast<double2 *>(img);
        for (int k = lane; k < g.stage_doubles / 2; k += 32) dst[k] = src[k];
    }
    double *const mat = matg + (size_t)c * N * MS;
    double *const com = img + g.o_com, *const hoek = img + g.o_hoek, *const energy = img + g.o_energy;
    double *const sites = img + g.o_sites;
    const double *const tabAB = img + g.o_tabAB, *const tabQ = img + g.o_tabQ;
    const int a = lane & 3, sub = lane >> 2;
    double *const wk = smem + g.stage_doubles;
    int *const list = reinterpret_cast<int *>(wk + g.o_list);
    double *const epart = wk + g.o_epart;
    double *const ring_thr = wk + g.spec_group_doubles + SPECX_DOUBLES;   // same shared-memory plan as k_run_spec, T = 1
    double *const ring_rv = ring_thr + 64;
    int *const ring_i = reinterpret_cast<int *>(ring_rv + 6 * 64);
    double *const rowsum = ring_rv + 6 * 64 + 32;
    unsigned short *const vidx = vb.idx + (size_t)c * N * VL_K;
    int *const vcnt = vb.cnt + (size_t)c * (N + 2);
    double *const vref = vb.ref + (size_t)c * 3 * Np;
    const double rl = g.cutoff + vb.skin, rl2 = rl * rl, h2 = 0.25 * vb.skin * vb.skin;
    ChainCtl s = ctl[c];
    for (int p = lane; p < nsolp; p += 32) list[p] = N + p;
    __syncwarp();
    for (int j = lane; j < N; j += 32) {   // row sums from the stored matrix (symmetric: column sums, coalesced)
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int i = 0;
        for (; i + 3 < N; i += 4) {
            a0 += mat[(size_t)i * MS + j]; a1 += mat[(size_t)(i + 1) * MS + j];
            a2 += mat[(size_t)(i + 2) * MS + j]; a3 += mat[(size_t)(i + 3) * MS + j];
        }
        for (; i < N; ++i) a0 += mat[(size_t)i * MS + j];
        rowsum[j] = (a0 + a1) + (a2 + a3);
    }
    __syncwarp();
    {
        double acc = 0.0;
        for (int i = lane; i < N; i += 32) acc += fma(0.5, rowsum[i], energy[i]);
        s.toteng = warp_sum(acc);
    }
    // lists are usable only if the reference's minimum image is a true metric: all centres of mass inside the box
    bool inside = true;
    for (int j = lane; j < N; j += 32) inside = inside && fabs(com[j]) <= g.boxl && fabs(com[Np + j]) <= g.boxl && fabs(com[2 * Np + j]) <= g.boxl;
    const bool lists_on = __all_sync(FULL, inside);
    bool need_rebuild = lists_on && !(vb.trust && vcnt[N + 1] == 1);

    const double relx = rel.r[3 * a], rely = rel.r[3 * a + 1], relz = rel.r[3 * a + 2];
    const int wpair = rel.n_sites * rel.n_sites, wsol = rel.n_sites * g.nsol;
    const double neg_inv_rt = -1000.0 / (8.314 * s.temperature);
    const int rr = min(max(lane - 1, 0), 5);
    const bool isang = rr >= 3;
    const int k = isang ? rr - 3 : rr;
    double *const base = (isang ? hoek : com) + k * Np;
    const double period = isang ? TAU_D : g.boxl2, half = isang ? PI_D : g.boxl;
    const double inv_half = isang ? 1.0 / PI_D : g.inv_boxl;
    int adj_count = (int)(s.step % s.nadj);
    int done = 0;
    unsigned nsuc = 0, nacc = (unsigned)s.naccept, racc = 0, nclamp = 0, uused = 0, prod_acc = 0, nstat = 0, nrebuild = 0;
    unsigned long long sitepairs = 0;
    const unsigned long long step0 = (unsigned long long)s.step;
    const int nsteps_i = (int)min(nsteps, (long long)0x3fffffff);
    const unsigned long long pos0 = s.rng_n;   // position in the uniform stream (steps since mcb_seed_chains), see k_run_spec
    unsigned long long gen_end = pos0;
    const long long pf = s.start_prod - (long long)step0 - 1;     // launch-local index of the first production step (:760)
    const int prod_from = pf <= 0 ? 0 : (pf > 0x3fffffff ? 0x7fffffff : (int)pf);
    // list of the NEXT step's molecule (known from the ring), fetched one step ahead: count, candidates, build position
    int p_cnt = -1, p_idx[2] = {0, 0};
    double p_rfx = 0.0, p_rfy = 0.0, p_rfz = 0.0;
    bool p_valid = false;

    while (done < nsteps_i) {
        if (need_rebuild) {
            // ---- (re)build every list from the current centres of mass, which become the reference positions
            for (int j = lane; j < 3 * Np; j += 32) vref[j] = com[j];
            for (int m = 0; m < N; ++m) {
                const double mx = com[m], my = com[Np + m], mz = com[2 * Np + m];
                int cnt = 0;
                for (int c0 = 0; c0 < N; c0 += 32) {
                    const int j = c0 + lane;
                    const bool in = j < N && j != m && !(pair_r2(com[j] - mx, com[Np + j] - my, com[2 * Np + j] - mz, g) > rl2);
                    const unsigned b = __ballot_sync(FULL, in);
                    const int pos = cnt + __popc(b & lt);
                    if (in && pos < VL_K) vidx[(size_t)m * VL_K + pos] = (unsigned short)j;
                    cnt += __popc(b);
                }
                if (lane == 0) vcnt[m] = cnt > VL_K ? -1 : cnt;
            }
            need_rebuild = false;
            p_valid = false;
            nrebuild++;
            __syncwarp();
        }
        while (pos0 + (unsigned)done + 2 > gen_end) {   // pre-draw the next 32 steps, one step per lane (as k_run_spec)
            const unsigned long long sidx = gen_end + (unsigned)lane;
            const int e = (int)(sidx & 63ULL);
            const double u0 = stream_uniform(s.rng_key, 8ULL * sidx);
            int ii = (int)(u0 * (double)N) + 1;                     // MCINIT, MonteCarlo.F90:697-698
            if (g.quirk31 && ii == 31) ii = 30;
            ring_i[e] = min(ii, N) - 1;
#pragma unroll
            for (int q = 0; q < 6; ++q) {
                const double rvu = (stream_uniform(s.rng_key, 8ULL * sidx + 1 + q) - 0.5) * 2.0;
                ring_rv[6 * e + q] = (q < 3 ? INV_SQRT3 : 1.0) * rvu;
            }
            ring_thr[e] = log(stream_uniform(s.rng_key, 8ULL * sidx + 7)) * (1.0 / LN_E_F32);
            gen_end += 32;
            __syncwarp();
        }
        const int ring_e = (int)((pos0 + (unsigned)done) & 63ULL);
        const int i = ring_i[ring_e];
        const double thr = ring_thr[ring_e];
        const double old_tot = rowsum[i] + energy[i];
        const double old_part = lane == 0 ? old_tot : 0.0;
        // ---- candidates of molecule i: fetched during the previous step (global memory, two dependent round trips
        //      otherwise); their old matrix entries are issued now and only needed by the commit
        if (lists_on && !p_valid) {
            p_cnt = vcnt[i];
            p_idx[0] = (int)vidx[(size_t)i * VL_K + lane]; p_idx[1] = (int)vidx[(size_t)i * VL_K + 32 + lane];
            p_rfx = vref[i]; p_rfy = vref[Np + i]; p_rfz = vref[2 * Np + i];
        }
        const int cnt = lists_on ? p_cnt : -1;
        int cj[2];
        double cold[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) cj[q] = lane + 32 * q < cnt ? p_idx[q] : -1;
        const double rfx = p_rfx, rfy = p_rfy, rfz = p_rfz;
#pragma unroll
        for (int q = 0; q < 2; ++q) cold[q] = cj[q] >= 0 ? mat[(size_t)i * MS + cj[q]] : 0.0;
        if (lists_on) {   // the next step's list (lists only change in a rebuild, which clears p_valid)
            const int in = ring_i[(int)((pos0 + (unsigned)done + 1ULL) & 63ULL)];
            p_cnt = vcnt[in];
            p_idx[0] = (int)vidx[(size_t)in * VL_K + lane]; p_idx[1] = (int)vidx[(size_t)in * VL_K + 32 + lane];
            p_rfx = vref[in]; p_rfy = vref[Np + in]; p_rfz = vref[2 * Np + in];
            p_valid = true;
        }
        // ---- trial move (MCINIT :699-712), RotMatrix of the own site
        const double rv = (isang ? s.dhoekmax : s.dposmax) * ring_rv[6 * ring_e + rr];
        const double nv = wrap_fma(base[i] + rv, period, half, inv_half);
        double sn, cs;
        sincos_small(nv, sn, cs);
        const double ncx = shfl_d(nv, 1), ncy = shfl_d(nv, 2), ncz = shfl_d(nv, 3);
        const double c1 = shfl_d(cs, 4), s1 = shfl_d(sn, 4), c2 = shfl_d(cs, 5), s2 = shfl_d(sn, 5), c3 = shfl_d(cs, 6), s3 = shfl_d(sn, 6);
        double sx, sy, sz;
        rotate_site(c1, s1, c2, s2, c3, s3, relx, rely, relz, ncx, ncy, ncz, sx, sy, sz);
        // is the trial position still covered by the list?  (same value on every lane)
        const bool far = !lists_on || pair_r2(ncx - rfx, ncy - rfy, ncz - rfz, g) > h2;
        const bool use_list = cnt >= 0 && !far;
        // ---- centre-of-mass gate (:804-813) on the candidates, or on everybody
        int nn = 0;
        bool cin[2] = {false, false};
        int cslot[2] = {0, 0};
        if (use_list) {
#pragma unroll
            for (int q = 0; q < 2; ++q) {   // every lane computes a distance (empty lanes on molecule 0), the guard selects
                const int jj = cj[q] >= 0 ? cj[q] : 0;
                const double r2 = pair_r2(com[jj] - ncx, com[Np + jj] - ncy, com[2 * Np + jj] - ncz, g);
                cin[q] = !((cj[q] >= 0 ? r2 : DBL_HUGE) > g.cut2);
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const unsigned b = __ballot_sync(FULL, cin[q]);
                cslot[q] = nsolp + nn + __popc(b & lt);
                if (cin[q]) list[cslot[q]] = cj[q];
                nn += __popc(b);
            }
        } else {
            for (int c0 = 0; c0 < N; c0 += 32) {
                const int j = c0 + lane;
                const bool in = j < N && j != i && !(pair_r2(com[j] - ncx, com[Np + j] - ncy, com[2 * Np + j] - ncz, g) > g.cut2);
                const unsigned b = __ballot_sync(FULL, in);
                if (in) list[nsolp + nn + __popc(b & lt)] = j;
                nn += __popc(b);
            }
        }
        __syncwarp();
        // ---- site pairs (calcLJ / CalcLJ_SV), staged
        const int nslots = nsolp + nn;
        double lane_tot = 0.0;
        for (int sb = 0; sb < nslots; sb += 2 * 8)   // the same number of passes on every lane
            lane_tot += items_staged<COUL, 2>(g, sites, tabAB, tabQ, list, nsolp, nslots, sb + sub, 8, a, sx, sy, sz, epart);
        double dE = lane_tot - old_part, esol = 0.0;
        __syncwarp();
        // what an accepted move would write for this lane's candidates (new matrix entry: the pair energy inside the
        // cutoff, 0 outside; the candidate's row sum) does not depend on dE: formed here, in the shadow of the
        // reduction, from the old entries fetched at the top of the step - the commit is then stores only
        double cval[2] = {0.0, 0.0}, crs[2] = {0.0, 0.0};
        bool cchg[2] = {false, false};
        if (use_list) {
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int sl = cin[q] ? cslot[q] : 0, jj = cj[q] >= 0 ? cj[q] : 0;
                const double2 e01 = *reinterpret_cast<const double2 *>(epart + 4 * sl), e23 = *reinterpret_cast<const double2 *>(epart + 4 * sl + 2);
                cval[q] = cin[q] ? (e01.x + e01.y) + (e23.x + e23.y) : 0.0;
                crs[q] = rowsum[jj] + (cval[q] - cold[q]);
                cchg[q] = cj[q] >= 0 && (cin[q] || cold[q] != 0.0);
            }
        }
        if (sub < nsolp) esol = epart[4 * sub + a];
        if (nsolp > 8) {
#pragma unroll 1
            for (int p = sub + 8; p < nsolp; p += 8) esol += epart[4 * p + a];
        }
        dE = warp_sum_mma(dE, lane); esol = warp_sum_mma(esol, lane);   // (fp64 tensor-core sums, as k_run_spec)
        // ---- METROPOLIS (:720-788) on DELTA = dE
        const double exponent = dE * neg_inv_rt;
        const bool clamped = !(dE <= DBL_HUGE);   // -Inf is not clamped (MonteCarlo.F90:421-428)
        const bool used8 = !clamped && exponent >= -75.0 && exponent < 0.0;
        const bool accept = !clamped && (exponent >= 0.0 || (used8 && thr <= exponent));
        if (accept) {
            if (use_list) {
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    if (cchg[q]) {   // SS(i,j) = SS(j,i) = EN inside the cutoff (:819-820), 0 for a partner that left it
                        const int j = cj[q];
                        rowsum[j] = crs[q];
                        mat[(size_t)i * MS + j] = cval[q]; mat[(size_t)j * MS + i] = cval[q];
                    }
            } else {
                // full scan: partners that left the cutoff, then the partners inside it
                for (int j = lane; j < N; j += 32) {
                    const double old = mat[(size_t)i * MS + j];
                    if (old != 0.0 && j != i && pair_r2(com[j] - ncx, com[Np + j] - ncy, com[2 * Np + j] - ncz, g) > g.cut2) {
                        rowsum[j] -= old;
                        mat[(size_t)i * MS + j] = 0.0; mat[(size_t)j * MS + i] = 0.0;
                    }
                }
                for (int sl = nsolp + lane; sl < nslots; sl += 32) {
                    const int j = list[sl];
                    const double e = (epart[4 * sl] + epart[4 * sl + 1]) + (epart[4 * sl + 2] + epart[4 * sl + 3]);
                    rowsum[j] += e - mat[(size_t)i * MS + j];
                    mat[(size_t)i * MS + j] = e; mat[(size_t)j * MS + i] = e;
                }
            }
            if (lane >= 1 && lane <= 6) base[i] = nv;
            if (lane < 4) {
                sites[SITE_DOUBLES * i + 3 * a] = sx; sites[SITE_DOUBLES * i + 3 * a + 1] = sy; sites[SITE_DOUBLES * i + 3 * a + 2] = sz;
            }
            if (lane == 0) { energy[i] = esol; rowsum[i] = (dE + old_tot) - esol; }
            if (lists_on && far) need_rebuild = true;   // the molecule now sits more than skin/2 from its build position
        }
        // ---- bookkeeping
        uused += 7 + (used8 ? 1 : 0);
        nclamp += clamped ? 1 : 0;
        sitepairs += (unsigned)(wpair * nn + wsol);
        s.last_delta = dE;
        const bool in_prod = done >= prod_from;                    // selects instead of branches: adding 0.0 changes nothing
        nsuc += accept; nacc += accept;
        s.toteng += accept ? dE : 0.0;
        s.production += accept && in_prod ? s.toteng : 0.0; prod_acc += accept && in_prod;
        const double te = in_prod ? s.toteng : 0.0;
        s.sum_e += te; s.sum_e2 = fma(te, te, s.sum_e2); nstat += in_prod;
        racc = nacc;
        if (++adj_count == s.nadj) {                                // mod(I, NADJ) == 0, :766-777
            adj_count = 0;
            const double ratio = (double)((float)nacc / (float)s.nadj);
            if (ratio > 0.5) { s.dposmax *= (1.0 + s.padj); s.dhoekmax *= (1.0 + s.padj); }
            else if (ratio < 0.5) { s.dposmax *= (1.0 - s.padj); s.dhoekmax *= (1.0 - s.padj); }
            s.dposmax = fmax(s.dposmax, s.dposmin); s.dhoekmax = fmax(s.dhoekmax, s.dhoekmin);
            nacc = 0;
        }
        ++done;
        __syncwarp();
    }
    {
        const int n_mut = g.o_sites + SITE_DOUBLES * N;
        for (int qq = lane; qq < n_mut; qq += 32) gimg[qq] = img[qq];
    }
    if (lane == 0) {
        s.step += done; s.rng_n += (unsigned long long)done; s.nsuc += nsuc; s.naccept = nacc; s.nclamp += nclamp; s.uused += uused; s.prod_acc += prod_acc;
        s.nstat += nstat; s.sitepairs += (long long)sitepairs;
        if (done > 0) s.ratio = (double)((float)racc / (float)s.nadj);
        ctl[c] = s;
        if (out_steps) out_steps[c] = done;
        vcnt[N] = (int)nrebuild;   // diagnostic: list rebuilds in this launch
        vcnt[N + 1] = lists_on && !need_rebuild ? 1 : 0;
    }
}

}  // namespace mcb

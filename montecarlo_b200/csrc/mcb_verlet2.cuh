// mcb_verlet2.cuh — k_run_vl2<W>: k_run_vl with W warps per chain and W consecutive steps in flight.
//
// A box whose matrix lives in global memory needs ~45-55 KB of shared memory per chain, so an SM holds four chains
// and, with one warp per chain, one warp per scheduler: every global-memory and shuffle latency of a step is
// exposed.  Here warp w evaluates step k+w of the same chain against the same committed state and posts dE, its
// flags and its trial position.  After ONE exchange barrier every warp answers, for all ordered pairs (ga < gb) of
// the batch at once (four lanes per pair, one distance test each), whether step ga's move would change anything
// step gb looked at: the same molecule; molecule i_ga inside the cutoff of gb's trial position before or after
// i_ga's own move; molecule i_ga inside the cutoff of gb's molecule as committed, before or after (that entry of
// the old row feeds dE).  Step k+w stands if all steps before it stand and no ACCEPTED one of them fired the test;
// the standing accepted steps touch disjoint molecules and matrix entries and are committed all at once, each by
// its own warp.  The first step that does not stand, and everything after it, is evaluated again in the next
// batch with the same uniforms, so the chain is the sequential one, decision for decision.  In a box of 240
// molecules a batch of four commits 3.78 steps on average.
//
// Every warp works in a small area (VL_K + nsolp partner slots: what a neighbour list can yield); a step that
// cannot use its list (trial position beyond skin/2, overflowed list, lists off) is evaluated by warp 0 only, with
// the full scan, in the chain's global work area - the batch is then that one step.
//
// With 512 chains on 148 SMs the SMs that hold four chains are ~88 % issue-bound (profiles/vl2_image_r02.txt), so
// the loop is written for few instructions rather than for short latency: work on the rarely used second half of a
// candidate list sits behind a warp-uniform flag, what a commit writes is only formed for committed steps, and the
// thread id and the matrix pointer are pinned in registers (the compiler re-derives them at every use otherwise).
#pragma once

#include "mcb_verlet.cuh"

namespace mcb {

struct Vl2X { double dE; int flags, nn; };   // flags: 1 accept, 2 used8, 4 clamped, 8 far, 16 evaluated
struct Vl2T { double ncx, ncy, ncz; int i, pad; };   // a step's molecule and trial centre of mass, for the dependency tests
constexpr int VL2_MAX_W = 8;

__host__ __device__ inline int vl2_small_slots(const Geo &g) { return VL_K + g.nsolp; }
__host__ __device__ inline int vl2_list_doubles(int slots) { return (((slots + 1) / 2) + 1) & ~1; }   // ints, padded to 16 bytes
__host__ __device__ inline int vl2_smem_doubles(const Geo &g, int w, bool img_smem = true)
{
    const int small = vl2_small_slots(g);
    return (img_smem ? g.stage_doubles : 0) + w * (vl2_list_doubles(small) + 4 * small) + 64 + 6 * 64 + 32 + 2 * VL2_MAX_W + 4 * VL2_MAX_W;
}

// IMG_SMEM = false: the chain's image (centres of mass, angles, site cache, parameter rows) stays in global memory and
// is read through L1/L2; shared memory then only holds the work areas and the ring (~6 KB per chain with one warp), so an
// SM holds up to 16 chains where the staged image allows 4 (N = 240) or 1 (N = 810).  That is the variant for MANY chains
// per GPU (more chains than all SMs can keep resident with staged images) and allows W = 1: one warp per chain, latency
// hidden by the other chains' warps instead of by steps in flight.  All warps of a chain sit in one block, so a commit
// through global memory is visible to the chain's other warps after the batch's closing barrier.
template <bool COUL, int W, bool IMG_SMEM = true>
__global__ void __launch_bounds__(32 * W, IMG_SMEM ? (W <= 2 ? 1 : 4) : 16 / W) k_run_vl2(Geo g, SiteRel rel, double *images, double *matg, double *workg,
                                                                    ChainCtl *ctl, long long nsteps, long long *out_steps, VerletBufs vb)
{
    static_assert(W >= 1 && W <= VL2_MAX_W && (IMG_SMEM ? W >= 2 : W <= 4), "1..8 warps per chain");
    constexpr int NT = 32 * W;
    extern __shared__ __align__(16) double smem[];
    int tid = threadIdx.x;
    asm volatile("" : "+r"(tid));   // (kept in a register: the compiler otherwise reads the special register again at every use)
    const int lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x;
    const int N = g.N, Np = g.Np, MS = g.MS, nsolp = g.nsolp;
    const unsigned lt = (1u << lane) - 1u;
    double *gimg = images + (size_t)c * g.image_doubles;
    double *const img = IMG_SMEM ? smem : gimg;
    const int stage = IMG_SMEM ? g.stage_doubles : 0;
    if (IMG_SMEM) {
        const double2 *src = reinterpret_cast<const double2 *>(gimg);
        double2 *dst = reinterpret_cast<double2 *>(smem);
        for (int k = tid; k < g.stage_doubles / 2; k += NT) dst[k] = src[k];
    }
    double *mat_ = matg + (size_t)c * N * MS;
    asm volatile("" : "+l"(mat_));    // (same: one 64-bit register pair instead of the multiply chain at every use)
    __builtin_assume(__isGlobal(mat_));
    double *const mat = mat_;
    double *const com = img + g.o_com, *const hoek = img + g.o_hoek, *const energy = img + g.o_energy;
    double *const sites = img + g.o_sites;
    const double *const tabAB = img + g.o_tabAB, *const tabQ = img + g.o_tabQ;
    const int a = lane & 3, sub = lane >> 2;
    // shared-memory plan behind the image: one small work area per warp (VL_K + nsolp partner slots: what a list can
    // yield), ring, row sums, exchange.  A step that needs the full scan runs on warp 0 in the chain's global work area.
    const int small = vl2_small_slots(g);
    double *p = smem + stage + warp * (vl2_list_doubles(small) + 4 * small);
    int *const list = reinterpret_cast<int *>(p);
    double *const epart = p + vl2_list_doubles(small);
    int *const list_g = reinterpret_cast<int *>(workg + (size_t)c * g.work_doubles + g.o_list);
    double *const epart_g = workg + (size_t)c * g.work_doubles + g.o_epart;
    p = smem + stage + W * (vl2_list_doubles(small) + 4 * small);
    double *const ring_thr = p; p += 64;
    double *const ring_rv = p; p += 6 * 64;
    int *const ring_i = reinterpret_cast<int *>(p); p += 32;
    Vl2X *const xch = reinterpret_cast<Vl2X *>(p); p += 2 * VL2_MAX_W;
    Vl2T *const tpos = reinterpret_cast<Vl2T *>(p);
    unsigned short *const vidx = vb.idx + (size_t)c * N * VL_K;
    int *const vcnt = vb.cnt + (size_t)c * (N + 2);
    double *const vref = vb.ref + (size_t)c * 3 * Np;
    const double rl = g.cutoff + vb.skin, rl2 = rl * rl, h2 = 0.25 * vb.skin * vb.skin;
    ChainCtl s = ctl[c];
    for (int q = lane; q < small; q += 32) list[q] = q < nsolp ? N + q : 0;   // solute slots first; the rest valid indices
    if (warp == 0) for (int q = lane; q < nsolp; q += 32) list_g[q] = N + q;
    __syncthreads();
    // TOTENG carries over from the previous launch (ChainCtl).  The O(N^2) pass over the stored matrix that re-anchors
    // it - so that rounding of the running sum cannot pile up for ever - only runs when this launch crosses a multiple of
    // 2^20 steps of the chain's stream: at 4000 steps per launch it used to be a quarter of the launch time.
    if (((s.rng_n + (unsigned long long)max(nsteps, 0LL)) >> 20) != (s.rng_n >> 20)) {
        double *const colsum = epart;   // (this warp's work area is free until the first step)
        double acc = 0.0;
        for (int j0 = 0; j0 < N; j0 += NT) {
            const int j = j0 + tid;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            if (j < N) {
                int i = 0;
                for (; i + 3 < N; i += 4) {
                    a0 += mat[(size_t)i * MS + j]; a1 += mat[(size_t)(i + 1) * MS + j];
                    a2 += mat[(size_t)(i + 2) * MS + j]; a3 += mat[(size_t)(i + 3) * MS + j];
                }
                for (; i < N; ++i) a0 += mat[(size_t)i * MS + j];
                acc += fma(0.5, (a0 + a1) + (a2 + a3), energy[j]);
            }
        }
        acc = warp_sum(acc);
        if (lane == 0) colsum[0] = acc;
        __syncthreads();
        double tot = 0.0;
        for (int w = 0; w < W; ++w) tot += (smem + stage + w * (vl2_list_doubles(small) + 4 * small) + vl2_list_doubles(small))[0];
        s.toteng = tot;
        __syncthreads();
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
    const int kk = isang ? rr - 3 : rr;
    double *const base = (isang ? hoek : com) + kk * Np;
    const double period = isang ? TAU_D : g.boxl2, half = isang ? PI_D : g.boxl;
    const double inv_half = isang ? 1.0 / PI_D : g.inv_boxl;
    int adj_count = (int)(s.step % s.nadj);
    int done = 0;
    unsigned nsuc = 0, nacc = (unsigned)s.naccept, racc = 0, nclamp = 0, n_used8 = 0, prod_acc = 0, nrebuild = 0;
    unsigned long long sum_nn = 0;
    const unsigned long long step0 = (unsigned long long)s.step;
    const int nsteps_i = (int)min(nsteps, (long long)0x3fffffff);
    int ring_hi = 0;                                               // the ring holds the launch-local steps [ring_hi - 64, ring_hi)
    const unsigned long long pos0 = s.rng_n;   // position in the uniform stream (steps since mcb_seed_chains), see k_run_spec
    const int ring0 = (int)(pos0 & 63ULL);
    const long long pf = s.start_prod - (long long)step0 - 1;     // launch-local index of the first production step (:760)
    const int prod_from = pf <= 0 ? 0 : (pf > 0x3fffffff ? 0x7fffffff : (int)pf);
    // list of a later step's molecule, fetched ahead: count, candidates, build position; p_step says for which step
    int p_cnt = -1, p_idx[2] = {0, 0}, p_step = -1;
    double p_rfx = 0.0, p_rfy = 0.0, p_rfz = 0.0;

    // the ordered pair (ga < gb) of steps whose dependency this lane tests after the exchange: pair index = lane / 4
    int dp_gb = 1;
    while (dp_gb * (dp_gb + 1) / 2 <= (lane >> 2)) ++dp_gb;
    const int dp_ga = (lane >> 2) - dp_gb * (dp_gb - 1) / 2;
    Vl2X pend[W];
    int pend_n = 0, pend_done = 0;
    auto sum_stats = [&]() {   // TOTENG, production and energy sums of the steps committed by the previous batch, in step order
#pragma unroll
        for (int gg = 0; gg < W; ++gg) {
            if (gg < pend_n) {
                const Vl2X x = pend[gg];
                const bool acc = (x.flags & 1) != 0;
                n_used8 += (x.flags >> 1) & 1;
                nclamp += (x.flags >> 2) & 1;
                sum_nn += (unsigned)x.nn;
                s.last_delta = x.dE;
                const bool in_prod = pend_done + gg >= prod_from;    // selects instead of branches: adding 0.0 changes nothing
                nsuc += acc;
                s.toteng += acc ? x.dE : 0.0;
                s.production += acc && in_prod ? s.toteng : 0.0; prod_acc += acc && in_prod;
                const double te = in_prod ? s.toteng : 0.0;
                s.sum_e += te; s.sum_e2 = fma(te, te, s.sum_e2);
            }
        }
        pend_n = 0;
    };
    while (done < nsteps_i) {
        if (warp == 0) sum_stats();
        if (need_rebuild) {
            // ---- (re)build every list from the current centres of mass, which become the reference positions
            for (int j = tid; j < 3 * Np; j += NT) vref[j] = com[j];
            for (int m = warp; m < N; m += W) {
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
            p_step = -1;
            nrebuild++;
            __syncthreads();
        }
        if (done + 2 * W > ring_hi) {                                   // pre-draw 32 steps, one per lane (as k_run_spec); warp 0 fills
            if (warp == 0) {
                const unsigned long long sidx = pos0 + (unsigned)(ring_hi + lane);
                const int e = (int)(sidx & 63ULL);
                const double u0 = stream_uniform(s.rng_key, 8ULL * sidx);
                int ii = (int)(u0 * (double)N) + 1;                 // MCINIT, MonteCarlo.F90:697-698
                if (g.quirk31 && ii == 31) ii = 30;
                ring_i[e] = min(ii, N) - 1;
#pragma unroll
                for (int q = 0; q < 6; ++q) {
                    const double rvu = (stream_uniform(s.rng_key, 8ULL * sidx + 1 + q) - 0.5) * 2.0;
                    ring_rv[6 * e + q] = (q < 3 ? INV_SQRT3 : 1.0) * rvu;
                }
                ring_thr[e] = log(stream_uniform(s.rng_key, 8ULL * sidx + 7)) * (1.0 / LN_E_F32);
            }
            ring_hi += 32;
            __syncthreads();
        }
        // a batch never runs past the launch or past a step-size adaptation
        const int limit = min(W, min(nsteps_i - done, s.nadj - adj_count));
        const bool active = warp < limit;
        const int k = done + warp;
        const int ring_e = (ring0 + k) & 63;
        const int i = ring_i[ring_e];
        const double thr = ring_thr[ring_e];
        const double e_old = energy[i];
        // ---- candidates of molecule i (fetched during an earlier batch if the guess of this step's index was right)
        if (lists_on && p_step != k) {
            p_cnt = vcnt[i];
            p_idx[0] = (int)vidx[(size_t)i * VL_K + lane]; p_idx[1] = (int)vidx[(size_t)i * VL_K + 32 + lane];
            p_rfx = vref[i]; p_rfy = vref[Np + i]; p_rfz = vref[2 * Np + i];
        }
        const int cnt = lists_on ? p_cnt : -1;
        // the list's second half (candidates 32..63) is rarely in use - a list holds ~24 - and the same on every lane:
        // everything that concerns it sits behind this flag instead of being computed on empty slots
        const bool two = cnt > 32;
        int cj[2];
        double cold[2];
        cj[0] = lane < cnt ? p_idx[0] : -1;
        cj[1] = two && lane + 32 < cnt ? p_idx[1] : -1;
        const double rfx = p_rfx, rfy = p_rfy, rfz = p_rfz;
        cold[0] = cj[0] >= 0 ? mat[(size_t)i * MS + cj[0]] : 0.0;
        cold[1] = 0.0;
        if (two) cold[1] = cj[1] >= 0 ? mat[(size_t)i * MS + cj[1]] : 0.0;
        if (lists_on) {   // this warp's step of the next batch if all steps of this one are committed
            const int in = ring_i[(ring0 + k + W) & 63];
            p_cnt = vcnt[in];
            p_idx[0] = (int)vidx[(size_t)in * VL_K + lane]; p_idx[1] = (int)vidx[(size_t)in * VL_K + 32 + lane];
            p_rfx = vref[in]; p_rfy = vref[Np + in]; p_rfz = vref[2 * Np + in];
            p_step = k + W;
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
        const bool far = !lists_on || pair_r2(ncx - rfx, ncy - rfy, ncz - rfz, g) > h2;
        const bool use_list = cnt >= 0 && !far;
        // warp 0 can always evaluate (its work area holds a full scan); warp 1 only from its list
        const bool evaluate = active && (warp == 0 || use_list);
        // ---- centre-of-mass gate (:804-813) on the candidates, or on everybody (warp 0 only)
        int nn = 0;
        bool cin[2] = {false, false};
        int cslot[2] = {0, 0};
        double dE = 0.0, esol = 0.0;
        double cval[2] = {0.0, 0.0};
        bool cchg[2] = {false, false};
        bool clamped = false, used8 = false, accept = false;
        if (evaluate) {
            if (use_list) {
#pragma unroll
                for (int q = 0; q < 2; ++q) {   // every lane computes a distance (empty lanes on molecule 0), the guard selects
                    if (q == 0 || two) {
                        const int jj = cj[q] >= 0 ? cj[q] : 0;
                        const double r2 = pair_r2(com[jj] - ncx, com[Np + jj] - ncy, com[2 * Np + jj] - ncz, g);
                        cin[q] = !((cj[q] >= 0 ? r2 : DBL_HUGE) > g.cut2);
                    }
                }
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    if (q == 0 || two) {
                        const unsigned b = __ballot_sync(FULL, cin[q]);
                        cslot[q] = nsolp + nn + __popc(b & lt);
                        if (cin[q]) list[cslot[q]] = cj[q];
                        nn += __popc(b);
                    }
                }
            } else {
                for (int c0 = 0; c0 < N; c0 += 32) {
                    const int j = c0 + lane;
                    const bool in = j < N && j != i && !(pair_r2(com[j] - ncx, com[Np + j] - ncy, com[2 * Np + j] - ncz, g) > g.cut2);
                    const unsigned b = __ballot_sync(FULL, in);
                    if (in) list_g[nsolp + nn + __popc(b & lt)] = j;
                    nn += __popc(b);
                }
            }
            __syncwarp();
            // ---- site pairs (calcLJ / CalcLJ_SV), staged
            const int nslots = nsolp + nn;
            double lane_tot = 0.0;
            if (use_list) {
                // passes of two items per lane, a remainder of at most 8 slots in a pass of one (as k_run_spec)
                int sb = 0;
                for (; nslots - sb > 8; sb += 2 * 8)
                    lane_tot += items_staged<COUL, 2>(g, sites, tabAB, tabQ, list, nsolp, nslots, sb + sub, 8, a, sx, sy, sz, epart);
                if (sb < nslots)
                    lane_tot += items_staged<COUL, 1>(g, sites, tabAB, tabQ, list, nsolp, nslots, sb + sub, 8, a, sx, sy, sz, epart);
            } else {
                for (int sb = 0; sb < nslots; sb += 2 * 8)
                    lane_tot += items_staged<COUL, 2>(g, sites, tabAB, tabQ, list_g, nsolp, nslots, sb + sub, 8, a, sx, sy, sz, epart_g);
            }
            // TOTENG_new - TOTENG_old = sum(new row) + ENERGY_new(i) - sum(old row) - ENERGY(i).  The old row's non-zero
            // entries all belong to listed candidates (a partner inside the cutoff is inside cutoff + skin), so with a
            // list the old row is summed from the entries fetched above; the full scan reads the whole row.
            double old_part = lane == 0 ? e_old : 0.0;
            if (use_list) old_part += two ? cold[0] + cold[1] : cold[0];
            else for (int j = lane; j < N; j += 32) old_part += mat[(size_t)i * MS + j];
            dE = lane_tot - old_part;
            __syncwarp();
            const double *const ep = use_list ? epart : epart_g;
            if (sub < nsolp) esol = ep[4 * sub + a];
            if (nsolp > 8) {
#pragma unroll 1
                for (int p2 = sub + 8; p2 < nsolp; p2 += 8) esol += ep[4 * p2 + a];
            }
            dE = warp_sum_mma(dE, lane); esol = warp_sum_mma(esol, lane);   // (fp64 tensor-core sums, as k_run_spec)
            // ---- METROPOLIS (:720-788) on DELTA = dE
            const double exponent = dE * neg_inv_rt;
            clamped = !(dE <= DBL_HUGE);   // -Inf is not clamped (MonteCarlo.F90:421-428)
            used8 = !clamped && exponent >= -75.0 && exponent < 0.0;
            accept = !clamped && (exponent >= 0.0 || (used8 && thr <= exponent));
        }
        if (lane == 0) {
            Vl2X x;
            x.dE = dE; x.nn = nn;
            x.flags = (accept ? 1 : 0) | (used8 ? 2 : 0) | (clamped ? 4 : 0) | (far ? 8 : 0) | (evaluate ? 16 : 0);
            xch[warp] = x;
            Vl2T tp;
            tp.ncx = ncx; tp.ncy = ncy; tp.ncz = ncz; tp.i = i; tp.pad = 0;
            tpos[warp] = tp;
        }
        __syncthreads();
        Vl2X xs[W];
#pragma unroll
        for (int q = 0; q < W; ++q) xs[q] = xch[q];
        // ---- would the move of an earlier step ga of this batch, if accepted, change anything a later step gb looked at?
        //      Every warp answers for all W(W-1)/2 ordered pairs at once, four lanes per pair, one distance test each, from
        //      the trial positions the warps have just exchanged and the committed coordinates: the same molecule; molecule
        //      i_ga inside the cutoff of gb's trial position before (t = 0) or after (t = 1) its own move; inside the cutoff
        //      of gb's molecule as committed - an entry of gb's old row - before (t = 2) or after (t = 3).  (Round 1 had
        //      each warp test its own step before the exchange, re-deriving the earlier trial positions: 17 % of the batch.)
        bool hit = false;
        if (dp_gb < W) {
            const Vl2T A = tpos[dp_ga], B = tpos[dp_gb];
            const double cax = com[A.i], cay = com[Np + A.i], caz = com[2 * Np + A.i];
            const double cbx = com[B.i], cby = com[Np + B.i], cbz = com[2 * Np + B.i];
            const int t = lane & 3;
            const double px = (t & 1) ? A.ncx : cax, py = (t & 1) ? A.ncy : cay, pz = (t & 1) ? A.ncz : caz;   // step ga's molecule: committed / trial
            const double qx = (t & 2) ? cbx : B.ncx, qy = (t & 2) ? cby : B.ncy, qz = (t & 2) ? cbz : B.ncz;   // step gb: trial position / its molecule as committed
            hit = A.i == B.i || !(pair_r2(px - qx, py - qy, pz - qz, g) > g.cut2);
        }
        const unsigned hb = __ballot_sync(FULL, hit);
        // step k+q stands if it was evaluated, every step before it stands, and no ACCEPTED earlier step touched what it looked at
        int progress = 1;
        unsigned acc_valid = xs[0].flags & 1u;
#pragma unroll
        for (int q = 1; q < W; ++q) {
            unsigned dm = 0u;   // bit a: step q depends on step a (pairs (a, q) sit at nibbles q(q-1)/2 + a of the ballot)
#pragma unroll
            for (int aa = 0; aa < q; ++aa) dm |= ((hb >> (4 * (q * (q - 1) / 2 + aa))) & 0xFu) ? (1u << aa) : 0u;
            if (progress == q && (xs[q].flags & 16) && (dm & acc_valid) == 0u) {
                progress = q + 1;
                acc_valid |= (unsigned)(xs[q].flags & 1) << q;
            }
        }
        const int nslots = nsolp + nn;
        // ---- commits, all at once: the standing accepted steps moved different molecules, none of which any other
        //      standing step looked at, so they write different matrix entries and different molecules' coordinates;
        //      nothing shared is read-modified-written (the old row is summed from the matrix entries, see above)
        if (warp < progress && accept) {
            if (use_list) {
                // what the move writes for this lane's candidates is only formed here, for the steps that are committed
                // (about half of them): the SMs that hold four chains are bound by the instructions they issue
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    if (q == 0 || two) {
                        const int sl = cin[q] ? cslot[q] : 0;
                        const double2 e01 = *reinterpret_cast<const double2 *>(epart + 4 * sl), e23 = *reinterpret_cast<const double2 *>(epart + 4 * sl + 2);
                        cval[q] = cin[q] ? (e01.x + e01.y) + (e23.x + e23.y) : 0.0;
                        cchg[q] = cj[q] >= 0 && (cin[q] || cold[q] != 0.0);
                    }
                }
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    if ((q == 0 || two) && cchg[q]) {   // SS(i,j) = SS(j,i) = EN inside the cutoff (:819-820), 0 for a partner that left it
                        const int j = cj[q];
                        mat[(size_t)i * MS + j] = cval[q]; mat[(size_t)j * MS + i] = cval[q];
                    }
            } else {
                // full scan: partners that left the cutoff, then the partners inside it
                for (int j = lane; j < N; j += 32) {
                    const double old = mat[(size_t)i * MS + j];
                    if (old != 0.0 && j != i && pair_r2(com[j] - ncx, com[Np + j] - ncy, com[2 * Np + j] - ncz, g) > g.cut2) {
                        mat[(size_t)i * MS + j] = 0.0; mat[(size_t)j * MS + i] = 0.0;
                    }
                }
                for (int sl = nsolp + lane; sl < nslots; sl += 32) {
                    const int j = list_g[sl];
                    const double e = (epart_g[4 * sl] + epart_g[4 * sl + 1]) + (epart_g[4 * sl + 2] + epart_g[4 * sl + 3]);
                    mat[(size_t)i * MS + j] = e; mat[(size_t)j * MS + i] = e;
                }
            }
            if (lane >= 1 && lane <= 6) base[i] = nv;
            if (lane < 4) {
                sites[SITE_DOUBLES * i + 3 * a] = sx; sites[SITE_DOUBLES * i + 3 * a + 1] = sy; sites[SITE_DOUBLES * i + 3 * a + 2] = sz;
            }
            if (lane == 0) energy[i] = esol;
        }
        // ---- bookkeeping in step order: every warp tracks what steers the chain (accepted moves of the adaptation
        //      window, list validity); the energy statistics only live on warp 0, whose lane 0 writes them back
        nacc += __popc(acc_valid);                                  // acc_valid: the accepted steps among the first `progress`
        {
            unsigned far_m = 0u;                                     // their molecule now sits more than skin/2 from its build position
#pragma unroll
            for (int gg = 0; gg < W; ++gg) far_m |= (xs[gg].flags & 8) ? 1u << gg : 0u;
            if (lists_on && (acc_valid & far_m)) need_rebuild = true;
        }
        // the energy statistics of this batch (warp 0) are summed at the top of the next one, off the stretch where
        // every other warp waits at the batch's last barrier
#pragma unroll
        for (int gg = 0; gg < W; ++gg) pend[gg] = xs[gg];
        pend_n = progress; pend_done = done;
        racc = nacc;
        adj_count += progress;
        if (adj_count == s.nadj) {                                  // mod(I, NADJ) == 0, :766-777
            adj_count = 0;
            const double ratio = (double)((float)nacc / (float)s.nadj);
            if (ratio > 0.5) { s.dposmax *= (1.0 + s.padj); s.dhoekmax *= (1.0 + s.padj); }
            else if (ratio < 0.5) { s.dposmax *= (1.0 - s.padj); s.dhoekmax *= (1.0 - s.padj); }
            s.dposmax = fmax(s.dposmax, s.dposmin); s.dhoekmax = fmax(s.dhoekmax, s.dhoekmin);
            nacc = 0;
        }
        done += progress;
        __syncthreads();
    }
    if (warp == 0) sum_stats();
    if (IMG_SMEM) {
        const int n_mut = g.o_sites + SITE_DOUBLES * N;
        for (int qq = tid; qq < n_mut; qq += NT) gimg[qq] = img[qq];
    }
    if (tid == 0) {
        s.step += done; s.rng_n += (unsigned long long)done; s.nsuc += nsuc; s.naccept = nacc; s.nclamp += nclamp; s.prod_acc += prod_acc;
        s.uused += 7LL * done + n_used8;
        s.nstat += done > prod_from ? done - prod_from : 0;
        s.sitepairs += (long long)wpair * (long long)sum_nn + (long long)wsol * done;
        if (done > 0) s.ratio = (double)((float)racc / (float)s.nadj);
        ctl[c] = s;
        if (out_steps) out_steps[c] = done;
        vcnt[N] = (int)nrebuild;
        vcnt[N + 1] = lists_on && !need_rebuild ? 1 : 0;
    }
}

}  // namespace mcb

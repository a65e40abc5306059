// mcb200.cu — host side of libmcb200.so: the C ABI declared in include/mcb200.h.
//
// Owns the device memory of one handle (per-chain images in HBM, optional global pair matrices, per-chain
// scalars, scratch), packs the caller's AoS TYPE(vector) arrays into the SoA planes the kernels read, and
// launches the kernels of mcb_kernels.cuh on the handle's stream.  There is no CPU path: every entry point
// that computes something launches a kernel or fails.
#include "../../include/mcb200.h"
#include "mcb_kernels.cuh"
#include "mcb_verlet.cuh"
#include "mcb_verlet2.cuh"
#include "mcb_cells.cuh"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace mcb;

struct mcb_comm_state;   // NCCL side of a handle, defined with the collective at the end of this file

struct mcb_handle {
    mcb_params prm{};
    Geo g{};
    SiteRel rel{};
    int n_chains = 0, device = 0, n_sites = 0;
    bool coul = false, can_step = false, dirty = true, energy_only = false;
    bool smem_fits = false;  // a chain's image fits in shared memory (every stepping kernel but the global-image k_run_vl2 needs that)
    double *big_e = nullptr, *big_sol = nullptr, *big_out = nullptr; long long *big_n = nullptr, *big_nout = nullptr;
    long long big_blocks = 0; int big_tiles = 0;
    CellGrid cells{};       // cell list of the matrix-free energy (mcb_cells.cuh); nc == 0: not in use
    int ncell = 0;
    VerletBufs vl{};        // neighbour lists of k_run_vl (large boxes); idx == nullptr: not in use
    int vl2_w = 0;          // warps per chain of k_run_vl2: 0 not chosen yet, -1 does not fit
    bool vl2_global = false; // k_run_vl2 reads the chain's image from global memory instead of staging it
    bool no_vl2 = false;    // MCB200_NO_VL2 (read once, at mcb_create)
    bool trial_pending = false;   // mcb_trial_delta has parked a trial that mcb_commit has not resolved yet
    cudaStream_t stream = nullptr;
    double *images = nullptr, *matg = nullptr, *work = nullptr;
    ChainCtl *ctl = nullptr;
    // scratch for entry-point inputs/outputs
    int *d_rsolv = nullptr; double *d_vec = nullptr; double *d_out = nullptr; long long *d_ll = nullptr;
    unsigned char *d_acc = nullptr; double *d_ubuf = nullptr; size_t ubuf_cap = 0;
    double *pin = nullptr; size_t pin_doubles = 0;   // pinned staging (solute tables)
    double *d_stage = nullptr; size_t stage_doubles = 0;   // device staging of the caller's AoS coordinate arrays
    std::vector<double> site_tab;                     // [4][3] padded
    std::vector<double> tab0AB, tab0Q;                // type-0 tables
    long long launches = 0;
    size_t smem_bytes = 0, smem_spec_bytes = 0;
    int lookahead = 4, spec_w = 1, spec_gl = 32;
    mcb_comm_state *comm = nullptr;   // NCCL communicator of the end-of-run collective (mcb_comm_init), or none
    std::string err;
};

static void comm_free(mcb_handle *h);

// error text of a failed mcb_create: one per host thread, so that concurrent creates do not write the same string
static thread_local std::string g_create_err;

#define CK(h, call)                                                                                   \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            (h)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                            \
            return -(int)e_;                                                                          \
        }                                                                                             \
    } while (0)

static int arg_err(mcb_t *h, const char *msg) { h->err = msg; return 1; }

static size_t round_up(size_t v, size_t m) { return (v + m - 1) / m * m; }

// largest x with sqrt(x) <= cutoff: `r2 > x` then decides exactly like `sqrt(r2) > cutoff` (MonteCarlo.F90:812)
static double exact_cut2(double cutoff)
{
    double x = cutoff * cutoff;
    while (std::sqrt(x) > cutoff) x = std::nextafter(x, 0.0);
    while (std::sqrt(std::nextafter(x, INFINITY)) <= cutoff) x = std::nextafter(x, INFINITY);
    return x;
}

// (A, B, QQ) of one site pair: LJ_FORMULA's E = sqrt(E1*E2), S = (S1+S2)/2 (LennardJones.f90:173-174)
static void pair_consts(const double *ta, const double *tb, double conv, double &A, double &B, double &Q)
{
    long double e = sqrtl((long double)ta[1] * tb[1]);
    long double s = ((long double)ta[2] + tb[2]) / 2.0L;
    long double s2 = s * s, s6 = s2 * s2 * s2;
    A = (double)(4.0L * e * s6 * s6);
    B = (double)(4.0L * e * s6);
    Q = (double)((long double)ta[0] * tb[0] * conv);
}

static void setup_geo(mcb_handle *h)
{
    const mcb_params &p = h->prm;
    Geo &g = h->g;
    g.N = p.n_mol; g.nsol = p.n_sol; g.nsolp = (p.n_sol + 3) / 4;
    // plane stride of the SoA coordinate arrays: even (16-byte alignment of what follows) and NOT a multiple of 16
    // doubles - with a stride of 32 doubles (N = 30 rounded up) the six planes of one molecule, read by six lanes of
    // one shared-memory load in MCINIT, all sit on the same bank
    g.Np = (int)round_up(p.n_mol, 2);
    if (g.Np % 16 == 0) g.Np += 2;
    g.MS = (g.N % 2) ? g.N : g.N + 1;
    g.o_com = 0; g.o_hoek = 3 * g.Np; g.o_energy = 6 * g.Np; g.o_sites = 7 * g.Np;
    g.o_tabAB = g.o_sites + SITE_DOUBLES * (g.N + g.nsolp);
    g.o_tabQ = g.o_tabAB + (1 + g.nsolp) * 4 * TAB_ROW;
    int end_tab = g.o_tabQ + (1 + g.nsolp) * 16;
    g.o_mat = end_tab;
    g.o_newrow = 0; g.o_list = g.Np;
    g.o_epart = (int)round_up(g.o_list + (g.Np + g.nsolp + 1) / 2, 2);
    g.o_pend = g.o_epart + 4 * (g.Np + g.nsolp);
    g.o_touch = (int)round_up(g.o_pend + (sizeof(Pending) + 7) / 8, 2);
    g.work_doubles = (int)round_up(g.o_touch + ((g.N + 31) / 32 + 1) / 2, 2);
    g.spec_group_doubles = g.work_doubles;
    size_t with_mat = round_up((size_t)end_tab + (size_t)g.N * g.MS, 2);
    // the matrix rides in shared memory while a chain's footprint leaves room for >= 7 chains per SM
    bool mat_smem = p.matrix_mode != 1 && (with_mat + 4 * (size_t)(g.work_doubles + SPECX_DOUBLES)) * 8 <= 32 * 1024;
    g.mat_in_image = mat_smem ? 1 : 0;
    g.image_doubles = (int)(mat_smem ? with_mat : (size_t)end_tab);
    g.stage_doubles = g.image_doubles;
    g.quirk31 = p.quirk_31;
    g.boxl = p.boxl; g.boxl2 = p.boxl * 2.0; g.inv_boxl = 1.0 / p.boxl; g.cut2 = exact_cut2(p.cutoff); g.conv = p.conv; g.cutoff = p.cutoff;
    // steps in flight per chain -> (warps per chain, lanes per step); MCB200_TUNE="W,GL" overrides (tuning only)
    h->lookahead = p.lookahead == 1 || p.lookahead == 2 || p.lookahead == 4 || p.lookahead == 8 ? p.lookahead : 1;
    switch (h->lookahead) {   // measured on configs[1] (profiles/tune_r01*.txt): once the scalar part of a step was made cheap, no look-ahead wins
        case 4: h->spec_w = 1; h->spec_gl = 8; break;
        case 8: h->spec_w = 2; h->spec_gl = 8; break;
        case 2: h->spec_w = 1; h->spec_gl = 16; break;
        default: h->spec_w = 1; h->spec_gl = 32; break;
    }
    if (const char *t = getenv("MCB200_TUNE")) {
        int w = 0, gl = 0;
        if (sscanf(t, "%d,%d", &w, &gl) == 2 && ((w == 1 && (gl == 8 || gl == 16 || gl == 32)) || (w == 2 && (gl == 8 || gl == 16 || gl == 32)) || (w == 4 && gl == 32))) {
            h->spec_w = w; h->spec_gl = gl; h->lookahead = w * 32 / gl;
        }
    }
    h->smem_bytes = ((size_t)g.stage_doubles + g.work_doubles) * 8;
    // image + look-ahead work areas + exchange + ring (threshold, moves, picks) + tot[] + sin/cos cache + ring of sin/cos + moved bits
    h->smem_spec_bytes = ((size_t)g.stage_doubles + (size_t)h->lookahead * (g.work_doubles + SPECX_DOUBLES) + 64 + 6 * 64 + 32 + g.Np + 6 * g.Np + 6 * 64 + (g.N + 63) / 64 + 2) * 8;
    // a box whose matrix lives in global memory steps with the neighbour-list kernels, which need less shared memory
    // than k_run_spec's plan (no sin/cos cache); k_run_spec is then only the list-free fall-back and may not fit
    h->smem_fits = (g.mat_in_image ? h->smem_spec_bytes : (size_t)vl_smem_doubles(g) * 8) <= 227 * 1024;
    // a larger box still steps with the built-in stream: k_run_vl2 with the image left in global memory (neighbour lists:
    // 32 < N < 65536), as long as the start-up kernel's per-warp row buffers fit
    const bool prep_fits = (size_t)prepare_smem_doubles(g, false) * 8 <= 227 * 1024;
    const bool lists_possible = !g.mat_in_image && g.N > 32 && g.N < 65536 && h->spec_w == 1 && h->spec_gl == 32 &&
                                !getenv("MCB200_NO_VERLET") && !getenv("MCB200_NO_VL2");
    h->can_step = h->smem_fits || (lists_possible && prep_fits);
}

extern "C" const char *mcb_last_error(const mcb_t *h) { return h ? h->err.c_str() : g_create_err.c_str(); }

extern "C" int mcb_destroy(mcb_t *h)
{
    if (!h) return 0;
    cudaSetDevice(h->device);
    comm_free(h);
    cudaFree(h->images); cudaFree(h->matg); cudaFree(h->work); cudaFree(h->ctl);
    cudaFree(h->d_rsolv); cudaFree(h->d_vec); cudaFree(h->d_out); cudaFree(h->d_ll); cudaFree(h->d_acc);
    cudaFree(h->d_ubuf); cudaFree(h->d_stage);
    cudaFree(h->vl.idx); cudaFree(h->vl.cnt); cudaFree(h->vl.ref);
    cudaFree(h->big_e); cudaFree(h->big_sol); cudaFree(h->big_out); cudaFree(h->big_n); cudaFree(h->big_nout);
    cudaFree(h->cells.count); cudaFree(h->cells.cell_of); cudaFree(h->cells.order);
    if (h->pin) cudaFreeHost(h->pin);
    delete h;
    return 0;
}

template <typename K>
static int set_smem_attr(mcb_handle *h, K kernel, size_t bytes)
{
    if (bytes > 48 * 1024)
        CK(h, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return 0;
}

extern "C" int mcb_create(mcb_t **out, const mcb_params *p)
{
    if (!out || !p) { g_create_err = "mcb_create: null argument"; return 1; }
    *out = nullptr;
    if (p->n_mol < 1 || p->n_chains < 1 || p->n_sites < 1 || p->n_sites > MCB_MAX_SITES || p->n_sol < 0 ||
        !(p->boxl > 0.0) || !(p->cutoff > 0.0) || !p->site_rel || !p->site_tab) {
        g_create_err = "mcb_create: bad parameter (n_mol, n_chains >= 1; 1 <= n_sites <= 4; boxl, cutoff > 0)";
        return 1;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_err = std::string("mcb_create: no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e);
        return -(int)(e != cudaSuccess ? e : cudaErrorNoDevice);
    }
    mcb_handle *h = new mcb_handle;
    h->prm = *p;
    h->n_chains = p->n_chains; h->n_sites = p->n_sites; h->coul = p->conv != 0.0;
    if (p->device >= 0) h->device = p->device; else cudaGetDevice(&h->device);
    auto fail = [&](int rc) { g_create_err = h->err; mcb_destroy(h); return rc; };
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail(-(int)e_); } } while (0)
    CKC(cudaSetDevice(h->device));
    setup_geo(h);
    const Geo &g = h->g;
    // site tables, padded to 4 sites with parameter-free sites (they contribute exactly 0)
    h->site_tab.assign(12, 0.0);
    memset(&h->rel, 0, sizeof(h->rel));
    h->rel.n_sites = p->n_sites;
    for (int a = 0; a < p->n_sites; ++a)
        for (int k = 0; k < 3; ++k) { h->rel.r[3 * a + k] = p->site_rel[3 * a + k]; h->site_tab[3 * a + k] = p->site_tab[3 * a + k]; }
    for (int a = p->n_sites; a < 4; ++a) { h->rel.r[3 * a] = 1.0e4 + 7.0 * a; h->rel.r[3 * a + 1] = 2.0e4; h->rel.r[3 * a + 2] = 3.0e4; }
    h->tab0AB.assign(4 * TAB_ROW, 0.0); h->tab0Q.assign(16, 0.0);
    for (int a = 0; a < p->n_sites; ++a)
        for (int b = 0; b < p->n_sites; ++b) {
            double A, B, Q;
            pair_consts(&h->site_tab[3 * a], &h->site_tab[3 * b], p->conv, A, B, Q);
            h->tab0AB[a * TAB_ROW + 2 * b] = A; h->tab0AB[a * TAB_ROW + 2 * b + 1] = B; h->tab0Q[a * 4 + b] = Q;
        }
    const size_t nc = (size_t)h->n_chains;
    CKC(cudaMalloc(&h->images, nc * g.image_doubles * sizeof(double)));
    CKC(cudaMemset(h->images, 0, nc * g.image_doubles * sizeof(double)));
    // energy-only mode: a chain that does not fit in shared memory, or pair matrices that do not fit in HBM
    // (BASELINE config 5).  Only mcb_total_energy (matrix-free kernels) is available then.
    h->energy_only = !h->can_step;
    if (!g.mat_in_image && !h->energy_only) {
        size_t bytes = nc * (size_t)g.N * g.MS * sizeof(double);
        size_t free_b = 0, total_b = 0;
        CKC(cudaMemGetInfo(&free_b, &total_b));
        if (bytes > free_b / 2) h->energy_only = true;
        else { CKC(cudaMalloc(&h->matg, bytes)); CKC(cudaMemset(h->matg, 0, bytes)); }
    }
    if (h->energy_only) {
        h->can_step = false;
        h->big_tiles = (g.N + BIG_TILE - 1) / BIG_TILE;
        h->big_blocks = (long long)h->big_tiles * (h->big_tiles + 1) / 2;
        CKC(cudaMalloc(&h->big_e, sizeof(double) * h->big_blocks));
        CKC(cudaMalloc(&h->big_n, sizeof(long long) * h->big_blocks));
        CKC(cudaMalloc(&h->big_sol, sizeof(double) * h->big_tiles));
        CKC(cudaMalloc(&h->big_out, sizeof(double) * 2));
        CKC(cudaMalloc(&h->big_nout, sizeof(long long)));
        // cells at least one cutoff wide (a hair more, so that the rounding of the binning cannot separate a pair
        // inside the cutoff by two cells); fewer than 3 per side: the all-pairs kernel stays
        int ncs = (int)floor(g.boxl2 / (g.cutoff * (1.0 + 1e-9)));
        if (ncs > CELL_NC_MAX) ncs = CELL_NC_MAX;
        if (ncs >= 3 && g.cutoff > 0.0 && !getenv("MCB200_NO_CELLS")) {
            h->cells.nc = ncs;
            h->cells.scale = (double)ncs / g.boxl2;
            h->ncell = ncs * ncs * ncs;
            // one allocation: count[ncell] fill[ncell] flag[1] pad start[ncell + 1]
            CKC(cudaMalloc(&h->cells.count, sizeof(int) * ((size_t)3 * h->ncell + 4)));
            h->cells.fill = h->cells.count + h->ncell;
            h->cells.flag = h->cells.fill + h->ncell;
            h->cells.start = h->cells.flag + 2;
            CKC(cudaMalloc(&h->cells.cell_of, sizeof(int) * (size_t)g.N));
            CKC(cudaMalloc(&h->cells.order, sizeof(int) * (size_t)g.N));
        }
    }
    // neighbour lists for every box whose pair matrix lives in global memory: with the next step's list fetched one
    // step ahead they win from the smallest such box on (N = 240: 6.0e10 vs 4.75e10 site pairs/s for the full scan,
    // N = 810: 1.2e10 vs 3e9; profiles/other_configs_r01.txt)
    int vl_min_n = 33;
    if (const char *mn = getenv("MCB200_VERLET_MIN_N")) vl_min_n = atoi(mn);   // tuning only
    h->no_vl2 = getenv("MCB200_NO_VL2") != nullptr;
    if (!h->smem_fits) vl_min_n = 33;   // (the lists are the only way such a box steps)
    if (!h->energy_only && !g.mat_in_image && g.N >= vl_min_n && g.N > 32 && g.N < 65536 && h->spec_w == 1 && h->spec_gl == 32 && !getenv("MCB200_NO_VERLET")) {
        // large box, one step per warp: neighbour lists for k_run_vl (mcb_verlet.cuh)
        h->vl.skin = 2.0;
        if (const char *sk = getenv("MCB200_SKIN")) { double v = atof(sk); if (v > 0.0) h->vl.skin = v; }   // tests shrink it to force rebuilds
        CKC(cudaMalloc(&h->vl.idx, nc * (size_t)g.N * VL_K * sizeof(unsigned short)));
        CKC(cudaMalloc(&h->vl.cnt, nc * (size_t)(g.N + 2) * sizeof(int)));
        CKC(cudaMalloc(&h->vl.ref, nc * (size_t)3 * g.Np * sizeof(double)));
        CKC(cudaMemset(h->vl.cnt, 0, nc * (size_t)(g.N + 2) * sizeof(int)));
    }
    CKC(cudaMalloc(&h->work, nc * g.work_doubles * sizeof(double)));
    CKC(cudaMemset(h->work, 0, nc * g.work_doubles * sizeof(double)));
    CKC(cudaMalloc(&h->ctl, nc * sizeof(ChainCtl)));
    CKC(cudaMalloc(&h->d_rsolv, nc * sizeof(int)));
    CKC(cudaMalloc(&h->d_vec, nc * 6 * sizeof(double)));
    CKC(cudaMalloc(&h->d_out, nc * 3 * sizeof(double)));
    CKC(cudaMalloc(&h->d_ll, nc * 2 * sizeof(long long)));
    CKC(cudaMalloc(&h->d_acc, nc));
    h->pin_doubles = nc * (size_t)std::max(6 * g.Np, g.image_doubles - g.o_sites - SITE_DOUBLES * g.N) + 64;
    CKC(cudaMallocHost(&h->pin, h->pin_doubles * sizeof(double)));
    // default Metropolis parameters = Tools/files/config0.ini:4-15; chain c is seeded with c+1
    std::vector<ChainCtl> ctl(nc);
    for (size_t c = 0; c < nc; ++c) {
        ChainCtl &s = ctl[c];
        memset(&s, 0, sizeof(s));
        s.temperature = 300.0; s.dposmax = 0.30; s.dposmin = 1e-10; s.dhoekmax = 0.16 * PI_D; s.dhoekmin = 1e-10 * PI_D;
        s.padj = 0.05; s.nadj = 100; s.start_prod = 0;
        unsigned long long x = (unsigned long long)(c + 1) + 0x9E3779B97F4A7C15ULL;
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL; x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL; x ^= x >> 31;
        s.rng_key = x; s.rng_n = 0;
    }
    CKC(cudaMemcpy(h->ctl, ctl.data(), nc * sizeof(ChainCtl), cudaMemcpyHostToDevice));
    if (!h->energy_only) {
        const size_t pb = (size_t)prepare_smem_doubles(g, prepare_can_stage(g)) * 8;
        if (prepare_can_stage(g) ? (set_smem_attr(h, k_prepare<true, true>, pb) || set_smem_attr(h, k_prepare<false, true>, pb))
                                 : (set_smem_attr(h, k_prepare<true, false>, pb) || set_smem_attr(h, k_prepare<false, false>, pb))) return fail(-1);
    }
    if (h->can_step && h->smem_fits) {
        int rc = 0;
        rc |= set_smem_attr(h, k_run_steps<false, false>, h->smem_bytes); rc |= set_smem_attr(h, k_run_steps<false, true>, h->smem_bytes);
        rc |= set_smem_attr(h, k_run_steps<true, false>, h->smem_bytes);  rc |= set_smem_attr(h, k_run_steps<true, true>, h->smem_bytes);
#define SPEC_ATTR1(GL, W, S) rc |= set_smem_attr(h, k_run_spec<false, false, GL, W, S>, h->smem_spec_bytes); rc |= set_smem_attr(h, k_run_spec<false, true, GL, W, S>, h->smem_spec_bytes); \
                            rc |= set_smem_attr(h, k_run_spec<true, false, GL, W, S>, h->smem_spec_bytes);  rc |= set_smem_attr(h, k_run_spec<true, true, GL, W, S>, h->smem_spec_bytes);
#define SPEC_ATTR(GL, W) SPEC_ATTR1(GL, W, true) SPEC_ATTR1(GL, W, false)
        if (h->smem_spec_bytes <= 227 * 1024) {
        SPEC_ATTR(32, 1)
#ifndef MCB_DEV_BUILD   // (-DMCB_DEV_BUILD: only the default variant, for a quick compile while tuning)
        SPEC_ATTR(16, 1) SPEC_ATTR(8, 1) SPEC_ATTR(16, 2) SPEC_ATTR(8, 2) SPEC_ATTR(32, 2) SPEC_ATTR(32, 4)
#endif
        }
#undef SPEC_ATTR
#undef SPEC_ATTR1
        if (rc) return fail(rc);
    }
#undef CKC
    *out = h;
    // an empty solute section still needs the type-0 table in every image
    return mcb_set_solute(h, 0, h->n_chains, nullptr, nullptr, 1);
}

extern "C" int mcb_set_stream(mcb_t *h, void *cuda_stream) { if (!h) return 1; h->stream = (cudaStream_t)cuda_stream; return 0; }
extern "C" int64_t mcb_launch_count(const mcb_t *h) { return h ? h->launches : 0; }

// ------------------------------------------------------------------------------------------------ state
static int check_range(mcb_t *h, int chain0, int nchains)
{
    if (!h) return 1;
    if (chain0 < 0 || nchains < 1 || chain0 + nchains > h->n_chains) return arg_err(h, "chain range outside [0, n_chains)");
    return 0;
}

extern "C" int mcb_set_solute(mcb_t *h, int chain0, int nchains, const double *solute, const double *sol_tab, int broadcast)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    if ((solute == nullptr) != (sol_tab == nullptr)) return arg_err(h, "mcb_set_solute: solute and sol_tab must both be given");
    h->trial_pending = false;
    const Geo &g = h->g;
    CK(h, cudaSetDevice(h->device));
    // the tail of an image: [solute pseudo-molecule sites][tabAB][tabQ], contiguous
    const int tail0 = g.o_sites + SITE_DOUBLES * g.N, tail = g.o_mat - tail0;
    if ((size_t)nchains * tail > h->pin_doubles) return arg_err(h, "internal: staging buffer too small");
    for (int c = 0; c < nchains; ++c) {
        double *t = h->pin + (size_t)c * tail;
        double *sites = t, *ab = t + (g.o_tabAB - tail0), *tq = t + (g.o_tabQ - tail0);
        const double *xyz = solute ? solute + (broadcast ? 0 : (size_t)c * g.nsol * 3) : nullptr;
        const double *tb = sol_tab ? sol_tab + (broadcast ? 0 : (size_t)c * g.nsol * 3) : nullptr;
        memset(t, 0, sizeof(double) * tail);
        memcpy(ab, h->tab0AB.data(), 4 * TAB_ROW * sizeof(double));
        memcpy(tq, h->tab0Q.data(), 16 * sizeof(double));
        for (int p = 0; p < g.nsolp; ++p)
            for (int b = 0; b < 4; ++b) {
                int atom = 4 * p + b;
                double *s = sites + SITE_DOUBLES * p + 3 * b;
                if (xyz && atom < g.nsol) {
                    s[0] = xyz[3 * atom]; s[1] = xyz[3 * atom + 1]; s[2] = xyz[3 * atom + 2];
                    for (int a = 0; a < h->n_sites; ++a) {
                        double A, B, Q;
                        pair_consts(&h->site_tab[3 * a], tb + 3 * atom, h->prm.conv, A, B, Q);
                        ab[((1 + p) * 4 + a) * TAB_ROW + 2 * b] = A; ab[((1 + p) * 4 + a) * TAB_ROW + 2 * b + 1] = B;
                        tq[((1 + p) * 4 + a) * 4 + b] = Q;
                    }
                } else {   // padding atom: no parameters, parked far outside any box
                    s[0] = 4.0e4 + 11.0 * atom; s[1] = 5.0e4; s[2] = 6.0e4;
                }
            }
    }
    CK(h, cudaMemcpy2DAsync(h->images + (size_t)chain0 * g.image_doubles + tail0, (size_t)g.image_doubles * 8, h->pin,
                            (size_t)tail * 8, (size_t)tail * 8, nchains, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    h->dirty = true;
    return 0;
}

// device staging for mcb_set_chain_state / mcb_get_chain_state: the caller's AoS arrays travel as they are (one
// contiguous copy each) and a kernel moves them into / out of the images' SoA planes
static int ensure_stage(mcb_t *h, size_t doubles)
{
    if (doubles <= h->stage_doubles) return 0;
    cudaFree(h->d_stage); h->d_stage = nullptr; h->stage_doubles = 0;
    CK(h, cudaMalloc(&h->d_stage, doubles * sizeof(double)));
    h->stage_doubles = doubles;
    return 0;
}

extern "C" int mcb_set_chain_state(mcb_t *h, int chain0, int nchains, const double *com, const double *hoek, int broadcast)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    if (!com || !hoek) return arg_err(h, "mcb_set_chain_state: null array");
    h->vl.trust = 0;   // new coordinates: the neighbour lists of k_run_vl are rebuilt at the next launch
    h->trial_pending = false;
    const Geo &g = h->g;
    CK(h, cudaSetDevice(h->device));
    const size_t per = (size_t)g.N * 3, n = broadcast ? per : per * nchains;
    if (int rc = ensure_stage(h, 2 * n)) return rc;
    CK(h, cudaMemcpyAsync(h->d_stage, com, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaMemcpyAsync(h->d_stage + n, hoek, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    k_state_pack<<<nchains, 128, 0, h->stream>>>(g, h->images, chain0, h->d_stage, h->d_stage + n, broadcast ? 1 : 0);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    h->dirty = true;
    return 0;
}

extern "C" int mcb_get_chain_state(mcb_t *h, int chain0, int nchains, double *com, double *hoek)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    const Geo &g = h->g;
    CK(h, cudaSetDevice(h->device));
    const size_t n = (size_t)g.N * 3 * nchains;
    if (int rc = ensure_stage(h, 2 * n)) return rc;
    k_state_unpack<<<nchains, 128, 0, h->stream>>>(g, h->images, chain0, com ? h->d_stage : nullptr, hoek ? h->d_stage + n : nullptr);
    h->launches++;
    CK(h, cudaGetLastError());
    if (com) CK(h, cudaMemcpyAsync(com, h->d_stage, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (hoek) CK(h, cudaMemcpyAsync(hoek, h->d_stage + n, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

// ------------------------------------------------------------------------------------------------ energies
static int launch_prepare(mcb_t *h, int chain0, int nchains, double *d_tot, double *d_esol, long long *d_pairs)
{
    const int staged = prepare_can_stage(h->g) ? 1 : 0;
    const size_t bytes = (size_t)prepare_smem_doubles(h->g, staged != 0) * 8;
#define PREP_LAUNCH(C, S) k_prepare<C, S><<<nchains, 32 * PREP_W, bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, chain0, d_tot, d_esol, d_pairs)
    if (h->coul) { if (staged) PREP_LAUNCH(true, true); else PREP_LAUNCH(true, false); }
    else { if (staged) PREP_LAUNCH(false, true); else PREP_LAUNCH(false, false); }
#undef PREP_LAUNCH
    h->launches++;
    CK(h, cudaGetLastError());
    return 0;
}

// matrix-free path: one chain at a time, partial sums through scratch (big_e / big_n / big_sol)
static int big_energy(mcb_t *h, int c, double *tot, double *esol, long long *pairs)
{
    const Geo &g = h->g;
    double *img = h->images + (size_t)c * g.image_doubles;
    double out[2]; long long n;
    k_big_sites<<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, h->rel, img);
    if (h->coul) k_big_solute<true><<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, img, h->big_sol);
    else k_big_solute<false><<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, img, h->big_sol);
    h->launches += 2;
    bool need_all_pairs = true;
    if (h->cells.nc) {
        // candidates from a cell list; the flag says whether the box allowed it (every centre of mass inside it)
        const CellGrid &cg = h->cells;
        CK(h, cudaMemsetAsync(cg.count, 0, sizeof(int) * ((size_t)2 * h->ncell + 2), h->stream));   // count, fill, flag
        k_cell_assign<<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, img, cg);
        k_cell_scan<<<1, 1024, 0, h->stream>>>(cg, h->ncell);
        k_cell_scatter<<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, cg);
        k_cell_sort<<<(h->ncell + BIG_TILE - 1) / BIG_TILE, BIG_TILE, 0, h->stream>>>(cg, h->ncell);
        if (h->coul) k_cell_pairs<true><<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, img, cg, h->big_e, h->big_n);
        else k_cell_pairs<false><<<h->big_tiles, BIG_TILE, 0, h->stream>>>(g, img, cg, h->big_e, h->big_n);
        k_big_reduce<<<1, BIG_TILE, 0, h->stream>>>(h->big_e, h->big_n, h->big_tiles, h->big_sol, h->big_tiles, h->big_out, h->big_nout);
        h->launches += 6;
        CK(h, cudaGetLastError());
        int flag = 0;
        CK(h, cudaMemcpyAsync(out, h->big_out, sizeof(out), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaMemcpyAsync(&n, h->big_nout, sizeof(n), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaMemcpyAsync(&flag, cg.flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
        need_all_pairs = flag != 0;
    }
    if (need_all_pairs) {
        if (h->coul) k_big_pairs<true><<<(unsigned)h->big_blocks, BIG_TILE, 0, h->stream>>>(g, img, h->big_tiles, h->big_e, h->big_n);
        else k_big_pairs<false><<<(unsigned)h->big_blocks, BIG_TILE, 0, h->stream>>>(g, img, h->big_tiles, h->big_e, h->big_n);
        k_big_reduce<<<1, BIG_TILE, 0, h->stream>>>(h->big_e, h->big_n, h->big_blocks, h->big_sol, h->big_tiles, h->big_out, h->big_nout);
        h->launches += 2;
        CK(h, cudaGetLastError());
        CK(h, cudaMemcpyAsync(out, h->big_out, sizeof(out), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaMemcpyAsync(&n, h->big_nout, sizeof(n), cudaMemcpyDeviceToHost, h->stream));
        CK(h, cudaStreamSynchronize(h->stream));
    }
    const double t = out[1] + out[0];
    CK(h, cudaMemcpyAsync(reinterpret_cast<char *>(h->ctl + c) + offsetof(ChainCtl, toteng), &t, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    if (tot) *tot = t;
    if (esol) *esol = out[1];
    if (pairs) *pairs = n;
    return 0;
}

static int ensure_prepared(mcb_t *h)
{
    if (!h->dirty) return 0;
    if (h->energy_only) {
        for (int c = 0; c < h->n_chains; ++c) if (int rc = big_energy(h, c, nullptr, nullptr, nullptr)) return rc;
    } else if (int rc = launch_prepare(h, 0, h->n_chains, nullptr, nullptr, nullptr)) return rc;
    h->dirty = false;
    return 0;
}

extern "C" int mcb_total_energy(mcb_t *h, int chain0, int nchains, double *toteng, double *e_solute_sum, int64_t *pairs_in_cutoff)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    CK(h, cudaSetDevice(h->device));
    if (h->energy_only) {
        for (int c = 0; c < nchains; ++c) {
            long long n = 0;
            if (int rc = big_energy(h, chain0 + c, toteng ? toteng + c : nullptr, e_solute_sum ? e_solute_sum + c : nullptr, &n)) return rc;
            if (pairs_in_cutoff) pairs_in_cutoff[c] = n;
        }
        if (chain0 == 0 && nchains == h->n_chains) h->dirty = false;
        return 0;
    }
    double *d_tot = h->d_out, *d_esol = h->d_out + h->n_chains;
    if (int rc = launch_prepare(h, chain0, nchains, d_tot, d_esol, h->d_ll)) return rc;
    std::vector<long long> pairs(nchains);
    if (toteng) CK(h, cudaMemcpyAsync(toteng, d_tot, sizeof(double) * nchains, cudaMemcpyDeviceToHost, h->stream));
    if (e_solute_sum) CK(h, cudaMemcpyAsync(e_solute_sum, d_esol, sizeof(double) * nchains, cudaMemcpyDeviceToHost, h->stream));
    if (pairs_in_cutoff) CK(h, cudaMemcpyAsync(pairs.data(), h->d_ll, sizeof(long long) * nchains, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    if (pairs_in_cutoff) for (int c = 0; c < nchains; ++c) pairs_in_cutoff[c] = pairs[c];
    if (chain0 == 0 && nchains == h->n_chains) h->dirty = false;
    return 0;
}

extern "C" int mcb_get_matrix(mcb_t *h, int chain, double *ss, double *energy)
{
    if (int rc = check_range(h, chain, 1)) return rc;
    if (h->energy_only) return arg_err(h, "mcb_get_matrix: this handle is in energy-only mode (no pair matrix is kept for a box this large)");
    if (int rc = ensure_prepared(h)) return rc;
    const Geo &g = h->g;
    const double *mat = g.mat_in_image ? h->images + (size_t)chain * g.image_doubles + g.o_mat : h->matg + (size_t)chain * g.N * g.MS;
    if (ss) CK(h, cudaMemcpy2DAsync(ss, (size_t)g.N * 8, mat, (size_t)g.MS * 8, (size_t)g.N * 8, g.N, cudaMemcpyDeviceToHost, h->stream));
    if (energy) CK(h, cudaMemcpyAsync(energy, h->images + (size_t)chain * g.image_doubles + g.o_energy, (size_t)g.N * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

static int launch_trial(mcb_t *h, int chain0, int nchains, int read_only)
{
    double *d_com = h->d_vec, *d_hoek = h->d_vec + 3 * (size_t)h->n_chains;
    double *d_dE = h->d_out, *d_en = h->d_out + h->n_chains;
    if (h->coul) k_trial<true><<<nchains, 32, 0, h->stream>>>(h->g, h->rel, h->images, h->matg, h->work, h->ctl, chain0, h->d_rsolv, d_com, d_hoek, d_dE, d_en, read_only);
    else k_trial<false><<<nchains, 32, 0, h->stream>>>(h->g, h->rel, h->images, h->matg, h->work, h->ctl, chain0, h->d_rsolv, d_com, d_hoek, d_dE, d_en, read_only);
    h->launches++;
    CK(h, cudaGetLastError());
    return 0;
}

extern "C" int mcb_trial_delta(mcb_t *h, int nchains, const int *rsolv, const double *new_com, const double *new_hoek,
                               double *d_e, double *e_new)
{
    if (int rc = check_range(h, 0, nchains)) return rc;
    if (!rsolv || !new_com || !new_hoek) return arg_err(h, "mcb_trial_delta: null array");
    if (h->energy_only) return arg_err(h, "mcb_trial_delta: this handle is in energy-only mode (box too large for the stepping kernels)");
    for (int c = 0; c < nchains; ++c)
        if (rsolv[c] < 1 || rsolv[c] > h->g.N) return arg_err(h, "mcb_trial_delta: rsolv outside 1..n_mol");
    CK(h, cudaSetDevice(h->device));
    if (int rc = ensure_prepared(h)) return rc;
    double *d_com = h->d_vec, *d_hoek = h->d_vec + 3 * (size_t)h->n_chains;
    CK(h, cudaMemcpyAsync(h->d_rsolv, rsolv, sizeof(int) * nchains, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaMemcpyAsync(d_com, new_com, sizeof(double) * 3 * nchains, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaMemcpyAsync(d_hoek, new_hoek, sizeof(double) * 3 * nchains, cudaMemcpyHostToDevice, h->stream));
    if (int rc = launch_trial(h, 0, nchains, 0)) return rc;
    h->trial_pending = true;
    if (d_e) CK(h, cudaMemcpyAsync(d_e, h->d_out, sizeof(double) * nchains, cudaMemcpyDeviceToHost, h->stream));
    if (e_new) CK(h, cudaMemcpyAsync(e_new, h->d_out + h->n_chains, sizeof(double) * nchains, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int mcb_commit(mcb_t *h, int nchains, const unsigned char *accept)
{
    if (int rc = check_range(h, 0, nchains)) return rc;
    if (!accept) return arg_err(h, "mcb_commit: null array");
    if (h->energy_only) return arg_err(h, "mcb_commit: this handle is in energy-only mode");
    // (new coordinates or a new solute since mcb_trial_delta drop the parked trial: committing it would apply a move
    //  that was evaluated against a state that no longer exists)
    if (!h->trial_pending) return arg_err(h, "mcb_commit: no trial is pending (call mcb_trial_delta first; mcb_set_chain_state / mcb_set_solute discard a pending trial)");
    h->vl.trust = 0;
    h->trial_pending = false;
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaMemcpyAsync(h->d_acc, accept, nchains, cudaMemcpyHostToDevice, h->stream));
    k_commit<<<nchains, 32, 0, h->stream>>>(h->g, h->images, h->matg, h->work, h->ctl, h->d_acc);
    h->launches++;
    CK(h, cudaGetLastError());
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int mcb_row_energy(mcb_t *h, int chain, int i, double *ss_row, double *e_i)
{
    if (int rc = check_range(h, chain, 1)) return rc;
    if (i < 1 || i > h->g.N) return arg_err(h, "mcb_row_energy: i outside 1..n_mol");
    if (h->energy_only) return arg_err(h, "mcb_row_energy: this handle is in energy-only mode");
    // the read-only evaluation uses the same per-chain work area as a parked trial and would drop it
    if (h->trial_pending) return arg_err(h, "mcb_row_energy: a trial from mcb_trial_delta is pending; call mcb_commit first");
    CK(h, cudaSetDevice(h->device));
    if (int rc = ensure_prepared(h)) return rc;
    const Geo &g = h->g;
    // a read-only "trial" that moves molecule i onto itself: new coordinates = current coordinates, copied on the
    // device from the chain's image (six values, one per SoA plane)
    double *d_com = h->d_vec, *d_hoek = h->d_vec + 3 * (size_t)h->n_chains;
    const double *img = h->images + (size_t)chain * g.image_doubles;
    CK(h, cudaMemcpyAsync(h->d_rsolv, &i, sizeof(int), cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaMemcpy2DAsync(d_com, 8, img + g.o_com + (i - 1), (size_t)g.Np * 8, 8, 3, cudaMemcpyDeviceToDevice, h->stream));
    CK(h, cudaMemcpy2DAsync(d_hoek, 8, img + g.o_hoek + (i - 1), (size_t)g.Np * 8, 8, 3, cudaMemcpyDeviceToDevice, h->stream));
    if (int rc = launch_trial(h, chain, 1, 1)) return rc;
    const double *wk = h->work + (size_t)chain * g.work_doubles;
    if (ss_row) CK(h, cudaMemcpyAsync(ss_row, wk + g.o_newrow, sizeof(double) * g.N, cudaMemcpyDeviceToHost, h->stream));
    if (e_i) CK(h, cudaMemcpyAsync(e_i, reinterpret_cast<const char *>(wk + g.o_pend) + offsetof(Pending, e_sol), sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

// ------------------------------------------------------------------------------------------------ batched mode
static int edit_ctl(mcb_t *h, int chain0, int nchains, std::vector<ChainCtl> &ctl)
{
    ctl.resize(nchains);
    CK(h, cudaSetDevice(h->device));
    CK(h, cudaMemcpyAsync(ctl.data(), h->ctl + chain0, sizeof(ChainCtl) * nchains, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

static int store_ctl(mcb_t *h, int chain0, int nchains, const std::vector<ChainCtl> &ctl)
{
    CK(h, cudaMemcpyAsync(h->ctl + chain0, ctl.data(), sizeof(ChainCtl) * nchains, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int mcb_set_mc_params(mcb_t *h, int chain0, int nchains, const mcb_mc_params *p, int broadcast)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    if (!p) return arg_err(h, "mcb_set_mc_params: null");
    for (int c = 0; c < (broadcast ? 1 : nchains); ++c)
        if (p[c].nadj < 1 || !(p[c].temperature > 0.0)) return arg_err(h, "mcb_set_mc_params: nadj >= 1 and temperature > 0 required");
    std::vector<ChainCtl> ctl;
    if (int rc = edit_ctl(h, chain0, nchains, ctl)) return rc;
    for (int c = 0; c < nchains; ++c) {
        const mcb_mc_params &q = p[broadcast ? 0 : c];
        ChainCtl &s = ctl[c];
        s.temperature = q.temperature; s.dposmax = q.dposmax; s.dposmin = q.dposmin; s.dhoekmax = q.dhoekmax;
        s.dhoekmin = q.dhoekmin; s.padj = q.padj; s.nadj = q.nadj; s.start_prod = q.start_prod;
    }
    return store_ctl(h, chain0, nchains, ctl);
}

extern "C" int mcb_seed_chains(mcb_t *h, int chain0, int nchains, const uint64_t *seeds)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    if (!seeds) return arg_err(h, "mcb_seed_chains: null");
    std::vector<ChainCtl> ctl;
    if (int rc = edit_ctl(h, chain0, nchains, ctl)) return rc;
    for (int c = 0; c < nchains; ++c) {
        unsigned long long x = seeds[c] + 0x9E3779B97F4A7C15ULL;
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL; x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL; x ^= x >> 31;
        ctl[c].rng_key = x; ctl[c].rng_n = 0; ctl[c].xn = 0;
    }
    return store_ctl(h, chain0, nchains, ctl);
}

extern "C" int mcb_reset_stats(mcb_t *h, int chain0, int nchains)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    std::vector<ChainCtl> ctl;
    if (int rc = edit_ctl(h, chain0, nchains, ctl)) return rc;
    for (auto &s : ctl) {
        s.production = s.sum_e = s.sum_e2 = s.ratio = s.last_delta = 0.0;
        s.step = s.nsuc = s.naccept = s.prod_acc = s.nclamp = s.sitepairs = s.uused = s.nstat = 0;
        s.xatt = s.xacc = 0;
    }
    return store_ctl(h, chain0, nchains, ctl);
}

extern "C" int mcb_get_stats(mcb_t *h, int chain0, int nchains, mcb_chain_stats *out)
{
    if (int rc = check_range(h, chain0, nchains)) return rc;
    if (!out) return arg_err(h, "mcb_get_stats: null");
    if (int rc = ensure_prepared(h)) return rc;
    std::vector<ChainCtl> ctl;
    if (int rc = edit_ctl(h, chain0, nchains, ctl)) return rc;
    for (int c = 0; c < nchains; ++c) {
        const ChainCtl &s = ctl[c];
        mcb_chain_stats &o = out[c];
        o.toteng = s.toteng; o.dposmax = s.dposmax; o.dhoekmax = s.dhoekmax; o.production = s.production;
        o.sum_e = s.sum_e; o.sum_e2 = s.sum_e2; o.ratio = s.ratio; o.last_delta = s.last_delta;
        o.step = s.step; o.nsuc = s.nsuc; o.naccept = s.naccept; o.prod_acc = s.prod_acc; o.n_clamped = s.nclamp;
        o.sitepairs = s.sitepairs; o.uniforms_used = s.uused; o.n_stat = s.nstat;
        o.exch_attempts = s.xatt; o.exch_accepted = s.xacc; o.temperature = s.temperature;
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------ replica exchange
// One thread per pair; the rule is METROPOLIS' (MonteCarlo.F90:728-744) applied to the parallel-tempering exponent.
__global__ void k_exchange(ChainCtl *ctl, int npairs, const int *pa, const int *pb, const double *u, unsigned char *acc)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    ChainCtl &A = ctl[pa[p]], &B = ctl[pb[p]];
    const double ba = 1000.0 / (8.314 * A.temperature), bb = 1000.0 / (8.314 * B.temperature);
    const double exponent = (ba - bb) * (A.toteng - B.toteng);
    bool accept;
    if (exponent < -75.0) accept = false;                       // KANS = 0, RV = 1: no draw (:730-733)
    else if (exponent >= 0.0) accept = true;                    // KANS = 1, RV = 0: no draw (:734-736)
    else {
        double rv;
        if (u) rv = u[p];
        else { rv = stream_uniform(mix64(A.rng_key + 0xD1B54A32D192ED03ULL), A.xn); A.xn += 1; }
        accept = rv <= fmin(exp(exponent * LN_E_F32), 1.0);    // KANS = E**EXPONENT, RV <= KANS (:738-744)
    }
    A.xatt += 1; B.xatt += 1;
    if (accept) {
        double t;
        t = A.temperature; A.temperature = B.temperature; B.temperature = t;
        t = A.dposmax; A.dposmax = B.dposmax; B.dposmax = t;
        t = A.dhoekmax; A.dhoekmax = B.dhoekmax; B.dhoekmax = t;
        A.xacc += 1; B.xacc += 1;
    }
    if (acc) acc[p] = accept ? 1 : 0;
}

extern "C" int mcb_exchange(mcb_t *h, int npairs, const int *pair_a, const int *pair_b, const double *uniforms, unsigned char *accepted)
{
    if (!h) return 1;
    if (npairs < 0 || (npairs > 0 && (!pair_a || !pair_b))) return arg_err(h, "mcb_exchange: null pair arrays");
    if (npairs == 0) return 0;
    std::vector<char> seen(h->n_chains, 0);
    for (int p = 0; p < npairs; ++p) {
        const int a = pair_a[p], b = pair_b[p];
        if (a < 0 || b < 0 || a >= h->n_chains || b >= h->n_chains || a == b || seen[a] || seen[b])
            return arg_err(h, "mcb_exchange: pairs must be distinct chains of this handle, each chain in at most one pair");
        seen[a] = seen[b] = 1;
    }
    CK(h, cudaSetDevice(h->device));
    if (int rc = ensure_prepared(h)) return rc;     // TOTENG must be current
    // inputs ride in the staging buffers of the single-step entry points: d_rsolv [n_chains] ints is too small for two
    // index arrays, so the pairs go through d_ll (2 x n_chains 64-bit words) and the uniforms through d_out
    int *d_a = reinterpret_cast<int *>(h->d_ll), *d_b = d_a + h->n_chains;
    CK(h, cudaMemcpyAsync(d_a, pair_a, sizeof(int) * npairs, cudaMemcpyHostToDevice, h->stream));
    CK(h, cudaMemcpyAsync(d_b, pair_b, sizeof(int) * npairs, cudaMemcpyHostToDevice, h->stream));
    if (uniforms) CK(h, cudaMemcpyAsync(h->d_out, uniforms, sizeof(double) * npairs, cudaMemcpyHostToDevice, h->stream));
    k_exchange<<<(npairs + 127) / 128, 128, 0, h->stream>>>(h->ctl, npairs, d_a, d_b, uniforms ? h->d_out : nullptr, accepted ? h->d_acc : nullptr);
    h->launches++;
    CK(h, cudaGetLastError());
    if (accepted) CK(h, cudaMemcpyAsync(accepted, h->d_acc, npairs, cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    return 0;
}

template <bool COUL>
static void launch_run_ubuf(mcb_t *h, long long nsteps, long long cap, long long *d_consumed, long long *d_steps)
{
    if (h->g.mat_in_image)
        k_run_steps<COUL, true><<<h->n_chains, 32, h->smem_bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, nsteps, h->d_ubuf, cap, d_consumed, d_steps);
    else
        k_run_steps<COUL, false><<<h->n_chains, 32, h->smem_bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, nsteps, h->d_ubuf, cap, d_consumed, d_steps);
}

template <bool COUL, int GL, int W, bool SMALL>
static void launch_run_spec2(mcb_t *h, long long nsteps, long long *d_steps)
{
    if (h->g.mat_in_image)
        k_run_spec<COUL, true, GL, W, SMALL><<<h->n_chains, 32 * W, h->smem_spec_bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, nsteps, d_steps);
    else
        k_run_spec<COUL, false, GL, W, SMALL><<<h->n_chains, 32 * W, h->smem_spec_bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, nsteps, d_steps);
}

template <bool COUL, int GL, int W>
static void launch_run_spec(mcb_t *h, long long nsteps, long long *d_steps)
{
    if (h->g.N <= 32) launch_run_spec2<COUL, GL, W, true>(h, nsteps, d_steps); else launch_run_spec2<COUL, GL, W, false>(h, nsteps, d_steps);
}

template <bool COUL>
static void launch_run_spec_any(mcb_t *h, long long nsteps, long long *d_steps)
{
    const int key = h->spec_gl * 10 + h->spec_w;
    switch (key) {
#ifdef MCB_DEV_BUILD
        default: launch_run_spec<COUL, 32, 1>(h, nsteps, d_steps); break;
#else
        case 321: launch_run_spec<COUL, 32, 1>(h, nsteps, d_steps); break;
        case 322: launch_run_spec<COUL, 32, 2>(h, nsteps, d_steps); break;
        case 324: launch_run_spec<COUL, 32, 4>(h, nsteps, d_steps); break;
        case 161: launch_run_spec<COUL, 16, 1>(h, nsteps, d_steps); break;
        case 162: launch_run_spec<COUL, 16, 2>(h, nsteps, d_steps); break;
        case 81: launch_run_spec<COUL, 8, 1>(h, nsteps, d_steps); break;
        default: launch_run_spec<COUL, 8, 2>(h, nsteps, d_steps); break;
#endif
    }
}

template <int W, bool IMG_SMEM>
static void launch_vl2(mcb_t *h, long long nsteps, long long *d_steps)
{
    const size_t bytes = (size_t)vl2_smem_doubles(h->g, W, IMG_SMEM) * 8;
    if (h->coul) k_run_vl2<true, W, IMG_SMEM><<<h->n_chains, 32 * W, bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->work, h->ctl, nsteps, d_steps, h->vl);
    else k_run_vl2<false, W, IMG_SMEM><<<h->n_chains, 32 * W, bytes, h->stream>>>(h->g, h->rel, h->images, h->matg, h->work, h->ctl, nsteps, d_steps, h->vl);
}

template <int W, bool IMG_SMEM>
static int vl2_blocks_per_sm(mcb_t *h)
{
    const size_t bytes = (size_t)vl2_smem_doubles(h->g, W, IMG_SMEM) * 8;
    if (bytes > 227 * 1024) return 0;
    if (set_smem_attr(h, k_run_vl2<true, W, IMG_SMEM>, bytes) || set_smem_attr(h, k_run_vl2<false, W, IMG_SMEM>, bytes)) return 0;
    int nb = 0;
    cudaError_t e = h->coul ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_run_vl2<true, W, IMG_SMEM>, 32 * W, bytes)
                            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_run_vl2<false, W, IMG_SMEM>, 32 * W, bytes);
    return e == cudaSuccess ? nb : 0;
}

// Which k_run_vl2 runs a handle: the chain's image staged in shared memory (W = 2-4 warps per chain) or left in global
// memory (W = 1-4), and how many warps per chain.  Staged: the widest variant that still keeps every chain resident at
// once; if none does, the one that puts most warps on an SM.  vl2_w = -1: nothing fits (k_run_vl is used).
// MCB200_VL2_W and MCB200_VL2_IMG=smem|global override (tuning).
static void pick_vl2_variant(mcb_t *h)
{
    int dev = 0, nsm = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    const int occ_s[5] = {0, 0, vl2_blocks_per_sm<2, true>(h), vl2_blocks_per_sm<3, true>(h), vl2_blocks_per_sm<4, true>(h)};
    const int occ_g[5] = {0, vl2_blocks_per_sm<1, false>(h), vl2_blocks_per_sm<2, false>(h), vl2_blocks_per_sm<3, false>(h), vl2_blocks_per_sm<4, false>(h)};
    const char *ti = getenv("MCB200_VL2_IMG"), *tw = getenv("MCB200_VL2_W");
    const int w_env = tw ? atoi(tw) : 0;
    if (ti && !strcmp(ti, "global")) {
        h->vl2_global = true;
        h->vl2_w = w_env >= 1 && w_env <= 4 && occ_g[w_env] > 0 ? w_env : 1;
        return;
    }
    h->vl2_global = false;
    if (w_env >= 2 && w_env <= 4 && occ_s[w_env] > 0) { h->vl2_w = w_env; return; }
    int best = -1, best_warps = 0;
    for (int w = 4; w >= 2 && best < 0; --w) if (occ_s[w] > 0 && (long long)occ_s[w] * nsm >= h->n_chains) best = w;
    if (best < 0) for (int w = 2; w <= 4; ++w) if (occ_s[w] * w > best_warps) { best_warps = occ_s[w] * w; best = w; }
    // Global image instead: with cps chains per SM it keeps min(cps, 16 / W) * W warps on an SM.  Measured on B200
    // (profiles/vl2_image_r02.txt): a warp reading its image through L1/L2 does about 0.6 of the work of one with the
    // image staged, so the global variant is taken when it at least doubles the resident warps (N = 810: 1 staged chain
    // per SM), never at N = 240 (4 staged chains per SM), and always when the image does not fit in shared memory.
    const int cps = (int)((h->n_chains + nsm - 1) / nsm);
    int w_g = std::min(4, std::max(1, 16 / std::max(cps, 1)));
    while (w_g > 1 && occ_g[w_g] == 0) --w_g;
    const int warps_g = std::min(cps, occ_g[w_g]) * w_g;
    const int warps_s = best > 0 ? std::min(cps, occ_s[best]) * best : 0;
    if (ti && !strcmp(ti, "smem") && best > 0) { h->vl2_w = best; return; }
    if (occ_g[w_g] > 0 && (best < 0 || !h->smem_fits || warps_g >= 2 * warps_s)) { h->vl2_global = true; h->vl2_w = w_g; return; }
    h->vl2_w = best;
}

extern "C" int mcb_run_steps(mcb_t *h, int64_t nsteps, const double *uniforms, int64_t cap, int64_t *consumed, int64_t *steps_done)
{
    if (!h) return 1;
    if (nsteps < 0) return arg_err(h, "mcb_run_steps: nsteps < 0");
    if (!h->can_step) return arg_err(h, "mcb_run_steps: a chain of this size does not fit in shared memory (energy-only mode)");
    if (uniforms && cap < 8) return arg_err(h, "mcb_run_steps: uniform buffer needs cap >= 8 per chain");
    if (uniforms && !h->smem_fits) return arg_err(h, "mcb_run_steps: a chain of this size does not fit in shared memory and steps with the built-in stream only (uniforms = NULL)");
    CK(h, cudaSetDevice(h->device));
    if (int rc = ensure_prepared(h)) return rc;
    const size_t nc = (size_t)h->n_chains;
    if (uniforms) {
        size_t need = nc * (size_t)cap;
        if (need > h->ubuf_cap) {
            cudaFree(h->d_ubuf); h->d_ubuf = nullptr; h->ubuf_cap = 0;
            CK(h, cudaMalloc(&h->d_ubuf, need * sizeof(double)));
            h->ubuf_cap = need;
        }
        CK(h, cudaMemcpyAsync(h->d_ubuf, uniforms, need * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    }
    long long *d_consumed = h->d_ll, *d_steps = h->d_ll + nc;
    if (nsteps > (1LL << 30)) return arg_err(h, "mcb_run_steps: at most 2^30 steps per call");
    if (uniforms) h->vl.trust = 0;   // k_run_steps moves molecules without looking at the lists
    if (uniforms) { if (h->coul) launch_run_ubuf<true>(h, nsteps, cap, d_consumed, d_steps); else launch_run_ubuf<false>(h, nsteps, cap, d_consumed, d_steps); }
    else {
        CK(h, cudaMemsetAsync(d_consumed, 0, sizeof(long long) * nc, h->stream));
        if (h->vl.idx && !h->vl2_w && !h->no_vl2) pick_vl2_variant(h);
        if (h->vl.idx && h->vl2_w > 0) {
            // W warps per chain, W consecutive steps in flight (mcb_verlet2.cuh)
            if (h->vl2_global) switch (h->vl2_w) {
                case 4: launch_vl2<4, false>(h, nsteps, d_steps); break;
                case 3: launch_vl2<3, false>(h, nsteps, d_steps); break;
                case 2: launch_vl2<2, false>(h, nsteps, d_steps); break;
                default: launch_vl2<1, false>(h, nsteps, d_steps); break;
            }
            else switch (h->vl2_w) {
                case 4: launch_vl2<4, true>(h, nsteps, d_steps); break;
                case 3: launch_vl2<3, true>(h, nsteps, d_steps); break;
                default: launch_vl2<2, true>(h, nsteps, d_steps); break;
            }
            h->vl.trust = 1;
        } else if (h->vl.idx) {
            if (!h->smem_fits) return arg_err(h, "mcb_run_steps: no stepping kernel fits a chain of this size");
            const size_t vb = (size_t)vl_smem_doubles(h->g) * 8;   // (its own size: k_run_spec's plan has grown caches this kernel does not use)
            set_smem_attr(h, k_run_vl<true>, vb); set_smem_attr(h, k_run_vl<false>, vb);
            if (h->coul) k_run_vl<true><<<h->n_chains, 32, vb, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, nsteps, d_steps, h->vl);
            else k_run_vl<false><<<h->n_chains, 32, vb, h->stream>>>(h->g, h->rel, h->images, h->matg, h->ctl, nsteps, d_steps, h->vl);
            h->vl.trust = 1;   // until an entry point other than this one changes the coordinates
        } else if (h->smem_spec_bytes > 227 * 1024) return arg_err(h, "mcb_run_steps: without neighbour lists a chain of this size does not fit in shared memory");
        else if (h->coul) launch_run_spec_any<true>(h, nsteps, d_steps); else launch_run_spec_any<false>(h, nsteps, d_steps);
    }
    h->launches++;
    CK(h, cudaGetLastError());
    // nothing to hand back: the launch stays asynchronous on the handle's stream (the next call that reads
    // results, e.g. mcb_get_stats, synchronises)
    if (!uniforms && !consumed && !steps_done) return 0;
    std::vector<long long> tmp;
    if (consumed || steps_done) {
        tmp.resize(2 * nc);
        CK(h, cudaMemcpyAsync(tmp.data(), h->d_ll, sizeof(long long) * 2 * nc, cudaMemcpyDeviceToHost, h->stream));
    }
    CK(h, cudaStreamSynchronize(h->stream));
    for (size_t c = 0; c < nc && !tmp.empty(); ++c) {
        if (consumed) consumed[c] = tmp[c];
        if (steps_done) steps_done[c] = tmp[nc + c];
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------ measurement
extern "C" int mcb_measure_fp64_peak(int device, double ms, double *tflops)
{
    if (!tflops) return 1;
    if (device >= 0 && cudaSetDevice(device) != cudaSuccess) return -1;
    cudaDeviceProp prop;
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return -1;
    const int blocks = prop.multiProcessorCount * 8, threads = 256;
    double *out = nullptr;
    if (cudaMalloc(&out, sizeof(double) * blocks * threads) != cudaSuccess) return -2;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 2000;
    double best = 0.0;
    float t = 0.f;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        k_fp64_peak<<<blocks, threads>>>(out, iters);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return -3; }
        cudaEventElapsedTime(&t, e0, e1);
        double flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
        double tf = flops / (t * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
        if (t < ms) iters = (int)std::min(2.0e6, iters * std::max(1.5, ms / std::max(t, 0.01f)));
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    *tflops = best;
    return 0;
}

#ifdef MCB_PHASE_CLOCKS
extern "C" int mcb_debug_phase_clocks(unsigned long long *out, int reset)
{
    if (out && cudaMemcpyFromSymbol(out, g_phase, sizeof(unsigned long long) * 8) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[8] = {0}; if (cudaMemcpyToSymbol(g_phase, z, sizeof(z)) != cudaSuccess) return -1; }
    return 0;
}
#endif

// ================================================================================================
// mcb_box_energy_noalloc — full-box energy for callers that must not allocate device memory.
//
// Some hosts cannot let a library call cudaMalloc: a driver that has handed every byte of the GPU to its own
// pool at start-up, a job running under a fixed per-process memory cap, or a timing-critical loop that cannot
// afford the device-wide synchronisation an allocation implies.  For them this one entry point computes the
// same number as mcb_total_energy's matrix-free path without creating a handle and without allocating
// anything, on the device or through any library: the caller passes device pointers for the coordinates and
// for the result, the small tables travel as kernel arguments, and the per-block partial sums — the only
// scratch the computation needs — live in the fixed array below.  Because that array is part of the loaded
// library, every GPU that loads libmcb200.so carries it (about 0.5 MB) until the process exits, and its size
// is what limits the box: MCB_NOALLOC_MAX_MOL molecules, MCB_NOALLOC_MAX_SOL solute atoms.
//
// The array is shared by every call and keeps whatever the last call left in it.  Nothing in it has to start
// at zero: each call writes every slot it later reads (one slot per tile pair / tile) before its own
// reduction kernel reads them, in stream order.  Two calls whose kernels overlap on the GPU (different
// streams, or several host threads) would write the same slots and race; callers must serialise them.
// ================================================================================================
namespace {

constexpr int NA_MAX_TILES = (MCB_NOALLOC_MAX_MOL + BIG_TILE - 1) / BIG_TILE;
constexpr int NA_MAX_BLOCKS = NA_MAX_TILES * (NA_MAX_TILES + 1) / 2;

struct NaScratch {
    double pair_e[NA_MAX_BLOCKS];
    long long pair_n[NA_MAX_BLOCKS];
    double sol_e[NA_MAX_TILES];
};

// In host code this name is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) NaScratch g_na_scratch;

struct NaArgs {                          // everything small travels by value: no device buffer needed
    int N, nsol, n_sites, pad;
    double boxl2, inv_boxl, cut2;
    double rel[NSITE * 3];
    double ab0[NSITE * NSITE * 2], q0[NSITE * NSITE];                      // solvent-solvent (A,B), q q conv
    double sol[MCB_NOALLOC_MAX_SOL * 3];
    double sol_ab[NSITE * MCB_NOALLOC_MAX_SOL * 2], sol_q[NSITE * MCB_NOALLOC_MAX_SOL];
};

__device__ __forceinline__ Geo na_geo(const NaArgs &a)
{
    Geo g{};
    g.boxl2 = a.boxl2; g.inv_boxl = a.inv_boxl; g.cut2 = a.cut2;
    return g;
}

// RotMatrix of one molecule into registers (MC_lib/lib.f90:27-83); com/hoek are the driver's AoS arrays
__device__ __forceinline__ void na_sites(const NaArgs &a, const double *com, const double *hoek, int j, double out[SITE_DOUBLES])
{
    double s1, c1, s2, c2, s3, c3;
    sincos(hoek[3 * (size_t)j], &s1, &c1); sincos(hoek[3 * (size_t)j + 1], &s2, &c2); sincos(hoek[3 * (size_t)j + 2], &s3, &c3);
    for (int k = 0; k < NSITE; ++k)
        rotate_site(c1, s1, c2, s2, c3, s3, a.rel[3 * k], a.rel[3 * k + 1], a.rel[3 * k + 2], com[3 * (size_t)j], com[3 * (size_t)j + 1],
                    com[3 * (size_t)j + 2], out[3 * k], out[3 * k + 1], out[3 * k + 2]);
}

template <bool COUL>
__global__ void __launch_bounds__(128) k_na_pairs(NaArgs a, const double *com, const double *hoek, NaScratch *sc)
{
    __shared__ double scom[3][BIG_TILE];
    __shared__ double red[BIG_TILE];
    const Geo g = na_geo(a);
    const long long b = blockIdx.x;
    int tj = (int)((sqrt(8.0 * (double)b + 1.0) - 1.0) * 0.5);
    while ((long long)tj * (tj + 1) / 2 > b) --tj;
    while ((long long)(tj + 1) * (tj + 2) / 2 <= b) ++tj;
    const int ti = (int)(b - (long long)tj * (tj + 1) / 2);
    const int i0 = ti * BIG_TILE, j = tj * BIG_TILE + threadIdx.x;
    {
        const int i = i0 + threadIdx.x;
        const bool ok = i < a.N;
        scom[0][threadIdx.x] = ok ? com[3 * (size_t)i] : 0.0; scom[1][threadIdx.x] = ok ? com[3 * (size_t)i + 1] : 0.0;
        scom[2][threadIdx.x] = ok ? com[3 * (size_t)i + 2] : 0.0;
    }
    __syncthreads();
    double e = 0.0, np = 0.0;
    if (j < a.N) {
        double mine[SITE_DOUBLES], other[SITE_DOUBLES];
        bool have_mine = false;
        const double cx = com[3 * (size_t)j], cy = com[3 * (size_t)j + 1], cz = com[3 * (size_t)j + 2];
        const int ni = min(BIG_TILE, min(a.N, j) - i0);
        for (int k = 0; k < ni; ++k) {
            if (pair_r2(scom[0][k] - cx, scom[1][k] - cy, scom[2][k] - cz, g) > g.cut2) continue;   // MonteCarlo.F90:806-813
            np += 1.0;
            if (!have_mine) { na_sites(a, com, hoek, j, mine); have_mine = true; }
            na_sites(a, com, hoek, i0 + k, other);
            double en = 0.0;
            for (int p = 0; p < NSITE; ++p)          // larger index j is MOL1 (outer), i is MOL2 (inner), LennardJones.f90:65-79
                for (int q = 0; q < NSITE; ++q)
                    en = site_pair_acc<COUL>(en, other[3 * q] - mine[3 * p], other[3 * q + 1] - mine[3 * p + 1], other[3 * q + 2] - mine[3 * p + 2],
                                             a.ab0[(p * 4 + q) * 2], a.ab0[(p * 4 + q) * 2 + 1], COUL ? a.q0[p * 4 + q] : 0.0, g);
            e += en;
        }
    }
    const double be = block_sum_128(e, red), bn = block_sum_128(np, red);
    if (threadIdx.x == 0) { sc->pair_e[blockIdx.x] = be; sc->pair_n[blockIdx.x] = (long long)bn; }
}

template <bool COUL>
__global__ void __launch_bounds__(128) k_na_solute(NaArgs a, const double *com, const double *hoek, NaScratch *sc)
{
    __shared__ double red[BIG_TILE];
    const Geo g = na_geo(a);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double e = 0.0;
    if (i < a.N && a.nsol > 0) {
        double mine[SITE_DOUBLES];
        na_sites(a, com, hoek, i, mine);
        for (int p = 0; p < a.n_sites; ++p)          // CalcLJ_SV, LennardJones.f90:131-143
            for (int q = 0; q < a.nsol; ++q)
                e = site_pair_acc<COUL>(e, a.sol[3 * q] - mine[3 * p], a.sol[3 * q + 1] - mine[3 * p + 1], a.sol[3 * q + 2] - mine[3 * p + 2],
                                        a.sol_ab[(p * MCB_NOALLOC_MAX_SOL + q) * 2], a.sol_ab[(p * MCB_NOALLOC_MAX_SOL + q) * 2 + 1],
                                        COUL ? a.sol_q[p * MCB_NOALLOC_MAX_SOL + q] : 0.0, g);
    }
    const double be = block_sum_128(e, red);
    if (threadIdx.x == 0) sc->sol_e[blockIdx.x] = be;
}

__global__ void __launch_bounds__(128) k_na_reduce(const NaScratch *sc, int npair_blocks, int ntiles, double *out)
{
    __shared__ double red[BIG_TILE];
    double e = 0.0, n = 0.0, es = 0.0;
    for (int k = threadIdx.x; k < npair_blocks; k += BIG_TILE) { e += sc->pair_e[k]; n += (double)sc->pair_n[k]; }
    for (int k = threadIdx.x; k < ntiles; k += BIG_TILE) es += sc->sol_e[k];
    const double te = block_sum_128(e, red), tn = block_sum_128(n, red), ts = block_sum_128(es, red);
    if (threadIdx.x == 0) { out[0] = ts + te; out[1] = ts; out[2] = tn; }
}

constexpr int NA_MAX_DEVICES = 64;
NaScratch *g_na_ptr[NA_MAX_DEVICES];      // device address of g_na_scratch, looked up once per device
std::mutex g_na_mutex;

void na_msg(char *buf, size_t len, const char *msg)
{
    if (buf && len) { snprintf(buf, len, "%s", msg); }
}

}  // namespace

extern "C" int mcb_box_energy_noalloc(const mcb_params *p, const double *d_com, const double *d_hoek, const double *solute,
                                      const double *sol_tab, double *d_out, void *cuda_stream, char *errbuf, size_t errbuf_len)
{
    if (!p || !d_com || !d_hoek || !d_out || !p->site_rel || !p->site_tab) { na_msg(errbuf, errbuf_len, "mcb_box_energy_noalloc: null argument"); return 1; }
    if (p->n_mol < 1 || p->n_mol > MCB_NOALLOC_MAX_MOL) {
        na_msg(errbuf, errbuf_len, "mcb_box_energy_noalloc: n_mol outside 1..MCB_NOALLOC_MAX_MOL (32768): the fixed scratch array holds "
                                   "one partial sum per 128x128 tile pair and cannot grow; use mcb_create + mcb_total_energy for larger boxes");
        return 2;
    }
    if (p->n_sol < 0 || p->n_sol > MCB_NOALLOC_MAX_SOL || (p->n_sol > 0 && (!solute || !sol_tab))) {
        na_msg(errbuf, errbuf_len, "mcb_box_energy_noalloc: n_sol outside 0..MCB_NOALLOC_MAX_SOL (24), or solute tables missing: "
                                   "the solute travels as a kernel argument of fixed size");
        return 2;
    }
    if (p->n_sites < 1 || p->n_sites > MCB_MAX_SITES || !(p->boxl > 0.0) || !(p->cutoff > 0.0)) {
        na_msg(errbuf, errbuf_len, "mcb_box_energy_noalloc: bad parameter (1 <= n_sites <= 4; boxl, cutoff > 0)");
        return 1;
    }
    // p->device >= 0 selects that GPU for this call only: the caller's current device is put back before returning
    int dev = 0, caller_dev = -1;
    cudaError_t e = cudaGetDevice(&caller_dev);
    if (e == cudaSuccess && p->device >= 0 && p->device != caller_dev) e = cudaSetDevice(p->device);
    if (e == cudaSuccess) e = cudaGetDevice(&dev);
    struct Restore { int d; bool on; ~Restore() { if (on) cudaSetDevice(d); } } restore{caller_dev, e == cudaSuccess && dev != caller_dev};
    if (e != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) { na_msg(errbuf, errbuf_len, e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal too large"); return -(int)(e != cudaSuccess ? e : cudaErrorInvalidDevice); }
    // an error some earlier call of the host left behind is the host's: report it as such instead of as a launch failure
    e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        if (errbuf && errbuf_len) snprintf(errbuf, errbuf_len, "mcb_box_energy_noalloc: the CUDA context already holds an error from an earlier call (%s); nothing was launched", cudaGetErrorString(e));
        return -(int)e;
    }
    NaScratch *sc = nullptr;
    {   // set-up, once per device, with that device current: resolve the device address of the scratch symbol
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_ptr[dev]) {
            void *ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, g_na_scratch);
            if (e != cudaSuccess || !ptr) { na_msg(errbuf, errbuf_len, cudaGetErrorString(e)); return -(int)(e != cudaSuccess ? e : cudaErrorInvalidSymbol); }
            g_na_ptr[dev] = static_cast<NaScratch *>(ptr);
        }
        sc = g_na_ptr[dev];
    }
    NaArgs a;
    memset(&a, 0, sizeof(a));
    a.N = p->n_mol; a.nsol = p->n_sol; a.n_sites = p->n_sites;
    a.boxl2 = 2.0 * p->boxl; a.inv_boxl = 1.0 / p->boxl; a.cut2 = exact_cut2(p->cutoff);
    double tab[NSITE * 3] = {0};
    for (int s = 0; s < p->n_sites; ++s)
        for (int k = 0; k < 3; ++k) { a.rel[3 * s + k] = p->site_rel[3 * s + k]; tab[3 * s + k] = p->site_tab[3 * s + k]; }
    for (int s = p->n_sites; s < NSITE; ++s) { a.rel[3 * s] = 1.0e4 + 7.0 * s; a.rel[3 * s + 1] = 2.0e4; a.rel[3 * s + 2] = 3.0e4; }   // parameter-free padding
    for (int s = 0; s < p->n_sites; ++s) {
        for (int t = 0; t < p->n_sites; ++t)
            pair_consts(&tab[3 * s], &tab[3 * t], p->conv, a.ab0[(s * 4 + t) * 2], a.ab0[(s * 4 + t) * 2 + 1], a.q0[s * 4 + t]);
        for (int q = 0; q < p->n_sol; ++q)
            pair_consts(&tab[3 * s], sol_tab + 3 * q, p->conv, a.sol_ab[(s * MCB_NOALLOC_MAX_SOL + q) * 2],
                        a.sol_ab[(s * MCB_NOALLOC_MAX_SOL + q) * 2 + 1], a.sol_q[s * MCB_NOALLOC_MAX_SOL + q]);
    }
    for (int q = 0; q < 3 * p->n_sol; ++q) a.sol[q] = solute[q];
    const int tiles = (a.N + BIG_TILE - 1) / BIG_TILE, blocks = tiles * (tiles + 1) / 2;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (p->conv != 0.0) { k_na_pairs<true><<<blocks, BIG_TILE, 0, st>>>(a, d_com, d_hoek, sc); k_na_solute<true><<<tiles, BIG_TILE, 0, st>>>(a, d_com, d_hoek, sc); }
    else { k_na_pairs<false><<<blocks, BIG_TILE, 0, st>>>(a, d_com, d_hoek, sc); k_na_solute<false><<<tiles, BIG_TILE, 0, st>>>(a, d_com, d_hoek, sc); }
    k_na_reduce<<<1, BIG_TILE, 0, st>>>(sc, blocks, tiles, d_out);
    e = cudaGetLastError();
    if (e != cudaSuccess) { na_msg(errbuf, errbuf_len, cudaGetErrorString(e)); return -(int)e; }
    return 0;
}

// ================================================================================================
// End-of-run collective: mcb_comm_* / mcb_reduce_stats / mcb_gather_chains.
//
// The reference runs its independent chains as separate OS processes and combines them afterwards by scraping
// their logs (PostProcess/PostProcess.F90:202-256, RECON and READLOG).  Here the chains of one job are spread over
// the GPUs of a box in contiguous blocks (one handle per GPU, nothing exchanged while sampling) and combined
// by ONE NCCL all-reduce of a small statistics vector and ONE all-gather of four numbers per chain, both over
// NVLink.  NCCL is bound at run time (dlopen of libnccl.so.2; the copy a host such as PyTorch has already loaded is
// reused), so single-GPU users of the library do not need it installed.
// ================================================================================================
#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    std::string err;
};

NcclApi g_nccl;
std::mutex g_nccl_mutex;

// nullptr on success, else why NCCL could not be bound
const char *nccl_bind()
{
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    NcclApi &n = g_nccl;
    if (n.lib) return nullptr;
    if (!n.err.empty()) return n.err.c_str();
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // the copy the host process already uses, if any
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { n.err = std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror(); return n.err.c_str(); }
    bool ok = true;
    auto sym = [&](const char *name) { void *p = dlsym(lib, name); if (!p) { ok = false; n.err = std::string("NCCL symbol missing: ") + name; } return p; };
    n.GetUniqueId = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
    n.CommInitRank = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
    n.CommDestroy = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
    n.AllReduce = reinterpret_cast<decltype(n.AllReduce)>(sym("ncclAllReduce"));
    n.AllGather = reinterpret_cast<decltype(n.AllGather)>(sym("ncclAllGather"));
    n.GroupStart = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
    n.GroupEnd = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
    n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
    n.GetVersion = reinterpret_cast<decltype(n.GetVersion)>(sym("ncclGetVersion"));
    if (!ok) return n.err.c_str();
    n.lib = lib;
    return nullptr;
}

constexpr int GS_D = 4, GS_I = 8;   // doubles and 64-bit counters of one mcb_global_stats

// Per-group partial sums over this rank's chains, one block per group, fixed order (no fp64 atomics): group g holds
// the global chains [g * per_group, (g + 1) * per_group); this rank owns [offset, offset + n_local).
__global__ void __launch_bounds__(128) k_stats_groups(const ChainCtl *ctl, int n_local, long long offset, long long per_group,
                                                      double *out_d, long long *out_i)
{
    __shared__ double red[BIG_TILE];
    const long long g0 = (long long)blockIdx.x * per_group, g1 = g0 + per_group;
    const long long lo = max(g0, offset) - offset, hi = min(g1, offset + (long long)n_local) - offset;
    double d[GS_D] = {0, 0, 0, 0}, c[GS_I] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long long k = lo + threadIdx.x; k < hi; k += BIG_TILE) {
        const ChainCtl &s = ctl[k];
        d[0] += s.sum_e; d[1] += s.sum_e2; d[2] += s.production; d[3] += s.toteng;
        c[0] += (double)s.nstat; c[1] += (double)s.prod_acc; c[2] += (double)s.nsuc; c[3] += (double)s.step;
        c[4] += (double)s.sitepairs; c[5] += (double)s.nclamp; c[6] += (double)s.uused; c[7] += 1.0;
    }
    // (the counters are summed as doubles inside the block: exact below 2^53, then stored as integers)
    for (int q = 0; q < GS_D; ++q) { const double v = block_sum_128(d[q], red); if (threadIdx.x == 0) out_d[(size_t)blockIdx.x * GS_D + q] = v; }
    for (int q = 0; q < GS_I; ++q) { const double v = block_sum_128(c[q], red); if (threadIdx.x == 0) out_i[(size_t)blockIdx.x * GS_I + q] = (long long)v; }
}

// what the all-gather carries per chain: final energy, mean energy, acceptance, production mean
__global__ void k_chain_finals(const ChainCtl *ctl, int n_local, int width, double *out)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= width) return;
    double v[4] = {0.0, 0.0, 0.0, 0.0};
    if (c < n_local) {
        const ChainCtl &s = ctl[c];
        v[0] = s.toteng;
        v[1] = s.nstat > 0 ? s.sum_e / (double)s.nstat : 0.0;
        v[2] = s.step > 0 ? (double)s.nsuc / (double)s.step : 0.0;
        v[3] = s.prod_acc > 0 ? s.production / (double)s.prod_acc : 0.0;   // PRODUCTION / PROD_ACC, MonteCarlo.F90:555-560
    }
    for (int q = 0; q < 4; ++q) out[(size_t)c * 4 + q] = v[q];
}

void block_of(int rank, int world, long long n_total, long long &start, long long &count)
{
    const long long base = n_total / world, rem = n_total % world;
    count = base + (rank < rem ? 1 : 0);
    start = (long long)rank * base + std::min<long long>(rank, rem);
}

}  // namespace

struct mcb_comm_state {
    ncclComm_t comm = nullptr;
    bool owned = false;
    int nranks = 1, rank = 0;
    long long n_total = 0, offset = 0;
    double *d_d = nullptr; long long *d_i = nullptr; size_t cap_groups = 0;
    double *d_send = nullptr, *d_recv = nullptr; size_t cap_width = 0;
};

static int nccl_err(mcb_t *h, const char *what, ncclResult_t r)
{
    h->err = std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "NCCL error");
    return -1000 - (int)r;
}

static void comm_free(mcb_t *h)
{
    mcb_comm_state *cs = h->comm;
    if (!cs) return;
    if (cs->comm && cs->owned && g_nccl.CommDestroy) g_nccl.CommDestroy(cs->comm);
    cudaFree(cs->d_d); cudaFree(cs->d_i); cudaFree(cs->d_send); cudaFree(cs->d_recv);
    delete cs;
    h->comm = nullptr;
}

extern "C" int mcb_comm_get_unique_id(char id[MCB_COMM_ID_BYTES])
{
    static_assert(sizeof(ncclUniqueId) == MCB_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    if (!id) return 1;
    if (nccl_bind()) return -1000;
    ncclUniqueId u;
    const ncclResult_t r = g_nccl.GetUniqueId(&u);
    if (r != ncclSuccess) return -1000 - (int)r;
    memcpy(id, &u, sizeof(u));
    return 0;
}

static int comm_setup(mcb_t *h, ncclComm_t comm, bool owned, int nranks, int rank, int64_t n_total_chains)
{
    long long start = 0, count = 0;
    block_of(rank, nranks, n_total_chains, start, count);
    if (count != h->n_chains) {
        if (owned && comm) g_nccl.CommDestroy(comm);
        return arg_err(h, "mcb_comm_init: this handle's n_chains is not rank's block of n_total_chains (contiguous blocks, the first n_total % nranks ranks hold one chain more)");
    }
    comm_free(h);
    mcb_comm_state *cs = new mcb_comm_state;
    cs->comm = comm; cs->owned = owned; cs->nranks = nranks; cs->rank = rank; cs->n_total = n_total_chains; cs->offset = start;
    h->comm = cs;
    return 0;
}

extern "C" int mcb_comm_init(mcb_t *h, int nranks, int rank, const char id[MCB_COMM_ID_BYTES], int64_t n_total_chains)
{
    if (!h) return 1;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks || n_total_chains < nranks) return arg_err(h, "mcb_comm_init: bad argument");
    if (const char *why = nccl_bind()) { h->err = why; return -1000; }
    CK(h, cudaSetDevice(h->device));
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    ncclComm_t comm = nullptr;
    const ncclResult_t r = g_nccl.CommInitRank(&comm, nranks, u, rank);
    if (r != ncclSuccess) return nccl_err(h, "ncclCommInitRank", r);
    return comm_setup(h, comm, true, nranks, rank, n_total_chains);
}

extern "C" int mcb_comm_attach(mcb_t *h, void *nccl_comm, int nranks, int rank, int64_t n_total_chains)
{
    if (!h) return 1;
    if (!nccl_comm || nranks < 1 || rank < 0 || rank >= nranks || n_total_chains < nranks) return arg_err(h, "mcb_comm_attach: bad argument");
    if (const char *why = nccl_bind()) { h->err = why; return -1000; }
    return comm_setup(h, static_cast<ncclComm_t>(nccl_comm), false, nranks, rank, n_total_chains);
}

extern "C" int mcb_comm_destroy(mcb_t *h)
{
    if (!h) return 1;
    comm_free(h);
    return 0;
}

extern "C" int mcb_reduce_stats(mcb_t *h, int64_t chains_per_group, mcb_global_stats *out, int n_groups)
{
    if (!h) return 1;
    if (!out || n_groups < 1) return arg_err(h, "mcb_reduce_stats: null output or n_groups < 1");
    mcb_comm_state local;          // without a communicator: this handle's chains alone (one rank)
    mcb_comm_state *cs = h->comm;
    if (!cs) { local.n_total = h->n_chains; cs = &local; }
    const long long per = chains_per_group > 0 ? chains_per_group : cs->n_total;
    if ((cs->n_total + per - 1) / per != n_groups) return arg_err(h, "mcb_reduce_stats: n_groups must be ceil(n_total_chains / chains_per_group)");
    CK(h, cudaSetDevice(h->device));
    if (int rc = ensure_prepared(h)) return rc;
    double *d_d = cs->d_d; long long *d_i = cs->d_i;
    if ((size_t)n_groups > cs->cap_groups) {
        cudaFree(cs->d_d); cudaFree(cs->d_i); cs->d_d = nullptr; cs->d_i = nullptr; cs->cap_groups = 0;
        CK(h, cudaMalloc(&cs->d_d, sizeof(double) * GS_D * n_groups));
        CK(h, cudaMalloc(&cs->d_i, sizeof(long long) * GS_I * n_groups));
        cs->cap_groups = n_groups;
        d_d = cs->d_d; d_i = cs->d_i;
    }
    k_stats_groups<<<n_groups, BIG_TILE, 0, h->stream>>>(h->ctl, h->n_chains, cs->offset, per, d_d, d_i);
    h->launches++;
    CK(h, cudaGetLastError());
    if (cs->comm && cs->nranks > 1) {
        ncclResult_t r = g_nccl.GroupStart();
        if (r == ncclSuccess) r = g_nccl.AllReduce(d_d, d_d, (size_t)GS_D * n_groups, ncclDouble, ncclSum, cs->comm, h->stream);
        if (r == ncclSuccess) r = g_nccl.AllReduce(d_i, d_i, (size_t)GS_I * n_groups, ncclInt64, ncclSum, cs->comm, h->stream);
        const ncclResult_t r2 = g_nccl.GroupEnd();
        if (r != ncclSuccess || r2 != ncclSuccess) return nccl_err(h, "ncclAllReduce", r != ncclSuccess ? r : r2);
    }
    std::vector<double> hd((size_t)GS_D * n_groups);
    std::vector<long long> hi((size_t)GS_I * n_groups);
    CK(h, cudaMemcpyAsync(hd.data(), d_d, sizeof(double) * hd.size(), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaMemcpyAsync(hi.data(), d_i, sizeof(long long) * hi.size(), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    if (cs == &local) { cudaFree(local.d_d); cudaFree(local.d_i); }
    for (int g = 0; g < n_groups; ++g) {
        mcb_global_stats &o = out[g];
        o.sum_e = hd[GS_D * g]; o.sum_e2 = hd[GS_D * g + 1]; o.production = hd[GS_D * g + 2]; o.sum_toteng = hd[GS_D * g + 3];
        o.n_stat = hi[GS_I * g]; o.prod_acc = hi[GS_I * g + 1]; o.nsuc = hi[GS_I * g + 2]; o.step = hi[GS_I * g + 3];
        o.sitepairs = hi[GS_I * g + 4]; o.n_clamped = hi[GS_I * g + 5]; o.uniforms_used = hi[GS_I * g + 6]; o.n_chains = hi[GS_I * g + 7];
    }
    return 0;
}

extern "C" int mcb_gather_chains(mcb_t *h, double *out, int64_t n_total_chains)
{
    if (!h) return 1;
    if (!out) return arg_err(h, "mcb_gather_chains: null output");
    mcb_comm_state local;
    mcb_comm_state *cs = h->comm;
    if (!cs) { local.n_total = h->n_chains; cs = &local; }
    if (n_total_chains != cs->n_total) return arg_err(h, "mcb_gather_chains: n_total_chains differs from the communicator's");
    CK(h, cudaSetDevice(h->device));
    if (int rc = ensure_prepared(h)) return rc;
    const size_t width = (size_t)((cs->n_total + cs->nranks - 1) / cs->nranks);   // widest block; shorter ones are padded
    if (width > cs->cap_width) {
        cudaFree(cs->d_send); cudaFree(cs->d_recv); cs->d_send = cs->d_recv = nullptr; cs->cap_width = 0;
        CK(h, cudaMalloc(&cs->d_send, sizeof(double) * 4 * width));
        CK(h, cudaMalloc(&cs->d_recv, sizeof(double) * 4 * width * cs->nranks));
        cs->cap_width = width;
    }
    k_chain_finals<<<(unsigned)((width + 127) / 128), 128, 0, h->stream>>>(h->ctl, h->n_chains, (int)width, cs->d_send);
    h->launches++;
    CK(h, cudaGetLastError());
    if (cs->comm && cs->nranks > 1) {
        const ncclResult_t r = g_nccl.AllGather(cs->d_send, cs->d_recv, 4 * width, ncclDouble, cs->comm, h->stream);
        if (r != ncclSuccess) return nccl_err(h, "ncclAllGather", r);
    } else CK(h, cudaMemcpyAsync(cs->d_recv, cs->d_send, sizeof(double) * 4 * width, cudaMemcpyDeviceToDevice, h->stream));
    std::vector<double> tmp(4 * width * cs->nranks);
    CK(h, cudaMemcpyAsync(tmp.data(), cs->d_recv, sizeof(double) * tmp.size(), cudaMemcpyDeviceToHost, h->stream));
    CK(h, cudaStreamSynchronize(h->stream));
    for (int r = 0; r < cs->nranks; ++r) {
        long long start = 0, count = 0;
        block_of(r, cs->nranks, cs->n_total, start, count);
        memcpy(out + 4 * start, tmp.data() + 4 * width * r, sizeof(double) * 4 * count);
    }
    if (cs == &local) { cudaFree(local.d_send); cudaFree(local.d_recv); }
    return 0;
}

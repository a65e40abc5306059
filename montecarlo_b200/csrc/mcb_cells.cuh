// mcb_cells.cuh — the matrix-free full-box energy (BASELINE config 5, ~10^5 molecules) with a cell list.
//
// k_big_pairs tests all N(N-1)/2 centre-of-mass pairs (5.1e9 for config 5) to find the ~8.5e5 inside the 7 A cutoff.
// Here the box is cut into nc^3 cells at least one cutoff wide, every molecule is binned, and a molecule only tests
// the molecules of its own and the 26 neighbouring cells (periodic).  The cells only pre-select CANDIDATES: every
// candidate still gets the reference's exact gate (MonteCarlo.F90:806-812) and every accepted pair the same
// calcLJ with the larger index as MOL1 (:357-364), so the set of pairs and each pair's energy are k_big_pairs's.
//
// That is only valid while the reference's minimum image R - 2*BOXL*AINT(R/BOXL) is a metric, i.e. every centre of
// mass lies inside [-BOXL, BOXL] (true for any box the driver has wrapped, :710-712), and with at least 3 cells
// per side.  k_cell_assign raises a flag otherwise (also when one cell holds more than CELL_MAX molecules, which
// would make the per-cell sort slow) and the host then runs the all-pairs kernel instead.
//
// Determinism: molecules are scattered into their cells with integer atomics and then SORTED by index inside each
// cell, partial sums go through the same fixed-order block and grid reductions as k_big_pairs; identical inputs
// give identical bits.
#pragma once

#include "mcb_kernels.cuh"

namespace mcb {

constexpr int CELL_MAX = 64;        // molecules per cell before the cell path gives up (liquid DMSO: ~3)
constexpr int CELL_NC_MAX = 64;     // cells per side (262 144 cells)

// count[ncell], fill[ncell], start[ncell + 1], cell_of[N], order[N], flag[1]
struct CellGrid { int nc; double scale; int *count, *fill, *start, *cell_of, *order, *flag; };

__global__ void __launch_bounds__(128) k_cell_assign(Geo g, const double *img, CellGrid cg)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N) return;
    const double *com = img + g.o_com;
    const double x = com[j], y = com[g.Np + j], z = com[2 * g.Np + j];
    if (!(fabs(x) <= g.boxl && fabs(y) <= g.boxl && fabs(z) <= g.boxl)) { atomicOr(cg.flag, 1); return; }
    const int cx = min(cg.nc - 1, max(0, (int)((x + g.boxl) * cg.scale)));
    const int cy = min(cg.nc - 1, max(0, (int)((y + g.boxl) * cg.scale)));
    const int cz = min(cg.nc - 1, max(0, (int)((z + g.boxl) * cg.scale)));
    const int cell = (cz * cg.nc + cy) * cg.nc + cx;
    cg.cell_of[j] = cell;
    if (atomicAdd(&cg.count[cell], 1) >= CELL_MAX) atomicOr(cg.flag, 2);
}

// exclusive prefix sum of count[] into start[] (one block; ncell <= 262 144)
__global__ void __launch_bounds__(1024) k_cell_scan(CellGrid cg, int ncell)
{
    __shared__ int part[1024];
    const int t = threadIdx.x;
    const int per = (ncell + 1023) / 1024, lo = t * per, hi = min(ncell, lo + per);
    int s = 0;
    for (int k = lo; k < hi; ++k) s += cg.count[k];
    part[t] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = t >= o ? part[t - o] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    int run = part[t] - s;
    for (int k = lo; k < hi; ++k) { cg.start[k] = run; run += cg.count[k]; }
    if (t == 1023) cg.start[ncell] = part[1023];
}

__global__ void __launch_bounds__(128) k_cell_scatter(Geo g, CellGrid cg)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= g.N || cg.flag[0] != 0) return;
    const int cell = cg.cell_of[j];
    cg.order[cg.start[cell] + atomicAdd(&cg.fill[cell], 1)] = j;
}

// ascending molecule index inside every cell: the order of the scatter's atomics must not reach the sums
__global__ void __launch_bounds__(128) k_cell_sort(CellGrid cg, int ncell)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell || cg.flag[0] != 0) return;
    int *o = cg.order + cg.start[c];
    const int n = cg.start[c + 1] - cg.start[c];
    for (int a = 1; a < n; ++a) {
        const int v = o[a];
        int b = a - 1;
        while (b >= 0 && o[b] > v) { o[b + 1] = o[b]; --b; }
        o[b + 1] = v;
    }
}

// one thread per molecule, taken in cell order so that a warp's threads look at the same few cells
template <bool COUL>
__global__ void __launch_bounds__(128) k_cell_pairs(Geo g, const double *img, CellGrid cg, double *scratch_e, long long *scratch_n)
{
    __shared__ double red[BIG_TILE];
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const double *com = img + g.o_com, *sites = img + g.o_sites, *tabAB = img + g.o_tabAB, *tabQ = img + g.o_tabQ;
    double e = 0.0;
    long long np = 0;
    if (p < g.N && cg.flag[0] == 0) {
        const int m = cg.order[p];
        const int cell = cg.cell_of[m], nc = cg.nc;
        const int cx = cell % nc, cy = (cell / nc) % nc, cz = cell / (nc * nc);
        const double mx = com[m], my = com[g.Np + m], mz = com[2 * g.Np + m];
        const double *s1 = sites + SITE_DOUBLES * (size_t)m;
        for (int dz = -1; dz <= 1; ++dz)
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    const int c2 = (((cz + dz + nc) % nc) * nc + (cy + dy + nc) % nc) * nc + (cx + dx + nc) % nc;
                    const int q1 = cg.start[c2 + 1];
                    for (int q = cg.start[c2]; q < q1; ++q) {
                        const int j = cg.order[q];
                        if (j >= m) break;                           // sorted: the rest of the cell is larger too
                        // R = CoM(J) - CoM(I) with I = m (MOL1, the larger index), J = j (MonteCarlo.F90:806-812)
                        if (pair_r2(com[j] - mx, com[g.Np + j] - my, com[2 * g.Np + j] - mz, g) > g.cut2) continue;
                        ++np;
                        const double *s2 = sites + SITE_DOUBLES * (size_t)j;
                        double en = 0.0;
                        for (int a = 0; a < NSITE; ++a)              // calcLJ: a outer over MOL1, bb inner over MOL2 (LennardJones.f90:65-79)
                            for (int bb = 0; bb < NSITE; ++bb)
                                en = site_pair_acc<COUL>(en, s2[3 * bb] - s1[3 * a], s2[3 * bb + 1] - s1[3 * a + 1], s2[3 * bb + 2] - s1[3 * a + 2],
                                                         tabAB[a * TAB_ROW + 2 * bb], tabAB[a * TAB_ROW + 2 * bb + 1], COUL ? tabQ[a * 4 + bb] : 0.0, g);
                        e += en;
                    }
                }
    }
    const double be = block_sum_128(e, red);
    const double bn = block_sum_128((double)np, red);
    if (threadIdx.x == 0) { scratch_e[blockIdx.x] = be; scratch_n[blockIdx.x] = (long long)bn; }
}

}  // namespace mcb

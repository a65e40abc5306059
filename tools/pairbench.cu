// Micro-benchmark (debug aid): cycles per site pair for ONE warp running the kernel's pair arithmetic with
// ILP independent pairs in flight, and the same with 2 and 4 warps per SM sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
#include "../montecarlo_b200/csrc/mcb_device.cuh"
using namespace mcb;

template <int ILP, int VARIANT> __global__ void k(Geo g, double *out, long long *cyc, int iters)
{
    double x[ILP], acc[ILP];
    for (int q = 0; q < ILP; ++q) { x[q] = 0.37 * q + threadIdx.x * 0.011; acc[q] = 0; }
    const double sx = 0.3, sy = -1.1, sz = 2.2, A = 1e6, B = 1e3;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int q = 0; q < ILP; ++q) {
            double dx = x[q] - sx, dy = x[q] * 0.5 - sy, dz = x[q] * 0.25 - sz;
            if (VARIANT == 0) acc[q] = site_pair_acc<false>(acc[q], dx, dy, dz, A, B, 0.0, g);
            if (VARIANT == 1) {   // trunc-based MIC for comparison
                dx = fma(-g.boxl2, trunc(dx * g.inv_boxl), dx); dy = fma(-g.boxl2, trunc(dy * g.inv_boxl), dy);
                dz = fma(-g.boxl2, trunc(dz * g.inv_boxl), dz);
                double r2 = fma(dz, dz, fma(dy, dy, dx * dx)), inv = rcp_nr(r2), s6 = inv * inv * inv;
                acc[q] = fma(s6, fma(A, s6, -B), acc[q]);
            }
            if (VARIANT == 2) {   // no MIC at all
                double r2 = fma(dz, dz, fma(dy, dy, dx * dx)), inv = rcp_nr(r2), s6 = inv * inv * inv;
                acc[q] = fma(s6, fma(A, s6, -B), acc[q]);
            }
            if (VARIANT == 3) { acc[q] += mic_abs(dx, g) + mic_abs(dy, g) + mic_abs(dz, g); }   // 3 sub + 9 MIC + 3 add
            if (VARIANT == 4) { acc[q] += rcp_nr(fma(dz, dz, fma(dy, dy, dx * dx))); }          // r2 + rcp
            if (VARIANT == 5) { double s6 = dx * dx * dx; acc[q] = fma(s6, fma(A, s6, -B), acc[q]); }
            if (VARIANT == 6) { double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(dx)); acc[q] += y; }
            if (VARIANT == 7) { double r2 = fma(dz, dz, fma(dy, dy, dx * dx)); double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(r2));
                                double e = fma(-r2, y, 1.0); acc[q] += fma(y, fma(e, e, e), y) ; }
            x[q] += 1e-3;
        }
    }
    long long t1 = clock64();
    double s = 0;
    for (int q = 0; q < ILP; ++q) s += acc[q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP> __global__ void k_staged(Geo g, double *out, long long *cyc, int iters)
{
    double x[ILP], acc[ILP];
    for (int q = 0; q < ILP; ++q) { x[q] = 0.37 * q + threadIdx.x * 0.011; acc[q] = 0; }
    const double sx = 0.3, sy = -1.1, sz = 2.2, A = 1e6, B = 1e3;
    const double M = 4503599627370496.0;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        double dx[ILP], dy[ILP], dz[ILP], r2[ILP], y[ILP], e[ILP];
#pragma unroll
        for (int q = 0; q < ILP; ++q) { dx[q] = fabs(x[q] - sx); dy[q] = fabs(x[q] * 0.5 - sy); dz[q] = fabs(x[q] * 0.25 - sz); }
        double tx[ILP], ty[ILP], tz[ILP];
#pragma unroll
        for (int q = 0; q < ILP; ++q) { tx[q] = __fma_rd(dx[q], g.inv_boxl, M); ty[q] = __fma_rd(dy[q], g.inv_boxl, M); tz[q] = __fma_rd(dz[q], g.inv_boxl, M); }
#pragma unroll
        for (int q = 0; q < ILP; ++q) { tx[q] -= M; ty[q] -= M; tz[q] -= M; }
#pragma unroll
        for (int q = 0; q < ILP; ++q) { dx[q] = fma(-g.boxl2, tx[q], dx[q]); dy[q] = fma(-g.boxl2, ty[q], dy[q]); dz[q] = fma(-g.boxl2, tz[q], dz[q]); }
#pragma unroll
        for (int q = 0; q < ILP; ++q) r2[q] = dx[q] * dx[q];
#pragma unroll
        for (int q = 0; q < ILP; ++q) r2[q] = fma(dy[q], dy[q], r2[q]);
#pragma unroll
        for (int q = 0; q < ILP; ++q) r2[q] = fma(dz[q], dz[q], r2[q]);
#pragma unroll
        for (int q = 0; q < ILP; ++q) asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y[q]) : "d"(r2[q]));
#pragma unroll
        for (int q = 0; q < ILP; ++q) e[q] = fma(-r2[q], y[q], 1.0);
#pragma unroll
        for (int q = 0; q < ILP; ++q) e[q] = fma(e[q], e[q], e[q]);
#pragma unroll
        for (int q = 0; q < ILP; ++q) y[q] = fma(y[q], e[q], y[q]);
#pragma unroll
        for (int q = 0; q < ILP; ++q) e[q] = y[q] * y[q];
#pragma unroll
        for (int q = 0; q < ILP; ++q) e[q] = e[q] * y[q];
#pragma unroll
        for (int q = 0; q < ILP; ++q) y[q] = fma(A, e[q], -B);
#pragma unroll
        for (int q = 0; q < ILP; ++q) { acc[q] = fma(e[q], y[q], acc[q]); x[q] += 1e-3; }
    }
    long long t1 = clock64();
    double s = 0;
    for (int q = 0; q < ILP; ++q) s += acc[q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int ILP> void run_staged(int warps_per_block, int blocks)
{
    Geo g{}; g.boxl = 7.78; g.boxl2 = 15.56; g.inv_boxl = 1 / 7.78;
    double *out; long long *cyc, h; int iters = 2000;
    cudaMalloc(&out, 8 * 32 * warps_per_block * blocks); cudaMalloc(&cyc, 8);
    k_staged<ILP><<<blocks, 32 * warps_per_block>>>(g, out, cyc, iters);
    k_staged<ILP><<<blocks, 32 * warps_per_block>>>(g, out, cyc, iters);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-26s ILP=%d warps/SM=%2d : %6.1f cycles per pair per warp\n", "pair STAGED", ILP, warps_per_block, (double)h / iters / ILP);
    cudaFree(out); cudaFree(cyc);
}

template <int ILP, int VARIANT> void run(const char *name, int warps_per_block, int blocks)
{
    Geo g{}; g.boxl = 7.78; g.boxl2 = 15.56; g.inv_boxl = 1 / 7.78;
    double *out; long long *cyc, h; int iters = 2000;
    cudaMalloc(&out, 8 * 32 * warps_per_block * blocks); cudaMalloc(&cyc, 8);
    k<ILP, VARIANT><<<blocks, 32 * warps_per_block>>>(g, out, cyc, iters);
    k<ILP, VARIANT><<<blocks, 32 * warps_per_block>>>(g, out, cyc, iters);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-26s ILP=%d warps/SM=%2d : %6.1f cycles per pair per warp\n", name, ILP, warps_per_block, (double)h / iters / ILP);
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    run<1, 0>("pair (floor-magic MIC)", 1, 1); run<4, 0>("pair (floor-magic MIC)", 1, 1); run<8, 0>("pair (floor-magic MIC)", 1, 1);
    run<8, 0>("pair (floor-magic MIC)", 4, 148); run<8, 0>("pair (floor-magic MIC)", 8, 148); run<8, 0>("pair (floor-magic MIC)", 16, 148);
    run<8, 1>("pair (trunc MIC)", 1, 1); run<8, 1>("pair (trunc MIC)", 8, 148);
    run<8, 2>("pair (no MIC)", 1, 1); run<8, 2>("pair (no MIC)", 8, 148);
    run<8, 3>("MIC x3 only (15 ops)", 1, 1); run<8, 4>("r2 + rcp_nr (3+3+1 ops, MUFU)", 1, 1); run<8, 5>("s6 + lj (4 ops)", 1, 1);
    run<8, 6>("3 sub + MUFU + add", 1, 1); run<8, 7>("r2 + inline rcp", 1, 1); run<4, 4>("r2 + rcp_nr ILP4", 1, 1); run<2, 4>("r2 + rcp_nr ILP2", 1, 1);
    run_staged<4>(1, 1); run_staged<8>(1, 1); run_staged<16>(1, 1); run_staged<8>(8, 148);
    return 0;
}

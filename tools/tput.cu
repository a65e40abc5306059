// Micro-benchmark (debug aid): single-warp throughput of fp64 ops with distinct register operands.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP, int ILP> __global__ void k(double *out, long long *cyc, int iters, double s0)
{
    double v[ILP], w[ILP], z[ILP];
    for (int q = 0; q < ILP; ++q) { v[q] = s0 + q * 0.01 + threadIdx.x * 1e-3; w[q] = 1.0 - 1e-9 * q; z[q] = 1e-7 * (q + 1); }
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int q = 0; q < ILP; ++q) {
                if (OP == 0) v[q] = fma(v[q], w[q], z[q]);
                if (OP == 1) v[q] = v[q] + z[q];
                if (OP == 2) v[q] = v[q] * w[q];
                if (OP == 3) v[q] = __fma_rd(fabs(v[q]), w[q], z[q]);
                if (OP == 4) v[q] = fma(v[q], w[q], 4503599627370496.0) - 4503599627370496.0;
                if (OP == 5) { v[q] = fma(v[q], w[q], z[q]); w[q] = fma(w[q], z[q], v[q]); }   // 2 ops, more operands
                if (OP == 6) v[q] = fma(v[q], v[q], v[q]);
            }
    }
    long long t1 = clock64();
    double s = 0;
    for (int q = 0; q < ILP; ++q) s += v[q] + w[q];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}
template <int OP, int ILP> void run(const char *name, int nops)
{
    double *out; long long *cyc, h; int iters = 1000;
    cudaMalloc(&out, 8 * 32); cudaMalloc(&cyc, 8);
    k<OP, ILP><<<1, 32>>>(out, cyc, iters, 0.5); k<OP, ILP><<<1, 32>>>(out, cyc, iters, 0.5);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-34s ILP=%2d : %5.2f cycles per fp64 op\n", name, ILP, (double)h / iters / 4 / ILP / nops);
}
int main()
{
    run<0, 8>("DFMA v=fma(v,w,z) distinct regs", 1); run<0, 16>("DFMA v=fma(v,w,z) distinct regs", 1);
    run<1, 8>("DADD v=v+z", 1); run<2, 8>("DMUL v=v*w", 1);
    run<3, 8>("DFMA.RM |v|", 1); run<4, 8>("DFMA imm + DADD imm", 2); run<5, 8>("2 x DFMA cross", 2); run<6, 8>("DFMA v=fma(v,v,v)", 1);
    return 0;
}

// Micro-benchmark (debug aid, not product): dependent-issue latency and single-warp throughput of the
// instructions the step kernel is made of.  nvcc -arch=sm_100a -O3 -o latency tools/latency.cu && ./latency
#include <cstdio>
#include <cuda_runtime.h>

#define REP 256
template <int OP> __device__ __forceinline__ double op(double x, double y, int lane)
{
    if (OP == 0) return fma(x, y, y);                       // DFMA
    if (OP == 1) return x + y;                              // DADD
    if (OP == 2) return x * y;                              // DMUL
    if (OP == 3) return trunc(x) + 0.0 * y;                 // FRND.TRUNC (+DFMA)
    if (OP == 4) { double r; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x)); return r; }  // MUFU.RCP64H
    if (OP == 5) return __shfl_xor_sync(0xffffffffu, x, 1);  // 2 SHFL
    if (OP == 6) return 1.0 / x;                            // full division
    if (OP == 7) return sqrt(x);
    if (OP == 8) { double s, c; sincos(x, &s, &c); return s + c; }
    if (OP == 9) return exp(x * 1e-3);
    if (OP == 10) return (double)(long long)(x);            // F2I + I2F
    if (OP == 11) return rsqrt(x);
    return x;
}

template <int OP, int ILP> __global__ void k(double *out, long long *cyc, double seed)
{
    double v[ILP];
    for (int q = 0; q < ILP; ++q) v[q] = seed + q * 0.01 + threadIdx.x * 1e-3;
    double y = 1.0000001;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < REP / 8; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int q = 0; q < ILP; ++q) v[q] = op<OP>(v[q], y, threadIdx.x);
    }
    long long t1 = clock64();
    double s = 0;
    for (int q = 0; q < ILP; ++q) s += v[q];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int OP, int ILP> void run(const char *name, double seed)
{
    double *out; long long *cyc, h;
    cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 8);
    k<OP, ILP><<<1, 32>>>(out, cyc, seed);
    k<OP, ILP><<<1, 32>>>(out, cyc, seed);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-22s ILP=%d  %7.1f cycles per op-round, %6.1f per op\n", name, ILP, (double)h / REP, (double)h / REP / ILP);
    cudaFree(out); cudaFree(cyc);
}

__global__ void k_lds(double *out, long long *cyc)
{
    __shared__ int next[64];
    __shared__ double val[64];
    for (int i = threadIdx.x; i < 64; i += 32) { next[i] = (i * 7 + 3) & 63; val[i] = i; }
    __syncwarp();
    int p = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 8
    for (int it = 0; it < REP; ++it) p = next[p];
    long long t1 = clock64();
    double acc = 0; int q = threadIdx.x;
    long long t2 = clock64();
#pragma unroll 8
    for (int it = 0; it < REP; ++it) { unsigned b = __ballot_sync(0xffffffffu, (q & 1)); q = __popc(b) + q; }
    long long t3 = clock64();
    out[threadIdx.x] = p + acc + q;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t3 - t2; }
}

int main()
{
    run<0, 1>("DFMA", 0.5); run<0, 4>("DFMA", 0.5); run<0, 8>("DFMA", 0.5);
    run<1, 1>("DADD", 0.5); run<2, 1>("DMUL", 1.0);
    run<3, 1>("FRND.TRUNC+DFMA", 5.5); run<3, 4>("FRND.TRUNC+DFMA", 5.5); run<3, 8>("FRND.TRUNC+DFMA", 5.5);
    run<4, 1>("MUFU.RCP64H", 1.5); run<4, 4>("MUFU.RCP64H", 1.5); run<4, 8>("MUFU.RCP64H", 1.5);
    run<5, 1>("SHFL x2 (double)", 1.5); run<5, 4>("SHFL x2 (double)", 1.5);
    run<6, 1>("1.0/x", 1.5); run<6, 4>("1.0/x", 1.5);
    run<7, 1>("sqrt", 1.5); run<8, 1>("sincos", 0.7); run<8, 4>("sincos", 0.7); run<9, 1>("exp", 0.7);
    run<10, 1>("F2I+I2F (64)", 5.5); run<11, 1>("rsqrt", 1.5); run<11, 4>("rsqrt", 1.5);
    double *out; long long *cyc, h[2];
    cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 16);
    k_lds<<<1, 32>>>(out, cyc); k_lds<<<1, 32>>>(out, cyc);
    cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
    printf("LDS dependent chain     %7.1f cycles per load\nballot+popc chain       %7.1f cycles\n", (double)h[0] / REP, (double)h[1] / REP);
    return 0;
}

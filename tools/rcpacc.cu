// Accuracy of the MUFU.RCP64H seed (rcp.approx.ftz.f64) and of one quadratic / one cubic correction step on it,
// over squared site distances (tools only).   nvcc -arch=sm_100a -o tools/rcpacc_bin tools/rcpacc.cu
#include <cstdio>
#include <cmath>
__global__ void k(double *out, int n)
{
    double ms = 0, mq = 0, mc = 0;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const double x = 0.5 + 400.0 * (double)t / (double)n + 1e-7 * (t % 977);
        double y;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
        const double ex = 1.0 / x;
        const double e = fma(-x, y, 1.0);
        const double yq = fma(y, e, y);
        const double yc = fma(y, fma(e, e, e), y);
        ms = fmax(ms, fabs(y - ex) / ex); mq = fmax(mq, fabs(yq - ex) / ex); mc = fmax(mc, fabs(yc - ex) / ex);
    }
    atomicMax((unsigned long long *)&out[0], __double_as_longlong(ms));
    atomicMax((unsigned long long *)&out[1], __double_as_longlong(mq));
    atomicMax((unsigned long long *)&out[2], __double_as_longlong(mc));
}
int main()
{
    double *d, h[3] = {0, 0, 0};
    cudaMalloc(&d, 24); cudaMemcpy(d, h, 24, cudaMemcpyHostToDevice);
    k<<<1184, 256>>>(d, 1 << 28);
    cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
    printf("max relative error over 2^28 arguments in [0.5, 400]: seed %.3e (2^%.1f)  +quadratic step %.3e  +cubic step %.3e\n", h[0], log2(h[0]), h[1], h[2]);
    return 0;
}

// FP64 (non-tensor) issue rate and latencies on one SM: what bounds the rotation kernel of the block Jacobi.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dfma_probe tools/dfma_probe.cu && ./tools/dfma_probe
#include <cstdio>
#include <cuda_runtime.h>
__global__ void thr(double* out, int iters, long long* cyc)
{
    double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    __syncthreads();
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> __global__ void lat(double* out, int iters, long long* cyc)
{
    double x = 1.5 + threadIdx.x * 1e-3;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (OP == 0) x = fma(x, 1.0000001, 1e-9);
        if (OP == 1) x = sqrt(x) + 1.0;
        if (OP == 2) x = 1.0 / x + 1.0;
        if (OP == 3) x = rsqrt(x) + 1.0;
        if (OP == 4) x = (double)rsqrtf((float)x) + 1.0;
    }
    const long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
int main()
{
    double* out; long long* cyc; long long h;
    cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
    const int iters = 4096;
    for (int warps : {1, 2, 4, 8, 16, 32}) {
        thr<<<1, warps * 32>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DFMA throughput, %2d warps on one SM: %.2f lane-FMA / clk / SM (%.1f cycles per warp instruction per SMSP)\n", warps,
               (double)iters * 8 * warps * 32 / h, (double)h / (iters * 8.0 * ((warps + 3) / 4)));
    }
    const char* names[] = {"dependent DFMA", "sqrt + add", "1/x + add", "rsqrt + add", "rsqrtf(float) + cvt + add"};
    lat<0><<<1, 32>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("%s: %.1f cycles\n", names[0], (double)h / iters);
    lat<1><<<1, 32>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("%s: %.1f cycles\n", names[1], (double)h / iters);
    lat<2><<<1, 32>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("%s: %.1f cycles\n", names[2], (double)h / iters);
    lat<3><<<1, 32>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("%s: %.1f cycles\n", names[3], (double)h / iters);
    lat<4><<<1, 32>>>(out, iters, cyc); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); printf("%s: %.1f cycles\n", names[4], (double)h / iters);
    return cudaDeviceSynchronize() != cudaSuccess;
}

// FP64 tensor-pipe (DMMA) issue-rate ceiling on this GPU: every warp runs NACC independent
// mma.sync.m8n8k4.f64 accumulation chains from registers (no memory traffic).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_peak tools/dmma_peak.cu
// Output: one JSON line {"dmma_peak_tflops": ..., "warps_per_sm": ..., ...} per configuration.
#include <cstdio>
#include <cuda_runtime.h>

template <int NACC>
__global__ void dmma_loop(double* out, int iters)
{
    double c0[NACC], c1[NACC];
#pragma unroll
    for (int i = 0; i < NACC; ++i) { c0[i] = 0.0; c1[i] = 0.0; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c0[i] + c1[i];
    if (s == 123.456) out[0] = s;
}

template <int NACC>
static void run(int sms, int threads, int ctas_per_sm)
{
    double* out;
    cudaMalloc(&out, 8);
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    dmma_loop<NACC><<<sms * ctas_per_sm, threads>>>(out, 100);
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        dmma_loop<NACC><<<sms * ctas_per_sm, threads>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double flops = (double)sms * ctas_per_sm * (threads / 32) * (double)iters * NACC * 512.0;
    printf("{\"dmma_peak_tflops\": %.3f, \"nacc\": %d, \"threads\": %d, \"ctas_per_sm\": %d, \"ms\": %.3f}\n",
           flops / (best * 1e-3) / 1e12, NACC, threads, ctas_per_sm, best);
    cudaFree(out);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    run<4>(sms, 128, 1); run<8>(sms, 128, 1); run<8>(sms, 256, 1); run<16>(sms, 256, 1);
    run<8>(sms, 512, 1); run<16>(sms, 512, 1); run<8>(sms, 1024, 1); run<8>(sms, 256, 4);
    return 0;
}

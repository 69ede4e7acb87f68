// probe_ffma2.cu -- stand-alone hardware probe (test infrastructure): throughput of packed fp32 FMA
// (fma.rn.f32x2 -> FFMA2) against scalar FFMA on sm_100a, per SM and clock, at 8 independent chains per thread.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_ffma2 tests/cuda/probe_ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long pack(float a, float b)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, int iters, float m, float c)
{
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = threadIdx.x * 0.001f + j;
    if (MODE == 0) {
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = fmaf(acc[j], m, c);
        }
    } else {
        unsigned long long a2[8];
        const unsigned long long m2 = pack(m, m), c2 = pack(c, c);
#pragma unroll
        for (int j = 0; j < 8; ++j) a2[j] = pack(acc[2 * j], acc[2 * j + 1]);
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int j = 0; j < 8; ++j) a2[j] = fma2(a2[j], m2, c2);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float lo, hi;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a2[j]));
            acc[2 * j] = lo; acc[2 * j + 1] = hi;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 8, iters = 20000;
    float *out;
    cudaMalloc(&out, blocks * 256 * sizeof(float));
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    for (int mode = 0; mode < 2; ++mode) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<blocks, 256>>>(out, iters, 0.999f, 0.001f);
            else k<1><<<blocks, 256>>>(out, iters, 0.999f, 0.001f);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
        }
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double fmas = (double)blocks * 256 * 16.0 * iters;
        printf("%s: %.3f ms, %.1f fp32 FMA / clk / SM at the nominal %d MHz (%.1f TFLOP/s)\n", mode ? "FFMA2 (f32x2)" : "FFMA scalar  ",
               ms, fmas / (ms * 1e-3) / sms / (khz * 1e3), khz / 1000, 2 * fmas / (ms * 1e-3) / 1e12);
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    return 0;
}

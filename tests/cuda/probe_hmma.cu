// probe_hmma.cu -- stand-alone hardware probe (test infrastructure): the warp-level mma.sync path on B200.
// The head's one-pass backward (32 -> 16 -> 8 -> 1) is a chain of tiny matrix products whose operands the tcgen05
// route has to stage through shared memory / TMEM; the warp-level MMA takes them from registers. Questions:
//   1. how many m16n8k8 tf32 MMAs per clock does an SM sustain (4 .. 16 warps, 8 independent accumulators each)?
//   2. does the 3xTF32 split with TRUNCATED high parts (hi = word as the tensor core reads it, lo = v - hi) reach
//      fp32 accuracy on this path as it does on tcgen05?
//   3. how does the fp32 accumulator round (tcgen05 adds with round-toward-zero)?
//   4. 8x8 transpose of 32-bit words in registers by two movmatrix on the 16-bit halves (L_M -> L_N fragment layouts)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_hmma tests/cuda/probe_hmma.cu
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int KIND, int ILP, int FMAS>
__global__ void __launch_bounds__(512) rate(float *out, int iters, long long *clk)
{
    float d[ILP][4];
    uint32_t a[4], b[2];
    for (int i = 0; i < 4; ++i) a[i] = __float_as_uint(1.0f + threadIdx.x * 1e-3f + i);
    for (int i = 0; i < 2; ++i) b[i] = __float_as_uint(0.5f + threadIdx.x * 1e-3f + i);
    for (int j = 0; j < ILP; ++j)
        for (int i = 0; i < 4; ++i) d[j][i] = 0.f;
    float f[8];
    for (int i = 0; i < 8; ++i) f[i] = threadIdx.x * 0.25f + i;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < ILP; ++j) {
            if (KIND == 0) mma_tf32(d[j], a, b);
            else mma_bf16(d[j], a, b);
#pragma unroll
            for (int q = 0; q < FMAS; ++q) f[q & 7] = fmaf(f[q & 7], 1.0001f, 0.5f);
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int j = 0; j < ILP; ++j)
        for (int i = 0; i < 4; ++i) s += d[j][i];
    for (int i = 0; i < 8; ++i) s += f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

// C[16 x 8] = A[16 x K] * B[K x 8] with 3xTF32 (truncation split), one warp; A row-major, B [k][n]
__global__ void numerics(const float *A, const float *B, float *C, int K, int passes)
{
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    float d[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k0 = 0; k0 < K; k0 += 8) {
        float av[4] = {A[g * K + k0 + t], A[(g + 8) * K + k0 + t], A[g * K + k0 + t + 4], A[(g + 8) * K + k0 + t + 4]};
        float bv[2] = {B[(k0 + t) * 8 + g], B[(k0 + t + 4) * 8 + g]};
        uint32_t ah[4], al[4], bh[2], bl[2];
        for (int i = 0; i < 4; ++i) {
            ah[i] = __float_as_uint(av[i]);
            al[i] = __float_as_uint(av[i] - __uint_as_float(ah[i] & 0xffffe000u));
        }
        for (int i = 0; i < 2; ++i) {
            bh[i] = __float_as_uint(bv[i]);
            bl[i] = __float_as_uint(bv[i] - __uint_as_float(bh[i] & 0xffffe000u));
        }
        if (passes == 3) {
            mma_tf32(d, al, bh);
            mma_tf32(d, ah, bl);
        }
        mma_tf32(d, ah, bh);
    }
    C[g * 8 + 2 * t] = d[0];
    C[g * 8 + 2 * t + 1] = d[1];
    C[(g + 8) * 8 + 2 * t] = d[2];
    C[(g + 8) * 8 + 2 * t + 1] = d[3];
}

// accumulator rounding: c = 1.0, product = p (exact); returns 1.0 (+) p as the tensor core adds it
__global__ void rounding(float *out)
{
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    const float ps[4] = {1.5f * 5.9604645e-8f, 0.75f * 5.9604645e-8f, -0.75f * 5.9604645e-8f, -0.25f * 5.9604645e-8f};   // x 2^-24
    for (int c = 0; c < 4; ++c) {
        uint32_t a[4] = {0, 0, 0, 0}, b[2] = {0, 0};
        if (t == 0) a[0] = __float_as_uint(1.0f), a[1] = __float_as_uint(1.0f);      // A[row][0] = 1
        if (t == 0) b[0] = __float_as_uint(ps[c]);                                      // B[0][n] = p
        float d[4] = {1.f, 1.f, 1.f, 1.f};
        mma_tf32(d, a, b);
        if (lane == 0) out[c] = d[0];
    }
    (void)g;
}

// 8x8 transpose of 32-bit words: in: thread (g,t) holds E[g][2t], E[g][2t+1]; out: E[2t][g], E[2t+1][g]
__device__ __forceinline__ uint32_t movm(uint32_t x)
{
    uint32_t y;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(y) : "r"(x));
    return y;
}
__global__ void transpose(uint32_t *out)
{
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    const uint32_t e0 = 0x01000000u * (g * 8 + 2 * t) + 0x00abc000u + (g * 8 + 2 * t);
    const uint32_t e1 = 0x01000000u * (g * 8 + 2 * t + 1) + 0x00abc000u + (g * 8 + 2 * t + 1);
    const uint32_t hi = __byte_perm(e0, e1, 0x7632), lo = __byte_perm(e0, e1, 0x5410);
    const uint32_t th = movm(hi), tl = movm(lo);
    out[lane * 2] = __byte_perm(tl, th, 0x5410);
    out[lane * 2 + 1] = __byte_perm(tl, th, 0x7632);
}

template <int KIND, int ILP, int FMAS>
static int run_rate(const char *name, int warps, int sms)
{
    float *out;
    long long *clk;
    CK(cudaMalloc(&out, sizeof(float) * sms * 512));
    CK(cudaMalloc(&clk, sizeof(long long) * sms));
    const int iters = 20000;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    rate<KIND, ILP, FMAS><<<sms, warps * 32>>>(out, 100, clk);
    CK(cudaEventRecord(e0));
    rate<KIND, ILP, FMAS><<<sms, warps * 32>>>(out, iters, clk);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<long long> h(sms);
    CK(cudaMemcpy(h.data(), clk, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
    const double mmas = (double)iters * ILP * warps;
    const double macs = mmas * (KIND == 0 ? 1024.0 : 2048.0);
    printf("%-28s warps %2d ilp %d fma/mma %d: %8.1f clk/mma/SM (%.3f mma/clk/SM), %7.1f MAC/clk/SM, %6.1f TFLOP/s chip (event %.3f ms)\n", name, warps, ILP, FMAS,
           h[0] / mmas, mmas / h[0], macs / h[0], 2.0 * macs * sms / (ms * 1e-3) * 1e-12, ms);
    CK(cudaFree(out));
    CK(cudaFree(clk));
    return 0;
}

int main()
{
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("%s, %d SMs\n", prop.name, sms);
    for (int w : {4, 8, 12, 16}) {
        if (run_rate<0, 8, 0>("tf32 m16n8k8", w, sms)) return 1;
    }
    for (int w : {4, 8, 16}) {
        if (run_rate<0, 4, 0>("tf32 m16n8k8", w, sms)) return 1;
    }
    for (int w : {8, 16}) {
        if (run_rate<0, 8, 4>("tf32 m16n8k8 + 4 ffma", w, sms)) return 1;
        if (run_rate<0, 8, 8>("tf32 m16n8k8 + 8 ffma", w, sms)) return 1;
    }
    for (int w : {4, 8, 16}) {
        if (run_rate<1, 8, 0>("bf16 m16n8k16", w, sms)) return 1;
    }
    // numerics
    {
        const int K = 32;
        std::vector<float> A(16 * K), B(K * 8), C(128), C1(128);
        srand(1);
        for (auto &v : A) v = (float)rand() / RAND_MAX * 2.f - 1.f;
        for (auto &v : B) v = (float)rand() / RAND_MAX * 2.f - 1.f;
        float *dA, *dB, *dC;
        CK(cudaMalloc(&dA, A.size() * 4));
        CK(cudaMalloc(&dB, B.size() * 4));
        CK(cudaMalloc(&dC, 128 * 4));
        CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
        for (int passes : {1, 3}) {
            numerics<<<1, 32>>>(dA, dB, dC, K, passes);
            CK(cudaMemcpy(C.data(), dC, 128 * 4, cudaMemcpyDeviceToHost));
            double emax = 0, peak = 0;
            for (int m = 0; m < 16; ++m)
                for (int n = 0; n < 8; ++n) {
                    double r = 0;
                    for (int k = 0; k < K; ++k) r += (double)A[m * K + k] * B[k * 8 + n];
                    emax = fmax(emax, fabs(r - C[m * 8 + n]));
                    peak = fmax(peak, fabs(r));
                }
            printf("numerics K=32, %d pass(es): max |err| / peak = %.3e\n", passes, emax / peak);
        }
        float *dr;
        CK(cudaMalloc(&dr, 16));
        rounding<<<1, 32>>>(dr);
        float r[4];
        CK(cudaMemcpy(r, dr, 16, cudaMemcpyDeviceToHost));
        printf("accumulator rounding: 1 + 1.5*2^-24 -> 1 + %.2f*2^-23 (RN: 1, RZ: 0); 1 + 0.75*2^-24 -> %.2f (RN: 0, RZ: 0... RU: 1); "
               "1 - 0.75*2^-24 -> 1 - %.2f*2^-24 (RN: 1, RZ: 1 [toward zero keeps 1-2^-24], RD: 1); 1 - 0.25*2^-24 -> 1 - %.2f*2^-24 (RN: 0, RZ: 1)\n",
               (r[0] - 1.f) / 1.1920929e-7f, (r[1] - 1.f) / 1.1920929e-7f, (1.f - r[2]) / 5.9604645e-8f, (1.f - r[3]) / 5.9604645e-8f);
        uint32_t *dt;
        CK(cudaMalloc(&dt, 64 * 4));
        transpose<<<1, 32>>>(dt);
        uint32_t ht[64];
        CK(cudaMemcpy(ht, dt, 64 * 4, cudaMemcpyDeviceToHost));
        int bad = 0;
        for (int lane = 0; lane < 32; ++lane) {
            const int g = lane >> 2, t = lane & 3;
            const uint32_t w0 = 0x01000000u * ((2 * t) * 8 + g) + 0x00abc000u + ((2 * t) * 8 + g);
            const uint32_t w1 = 0x01000000u * ((2 * t + 1) * 8 + g) + 0x00abc000u + ((2 * t + 1) * 8 + g);
            if (ht[lane * 2] != w0 || ht[lane * 2 + 1] != w1) ++bad;
        }
        printf("movmatrix 32-bit transpose: %s (%d lanes differ; lane 1 got %08x %08x)\n", bad ? "MISMATCH" : "ok", bad, ht[2], ht[3]);
    }
    return 0;
}

// probe_sw128.cu -- stand-alone hardware probe (test infrastructure): 128-byte-swizzled operand
// layouts for tcgen05.mma kind::tf32 on sm_100a, as used by the fused backward kernel.
//   a_mode 0: A from TMEM (thread = row = lane), 1: A smem K-major SW128, 2: A smem MN-major SW128
//   b_mode 1: B smem K-major SW128, 2: B smem MN-major SW128
// K-major SW128 image of X[R][K]:  k-block kb = k/32 at kb*R*128; row r at r*128; 16-byte chunk
// (k%32)/4 XOR (r%8).  MN-major SW128 image of X[R][K]: mn-block r/32 at (r/32)*K*128; k at k*128;
// chunk (r%32)/4 XOR (k%8).  (The K-major image of X is the MN-major image of X^T.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_sw128 tests/cuda/probe_sw128.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../bts_b200/csrc/tc.cuh"

using namespace bts::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __host__ inline uint32_t kmaj_off(int r, int k, int R) { return (k / 32) * R * 128 + r * 128 + ((((k % 32) / 4) ^ (r % 8)) * 16) + (k % 4) * 4; }
__device__ __host__ inline uint32_t mnmaj_off(int r, int k, int K) { return (r / 32) * K * 128 + k * 128 + ((((r % 32) / 4) ^ (k % 8)) * 16) + (r % 4) * 4; }

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr, uint32_t lbo, uint32_t sbo)
{
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
           (1ull << 46) | (2ull << 61);
}

template <int M, int N, int K>
__global__ void __launch_bounds__(128) probe(const float *A, const float *Bm, float *D, int a_mode, int b_mode)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned char *sa = smem;
    unsigned char *sb = smem + ((M * K * 4 + 1023) / 1024) * 1024;
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    const int t = threadIdx.x, warp = t >> 5;
    if (a_mode != 0)
        for (int e = t; e < M * K; e += 128) {
            const int r = e / K, k = e % K;
            *reinterpret_cast<float *>(sa + (a_mode == 1 ? kmaj_off(r, k, M) : mnmaj_off(r, k, K))) = A[e];
        }
    for (int e = t; e < N * K; e += 128) {
        const int r = e / K, k = e % K;
        *reinterpret_cast<float *>(sb + (b_mode == 1 ? kmaj_off(r, k, N) : mnmaj_off(r, k, K))) = Bm[e];
    }
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&tbase_slot, 512);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = tbase_slot;
    const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    {
        uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int n0 = 0; n0 < N; n0 += 8) tmem_st<8>(lane_base + 256 + n0, z);
    }
    if (a_mode == 0) {
        for (int k0 = 0; k0 < K; k0 += 8) {
            uint32_t hi[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) hi[j] = __float_as_uint(A[(size_t)t * K + k0 + j]);
            tmem_st<8>(lane_base + k0, hi);
        }
    }
    wait_st();
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
        fence_after_sync();
        const uint32_t idesc = idesc_tf32(M, N, a_mode == 2, b_mode == 2);
        for (int ks = 0; ks < K / 8; ++ks) {
            uint64_t ad = 0, bd;
            if (a_mode == 1) ad = desc_sw128(smem_u32(sa) + (ks / 4) * M * 128 + (ks % 4) * 32, 16, 1024);
            if (a_mode == 2) ad = desc_sw128(smem_u32(sa) + ks * 1024, K * 128, 1024);
            if (b_mode == 1) bd = desc_sw128(smem_u32(sb) + (ks / 4) * N * 128 + (ks % 4) * 32, 16, 1024);
            else bd = desc_sw128(smem_u32(sb) + ks * 1024, K * 128, 1024);
            if (a_mode == 0) mma_tf32_ts(tbase + 256, tbase + ks * 8, bd, idesc, ks > 0);
            else mma_tf32_ss(tbase + 256, ad, bd, idesc, ks > 0);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after_sync();
    for (int n0 = 0; n0 < N; n0 += 8) {
        uint32_t r[8];
        tmem_ld<8>(lane_base + 256 + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) D[(size_t)t * N + n0 + j] = __uint_as_float(r[j]);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 512);
}

static float trunc_tf32(float v)
{
    uint32_t u;
    memcpy(&u, &v, 4);
    u &= 0xFFFFE000u;
    memcpy(&v, &u, 4);
    return v;
}

template <int M, int N, int K>
static int run(int a_mode, int b_mode)
{
    std::vector<float> A(M * K), B(N * K), D(128 * N);
    srand(7);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2 - 1;
    for (auto &v : B) v = (float)rand() / RAND_MAX * 2 - 1;
    float *dA, *dB, *dD;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0, D.size() * 4));
    const size_t smem = (size_t)(M + N) * K * 4 + 4096;
    CK(cudaFuncSetAttribute(probe<M, N, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe<M, N, K><<<1, 128, smem>>>(dA, dB, dD, a_mode, b_mode);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double worst = 0, peak = 0;
    for (int m = 0; m < M; ++m) {
        const int lane = (M == 64) ? (m % 16) + 32 * (m / 16) : m;
        for (int n = 0; n < N; ++n) {
            double st = 0;
            for (int k = 0; k < K; ++k) st += (double)trunc_tf32(A[m * K + k]) * trunc_tf32(B[n * K + k]);
            worst = fmax(worst, fabs(D[lane * N + n] - st));
            peak = fmax(peak, fabs(st));
        }
    }
    static const char *an[] = {"TMEM", "smem K-SW128", "smem MN-SW128"};
    static const char *bn[] = {"", "K-SW128", "MN-SW128"};
    printf("A=%-13s B=%-8s M=%3d N=%3d K=%3d: max err/peak=%.3e  D[0][0]=%g\n", an[a_mode], bn[b_mode], M, N, K,
           worst / peak, D[0]);
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
    return 0;
}

int main()
{
    // forward view of a swizzled weight image (TS mode, B K-major)
    if (run<128, 16, 32>(0, 1)) return 1;
    if (run<128, 32, 64>(0, 1)) return 1;
    if (run<128, 64, 128>(0, 1)) return 1;
    if (run<128, 128, 128>(0, 1)) return 1;
    // dgrad view of the same image (TS mode, B MN-major)
    if (run<128, 32, 16>(0, 2)) return 1;
    if (run<128, 64, 32>(0, 2)) return 1;
    if (run<128, 128, 64>(0, 2)) return 1;
    if (run<128, 128, 128>(0, 2)) return 1;
    // wgrad staging: both operands K-major SW128 with the pixel axis as K
    if (run<64, 32, 32>(1, 1)) return 1;
    if (run<128, 64, 32>(1, 1)) return 1;
    if (run<64, 32, 128>(1, 1)) return 1;
    if (run<128, 16, 128>(1, 1)) return 1;
    // A MN-major from shared memory (a TMA-loaded NCHW tile)
    if (run<128, 16, 32>(2, 1)) return 1;
    if (run<128, 32, 64>(2, 1)) return 1;
    if (run<64, 32, 32>(2, 2)) return 1;
    printf("probe_sw128 done\n");
    return 0;
}

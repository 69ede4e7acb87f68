// probe_umma.cu -- stand-alone hardware probe for bts_b200/csrc/tc.cuh (test infrastructure).
// Checks, on a real B200, the tcgen05 building blocks the chain kernels rely on:
//   T1  TS-mode  D[128xN] = A[128xK](TMEM) * B[NxK]^T (smem, K-major, no swizzle)
//   T2  SS-mode  D[MxN]   = A[MxK](smem, K-major, padded LBO) * B[NxK]^T (smem, K-major, padded LBO)
//   T3  3xTF32 accuracy vs fp64 (hi*hi + lo*hi + hi*lo)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_umma tests/cuda/probe_umma.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../bts_b200/csrc/tc.cuh"

using namespace bts::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

// ---- T1: A via tcgen05.st (hi / lo halves), B K-major no-swizzle with lbo=128, sbo=(K/4)*128
template <int N, int K, int SPLIT>
__global__ void __launch_bounds__(128) probe_ts(const float *A, const float *Bm, float *D)
{
    extern __shared__ __align__(128) unsigned char smem[];
    float *bhi = reinterpret_cast<float *>(smem);
    float *blo = bhi + N * K;
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    const int t = threadIdx.x, warp = t >> 5;
    constexpr uint32_t LBO = 128, SBO = (K / 4) * 128;
    for (int e = t; e < N * K; e += 128) {
        const int n = e / K, k = e % K;
        const uint32_t off = (n % 8) * 16 + (n / 8) * SBO + (k / 4) * LBO + (k % 4) * 4;
        const float v = Bm[e];
        bhi[off / 4] = v;
        blo[off / 4] = tf32_lo(v);
    }
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&tbase_slot, 512);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = tbase_slot;
    const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    // A hi at columns [0,K), lo at [K,2K), D at [256, 256+N)
    for (int k0 = 0; k0 < K; k0 += 8) {
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float v = A[(size_t)t * K + k0 + j];
            hi[j] = __float_as_uint(v);
            lo[j] = __float_as_uint(tf32_lo(v));
        }
        tmem_st<8>(lane_base + k0, hi);
        tmem_st<8>(lane_base + K + k0, lo);
    }
    wait_st();
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
        fence_after_sync();
        const uint32_t idesc = idesc_tf32(128, N, 0, 0);
        const uint32_t d = tbase + 256;
        uint32_t acc = 0;
        for (int pass = 0; pass < (SPLIT ? 3 : 1); ++pass) {
            // pass 0: lo*hi, pass 1: hi*lo, pass 2: hi*hi (small terms first); SPLIT=0: hi*hi only
            const uint32_t a_off = (SPLIT && pass == 0) ? K : 0;
            const float *bsrc = (SPLIT && pass == 1) ? blo : bhi;
            for (int k0 = 0; k0 < K; k0 += 8) {
                const uint64_t bd = smem_desc(smem_u32(bsrc) + (k0 / 4) * LBO, LBO, SBO);
                mma_tf32_ts(d, tbase + a_off + k0, bd, idesc, acc);
                acc = 1;
            }
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

// ---- T2: SS mode, both operands K-major with padded LBO = 144 (conflict-free pixel-major writes)
template <int M, int N, int K>
__global__ void __launch_bounds__(128) probe_ss(const float *A, const float *Bm, float *D /* [128 lanes][N] raw dump */)
{
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr uint32_t LBO = 144, SBO = (K / 4) * LBO;
    float *sa = reinterpret_cast<float *>(smem);
    float *sb = sa + (M / 8) * SBO / 4;
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    const int t = threadIdx.x, warp = t >> 5;
    for (int e = t; e < M * K; e += 128) {
        const int m = e / K, k = e % K;
        sa[((m % 8) * 16 + (m / 8) * SBO + (k / 4) * LBO + (k % 4) * 4) / 4] = A[e];
    }
    for (int e = t; e < N * K; e += 128) {
        const int n = e / K, k = e % K;
        sb[((n % 8) * 16 + (n / 8) * SBO + (k / 4) * LBO + (k % 4) * 4) / 4] = Bm[e];
    }
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&tbase_slot, 32 > N ? 32 : N);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = tbase_slot;
    // zero D first so untouched lanes are recognisable
    {
        uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int n0 = 0; n0 < N; n0 += 8) tmem_st<8>(tbase + ((uint32_t)(warp * 32) << 16) + n0, z);
        wait_st();
    }
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
        fence_after_sync();
        const uint32_t idesc = idesc_tf32(M, N, 0, 0);
        for (int k0 = 0; k0 < K; k0 += 8) {
            const uint64_t ad = smem_desc(smem_u32(sa) + (k0 / 4) * LBO, LBO, SBO);
            const uint64_t bd = smem_desc(smem_u32(sb) + (k0 / 4) * LBO, LBO, SBO);
            mma_tf32_ss(tbase, ad, bd, idesc, k0 > 0);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after_sync();
    for (int n0 = 0; n0 < N; n0 += 8) {
        uint32_t r[8];
        tmem_ld<8>(tbase + ((uint32_t)(warp * 32) << 16) + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) D[(size_t)t * N + n0 + j] = __uint_as_float(r[j]);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 32 > N ? 32 : N);
}

// ---- T4: SS mode, both operands MN-major images as the wgrad kernels build them:
//   element (ch, px) at byte (px%8)*16 + (px/8)*128 + (ch%4)*4 + (ch/4)*SBO, SBO = (P/8)*128
//   D[M=acts ch][N=delta ch] = sum_px A[ch_a][px] * B[ch_d][px]
template <int M, int N, int P>
__global__ void __launch_bounds__(128) probe_ss_mn(const float *A /*[M][P]*/, const float *Bm /*[N][P]*/, float *D, int swap)
{
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr uint32_t LBO = 128, SBO = (P / 8) * 128;
    float *sa = reinterpret_cast<float *>(smem);
    float *sb = sa + (M / 4) * SBO / 4;
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    const int t = threadIdx.x, warp = t >> 5;
    for (int e = t; e < M * P; e += 128) {
        const int c = e / P, px = e % P;
        sa[((px % 8) * 16 + (px / 8) * LBO + (c % 4) * 4 + (c / 4) * SBO) / 4] = A[e];
    }
    for (int e = t; e < N * P; e += 128) {
        const int c = e / P, px = e % P;
        sb[((px % 8) * 16 + (px / 8) * LBO + (c % 4) * 4 + (c / 4) * SBO) / 4] = Bm[e];
    }
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&tbase_slot, 32 > N ? 32 : N);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = tbase_slot;
    {
        uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int n0 = 0; n0 < N; n0 += 8) tmem_st<8>(tbase + ((uint32_t)(warp * 32) << 16) + n0, z);
        wait_st();
    }
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
        fence_after_sync();
        const uint32_t idesc = idesc_tf32(M, N, 1, 1);
        for (int k0 = 0; k0 < P; k0 += 8) {
            const uint64_t ad = swap ? smem_desc(smem_u32(sa) + (k0 / 8) * LBO, SBO, LBO) : smem_desc(smem_u32(sa) + (k0 / 8) * LBO, LBO, SBO);
            const uint64_t bd = swap ? smem_desc(smem_u32(sb) + (k0 / 8) * LBO, SBO, LBO) : smem_desc(smem_u32(sb) + (k0 / 8) * LBO, LBO, SBO);
            mma_tf32_ss(tbase, ad, bd, idesc, k0 > 0);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after_sync();
    for (int n0 = 0; n0 < N; n0 += 8) {
        uint32_t r[8];
        tmem_ld<8>(tbase + ((uint32_t)(warp * 32) << 16) + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) D[(size_t)t * N + n0 + j] = __uint_as_float(r[j]);
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 32 > N ? 32 : N);
}

// ---- T5: TS mode with B read MN-major from the FORWARD weight image (dgrad form):
//   W [NO][KI] packed K-major as for the forward (lbo=128, sbo=(KI/4)*128); D[128 x KI] = A[128 x NO] * W
template <int NO, int KI>
__global__ void __launch_bounds__(128) probe_ts_mn(const float *A /*[128][NO]*/, const float *W /*[NO][KI]*/, float *D /*[128][KI]*/)
{
    extern __shared__ __align__(128) unsigned char smem[];
    float *sw = reinterpret_cast<float *>(smem);
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_slot;
    const int t = threadIdx.x, warp = t >> 5;
    constexpr uint32_t LBO_F = 128, SBO_F = (KI / 4) * 128;
    for (int e = t; e < NO * KI; e += 128) {
        const int n = e / KI, k = e % KI;
        sw[((n % 8) * 16 + (n / 8) * SBO_F + (k / 4) * LBO_F + (k % 4) * 4) / 4] = W[e];
    }
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&tbase_slot, 512);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = tbase_slot;
    const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
    for (int k0 = 0; k0 < NO; k0 += 8) {
        uint32_t hi[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) hi[j] = __float_as_uint(A[(size_t)t * NO + k0 + j]);
        tmem_st<8>(lane_base + k0, hi);
    }
    wait_st();
    fence_before_sync();
    __syncthreads();
    if (t == 0) {
        fence_after_sync();
        const uint32_t idesc = idesc_tf32(128, KI, 0, 1);
        // MN-major view of the same bytes: mn = input channel c, k = output channel o
        //   byte = (o%8)*16 + (o/8)*SBO_F + (c%4)*4 + (c/4)*LBO_F  => lbo' = SBO_F (k groups), sbo' = LBO_F (mn groups)
        for (int k0 = 0; k0 < NO; k0 += 8) {
            const uint64_t bd = smem_desc(smem_u32(sw) + (k0 / 8) * SBO_F, SBO_F, LBO_F);
            mma_tf32_ts(tbase + 256, tbase + k0, bd, idesc, k0 > 0);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after_sync();
    for (int n0 = 0; n0 < KI; n0 += 8) {
        uint32_t r[8];
        tmem_ld<8>(lane_base + 256 + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) D[(size_t)t * KI + n0 + j] = __uint_as_float(r[j]);
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

template <int N, int K, int SPLIT>
static int run_ts(const char *name)
{
    std::vector<float> A(128 * K), B(N * K), D(128 * N);
    srand(1);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2 - 1;
    for (auto &v : B) v = (float)rand() / RAND_MAX * 2 - 1;
    float *dA, *dB, *dD;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    const size_t smem = 2 * N * K * 4;
    CK(cudaFuncSetAttribute(probe_ts<N, K, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_ts<N, K, SPLIT><<<1, 128, smem>>>(dA, dB, dD);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double err_exact = 0, err_trunc = 0, peak = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double se = 0, st = 0;
            for (int k = 0; k < K; ++k) {
                se += (double)A[m * K + k] * B[n * K + k];
                st += (double)trunc_tf32(A[m * K + k]) * trunc_tf32(B[n * K + k]);
            }
            err_exact = fmax(err_exact, fabs(D[m * N + n] - se));
            err_trunc = fmax(err_trunc, fabs(D[m * N + n] - st));
            peak = fmax(peak, fabs(se));
        }
    printf("%s N=%d K=%d split=%d: max|D-exact|/peak=%.3e  max|D-tf32(trunc)|/peak=%.3e\n", name, N, K, SPLIT,
           err_exact / peak, err_trunc / peak);
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
    return 0;
}

template <int M, int N, int K>
static int run_ss(const char *name)
{
    std::vector<float> A(M * K), B(N * K), D(128 * N);
    srand(2);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2 - 1;
    for (auto &v : B) v = (float)rand() / RAND_MAX * 2 - 1;
    float *dA, *dB, *dD;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    const size_t smem = (size_t)(M / 8 + N / 8 + 1) * (K / 4) * 144 + 256;
    CK(cudaFuncSetAttribute(probe_ss<M, N, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_ss<M, N, K><<<1, 128, smem>>>(dA, dB, dD);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    // find which TMEM lane holds each row
    int lane_of_row[128];
    double worst = 0, peak = 0;
    for (int m = 0; m < M; ++m) {
        double best = 1e30; int bl = -1;
        for (int l = 0; l < 128; ++l) {
            double e = 0;
            for (int n = 0; n < N; ++n) {
                double st = 0;
                for (int k = 0; k < K; ++k) st += (double)trunc_tf32(A[m * K + k]) * trunc_tf32(B[n * K + k]);
                e = fmax(e, fabs(D[l * N + n] - st));
                peak = fmax(peak, fabs(st));
            }
            if (e < best) { best = e; bl = l; }
        }
        lane_of_row[m] = bl;
        worst = fmax(worst, best);
    }
    printf("%s M=%d N=%d K=%d: worst row error/peak=%.3e; row->lane:", name, M, N, K, worst / peak);
    for (int m = 0; m < M; m += (M >= 64 ? 8 : 1)) printf(" %d->%d", m, lane_of_row[m]);
    printf("\n");
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
    return 0;
}

template <int M, int N, int P>
static int run_ss_mn(const char *name, int swap)
{
    std::vector<float> A(M * P), B(N * P), D(128 * N);
    srand(3);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2 - 1;
    for (auto &v : B) v = (float)rand() / RAND_MAX * 2 - 1;
    float *dA, *dB, *dD;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    const size_t smem = (size_t)(M + N) * P * 4 + 256;
    CK(cudaFuncSetAttribute(probe_ss_mn<M, N, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_ss_mn<M, N, P><<<1, 128, smem>>>(dA, dB, dD, swap);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double worst = 0, peak = 0;
    for (int m = 0; m < M; ++m) {
        const int lane = (M == 64) ? (m % 16) + 32 * (m / 16) : m;
        for (int n = 0; n < N; ++n) {
            double st = 0;
            for (int k = 0; k < P; ++k) st += (double)trunc_tf32(A[m * P + k]) * trunc_tf32(B[n * P + k]);
            worst = fmax(worst, fabs(D[lane * N + n] - st));
            peak = fmax(peak, fabs(st));
        }
    }
    printf("%s M=%d N=%d P=%d swap=%d (both MN-major): max err/peak=%.3e  D[0][0..3]=%g %g %g %g D[1][0]=%g\n", name, M, N, P, swap, worst / peak, D[0], D[1], D[2], D[3], D[N]);
    { double e[4] = {0,0,0,0}; for (int n = 0; n < 4; ++n) for (int k = 0; k < P; ++k) e[n] += (double)trunc_tf32(A[k]) * trunc_tf32(B[n * P + k]);
      double e10 = 0; for (int k = 0; k < P; ++k) e10 += (double)trunc_tf32(A[P + k]) * trunc_tf32(B[k]);
      printf("      expected %g %g %g %g | %g\n", e[0], e[1], e[2], e[3], e10); }
    cudaFree(dA); cudaFree(dB); cudaFree(dD);
    return 0;
}

template <int NO, int KI>
static int run_ts_mn(const char *name)
{
    std::vector<float> A(128 * NO), W(NO * KI), D(128 * KI);
    srand(4);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2 - 1;
    for (auto &v : W) v = (float)rand() / RAND_MAX * 2 - 1;
    float *dA, *dW, *dD;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dW, W.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dW, W.data(), W.size() * 4, cudaMemcpyHostToDevice));
    const size_t smem = (size_t)NO * KI * 4 + 256;
    CK(cudaFuncSetAttribute(probe_ts_mn<NO, KI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    probe_ts_mn<NO, KI><<<1, 128, smem>>>(dA, dW, dD);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double worst = 0, peak = 0;
    for (int m = 0; m < 128; ++m)
        for (int c = 0; c < KI; ++c) {
            double st = 0;
            for (int o = 0; o < NO; ++o) st += (double)trunc_tf32(A[m * NO + o]) * trunc_tf32(W[o * KI + c]);
            worst = fmax(worst, fabs(D[m * KI + c] - st));
            peak = fmax(peak, fabs(st));
        }
    printf("%s dgrad form NO=%d KI=%d (B MN-major view of fwd image): max err/peak=%.3e\n", name, NO, KI, worst / peak);
    cudaFree(dA); cudaFree(dW); cudaFree(dD);
    return 0;
}

int main()
{
    if (run_ss_mn<128, 64, 64>("T4 ss-mn", 0)) return 1;
    if (run_ss_mn<128, 64, 64>("T4 ss-mn", 1)) return 1;
    if (run_ss_mn<64, 32, 128>("T4 ss-mn", 0)) return 1;
    if (run_ss_mn<64, 32, 128>("T4 ss-mn", 1)) return 1;
    if (run_ss_mn<128, 16, 8>("T4 ss-mn", 0)) return 1;
    if (run_ss_mn<128, 16, 8>("T4 ss-mn", 1)) return 1;
    if (run_ts_mn<16, 32>("T5 ts-mn")) return 1;
    if (run_ts_mn<128, 128>("T5 ts-mn")) return 1;
    if (run_ts_mn<32, 64>("T5 ts-mn")) return 1;
    if (run_ts_mn<64, 128>("T5 ts-mn")) return 1;
    if (run_ts<16, 32, 0>("T1 ts")) return 1;
    if (run_ts<128, 128, 0>("T1 ts")) return 1;
    if (run_ts<64, 128, 0>("T1 ts")) return 1;
    if (run_ts<32, 64, 0>("T1 ts")) return 1;
    if (run_ts<16, 32, 1>("T3 3xtf32")) return 1;
    if (run_ts<128, 128, 1>("T3 3xtf32")) return 1;
    if (run_ss<128, 16, 128>("T2 ss")) return 1;
    if (run_ss<64, 16, 128>("T2 ss")) return 1;
    if (run_ss<64, 32, 64>("T2 ss")) return 1;
    printf("probe done\n");
    return 0;
}

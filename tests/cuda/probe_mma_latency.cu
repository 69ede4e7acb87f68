// probe_mma_latency.cu -- stand-alone hardware probe (test infrastructure): how long does a short chain of small
// tcgen05.mma kind::tf32 instructions take from issue to the commit's mbarrier arrival, as seen by the issuing warp?
// The chain kernels issue 6..48 MMAs of shape 128 x N x 8 (N = 16..128) per hand-off, all accumulating into the same
// TMEM tile; this measures whether such a chain costs per-instruction latency or per-instruction throughput.
//   mode 0: n MMAs accumulate into ONE D tile (what the kernels do)      mode 1: n MMAs into n/2 alternating D tiles
// Output: cycles (clock64) from before the first issue to after the mbarrier wait, best of 20.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_mma_latency tests/cuda/probe_mma_latency.cu
#include <cstdio>
#include <cstdlib>

#include "../../bts_b200/csrc/tc.cuh"

using namespace bts::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

template <int N, int NMMA, int MODE, int SS>
__global__ void __launch_bounds__(128) probe(long long *out)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int t = threadIdx.x, warp = t >> 5;
    for (int e = t; e < 16384; e += 128) reinterpret_cast<float *>(smem)[e] = 1.0f + (e & 7) * 0.125f;
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&slot, 512);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = (uint32_t)uniform((int)slot);
    {
        uint32_t z[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) z[j] = 0x3f800000u;
        for (int c0 = 0; c0 < 512; c0 += 32) tmem_st<32>(tbase + ((uint32_t)(warp * 32) << 16) + c0, z);
        wait_st();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    long long best = 1ll << 60;
    uint32_t phase = 0;
    if (warp == 0) {
        constexpr uint32_t idesc = idesc_tf32(128, N, 0, 0);
        const uint32_t sb = smem_u32(smem);
        for (int rep = 0; rep < 20; ++rep) {
            __syncwarp();
            const long long t0 = clock64();
            if (elect_one()) {
#pragma unroll
                for (int i = 0; i < NMMA; ++i) {
                    const uint32_t d = tbase + 256 + (MODE == 1 ? (i & 1) * 128 : 0);
                    if (SS) mma_tf32_ss(d, smem_desc(sb + 32768 + (i % 4) * 256, 128, 1024), smem_desc(sb + (i % 4) * 256, 128, 1024), idesc, 1);
                    else mma_tf32_ts(d, tbase + (i % 8) * 8, smem_desc(sb + (i % 4) * 256, 128, 1024), idesc, 1);
                }
                mma_commit(&bar);
            }
            __syncwarp();
            mbar_wait_spin(&bar, phase);
            phase ^= 1;
            const long long t1 = clock64();
            if (t1 - t0 < best) best = t1 - t0;
        }
        if (t == 0) *out = best;
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 512);
}

template <int N, int NMMA, int MODE, int SS>
static int run(const char *what)
{
    long long *d, h = 0;
    CK(cudaMalloc(&d, 8));
    CK(cudaFuncSetAttribute(probe<N, NMMA, MODE, SS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000));
    probe<N, NMMA, MODE, SS><<<1, 128, 70000>>>(d);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost));
    printf("%-10s N=%3d  %2d MMAs  %s  %6lld cycles  (%.1f per MMA)\n", SS ? "SS" : "TS", N, NMMA, what, h, (double)h / NMMA);
    CK(cudaFree(d));
    return 0;
}

int main()
{
    int rc = 0;
    rc |= run<16, 1, 0, 0>("one D ");
    rc |= run<16, 4, 0, 0>("one D ");
    rc |= run<16, 12, 0, 0>("one D ");
    rc |= run<16, 12, 1, 0>("two Ds");
    rc |= run<32, 12, 0, 0>("one D ");
    rc |= run<32, 24, 0, 0>("one D ");
    rc |= run<64, 24, 0, 0>("one D ");
    rc |= run<128, 24, 0, 0>("one D ");
    rc |= run<128, 48, 0, 0>("one D ");
    rc |= run<64, 1, 0, 1>("one D ");
    rc |= run<64, 8, 0, 1>("one D ");
    rc |= run<64, 8, 1, 1>("two Ds");
    rc |= run<128, 8, 0, 1>("one D ");
    return rc;
}

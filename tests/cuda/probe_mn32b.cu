// probe_mn32b.cu -- stand-alone hardware probe (test infrastructure): which shared-memory words does tcgen05.mma
// kind::tf32 read for an MN-major A operand under layout type 1 (SWIZZLE_128B_BASE32B)?  The vendored CUTLASS headers
// say this is the only MN-major layout for 32-bit operands; probe_sw128.cu showed that the plain SWIZZLE_128B MN-major
// descriptor returns zeros. Instead of guessing the layout, the probe READS IT OUT: every 32-bit word of the A region
// holds its own word index (exact in tf32 below 2048), B (K-major SWIZZLE_128B, known good) selects one k, so
// D[m][0] = index of the word the hardware fetched for A(m, k).  Printed as a table m x k for one K = 8 step.
//   argv[1] = layout type (default 1), argv[2] = LBO bytes (default 1024), argv[3] = SBO bytes (default 512)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_mn32b tests/cuda/probe_mn32b.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../bts_b200/csrc/tc.cuh"

using namespace bts::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __host__ inline uint32_t kmaj_off(int r, int k, int R) { return (k / 32) * R * 128 + r * 128 + ((((k % 32) / 4) ^ (r % 8)) * 16) + (k % 4) * 4; }

__global__ void __launch_bounds__(128) probe(float *D, int k0, int a_is_b, uint32_t ltype, uint32_t lbo, uint32_t sbo)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    float *sa = reinterpret_cast<float *>(smem);              // 16 KB: word w holds the value w % 2048
    unsigned char *sb = smem + 16384;                          // K-major SW128 selector, 128 rows x 32 k
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int t = threadIdx.x, warp = t >> 5;
    for (int w = t; w < 4096; w += 128) sa[w] = (float)(w % 2048);
    for (int e = t; e < 128 * 32; e += 128) {
        const int r = e / 32, k = e % 32;
        *reinterpret_cast<float *>(sb + kmaj_off(r, k, 128)) = (r == 0 && k == k0) ? 1.0f : 0.0f;
    }
    fence_smem_to_async();
    if (warp == 0) tmem_alloc(&slot, 512);
    if (t == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = (uint32_t)uniform((int)slot);
    if (warp == 0) {
        if (elect_one()) {
            const uint64_t mn = (uint64_t)((smem_u32(sa) & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) |
                                ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | ((uint64_t)ltype << 61);
            const uint64_t km = smem_desc_sw128(smem_u32(sb));      // first K = 8 step of the selector
            if (!a_is_b) {   // A = probed MN-major operand (M = 128), B = selector (N = 16)
                mma_tf32_ss(tbase, mn, km, idesc_tf32(128, 16, 1, 0), 0);
            } else {         // A = selector (M = 128, row 0 selects), B = probed MN-major operand (N = 128): D[0][n] = B(n, k0)
                mma_tf32_ss(tbase, km, mn, idesc_tf32(128, 128, 0, 1), 0);
            }
            mma_commit(&bar);
        }
        __syncwarp();
    }
    mbar_wait(&bar, 0);
    fence_after_sync();
    {
        uint32_t r[32];
        for (int c0 = 0; c0 < 128; c0 += 32) {
            tmem_ld<32>(tbase + ((uint32_t)(warp * 32) << 16) + c0, r);
            wait_ld();
            for (int j = 0; j < 32; ++j) D[(size_t)t * 128 + c0 + j] = __uint_as_float(r[j]);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 512);
}

int main(int argc, char **argv)
{
    const uint32_t ltype = argc > 1 ? (uint32_t)atoi(argv[1]) : 1u;
    const uint32_t lbo = argc > 2 ? (uint32_t)atoi(argv[2]) : 1024u;
    const uint32_t sbo = argc > 3 ? (uint32_t)atoi(argv[3]) : 512u;
    float *dD;
    CK(cudaMalloc(&dD, 128 * 128 * 4));
    std::vector<float> D(128 * 128);
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000));
    for (int a_is_b = 0; a_is_b < 2; ++a_is_b) {
        printf("# %s operand MN-major, layout type %u, LBO %u, SBO %u: word index fetched for (mn, k); rows mn = 0..127 step shown, columns k = 0..7\n",
               a_is_b ? "B" : "A", ltype, lbo, sbo);
        std::vector<std::vector<int>> tab(128, std::vector<int>(8, -1));
        for (int k0 = 0; k0 < 8; ++k0) {
            CK(cudaMemset(dD, 0, 128 * 128 * 4));
            probe<<<1, 128, 40000>>>(dD, k0, a_is_b, ltype, lbo, sbo);
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
            for (int mn = 0; mn < 128; ++mn) tab[mn][k0] = (int)(a_is_b ? D[mn] : D[(size_t)mn * 128]);
        }
        for (int mn = 0; mn < 128; ++mn) {
            printf("mn %3d:", mn);
            for (int k = 0; k < 8; ++k) printf(" %5d", tab[mn][k]);
            printf("\n");
        }
    }
    return 0;
}

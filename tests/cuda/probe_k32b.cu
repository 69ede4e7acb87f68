// probe_k32b.cu -- stand-alone hardware probe (test infrastructure), two questions about ONE shared-memory image
// serving several roles (rows of 128 bytes = 32 fp32 pixels of one channel, 32-byte chunks XOR-ed with row % 4):
//   (1) does tcgen05.mma kind::tf32 read a K-MAJOR operand under layout type 1 (SWIZZLE_128B_BASE32B), and which words?
//       (probe_mn32b.cu read out the MN-major layout of the same type; if the K-major one addresses the same bytes,
//       a [channel][pixel] tile is at once the MN-major operand of the chain MMAs and the K-major operand of the
//       weight-gradient MMAs, and a raw [out][in] weight matrix serves forward (K-major) and dgrad (MN-major) alike)
//   (2) where does TMA put element (row, col) of a {32 floats, R rows} box under CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B?
// Method as in probe_mn32b.cu: every word of the probed region holds its own index, a known-good K-major SWIZZLE_128B
// selector picks one k, D[m][0] = index of the word fetched for (m, k).
//   argv: ltype (1) lbo (1024) sbo (1024) start_off_bytes (0) as_b (0)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o probe_k32b tests/cuda/probe_k32b.cu -lcuda
#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../bts_b200/csrc/tc.cuh"

using namespace bts::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __host__ inline uint32_t kmaj_off(int r, int k, int R) { return (k / 32) * R * 128 + r * 128 + ((((k % 32) / 4) ^ (r % 8)) * 16) + (k % 4) * 4; }

__global__ void __launch_bounds__(128) probe(float *D, int k0, int as_b, uint32_t ltype, uint32_t lbo, uint32_t sbo, uint32_t off)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    float *sa = reinterpret_cast<float *>(smem);              // 32 KB: word w holds the value w % 2048
    unsigned char *sb = smem + 32768;                          // K-major SW128 selector, 128 rows x 32 k
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int t = threadIdx.x, warp = t >> 5;
    for (int w = t; w < 8192; w += 128) sa[w] = (float)(w % 2048);
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
            const uint64_t pr = (uint64_t)(((smem_u32(sa) + off) & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) |
                                ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | ((uint64_t)ltype << 61);
            const uint64_t km = smem_desc_sw128(smem_u32(sb));
            if (!as_b) mma_tf32_ss(tbase, pr, km, idesc_tf32(128, 16, 0, 0), 0);     // A probed (K-major), B selector
            else mma_tf32_ss(tbase, km, pr, idesc_tf32(128, 128, 0, 0), 0);          // A selector, B probed (K-major)
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

// ---- TMA landing pattern: box {32 floats, rows} of a [rows][64] fp32 tensor whose element (r, c) holds r * 64 + c
__global__ void __launch_bounds__(32) tma_probe(const __grid_constant__ CUtensorMap map, float *out, int rows, int c0)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        mbar_arrive_expect_tx(&bar, (uint32_t)rows * 128u);
        tma_load_2d(smem_u32(smem), &map, c0, 0, &bar);
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    for (int w = threadIdx.x; w < rows * 32; w += 32) out[w] = reinterpret_cast<float *>(smem)[w];
}

int main(int argc, char **argv)
{
    const uint32_t ltype = argc > 1 ? (uint32_t)atoi(argv[1]) : 1u;
    const uint32_t lbo = argc > 2 ? (uint32_t)atoi(argv[2]) : 1024u;
    const uint32_t sbo = argc > 3 ? (uint32_t)atoi(argv[3]) : 1024u;
    const uint32_t off = argc > 4 ? (uint32_t)atoi(argv[4]) : 0u;
    const int as_b = argc > 5 ? atoi(argv[5]) : 0;
    float *dD;
    CK(cudaMalloc(&dD, 128 * 128 * 4));
    std::vector<float> D(128 * 128);
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000));
    printf("# %s operand K-major, layout type %u, LBO %u, SBO %u, start offset %u B: word index fetched for (row, k), k = 0..7\n",
           as_b ? "B" : "A", ltype, lbo, sbo, off);
    std::vector<std::vector<int>> tab(128, std::vector<int>(8, -1));
    for (int k0 = 0; k0 < 8; ++k0) {
        CK(cudaMemset(dD, 0, 128 * 128 * 4));
        probe<<<1, 128, 60000>>>(dD, k0, as_b, ltype, lbo, sbo, off);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
        for (int m = 0; m < 128; ++m) tab[m][k0] = (int)(as_b ? D[m] : D[(size_t)m * 128]);
    }
    const int show[] = {0, 1, 2, 3, 4, 5, 7, 8, 9, 12, 15, 16, 31, 32, 64, 127};
    for (int m : show) {
        printf("row %3d:", m);
        for (int k = 0; k < 8; ++k) printf(" %5d", tab[m][k]);
        printf("\n");
    }
    if (argc > 6) return 0;

    // ---- TMA
    void *fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
        printf("cuTensorMapEncodeTiled unavailable\n");
        return 0;
    }
    auto enc = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(fp);
    const int rows = 16, cols = 64;
    std::vector<float> h(rows * cols);
    for (int i = 0; i < rows * cols; ++i) h[i] = (float)i;
    float *dsrc, *dout;
    CK(cudaMalloc(&dsrc, h.size() * 4));
    CK(cudaMalloc(&dout, rows * 32 * 4));
    CK(cudaMemcpy(dsrc, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    for (int mode = 0; mode < 2; ++mode) {
        CUtensorMap m;
        const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
        const cuuint64_t str[1] = {(cuuint64_t)cols * 4};
        const cuuint32_t box[2] = {32, (cuuint32_t)rows};
        const cuuint32_t es[2] = {1, 1};
        const CUtensorMapSwizzle sw = mode == 0 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
        const CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dsrc, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("# TMA box {32, %d} swizzle %s: encode result %d; smem word -> source (row, col)\n", rows,
               mode == 0 ? "128B_ATOM_32B" : "128B", (int)r);
        if (r != CUDA_SUCCESS) continue;
        CK(cudaFuncSetAttribute(tma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192));
        tma_probe<<<1, 32, 8192>>>(m, dout, rows, 0);
        CK(cudaDeviceSynchronize());
        std::vector<float> o(rows * 32);
        CK(cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost));
        for (int rr = 0; rr < 9; ++rr) {
            printf("smem row %2d (bytes %4d..): chunk(32B) sources:", rr, rr * 128);
            for (int c = 0; c < 32; c += 4) printf(" (%d,%2d)", (int)o[rr * 32 + c] / cols, (int)o[rr * 32 + c] % cols);
            printf("\n");
        }
    }
    return 0;
}

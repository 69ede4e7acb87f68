// chain_wgrad_tma.cu -- weight gradients of the WIDE chains (first layer > 64 channels) on sm_100a.
//
// dW_l = sum over pixels of delta_{l+1} (x) a_l  (SURVEY.md 8(a) row a8). The multi-kernel backward
// (chain_tc_bwd.cu) leaves a_l and delta_l as [channel][pixel] fp32 rows in an L2-sized workspace; x is
// NCHW. With the PIXEL axis as the MMA K dimension those rows are K-major operands as they lie, so the
// TMA engine drops [8 channels][32 pixels] boxes straight into SWIZZLE_128B shared-memory images
// (cp.async.bulk.tensor, zero-filled out of bounds) -- no register staging, no transposition:
//   warp 0      producer : per (32-pixel chunk, job) one transaction barrier + nops box loads
//   warps 2..5  splitters: remainder images lo = v - tf32(v) into one of two remainder buffers (element-wise, 16 B)
//   warp 1      issuer   : per K-step of 8 pixels  D[:, :nb] += A_hi * B_hi^T + A_lo * B_hi^T,  D[:, nb:] += A_hi * B_lo^T
//                          (tcgen05.mma kind::tf32, both operands from shared memory, fp32 accumulators
//                          resident in TMEM for the whole kernel), tcgen05.commit frees the landing slot and
//                          the remainder buffer
// A job is one wide layer (A = a_l rows, B = delta_{l+1} rows) or the concatenated narrow layers
// (block-diagonal). Per-CTA partial sums go to the workspace and are reduced deterministically.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdlib>
#include <cstring>

#include "chain_tc.cuh"

namespace bts {

constexpr int kWtMaxSlots = 8;
constexpr int kWtRingBytes = 225 * 1024;   // all of the 227 KB a CTA may have, less barriers and alignment slack
// Shared memory: a ring of LANDING slots the TMA producer fills (A_hi 16K | B_hi nb rows, <= 16K: raw fp32 = the hi
// operand) and TWO remainder buffers of the same shape (A_lo | B_lo) the splitter warps write. The bytes in flight
// per SM are what the producer runs ahead of the splitters and the MMAs, about (landing slots - 2) slots. The first
// version kept hi and lo images of a job in one 64 KB slot, three slots: ~32 KB in flight per SM = 4.7 MB on the
// chip against the ~10 MB the HBM needs, and the kernel sat at 4 TB/s. The slot is sized to the widest job of the
// plan: 5 landing slots for the 8x8 chain, 7 for the 4x4 chain.
constexpr int kWtMaxOps = 3;                 // box loads per job: [x] [activation rows] [delta rows]
constexpr int kWtMaxMaps = 6;                // distinct (tensor, box height) pairs
constexpr int kWtThreads = 320;              // producer, issuer, 4 splitter warps, 4 drainer warps
// The tensor core accumulates into TMEM with round-toward-zero (see kWgradDrain in chain_bwd_fused.cu): an accumulator
// loses ~3.4e-8 of its magnitude per accumulate step, always toward zero. A CTA therefore works in SEGMENTS of
// kWtSegChunks 32-pixel chunks (8 steps each into the widest accumulator block; measured at KITTI b16: grad_W error
// <= 4.7e-6 / 6.7e-6 / 8.6e-6 of peak for 16 / 32 / 64 chunks, where one segment per kernel gave 2.5e-5): at a
// segment's end the splitter warps add the accumulators to this CTA's partial sums in the workspace (every element has
// one owner thread: plain read-modify-write, L2-resident), and the next segment's first MMAs overwrite the
// accumulators (accumulate flag 0) instead of adding to them.
#ifndef BTS_WT_SEG_CHUNKS
#define BTS_WT_SEG_CHUNKS 16
#endif
constexpr int kWtSegChunks = BTS_WT_SEG_CHUNKS;

struct WtOp {
    unsigned short src_row;                  // first source row (channel) of the box
    unsigned short dst_row;                  // first row in the A image, or 128 + first row in the B image
    unsigned short rows;                     // box height
    unsigned char map;                       // index into WtMaps
    unsigned char is_x;                      // 3-D map over the NCHW input
};
struct WtJob {
    WtOp ops[kWtMaxOps];
    int nops, a_rows, nb, tmem_col, tx_bytes;
};
struct WtMaps {
    CUtensorMap m[kWtMaxMaps];
};
struct WtExtract {                           // layer l: D_job[row0 + k][col0 + n] -> partial[dst_off + n*K + k]
    int job, row0, col0, K, N, dst_off;
};
struct WtPlan {
    WtJob jobs[3];
    WtExtract ex[BTS_MAX_LAYERS];
    int njobs, nlayers, total_w, hw_pad, nmaps, nslots, slot_bytes, seg_chunks;
    int set_cols;                            // > 0: two accumulator sets this many TMEM columns apart, alternating per segment
    long long nchunks;                       // 32-pixel chunks of the padded pixel axis
    float *partials;                         // [gridDim.x][total_w]
};

__global__ void __launch_bounds__(kWtThreads, 1)
chain_wgrad_tma_kernel(const __grid_constant__ WtMaps maps, const __grid_constant__ WtPlan pl)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char *slots = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int nslots = pl.nslots;
    const uint32_t slot_bytes = (uint32_t)pl.slot_bytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(slots + kWtRingBytes);
    uint64_t *full = bars, *split = bars + kWtMaxSlots, *empty = bars + 2 * kWtMaxSlots, *fin = bars + 3 * kWtMaxSlots;
    uint64_t *lofree = fin + 2;                                   // [2]: the MMAs that read a remainder buffer are done
    uint64_t *drained = lofree + 2;                               // [2]: the splitters have read an accumulator set
    uint32_t *slot_tmem = reinterpret_cast<uint32_t *>(drained + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < nslots; ++s) {
            mbar_init(full + s, 1);
            mbar_init(split + s, 128);
            mbar_init(empty + s, 1);
        }
        mbar_init(fin, 1);
        mbar_init(fin + 1, 1);
        mbar_init(lofree, 1);
        mbar_init(lofree + 1, 1);
        mbar_init(drained, 128);
        mbar_init(drained + 1, 128);
        fence_mbar_init();
        for (int i = 0; i < pl.nmaps; ++i) tma_prefetch_desc(&maps.m[i]);
    }
    if (warp == 0) tmem_alloc(slot_tmem, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = (uint32_t)uniform((int)*slot_tmem);
    const uint32_t lane_addr = tbase + ((uint32_t)((warp & 3) * 32) << 16);   // a warp reaches TMEM lane quarter warp % 4
    // (no zero fill of the accumulators: the first MMAs of every segment run with the accumulate flag off)
    pdl_trigger();
    pdl_wait();                                   // the delta rows come from the dgrad kernel launched just before
    const uint32_t slot0 = smem_u32(slots);
    // balanced segments: a CTA with 33 chunks and a limit of 32 runs 17 + 16, not 32 + 1 (a drain on the kernel's tail)
    const int my_chunks = (int)((pl.nchunks - blockIdx.x + gridDim.x - 1) / gridDim.x);
    const int my_segs = (my_chunks + pl.seg_chunks - 1) / pl.seg_chunks;
    const int seg_len = my_segs > 0 ? (my_chunks + my_segs - 1) / my_segs : 1;
    if (warp == 0) {
        if (lane == 0) {                                          // ---- producer
            int s = 0;
            uint32_t n = 0;                                       // ring position: slot s, n-th pass over the ring
            for (long long ck = blockIdx.x; ck < pl.nchunks; ck += gridDim.x) {
                // chunks are visited LAST to FIRST: the dgrad kernel has just written the delta rows (and read the
                // activation rows) in ascending pixel order, so the highest pixels are what the L2 still holds
                const long long gp = (pl.nchunks - 1 - ck) * 32;
                const int b = (int)(gp / pl.hw_pad), p0 = (int)(gp - (long long)b * pl.hw_pad);
                for (int j = 0; j < pl.njobs; ++j) {
                    if (n > 0) mbar_wait(empty + s, (n - 1) & 1);
                    const WtJob &job = pl.jobs[j];
                    mbar_arrive_expect_tx(full + s, (uint32_t)job.tx_bytes);
                    const uint32_t base = slot0 + s * slot_bytes;
                    for (int o = 0; o < job.nops; ++o) {
                        const WtOp op = job.ops[o];
                        const uint32_t dst = base + (op.dst_row < 128 ? op.dst_row * 128u : 16384u + (op.dst_row - 128) * 128u);
                        if (op.is_x) tma_load_3d(dst, &maps.m[op.map], p0, op.src_row, b, full + s);
                        else tma_load_2d(dst, &maps.m[op.map], (int)gp, op.src_row, full + s);
                    }
                    if (++s == nslots) { s = 0; ++n; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        {                                                         // ---- MMA issuer (whole warp waits, one elected lane issues)
            int s = 0, ci = 0;
            uint32_t n = 0, it = 0;
            // Segment g accumulates into set g % nsets (nsets = 2 when two sets fit the 512 TMEM columns, else 1). Before
            // it overwrites a set the issuer waits until the splitters have read the segment that used it last.
            const int nsets = pl.set_cols > 0 ? 2 : 1;
            int seg = 0;
            for (long long ck = blockIdx.x; ck < pl.nchunks; ck += gridDim.x) {
                const uint32_t keep = (ci % seg_len) != 0;   // 0: first chunk of a segment overwrites the accumulators
                const int set = seg % nsets;
                if (!keep && seg >= nsets) mbar_wait(drained + set, (uint32_t)(seg / nsets - 1) & 1);
                const uint32_t tset = tbase + (uint32_t)(set * pl.set_cols);
                for (int j = 0; j < pl.njobs; ++j, ++it) {
                    mbar_wait(split + s, n & 1);
                    fence_after_sync();
                    if (elect_one()) {
                        const WtJob &job = pl.jobs[j];
                        const uint32_t base = slot0 + s * slot_bytes, lob = slot0 + (nslots + (it & 1)) * slot_bytes;
                        const uint32_t a_hi = base, b_hi = base + 16384u, a_lo = lob, b_lo = lob + 16384u;
                        const uint32_t d = tset + job.tmem_col;
                        const uint32_t id = idesc_tf32(128, job.nb, 0, 0);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {          // D[:, 0:nb] = A_hi B_hi + A_lo B_hi, D[:, nb:2nb] = A_hi B_lo
                            const uint32_t acc = ks == 0 ? keep : 1u;
                            mma_tf32_ss(d, smem_desc_sw128(a_hi + ks * 32), smem_desc_sw128(b_hi + ks * 32), id, acc);
                            mma_tf32_ss(d + job.nb, smem_desc_sw128(a_hi + ks * 32), smem_desc_sw128(b_lo + ks * 32), id, acc);
                            mma_tf32_ss(d, smem_desc_sw128(a_lo + ks * 32), smem_desc_sw128(b_hi + ks * 32), id, 1);
                        }
                        mma_commit(empty + s);
                        mma_commit(lofree + (it & 1));
                    }
                    __syncwarp();
                    if (++s == nslots) { s = 0; ++n; }
                }
                ++ci;
                const bool last = ck + gridDim.x >= pl.nchunks;
                if (last || ci % seg_len == 0) {                  // segment end: hand the set to the splitters
                    if (elect_one()) mma_commit(fin + set);
                    __syncwarp();
                    ++seg;
                }
            }
        }
        __syncwarp();
    } else if (warp >= 6) {                                       // ---- drainers (warps 6..9 cover the four TMEM lane quarters)
        // They sleep on the segment barriers, so the splitters keep splitting and the MMAs run on into the other
        // accumulator set while a finished segment is added to the partial sums.
        float *part = pl.partials + (size_t)blockIdx.x * pl.total_w;
        const int m = (warp & 3) * 32 + lane;
        // accumulators -> this CTA's partial sums (row m of a job = TMEM lane m); segments after the first add
        const int nsets = pl.set_cols > 0 ? 2 : 1;
        auto drain = [&](bool add, uint32_t lane_addr) {
            for (int l = 0; l < pl.nlayers; ++l) {
                const WtExtract ex = pl.ex[l];
                const WtJob &job = pl.jobs[ex.job];
                const int k = m - ex.row0;
                const bool mine = k >= 0 && k < ex.K;
                auto put = [&](int n, float v) {
                    float *dst = part + ex.dst_off + (size_t)n * ex.K + k;
                    // later segments add with a fire-and-forget reduction (no L2 round trip on the critical path).
                    // One owner thread per element and in-order segments: the sum's order is fixed.
                    if (add) asm volatile("red.global.add.f32 [%0], %1;" ::"l"(dst), "f"(v) : "memory");
                    else *dst = v;
                };
                int n0 = 0;
                for (; n0 + 32 <= ex.N; n0 += 32) {               // 32 columns per TMEM round trip
                    uint32_t r0[32], r1[32];
                    tmem_ld<32>(lane_addr + job.tmem_col + ex.col0 + n0, r0);
                    tmem_ld<32>(lane_addr + job.tmem_col + job.nb + ex.col0 + n0, r1);
                    wait_ld();
                    if (mine) {
#pragma unroll
                        for (int jn = 0; jn < 32; ++jn) put(n0 + jn, __uint_as_float(r0[jn]) + __uint_as_float(r1[jn]));
                    }
                }
                for (; n0 < ex.N; n0 += 8) {
                    uint32_t r0[8], r1[8];
                    tmem_ld<8>(lane_addr + job.tmem_col + ex.col0 + n0, r0);
                    tmem_ld<8>(lane_addr + job.tmem_col + job.nb + ex.col0 + n0, r1);
                    wait_ld();
                    if (mine) {
#pragma unroll
                        for (int jn = 0; jn < 8; ++jn)
                            if (n0 + jn < ex.N) put(n0 + jn, __uint_as_float(r0[jn]) + __uint_as_float(r1[jn]));
                    }
                }
            }
        };
        for (int seg = 0; seg < my_segs; ++seg) {
            const int set = seg % nsets;
            mbar_wait(fin + set, (uint32_t)(seg / nsets) & 1);
            fence_after_sync();
            drain(seg > 0, lane_addr + (uint32_t)(set * pl.set_cols));
            fence_before_sync();
            mbar_arrive(drained + set);                           // the issuer may overwrite this set now
        }
    } else {                                                      // ---- splitters (warps 2..5)
        const int st = tid - 64;
        int s = 0;
        uint32_t n = 0, it = 0;
        for (long long ck = blockIdx.x; ck < pl.nchunks; ck += gridDim.x) {
            for (int j = 0; j < pl.njobs; ++j, ++it) {
                mbar_wait(full + s, n & 1);
                if (it >= 2) mbar_wait(lofree + (it & 1), ((it >> 1) - 1) & 1);   // the job two back has read this buffer
                const WtJob &job = pl.jobs[j];
                float4 *a_hi = reinterpret_cast<float4 *>(slots + (size_t)s * slot_bytes);
                float4 *a_lo = reinterpret_cast<float4 *>(slots + (size_t)(nslots + (it & 1)) * slot_bytes);
                float4 *b_hi = a_hi + 1024, *b_lo = a_lo + 1024;
                // remainder images, four independent 16-byte units in flight per thread
                auto split_image = [&](const float4 *hi, float4 *lo, int units) {
#pragma unroll 1
                    for (int i = st; i < units; i += 512) {
                        float4 v[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (i + u * 128 < units) v[u] = hi[i + u * 128];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (i + u * 128 < units) {
                                float4 r;
                                tf32_lo2(v[u].x, v[u].y, r.x, r.y);
                                tf32_lo2(v[u].z, v[u].w, r.z, r.w);
                                lo[i + u * 128] = r;
                            }
                        }
                    }
                };
                split_image(a_hi, a_lo, job.a_rows * 8);
                split_image(b_hi, b_lo, job.nb * 8);
                fence_smem_to_async();
                mbar_arrive(split + s);
                if (++s == nslots) { s = 0; ++n; }
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tbase, 512);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
int launch_wreduce(const ChainParams &cp, const float *partials, int nparts, int total_w, cudaStream_t s);

static PFN_cuTensorMapEncodeTiled encode_fn()
{
    static PFN_cuTensorMapEncodeTiled fn = nullptr;     // resolved once; the pointer is immutable afterwards
    if (fn == nullptr) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    }
    return fn;
}

// fp32 tensor, innermost dimension = pixels, boxes of [box_rows][32 pixels], 128-byte swizzle, zero fill
static int make_map(CUtensorMap *m, const void *base, int rank, const cuuint64_t *dims, const cuuint64_t *strides_bytes,
                    int box_rows)
{
    PFN_cuTensorMapEncodeTiled fn = encode_fn();
    if (fn == nullptr) return set_error(BTS_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available");
    const cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void *>(base), dims, strides_bytes, box,
                          estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(BTS_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return BTS_OK;
}

// usable when every layer is at most 128 wide and the NCHW plane stride is a multiple of 16 bytes
bool wgrad_tma_supported(const ChainParams &cp, int hw, const void *x)
{
    if (hw % 4 != 0 || (reinterpret_cast<uintptr_t>(x) & 15u) != 0) return false;
    for (int l = 0; l <= cp.n_layers; ++l)
        if (cp.dims[l] > 128) return false;
    return encode_fn() != nullptr;
}

// act = [ra][npp] rows a_1 .. a_{L-1}; del = [rd][npp] rows delta_1 .. delta_L (chain order); the pixel
// axis is padded per image to hw_pad (a multiple of 32), npp = padded total (a multiple of 128).
// x_rows (bf16 twin): the features as fp32 rows [dims[0]][npp] in the stash, read through a 2-D map like the other rows
int launch_wgrad_tma(const ChainParams &cp, const float *x, const float *act, const float *del, float *partials,
                     int B, int hw, int hw_pad, long long npp, int sms, cudaStream_t s, const float *x_rows)
{
    const int L = cp.n_layers;
    WtPlan pl;
    memset(&pl, 0, sizeof(pl));
    int prefix[BTS_MAX_LAYERS + 2];                     // prefix[l] = sum_{j=1}^{l-1} dims[j]: first row of a_l / delta_l
    prefix[0] = prefix[1] = 0;
    for (int l = 2; l <= L + 1; ++l) prefix[l] = prefix[l - 1] + cp.dims[l - 1];
    const int ra = prefix[L], rd = prefix[L + 1];
    int woffs[BTS_MAX_LAYERS], woff = 0;
    for (int l = 0; l < L; ++l) { woffs[l] = woff; woff += cp.dims[l] * cp.dims[l + 1]; }
    // trailing layers are concatenated into one block-diagonal job while their inputs fit 128 rows
    int fc = L, sum = 0;
    while (fc > 0 && sum + cp.dims[fc - 1] <= 128) sum += cp.dims[--fc];
    if (fc == L) return set_error(BTS_ERR_UNSUPPORTED, "layer wider than 128 channels");
    pl.njobs = fc + 1;
    if (pl.njobs > 3) return set_error(BTS_ERR_UNSUPPORTED, "too many weight-gradient jobs");
    // tensor maps are created on demand, one per (tensor, box height)
    WtMaps maps;
    memset(&maps, 0, sizeof(maps));
    int map_src[kWtMaxMaps], map_rows[kWtMaxMaps];
    auto get_map = [&](int src, int rows) -> int {       // src 0: x, 1: activation rows, 2: delta rows
        for (int i = 0; i < pl.nmaps; ++i)
            if (map_src[i] == src && map_rows[i] == rows) return i;
        if (pl.nmaps == kWtMaxMaps) return -1;
        const int i = pl.nmaps;
        int e;
        if (src == 0 && x_rows != nullptr) {
            const cuuint64_t dims[2] = {(cuuint64_t)npp, (cuuint64_t)cp.dims[0]};
            const cuuint64_t str[1] = {(cuuint64_t)npp * 4};
            e = make_map(&maps.m[i], x_rows, 2, dims, str, rows);
        } else if (src == 0) {
            const cuuint64_t dims[3] = {(cuuint64_t)hw, (cuuint64_t)cp.dims[0], (cuuint64_t)B};
            const cuuint64_t str[2] = {(cuuint64_t)hw * 4, (cuuint64_t)hw * 4 * cp.dims[0]};
            e = make_map(&maps.m[i], x, 3, dims, str, rows);
        } else {
            const cuuint64_t dims[2] = {(cuuint64_t)npp, (cuuint64_t)(src == 1 ? ra : rd)};
            const cuuint64_t str[1] = {(cuuint64_t)npp * 4};
            e = make_map(&maps.m[i], src == 1 ? act : del, 2, dims, str, rows);
        }
        if (e) return -1;
        map_src[i] = src; map_rows[i] = rows;
        return pl.nmaps++;
    };
    int col = 0;
    for (int j = 0; j < pl.njobs; ++j) {
        WtJob &job = pl.jobs[j];
        const int l0 = j < fc ? j : fc, l1 = j < fc ? j + 1 : L;
        int arow = 0, brow = 0;
        for (int l = l0; l < l1; ++l) {
            const int K = cp.dims[l], N = cp.dims[l + 1];
            // the rows of consecutive layers follow each other in the workspace exactly as in the images
            if (arow % 8 || brow % 8 || K % 8) return set_error(BTS_ERR_UNSUPPORTED, "layer width not a multiple of 8");
            WtExtract &ex = pl.ex[l];
            ex.job = j; ex.row0 = arow; ex.col0 = brow; ex.K = K; ex.N = N; ex.dst_off = woffs[l];
            arow += K;
            brow += N;
        }
        if (arow > 128 || brow > 128) return set_error(BTS_ERR_UNSUPPORTED, "job exceeds 128 rows");
        job.a_rows = arow;
        job.nb = (brow + 15) / 16 * 16;
        job.tmem_col = col;
        col += 2 * job.nb;
        // box loads: [x] [a_l0' .. a_(l1-1)] [delta_(l0+1) .. delta_l1]; rows past a tensor's end are zero-filled
        int act_first = l0, act_dst = 0;
        if (l0 == 0) {
            const int m = get_map(0, cp.dims[0]);
            if (m < 0) return set_error(BTS_ERR_UNSUPPORTED, "tensor map for x");
            job.ops[job.nops++] = WtOp{0, 0, (unsigned short)cp.dims[0], (unsigned char)m, (unsigned char)(x_rows == nullptr ? 1 : 0)};
            act_first = 1;
            act_dst = cp.dims[0];
        }
        if (act_first < l1) {
            const int rows = arow - act_dst;
            const int m = get_map(1, rows);
            if (m < 0) return set_error(BTS_ERR_UNSUPPORTED, "tensor map for the activation rows");
            job.ops[job.nops++] = WtOp{(unsigned short)prefix[act_first], (unsigned short)act_dst, (unsigned short)rows, (unsigned char)m, 0};
        }
        {
            const int m = get_map(2, job.nb);
            if (m < 0) return set_error(BTS_ERR_UNSUPPORTED, "tensor map for the delta rows");
            job.ops[job.nops++] = WtOp{(unsigned short)prefix[l0 + 1], 128, (unsigned short)job.nb, (unsigned char)m, 0};
        }
        job.tx_bytes = 0;
        for (int o = 0; o < job.nops; ++o) job.tx_bytes += job.ops[o].rows * 128;
    }
    if (col > 512) return set_error(BTS_ERR_UNSUPPORTED, "weight-gradient accumulators exceed TMEM");
    pl.set_cols = col <= 256 ? 256 : 0;
    pl.nlayers = L;
    pl.total_w = woff;
    pl.hw_pad = hw_pad;
    pl.nchunks = npp / 32;
    pl.partials = partials;
    static const int seg_env = getenv("BTS_WT_SEG_CHUNKS") ? atoi(getenv("BTS_WT_SEG_CHUNKS")) : 0;   // tuning hook
    pl.seg_chunks = seg_env > 0 ? seg_env : kWtSegChunks;

    int max_nb = 8;
    for (int j = 0; j < pl.njobs; ++j) max_nb = pl.jobs[j].nb > max_nb ? pl.jobs[j].nb : max_nb;
    pl.slot_bytes = 16384 + max_nb * 128;
    pl.nslots = kWtRingBytes / pl.slot_bytes - 2;                 // two slots' worth of room are the remainder buffers
    if (pl.nslots > kWtMaxSlots) pl.nslots = kWtMaxSlots;
    const size_t smem = 1024 + (size_t)kWtRingBytes + 8 * (3 * kWtMaxSlots + 6) + 16;
    cudaError_t e = cudaFuncSetAttribute(chain_wgrad_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    long long grid = pl.nchunks < sms ? pl.nchunks : sms;
    // always follows the dgrad (or K1) kernel of the same call
    launch_dependent(chain_wgrad_tma_kernel, dim3((unsigned)grid), dim3(kWtThreads), smem, s, maps, pl);
    if (int e2 = check_launch("chain_wgrad_tma")) return e2;
    return launch_wreduce(cp, partials, (int)grid, pl.total_w, s);
}

}  // namespace bts

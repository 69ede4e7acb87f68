// chain_tc.cuh -- shared pieces of the tensor-core chain kernels (chain_tc_fwd.cu, chain_tc_bwd.cu).
//
// A reduction_1x1 chain  D0 -> D1 -> ... -> 16 -> 8 -> (3 | 1)  is split in two parts:
//   * "wide" layers (every layer whose output width is >= 16) run on the 5th-gen tensor cores as
//     3xTF32 tcgen05.mma: M = 128 low-res pixels (one per TMEM lane / per thread of a warpgroup),
//     N = output channels, K = input channels. Activations live in TMEM (A operand, hi and lo
//     halves), weights in shared memory (B operand, K-major, no swizzle, hi and lo halves).
//   * the narrow tail 16 -> 8 -> (3|1), the ELUs, the plane parameterisation and the LPG expansion
//     run as SIMT code, one thread per low-res pixel, on registers.
// 3xTF32: x = hi + lo with hi = the fp32 word as the tensor core truncates it; D = lo*Whi + hi*Wlo
// + hi*Whi accumulated in fp32 (error ~5e-7 of peak, measured by tests/cuda/probe_umma.cu).
#pragma once
#include "common.cuh"
#include "tc.cuh"

namespace bts {

constexpr int kTile = 128;              // low-res pixels per tile = TMEM lanes = threads per warpgroup
constexpr int kTailIn = 16;             // width at which the SIMT tail takes over
constexpr int kTailMid = 8;

// Widths of the tensor-core part: D0 -> D1 [-> D2 [-> D3 [-> D4]]], the last one being 16.
template <int NW_, int D0_, int D1_, int D2_ = 0, int D3_ = 0, int D4_ = 0>
struct TcCfg {
    static constexpr int NW = NW_;
    __host__ __device__ static constexpr int dim(int l)
    {
        return l == 0 ? D0_ : l == 1 ? D1_ : l == 2 ? D2_ : l == 3 ? D3_ : D4_;
    }
    __host__ __device__ static constexpr int K(int l) { return dim(l); }
    __host__ __device__ static constexpr int N(int l) { return dim(l + 1); }
    __host__ __device__ static constexpr int kmax()
    {
        int m = 0;
        for (int l = 0; l < NW; ++l) m = K(l) > m ? K(l) : m;
        return m;
    }
    __host__ __device__ static constexpr int nmax()
    {
        int m = 0;
        for (int l = 0; l < NW; ++l) m = N(l) > m ? N(l) : m;
        return m;
    }
    // packed weight image (floats): per wide layer [hi image | lo image], then the raw tail weights
    __host__ __device__ static constexpr int w_off(int l)
    {
        int o = 0;
        for (int j = 0; j < l; ++j) o += 2 * N(j) * K(j);
        return o;
    }
    __host__ __device__ static constexpr int tail_off() { return w_off(NW); }
    __host__ __device__ static constexpr int pack_floats(int last) { return tail_off() + kTailMid * kTailIn + last * kTailMid; }
    static_assert(NW_ >= 1 && NW_ <= 4, "1..4 tensor-core layers");
};

// canonical K-major, no-swizzle operand: element (row, k) of an [rows][K] fp32 matrix
__host__ __device__ constexpr uint32_t kmajor_off_floats(int row, int k, int K)
{
    return (uint32_t)((row % 8) * 4 + (row / 8) * (K / 4) * 32 + (k / 4) * 32 + (k % 4));
}

// ELU with the hardware exponential (ex2.approx): absolute error <= 2e-7 on the negative branch.
__device__ __forceinline__ float elu_fast(float t)
{
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t * 1.4426950408889634f));
    return t > 0.f ? t : e - 1.f;
}

// ---- bf16 storage (the optional bf16 twin of the path: big tensors in bf16, arithmetic unchanged). A bf16 value is
// the top 16 bits of the fp32 word, hence exactly representable in tf32: the 3xTF32 remainder of a bf16 input is zero.
__device__ __forceinline__ float bf16_to_f32(unsigned short v) { return __uint_as_float((uint32_t)v << 16); }
__device__ __forceinline__ unsigned short f32_to_bf16(float v)        // round to nearest even (one hardware instruction)
{
    unsigned short r;
    asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi)   // {hi : lo}, both rounded to nearest even
{
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
// element i of a tensor stored as T (float or bf16 = unsigned short)
template <class T> __device__ __forceinline__ float io_load(const T *p);
template <> __device__ __forceinline__ float io_load<float>(const float *p) { return __ldg(p); }
template <> __device__ __forceinline__ float io_load<unsigned short>(const unsigned short *p) { return bf16_to_f32(__ldg(p)); }
template <class T> __device__ __forceinline__ void io_store(T *p, float v);
template <> __device__ __forceinline__ void io_store<float>(float *p, float v) { *p = v; }
template <> __device__ __forceinline__ void io_store<unsigned short>(unsigned short *p, float v) { *p = f32_to_bf16(v); }
// 4 / 2 consecutive elements (16- / 8-byte aligned for float, 8 / 4 for bf16)
__device__ __forceinline__ void io_store4(float *p, float a, float b, float c, float d) { *reinterpret_cast<float4 *>(p) = make_float4(a, b, c, d); }
__device__ __forceinline__ void io_store4(unsigned short *p, float a, float b, float c, float d)
{
    *reinterpret_cast<uint2 *>(p) = make_uint2(pack_bf16x2(a, b), pack_bf16x2(c, d));
}
__device__ __forceinline__ void io_store2(float *p, float a, float b) { *reinterpret_cast<float2 *>(p) = make_float2(a, b); }
__device__ __forceinline__ void io_store2(unsigned short *p, float a, float b) { *reinterpret_cast<uint32_t *>(p) = pack_bf16x2(a, b); }

// Parameters shared by the forward and backward tensor-core kernels.
struct TcArgs {
    const float *x;          // [B][D0][hw]
    const float *W[BTS_MAX_LAYERS];  // raw weights [out][in] of every layer; each CTA builds its operand images itself
    float *out;              // forward: plane_eq [B][4][hw] (may be null) or head output [B][1][hw]
    float *depth;            // forward fused: [B][h*r][w*r] (null when not fused)
    float *abs_min;
    const float *grad;       // backward: grad_depth (fused) / grad_out
    float *grad_x;
    float *gw[BTS_MAX_LAYERS];
    int B, h, w, hw;
    int x_vec16;             // hw % 4 == 0 (% 8 for bf16) and x 16-byte aligned: 128-bit cp.async of a pixel group is legal
    int bf16_io;             // bf16 twin: x, out, depth, grad, grad_x point at bf16 tensors
    long long npx;           // B*hw
    float max_depth, eps;
    // activation stash (forward writes, backward reads; null = recompute): rows [channel][stash_npp] of
    // a_1 .. a_NW (= a16), a8 on a pixel axis padded per image to stash_hw_pad, and plane_eq [B][4][hw]
    float *stash_act;
    float *stash_pe;
    float *stash_x;          // bf16 twin only: the features as fp32 rows [D0][stash_npp] (the TMA weight-gradient kernel's operand)
    long long stash_npp;
    int stash_hw_pad;
    // post-LPG glue (pytorch/bts.py:255-256, 269-270, 283) folded into the fused stage: the depth map is divided
    // by depth_div (max_depth; 1 = plain depth), images are depth_bstride elements apart (a channel slice of a
    // concat buffer; h*R*w*R when dense), and ds_out [B][2h][2w] (nullable, upratio 8 / 4 only) receives the
    // nearest-neighbour down-sampling by R/2 = rows/cols 0 and R/2 of every block. Backward twins: grad of the
    // scaled map with its batch stride, grad of the down-sampled map (nullable).
    float depth_div;
    long long depth_bstride;
    float *ds_out;
    long long ds_bstride;
    long long grad_bstride;
    const float *grad_ds;
    long long grad_ds_bstride;
};

// Chains whose backward is the multi-kernel one (chain_tc_bwd.cu) can leave their activations in a
// caller-provided stash so that backward skips the recompute; the two narrow chains are recomputed
// inside the one-pass backward (chain_bwd_fused.cu) and never stash.
template <class CFG>
constexpr bool kStashCfg = !(CFG::NW <= 2 && CFG::dim(0) >= 2 * CFG::dim(1));

// first row of a_l / delta_l (l = 1..NW) in the backward workspace / the stash; the tail rows follow
template <class CFG>
struct BwdRows {
    __host__ __device__ static constexpr int row(int l)
    {
        int r = 0;
        for (int j = 1; j < l; ++j) r += CFG::dim(j);
        return r;
    }
    __host__ __device__ static constexpr int row8() { return row(CFG::NW + 1); }
    __host__ __device__ static constexpr int rowz() { return row8() + kTailMid; }
};

using namespace tc;

// ---- one wide layer: 3 x (K/8) MMAs issued by a single thread, then commit -> mbarrier
template <class CFG, int L>
__device__ __forceinline__ void tc_issue_layer(uint32_t wg_base, const float *sw, uint64_t *bar)
{
    constexpr int K = CFG::K(L), N = CFG::N(L), KMAX = CFG::kmax();
    constexpr uint32_t LBO = 128, SBO = (K / 4) * 128;
    constexpr uint32_t idesc = idesc_tf32(kTile, N, 0, 0);
    const uint32_t whi = smem_u32(sw + CFG::w_off(L));
    const uint32_t wlo = whi + N * K * 4;
    const uint32_t d = wg_base + 2 * KMAX;
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 8)   // lo(a) * hi(W)
        mma_tf32_ts(d, wg_base + KMAX + k0, smem_desc(whi + (k0 / 4) * LBO, LBO, SBO), idesc, k0 > 0);
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 8)   // hi(a) * lo(W)
        mma_tf32_ts(d, wg_base + k0, smem_desc(wlo + (k0 / 4) * LBO, LBO, SBO), idesc, 1);
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 8)   // hi(a) * hi(W)
        mma_tf32_ts(d, wg_base + k0, smem_desc(whi + (k0 / 4) * LBO, LBO, SBO), idesc, 1);
    mma_commit(bar);
}

// ---- hand a finished A operand to the tensor core and wait for the layer's accumulator
template <class CFG, int L>
__device__ __forceinline__ void tc_run_layer(uint32_t wg_base, const float *sw, uint64_t *bar, uint32_t &phase,
                                             int wg, int t)
{
    wait_st();
    fence_before_sync();
    named_bar_sync(1 + wg, kTile);
    if (t == 0) {                                                 // t: warp index inside the warpgroup, made warp-uniform once by the kernel
        if (elect_one()) {
            fence_after_sync();
            tc_issue_layer<CFG, L>(wg_base, sw, bar);
        }
        __syncwarp();
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    fence_after_sync();
}

// ---- epilogue of a wide layer that feeds another wide layer: D -> ELU -> (hi, lo) -> A
template <class CFG, int L>
__device__ __forceinline__ void tc_epilogue_to_a(uint32_t lane_addr)
{
    constexpr int N = CFG::N(L), KMAX = CFG::kmax();
#pragma unroll
    for (int n0 = 0; n0 < N; n0 += 16) {
        uint32_t r[16], lo[16];
        tmem_ld<16>(lane_addr + 2 * KMAX + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float v = elu_fast(__uint_as_float(r[j]));
            r[j] = __float_as_uint(v);
            lo[j] = __float_as_uint(tf32_lo(v));
        }
        tmem_st<16>(lane_addr + n0, r);
        tmem_st<16>(lane_addr + KMAX + n0, lo);
    }
}

// same, and a_{L+1} also goes to its [channel][pixel] rows (row stride npx_pad) when `store`
template <class CFG, int L>
__device__ __forceinline__ void tc_epilogue_to_a_spill(uint32_t lane_addr, float *act_gp, long long npx_pad, bool store = true)
{
    constexpr int N = CFG::N(L), KMAX = CFG::kmax();
    float *dst = act_gp + (size_t)BwdRows<CFG>::row(L + 1) * npx_pad;
#pragma unroll
    for (int n0 = 0; n0 < N; n0 += 16) {
        uint32_t r[16], lo[16];
        tmem_ld<16>(lane_addr + 2 * KMAX + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float v = elu_fast(__uint_as_float(r[j]));
            if (store) dst[(size_t)(n0 + j) * npx_pad] = v;
            r[j] = __float_as_uint(v);
            lo[j] = __float_as_uint(tf32_lo(v));
        }
        tmem_st<16>(lane_addr + n0, r);
        tmem_st<16>(lane_addr + KMAX + n0, lo);
    }
}

// ---- chunked variants: the A operand region is AW columns wide (hi at [0,AW), lo at [AW,2AW), D at 2AW),
// so a layer with K > AW input channels runs as K/AW rounds of MMAs that accumulate into the same D.
// With AW = 64 a 128-channel chain needs 256 TMEM columns per warpgroup and two warpgroups fit.
// NOLO: the A operand has no remainder part (bf16 features): the lo(a) * hi(W) pass is skipped
template <class CFG, int L, int AW, int KOFF, int KC, bool NOLO = false>
__device__ __forceinline__ void tc_issue_chunk(uint32_t wg_base, const float *sw, uint64_t *bar)
{
    constexpr int K = CFG::K(L), N = CFG::N(L);
    constexpr uint32_t LBO = 128, SBO = (K / 4) * 128;
    constexpr uint32_t idesc = idesc_tf32(kTile, N, 0, 0);
    const uint32_t whi = smem_u32(sw + CFG::w_off(L));
    const uint32_t wlo = whi + N * K * 4;
    const uint32_t d = wg_base + 2 * AW;
    if constexpr (!NOLO) {
#pragma unroll
        for (int k0 = 0; k0 < KC; k0 += 8)   // lo(a) * hi(W)
            mma_tf32_ts(d, wg_base + AW + k0, smem_desc(whi + ((KOFF + k0) / 4) * LBO, LBO, SBO), idesc, (KOFF + k0) > 0);
    }
#pragma unroll
    for (int k0 = 0; k0 < KC; k0 += 8)   // hi(a) * lo(W)
        mma_tf32_ts(d, wg_base + k0, smem_desc(wlo + ((KOFF + k0) / 4) * LBO, LBO, SBO), idesc, NOLO ? (KOFF + k0) > 0 : 1);
#pragma unroll
    for (int k0 = 0; k0 < KC; k0 += 8)   // hi(a) * hi(W)
        mma_tf32_ts(d, wg_base + k0, smem_desc(whi + ((KOFF + k0) / 4) * LBO, LBO, SBO), idesc, 1);
    mma_commit(bar);
}

// SUBS: threads per pixel. With SUBS = 2 a tile group is 256 threads: warps w and w + 4 share TMEM lane quarter w and
// split every epilogue's columns (sub = 0 / 1), which halves the SIMT latency of a tile and doubles the warps per SM.
template <class CFG, int L, int AW, int KOFF, int KC, int SUBS = 1, bool NOLO = false>
__device__ __forceinline__ void tc_run_chunk(uint32_t wg_base, const float *sw, uint64_t *bar, uint32_t &phase,
                                             int wg, int t)
{
    wait_st();
    fence_before_sync();
    named_bar_sync(1 + wg, SUBS * kTile);
    if (t == 0) {                                                 // t: warp index inside the warpgroup, made warp-uniform once by the kernel
        if (elect_one()) {
            fence_after_sync();
            tc_issue_chunk<CFG, L, AW, KOFF, KC, NOLO>(wg_base, sw, bar);
        }
        __syncwarp();
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    fence_after_sync();
}

// channels [C0, C0+KC) of a_{L+1} = ELU(D[C0 ..]) -> (hi, lo) -> A columns [0, KC); optionally to the stash rows
template <class CFG, int L, int AW, int C0, int KC, bool SPILL, int SUBS = 1>
__device__ __forceinline__ void tc_epilogue_chunk(uint32_t lane_addr, float *act_gp, long long npx_pad, bool store, int sub = 0)
{
    static_assert((KC / SUBS) % 16 == 0, "every sub-thread takes whole 16-column steps");
    float *dst = SPILL ? act_gp + (size_t)(BwdRows<CFG>::row(L + 1) + C0) * npx_pad : nullptr;
    // rolled: the wide chains run two to five tiles per warpgroup, so their straight-line code is fetched cold
    // (instruction-cache misses were the top stall of the 8x8 stage's forward when this was unrolled)
    const int nb = sub * (KC / SUBS), ne = nb + KC / SUBS;
#pragma unroll 1
    for (int n0 = nb; n0 < ne; n0 += 16) {
        uint32_t r[16], lo[16];
        tmem_ld<16>(lane_addr + 2 * AW + C0 + n0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
            const float v0 = elu_fast(__uint_as_float(r[j])), v1 = elu_fast(__uint_as_float(r[j + 1]));
            if constexpr (SPILL) {
                if (store) {
                    dst[(size_t)(n0 + j) * npx_pad] = v0;
                    dst[(size_t)(n0 + j + 1) * npx_pad] = v1;
                }
            }
            r[j] = __float_as_uint(v0);
            r[j + 1] = __float_as_uint(v1);
            tf32_lo2(v0, v1, lo[j], lo[j + 1]);
        }
        tmem_st<16>(lane_addr + n0, r);
        tmem_st<16>(lane_addr + AW + n0, lo);
    }
}

// layer L >= 1 from the previous layer's accumulator: K/AW rounds of (epilogue chunk, MMAs). The MMAs of
// round 0 overwrite D[0, N): legal because round 0's epilogue has consumed D[0, AW) and N <= AW.
template <class CFG, int L, int AW, bool SPILL, int SUBS = 1>
__device__ __forceinline__ void tc_layer_from_d(uint32_t wg_base, uint32_t lane_addr, const float *sw, uint64_t *bar,
                                                uint32_t &phase, int wg, int t, float *act_gp, long long npx_pad, bool store,
                                                int sub = 0)
{
    constexpr int K = CFG::K(L);
    if constexpr (K <= AW) {
        tc_epilogue_chunk<CFG, L - 1, AW, 0, K, SPILL, SUBS>(lane_addr, act_gp, npx_pad, store, sub);
        tc_run_chunk<CFG, L, AW, 0, K, SUBS>(wg_base, sw, bar, phase, wg, t);
    } else {
        static_assert(K == 2 * AW && CFG::N(L) <= AW, "two rounds, accumulator no wider than one round's inputs");
        tc_epilogue_chunk<CFG, L - 1, AW, 0, AW, SPILL, SUBS>(lane_addr, act_gp, npx_pad, store, sub);
        tc_run_chunk<CFG, L, AW, 0, AW, SUBS>(wg_base, sw, bar, phase, wg, t);
        tc_epilogue_chunk<CFG, L - 1, AW, AW, AW, SPILL, SUBS>(lane_addr, act_gp, npx_pad, store, sub);
        tc_run_chunk<CFG, L, AW, AW, AW, SUBS>(wg_base, sw, bar, phase, wg, t);
    }
}

// ---- LPG expansion of this thread's plane (n1..n4) into depth; returns the abs_min key.
// The FUSED stage is held to 1e-5 of the reference, not to bit-exactness (the standalone kernels in lpg.cu are, and
// keep IEEE division): per output pixel this is 2 adds for den in the reference's order, one min for abs_min, the clamp
// and  depth = n4 * rcp(den)  with the hardware reciprocal (relative error 2e-7) -- 10 instructions where IEEE
// division, integer min keys and a NaN test per pixel took 22; the expansion was 27 % of the 8x8 forward's instructions.
// NaN: den is NaN only when a plane coefficient is (they are sines / cosines of finite angles otherwise), so NaN is
// looked for once per low-res pixel.
__device__ __forceinline__ float rcp_fast(float v)
{
    float r;
    asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}
// clamp of bts.py:168-171 on |den| (den == 0 stays 0 -> inf), then n4 / den
__device__ __forceinline__ float lpg_depth_fast(float n4, float den, float eps)
{
    const float ad = fabsf(den);
    const float dc = (ad < eps && ad > 0.f) ? copysignf(eps, den) : den;
    return n4 * rcp_fast(dc);
}
__device__ __forceinline__ int lpg_key(float amin, bool nan) { return nan ? (int)0xFFC00000 : __float_as_int(amin); }

// SUBS = 2: sub-thread `sub` writes half of the block: the 16 pixels of half `sub` (R = 8), rows [sub R/2, (sub+1) R/2)
template <int R, int SUBS = 1, class TO = float>
__device__ __forceinline__ int lpg_expand(const TcArgs &a, float n1, float n2, float n3, float n4,
                                          unsigned gp, int b_own, int p_own, bool valid, int lane, int sub = 0)
{
    TO *const depth_base = reinterpret_cast<TO *>(a.depth);
    static_assert(SUBS == 1 || (R % 2) == 0, "two sub-threads split an even number of rows");
    // all index arithmetic is 32-bit: the C ABI rejects more than 2^31 - 1 low-res pixels
    float amin = __int_as_float(0x7f800000);
    bool nan = false;
    if (a.depth_div != 1.f) n4 = __fdiv_rn(n4, a.depth_div);       // depth / max_depth = (n4 / max_depth) / den
    if constexpr (R == 8) {
        // two 512 B stores per output row and warp: lane L writes columns (L&1)*4..+3 of pixel 16*half + L/2
        const unsigned gp0 = gp - (unsigned)lane;
#pragma unroll 1
        for (int half = (SUBS == 2 ? sub : 0); half < (SUBS == 2 ? sub + 1 : 2); ++half) {
            const int src = half * 16 + (lane >> 1);
            const float m1 = __shfl_sync(0xffffffffu, n1, src), m2 = __shfl_sync(0xffffffffu, n2, src);
            const float m3 = __shfl_sync(0xffffffffu, n3, src), m4 = __shfl_sync(0xffffffffu, n4, src);
            const unsigned gs = gp0 + (unsigned)src;
            if (gs < (unsigned)a.npx) {
                nan |= (m1 != m1) | (m2 != m2) | (m3 != m3);
                const int b = (int)(gs / (unsigned)a.hw);
                const int rem = (int)(gs - (unsigned)b * (unsigned)a.hw);
                const int i = rem / a.w, j = rem - i * a.w;
                const int q0 = (lane & 1) * 4;
                const int Wd = a.w * 8;
                TO *out = depth_base + (size_t)b * a.depth_bstride + (size_t)i * 8 * (size_t)Wd + (size_t)j * 8 + q0;
                float *ds = a.ds_out ? a.ds_out + (size_t)b * a.ds_bstride + (size_t)(2 * i) * (2 * a.w) + 2 * j + (lane & 1) : nullptr;
                float n1u[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) n1u[k] = __fmul_rn(m1, lpg_offset(q0 + k, 8));
#pragma unroll 2
                for (int p = 0; p < 8; ++p) {
                    const float n2v = __fmul_rn(m2, lpg_offset(p, 8));
                    float4 o;
                    float *op = reinterpret_cast<float *>(&o);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float den = lpg_den(n1u[k], n2v, m3);
                        amin = fminf(amin, fabsf(den));
                        op[k] = lpg_depth_fast(m4, den, a.eps);
                    }
                    io_store4(out + (size_t)p * Wd, o.x, o.y, o.z, o.w);
                    if ((p & 3) == 0 && ds != nullptr) ds[(size_t)(p >> 2) * (2 * a.w)] = op[0];   // rows 0, 4; cols 0, 4
                }
            }
        }
    } else if (valid) {
        nan = (n1 != n1) | (n2 != n2) | (n3 != n3);
        const int b = b_own, rem = p_own;
        const int i = rem / a.w, j = rem - i * a.w;
        const int Wd = a.w * R;
        TO *out = depth_base + (size_t)b * a.depth_bstride + (size_t)i * R * (size_t)Wd + (size_t)j * R;
        float n1u[R];
#pragma unroll
        for (int k = 0; k < R; ++k) n1u[k] = __fmul_rn(n1, lpg_offset(k, R));
#pragma unroll
        for (int p = 0; p < R; ++p) {
            if (SUBS == 2 && (p / (R / 2 > 0 ? R / 2 : 1)) != sub) continue;
            const float n2v = __fmul_rn(n2, lpg_offset(p, R));
            float o[R];
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const float den = lpg_den(n1u[k], n2v, n3);
                amin = fminf(amin, fabsf(den));
                o[k] = lpg_depth_fast(n4, den, a.eps);
            }
            TO *row = out + (size_t)p * Wd;
            if constexpr (R == 4) {
                if ((p & 1) == 0 && a.ds_out != nullptr) {        // rows 0, 2; cols 0, 2
                    float *ds = a.ds_out + (size_t)b * a.ds_bstride + (size_t)(2 * i + (p >> 1)) * (2 * a.w) + 2 * j;
                    ds[0] = o[0]; ds[1] = o[2];
                }
            }
            if constexpr (R == 4) io_store4(row, o[0], o[1], o[2], o[3]);
            else if constexpr (R == 2) io_store2(row, o[0], o[1]);
            else {
#pragma unroll
                for (int k = 0; k < R; ++k) io_store(row + k, o[k]);
            }
        }
    }
    return lpg_key(amin, nan);
}

// ---- upstream gradient of one low-res pixel in fused mode: reduce grad_depth over its R x R block
template <int R>
__device__ __forceinline__ void lpg_block_grad(const TcArgs &a, const PlaneFwd &pl, int b, int rem, float g[4])
{
    const int i = rem / a.w, j = rem - i * a.w;
    const int Wd = a.w * R;
    const float *gp0 = a.grad + (size_t)b * a.grad_bstride + (size_t)i * R * (size_t)Wd + (size_t)j * R;
    LpgAcc acc{0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int p = 0; p < R; ++p) {
        float gv[R];
        const float *row = gp0 + (size_t)p * Wd;
        if constexpr (R % 4 == 0) {
#pragma unroll
            for (int k = 0; k < R; k += 4) {
                const float4 t4 = __ldcs(reinterpret_cast<const float4 *>(row + k));
                gv[k] = t4.x; gv[k + 1] = t4.y; gv[k + 2] = t4.z; gv[k + 3] = t4.w;
            }
        } else if constexpr (R == 2) {
            const float2 t2 = __ldcs(reinterpret_cast<const float2 *>(row));
            gv[0] = t2.x; gv[1] = t2.y;
        } else {
#pragma unroll
            for (int k = 0; k < R; ++k) gv[k] = __ldcs(row + k);
        }
        if constexpr (R == 4 || R == 8) {                          // gradient of the down-sampled map lands on rows/cols 0, R/2
            if (a.grad_ds != nullptr && (p % (R / 2)) == 0) {
                const float *gd = a.grad_ds + (size_t)b * a.grad_ds_bstride + (size_t)(2 * i + p / (R / 2)) * (2 * a.w) + 2 * j;
                gv[0] += __ldcs(gd);
                gv[R / 2] += __ldcs(gd + 1);
            }
        }
        const float v = lpg_offset(p, R);
        const float n2v = __fmul_rn(pl.n2, v);
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const float u = lpg_offset(k, R);
            lpg_bwd_accum(acc, gv[k], __fmul_rn(pl.n1, u), n2v, pl.n3, u, v, a.eps);
        }
    }
    const float sc = a.depth_div != 1.f ? __fdiv_rn(1.f, a.depth_div) : 1.f;   // d(depth / max_depth) / d depth
    g[0] = -pl.n4 * acc.tu * sc; g[1] = -pl.n4 * acc.tv * sc; g[2] = -pl.n4 * acc.t * sc; g[3] = acc.s4 * sc;
}

// ---- last layer 8 -> (3|1): logits from a8 (shared by the forward tail and the stash backward, so both
// see bit-identical logits). Packed FMAs: the even and the odd input channels accumulate in the two halves.
template <int LAST>
__device__ __forceinline__ void tail_logits(const float *w5, const float (&a8)[kTailMid], float (&z)[LAST])
{
    f32x2 a2[kTailMid / 2];
#pragma unroll
    for (int c = 0; c < kTailMid; c += 2) a2[c / 2] = pack2(a8[c], a8[c + 1]);
#pragma unroll
    for (int o = 0; o < LAST; ++o) {
        f32x2 s = 0ull;
#pragma unroll
        for (int c = 0; c < kTailMid; c += 4) {
            const ulonglong2 w4 = *reinterpret_cast<const ulonglong2 *>(w5 + o * kTailMid + c);
            s = fma2(w4.x, a2[c / 2], s);
            s = fma2(w4.y, a2[c / 2 + 1], s);
        }
        z[o] = hsum2(s);
    }
}

// ---- the 16 -> 8 -> (3|1) tail on registers; weights broadcast from shared memory
// PACKED = false keeps scalar FMAs: the packed operands cost 16 more live registers, which the 2 x 2 stage's
// one-pass backward (168-register cap) pays for in spills.
template <int LAST, bool PACKED = true, bool FENCE = true>
__device__ __forceinline__ void tail_forward(const float *tw, const float (&a16)[kTailIn], float (&a8)[kTailMid],
                                             float (&z)[LAST])
{
    if constexpr (!PACKED) {
#pragma unroll
        for (int o = 0; o < kTailMid; ++o) {
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < kTailIn; c += 4) {
                const float4 w4 = *reinterpret_cast<const float4 *>(tw + o * kTailIn + c);
                s = fmaf(w4.x, a16[c], s); s = fmaf(w4.y, a16[c + 1], s);
                s = fmaf(w4.z, a16[c + 2], s); s = fmaf(w4.w, a16[c + 3], s);
            }
            a8[o] = elu_fast(s);
            asm volatile("" ::: "memory");
        }
        tail_logits<LAST>(tw + kTailMid * kTailIn, a8, z);
        return;
    }
    f32x2 a2[kTailIn / 2];
#pragma unroll
    for (int c = 0; c < kTailIn; c += 2) a2[c / 2] = pack2(a16[c], a16[c + 1]);
#pragma unroll
    for (int o = 0; o < kTailMid; ++o) {
        f32x2 s = 0ull;
#pragma unroll
        for (int c = 0; c < kTailIn; c += 4) {
            const ulonglong2 w4 = *reinterpret_cast<const ulonglong2 *>(tw + o * kTailIn + c);
            s = fma2(w4.x, a2[c / 2], s);
            s = fma2(w4.y, a2[c / 2 + 1], s);
        }
        a8[o] = elu_fast(hsum2(s));
        if constexpr (FENCE) asm volatile("" ::: "memory");   // keep the weight loads of later rows from being hoisted (register pressure)
    }
    tail_logits<LAST>(tw + kTailMid * kTailIn, a8, z);
}

// dl16 = W4^T d8 (16 <- 8) with packed FMAs: channel pairs (c, c+1) share an accumulator register pair
template <bool PACKED = true, int FENCE = 1>   // FENCE: scheduling fence after every FENCE-th row of W4 (0 = none)
__device__ __forceinline__ void tail_dgrad16(const float *tw, const float (&d8)[kTailMid], float (&dl16)[kTailIn])
{
    if constexpr (!PACKED) {
#pragma unroll
        for (int c = 0; c < kTailIn; ++c) dl16[c] = 0.f;
#pragma unroll
        for (int o = 0; o < kTailMid; ++o) {
#pragma unroll
            for (int c = 0; c < kTailIn; c += 4) {
                const float4 w4 = *reinterpret_cast<const float4 *>(tw + o * kTailIn + c);
                dl16[c] = fmaf(w4.x, d8[o], dl16[c]); dl16[c + 1] = fmaf(w4.y, d8[o], dl16[c + 1]);
                dl16[c + 2] = fmaf(w4.z, d8[o], dl16[c + 2]); dl16[c + 3] = fmaf(w4.w, d8[o], dl16[c + 3]);
            }
            asm volatile("" ::: "memory");
        }
        return;
    }
    f32x2 acc[kTailIn / 2];
#pragma unroll
    for (int c = 0; c < kTailIn / 2; ++c) acc[c] = 0ull;
#pragma unroll
    for (int o = 0; o < kTailMid; ++o) {
        const f32x2 d2 = pack2(d8[o], d8[o]);
#pragma unroll
        for (int c = 0; c < kTailIn; c += 4) {
            const ulonglong2 w4 = *reinterpret_cast<const ulonglong2 *>(tw + o * kTailIn + c);
            acc[c / 2] = fma2(w4.x, d2, acc[c / 2]);
            acc[c / 2 + 1] = fma2(w4.y, d2, acc[c / 2 + 1]);
        }
        if constexpr (FENCE > 0) {
            if ((o + 1) % FENCE == 0) asm volatile("" ::: "memory");   // bound the hoisting of weight loads
        }
    }
#pragma unroll
    for (int c = 0; c < kTailIn; c += 2) unpack2(acc[c / 2], dl16[c], dl16[c + 1]);
}

// ---- weight packing, done by every CTA for itself straight from the raw fp32 weights (they are a few
// hundred KB at most and L2-resident; no pack launch, no packed copy in HBM).
// Canonical K-major (hi | lo) operand images: with float4 unit u <-> (row%8 = u%8, k/4 = (u/8)%(K/4),
// row/8 = u/(2K)) the unit lands at float offset 4u (see kmajor_off_floats), so shared-memory stores are linear.
__device__ __forceinline__ float4 tf32_lo4(float4 v)
{
    return make_float4(tc::tf32_lo(v.x), tc::tf32_lo(v.y), tc::tf32_lo(v.z), tc::tf32_lo(v.w));
}

#ifndef BTS_PACK_PF
#define BTS_PACK_PF 8
#endif
// forward images (rows = output channel n, K = input channel) + raw tail weights W4 [8][16], W5 [last][8]
template <class CFG>
__device__ __forceinline__ void tc_pack_to_smem(const float *const *W, float *sw, int last, int tid, int nth)
{
    constexpr int PF = BTS_PACK_PF;   // loads in flight per thread: the packing is the kernel's prologue and pure L2 latency otherwise
#pragma unroll
    for (int l = 0; l < CFG::NW; ++l) {
        const int K = CFG::K(l), N = CFG::N(l);
        float *hi = sw + CFG::w_off(l), *lo = hi + N * K;
        const float *w = W[l];
        for (int u0 = tid; u0 < N * K / 4; u0 += PF * nth) {
            float4 v[PF];
#pragma unroll
            for (int i = 0; i < PF; ++i) {
                const int u = u0 + i * nth;
                if (u < N * K / 4) {
                    const int n = (u / (2 * K)) * 8 + (u & 7), kq = (u >> 3) % (K / 4);
                    v[i] = __ldg(reinterpret_cast<const float4 *>(w + (size_t)n * K + 4 * kq));
                }
            }
#pragma unroll
            for (int i = 0; i < PF; ++i) {
                const int u = u0 + i * nth;
                if (u < N * K / 4) {
                    *reinterpret_cast<float4 *>(hi + 4 * u) = v[i];
                    *reinterpret_cast<float4 *>(lo + 4 * u) = tf32_lo4(v[i]);
                }
            }
        }
    }
    float *tw = sw + CFG::tail_off();
    for (int e = tid; e < kTailMid * kTailIn; e += nth) tw[e] = __ldg(W[CFG::NW] + e);
    for (int e = tid; e < last * kTailMid; e += nth) tw[kTailMid * kTailIn + e] = __ldg(W[CFG::NW + 1] + e);
}

// transposed images for dgrad: layer l, rows = input channel c, K' = output channel o
template <class CFG>
__device__ __forceinline__ void tc_pack_t_to_smem(const float *const *W, float *swt, int tid, int nth)
{
#pragma unroll
    for (int l = 0; l < CFG::NW; ++l) {
        const int K = CFG::K(l), N = CFG::N(l);
        float *hi = swt + CFG::w_off(l), *lo = hi + N * K;
        const float *w = W[l];
        constexpr int PF = BTS_PACK_PF / 2;   // four scalar loads each
        for (int u0 = tid; u0 < N * K / 4; u0 += PF * nth) {
            float4 v[PF];
#pragma unroll
            for (int i = 0; i < PF; ++i) {
                const int u = u0 + i * nth;
                if (u < N * K / 4) {
                    const int c = (u / (2 * N)) * 8 + (u & 7), oq = (u >> 3) % (N / 4);
                    const float *src = w + (size_t)(4 * oq) * K + c;
                    v[i] = make_float4(__ldg(src), __ldg(src + K), __ldg(src + 2 * K), __ldg(src + 3 * K));
                }
            }
#pragma unroll
            for (int i = 0; i < PF; ++i) {
                const int u = u0 + i * nth;
                if (u < N * K / 4) {
                    *reinterpret_cast<float4 *>(hi + 4 * u) = v[i];
                    *reinterpret_cast<float4 *>(lo + 4 * u) = tf32_lo4(v[i]);
                }
            }
        }
    }
}

inline void tc_set_weights(TcArgs &a, const ChainParams &cp)
{
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) a.W[l] = l < cp.n_layers ? cp.W[l] : nullptr;
}

// chain signatures served by the tensor-core path (bts_size = 512 and 256 decoders)
using Cfg128_128 = TcCfg<4, 128, 128, 64, 32, 16>;
using Cfg128_64 = TcCfg<3, 128, 64, 32, 16>;
using Cfg64_64 = TcCfg<3, 64, 64, 32, 16>;
using Cfg64_32 = TcCfg<2, 64, 32, 16>;
using Cfg32_32 = TcCfg<2, 32, 32, 16>;
using Cfg32_16 = TcCfg<1, 32, 16>;

template <class CFG>
static bool cfg_matches(const ChainParams &cp)
{
    if (cp.n_layers != CFG::NW + 2) return false;
    for (int l = 0; l <= CFG::NW; ++l)
        if (cp.dims[l] != CFG::dim(l)) return false;
    return cp.dims[CFG::NW + 1] == kTailMid;
}

}  // namespace bts

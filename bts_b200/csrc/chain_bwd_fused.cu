// chain_bwd_fused.cu -- ONE-PASS backward of a narrow BTS decoder stage on sm_100a.
//
// Autograd backward of reduction_1x1 (+ local_planar_guidance) (pytorch/bts.py:124-173; SURVEY.md
// 8(a) row a8) for the chains whose first layer is at most 64 channels wide -- the full-resolution
// head 32->16->8->1 and the half-resolution stage 64->32->16->8->3, which together move 81 % of the
// decoder's bytes. x and grad are read once, grad_x is written once, nothing else touches HBM:
//   recompute the chain (3xTF32 tcgen05, A in TMEM)  ->  upstream gradient (sigmoid head / plane
//   parameterisation / LPG block reduction)  ->  SIMT tail backward  ->  dgrad through the wide layers
//   (tcgen05, transposed weight images)  ->  grad_x;   weight gradients accumulate in TMEM for the whole
//   kernel: every 64-pixel half tile is staged once as K-major SWIZZLE_128B operands with the PIXEL
//   axis as K, activations and their tf32 remainders stacked along M ([a_hi ; a_lo], M = 128), deltas
//   stacked along N ([d_hi ; d_lo], N = 64), so ONE MMA per 8 pixels yields hi*hi + hi*lo + lo*hi (+lo*lo).
//
// Roles: NWG compute warpgroups (thread = low-res pixel = TMEM lane) + one MMA-issuing warp per warpgroup. Hand-offs
// are mbarriers only (no CTA-wide barrier inside the tile loop):
//   ready[g]  4 warp arrivals compute -> issuer   "A operand for the next chain MMA is in TMEM"
//   done[g]   tcgen05.commit issuer -> compute  "chain MMA finished, D is readable"
//   stg[g][h] 2 warp arrivals compute -> issuer   "pair h's 64-pixel half tile is staged in shared memory"
//   wfree[g][h] tcgen05.commit issuer -> compute "the weight-gradient MMAs of pair h's half finished"
// The two warp pairs of a warpgroup share one staging buffer and take turns: pair 1 waits for
// wfree[g][0], pair 0 for wfree[g][1] (one waiter group per mbarrier, so nobody lags two phases).
#include <cstdlib>
#include <type_traits>

#include "chain_tc.cuh"

namespace bts {

constexpr int kBLo = 32 * 128;                         // byte offset of the remainder rows in the delta image

// MMA issue: 1 = inline, by the first warp of the warpgroup / warp pair that hands the operands over (split hardware
// barriers: the other warps bar.arrive and go on, the issuing warp bar.syncs) -- no issuer warps, so the CTA is
// NWG x 128 threads (255 registers at two warpgroups) and nobody polls; 0 = one dedicated issuer warp per warpgroup.
#ifndef BTS_FUSED_INLINE_ISSUE
#define BTS_FUSED_INLINE_ISSUE 1
#endif
constexpr bool kInlineIssue = BTS_FUSED_INLINE_ISSUE != 0;

// The tensor core adds into its fp32 TMEM accumulator with round-toward-zero: a weight-gradient accumulator that lives
// in TMEM for the whole kernel loses ~3.4e-8 of its magnitude per accumulate step, ALWAYS toward zero, so the error is a
// bias that grows with the pixels a CTA sees (measured against the fp64 oracle: 2.6e-5 at 416x544 b4, where a CTA issues
// ~760 steps per accumulator; it would be 2e-4 at KITTI b16). Every kWgradDrain iterations of the tile loop the CTA
// therefore drains the accumulators into fp32 master sums (round-to-nearest adds; spare TMEM columns or registers)
// and zeroes them. Measured at KITTI b16 with a drain every 4 iterations: grad_W within 2.6e-6 (head, one accumulator
// set per warpgroup: 64 steps per chain) / 2.8e-6 (2x2 stage, one shared set: 128 steps) of the fp64 oracle's peak.
#ifndef BTS_WGRAD_DRAIN
#define BTS_WGRAD_DRAIN 4
#endif
constexpr int kWgradDrain = BTS_WGRAD_DRAIN;
// Head with private staging buffers: write the feature rows of the weight-gradient operand at the START of the tile, from
// the registers that hold them anyway, instead of reading them back from TMEM at its end (32 tcgen05.ld per thread
// less, and the tile's tail -- everything between the last MMA hand-off and the next tile -- gets shorter).
#ifndef BTS_STAGE_X_EARLY
#define BTS_STAGE_X_EARLY 0   // measured: 199.3 vs 190.2 us (the wait for the previous tile's MMAs at the top of the tile costs more than it saves)
#endif
#ifndef BTS_DRAIN_DEFERRED
#define BTS_DRAIN_DEFERRED 0
#endif
constexpr bool kDrainDeferred = BTS_DRAIN_DEFERRED != 0;   // drain one iteration late (see the tile loop) or at the iteration's end
constexpr int fused_threads(int nwg) { return nwg * (kTile + (kInlineIssue ? 0 : 32)); }

struct FusedArgs {
    TcArgs a;
    float *partials;        // [gridDim.x][total_w] (deterministic route: fixed-order reduction kernel afterwards)
    int total_w;
    int atomic_w;           // 1: every CTA adds its partial sums to a.gw[] with red.global.add (no reduction launch)
    int drain_every;        // tile-loop iterations between drains of the weight-gradient accumulators (kWgradDrain)
};

// D[128 x N](TMEM, col d) = A[128 x K](TMEM; hi at a_hi, lo at a_lo) * B[N x K]^T (smem image at w: hi | lo)
// NOLO: the A operand has no remainder part (bf16 features are exact in tf32): two passes instead of three
template <int K, int N, bool NOLO = false>
__device__ __forceinline__ void issue_ts3(uint32_t d, uint32_t a_hi, uint32_t a_lo, const float *w)
{
    constexpr uint32_t LBO = 128, SBO = (K / 4) * 128;
    constexpr uint32_t idesc = idesc_tf32(kTile, N, 0, 0);
    const uint32_t whi = smem_u32(w), wlo = whi + N * K * 4;
    if constexpr (!NOLO) {
#pragma unroll
        for (int k0 = 0; k0 < K; k0 += 8)
            mma_tf32_ts(d, a_lo + k0, smem_desc(whi + (k0 / 4) * LBO, LBO, SBO), idesc, k0 > 0);
    }
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 8)
        mma_tf32_ts(d, a_hi + k0, smem_desc(wlo + (k0 / 4) * LBO, LBO, SBO), idesc, NOLO ? k0 > 0 : 1);
#pragma unroll
    for (int k0 = 0; k0 < K; k0 += 8)
        mma_tf32_ts(d, a_hi + k0, smem_desc(whi + (k0 / 4) * LBO, LBO, SBO), idesc, 1);
}

// weight-gradient MMAs of one staged half tile (two 32-pixel sub-buffers): D[128 x 64] += A_cat * B_cat^T
template <int kSubBytes, int kSubB>
__device__ __forceinline__ void issue_wgrad_half(uint32_t d, uint32_t sub0)
{
    constexpr uint32_t idesc = idesc_tf32(128, 64, 0, 0);
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
            mma_tf32_ss(d, smem_desc_sw128(sub0 + s * kSubBytes + ks * 32),
                        smem_desc_sw128(sub0 + s * kSubBytes + kSubB + ks * 32), idesc, 1);
}

template <class CFG, int MODE, int R, int NWG, bool BF = false>
struct FusedLayout {
    static constexpr bool FINAL = MODE == 0;
    static constexpr int LAST = FINAL ? 1 : 3;
    static constexpr int NW = CFG::NW;
    static constexpr int K0 = CFG::dim(0);
    static constexpr int NJOBS = NW;                       // tail round (+ the x round when NW == 2)
    static constexpr int ACC = NJOBS * 64;                 // TMEM columns of the weight-gradient accumulators
    static constexpr int WGCOLS = 3 * K0;                  // per warpgroup: x_hi | x_lo / A2 | D
    static constexpr int PACK = CFG::pack_floats(LAST);
    static constexpr int IMG = CFG::tail_off();
    static constexpr int NCMD = NW == 1 ? 4 : 8;
    // staging: one sub-buffer per 32 pixels = activation image [hi rows | lo rows at ALO] + delta image
    // [hi | lo at +32 rows]. PRIV (head with <= 2 warpgroups): every warp pair owns its two sub-buffers, the
    // remainder rows start right after the 56 used rows; otherwise the pairs of a warpgroup share two.
    static constexpr bool PRIV = NW == 1 && NWG <= 2;
    static constexpr int ALO_ROWS = PRIV ? 56 : 64;
    static constexpr int ALO = ALO_ROWS * 128;
    static constexpr int SUBB = PRIV ? 112 * 128 : 128 * 128;
    static constexpr int SUB = SUBB + 64 * 128;
    static constexpr int NSUB = PRIV ? 4 : 2;
    static constexpr size_t stage_bytes = (size_t)NWG * NSUB * SUB;
    // next tile's features [channel][pixel] + 4 upstream-gradient words per pixel (cp.async)
    // (bf16 twin: the feature rows are 2-byte elements; the 4 gradient words per pixel stay fp32)
    static constexpr size_t xs_wg_bytes = (size_t)K0 * kTile * (BF ? 2 : 4) + 4 * kTile * sizeof(float);
    static constexpr size_t xs_bytes = (size_t)NWG * xs_wg_bytes;
    static constexpr size_t smem_bytes = 1024 + stage_bytes + xs_bytes + sizeof(float) * (PACK + IMG) + 8 * 6 * NWG + 16;
    static_assert(NW == 1 || NW == 2, "one or two tensor-core layers");
    static_assert(CFG::dim(NW) == kTailIn && CFG::dim(NW - 1) == 32, "chain must end ... -> 32 -> 16 -> 8");
    // weight-gradient accumulators: one set per warpgroup when the 512 TMEM columns allow it (the head), so that a
    // warpgroup drains its own set behind its own barrier; else one set shared by the CTA (2x2 stage), drained CTA-wide
#ifndef BTS_FUSED_PERWG
#define BTS_FUSED_PERWG 1
#endif
    static constexpr bool PERWG = BTS_FUSED_PERWG && NWG * ACC + NWG * WGCOLS <= 512;
    static constexpr int NACC = PERWG ? NWG : 1;
    static constexpr int WG0 = NACC * ACC;                 // first TMEM column of the warpgroups' own regions
    static constexpr int DRC = PERWG ? 32 : 16;            // summed accumulator columns a thread drains per job
    // where the drained sums live between drains: spare TMEM columns when there are some (the head: no registers are
    // held across the tile loop, whose unfenced SIMT tails want them all), else registers
    static constexpr bool TMASTER = PERWG && WG0 + NWG * WGCOLS + NWG * NJOBS * 32 <= 512;
    static constexpr int MASTER0 = WG0 + NWG * WGCOLS;     // first master column
    static_assert(WG0 + NWG * WGCOLS <= 512, "TMEM budget");
};

// ---- the MMA-issuing warps: one per warpgroup, each walks its warpgroup's fixed command list in order.
//        NW == 1: F0, DG0, WT(h0), WT(h1).     NW == 2: F0, F1, DG1, WT(h0), DG0, WT(h1), WX(h0), WX(h1).
// The whole warp runs the control flow (blocking mbarrier waits); the MMAs and the commit sit behind elect.sync
// with warp-uniform operands, so they compile to back-to-back UTCHMMA. (A single polling thread behind
// `if (tid == k)` cost ~8 instructions / ~45 cycles per MMA and was ~80 % busy: it bounded the kernel.)
template <class CFG, int MODE, int R, int NWG>
__device__ __noinline__ void fused_issuer(int g, const float *sw, const float *swt, uint32_t stage_addr, uint64_t *bars,
                                          uint32_t tbase, int ntiles)
{
    using L = FusedLayout<CFG, MODE, R, NWG>;
    constexpr int K0 = L::K0, NW = L::NW;
    const int stride = (int)gridDim.x * NWG;
    const int first = (int)blockIdx.x * NWG + g;
    const int mine = first < ntiles ? (ntiles - first + stride - 1) / stride : 0;
    uint64_t *ready = bars + 6 * g, *done = ready + 1, *stg0 = ready + 2, *stg1 = ready + 3, *wfree0 = ready + 4, *wfree1 = ready + 5;
    const uint32_t wb = tbase + L::WG0 + g * L::WGCOLS;           // x_hi
    const uint32_t tacc = tbase + (L::PERWG ? g * L::ACC : 0);
    const uint32_t xl = wb + K0, dd = wb + 2 * K0;
    // shared staging: both pairs use sub-buffers 0,1; private: pair h uses 2h, 2h+1
    const uint32_t subA = stage_addr + g * L::NSUB * L::SUB;
    const uint32_t subB = subA + (L::PRIV ? 2 * L::SUB : 0);
    static_assert(!L::PRIV || NW == 1, "private staging is laid out for the single-round chains");
    uint32_t rph = 0, sph0 = 0, sph1 = 0;
#pragma unroll 1
    for (int i = 0; i < mine; ++i) {
#pragma unroll
        for (int c = 0; c < L::NCMD; ++c) {
            const bool is_w = (NW == 1) ? (c >= 2) : (c == 3 || c >= 5);
            const bool h1 = is_w && ((NW == 1) ? (c == 3) : (c == 5 || c == 7));   // which pair's half this command multiplies
            if (!is_w) {
                mbar_wait(ready, rph);
                rph ^= 1;
            } else if (!h1) {
                mbar_wait(stg0, sph0);
                sph0 ^= 1;
            } else {
                mbar_wait(stg1, sph1);
                sph1 ^= 1;
            }
            fence_after_sync();
            if (elect_one()) {
                const uint32_t sub0 = h1 ? subB : subA;
                if constexpr (NW == 1) {
                    if (c == 0) { issue_ts3<CFG::K(0), CFG::N(0)>(dd, wb, xl, sw + CFG::w_off(0)); mma_commit(done); }
                    else if (c == 1) { issue_ts3<CFG::N(0), CFG::K(0)>(dd, xl, xl + 16, swt + CFG::w_off(0)); mma_commit(done); }
                    else { issue_wgrad_half<L::SUB, L::SUBB>(tacc, sub0); mma_commit(h1 ? wfree1 : wfree0); }
                } else {
                    if (c == 0) { issue_ts3<CFG::K(0), CFG::N(0)>(dd, wb, xl, sw + CFG::w_off(0)); mma_commit(done); }
                    else if (c == 1) { issue_ts3<CFG::K(1), CFG::N(1)>(dd, xl, xl + 32, sw + CFG::w_off(1)); mma_commit(done); }
                    else if (c == 2) { issue_ts3<CFG::N(1), CFG::K(1)>(dd, xl + 32, xl + 48, swt + CFG::w_off(1)); mma_commit(done); }
                    else if (c == 4) { issue_ts3<CFG::N(0), CFG::K(0)>(dd, xl, xl + 32, swt + CFG::w_off(0)); mma_commit(done); }
                    else if (c == 3 || c == 5) { issue_wgrad_half<L::SUB, L::SUBB>(tacc, sub0); mma_commit(h1 ? wfree1 : wfree0); }
                    else { issue_wgrad_half<L::SUB, L::SUBB>(tacc + 64, sub0); mma_commit(h1 ? wfree1 : wfree0); }
                }
            }
            __syncwarp();
        }
    }
}

// ---- staging helpers. A thread owns pixel `lane` of its warp's 32-pixel sub-buffer; row r of an image
// sits at r*128 bytes, its 16-byte chunks XOR-swizzled with r%8 (a warp's 32 stores of one row cover
// all 32 banks). `coff[q]` is the thread's byte offset inside a row with r%8 == q. Callers pass
// compile-time rows (fully unrolled loops), so coff[] stays in registers.
struct StageCtx {
    unsigned char *sub;     // this warp's sub-buffer
    uint32_t coff[8];
};
// two consecutive rows at once (packed remainder)
template <int kALo>
__device__ __forceinline__ void stage_a2(const StageCtx &s, int row, float v0, float v1)
{
    unsigned char *p0 = s.sub + row * 128 + s.coff[row & 7];
    unsigned char *p1 = s.sub + (row + 1) * 128 + s.coff[(row + 1) & 7];
    float l0, l1;
    tf32_lo2(v0, v1, l0, l1);
    *reinterpret_cast<float *>(p0) = v0;
    *reinterpret_cast<float *>(p1) = v1;
    *reinterpret_cast<float *>(p0 + kALo) = l0;
    *reinterpret_cast<float *>(p1 + kALo) = l1;
}
template <int kSubB>
__device__ __forceinline__ void stage_b2(const StageCtx &s, int row, float v0, float v1)
{
    unsigned char *p0 = s.sub + kSubB + row * 128 + s.coff[row & 7];
    unsigned char *p1 = s.sub + kSubB + (row + 1) * 128 + s.coff[(row + 1) & 7];
    float l0, l1;
    tf32_lo2(v0, v1, l0, l1);
    *reinterpret_cast<float *>(p0) = v0;
    *reinterpret_cast<float *>(p1) = v1;
    *reinterpret_cast<float *>(p0 + kBLo) = l0;
    *reinterpret_cast<float *>(p1 + kBLo) = l1;
}
template <int kALo>
__device__ __forceinline__ void stage_a(const StageCtx &s, int row, float v)
{
    unsigned char *p = s.sub + row * 128 + s.coff[row & 7];
    *reinterpret_cast<float *>(p) = v;
    *reinterpret_cast<float *>(p + kALo) = tf32_lo(v);
}
template <int kSubB>
__device__ __forceinline__ void stage_b(const StageCtx &s, int row, float v)
{
    unsigned char *p = s.sub + kSubB + row * 128 + s.coff[row & 7];
    *reinterpret_cast<float *>(p) = v;
    *reinterpret_cast<float *>(p + kBLo) = tf32_lo(v);
}
// rows row0 .. row0+31 of the A image <- 32 TMEM columns of this thread's lane (raw fp32 words)
// LOMODE 0: remainder rows computed and stored; 1: skipped (bf16 features: they are zero and were cleared once);
// 2: zeros stored (bf16 features in a buffer whose remainder rows another job overwrites)
template <int kALo, int LOMODE = 0>
__device__ __forceinline__ void stage_a_from_tmem32(const StageCtx &s, int row0, uint32_t taddr)
{
#pragma unroll
    for (int c0 = 0; c0 < 32; c0 += 16) {
        uint32_t r[16];
        tmem_ld<16>(taddr + c0, r);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
            if constexpr (LOMODE == 0) {
                stage_a2<kALo>(s, row0 + c0 + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]));
            } else {
                const int row = row0 + c0 + j;
                unsigned char *p0 = s.sub + row * 128 + s.coff[row & 7];
                unsigned char *p1 = s.sub + (row + 1) * 128 + s.coff[(row + 1) & 7];
                *reinterpret_cast<uint32_t *>(p0) = r[j];
                *reinterpret_cast<uint32_t *>(p1) = r[j + 1];
                if constexpr (LOMODE == 2) {
                    *reinterpret_cast<uint32_t *>(p0 + kALo) = 0u;
                    *reinterpret_cast<uint32_t *>(p1 + kALo) = 0u;
                }
            }
        }
    }
}

// wait until `n` completions of `bar` have been observed by this thread
__device__ __forceinline__ void wait_count(uint64_t *bar, uint32_t &seen, uint32_t n)
{
    while (seen < n) {
        mbar_wait(bar, seen & 1);
        ++seen;
    }
}

// TMEM part of a drain with TMEM-resident master sums: master[job] += acc[job][0:32) + acc[job][32:64), acc = 0.
// Out of line on purpose: the tile loop's body is one long unrolled schedule, and this runs once in several iterations.
#ifndef BTS_DRAIN_NOINLINE
#define BTS_DRAIN_NOINLINE 0
#endif
#if BTS_DRAIN_NOINLINE
#define BTS_DRAIN_ATTR __noinline__
#else
#define BTS_DRAIN_ATTR __forceinline__
#endif
template <int NJOBS>
__device__ BTS_DRAIN_ATTR void drain_to_tmem_master(uint32_t la, uint32_t tmaster)
{
#pragma unroll
    for (int job = 0; job < NJOBS; ++job) {
        uint32_t r0[32], r1[32], m[32];
        tmem_ld<32>(la + job * 64, r0);
        tmem_ld<32>(la + job * 64 + 32, r1);
        tmem_ld<32>(tmaster + job * 32, m);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            m[j] = __float_as_uint(__uint_as_float(m[j]) + (__uint_as_float(r0[j]) + __uint_as_float(r1[j])));
            r0[j] = 0u;
        }
        tmem_st<32>(tmaster + job * 32, m);
        tmem_st<32>(la + job * 64, r0);
        tmem_st<32>(la + job * 64 + 32, r0);
    }
    wait_st();
}

// MODE 0: sigmoid head (grad [B][1][hw]); 1: plane parameters (grad [B][4][hw]); 2: fused LPG (grad_depth, upratio R)
template <class CFG, int MODE, int R, int NWG, bool BF>
__global__ void __launch_bounds__(fused_threads(NWG), 1) chain_bwd_fused_kernel(const __grid_constant__ FusedArgs fa)
{
    using L = FusedLayout<CFG, MODE, R, NWG, BF>;
    using TIO = std::conditional_t<BF, unsigned short, float>;   // storage type of x, grad and grad_x (bf16 twin)
    constexpr int PXG = BF ? 8 : 4, NGRP = kTile / PXG;
    constexpr int LAST = L::LAST, NW = L::NW, K0 = L::K0;
    constexpr int NT = fused_threads(NWG);                         // NWG compute warpgroups (+ NWG issuer warps)
    constexpr bool GPRE = MODE != 2 || R <= 2;
    // head with >= 168 registers: unfenced tails (measured 228 -> 207 us). The 2 x 2 stage is faster with scalar,
    // fenced tails even at 255 registers (141 vs 147 us).
    constexpr bool RICH = NW == 1 && (kInlineIssue ? NWG <= 2 : L::PRIV);
    const TcArgs &a = fa.a;

    extern __shared__ unsigned char smem_raw[];
    // round the staging area up to the 1024-byte alignment of a SWIZZLE_128B atom (offset arithmetic on
    // the shared array keeps the address space known to the compiler)
    unsigned char *stage = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float *xs_all = reinterpret_cast<float *>(stage + L::stage_bytes);
    float *sw = reinterpret_cast<float *>(stage + L::stage_bytes + L::xs_bytes);
    float *swt = sw + L::PACK;
    uint64_t *bars = reinterpret_cast<uint64_t *>(swt + L::IMG);
    uint32_t *slot = reinterpret_cast<uint32_t *>(bars + 6 * NWG);

    const int tid = threadIdx.x;
    // weight images, barriers, TMEM and zeroed weight-gradient accumulators. With inline issue the compute threads
    // run it AFTER their first tile's copies have taken off, so the DRAM round trip hides behind the packing.
    uint32_t tbase = 0;
    auto prologue = [&]() {
        if constexpr (BF) {                                       // remainder rows of the bf16 features stay zero for the whole kernel
            for (size_t i = (size_t)tid * 16; i < L::stage_bytes; i += (size_t)NT * 16)
                *reinterpret_cast<uint4 *>(stage + i) = make_uint4(0u, 0u, 0u, 0u);
        }
        tc_pack_to_smem<CFG>(a.W, sw, LAST, tid, NT);
        tc_pack_t_to_smem<CFG>(a.W, swt, tid, NT);
        fence_smem_to_async();
        if (tid < 32) tmem_alloc(slot, 512);
        if (tid == 0) {
            for (int g = 0; g < NWG; ++g) {
                mbar_init(bars + 6 * g, kTile / 32);                  // one arrival per warp
                mbar_init(bars + 6 * g + 1, 1);
                mbar_init(bars + 6 * g + 2, kTile / 64);
                mbar_init(bars + 6 * g + 3, kTile / 64);
                mbar_init(bars + 6 * g + 4, 1);
                mbar_init(bars + 6 * g + 5, 1);
            }
            fence_mbar_init();
        }
        fence_before_sync();
        __syncthreads();
        fence_after_sync();
        tbase = (uint32_t)uniform((int)*slot);
        if (tid < kTile) {   // weight-gradient accumulators start at zero; every MMA into them accumulates
            uint32_t z[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) z[j] = 0u;
#pragma unroll
            for (int c0 = 0; c0 < L::WG0; c0 += 32) tmem_st<32>(tbase + ((uint32_t)((tid >> 5) * 32) << 16) + c0, z);
            if constexpr (L::TMASTER) {
#pragma unroll
                for (int c0 = 0; c0 < NWG * L::NJOBS * 32; c0 += 32)
                    tmem_st<32>(tbase + ((uint32_t)((tid >> 5) * 32) << 16) + L::MASTER0 + c0, z);
            }
            wait_st();
        }
        fence_before_sync();
        __syncthreads();
        fence_after_sync();
    };

    // 32-bit index arithmetic throughout (the C ABI rejects more than 2^31 - 1 low-res pixels)
    const unsigned npx = (unsigned)a.npx, hw = (unsigned)a.hw;
    const int ntiles = (int)((npx + kTile - 1) / kTile);
    const int stride = (int)gridDim.x * NWG;

    if constexpr (!kInlineIssue) prologue();
    if (!kInlineIssue && tid >= NWG * kTile) {
        if constexpr (!kInlineIssue)
            fused_issuer<CFG, MODE, R, NWG>(uniform((tid - NWG * kTile) >> 5), sw, swt, smem_u32(stage), bars, tbase, ntiles);
    } else {
        // g, wj, half are warp-uniform; routed through a shuffle once so that the compiler knows it too (the MMA
        // operands derived from them must sit in uniform registers)
        const int g = uniform(tid / kTile), t = tid % kTile, lane = tid & 31, wj = uniform((tid % kTile) >> 5), half = wj >> 1;
        unsigned char *xs_raw = reinterpret_cast<unsigned char *>(xs_all) + (size_t)g * L::xs_wg_bytes;
        TIO *xs_wg = reinterpret_cast<TIO *>(xs_raw);             // [channel][pixel] of the staged tile
        TIO *xs = xs_wg + t;                                      // this thread's pixel column
        float *gs = reinterpret_cast<float *>(xs_raw + (size_t)K0 * kTile * sizeof(TIO)) + t * 4;   // this pixel's gradient words
        const TIO *const xg = reinterpret_cast<const TIO *>(a.x);
        const TIO *const gg_in = reinterpret_cast<const TIO *>(a.grad);
        // (image, pixel) of this thread's own pixel and of the 4-pixel group it copies, for the tile the NEXT prefetch
        // fetches. They advance by the same number of pixels every iteration, so the two divisions by hw per
        // position and tile become one division per kernel (dq, dr) and an add with a wrap-around.
        struct PixPos {
            unsigned g;
            int b, p;
        };
        const unsigned step = (unsigned)stride * kTile;
        const unsigned dq = step / hw, dr = step - dq * hw;
        auto pos_at = [&](unsigned g0) {
            PixPos s;
            s.g = g0;
            s.b = g0 < npx ? (int)(g0 / hw) : 0;
            s.p = g0 < npx ? (int)(g0 - (unsigned)s.b * hw) : 0;
            return s;
        };
        auto pos_advance = [&](PixPos &s) {
            s.g += step;
            s.b += (int)dq;
            s.p += (int)dr;
            if ((unsigned)s.p >= hw) {
                s.p -= (int)hw;
                ++s.b;
            }
        };
        const int tile0 = (int)blockIdx.x * NWG + g;
        PixPos own = pos_at((unsigned)tile0 * kTile + t), quad = pos_at((unsigned)tile0 * kTile + PXG * (t & (NGRP - 1)));
        [[maybe_unused]] float gnext[4] = {0.f, 0.f, 0.f, 0.f};   // bf16 twin: the next tile's gradient words
        auto prefetch = [&](int tl) {
            const unsigned gp = own.g;
            const bool ok = tl < ntiles && gp < npx;
            [[maybe_unused]] const int b = ok ? own.b : 0;
            [[maybe_unused]] const int p = ok ? own.p : 0;
            {                                                     // 128-bit copies (host checks x_vec16), see chain_tc_fwd.cu
                constexpr int CWN = kTile / NGRP;                 // channels per pass of the warpgroup
                const int q = t & (NGRP - 1), cw = t / NGRP;
                const bool okv = tl < ntiles && quad.g < npx;
                const int bn = okv ? quad.b : 0;
                const int pn = okv ? quad.p : 0;
                const TIO *xq = xg + ((size_t)bn * K0 + cw) * a.hw + pn;
                const uint32_t dst = smem_u32(xs_wg + cw * kTile + PXG * q);
#pragma unroll
                for (int i = 0; i < K0 / CWN; ++i) cp_async16(dst + i * CWN * kTile * (int)sizeof(TIO), xq + (size_t)i * CWN * a.hw, okv);
            }
            if constexpr (GPRE && !BF) {
                const uint32_t gd = smem_u32(gs);
                if constexpr (MODE == 0) {
                    cp_async4(gd, a.grad + (ok ? gp : 0), ok);
                } else if constexpr (MODE == 2 && R == 1) {
                    cp_async4(gd, a.grad + (size_t)b * a.grad_bstride + p, ok);
                } else if constexpr (MODE == 1) {
                    const float *gq = a.grad + (size_t)b * 4 * a.hw + p;
#pragma unroll
                    for (int j = 0; j < 4; ++j) cp_async4(gd + 4 * j, gq + (size_t)j * a.hw, ok);
                } else {                                          // R == 2: the 2 x 2 block of grad_depth
                    const int i = p / a.w, j = p - i * a.w;
                    const int Wd = a.w * 2;
                    const float *g0 = a.grad + (size_t)b * a.grad_bstride + (size_t)i * 2 * (size_t)Wd + (size_t)j * 2;
                    cp_async8(gd, g0, ok);
                    cp_async8(gd + 8, g0 + Wd, ok);
                }
            }
            if constexpr (GPRE && BF) {
                // bf16 gradients are 2-byte elements, below cp.async's minimum: plain loads one tile ahead, parked in the
                // pixel's gradient words (the loads are in flight while the current tile computes)
                float g4[4] = {0.f, 0.f, 0.f, 0.f};
                if (ok) {
                    if constexpr (MODE == 0) {
                        g4[0] = io_load<TIO>(gg_in + gp);
                    } else if constexpr (MODE == 2 && R == 1) {
                        g4[0] = io_load<TIO>(gg_in + (size_t)b * a.grad_bstride + p);
                    } else if constexpr (MODE == 1) {
                        const TIO *gq = gg_in + (size_t)b * 4 * a.hw + p;
#pragma unroll
                        for (int j = 0; j < 4; ++j) g4[j] = io_load<TIO>(gq + (size_t)j * a.hw);
                    } else {
                        const int i = p / a.w, j = p - i * a.w;
                        const int Wd = a.w * 2;
                        const TIO *g0 = gg_in + (size_t)b * a.grad_bstride + (size_t)i * 2 * (size_t)Wd + (size_t)j * 2;
                        const uint32_t r0 = __ldg(reinterpret_cast<const uint32_t *>(g0)), r1 = __ldg(reinterpret_cast<const uint32_t *>(g0 + Wd));
                        g4[0] = bf16_to_f32((unsigned short)(r0 & 0xffffu)); g4[1] = bf16_to_f32((unsigned short)(r0 >> 16));
                        g4[2] = bf16_to_f32((unsigned short)(r1 & 0xffffu)); g4[3] = bf16_to_f32((unsigned short)(r1 >> 16));
                    }
                }
                gnext[0] = g4[0]; gnext[1] = g4[1]; gnext[2] = g4[2]; gnext[3] = g4[3];
            }
            cp_async_commit();
        };

        prefetch(tile0);
        if constexpr (kInlineIssue) prologue();
        uint64_t *ready = bars + 6 * g, *done = ready + 1, *stg = ready + 2 + half;
        // inline issue: hardware barrier ids of this warpgroup, warpgroup-level TMEM / staging addresses, hand-off
        const uint32_t bid = 1 + 5 * g;                           // +0 tile, +1 / +2 chain hand-offs (alternating), +3 / +4 pairs
        const bool lead_wg = wj == 0, lead_pair = (wj & 1) == 0;
        const uint32_t wbu = tbase + L::WG0 + g * L::WGCOLS, xlu = wbu + K0, ddu = wbu + 2 * K0;
        const uint32_t tacc = tbase + (L::PERWG ? g * L::ACC : 0);   // this warpgroup's (or the CTA's) accumulators
        const uint32_t subu = smem_u32(stage) + g * L::NSUB * L::SUB + ((L::PRIV && half) ? 2 * L::SUB : 0);
        uint64_t *wdone = ready + 4 + half;                       // completion of this pair's weight-gradient MMAs
        // the first warp of the group waits for the others (who only signal) and issues; with dedicated issuer
        // warps the hand-off is an mbarrier arrival instead
        auto handoff = [&](uint32_t id, uint32_t nthr, bool lead, uint64_t *mbar, auto &&issue) {
            if constexpr (kInlineIssue) {
                if (lead) {
                    named_bar_sync(id, nthr);
                    fence_after_sync();
                    if (elect_one()) issue();
                    __syncwarp();
                } else {
                    named_bar_arrive(id, nthr);
                }
            } else {
                __syncwarp();
                if (lane == 0) mbar_arrive(mbar);
            }
        };
        auto issue_wt = [&](uint32_t acc) {
            issue_wgrad_half<L::SUB, L::SUBB>(acc, subu);
            mma_commit(wdone);
        };
        // shared staging: the OTHER pair's completions free the buffer for me; private: my own
        uint64_t *wfree = ready + 4 + (L::PRIV ? half : (half ^ 1));
        const uint32_t wlag = L::PRIV ? 0u : (uint32_t)half;
        const uint32_t xh = tbase + L::WG0 + g * L::WGCOLS + ((uint32_t)(wj * 32) << 16);
        const uint32_t xl = xh + K0, dd = xh + 2 * K0;
        StageCtx sc;
        sc.sub = stage + (size_t)(g * L::NSUB + (L::PRIV ? wj : (wj & 1))) * L::SUB;
#pragma unroll
        for (int q = 0; q < 8; ++q) sc.coff[q] = (uint32_t)((((lane >> 2) ^ q) << 4) | ((lane & 3) << 2));
        const float *tw = sw + CFG::tail_off();
        const float *w5 = tw + kTailMid * kTailIn;
        uint32_t done_seen = 0, wseen = 0, nstage = 0;
        auto warp_arrive = [&](uint64_t *bar) {                   // the warp's writes are ordered before lane 0's arrival
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };
        auto done_wait = [&]() {
            mbar_wait(done, done_seen & 1);
            ++done_seen;
            fence_after_sync();
        };

        // ---- drain of the weight-gradient accumulators (see kWgradDrain). Columns j and j + 32 of a job (hi and remainder
        // halves of the deltas) are summed on the way. Per-warpgroup sets: every thread takes its TMEM lane t, all 32
        // summed columns, behind the warpgroup's own barrier. Shared set: warpgroup 0 takes j = 0..15, warpgroup 1
        // j = 16..31 (both can read every TMEM lane quarter), behind a CTA-wide barrier.
        static_assert(NWG >= 2, "the shared-set drain splits the accumulator columns over two warpgroups");
        constexpr int DRC = L::DRC;
        float wacc[L::TMASTER ? 1 : L::NJOBS][L::TMASTER ? 1 : DRC];
        if constexpr (!L::TMASTER) {
#pragma unroll
            for (int job = 0; job < L::NJOBS; ++job)
#pragma unroll
                for (int j = 0; j < DRC; ++j) wacc[job][j] = 0.f;
        }
        const uint32_t tmaster = tbase + ((uint32_t)(wj * 32) << 16) + L::MASTER0 + g * L::NJOBS * 32;
        auto drain_sync = [&]() {
            if constexpr (L::PERWG) named_bar_sync(kInlineIssue ? bid : 1 + g, kTile);
            else named_bar_sync(0, NWG * kTile);
        };
        auto drain = [&]() {
            wait_count(wfree, wseen, nstage);                     // every weight-gradient MMA this warp can vouch for is done
            fence_before_sync();
            drain_sync();                                         // ... and so is everybody else's
            fence_after_sync();
            if constexpr (L::TMASTER) {
                drain_to_tmem_master<L::NJOBS>(tacc + ((uint32_t)(wj * 32) << 16), tmaster);
            } else if (L::PERWG || g < 2) {
                const uint32_t la = tacc + ((uint32_t)(wj * 32) << 16) + (L::PERWG ? 0u : (uint32_t)g * 16);
#pragma unroll
                for (int job = 0; job < L::NJOBS; ++job) {
                    uint32_t r0[DRC], r1[DRC];
                    tmem_ld<DRC>(la + job * 64, r0);
                    tmem_ld<DRC>(la + job * 64 + 32, r1);
                    if constexpr (L::TMASTER) {
                        uint32_t m[DRC];
                        tmem_ld<DRC>(tmaster + job * 32, m);
                        wait_ld();
#pragma unroll
                        for (int j = 0; j < DRC; ++j) {
                            m[j] = __float_as_uint(__uint_as_float(m[j]) + (__uint_as_float(r0[j]) + __uint_as_float(r1[j])));
                            r0[j] = 0u;
                        }
                        tmem_st<DRC>(tmaster + job * 32, m);
                    } else {
                        wait_ld();
#pragma unroll
                        for (int j = 0; j < DRC; ++j) {
                            wacc[job][j] += __uint_as_float(r0[j]) + __uint_as_float(r1[j]);
                            r0[j] = 0u;
                        }
                    }
                    tmem_st<DRC>(la + job * 64, r0);
                    tmem_st<DRC>(la + job * 64 + 32, r0);
                }
                wait_st();
            }
            fence_before_sync();
            drain_sync();                                         // accumulators are zero again before anybody issues into them
            fence_after_sync();
        };
        // CTA-uniform trip count (warpgroup 0 has the most tiles): every warpgroup joins every drain
        const int iters_cta = (ntiles - (int)blockIdx.x * NWG + stride - 1) / stride;   // warpgroup 0's tiles: the most
        bool drain_due = false;
        int tile = tile0, since_drain = 0, drains_done = 0;
        // The plain loop over my own tiles (a CTA-uniform trip count with the body under `if (tile < ntiles)` compiled
        // to a 3 % slower kernel). Shared accumulator set: every warpgroup must join every CTA-wide drain, so the one
        // with fewer tiles catches up after its loop.
        for (; tile < ntiles; tile += stride) {
          {
            const unsigned gp = own.g;                            // == tile * kTile + t
            const bool valid = gp < npx;
            const int b = valid ? own.b : 0;
            const int p = valid ? own.p : 0;
            pos_advance(own);                                     // own / quad now describe tile + stride
            pos_advance(quad);
            // ---- x -> (hi, lo) -> TMEM A operand of layer 0; then the next tile's loads take off
            constexpr bool XEARLY = BTS_STAGE_X_EARLY && NW == 1 && L::PRIV;
            cp_async_wait_all();
            named_bar_sync(kInlineIssue ? bid : 1 + g, kTile);     // the tile was copied by all threads of the warpgroup
            if constexpr (XEARLY) wait_count(wfree, wseen, nstage + wlag);   // my pair's last weight-gradient MMAs have read the buffer
#pragma unroll
            for (int c0 = 0; c0 < K0; c0 += 16) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    float v0, v1;
                    if constexpr (BF) {                           // bf16 is exact in tf32: the remainder is zero
                        v0 = bf16_to_f32(xs[(c0 + j) * kTile]);
                        v1 = bf16_to_f32(xs[(c0 + j + 1) * kTile]);
                        lo[j] = lo[j + 1] = 0u;
                    } else {
                        v0 = xs[(c0 + j) * kTile];
                        v1 = xs[(c0 + j + 1) * kTile];
                        tf32_lo2(v0, v1, lo[j], lo[j + 1]);
                    }
                    hi[j] = __float_as_uint(v0);
                    hi[j + 1] = __float_as_uint(v1);
                    if constexpr (XEARLY) {                       // rows c0 + j, c0 + j + 1 of the A image (+ remainder rows)
                        const int row = c0 + j;
                        unsigned char *p0 = sc.sub + row * 128 + sc.coff[row & 7];
                        unsigned char *p1 = sc.sub + (row + 1) * 128 + sc.coff[(row + 1) & 7];
                        *reinterpret_cast<float *>(p0) = v0;
                        *reinterpret_cast<float *>(p1) = v1;
                        if constexpr (!BF) {
                            *reinterpret_cast<uint32_t *>(p0 + L::ALO) = lo[j];
                            *reinterpret_cast<uint32_t *>(p1 + L::ALO) = lo[j + 1];
                        }
                    }
                }
                tmem_st<16>(xh + c0, hi);
                if constexpr (!BF) tmem_st<16>(xl + c0, lo);      // (bf16: layer 0 skips the remainder pass)
            }
            float gcur[4] = {0.f, 0.f, 0.f, 0.f};
            if constexpr (GPRE && !BF) {
                const float4 g4 = *reinterpret_cast<const float4 *>(gs);
                gcur[0] = g4.x; gcur[1] = g4.y; gcur[2] = g4.z; gcur[3] = g4.w;
            }
            if constexpr (GPRE && BF) {
                gcur[0] = gnext[0]; gcur[1] = gnext[1]; gcur[2] = gnext[2]; gcur[3] = gnext[3];
            }
            wait_st();
            fence_before_sync();
            if constexpr (!kInlineIssue) warp_arrive(ready);      // F0
            named_bar_sync(kInlineIssue ? bid : 1 + g, kTile);    // everybody has read the staged tile (and written its A operand)
            if constexpr (kInlineIssue) {
                if (lead_wg) {                                    // F0
                    fence_after_sync();
                    if (elect_one()) {
                        issue_ts3<CFG::K(0), CFG::N(0), BF>(ddu, wbu, xlu, sw + CFG::w_off(0));
                        mma_commit(done);
                    }
                    __syncwarp();
                }
            }
            prefetch(tile + stride);
            done_wait();
            float a1[NW == 2 ? 32 : 1];
            if constexpr (NW == 2) {                              // a1 = ELU(D) -> A operand of layer 1
#pragma unroll
                for (int n0 = 0; n0 < 32; n0 += 16) {
                    uint32_t r[16], lo[16];
                    tmem_ld<16>(dd + n0, r);
                    wait_ld();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        a1[n0 + j] = elu_fast(__uint_as_float(r[j]));
                        r[j] = __float_as_uint(a1[n0 + j]);
                        lo[j] = __float_as_uint(tf32_lo(a1[n0 + j]));
                    }
                    tmem_st<16>(xl + n0, r);
                    tmem_st<16>(xl + 32 + n0, lo);
                }
                wait_st();
                fence_before_sync();
                handoff(bid + 1, kTile, lead_wg, ready, [&]() {   // F1
                    issue_ts3<CFG::K(1), CFG::N(1)>(ddu, xlu, xlu + 32, sw + CFG::w_off(1));
                    mma_commit(done);
                });
                done_wait();
            }
            float a16[kTailIn], a8[kTailMid], z[LAST];
            {
                uint32_t r[16];
                tmem_ld<16>(dd, r);
                wait_ld();
#pragma unroll
                for (int j = 0; j < 16; ++j) a16[j] = elu_fast(__uint_as_float(r[j]));
            }
            tail_forward<LAST, NW == 1 || RICH, !RICH>(tw, a16, a8, z);   // registers to spare: packed FMAs, weight loads float
            // ---- upstream gradient -> dz
            float dz[3] = {0.f, 0.f, 0.f};
            if constexpr (MODE == 0) {
                const float s = sigmoidf(z[0]);
                dz[0] = gcur[0] * s * (1.f - s);
            } else {
                const PlaneFwd pl = plane_from_logits(z[0], z[1], z[2], a.max_depth);
                if (valid) {
                    float gg[4];
                    if constexpr (MODE == 1) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) gg[j] = gcur[j];
                    } else if constexpr (R <= 2) {
                        LpgAcc acc{0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (int pp = 0; pp < R; ++pp) {
                            const float v = lpg_offset(pp, R);
                            const float n2v = __fmul_rn(pl.n2, v);
#pragma unroll
                            for (int k = 0; k < R; ++k) {
                                const float u = lpg_offset(k, R);
                                lpg_bwd_accum(acc, gcur[pp * R + k], __fmul_rn(pl.n1, u), n2v, pl.n3, u, v, a.eps);
                            }
                        }
                        const float sc = a.depth_div != 1.f ? __fdiv_rn(1.f, a.depth_div) : 1.f;
                        gg[0] = -pl.n4 * acc.tu * sc; gg[1] = -pl.n4 * acc.tv * sc; gg[2] = -pl.n4 * acc.t * sc; gg[3] = acc.s4 * sc;
                    } else {
                        lpg_block_grad<R>(a, pl, b, p, gg);
                    }
                    plane_bwd(pl, gg[0], gg[1], gg[2], gg[3], a.max_depth, dz);
                }
            }
            // ---- tail backward: d8 = (W5^T dz) * ELU'(a8);  dl16 = (W4^T d8) * ELU'(a16)
            float d8[kTailMid], dl16[kTailIn];
#pragma unroll
            for (int o = 0; o < kTailMid; ++o) {
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < LAST; ++j) s = fmaf(w5[j * kTailMid + o], dz[j], s);
                d8[o] = s * elu_grad_from_out(a8[o]);
            }
            tail_dgrad16<NW == 1 || RICH, RICH ? 4 : 1>(tw, d8, dl16);
            {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    dl16[j] *= elu_grad_from_out(a16[j]);
                    dl16[j + 1] *= elu_grad_from_out(a16[j + 1]);
                    hi[j] = __float_as_uint(dl16[j]);
                    hi[j + 1] = __float_as_uint(dl16[j + 1]);
                    tf32_lo2(dl16[j], dl16[j + 1], lo[j], lo[j + 1]);
                }
                constexpr int OFF = NW == 1 ? 0 : 32;
                tmem_st<16>(xl + OFF, hi);
                tmem_st<16>(xl + OFF + 16, lo);
            }
            wait_st();
            fence_before_sync();
            handoff(bid + 2, kTile, lead_wg, ready, [&]() {       // DG(NW-1)
                if constexpr (NW == 1) issue_ts3<CFG::N(0), CFG::K(0)>(ddu, xlu, xlu + 16, swt + CFG::w_off(0));
                else issue_ts3<CFG::N(1), CFG::K(1)>(ddu, xlu + 32, xlu + 48, swt + CFG::w_off(1));
                mma_commit(done);
            });
            // A due drain runs HERE, one iteration late: the previous tile's weight-gradient MMAs were issued more
            // than a microsecond ago (waiting for them at the end of their own iteration cost ~0.7 us per drain), the
            // dgrad MMAs issued above run underneath, and nobody issues into the accumulators before stage_tail.
            if (kDrainDeferred && drain_due) {
                drain();
                drain_due = false;
            }

            auto stage_tail = [&]() {
                wait_count(wfree, wseen, nstage + wlag);
                ++nstage;
                if constexpr (NW == 2) {
#pragma unroll
                    for (int j = 0; j < 32; j += 2) stage_a2<L::ALO>(sc, j, a1[j], a1[j + 1]);
                }
#pragma unroll
                for (int j = 0; j < kTailIn; j += 2) stage_a2<L::ALO>(sc, 32 + j, a16[j], a16[j + 1]);
#pragma unroll
                for (int j = 0; j < kTailMid; j += 2) stage_a2<L::ALO>(sc, 48 + j, a8[j], a8[j + 1]);
#pragma unroll
                for (int j = 0; j < kTailIn; j += 2) stage_b2<L::SUBB>(sc, j, dl16[j], dl16[j + 1]);
#pragma unroll
                for (int j = 0; j < kTailMid; j += 2) stage_b2<L::SUBB>(sc, 16 + j, d8[j], d8[j + 1]);
#pragma unroll
                for (int j = 0; j < LAST; ++j) stage_b<L::SUBB>(sc, 24 + j, dz[j]);
                if constexpr (NW == 1 && !XEARLY) stage_a_from_tmem32<L::ALO, (BF ? (L::PRIV ? 1 : 2) : 0)>(sc, 0, xh);
                fence_smem_to_async();
                handoff(bid + 3 + half, 64, lead_pair, stg, [&]() { issue_wt(tacc); });           // WT(half)
            };
            auto gradx_epilogue = [&]() {
                done_wait();
                TIO *gx = reinterpret_cast<TIO *>(a.grad_x) + (size_t)b * K0 * a.hw + p;
#pragma unroll
                for (int n0 = 0; n0 < K0; n0 += 16) {
                    uint32_t r[16];
                    tmem_ld<16>(dd + n0, r);
                    wait_ld();
                    if (valid && a.grad_x != nullptr) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) io_store(gx + (size_t)(n0 + j) * a.hw, __uint_as_float(r[j]));
                    }
                }
            };
            if constexpr (NW == 1 && L::PRIV) {
                // private staging buffers: nobody takes turns, so every warp stages first and the dgrad MMAs
                // (issued when the last warp arrived above) complete under the staging stores
                stage_tail();
                gradx_epilogue();
            } else if constexpr (NW == 1) {
#pragma unroll 1
                for (int ph = 0; ph < 2; ++ph) {                  // pair 0 stages first, pair 1 second
                    if (ph == half) stage_tail();
                    else gradx_epilogue();
                }
            } else {
                // delta_1 lives only in TMEM (hi words at xl[0:32) until the next tile): not in registers
                auto dg1_epilogue = [&]() {                       // delta_1 = (delta_16 W1) * ELU'(a1) -> A operand of DG0
                    done_wait();
#pragma unroll
                    for (int n0 = 0; n0 < 32; n0 += 16) {
                        uint32_t r[16], lo[16];
                        tmem_ld<16>(dd + n0, r);
                        wait_ld();
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float v = __uint_as_float(r[j]) * elu_grad_from_out(a1[n0 + j]);
                            r[j] = __float_as_uint(v);
                            lo[j] = __float_as_uint(tf32_lo(v));
                        }
                        tmem_st<16>(xl + n0, r);
                        tmem_st<16>(xl + 32 + n0, lo);
                    }
                    wait_st();
                    fence_before_sync();
                    handoff(bid + 1, kTile, lead_wg, ready, [&]() {   // DG0
                        issue_ts3<CFG::N(0), CFG::K(0)>(ddu, xlu, xlu + 32, swt + CFG::w_off(0));
                        mma_commit(done);
                    });
                };
                auto stage_x = [&]() {
                    wait_count(wfree, wseen, nstage + wlag);
                    ++nstage;
#pragma unroll
                    for (int n0 = 0; n0 < 32; n0 += 16) {
                        uint32_t r[16];
                        tmem_ld<16>(xl + n0, r);
                        wait_ld();
#pragma unroll
                        for (int j = 0; j < 16; j += 2) stage_b2<L::SUBB>(sc, n0 + j, __uint_as_float(r[j]), __uint_as_float(r[j + 1]));
                    }
                    stage_a_from_tmem32<L::ALO, (BF ? 2 : 0)>(sc, 0, xh);
                    stage_a_from_tmem32<L::ALO, (BF ? 2 : 0)>(sc, 32, xh + 32);
                    fence_smem_to_async();
                    handoff(bid + 3 + half, 64, lead_pair, stg, [&]() { issue_wt(tacc + 64); });    // WX(half)
                };
#pragma unroll 1
                for (int ph = 0; ph < 2; ++ph) {
                    if (ph == half) stage_tail();
                    else dg1_epilogue();
                }
#pragma unroll 1
                for (int ph = 0; ph < 2; ++ph) {
                    if (ph == half) stage_x();
                    else gradx_epilogue();
                }
            }
          }
          if (++since_drain == fa.drain_every) {
            since_drain = 0;
            ++drains_done;
            if (kDrainDeferred) drain_due = true;
            else drain();
          }
        }
        if (kDrainDeferred && drain_due) drain();
        if constexpr (!L::PERWG) {
            for (; drains_done < iters_cta / fa.drain_every; ++drains_done) drain();   // warpgroup 0 ran that many
        }
        drain();                                                  // whatever the last iterations accumulated
        // ---- drained sums -> shared memory (the staging buffers are idle: the last drain waited for every MMA).
        // Row m of a job sat in TMEM lane m; the hi rows and the remainder rows of the same channel are summed below.
        if (L::PERWG || g < 2) {
            // per-warpgroup sets: one [NJOBS][128][33] block per warpgroup, summed by the extraction below
            // (a warpgroup may get here while another one still stages: each writes into its OWN idle staging buffers)
            float *scr = reinterpret_cast<float *>(stage + (L::PERWG ? (size_t)g * L::NSUB * L::SUB : 0));
#pragma unroll
            for (int job = 0; job < L::NJOBS; ++job) {
                if constexpr (L::TMASTER) {
                    uint32_t m[DRC];
                    tmem_ld<DRC>(tmaster + job * 32, m);
                    wait_ld();
#pragma unroll
                    for (int j = 0; j < DRC; ++j) scr[(job * kTile + t) * 33 + j] = __uint_as_float(m[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < DRC; ++j) scr[(job * kTile + t) * 33 + (L::PERWG ? 0 : g * 16) + j] = wacc[job][j];
                }
            }
        }
    }
    constexpr int SP = 33;
    constexpr int SCR_SETS = L::PERWG ? NWG : 1, SCR_SET = L::PERWG ? L::NSUB * L::SUB / 4 : 0;   // floats between the sets
    static_assert((size_t)L::NJOBS * kTile * SP * sizeof(float) <= (size_t)L::NSUB * L::SUB, "a set fits one warpgroup's staging");
    float *scratch = reinterpret_cast<float *>(stage);            // [SCR_SETS][NJOBS][128][SP]
    fence_before_sync();
    __syncthreads();
    if (fa.partials != nullptr || fa.atomic_w) {
        float *part = fa.partials + (size_t)blockIdx.x * fa.total_w;
        // layer l of the chain: (job, first row, first column, K = inputs, N = outputs)
        constexpr int NL = NW + 2;
        int off = 0;
#pragma unroll
        for (int l = 0; l < NL; ++l) {
            const int t_l = l - (NW - 1);                         // 0: 32->16 layer, 1: 16->8, 2: 8->LAST; <0: x round
            const int job = t_l < 0 ? 1 : 0;
            const int row0 = t_l <= 0 ? 0 : (t_l == 1 ? 32 : 48);
            const int col0 = t_l <= 0 ? 0 : (t_l == 1 ? 16 : 24);
            const int K = t_l < 0 ? K0 : (t_l == 0 ? 32 : (t_l == 1 ? kTailIn : kTailMid));
            const int N = t_l < 0 ? 32 : (t_l == 0 ? kTailIn : (t_l == 1 ? kTailMid : LAST));
            // atomic route: CTAs start at different elements so that they do not all hit the same address at once
            const int shift = fa.atomic_w ? (int)((blockIdx.x * 67u) % (unsigned)(K * N)) : 0;
            for (int i = tid; i < K * N; i += NT) {
                int e = i + shift;
                if (e >= K * N) e -= K * N;
                const int o = e / K, c = e - o * K;
                const float *s = scratch + (job * kTile + row0 + c) * SP + col0 + o;
                float v = s[0] + s[L::ALO_ROWS * SP];
#pragma unroll
                for (int q = 1; q < SCR_SETS; ++q) v += s[q * SCR_SET] + s[q * SCR_SET + L::ALO_ROWS * SP];
                if (fa.atomic_w) atomicAdd(a.gw[l] + e, v);
                else part[off + e] = v;
            }
            off += K * N;
        }
    }
    __syncthreads();
    if (tid < 32) tmem_dealloc(tbase, 512);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
int launch_wreduce(const ChainParams &cp, const float *partials, int nparts, int total_w, cudaStream_t s);
size_t tc_pack_bytes();
bool head_bwd_mma_wanted(const ChainParams &cp, const TcArgs &a, int mode);        // chain_head_mma.cu
int launch_head_bwd_mma(const ChainParams &cp, const TcArgs &a, int sms, cudaStream_t s);

template <class CFG, int MODE, int R, int NWG, bool BF>
static int launch_fused_cfg2(const ChainParams &cp, const TcArgs &a, char *ws, int sms, cudaStream_t s)
{
    using L = FusedLayout<CFG, MODE, R, NWG, BF>;
    float *partials = reinterpret_cast<float *>(ws + 2 * tc_pack_bytes());
    int total_w = 0;
    for (int l = 0; l < cp.n_layers; ++l) total_w += cp.dims[l] * cp.dims[l + 1];

    FusedArgs fa;
    fa.a = a;
    tc_set_weights(fa.a, cp);
    // weight gradients: by default every CTA adds its partial sums to grad_W with fp32 atomics (grad_W is an
    // accumulate target by contract); BTS_DETERMINISTIC=1 keeps per-CTA partials + the fixed-order reduction kernel
    static const bool deterministic = getenv("BTS_DETERMINISTIC") != nullptr && atoi(getenv("BTS_DETERMINISTIC")) != 0;
    const bool want_w = cp.gW[0] != nullptr;
    fa.atomic_w = want_w && !deterministic ? 1 : 0;
    fa.partials = want_w && deterministic ? partials : nullptr;
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) fa.a.gw[l] = l < cp.n_layers ? cp.gW[l] : nullptr;
    fa.total_w = total_w;
    static const int drain_env = getenv("BTS_WGRAD_DRAIN") ? atoi(getenv("BTS_WGRAD_DRAIN")) : 0;   // tuning hook
    fa.drain_every = drain_env > 0 ? drain_env : kWgradDrain;
    auto kern = chain_bwd_fused_kernel<CFG, MODE, R, NWG, BF>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::smem_bytes);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long ntiles = (a.npx + kTile - 1) / kTile;
    long long grid = (ntiles + NWG - 1) / NWG;
    if (grid > sms) grid = sms;
    kern<<<(unsigned)grid, fused_threads(NWG), L::smem_bytes, s>>>(fa);
    if (int e2 = check_launch("chain_bwd_fused")) return e2;
    if (fa.partials != nullptr) {
        if (int e2 = launch_wreduce(cp, partials, (int)grid, total_w, s)) return e2;
    }
    return BTS_OK;
}

template <class CFG, int MODE, int R, int NWG>
static int launch_fused_cfg(const ChainParams &cp, const TcArgs &a, char *ws, int sms, cudaStream_t s)
{
    return a.bf16_io ? launch_fused_cfg2<CFG, MODE, R, NWG, true>(cp, a, ws, sms, s) : launch_fused_cfg2<CFG, MODE, R, NWG, false>(cp, a, ws, sms, s);
}

template <class CFG, int NWG>
static int dispatch_fused(const ChainParams &cp, const TcArgs &a, int mode, int r, char *ws, int sms, cudaStream_t s)
{
    if (mode == 0) return launch_fused_cfg<CFG, 0, 1, NWG>(cp, a, ws, sms, s);
    if (mode == 1) return launch_fused_cfg<CFG, 1, 1, NWG>(cp, a, ws, sms, s);
    switch (r) {
    case 8: return launch_fused_cfg<CFG, 2, 8, NWG>(cp, a, ws, sms, s);
    case 4: return launch_fused_cfg<CFG, 2, 4, NWG>(cp, a, ws, sms, s);
    case 2: return launch_fused_cfg<CFG, 2, 2, NWG>(cp, a, ws, sms, s);
    default: return launch_fused_cfg<CFG, 2, 1, NWG>(cp, a, ws, sms, s);
    }
}

// (the caller also requires a.x_vec16: the feature prefetch moves 4 pixels per 128-bit cp.async)
bool fused_bwd_supported(const ChainParams &cp) { return cfg_matches<Cfg32_16>(cp) || cfg_matches<Cfg64_32>(cp); }

// mode: 0 head (grad [B,1,hw]), 1 plane parameters (grad [B,4,hw]), 2 fused LPG (grad_depth, upratio r)
int launch_chain_bwd_fused(const ChainParams &cp, const TcArgs &a, int mode, int r, void *ws, int sms, cudaStream_t s)
{
    char *w = static_cast<char *>(ws);
    // the sigmoid head in fp32: register-operand kernel on warp-level MMAs (chain_head_mma.cu; BTS_HEAD_MMA=0 disables)
    if (head_bwd_mma_wanted(cp, a, mode)) return launch_head_bwd_mma(cp, a, sms, s);
    if (cfg_matches<Cfg32_16>(cp)) {
        // two warpgroups with private staging buffers per warp pair (168 registers, no spills, no turn-taking between
        // the pairs) beat three warpgroups that share staging (128 registers, ~200 bytes of spills): 242 vs 248-256 us
        // at nyu_b8. BTS_HEAD_NWG=3 selects the other variant for tuning.
        static const int head_nwg = getenv("BTS_HEAD_NWG") ? atoi(getenv("BTS_HEAD_NWG")) : 2;
        if (head_nwg == 3) return dispatch_fused<Cfg32_16, 3>(cp, a, mode, r, w, sms, s);
        return dispatch_fused<Cfg32_16, 2>(cp, a, mode, r, w, sms, s);
    }
    if (cfg_matches<Cfg64_32>(cp)) return dispatch_fused<Cfg64_32, 2>(cp, a, mode, r, w, sms, s);
    return set_error(BTS_ERR_UNSUPPORTED, "chain not covered by the fused backward");
}

}  // namespace bts

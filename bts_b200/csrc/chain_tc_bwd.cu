// chain_tc_bwd.cu -- backward of a BTS decoder stage on sm_100a tensor cores.
//
// Autograd backward of reduction_1x1 (+ LPG) (pytorch/bts.py:124-173), with
// the chain recomputed from x instead of saved. Four launches:
//   K1 bwd1   : recompute the chain (3xTF32 tcgen05, as the forward), evaluate the upstream gradient
//               (sigmoid head / plane parameterisation / LPG block reduction), back-propagate through
//               the SIMT tail; activations a_l and deltas of the tail go to an L2-sized workspace
//   K2 dgrad  : delta_l = (delta_{l+1} W_l) * ELU'(a_l) through the wide layers (3xTF32 tcgen05, the
//               weights as a transposed K-major image), ending in grad_x
//   K3 wgrad  : dW_l = sum_px delta_{l+1} (x) a_l as tcgen05 SS-mode MMAs with the pixel axis as K;
//               per-CTA partial sums (deterministic, no atomics)
//   K4 reduce : grad_W += sum of partials
// Workspace rows are [channel][pixel] fp32 matrices padded to a multiple of 128 pixels.
#include <cstring>
#include <type_traits>

#include "chain_tc.cuh"

namespace bts {

struct BwdArgs {
    TcArgs a;              // x, raw weights, grad, grad_x, sizes
    float *act;            // [RA][npx_pad]   a_1 .. a_NW (=a16), a8
    float *del;            // [RD][npx_pad]   delta_1 .. delta_NW (=d16), d8, dz
    long long npx_pad;
    int hw_pad;            // pixels per image on the workspace's pixel axis (hw, or hw rounded up to 32)
    const float *w4, *w5;  // stash backward: raw tail weights [8][16], [LAST][8] (device)
};

// MODE 0: sigmoid head (grad [B][1][hw]); 1: plane parameters (grad [B][4][hw]); 2: fused LPG (grad_depth)
// AW: width of the A operand region in TMEM, as in the forward kernel (the recompute follows the forward's
// accumulation order exactly, so both see bit-identical activations)
template <class CFG, int MODE, int R, int NWG, int AW>
__global__ void __launch_bounds__(NWG * kTile, 1) chain_tc_bwd1_kernel(BwdArgs ba)
{
    constexpr bool FINAL = (MODE == 0);
    constexpr int LAST = FINAL ? 1 : 3;
    constexpr int PACK = CFG::pack_floats(LAST);
    constexpr int NMAX = CFG::nmax();
    constexpr int COLS = 2 * AW + NMAX;
    constexpr int D0 = CFG::dim(0);
    constexpr int KC0 = D0 < AW ? D0 : AW;
    static_assert(NWG * COLS <= 512, "TMEM budget");
    const TcArgs &a = ba.a;

    extern __shared__ __align__(128) float smem[];
    float *sw = smem;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + PACK);
    uint32_t *slot = reinterpret_cast<uint32_t *>(bars + NWG);

    const int tid = threadIdx.x, wg = uniform(tid / kTile), t = tid % kTile;
    const int lt = uniform(t >> 5);                               // warp index inside the warpgroup; warp 0 issues the MMAs
    tc_pack_to_smem<CFG>(a.W, sw, LAST, tid, NWG * kTile);
    fence_smem_to_async();
    if (tid < 32) tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int i = 0; i < NWG; ++i) mbar_init(&bars[i], 1);
        fence_mbar_init();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = (uint32_t)uniform((int)*slot);
    const uint32_t wg_base = tbase + wg * COLS;
    const uint32_t lane_addr = wg_base + ((uint32_t)((t >> 5) * 32) << 16);
    uint64_t *bar = &bars[wg];
    uint32_t phase = 0;
    const float *tw = sw + CFG::tail_off();
    const float *w5 = tw + kTailMid * kTailIn;
    const long long npp = ba.npx_pad;

    const long long ntiles = npp / kTile;
    for (long long tile = (long long)blockIdx.x * NWG + wg; tile < ntiles; tile += (long long)gridDim.x * NWG) {
        const long long gp = tile * kTile + t;                    // index on the (per image padded) workspace pixel axis
        const int bb = (int)((unsigned)gp / (unsigned)ba.hw_pad), pp = (int)((unsigned)gp - (unsigned)bb * (unsigned)ba.hw_pad);
        const bool valid = bb < a.B && pp < a.hw;
        const int b = valid ? bb : 0;
        const int p = valid ? pp : 0;
        const long long gpt = (long long)b * a.hw + p;           // index in the tensors
        const float *xp = a.x + (size_t)b * D0 * a.hw + p;
        float *act_gp = ba.act + gp;
        float *del_gp = ba.del + gp;
        auto load_x = [&](auto c0_tag) {
            constexpr int C0 = decltype(c0_tag)::value;
#pragma unroll 1
            for (int c0 = 0; c0 < KC0; c0 += 32) {
                uint32_t hi[32], lo[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) hi[j] = valid ? __float_as_uint(__ldg(xp + (size_t)(C0 + c0 + j) * a.hw)) : 0u;
#pragma unroll
                for (int j = 0; j < 32; ++j) lo[j] = __float_as_uint(tf32_lo(__uint_as_float(hi[j])));
                tmem_st<32>(lane_addr + c0, hi);
                tmem_st<32>(lane_addr + AW + c0, lo);
            }
        };
        load_x(std::integral_constant<int, 0>{});
        if constexpr (D0 > KC0) {
            tc_run_chunk<CFG, 0, AW, 0, KC0>(wg_base, sw, bar, phase, wg, lt);
            load_x(std::integral_constant<int, KC0>{});
        }
        tc_run_chunk<CFG, 0, AW, D0 - KC0, KC0>(wg_base, sw, bar, phase, wg, lt);
        if constexpr (CFG::NW > 1) tc_layer_from_d<CFG, 1, AW, true>(wg_base, lane_addr, sw, bar, phase, wg, lt, act_gp, npp, true);
        if constexpr (CFG::NW > 2) tc_layer_from_d<CFG, 2, AW, true>(wg_base, lane_addr, sw, bar, phase, wg, lt, act_gp, npp, true);
        if constexpr (CFG::NW > 3) tc_layer_from_d<CFG, 3, AW, true>(wg_base, lane_addr, sw, bar, phase, wg, lt, act_gp, npp, true);
        float a16[kTailIn], a8[kTailMid], z[LAST];
        {
            uint32_t r[16];
            tmem_ld<16>(lane_addr + 2 * AW, r);
            wait_ld();
            float *dst = act_gp + (size_t)BwdRows<CFG>::row(CFG::NW) * npp;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                a16[j] = elu_fast(__uint_as_float(r[j]));
                dst[(size_t)j * npp] = a16[j];
            }
        }
        tail_forward<LAST>(tw, a16, a8, z);
        {
            float *dst = act_gp + (size_t)BwdRows<CFG>::row8() * npp;
#pragma unroll
            for (int j = 0; j < kTailMid; ++j) dst[(size_t)j * npp] = a8[j];
        }
        // upstream gradient -> dz
        float dz[3] = {0.f, 0.f, 0.f};
        if constexpr (MODE == 0) {
            const float s = sigmoidf(z[0]);
            const float g = valid ? __ldcs(a.grad + gpt) : 0.f;
            dz[0] = g * s * (1.f - s);
        } else {
            const PlaneFwd pl = plane_from_logits(z[0], z[1], z[2], a.max_depth);
            float g[4] = {0.f, 0.f, 0.f, 0.f};
            if (valid) {
                if constexpr (MODE == 1) {
                    const float *gq = a.grad + (size_t)b * 4 * a.hw + p;
                    g[0] = __ldcs(gq); g[1] = __ldcs(gq + a.hw); g[2] = __ldcs(gq + (size_t)2 * a.hw); g[3] = __ldcs(gq + (size_t)3 * a.hw);
                } else {
                    lpg_block_grad<R>(a, pl, b, p, g);
                }
                plane_bwd(pl, g[0], g[1], g[2], g[3], a.max_depth, dz);
            }
        }
        // tail backward: d8 = (W5^T dz) * ELU'(a8);  d16 = (W4^T d8) * ELU'(a16)
        float d8[kTailMid], d16[kTailIn];
#pragma unroll
        for (int o = 0; o < kTailMid; ++o) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < LAST; ++j) s = fmaf(w5[j * kTailMid + o], dz[j], s);
            d8[o] = s * elu_grad_from_out(a8[o]);
        }
            tail_dgrad16(tw, d8, d16);
        {
            float *dz_dst = del_gp + (size_t)BwdRows<CFG>::rowz() * npp;
#pragma unroll
            for (int j = 0; j < LAST; ++j) dz_dst[(size_t)j * npp] = dz[j];
            float *d8_dst = del_gp + (size_t)BwdRows<CFG>::row8() * npp;
#pragma unroll
            for (int j = 0; j < kTailMid; ++j) d8_dst[(size_t)j * npp] = d8[j];
            float *d16_dst = del_gp + (size_t)BwdRows<CFG>::row(CFG::NW) * npp;
#pragma unroll
            for (int j = 0; j < kTailIn; ++j) d16_dst[(size_t)j * npp] = d16[j] * elu_grad_from_out(a16[j]);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (tid < 32) tmem_dealloc(tbase, 512);
}

// ------------------------------------------------------------------------------------------------
// K2: data gradient through the wide layers. A = delta_{l+1} (TMEM, hi/lo), B = W_l^T image
// (rows = input channel c, K' = output channel o, K-major), D = delta_{l+1} W_l  [128 px x K_l].
// ------------------------------------------------------------------------------------------------
template <class CFG>
struct DgradCfg {
    __host__ __device__ static constexpr int kmax()   // widest A operand: N_l = dim(l+1)
    {
        int m = 0;
        for (int l = 0; l < CFG::NW; ++l) m = CFG::N(l) > m ? CFG::N(l) : m;
        return m;
    }
    __host__ __device__ static constexpr int nmax() { return CFG::kmax(); }
};

// A operand region width of the dgrad kernel: 64 columns at most, so that a 128-channel chain needs
// 2*64 + 128 = 256 TMEM columns per warpgroup and two warpgroups fit (the 128-wide delta goes in two rounds)
template <class CFG>
constexpr int kDgradAW = DgradCfg<CFG>::kmax() > 64 ? 64 : DgradCfg<CFG>::kmax();

// MMAs over delta channels [KOFF, KOFF+KC) of layer L's input gradient; A columns [0,KC) hi, [AW,AW+KC) lo
template <class CFG, int L, int KOFF, int KC>
__device__ __forceinline__ void tc_issue_dgrad(uint32_t wg_base, const float *swt, uint64_t *bar)
{
    constexpr int KP = CFG::N(L), NP = CFG::K(L), AW = kDgradAW<CFG>;
    constexpr uint32_t LBO = 128, SBO = (KP / 4) * 128;
    constexpr uint32_t idesc = idesc_tf32(kTile, NP, 0, 0);
    const uint32_t whi = smem_u32(swt + CFG::w_off(L));
    const uint32_t wlo = whi + NP * KP * 4;
    const uint32_t d = wg_base + 2 * AW;
#pragma unroll
    for (int k0 = 0; k0 < KC; k0 += 8)
        mma_tf32_ts(d, wg_base + AW + k0, smem_desc(whi + ((KOFF + k0) / 4) * LBO, LBO, SBO), idesc, (KOFF + k0) > 0);
#pragma unroll
    for (int k0 = 0; k0 < KC; k0 += 8)
        mma_tf32_ts(d, wg_base + k0, smem_desc(wlo + ((KOFF + k0) / 4) * LBO, LBO, SBO), idesc, 1);
#pragma unroll
    for (int k0 = 0; k0 < KC; k0 += 8)
        mma_tf32_ts(d, wg_base + k0, smem_desc(whi + ((KOFF + k0) / 4) * LBO, LBO, SBO), idesc, 1);
    mma_commit(bar);
}

// `keep` carries the second half of a 128-wide delta from one layer's epilogue to the next layer's second
// round of MMAs (the first round overwrites the accumulator the values came from).
template <class CFG, int L>
__device__ __forceinline__ void dgrad_layer(const BwdArgs &ba, uint32_t wg_base, uint32_t lane_addr, const float *swt,
                                            uint64_t *bar, uint32_t &phase, int wg, int t, long long gp, bool valid,
                                            int b, int p, float (&keep)[64])
{
    constexpr int KP = CFG::N(L), NP = CFG::K(L), AW = kDgradAW<CFG>;
    constexpr int KC = KP < AW ? KP : AW;
    static_assert(KP == KC || KP == 2 * AW, "one or two rounds");
    const long long npp = ba.npx_pad;
    wait_st();
    fence_before_sync();
    named_bar_sync(1 + wg, kTile);
    if (t == 0) {                                                 // t: warp index inside the warpgroup (warp-uniform)
        if (elect_one()) {
            fence_after_sync();
            tc_issue_dgrad<CFG, L, 0, KC>(wg_base, swt, bar);
        }
        __syncwarp();
    }
    float av[16], nx[16];
    if constexpr (L > 0) {                                        // the first two chunks of a_l travel under the MMAs
        const float *act0 = ba.act + (size_t)BwdRows<CFG>::row(L) * npp + gp;
#pragma unroll
        for (int j = 0; j < 16; ++j) av[j] = __ldcg(act0 + (size_t)j * npp);
        if constexpr (NP > 16) {
#pragma unroll
            for (int j = 0; j < 16; ++j) nx[j] = __ldcg(act0 + (size_t)(16 + j) * npp);
        }
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    fence_after_sync();
    if constexpr (KP > KC) {                                      // second round: channels AW .. 2AW-1 of delta
#pragma unroll
        for (int n0 = 0; n0 < AW; n0 += 16) {
            uint32_t r[16], lo[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                r[j] = __float_as_uint(keep[n0 + j]);
                lo[j] = __float_as_uint(tf32_lo(keep[n0 + j]));
            }
            tmem_st<16>(lane_addr + n0, r);
            tmem_st<16>(lane_addr + AW + n0, lo);
        }
        wait_st();
        fence_before_sync();
        named_bar_sync(1 + wg, kTile);
        if (t == 0) {                                                 // t: warp index inside the warpgroup (warp-uniform)
            if (elect_one()) {
                fence_after_sync();
                tc_issue_dgrad<CFG, L, AW, AW>(wg_base, swt, bar);
            }
            __syncwarp();
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        fence_after_sync();
    }
    if constexpr (L == 0) {
        float *gx = ba.a.grad_x + (size_t)b * NP * ba.a.hw + p;
        unsigned short *gxh = reinterpret_cast<unsigned short *>(ba.a.grad_x) + (size_t)b * NP * ba.a.hw + p;   // bf16 twin
        const bool bf = ba.a.bf16_io != 0;
#pragma unroll 1
        for (int n0 = 0; n0 < NP; n0 += 16) {
            uint32_t r[16];
            tmem_ld<16>(lane_addr + 2 * AW + n0, r);
            wait_ld();
            if (valid) {
                if (bf) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) gxh[(size_t)(n0 + j) * ba.a.hw] = f32_to_bf16(__uint_as_float(r[j]));
                } else {
#pragma unroll
                    for (int j = 0; j < 16; ++j) gx[(size_t)(n0 + j) * ba.a.hw] = __uint_as_float(r[j]);
                }
            }
        }
    } else {
        const float *act = ba.act + (size_t)BwdRows<CFG>::row(L) * npp + gp;
        float *del = ba.del + (size_t)BwdRows<CFG>::row(L) * npp + gp;
        // a_l chunks are fetched TWO chunks ahead of their use (one chunk of processing is ~0.2 us, a DRAM round trip
        // ~0.8 us, and there are only eight warps per SM to hide it). The first AW channels go back to TMEM in a ROLLED
        // loop (these kernels run two to five tiles per warpgroup: straight-line code is fetched cold); the rest
        // stay in `keep` for the second round and need compile-time indices.
        constexpr int N1 = NP < AW ? NP : AW;
#pragma unroll 1
        for (int n0 = 0; n0 < N1; n0 += 16) {
            uint32_t r[16], lo[16];
            float nx2[16];
            if (n0 + 32 < NP) {
#pragma unroll
                for (int j = 0; j < 16; ++j) nx2[j] = __ldcg(act + (size_t)(n0 + 32 + j) * npp);
            }
            tmem_ld<16>(lane_addr + 2 * AW + n0, r);
            wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float v = __uint_as_float(r[j]) * elu_grad_from_out(av[j]);
                del[(size_t)(n0 + j) * npp] = v;
                r[j] = __float_as_uint(v);
                lo[j] = __float_as_uint(tf32_lo(v));
            }
            tmem_st<16>(lane_addr + n0, r);
            tmem_st<16>(lane_addr + AW + n0, lo);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                av[j] = nx[j];
                nx[j] = nx2[j];
            }
        }
#pragma unroll
        for (int n0 = N1; n0 < NP; n0 += 16) {
            uint32_t r[16];
            float nx2[16];
            if (n0 + 32 < NP) {
#pragma unroll
                for (int j = 0; j < 16; ++j) nx2[j] = __ldcg(act + (size_t)(n0 + 32 + j) * npp);
            }
            tmem_ld<16>(lane_addr + 2 * AW + n0, r);
            wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float v = __uint_as_float(r[j]) * elu_grad_from_out(av[j]);
                del[(size_t)(n0 + j) * npp] = v;
                keep[n0 - AW + j] = v;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                av[j] = nx[j];
                nx[j] = nx2[j];
            }
        }
    }
}

// TM (tail mode): -1 = delta_16 was left in the workspace by K1; 0 / 1 = stash backward -- the kernel
// starts from the stashed a16, a8 and the gradient of the head output (0: [B][1][hw]) or of the plane
// parameters (1: [B][4][hw]) and back-propagates through the SIMT tail itself.
template <class CFG, int NWG, int TM>
__global__ void __launch_bounds__(NWG * kTile, 1) chain_tc_dgrad_kernel(BwdArgs ba)
{
    constexpr int AW = kDgradAW<CFG>, NMAX = DgradCfg<CFG>::nmax();
    constexpr int COLS = 2 * AW + NMAX;
    constexpr int IMG = CFG::tail_off();          // floats of transposed images
    static_assert(NWG * COLS <= 512, "TMEM budget");
    extern __shared__ __align__(128) float smem[];
    constexpr int TAILW = kTailMid * kTailIn + 3 * kTailMid;   // raw W4 [8][16], W5 [LAST][8]
    float *swt = smem;
    float *stw = smem + IMG;
    uint64_t *bars = reinterpret_cast<uint64_t *>(stw + TAILW);
    uint32_t *slot = reinterpret_cast<uint32_t *>(bars + NWG);
    const int tid = threadIdx.x, wg = uniform(tid / kTile), t = tid % kTile;
    const int lt = uniform(t >> 5);                               // warp index inside the warpgroup; warp 0 issues the MMAs
    tc_pack_t_to_smem<CFG>(ba.a.W, swt, tid, NWG * kTile);
    if constexpr (TM >= 0) {
        constexpr int LAST = TM == 0 ? 1 : 3;
        for (int i = tid; i < kTailMid * kTailIn; i += NWG * kTile) stw[i] = __ldg(ba.w4 + i);
        for (int i = tid; i < LAST * kTailMid; i += NWG * kTile) stw[kTailMid * kTailIn + i] = __ldg(ba.w5 + i);
    }
    fence_smem_to_async();
    if (tid < 32) tmem_alloc(slot, 512);
    if (tid == 0) {
        for (int i = 0; i < NWG; ++i) mbar_init(&bars[i], 1);
        fence_mbar_init();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    pdl_trigger();                                // the weight-gradient kernel may take my SM as soon as I leave it
    pdl_wait();                                   // weights are packed; now the LPG reduction's output is needed
    const uint32_t tbase = (uint32_t)uniform((int)*slot);
    const uint32_t wg_base = tbase + wg * COLS;
    const uint32_t lane_addr = wg_base + ((uint32_t)((t >> 5) * 32) << 16);
    uint64_t *bar = &bars[wg];
    uint32_t phase = 0;
    const long long npp = ba.npx_pad;
    const long long ntiles = npp / kTile;
    // stash backward: the next tile's a16 / a8 rows and upstream gradient are requested BEFORE this tile's wide
    // layers, so their DRAM round trip hides behind the MMA hand-offs instead of opening every tile
    [[maybe_unused]] float na16[kTailIn], na8[kTailMid], ng[4];
    [[maybe_unused]] auto fetch_tail = [&](long long tl) {
        if constexpr (TM >= 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) ng[j] = 0.f;
            if (tl >= ntiles) return;
            const long long gq_ = tl * kTile + t;
            const int bb = (int)((unsigned)gq_ / (unsigned)ba.hw_pad), pp = (int)((unsigned)gq_ - (unsigned)bb * (unsigned)ba.hw_pad);
            const bool ok = bb < ba.a.B && pp < ba.a.hw;
            const float *a16p = ba.act + (size_t)BwdRows<CFG>::row(CFG::NW) * npp + gq_;
            const float *a8p = ba.act + (size_t)BwdRows<CFG>::row8() * npp + gq_;
#pragma unroll
            for (int j = 0; j < kTailIn; ++j) na16[j] = __ldcg(a16p + (size_t)j * npp);
#pragma unroll
            for (int j = 0; j < kTailMid; ++j) na8[j] = __ldcg(a8p + (size_t)j * npp);
            if (ok) {
                const TcArgs &a = ba.a;
                if constexpr (TM == 0) {
                    ng[0] = __ldcs(a.grad + (size_t)bb * a.hw + pp);
                } else {
                    const float *gq = a.grad + (size_t)bb * 4 * a.hw + pp;
                    ng[0] = __ldcg(gq); ng[1] = __ldcg(gq + a.hw); ng[2] = __ldcg(gq + (size_t)2 * a.hw); ng[3] = __ldcg(gq + (size_t)3 * a.hw);
                }
            }
        }
    };
    const long long tstride = (long long)gridDim.x * NWG;
    fetch_tail((long long)blockIdx.x * NWG + wg);
    for (long long tile = (long long)blockIdx.x * NWG + wg; tile < ntiles; tile += tstride) {
        const long long gp = tile * kTile + t;
        const int bb = (int)((unsigned)gp / (unsigned)ba.hw_pad), pp = (int)((unsigned)gp - (unsigned)bb * (unsigned)ba.hw_pad);
        const bool valid = bb < ba.a.B && pp < ba.a.hw;
        const int b = valid ? bb : 0;
        const int p = valid ? pp : 0;
        if constexpr (TM < 0) {
            const float *d16 = ba.del + (size_t)BwdRows<CFG>::row(CFG::NW) * npp + gp;
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) hi[j] = __float_as_uint(__ldcg(d16 + (size_t)j * npp));
#pragma unroll
            for (int j = 0; j < 16; ++j) lo[j] = __float_as_uint(tf32_lo(__uint_as_float(hi[j])));
            tmem_st<16>(lane_addr, hi);
            tmem_st<16>(lane_addr + AW, lo);
        } else {
            constexpr int LAST = TM == 0 ? 1 : 3;
            const TcArgs &a = ba.a;
            const float *tw = stw, *w5 = stw + kTailMid * kTailIn;
            float a16[kTailIn], a8[kTailMid], z[LAST], g[4];
#pragma unroll
            for (int j = 0; j < kTailIn; ++j) a16[j] = na16[j];
#pragma unroll
            for (int j = 0; j < kTailMid; ++j) a8[j] = na8[j];
#pragma unroll
            for (int j = 0; j < 4; ++j) g[j] = ng[j];
            tail_logits<LAST>(w5, a8, z);
            float dz[3] = {0.f, 0.f, 0.f};
            if constexpr (TM == 0) {
                const float sg = sigmoidf(z[0]);
                dz[0] = g[0] * sg * (1.f - sg);
            } else {
                const PlaneFwd pl = plane_from_logits(z[0], z[1], z[2], a.max_depth);
                if (valid) plane_bwd(pl, g[0], g[1], g[2], g[3], a.max_depth, dz);
            }
            float d8[kTailMid], d16[kTailIn];
#pragma unroll
            for (int o = 0; o < kTailMid; ++o) {
                float sm = 0.f;
#pragma unroll
                for (int j = 0; j < LAST; ++j) sm = fmaf(w5[j * kTailMid + o], dz[j], sm);
                d8[o] = sm * elu_grad_from_out(a8[o]);
            }
            tail_dgrad16<true, 0>(tw, d8, d16);
            float *del_gp = ba.del + gp;
            float *dz_dst = del_gp + (size_t)BwdRows<CFG>::rowz() * npp;
#pragma unroll
            for (int j = 0; j < LAST; ++j) dz_dst[(size_t)j * npp] = dz[j];
            float *d8_dst = del_gp + (size_t)BwdRows<CFG>::row8() * npp;
#pragma unroll
            for (int j = 0; j < kTailMid; ++j) d8_dst[(size_t)j * npp] = d8[j];
            float *d16_dst = del_gp + (size_t)BwdRows<CFG>::row(CFG::NW) * npp;
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int j = 0; j < kTailIn; ++j) {
                const float v = d16[j] * elu_grad_from_out(a16[j]);
                d16_dst[(size_t)j * npp] = v;
                hi[j] = __float_as_uint(v);
                lo[j] = __float_as_uint(tf32_lo(v));
            }
            tmem_st<16>(lane_addr, hi);
            tmem_st<16>(lane_addr + AW, lo);
            fetch_tail(tile + tstride);
        }
        float keep[64];
        if constexpr (CFG::NW > 3) dgrad_layer<CFG, 3>(ba, wg_base, lane_addr, swt, bar, phase, wg, lt, gp, valid, b, p, keep);
        if constexpr (CFG::NW > 2) dgrad_layer<CFG, 2>(ba, wg_base, lane_addr, swt, bar, phase, wg, lt, gp, valid, b, p, keep);
        if constexpr (CFG::NW > 1) dgrad_layer<CFG, 1>(ba, wg_base, lane_addr, swt, bar, phase, wg, lt, gp, valid, b, p, keep);
        if (ba.a.grad_x != nullptr) dgrad_layer<CFG, 0>(ba, wg_base, lane_addr, swt, bar, phase, wg, lt, gp, valid, b, p, keep);
    }
    fence_before_sync();
    __syncthreads();
    if (tid < 32) tmem_dealloc(tbase, 512);
}

// ------------------------------------------------------------------------------------------------
// K3: weight gradients. For a tile of 64 pixels the activations (A operand) and deltas (B operand)
// of a "job" are staged as K-major images with the PIXEL axis as K (thread = pixel, chunk stride
// padded to 144 B so the 4-byte stores of a warp hit 32 different banks), hi and lo halves, and
// multiplied by 3 x 8 SS-mode MMAs into a TMEM accumulator that lives for the whole kernel.
// The plan (which rows go where) is data, built on the host from the chain description.
// ------------------------------------------------------------------------------------------------
constexpr int kWP = 64;                      // pixels per staging tile (K of one MMA batch)
constexpr int kWLbo = 144;                   // bytes between 4-pixel chunks
constexpr int kWSbo = (kWP / 4) * kWLbo;     // bytes between 8-row groups
constexpr int kWMaxBlk = 5;

struct WBlock {
    const float *src;   // [rows][npx_pad] workspace matrix, or the NCHW input when is_x
    int rows;
    int is_x;
};
struct WJob {
    WBlock a[kWMaxBlk], b[kWMaxBlk];
    int na, nb;
    int MA, NB;         // MMA shape (rows padded: 64|128, multiple of 8|16)
    int tmem_col;
};
struct WExtract {       // layer l: D_job[row0 + k][col0 + n] -> partial[dst_off + n*K + k]
    int job, row0, col0, K, N, dst_off;
};
struct WPlan {
    WJob jobs[4];
    WExtract ex[BTS_MAX_LAYERS];
    int njobs, nlayers, total_w, tmem_cols;
    int imgA_floats, imgB_floats;           // per half (hi or lo)
    long long npx, npx_pad;
    int hw, C0;
    float *partials;                         // [gridDim.x][total_w]
};

__device__ __forceinline__ void wgrad_stage_block(const WBlock &blk, int row_base, float *hi, float *lo,
                                                  int px, int rq, long long gp, bool valid, int b, int p,
                                                  const WPlan &pl)
{
    const int chunk = (px >> 2) * (kWLbo / 4) + (px & 3);
    if (blk.is_x) {
        const float *src = blk.src + (size_t)b * pl.C0 * pl.hw + p;
#pragma unroll 4
        for (int r = rq; r < blk.rows; r += 4) {
            const float v = valid ? __ldg(src + (size_t)r * pl.hw) : 0.f;
            const int row = row_base + r;
            const int off = (row & 7) * 4 + (row >> 3) * (kWSbo / 4) + chunk;
            hi[off] = v;
            lo[off] = tc::tf32_lo(v);
        }
    } else {
        const float *src = blk.src + gp;
#pragma unroll 4
        for (int r = rq; r < blk.rows; r += 4) {
            const float v = __ldcg(src + (size_t)r * pl.npx_pad);
            const int row = row_base + r;
            const int off = (row & 7) * 4 + (row >> 3) * (kWSbo / 4) + chunk;
            hi[off] = v;
            lo[off] = tc::tf32_lo(v);
        }
    }
}

__global__ void __launch_bounds__(256, 1) chain_tc_wgrad_kernel(const __grid_constant__ WPlan pl)
{
    extern __shared__ __align__(128) float smem[];
    float *a_hi = smem, *a_lo = a_hi + pl.imgA_floats;
    float *b_hi = a_lo + pl.imgA_floats, *b_lo = b_hi + pl.imgB_floats;
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int tid = threadIdx.x;
    if (tid < 32) tmem_alloc(&slot, pl.tmem_cols);
    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tbase = slot;
    const int px = tid & (kWP - 1), rq = tid >> 6;
    uint32_t phase = 0;
    bool pending = false, first = true;
    const long long ntiles = pl.npx_pad / kWP;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long gp = tile * kWP + px;
        const bool valid = gp < pl.npx;
        const int b = valid ? (int)(gp / pl.hw) : 0;
        const int p = valid ? (int)(gp - (long long)b * pl.hw) : 0;
        for (int j = 0; j < pl.njobs; ++j) {
            const WJob &job = pl.jobs[j];
            if (pending) {                    // the previous MMAs still read the images
                mbar_wait(&bar, phase);
                phase ^= 1;
                pending = false;
            }
            int rb = 0;
            for (int i = 0; i < job.na; ++i) {
                wgrad_stage_block(job.a[i], rb, a_hi, a_lo, px, rq, gp, valid, b, p, pl);
                rb += job.a[i].rows;
            }
            rb = 0;
            for (int i = 0; i < job.nb; ++i) {
                wgrad_stage_block(job.b[i], rb, b_hi, b_lo, px, rq, gp, valid, b, p, pl);
                rb += job.b[i].rows;
            }
            fence_smem_to_async();
            __syncthreads();
            if (uniform(tid >> 5) == 0 && elect_one()) {
                fence_after_sync();
                const uint32_t idesc = idesc_tf32(job.MA, job.NB, 0, 0);
                const uint32_t d = tbase + job.tmem_col;
                const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
                uint32_t acc = first ? 0u : 1u;
#pragma unroll 1
                for (int pass = 0; pass < 3; ++pass) {
                    const uint32_t sa = pass == 0 ? al : ah, sb = pass == 1 ? bl : bh;
#pragma unroll
                    for (int ks = 0; ks < kWP / 8; ++ks) {
                        mma_tf32_ss(d, smem_desc(sa + ks * 2 * kWLbo, kWLbo, kWSbo),
                                    smem_desc(sb + ks * 2 * kWLbo, kWLbo, kWSbo), idesc, acc);
                        acc = 1u;
                    }
                }
                mma_commit(&bar);
            }
            pending = true;
        }
        first = false;
    }
    if (pending) {
        mbar_wait(&bar, phase);
        phase ^= 1;
    }
    fence_after_sync();
    // accumulators -> this CTA's partial sums
    float *part = pl.partials + (size_t)blockIdx.x * pl.total_w;
    if (first) {                              // CTA had no tile: contribute zeros
        for (int e = tid; e < pl.total_w; e += blockDim.x) part[e] = 0.f;
    } else if (tid < 128) {
        const int lane_row_base = (tid >> 5) * 32;
        const uint32_t lane_addr = tbase + ((uint32_t)lane_row_base << 16);
        for (int l = 0; l < pl.nlayers; ++l) {
            const WExtract &ex = pl.ex[l];
            const int MA = pl.jobs[ex.job].MA;
            // TMEM lane -> accumulator row: identity for M=128; M=64 keeps 16 rows per 32-lane quarter
            const int li = tid & 31;
            const int m = (MA == 128) ? tid : ((li < 16) ? (tid >> 5) * 16 + li : -1);
            const int k = m - ex.row0;
            const bool mine = (m >= 0) && k >= 0 && k < ex.K;
            for (int n0 = 0; n0 < ex.N; n0 += 8) {
                uint32_t r[8];
                tmem_ld<8>(lane_addr + pl.jobs[ex.job].tmem_col + ex.col0 + n0, r);
                wait_ld();
                if (mine) {
#pragma unroll
                    for (int jn = 0; jn < 8; ++jn)
                        if (n0 + jn < ex.N) part[ex.dst_off + (size_t)(n0 + jn) * ex.K + k] = __uint_as_float(r[jn]);
                }
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (tid < 32) tmem_dealloc(tbase, pl.tmem_cols);
}

// K4: grad_W[l] += sum over CTAs of the partial sums. 32 elements x 8 slices of the partial list per
// block (coalesced 128-byte rows, 8-way parallel over the list), fixed summation order.
__global__ void __launch_bounds__(256) chain_tc_wreduce_kernel(ChainParams cp, const float *partials, int nparts, int total_w)
{
    __shared__ float red[8][33];
    pdl_wait();
    const int el = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int e = blockIdx.x * 32 + el;
    float s = 0.f;
    if (e < total_w) {
        // eight independent loads in flight per thread, summed in the same (ascending) order as a plain loop
        int c = sl;
        for (; c + 56 < nparts; c += 64) {
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldcg(partials + (size_t)(c + 8 * i) * total_w + e);
#pragma unroll
            for (int i = 0; i < 8; ++i) s += v[i];
        }
        for (; c < nparts; c += 8) s += __ldcg(partials + (size_t)c * total_w + e);
    }
    red[sl][el] = s;
    __syncthreads();
    if (sl != 0 || e >= total_w) return;
#pragma unroll
    for (int j = 1; j < 8; ++j) s += red[j][el];
    int off = 0;
    for (int l = 0; l < cp.n_layers; ++l) {
        const int n = cp.dims[l] * cp.dims[l + 1];
        if (e < off + n) {
            cp.gW[l][e - off] += s;
            return;
        }
        off += n;
    }
}

int launch_wreduce(const ChainParams &cp, const float *partials, int nparts, int total_w, cudaStream_t s)
{
    // always follows a kernel of this library in the same call (weight-gradient or one-pass backward)
    launch_dependent(chain_tc_wreduce_kernel, dim3((total_w + 31) / 32), dim3(256), 0, s, cp, partials, nparts, total_w);
    return check_launch("chain_tc_wreduce");
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
size_t tc_pack_bytes();

static long long pad128(long long n) { return (n + kTile - 1) / kTile * kTile; }

static void bwd_rows(const ChainParams &cp, int *ra, int *rd, int *total_w)
{
    int a = 0, d = 0, w = 0;
    for (int l = 1; l < cp.n_layers; ++l) a += cp.dims[l];
    for (int l = 1; l <= cp.n_layers; ++l) d += cp.dims[l];
    for (int l = 0; l < cp.n_layers; ++l) w += cp.dims[l] * cp.dims[l + 1];
    *ra = a; *rd = d; *total_w = w;
}

// bytes the tensor-core backward needs after the two packed-weight regions
static int pad32(int n) { return (n + 31) / 32 * 32; }

size_t tc_bwd_workspace_bytes(const ChainParams &cp, int B, int hw, int sms)
{
    int ra, rd, tw;
    bwd_rows(cp, &ra, &rd, &tw);
    return sizeof(float) * ((size_t)(ra + rd) * (size_t)pad128((long long)B * pad32(hw)) + (size_t)sms * tw);
}

bool wgrad_tma_supported(const ChainParams &cp, int hw, const void *x);
int launch_wgrad_tma(const ChainParams &cp, const float *x, const float *act, const float *del, float *partials,
                     int B, int hw, int hw_pad, long long npp, int sms, cudaStream_t s, const float *x_rows = nullptr);
int launch_lpg_bwd_bf16_to_f32(const void *, const float *, float *, int, int, int, int, float, float, cudaStream_t);
int launch_cvt_bf16_to_f32(const void *, float *, long long, cudaStream_t);

static int build_wgrad_plan(const ChainParams &cp, WPlan &pl, const float *x, float *act, float *del,
                            long long npx, long long npp, int hw)
{
    memset(&pl, 0, sizeof(pl));
    const int L = cp.n_layers;
    int fc = L, sum = 0;
    while (fc > 0 && sum + cp.dims[fc - 1] <= 128) sum += cp.dims[--fc];
    if (fc == L) return set_error(BTS_ERR_UNSUPPORTED, "layer wider than 128 channels");
    int prefix[BTS_MAX_LAYERS + 2];                     // prefix[l] = sum_{j=1}^{l-1} dims[j]
    prefix[0] = prefix[1] = 0;
    for (int l = 2; l <= L + 1; ++l) prefix[l] = prefix[l - 1] + cp.dims[l - 1];
    int col = 0, maxMA = 0, maxNB = 0, woff = 0;
    int woffs[BTS_MAX_LAYERS];
    for (int l = 0; l < L; ++l) { woffs[l] = woff; woff += cp.dims[l] * cp.dims[l + 1]; }
    pl.njobs = fc + 1;
    if (pl.njobs > 4) return set_error(BTS_ERR_UNSUPPORTED, "too many weight-gradient jobs");
    for (int j = 0; j < pl.njobs; ++j) {
        WJob &job = pl.jobs[j];
        const int l0 = j < fc ? j : fc, l1 = j < fc ? j + 1 : L;
        if (l1 - l0 > kWMaxBlk) return set_error(BTS_ERR_UNSUPPORTED, "too many layers in one job");
        int ra = 0, rb = 0;
        for (int l = l0; l < l1; ++l) {
            if (cp.dims[l] > 128) return set_error(BTS_ERR_UNSUPPORTED, "layer wider than 128 channels");
            WBlock &ba = job.a[job.na++];
            ba.is_x = (l == 0);
            ba.src = (l == 0) ? x : act + (size_t)prefix[l] * npp;
            ba.rows = cp.dims[l];
            WBlock &bb = job.b[job.nb++];
            bb.is_x = 0;
            bb.src = del + (size_t)prefix[l + 1] * npp;
            bb.rows = cp.dims[l + 1];
            WExtract &ex = pl.ex[l];
            ex.job = j; ex.row0 = ra; ex.col0 = rb; ex.K = cp.dims[l]; ex.N = cp.dims[l + 1]; ex.dst_off = woffs[l];
            ra += cp.dims[l];
            rb += cp.dims[l + 1];
        }
        job.MA = ra <= 64 ? 64 : 128;
        const int g = job.MA == 128 ? 16 : 8;
        job.NB = (rb + g - 1) / g * g;
        if (job.NB < g) job.NB = g;
        if (job.NB > 256) return set_error(BTS_ERR_UNSUPPORTED, "delta block wider than 256");
        job.tmem_col = col;
        col += job.NB;
        if (job.MA > maxMA) maxMA = job.MA;
        if (job.NB > maxNB) maxNB = job.NB;
    }
    if (col > 512) return set_error(BTS_ERR_UNSUPPORTED, "weight-gradient accumulators exceed TMEM");
    int cols = 32;
    while (cols < col) cols <<= 1;
    pl.tmem_cols = cols;
    pl.nlayers = L;
    pl.total_w = woff;
    pl.imgA_floats = (maxMA / 8) * (kWSbo / 4);
    pl.imgB_floats = ((maxNB + 7) / 8) * (kWSbo / 4);
    pl.npx = npx; pl.npx_pad = npp; pl.hw = hw; pl.C0 = cp.dims[0];
    return BTS_OK;
}

template <class CFG, int NWG>
static constexpr size_t dgrad_smem_bytes()
{
    return sizeof(float) * (CFG::tail_off() + kTailMid * kTailIn + 3 * kTailMid) + 8 * NWG + 16;
}

template <class CFG, int MODE, int R, int NWG1, int NWG2, int AW>
static int launch_bwd_cfg(const ChainParams &cp, const TcArgs &a, char *ws, int sms, cudaStream_t s)
{
    constexpr int LAST = MODE == 0 ? 1 : 3;
    const bool tma = wgrad_tma_supported(cp, a.hw, a.x);
    const int hw_pad = tma ? pad32(a.hw) : a.hw;
    const long long npp = pad128((long long)a.B * hw_pad);
    int ra, rd, tw;
    bwd_rows(cp, &ra, &rd, &tw);
    float *act = reinterpret_cast<float *>(ws + 2 * tc_pack_bytes());
    float *del = act + (size_t)ra * npp;
    float *partials = del + (size_t)rd * npp;

    BwdArgs ba;
    ba.a = a;
    tc_set_weights(ba.a, cp);
    ba.act = act; ba.del = del; ba.npx_pad = npp; ba.hw_pad = hw_pad;
    const long long ntiles = npp / kTile;
    {
        auto k1 = chain_tc_bwd1_kernel<CFG, MODE, R, NWG1, AW>;
        const size_t smem = sizeof(float) * CFG::pack_floats(LAST) + 8 * NWG1 + 16;
        cudaError_t e = cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        long long grid = (ntiles + NWG1 - 1) / NWG1;
        if (grid > sms) grid = sms;
        k1<<<(unsigned)grid, NWG1 * kTile, smem, s>>>(ba);
        if (int e2 = check_launch("chain_tc_bwd1")) return e2;
    }
    {
        auto k2 = chain_tc_dgrad_kernel<CFG, NWG2, -1>;
        const size_t smem = dgrad_smem_bytes<CFG, NWG2>();
        cudaError_t e = cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        long long grid = (ntiles + NWG2 - 1) / NWG2;
        if (grid > sms) grid = sms;
        // with no grad_x and a single wide layer there is nothing for K2 to do
        if (a.grad_x != nullptr || CFG::NW > 1) {
            k2<<<(unsigned)grid, NWG2 * kTile, smem, s>>>(ba);
            if (int e2 = check_launch("chain_tc_dgrad")) return e2;
        }
    }
    if (cp.gW[0] != nullptr && tma) {
        if (int e = launch_wgrad_tma(cp, a.x, act, del, partials, a.B, a.hw, hw_pad, npp, sms, s)) return e;
    } else if (cp.gW[0] != nullptr) {
        WPlan pl;
        if (int e = build_wgrad_plan(cp, pl, a.x, act, del, a.npx, npp, a.hw)) return e;
        pl.partials = partials;
        const size_t smem = sizeof(float) * 2 * (size_t)(pl.imgA_floats + pl.imgB_floats);
        cudaError_t e = cudaFuncSetAttribute(chain_tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        long long grid = npp / kWP;
        if (grid > sms) grid = sms;
        chain_tc_wgrad_kernel<<<(unsigned)grid, 256, smem, s>>>(pl);
        if (int e2 = check_launch("chain_tc_wgrad")) return e2;
        if (int e2 = launch_wreduce(cp, partials, (int)grid, pl.total_w, s)) return e2;
    }
    return BTS_OK;
}

// ---- stash backward: the forward left a_1 .. a8 (and plane_eq for the fused stage) in the caller's stash,
// so K1 disappears: [LPG block reduction at full parallelism ->] tail backward + dgrad -> wgrad -> reduce.
int launch_lpg_bwd_glue(const float *, const float *, float *, int, int, int, int, float, const LpgGlue &, cudaStream_t);
bool fused_bwd_supported(const ChainParams &);
bool tc_supported(const ChainParams &);

bool tc_stash_supported(const ChainParams &cp, int hw, const void *x)
{
    return tc_supported(cp) && !fused_bwd_supported(cp) && wgrad_tma_supported(cp, hw, x);
}

static size_t stash_act_rows(const ChainParams &cp)
{
    size_t ra = 0;
    for (int l = 1; l < cp.n_layers; ++l) ra += cp.dims[l];
    return ra;
}

size_t tc_stash_bytes(const ChainParams &cp, int B, int hw)
{
    const long long npp = pad128((long long)B * pad32(hw));
    return sizeof(float) * (stash_act_rows(cp) * (size_t)npp + (cp.is_final ? 0 : (size_t)4 * B * hw));
}
// bf16 twin: the stash also holds the features as fp32 rows [D0][npp] (the TMA weight-gradient kernel's operand)
size_t tc_stash_bytes_bf16(const ChainParams &cp, int B, int hw)
{
    const long long npp = pad128((long long)B * pad32(hw));
    return tc_stash_bytes(cp, B, hw) + sizeof(float) * (size_t)cp.dims[0] * (size_t)npp;
}

// point the TcArgs at the stash; with `clear`, zero the activation rows when the padded pixel axis has
// columns no forward thread writes (the weight-gradient MMAs read whole rows)
int tc_stash_bind(const ChainParams &cp, TcArgs &a, void *stash, bool clear, cudaStream_t s)
{
    a.stash_hw_pad = pad32(a.hw);
    a.stash_npp = pad128((long long)a.B * a.stash_hw_pad);
    a.stash_act = static_cast<float *>(stash);
    a.stash_pe = cp.is_final ? nullptr : a.stash_act + stash_act_rows(cp) * (size_t)a.stash_npp;
    a.stash_x = a.bf16_io ? a.stash_act + stash_act_rows(cp) * (size_t)a.stash_npp + (cp.is_final ? 0 : (size_t)4 * a.B * a.hw) : nullptr;
    if (clear && a.stash_npp != a.npx) {
        cudaError_t e = cudaMemsetAsync(stash, 0, sizeof(float) * stash_act_rows(cp) * (size_t)a.stash_npp, s);
        if (e != cudaSuccess) return set_error((int)e, "cudaMemsetAsync(stash): %s", cudaGetErrorString(e));
        if (a.stash_x != nullptr) {
            e = cudaMemsetAsync(a.stash_x, 0, sizeof(float) * (size_t)cp.dims[0] * (size_t)a.stash_npp, s);
            if (e != cudaSuccess) return set_error((int)e, "cudaMemsetAsync(stash x rows): %s", cudaGetErrorString(e));
        }
    }
    return BTS_OK;
}

template <class CFG, int TM, int NWG2>
static int launch_bwd_stash_cfg(const ChainParams &cp, const TcArgs &a, int r, char *ws, int sms, cudaStream_t s)
{
    if constexpr (!kStashCfg<CFG>) {
        return set_error(BTS_ERR_UNSUPPORTED, "chain does not stash");
    } else {
        const long long npp = a.stash_npp;
        int ra, rd, tw;
        bwd_rows(cp, &ra, &rd, &tw);
        float *scratch = reinterpret_cast<float *>(ws + 2 * tc_pack_bytes());   // the recompute path's a_l rows: free here
        float *del = scratch + (size_t)ra * npp;
        float *partials = del + (size_t)rd * npp;

        BwdArgs ba;
        ba.a = a;
        tc_set_weights(ba.a, cp);
        ba.act = a.stash_act; ba.del = del; ba.npx_pad = npp; ba.hw_pad = a.stash_hw_pad;
        ba.w4 = cp.W[CFG::NW]; ba.w5 = cp.W[CFG::NW + 1];
        bool chained = TM == 1 && r > 0;     // the dgrad kernel follows a kernel of this call (programmatic dependent launch)
        if (TM == 1 && r > 0) {              // grad_depth -> gradient of the plane parameters [B][4][hw]
            if (a.bf16_io) {
                if (int e = launch_lpg_bwd_bf16_to_f32(a.grad, a.stash_pe, scratch, a.B, a.h, a.w, r, a.eps, a.depth_div, s)) return e;
            } else {
                const LpgGlue gl{a.depth_div, a.grad_bstride, nullptr, a.grad_ds, a.grad_ds_bstride};
                if (int e = launch_lpg_bwd_glue(a.grad, a.stash_pe, scratch, a.B, a.h, a.w, r, a.eps, gl, s)) return e;
            }
            ba.a.grad = scratch;
        } else if (a.bf16_io) {              // bf16 gradient of the head output / plane parameters -> fp32 scratch
            if (int e = launch_cvt_bf16_to_f32(a.grad, scratch, (long long)(TM == 0 ? 1 : 4) * a.B * a.hw, s)) return e;
            ba.a.grad = scratch;
            chained = true;
        }
        const long long ntiles = npp / kTile;
        auto k2 = chain_tc_dgrad_kernel<CFG, NWG2, TM>;
        const size_t smem = dgrad_smem_bytes<CFG, NWG2>();
        cudaError_t e = cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        long long grid = (ntiles + NWG2 - 1) / NWG2;
        if (grid > sms) grid = sms;
        if (chained) launch_dependent(k2, dim3((unsigned)grid), dim3(NWG2 * kTile), smem, s, ba);
        else k2<<<(unsigned)grid, NWG2 * kTile, smem, s>>>(ba);
        if (int e2 = check_launch("chain_tc_dgrad")) return e2;
        if (cp.gW[0] != nullptr)
            return launch_wgrad_tma(cp, a.x, a.stash_act, del, partials, a.B, a.hw, a.stash_hw_pad, npp, sms, s,
                                    a.bf16_io ? a.stash_x : nullptr);
        return BTS_OK;
    }
}

template <class CFG, int NWG2>
static int dispatch_bwd_stash(const ChainParams &cp, const TcArgs &a, int mode, int r, char *ws, int sms, cudaStream_t s)
{
    if (mode == 0) return launch_bwd_stash_cfg<CFG, 0, NWG2>(cp, a, 0, ws, sms, s);
    return launch_bwd_stash_cfg<CFG, 1, NWG2>(cp, a, mode == 2 ? r : 0, ws, sms, s);
}

// same modes as launch_chain_tc_bwd; a.stash_* must be bound to the buffer the forward filled
int launch_chain_tc_bwd_stash(const ChainParams &cp, const TcArgs &a, int mode, int r, void *ws, int sms, cudaStream_t s)
{
    char *w = static_cast<char *>(ws);
    if (cfg_matches<Cfg128_128>(cp)) return dispatch_bwd_stash<Cfg128_128, 2>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg128_64>(cp)) return dispatch_bwd_stash<Cfg128_64, 2>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg64_64>(cp)) return dispatch_bwd_stash<Cfg64_64, 2>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg32_32>(cp)) return dispatch_bwd_stash<Cfg32_32, 3>(cp, a, mode, r, w, sms, s);
    return set_error(BTS_ERR_UNSUPPORTED, "chain does not stash");
}

template <class CFG, int NWG1, int NWG2, int AW = CFG::kmax()>
static int dispatch_bwd_mode(const ChainParams &cp, const TcArgs &a, int mode, int r, char *ws, int sms, cudaStream_t s)
{
    if (mode == 0) return launch_bwd_cfg<CFG, 0, 1, NWG1, NWG2, AW>(cp, a, ws, sms, s);
    if (mode == 1) return launch_bwd_cfg<CFG, 1, 1, NWG1, NWG2, AW>(cp, a, ws, sms, s);
    switch (r) {
    case 8: return launch_bwd_cfg<CFG, 2, 8, NWG1, NWG2, AW>(cp, a, ws, sms, s);
    case 4: return launch_bwd_cfg<CFG, 2, 4, NWG1, NWG2, AW>(cp, a, ws, sms, s);
    case 2: return launch_bwd_cfg<CFG, 2, 2, NWG1, NWG2, AW>(cp, a, ws, sms, s);
    default: return launch_bwd_cfg<CFG, 2, 1, NWG1, NWG2, AW>(cp, a, ws, sms, s);
    }
}

// mode: 0 head (grad [B,1,hw]), 1 plane parameters (grad [B,4,hw]), 2 fused LPG (grad_depth, upratio r)
int launch_chain_tc_bwd(const ChainParams &cp, const TcArgs &a, int mode, int r, void *ws, int sms, cudaStream_t s)
{
    char *w = static_cast<char *>(ws);
    if (cfg_matches<Cfg128_128>(cp)) return dispatch_bwd_mode<Cfg128_128, 2, 2, 64>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg128_64>(cp)) return dispatch_bwd_mode<Cfg128_64, 2, 2, 64>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg64_64>(cp)) return dispatch_bwd_mode<Cfg64_64, 2, 2>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg64_32>(cp)) return dispatch_bwd_mode<Cfg64_32, 3, 3>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg32_32>(cp)) return dispatch_bwd_mode<Cfg32_32, 3, 3>(cp, a, mode, r, w, sms, s);
    if (cfg_matches<Cfg32_16>(cp)) return dispatch_bwd_mode<Cfg32_16, 3, 3>(cp, a, mode, r, w, sms, s);
    return set_error(BTS_ERR_UNSUPPORTED, "chain not covered by the tensor-core path");
}

}  // namespace bts

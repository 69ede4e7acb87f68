// chain_tc_fwd.cu -- fused forward of one BTS decoder stage on sm_100a tensor cores:
//   reduction_1x1 chain (pytorch/bts.py:124-136)  ->  plane parameterisation (127-134)
//   -> [normalise/cat glue 250-253: identity on a unit normal]  ->  LPG expansion (149-173)
// in ONE persistent kernel. Features are read once from HBM, depth is written once; no intermediate
// ever leaves the SM. See chain_tc.cuh for the split between tcgen05 (wide layers) and SIMT (tail).
//
// Work decomposition: a warpgroup (128 threads) owns one tile of 128 consecutive low-res pixels;
// thread t <-> pixel t <-> TMEM lane t. NWG independent warpgroups per CTA (each with its own TMEM
// columns, mbarrier and MMA-issuing thread) keep the tensor core, the LSU and the FMA pipes busy at
// the same time: while one warpgroup waits for its MMAs another one loads or runs its epilogue.
#include <type_traits>

#include "chain_tc.cuh"

namespace bts {

#ifndef BTS_WIDE_SUBS
#define BTS_WIDE_SUBS 2
#endif
constexpr int kWideSubs = BTS_WIDE_SUBS;   // threads per pixel in the 128-channel chains' forward

// PREFETCH: the next tile's features travel global -> shared with cp.async while the current tile is
// computed (needs D0*512 B of shared memory per warpgroup next to the weight images).
// SUBS: threads per pixel (chain_tc.cuh): 2 for the 128-channel chains, whose tiles are few and long.
// BF: bf16 twin -- x, out and depth are bf16 tensors (TcArgs carries them as float pointers), arithmetic unchanged.
template <class CFG, int R, bool FINAL, int NWG, bool PREFETCH, int AW, int SUBS, bool BF>
__global__ void __launch_bounds__(NWG * SUBS * kTile, 1) chain_tc_fwd_kernel(TcArgs a)
{
    using TIO = std::conditional_t<BF, unsigned short, float>;
    constexpr int PXG = BF ? 8 : 4;                               // pixels per 16-byte copy
    constexpr int NGRP = kTile / PXG;                             // copy groups per tile row
    constexpr int GT = SUBS * kTile;                              // threads of one tile group
    constexpr int LAST = FINAL ? 1 : 3;
    constexpr int PACK = CFG::pack_floats(LAST);
    constexpr int NMAX = CFG::nmax();
    constexpr int COLS = 2 * AW + NMAX;                           // A hi | A lo | D
    constexpr int D0 = CFG::dim(0);
    constexpr int KC0 = D0 < AW ? D0 : AW;                        // input channels per round of layer 0
    static_assert(NWG * COLS <= 512, "TMEM budget");
    static_assert(D0 % KC0 == 0 && D0 / KC0 <= 2, "layer 0 runs in one or two rounds");
    static_assert(CFG::dim(CFG::NW) == kTailIn, "tensor-core part must end at 16 channels");
    constexpr bool SPILL = kStashCfg<CFG>;

    extern __shared__ __align__(128) float smem[];
    float *sw = smem;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + PACK);
    uint32_t *slot = reinterpret_cast<uint32_t *>(bars + NWG);
    float *xs_all = smem + PACK + (8 * NWG + 16 + 15) / 16 * 4;   // [NWG][D0][128] when PREFETCH (16-byte aligned)

    const int tid = threadIdx.x, wg = uniform(tid / GT), tt = tid % GT, t = tt % kTile, lane = tid & 31;
    const int sub = uniform(tt / kTile);                          // which half of the columns / rows this thread takes
    const int lt = uniform(tt >> 5);                              // warp index inside the tile group; warp 0 issues the MMAs
    uint64_t *bar = &bars[wg];
    uint32_t phase = 0;
    const float *tw = sw + CFG::tail_off();
    int key = 0x7f800000;

    // 32-bit index arithmetic throughout (the C ABI rejects more than 2^31 - 1 low-res pixels)
    const unsigned npx = (unsigned)a.npx, hw = (unsigned)a.hw;
    const int ntiles = (int)((npx + kTile - 1) / kTile);
    const int stride = (int)gridDim.x * NWG;
    TIO *xs_wg = reinterpret_cast<TIO *>(xs_all) + (size_t)wg * D0 * kTile;   // [channel][pixel] of the staged tile
    TIO *xs = xs_wg + t;                                          // this thread's pixel column
    const TIO *const xg = reinterpret_cast<const TIO *>(a.x);
    // 128-bit copies: 4 pixels never straddle two images because hw % 4 == 0 (the launcher picks the
    // non-prefetching kernel otherwise). Two mappings, chosen by measurement: WARP_PRIVATE (32-channel head) -- each
    // warp copies its own 32 pixels (lane l: pixels 4(l%8)..+3 of channels l/8, l/8+4, ...) and the staged tile
    // needs no barrier wider than the warp; otherwise a warp copies 512 contiguous bytes of one channel per
    // instruction (thread t: pixels 4(t%32)..+3 of channels t/32, t/32+4, ...), which DRAM prefers for wider
    // chains, and the warpgroup synchronises on a named barrier.
    constexpr bool WARP_PRIVATE = D0 <= 32 && SUBS == 1;
    auto tile_sync = [&]() {
        if constexpr (WARP_PRIVATE) __syncwarp();
        else named_bar_sync(1 + wg, GT);
    };
    // (image, pixel) of this thread's own pixel and of the 4-pixel group it copies, advanced by the same number of
    // pixels every iteration: one division per kernel (dq, dr) instead of two per position and tile
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
    const int tile0 = (int)blockIdx.x * NWG + wg;
    // copy group of this thread: WARP_PRIVATE -> one of the NGRP / 4 groups of its own warp's 32 pixels
    const int pq = WARP_PRIVATE ? ((t >> 5) * (NGRP / 4) + (lane & (NGRP / 4 - 1))) : (tt & (NGRP - 1));
    PixPos own = pos_at((unsigned)tile0 * kTile + t), quad = pos_at((unsigned)tile0 * kTile + PXG * pq);
    auto prefetch = [&](int tl) {                                 // fetches the tile `quad` points at
        // channels copied per pass: a warp's 32 lanes cover NGRP / 4 groups x (128 / NGRP) channels when WARP_PRIVATE,
        // else the tile group's GT threads cover NGRP groups x GT / NGRP channels
        constexpr int CWN = WARP_PRIVATE ? 32 / (NGRP / 4) : GT / NGRP;
        const int q = pq, cw = WARP_PRIVATE ? lane / (NGRP / 4) : tt / NGRP;
        const bool ok = tl < ntiles && quad.g < npx;
        const int bn = ok ? quad.b : 0;
        const int pn = ok ? quad.p : 0;
        const TIO *xq = xg + ((size_t)bn * D0 + cw) * a.hw + pn;
        const uint32_t dst = smem_u32(xs_wg + cw * kTile + PXG * q);
#pragma unroll
        for (int i = 0; i < D0 / CWN; ++i) cp_async16(dst + i * CWN * kTile * (int)sizeof(TIO), xq + (size_t)i * CWN * a.hw, ok);
        cp_async_commit();
    };
    if constexpr (PREFETCH) prefetch(tile0);                      // in flight while the weight images are packed
    tc_pack_to_smem<CFG>(a.W, sw, LAST, tid, NWG * GT);
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
    for (int tile = tile0; tile < ntiles; tile += stride) {
        const unsigned gp = own.g;                                // == tile * kTile + t
        const bool valid = gp < npx;
        const int b = valid ? own.b : 0;
        const int p = valid ? own.p : 0;
        pos_advance(own);
        if constexpr (PREFETCH) pos_advance(quad);                // now the tile the next prefetch fetches
        const TIO *xp = xg + (size_t)b * D0 * a.hw + p;
        bool stash = false;
        float *act_gp = nullptr;
        if constexpr (SPILL) {
            stash = valid && a.stash_act != nullptr;
            act_gp = a.stash_act + (size_t)b * a.stash_hw_pad + p;
        }
        // channels [C0, C0+KC0) of the features -> (hi, lo) -> TMEM A operand
        // bf16 twin, wide chains: the TMA weight-gradient kernel wants fp32 rows, so the features go to the stash too
        [[maybe_unused]] float *xrow = nullptr;
        if constexpr (BF && SPILL) {
            if (stash && a.stash_x != nullptr) xrow = a.stash_x + (size_t)b * a.stash_hw_pad + p;
        }
        // XS: also write the features to the stash as fp32 rows (bf16 twin of a stashing chain, training only). A template
        // flag chosen by ONE kernel-uniform branch per round: predicated-off stores in the inner loops cost 25 % of the
        // inference forward.
        auto load_x_impl = [&](auto c0_tag, auto xs_tag) {
            constexpr int C0 = decltype(c0_tag)::value;
            constexpr bool XS = decltype(xs_tag)::value;
            constexpr int KS = KC0 / SUBS;                        // channels of this round per sub-thread
            const int cb = sub * KS;
            if constexpr (PREFETCH) {
#pragma unroll 1
                for (int c0 = cb; c0 < cb + KS; c0 += 16) {
                    uint32_t hi[16], lo[16];
#pragma unroll
                    for (int j = 0; j < 16; j += 2) {
                        float v0, v1;
                        if constexpr (BF) {
                            v0 = bf16_to_f32(xs[(C0 + c0 + j) * kTile]);
                            v1 = bf16_to_f32(xs[(C0 + c0 + j + 1) * kTile]);
                        } else {
                            v0 = xs[(C0 + c0 + j) * kTile];
                            v1 = xs[(C0 + c0 + j + 1) * kTile];
                            tf32_lo2(v0, v1, lo[j], lo[j + 1]);
                        }
                        hi[j] = __float_as_uint(v0);
                        hi[j + 1] = __float_as_uint(v1);
                        if constexpr (XS) {
                            if (xrow != nullptr) {                // (per thread: pixels past the end have no row)
                                xrow[(size_t)(C0 + c0 + j) * a.stash_npp] = v0;
                                xrow[(size_t)(C0 + c0 + j + 1) * a.stash_npp] = v1;
                            }
                        }
                    }
                    tmem_st<16>(lane_addr + c0, hi);
                    if constexpr (!BF) tmem_st<16>(lane_addr + AW + c0, lo);
                }
            } else {
                // up to 64 independent coalesced loads in flight per thread (a whole K round of the 128-channel chains:
                // with few warps per SM every DRAM round trip saved is ~1 us of the tile)
                constexpr int LB = KS >= 64 ? 64 : 32;
                static_assert(KS % 32 == 0, "whole 32-channel steps per sub-thread");
#pragma unroll 1
                for (int c0 = cb; c0 < cb + KS; c0 += LB) {
                    uint32_t hi[LB];
#pragma unroll
                    for (int j = 0; j < LB; ++j) hi[j] = valid ? __float_as_uint(io_load<TIO>(xp + (size_t)(C0 + c0 + j) * a.hw)) : 0u;
#pragma unroll
                    for (int h = 0; h < LB; h += 32) {
                        uint32_t h32[32], lo[32];
#pragma unroll
                        for (int j = 0; j < 32; j += 2) {
                            h32[j] = hi[h + j];
                            h32[j + 1] = hi[h + j + 1];
                            if constexpr (!BF) tf32_lo2(__uint_as_float(hi[h + j]), __uint_as_float(hi[h + j + 1]), lo[j], lo[j + 1]);
                            if constexpr (XS) {
                                if (xrow != nullptr) {
                                    xrow[(size_t)(C0 + c0 + h + j) * a.stash_npp] = __uint_as_float(hi[h + j]);
                                    xrow[(size_t)(C0 + c0 + h + j + 1) * a.stash_npp] = __uint_as_float(hi[h + j + 1]);
                                }
                            }
                        }
                        tmem_st<32>(lane_addr + c0 + h, h32);
                        if constexpr (!BF) tmem_st<32>(lane_addr + AW + c0 + h, lo);
                    }
                }
            }
        };
        auto load_x = [&](auto c0_tag) {
            if constexpr (BF && SPILL) {
                // KERNEL-uniform condition (the loads end in warp-collective tcgen05.st: no divergence around them)
                if (a.stash_x != nullptr && a.stash_act != nullptr) {
                    load_x_impl(c0_tag, std::true_type{});
                    return;
                }
            }
            load_x_impl(c0_tag, std::false_type{});
        };
        if constexpr (PREFETCH) {
            cp_async_wait_all();
            tile_sync();                                          // the staged pixels this thread reads have landed
        }
        load_x(std::integral_constant<int, 0>{});
        if constexpr (D0 > KC0) {
            tc_run_chunk<CFG, 0, AW, 0, KC0, SUBS, BF>(wg_base, sw, bar, phase, wg, lt);
            load_x(std::integral_constant<int, KC0>{});
        }
        if constexpr (PREFETCH) {
            tile_sync();                                          // the staged tile has been read by everyone who shares it
            prefetch(tile + stride);
        }
        tc_run_chunk<CFG, 0, AW, D0 - KC0, KC0, SUBS, BF>(wg_base, sw, bar, phase, wg, lt);
        // wide layers
        if constexpr (CFG::NW > 1) tc_layer_from_d<CFG, 1, AW, SPILL, SUBS>(wg_base, lane_addr, sw, bar, phase, wg, lt, act_gp, a.stash_npp, stash, sub);
        if constexpr (CFG::NW > 2) tc_layer_from_d<CFG, 2, AW, SPILL, SUBS>(wg_base, lane_addr, sw, bar, phase, wg, lt, act_gp, a.stash_npp, stash, sub);
        if constexpr (CFG::NW > 3) tc_layer_from_d<CFG, 3, AW, SPILL, SUBS>(wg_base, lane_addr, sw, bar, phase, wg, lt, act_gp, a.stash_npp, stash, sub);
        float a16[kTailIn], a8[kTailMid], z[LAST];
        {
            uint32_t r[16];
            tmem_ld<16>(lane_addr + 2 * AW, r);
            wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) a16[j] = elu_fast(__uint_as_float(r[j]));
        }
        tail_forward<LAST, true, false>(tw, a16, a8, z);           // no scheduling fences: the weight loads float above the FMAs
        if constexpr (kStashCfg<CFG>) {
            if (stash && (SUBS == 1 || sub == 0)) {              // both sub-threads hold the tail: one stores a16, the other a8
                float *d16 = act_gp + (size_t)BwdRows<CFG>::row(CFG::NW) * a.stash_npp;
#pragma unroll
                for (int j = 0; j < kTailIn; ++j) d16[(size_t)j * a.stash_npp] = a16[j];
            }
            if (stash && (SUBS == 1 || sub == 1)) {
                float *d8 = act_gp + (size_t)BwdRows<CFG>::row8() * a.stash_npp;
#pragma unroll
                for (int j = 0; j < kTailMid; ++j) d8[(size_t)j * a.stash_npp] = a8[j];
            }
        }
        if constexpr (FINAL) {
            if (valid && sub == 0) io_store(reinterpret_cast<TIO *>(a.out) + gp, sigmoidf(z[0]));
        } else {
            const PlaneFwd pl = plane_from_logits(z[0], z[1], z[2], a.max_depth);
            if (a.out != nullptr && valid && sub == 0) {
                TIO *o = reinterpret_cast<TIO *>(a.out) + (size_t)b * 4 * a.hw + p;
                io_store(o, pl.n1); io_store(o + (size_t)a.hw, pl.n2); io_store(o + (size_t)2 * a.hw, pl.n3); io_store(o + (size_t)3 * a.hw, pl.n4);
            }
            if constexpr (kStashCfg<CFG>) {
                if (stash && a.stash_pe != nullptr && sub == SUBS - 1) {
                    float *o = a.stash_pe + (size_t)b * 4 * a.hw + p;
                    o[0] = pl.n1; o[(size_t)a.hw] = pl.n2; o[(size_t)2 * a.hw] = pl.n3; o[(size_t)3 * a.hw] = pl.n4;
                }
            }
            if (a.depth != nullptr) key = min(key, lpg_expand<R, (R % 2 == 0 ? SUBS : 1), TIO>(a, pl.n1, pl.n2, pl.n3, pl.n4, gp, b, p, valid && (R % 2 == 0 || sub == 0), lane, sub));
        }
    }
    if constexpr (!FINAL) {
        if (a.depth != nullptr && a.abs_min != nullptr) {
            key = warp_min_int(key);
            int *addr = reinterpret_cast<int *>(a.abs_min);
            if (lane == 0 && key < *reinterpret_cast<volatile int *>(addr)) atomicMin(addr, key);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (tid < 32) tmem_dealloc(tbase, 512);
}

template <class CFG, int R, bool FINAL, int NWG, int AW, int SUBS, bool BF>
static int launch_one(const ChainParams &cp, const TcArgs &a, int sms, cudaStream_t s)
{
    constexpr int LAST = FINAL ? 1 : 3;
    constexpr size_t base = sizeof(float) * CFG::pack_floats(LAST) + (8 * NWG + 16 + 15) / 16 * 16;
    constexpr size_t xs_bytes = (size_t)NWG * CFG::dim(0) * kTile * (BF ? 2 : 4);
    constexpr bool CAN_PREFETCH = base + xs_bytes <= 227 * 1024;
    const bool prefetch = CAN_PREFETCH && a.x_vec16;
    const size_t smem = base + (prefetch ? xs_bytes : 0);
    auto kern = prefetch ? chain_tc_fwd_kernel<CFG, R, FINAL, NWG, CAN_PREFETCH, AW, SUBS, BF> : chain_tc_fwd_kernel<CFG, R, FINAL, NWG, false, AW, SUBS, BF>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long ntiles = (a.npx + kTile - 1) / kTile;
    long long grid = (ntiles + NWG - 1) / NWG;
    if (grid > sms) grid = sms;
    TcArgs aa = a;
    tc_set_weights(aa, cp);
    kern<<<(unsigned)grid, NWG * SUBS * kTile, smem, s>>>(aa);
    return check_launch("chain_tc_fwd");
}

// header bytes at the start of the workspace (nothing is packed in HBM any more: every CTA builds its operand images
// in shared memory); a small aligned prefix so that the scratch rows start 256-byte aligned
size_t tc_pack_bytes() { return 128; }

bool tc_supported(const ChainParams &cp)
{
    return cfg_matches<Cfg128_128>(cp) || cfg_matches<Cfg128_64>(cp) || cfg_matches<Cfg64_64>(cp) ||
           cfg_matches<Cfg64_32>(cp) || cfg_matches<Cfg32_32>(cp) || cfg_matches<Cfg32_16>(cp);
}

template <class CFG, int NWG, int AW, int SUBS, bool BF>
static int dispatch_r2(const ChainParams &cp, const TcArgs &a, int r, int sms, cudaStream_t s)
{
    if (cp.is_final) return launch_one<CFG, 1, true, NWG, AW, SUBS, BF>(cp, a, sms, s);
    switch (r) {
    case 8: return launch_one<CFG, 8, false, NWG, AW, SUBS, BF>(cp, a, sms, s);
    case 4: return launch_one<CFG, 4, false, NWG, AW, SUBS, BF>(cp, a, sms, s);
    case 2: return launch_one<CFG, 2, false, NWG, AW, SUBS, BF>(cp, a, sms, s);
    default: return launch_one<CFG, 1, false, NWG, AW, SUBS, BF>(cp, a, sms, s);   // r == 1, or plane_eq only
    }
}
template <class CFG, int NWG, int AW = CFG::kmax(), int SUBS = 1>
static int dispatch_r(const ChainParams &cp, const TcArgs &a, int r, int sms, cudaStream_t s)
{
    return a.bf16_io ? dispatch_r2<CFG, NWG, AW, SUBS, true>(cp, a, r, sms, s) : dispatch_r2<CFG, NWG, AW, SUBS, false>(cp, a, r, sms, s);
}

// r = 0: no LPG (reduction only). Returns BTS_ERR_UNSUPPORTED when the chain is not covered.
bool head_fwd_mma_wanted(const ChainParams &cp, const TcArgs &a);                   // chain_head_mma.cu
int launch_head_fwd_mma(const ChainParams &cp, const TcArgs &a, int sms, cudaStream_t s);

int launch_chain_tc_fwd(const ChainParams &cp, const TcArgs &a, int r, int sms, cudaStream_t s)
{
    // the sigmoid head in fp32: register-operand kernel on warp-level MMAs (chain_head_mma.cu; BTS_HEAD_MMA=0 disables)
    if (r == 0 && head_fwd_mma_wanted(cp, a)) return launch_head_fwd_mma(cp, a, sms, s);
    if (cfg_matches<Cfg128_128>(cp)) return dispatch_r<Cfg128_128, 2, 64, kWideSubs>(cp, a, r, sms, s);
    if (cfg_matches<Cfg128_64>(cp)) return dispatch_r<Cfg128_64, 2, 64, kWideSubs>(cp, a, r, sms, s);
    if (cfg_matches<Cfg64_64>(cp)) return dispatch_r<Cfg64_64, 2>(cp, a, r, sms, s);
    if (cfg_matches<Cfg64_32>(cp)) return dispatch_r<Cfg64_32, 3>(cp, a, r, sms, s);
    if (cfg_matches<Cfg32_32>(cp)) return dispatch_r<Cfg32_32, 4>(cp, a, r, sms, s);
    if (cfg_matches<Cfg32_16>(cp)) return dispatch_r<Cfg32_16, 4>(cp, a, r, sms, s);
    return set_error(BTS_ERR_UNSUPPORTED, "chain not covered by the tensor-core path");
}

}  // namespace bts

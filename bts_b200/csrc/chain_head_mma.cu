// chain_head_mma.cu -- the full-resolution head 32 -> 16 -> 8 -> 1 (+ sigmoid) on sm_100a with its operands in
// REGISTERS: warp-level tensor-core MMAs (mma.sync m16n8k8 tf32 + m16n8k16 bf16) instead of tcgen05. ONE-PASS backward
// (head_bwd_mma_kernel) and forward (head_fwd_mma_kernel).
//
// reduction_1x1(is_final=True) and its autograd backward (pytorch/bts.py:108-110, 124-126, 136; SURVEY.md 8(a) rows
// a3 / a8): the backward is the dominant call of the path (260 B per full-resolution pixel: x and grad read once, grad_x
// written once), the forward the second largest (132 B per pixel).
//
// Why not tcgen05 here (chain_bwd_fused.cu is that kernel, and stays as the alternative route): the chain's matrices are
// 32x16, 16x8 and 8x1. tcgen05 wants M >= 64 rows per operand and operands in shared memory / TMEM, so every pixel's
// activations and deltas are staged (swizzled 4-byte stores, TMEM round trips, hand-offs between warps) -- 34 % of that
// kernel's samples -- at 8 warps per SM. The warp-level MMA takes A and B from registers and its accumulator layout IS the
// next product's operand layout:
//   * px-on-M products (rows = 16 pixels): C[px g][ch 2t, 2t+1] feeds the next layer as A[px g][k = t, t+4] once the
//     weights' K rows are permuted to (2t, 2t+1) -- forward chain and dgrad run register to register;
//   * the same C fragment is a B operand [k = ch][n = px] of a px-on-N product: grad_x^T = W0^T * delta1^T comes out as
//     [channel g][pixel 2t, 2t+1] = 8-byte stores along the pixel axis of NCHW;
//   * weight gradients need the pixel axis as K: an 8x8 register transpose of 32-bit words (two movmatrix on the 16-bit
//     halves, tests/cuda/probe_hmma.cu) turns C[px g][ch 2t..] into [ch g][px 2t, 2t+1], and the x operand comes from the
//     warp's shared-memory tile with one 64-bit load.
// No TMEM, no staging buffers, no barrier inside the loop: a warp owns 32 pixels at a time, its tile arrives by cp.async
// into a warp-private ring, 16 warps per SM. Measured (probe_hmma): 0.5 m16n8k8 tf32 MMAs / clock / SM; the accumulator
// adds truncated addends (like tcgen05), so weight-gradient sums leave the tensor core after every 32 pixels (8 steps)
// and continue in fp32 registers with round-to-nearest adds. Products are fp32-accurate in TWO instructions (FragA / FragB
// below); the bf16 twin (BF) reads bf16 tiles and writes bf16 tensors with the same arithmetic.
// nyu_b8: backward 159 us = 62 % of the measured HBM peak (tcgen05 kernel: 190 us), forward 62 us = 81 % (72 us).
#include <cstdlib>
#include <type_traits>

#include "chain_tc.cuh"

namespace bts {
namespace {

constexpr int kGP = 32;                    // pixels per warp and iteration
constexpr int kXS = 40;                    // floats between channel rows of a tile: 32 pixels + 8 (conflict-free fragments)
constexpr int kTileFloats = 32 * kXS;
constexpr int kNW0 = 16 * 32, kNW1 = 8 * 16, kNW2 = 8;
constexpr int kSumFloats = 656;            // 648 weight-gradient sums, padded to 16 bytes

struct MmaArgs {
    TcArgs a;
    int atomic_w;                          // add the CTA's sums to a.gw[] (null gw: skip)
};

__device__ __forceinline__ void hmma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// m16n8k16 bf16: twice the K of the tf32 shape at the same issue rate (probe_hmma: 0.5 / clock / SM for both)
__device__ __forceinline__ void hmma_bf(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// d = a * b (no accumulator input: the zeros are the zero register, no moves)
__device__ __forceinline__ void hmma_bf_z(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.f));
}
// {hi : lo} bf16 pair, both rounded to nearest even
__device__ __forceinline__ uint32_t bfpair(uint32_t lo, uint32_t hi)
{
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(__uint_as_float(hi)), "f"(__uint_as_float(lo)));
    return r;
}
// ---- fp32-accurate products out of two tensor-core instructions per K = 8.
// v = hi + lo, hi = v with the 13 bits the tf32 tensor core ignores cleared, lo = v - hi (exact). a * b = hi_a hi_b
// (one tf32 MMA) + lo_a hi_b + hi_a lo_b (+ lo_a lo_b ~ 2^-22, dropped). The two small products only need ~8 bits
// (they are 2^-11 of the result each), so they run as ONE bf16 m16n8k16 MMA: its 16 K slots hold the 8 lo_a x hi_b
// pairs and the 8 hi_a x lo_b pairs (k = 2t, 2t+1 <-> tf32 k = t, t+4; k = 2t+8, 2t+9 likewise for the second product).
// Error of the bf16 roundings: 2^-9 * 2^-11 = 2^-20 of a term, unbiased.
struct FragA {
    uint32_t h[4];                         // tf32 A fragment (row g | g+8, k = t | t+4) of the hi parts
    uint32_t s[4];                         // bf16 A fragment: {lo(k=t), lo(k=t+4)} rows g, g+8; {hi, hi} rows g, g+8
    uint32_t l[4];                         // tf32 A fragment of the remainders (the exact three-instruction variant only)
};
struct FragB {
    uint32_t h0, h1;                       // tf32 B fragment (k = t | t+4, n = g) of the hi parts
    uint32_t s0, s1;                       // bf16 B fragment: {hi(k=t), hi(k=t+4)}, {lo, lo}
    uint32_t l0, l1;                       // tf32 B fragment of the remainders (exact variant only)
};
__device__ __forceinline__ void split1(float v, uint32_t &h, uint32_t &l)
{
    h = __float_as_uint(v) & 0xFFFFE000u;
    l = __float_as_uint(v - __uint_as_float(h));
}
// (v0 .. v3) in fragment order: (row g, k=t), (row g+8, k=t), (row g, k=t+4), (row g+8, k=t+4)
__device__ __forceinline__ void make_a(float v0, float v1, float v2, float v3, FragA &f)
{
    split1(v0, f.h[0], f.l[0]);
    split1(v1, f.h[1], f.l[1]);
    split1(v2, f.h[2], f.l[2]);
    split1(v3, f.h[3], f.l[3]);
    f.s[0] = bfpair(f.l[0], f.l[2]);
    f.s[1] = bfpair(f.l[1], f.l[3]);
    f.s[2] = bfpair(f.h[0], f.h[2]);
    f.s[3] = bfpair(f.h[1], f.h[3]);
}
__device__ __forceinline__ void make_b(float v0, float v1, FragB &f)
{
    split1(v0, f.h0, f.l0);
    split1(v1, f.h1, f.l1);
    f.s0 = bfpair(f.h0, f.h1);
    f.s1 = bfpair(f.l0, f.l1);
}
__device__ __forceinline__ FragB frag_b(const float4 &w)
{
    FragB f;
    f.h0 = __float_as_uint(w.x); f.h1 = __float_as_uint(w.y); f.s0 = __float_as_uint(w.z); f.s1 = __float_as_uint(w.w);
    f.l0 = f.l1 = 0u;
    return f;
}
// exact variant: the table entry carries the remainder words instead of the bf16 pairs
__device__ __forceinline__ FragB frag_b_exact(const float4 &w)
{
    FragB f;
    f.h0 = __float_as_uint(w.x); f.h1 = __float_as_uint(w.y); f.l0 = __float_as_uint(w.z); f.l1 = __float_as_uint(w.w);
    f.s0 = f.s1 = 0u;
    return f;
}
__device__ __forceinline__ float4 with_lo_exact(float b0, float b1) { return make_float4(tf32_hi(b0), tf32_hi(b1), tf32_lo(b0), tf32_lo(b1)); }
// d (+)= a * b in three tf32 instructions (plain 3xTF32: ~5e-7 of peak instead of ~3e-6), small products first
template <bool ZERO = false>
__device__ __forceinline__ void mma3_small(float (&d)[4], const FragA &a, const FragB &b)
{
    if constexpr (ZERO) {
        d[0] = d[1] = d[2] = d[3] = 0.f;
    }
    hmma(d, a.l, b.h0, b.h1);
    hmma(d, a.h, b.l0, b.l1);
}
// d (+)= a * b in two instructions, small products first. Callers issue the first halves of several independent
// accumulators, then the second halves: back-to-back MMAs into one accumulator wait for each other (the asm statements
// are volatile, the compiler keeps their order).
template <bool ZERO = false>
__device__ __forceinline__ void mma2_small(float (&d)[4], const FragA &a, const FragB &b)
{
    if constexpr (ZERO) hmma_bf_z(d, a.s, b.s0, b.s1);
    else hmma_bf(d, a.s, b.s0, b.s1);
}
__device__ __forceinline__ void mma2_main(float (&d)[4], const FragA &a, const FragB &b) { hmma(d, a.h, b.h0, b.h1); }
// sigmoid with the hardware exponential and reciprocal (relative error ~2e-7; the backward is held to 1e-5)
__device__ __forceinline__ float sigmoid_fast(float t)
{
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t * -1.4426950408889634f));
    return rcp_fast(1.f + e);
}
// d * ELU'(t) through a = ELU(t): a > 0 ? d : d * (a + 1)
__device__ __forceinline__ float mul_elu_grad(float d, float a) { return a > 0.f ? d : fmaf(d, a, d); }
// 8x8 transpose of 32-bit words: in: thread (g, t) holds E[g][2t], E[g][2t+1]; out: E[2t][g], E[2t+1][g]
__device__ __forceinline__ void tr8x8(float e0, float e1, float &w0, float &w1)
{
    const uint32_t u0 = __float_as_uint(e0), u1 = __float_as_uint(e1);
    const uint32_t hi = __byte_perm(u0, u1, 0x7632), lo = __byte_perm(u0, u1, 0x5410);
    uint32_t th, tl;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(th) : "r"(hi));
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(tl) : "r"(lo));
    w0 = __uint_as_float(__byte_perm(tl, th, 0x5410));
    w1 = __uint_as_float(__byte_perm(tl, th, 0x7632));
}
__device__ __forceinline__ uint32_t movm(uint32_t v)                 // 8x8 transpose of 16-bit elements (a bf16 pair per thread)
{
    uint32_t r;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ void tr8x8(uint32_t e0, uint32_t e1, uint32_t &w0, uint32_t &w1)
{
    float a, b;
    tr8x8(__uint_as_float(e0), __uint_as_float(e1), a, b);
    w0 = __float_as_uint(a);
    w1 = __float_as_uint(b);
}
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ float4 with_lo(float b0, float b1)
{
    FragB f;
    make_b(b0, b1, f);
    return make_float4(__uint_as_float(f.h0), __uint_as_float(f.h1), __uint_as_float(f.s0), __uint_as_float(f.s1));
}

// ---- tile / tensor access for fp32 and for the bf16 twin (BF: x, grad, grad_x, out are bf16 tensors; a tile row is 32
// two-byte pixels at a stride of kXS elements, both fragment patterns stay bank-conflict-free). A bf16 value is exact in
// tf32: its remainder is zero (the remainder halves of the bf16 operands are zero registers) and its bf16 pair is a
// byte permute of the raw halves.
template <bool BF>
__device__ __forceinline__ void x_frag_a(const void *tile, int ch, int px, FragA &f)   // rows px, px+8; k = ch, ch+4
{
    if constexpr (!BF) {
        const float *q = static_cast<const float *>(tile) + ch * kXS + px;
        make_a(q[0], q[8], q[4 * kXS], q[4 * kXS + 8], f);
    } else {
        const unsigned short *q = static_cast<const unsigned short *>(tile) + ch * kXS + px;
        f.h[0] = (uint32_t)q[0] << 16; f.h[1] = (uint32_t)q[8] << 16; f.h[2] = (uint32_t)q[4 * kXS] << 16; f.h[3] = (uint32_t)q[4 * kXS + 8] << 16;
        f.l[0] = f.l[1] = f.l[2] = f.l[3] = 0u;
        f.s[0] = f.s[1] = 0u;
        f.s[2] = __byte_perm(f.h[0], f.h[2], 0x7632);
        f.s[3] = __byte_perm(f.h[1], f.h[3], 0x7632);
    }
}
template <bool BF>
__device__ __forceinline__ void x_frag_b(const void *tile, int ch, int px, FragB &f)   // k = px, px+1 (a pair); n = ch
{
    if constexpr (!BF) {
        const float2 xv = *reinterpret_cast<const float2 *>(static_cast<const float *>(tile) + ch * kXS + px);
        make_b(xv.x, xv.y, f);
    } else {
        const uint32_t w = *reinterpret_cast<const uint32_t *>(static_cast<const unsigned short *>(tile) + ch * kXS + px);
        f.h0 = w << 16; f.h1 = w & 0xFFFF0000u;
        f.l0 = f.l1 = 0u;
        f.s0 = w;
        f.s1 = 0u;
    }
}

template <bool EXACT, bool ZERO>
__device__ __forceinline__ void mma_small(float (&d)[4], const FragA &a, const FragB &b)
{
    if constexpr (EXACT) mma3_small<ZERO>(d, a, b);
    else mma2_small<ZERO>(d, a, b);
}

// Tuning / accuracy switches (A/B builds): the forward-chain products of the backward in plain 3xTF32.
#ifndef BTS_MMA_EXACT_F
#define BTS_MMA_EXACT_F 0
#endif
#if BTS_MMA_EXACT_F
#define MMA_F_TABLE with_lo_exact
#define MMA_F_FRAG frag_b_exact
#define MMA_F_SMALL mma3_small
#else
#define MMA_F_TABLE with_lo
#define MMA_F_FRAG frag_b
#define MMA_F_SMALL mma2_small
#endif

template <int WARPS, int NSTAGE, int MINB, bool BF>
__global__ void __launch_bounds__(WARPS * 32, MINB) head_bwd_mma_kernel(const __grid_constant__ MmaArgs ma)
{
    const TcArgs &a = ma.a;
    extern __shared__ __align__(16) float smem[];
    // weight fragments in the order the lanes read them (one 128-bit load per fragment pair, hi and remainder together)
    float4 *f0b = reinterpret_cast<float4 *>(smem);               // [4 k-steps][2 n-tiles][32 lanes]  W0 as B of x * W0^T
    float4 *f1b = f0b + 4 * 2 * 32;                               // [2][32]                            W1 as B of a1 * W1^T
    float4 *b1b = f1b + 2 * 32;                                   // [2 n-tiles][32]                    W1 as B of d2 * W1
    float4 *b0a = b1b + 2 * 32;                                   // [2 m-tiles][2 k-steps][32][hi, lo] W0^T as A of W0^T * d1^T
    float *sums = reinterpret_cast<float *>(b0a + 2 * 2 * 32 * 2);   // the CTA's weight-gradient sums
    float *tiles = sums + kSumFloats;                             // [WARPS][NSTAGE][32 channels][kXS]

    const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int warp = uniform(tid >> 5);
    const float *W0 = a.W[0], *W1 = a.W[1], *W2 = a.W[2];
    const unsigned npx = (unsigned)a.npx, hw = (unsigned)a.hw;
    const int ngroups = (int)((npx + kGP - 1) / kGP);
    const int stride = (int)gridDim.x * WARPS;
    const int first = (int)blockIdx.x * WARPS + warp;
    float *mytiles = tiles + (size_t)warp * NSTAGE * kTileFloats;

    // Positions advance by the same number of pixels every iteration: one division per kernel, then adds with a wrap.
    struct Pos {
        unsigned gp, b, p;                                        // flat pixel, image, pixel inside the image
    };
    const unsigned step = (unsigned)stride * kGP;
    const unsigned dq = step / hw, dr = step - dq * hw;
    auto pos_at = [&](unsigned gp) {
        Pos s;
        s.gp = gp;
        s.b = gp / hw;
        s.p = gp - s.b * hw;
        return s;
    };
    auto pos_advance = [&](Pos &s) {
        s.gp += step;
        s.b += dq;
        s.p += dr;
        if (s.p >= hw) {
            s.p -= hw;
            ++s.b;
        }
    };
    // the tile of the group at `q` (this lane's pixel group) -> ring slot st. fp32: pixels 4 (lane % 8).. of channels
    // lane / 8 + 4 i; bf16: pixels 8 (lane % 4).. of channels lane / 4 + 8 i (16 bytes per copy either way)
    using TIO = std::conditional_t<BF, unsigned short, float>;
    constexpr int PXG = BF ? 8 : 4, NCP = BF ? 4 : 8, LPR = 32 / PXG;      // pixels per copy, copies per lane, lanes per row
    const TIO *const xin = reinterpret_cast<const TIO *>(a.x);
    const size_t hwrow = (size_t)hw * (32 / NCP);                          // elements between this lane's successive rows
    auto prefetch = [&](const Pos &q, int st) {
        const bool ok = q.gp < npx;
        const TIO *src = xin + ((size_t)(ok ? q.b : 0u) * 32 + (lane / LPR)) * hw + (ok ? q.p : 0u);
        const uint32_t dst = smem_u32(reinterpret_cast<TIO *>(mytiles + st * kTileFloats) + (lane / LPR) * kXS + PXG * (lane % LPR));
#pragma unroll
        for (int i = 0; i < NCP; ++i) cp_async16(dst + i * (32 / NCP) * kXS * (int)sizeof(TIO), src + i * hwrow, ok);
        cp_async_commit();
    };
    // upstream gradient of pixels 8q + g of the group that starts at gp0 (0 beyond the end: such pixels add nothing anywhere)
    const TIO *const gin = reinterpret_cast<const TIO *>(a.grad);
    auto load_grad = [&](unsigned gp0, float (&gv)[4]) {
        if (gp0 + kGP <= npx) {
            const TIO *gsrc = gin + gp0 + g;
#pragma unroll
            for (int q = 0; q < 4; ++q) gv[q] = io_load<TIO>(gsrc + 8 * q);
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const unsigned gp = gp0 + 8u * q + g;
                gv[q] = (gp0 < npx && gp < npx) ? io_load<TIO>(gin + gp) : 0.f;
            }
        }
    };

    Pos quad = pos_at((unsigned)first * kGP + (unsigned)(PXG * (lane % LPR)));  // this lane's quad of the next group to fetch
    Pos base = pos_at((unsigned)first * kGP);                     // first pixel of the current group
#pragma unroll
    for (int k = 0; k < NSTAGE - 1; ++k) {
        prefetch(quad, k);
        pos_advance(quad);
    }
    float gn[4];
    load_grad(base.gp, gn);

    // ---- fragment tables (while the first tiles are in flight)
    for (int e = tid; e < 4 * 2 * 32; e += WARPS * 32) {
        const int ln = e & 31, j = (e >> 5) & 1, s = e >> 6, gg = ln >> 2, tt = ln & 3;
        f0b[e] = MMA_F_TABLE(__ldg(W0 + (8 * j + gg) * 32 + 8 * s + tt), __ldg(W0 + (8 * j + gg) * 32 + 8 * s + tt + 4));
    }
    for (int e = tid; e < 2 * 32; e += WARPS * 32) {
        const int ln = e & 31, s = e >> 5, gg = ln >> 2, tt = ln & 3;
        f1b[e] = MMA_F_TABLE(__ldg(W1 + gg * 16 + 8 * s + 2 * tt), __ldg(W1 + gg * 16 + 8 * s + 2 * tt + 1));
        b1b[e] = with_lo(__ldg(W1 + (2 * tt) * 16 + 8 * s + gg), __ldg(W1 + (2 * tt + 1) * 16 + 8 * s + gg));
    }
    for (int e = tid; e < 2 * 2 * 32; e += WARPS * 32) {
        const int ln = e & 31, s = (e >> 5) & 1, i = e >> 6, gg = ln >> 2, tt = ln & 3;
        const float v0 = __ldg(W0 + (8 * s + 2 * tt) * 32 + 16 * i + gg), v1 = __ldg(W0 + (8 * s + 2 * tt) * 32 + 16 * i + gg + 8);
        const float v2 = __ldg(W0 + (8 * s + 2 * tt + 1) * 32 + 16 * i + gg), v3 = __ldg(W0 + (8 * s + 2 * tt + 1) * 32 + 16 * i + gg + 8);
        FragA f;
        make_a(v0, v1, v2, v3, f);
        b0a[2 * e] = make_float4(__uint_as_float(f.h[0]), __uint_as_float(f.h[1]), __uint_as_float(f.h[2]), __uint_as_float(f.h[3]));
        b0a[2 * e + 1] = make_float4(__uint_as_float(f.s[0]), __uint_as_float(f.s[1]), __uint_as_float(f.s[2]), __uint_as_float(f.s[3]));
    }
    for (int e = tid; e < kSumFloats; e += WARPS * 32) sums[e] = 0.f;
    const float w2a = __ldg(W2 + 2 * t), w2b = __ldg(W2 + 2 * t + 1);
    __syncthreads();

    // weight-gradient master sums (fp32, round-to-nearest): dW0[o][c], dW1[o][c], dW2[c] in fragment order
    float m0[4][4], m1[4], m2[2];
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int q = 0; q < 4; ++q) m0[n][q] = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) m1[q] = 0.f;
    m2[0] = m2[1] = 0.f;

    int slot = 0;
#pragma unroll 1
    for (int G = first; G < ngroups; G += stride) {
        cp_async_wait<NSTAGE - 2>();
        __syncwarp();
        {
            int pst = slot + NSTAGE - 1;
            if (pst >= NSTAGE) pst -= NSTAGE;
            prefetch(quad, pst);
            pos_advance(quad);
        }
        float gc[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) gc[q] = gn[q];
        load_grad(base.gp + step, gn);
        const void *xt = mytiles + slot * kTileFloats;
        if (++slot == NSTAGE) slot = 0;

        // ---- layer 0 (px on M): pre1[16 px x 16] per m-tile = x[16 px x 32] * W0^T; a1 = ELU
        float a1[2][2][4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            FragA fa[2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) x_frag_a<BF>(xt, 8 * s + t, 16 * mt + g, fa[mt]);
            FragB fb[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) fb[j] = MMA_F_FRAG(f0b[(s * 2 + j) * 32 + lane]);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    if (s == 0) MMA_F_SMALL<true>(a1[mt][j], fa[mt], fb[j]);
                    else MMA_F_SMALL(a1[mt][j], fa[mt], fb[j]);
                }
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma2_main(a1[mt][j], fa[mt], fb[j]);
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) a1[mt][j][q] = elu_fast(a1[mt][j][q]);

        // ---- layer 1: the accumulator fragment (px g / g+8, channels 8s + 2t, 2t+1) is the A fragment (k = t, t+4)
        float a2[2][4];
        FragA fa1[2][2];                                          // a1 as A operand [k-step = its n-tile][m-tile]: kept for dW1
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const FragB fb = MMA_F_FRAG(f1b[s * 32 + lane]);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) make_a(a1[mt][s][0], a1[mt][s][2], a1[mt][s][1], a1[mt][s][3], fa1[s][mt]);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                if (s == 0) MMA_F_SMALL<true>(a2[mt], fa1[s][mt], fb);
                else MMA_F_SMALL(a2[mt], fa1[s][mt], fb);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) mma2_main(a2[mt], fa1[s][mt], fb);
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int q = 0; q < 4; ++q) a2[mt][q] = elu_fast(a2[mt][q]);

        // ---- layer 2 + sigmoid on the SIMT side: z = W2 . a2 (sum over the four t lanes), dz = g * s * (1 - s)
        float d2[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                float zp = fmaf(a2[mt][2 * r + 1], w2b, a2[mt][2 * r] * w2a);
                zp += __shfl_xor_sync(0xffffffffu, zp, 1);
                zp += __shfl_xor_sync(0xffffffffu, zp, 2);
                const float sg = sigmoid_fast(zp);
                const float dz = gc[2 * mt + r] * (sg - sg * sg);
                m2[0] = fmaf(dz, a2[mt][2 * r], m2[0]);
                m2[1] = fmaf(dz, a2[mt][2 * r + 1], m2[1]);
                d2[mt][2 * r] = mul_elu_grad(dz * w2a, a2[mt][2 * r]);
                d2[mt][2 * r + 1] = mul_elu_grad(dz * w2b, a2[mt][2 * r + 1]);
            }

        // The weight gradients need the pixel axis as K: operands [channel g][pixel 2t, 2t+1] = 8x8 transposes of what the
        // chain's fragments hold as [pixel g][channel 2t, 2t+1]. Nothing is split twice: the fragments built for the chain
        // products already carry the hi words and the bf16 pairs {ch 2t, ch 2t+1} of the hi parts and of the remainders;
        // a pair register IS an 8x8 matrix of 16-bit elements, so ONE movmatrix yields the transposed pair {px 2t, px 2t+1},
        // and the hi words take the 32-bit transpose (two movmatrix on the halves).
        // (mt, r): the eight pixels of row half r of m-tile mt = one k-step.
        float g0[4][4], g1[4];

        // ---- delta1 = (delta2 * W1) . ELU'(a1)    (px on M, K = 8: one k-step)
        float d1[2][2][4];
        {
            FragA fa2[2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) make_a(d2[mt][0], d2[mt][2], d2[mt][1], d2[mt][3], fa2[mt]);
            FragB fb[2];
            float acc[2][2][4];
#pragma unroll
            for (int j = 0; j < 2; ++j) fb[j] = frag_b(b1b[j * 32 + lane]);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma2_small<true>(acc[mt][j], fa2[mt], fb[j]);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma2_main(acc[mt][j], fa2[mt], fb[j]);
            // ---- dW1[o][c] += delta2[px][o] * a1[px][c]: A = a1^T (m = c: rows g from n-tile 0, g+8 from n-tile 1), B = delta2^T
            // (fragment order of make_a(c0, c2, c1, c3): h[r], h[2+r] = channels 2t, 2t+1 of row half r; s[r] / s[2+r] = the
            // remainder / hi pair of row half r)
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    FragA fc;
                    FragB fe;
                    tr8x8(fa1[0][mt].h[r], fa1[0][mt].h[2 + r], fc.h[0], fc.h[2]);
                    tr8x8(fa1[1][mt].h[r], fa1[1][mt].h[2 + r], fc.h[1], fc.h[3]);
                    fc.s[0] = movm(fa1[0][mt].s[r]);
                    fc.s[1] = movm(fa1[1][mt].s[r]);
                    fc.s[2] = movm(fa1[0][mt].s[2 + r]);
                    fc.s[3] = movm(fa1[1][mt].s[2 + r]);
                    tr8x8(fa2[mt].h[r], fa2[mt].h[2 + r], fe.h0, fe.h1);
                    fe.s0 = movm(fa2[mt].s[2 + r]);
                    fe.s1 = movm(fa2[mt].s[r]);
                    if (mt == 0 && r == 0) mma2_small<true>(g1, fc, fe);
                    else mma2_small(g1, fc, fe);
                    mma2_main(g1, fc, fe);
                }
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                    for (int q = 0; q < 4; ++q) d1[mt][j][q] = mul_elu_grad(acc[mt][j][q], a1[mt][j][q]);
        }

        // ---- grad_x^T[32 ch x 32 px] = W0^T[32 x 16] * delta1^T (px on N): delta1's accumulator fragment is the B fragment
        FragB fd[2][2][2];                                        // delta1 as B operand: [m-tile][channel k-step][row half]
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                make_b(d1[mt][j][0], d1[mt][j][1], fd[mt][j][0]);
                make_b(d1[mt][j][2], d1[mt][j][3], fd[mt][j][1]);
            }
        if (a.grad_x != nullptr) {
            // whole group inside one image and inside the tensor: four row pointers, immediate offsets
            const bool fast = base.gp + kGP <= npx && base.p + kGP <= hw;
            TIO *const gxo = reinterpret_cast<TIO *>(a.grad_x);
            TIO *row0 = gxo + ((size_t)base.b * 32 + g) * hw + base.p + 2 * t;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                FragA fw[2];
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    const float4 vh = b0a[((i * 2 + s) * 32 + lane) * 2], vs = b0a[((i * 2 + s) * 32 + lane) * 2 + 1];
                    fw[s].h[0] = __float_as_uint(vh.x); fw[s].h[1] = __float_as_uint(vh.y); fw[s].h[2] = __float_as_uint(vh.z); fw[s].h[3] = __float_as_uint(vh.w);
                    fw[s].s[0] = __float_as_uint(vs.x); fw[s].s[1] = __float_as_uint(vs.y); fw[s].s[2] = __float_as_uint(vs.z); fw[s].s[3] = __float_as_uint(vs.w);
                }
                float acc[4][4];
                // n-tile n: 8 pixels = rows r = n % 2 of m-tile n / 2; four independent accumulators per pass
#pragma unroll
                for (int n = 0; n < 4; ++n) hmma_bf_z(acc[n], fw[0].s, fd[n >> 1][0][n & 1].s0, fd[n >> 1][0][n & 1].s1);
#pragma unroll
                for (int n = 0; n < 4; ++n) hmma_bf(acc[n], fw[1].s, fd[n >> 1][1][n & 1].s0, fd[n >> 1][1][n & 1].s1);
#pragma unroll
                for (int n = 0; n < 4; ++n) hmma(acc[n], fw[0].h, fd[n >> 1][0][n & 1].h0, fd[n >> 1][0][n & 1].h1);
#pragma unroll
                for (int n = 0; n < 4; ++n) hmma(acc[n], fw[1].h, fd[n >> 1][1][n & 1].h0, fd[n >> 1][1][n & 1].h1);
                if (fast) {
                    TIO *ra = row0 + (size_t)(16 * i) * hw, *rb = ra + (size_t)8 * hw;
#pragma unroll
                    for (int n = 0; n < 4; ++n) {
                        io_store2(ra + 8 * n, acc[n][0], acc[n][1]);
                        io_store2(rb + 8 * n, acc[n][2], acc[n][3]);
                    }
                } else {
#pragma unroll 1
                    for (int n = 0; n < 4; ++n) {
                        const unsigned off = 8u * n + 2u * t;
                        if (base.gp + off < npx) {
                            unsigned b = base.b, p = base.p + off;
                            if (p >= hw) {
                                b += p / hw;
                                p %= hw;
                            }
                            TIO *dst = gxo + ((size_t)b * 32 + 16 * i + g) * hw + p;
                            // (acc[n] with a run-time n: select instead of indexing, the array stays in registers)
                            float v0 = acc[0][0], v1 = acc[0][1], v2 = acc[0][2], v3 = acc[0][3];
#pragma unroll
                            for (int m = 1; m < 4; ++m)
                                if (n == m) { v0 = acc[m][0]; v1 = acc[m][1]; v2 = acc[m][2]; v3 = acc[m][3]; }
                            io_store2(dst, v0, v1);
                            io_store2(dst + (size_t)8 * hw, v2, v3);
                        }
                    }
                }
            }
        }

        // ---- dW0[o][c] += delta1[px][o] * x[px][c]: A = delta1^T (m = o: rows g from n-tile 0, g+8 from n-tile 1) out of the
        // fragments of the product above, B = x (n = c) from the tile, one 64-bit load per fragment
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                FragA fa;
                FragB fb[4];
                tr8x8(fd[mt][0][r].h0, fd[mt][0][r].h1, fa.h[0], fa.h[2]);
                tr8x8(fd[mt][1][r].h0, fd[mt][1][r].h1, fa.h[1], fa.h[3]);
                fa.s[0] = movm(fd[mt][0][r].s1);
                fa.s[1] = movm(fd[mt][1][r].s1);
                fa.s[2] = movm(fd[mt][0][r].s0);
                fa.s[3] = movm(fd[mt][1][r].s0);
#pragma unroll
                for (int n = 0; n < 4; ++n) x_frag_b<BF>(xt, 8 * n + g, 16 * mt + 8 * r + 2 * t, fb[n]);
#pragma unroll
                for (int n = 0; n < 4; ++n) {
                    if (mt == 0 && r == 0) mma2_small<true>(g0[n], fa, fb[n]);
                    else mma2_small(g0[n], fa, fb[n]);
                }
#pragma unroll
                for (int n = 0; n < 4; ++n) mma2_main(g0[n], fa, fb[n]);
            }
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 4; ++q) m0[n][q] += g0[n][q];
#pragma unroll
        for (int q = 0; q < 4; ++q) m1[q] += g1[q];
        pos_advance(base);
    }
    cp_async_wait<0>();

    // ---- the warps' sums -> the CTA's -> grad_W (fp32 atomics; grad_W is an accumulate target by contract)
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int q = 0; q < 4; ++q) atomicAdd(sums + (g + (q >> 1) * 8) * 32 + 8 * n + 2 * t + (q & 1), m0[n][q]);
#pragma unroll
    for (int q = 0; q < 4; ++q) atomicAdd(sums + kNW0 + (2 * t + (q & 1)) * 16 + g + (q >> 1) * 8, m1[q]);
#pragma unroll
    for (int e = 0; e < 2; ++e) {
        float v = m2[e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (g == 0) atomicAdd(sums + kNW0 + kNW1 + 2 * t + e, v);
    }
    __syncthreads();
    if (ma.atomic_w) {
        // CTAs start at different elements so that they do not all hit the same address at once
        const int shift = (int)((blockIdx.x * 67u) % (unsigned)(kNW0 + kNW1 + kNW2));
        for (int i = tid; i < kNW0 + kNW1 + kNW2; i += WARPS * 32) {
            int e = i + shift;
            if (e >= kNW0 + kNW1 + kNW2) e -= kNW0 + kNW1 + kNW2;
            float *dst = e < kNW0 ? a.gw[0] + e : (e < kNW0 + kNW1 ? a.gw[1] + (e - kNW0) : a.gw[2] + (e - kNW0 - kNW1));
            atomicAdd(dst, sums[e]);
        }
    }
}


// ---- forward: out[px] = sigmoid(W2 . ELU(W1 ELU(W0 x))). The first half of the backward's iteration; memory-bound
// (~40 MMAs and ~400 other instructions per 32 pixels), 16 warps x one tile ahead. EXACT: plain three-instruction 3xTF32
// products (output within 8e-7 of the fp64 oracle's peak instead of 3.6e-6, 72 instead of 61 us at nyu_b8: the kernel is
// latency-bound on its dependent MMA chain, not idle under the copies); BTS_HEAD_EXACT=1 selects it.
template <int WARPS, int NSTAGE, int MINB, bool EXACT, bool BF>
__global__ void __launch_bounds__(WARPS * 32, MINB) head_fwd_mma_kernel(const __grid_constant__ MmaArgs ma)
{
    const TcArgs &a = ma.a;
    extern __shared__ __align__(16) float smem[];
    float4 *f0b = reinterpret_cast<float4 *>(smem);               // [4 k-steps][2 n-tiles][32 lanes]
    float4 *f1b = f0b + 4 * 2 * 32;                               // [2][32]
    float *tiles = reinterpret_cast<float *>(f1b + 2 * 32);       // [WARPS][NSTAGE][32 channels][kXS]

    const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int warp = uniform(tid >> 5);
    const float *W0 = a.W[0], *W1 = a.W[1], *W2 = a.W[2];
    const unsigned npx = (unsigned)a.npx, hw = (unsigned)a.hw;
    const int ngroups = (int)((npx + kGP - 1) / kGP);
    const int stride = (int)gridDim.x * WARPS;
    const int first = (int)blockIdx.x * WARPS + warp;
    float *mytiles = tiles + (size_t)warp * NSTAGE * kTileFloats;

    const unsigned step = (unsigned)stride * kGP;
    const unsigned dq = step / hw, dr = step - dq * hw;
    using TIO = std::conditional_t<BF, unsigned short, float>;
    constexpr int PXG = BF ? 8 : 4, NCP = BF ? 4 : 8, LPR = 32 / PXG;      // pixels per copy, copies per lane, lanes per row
    unsigned q_gp = (unsigned)first * kGP + (unsigned)(PXG * (lane % LPR)), q_b = q_gp / hw, q_p = q_gp - q_b * hw;
    const TIO *const xin = reinterpret_cast<const TIO *>(a.x);
    const size_t hwrow = (size_t)hw * (32 / NCP);
    auto prefetch = [&](int st) {
        const bool ok = q_gp < npx;
        const TIO *src = xin + ((size_t)(ok ? q_b : 0u) * 32 + (lane / LPR)) * hw + (ok ? q_p : 0u);
        const uint32_t dst = smem_u32(reinterpret_cast<TIO *>(mytiles + st * kTileFloats) + (lane / LPR) * kXS + PXG * (lane % LPR));
#pragma unroll
        for (int i = 0; i < NCP; ++i) cp_async16(dst + i * (32 / NCP) * kXS * (int)sizeof(TIO), src + i * hwrow, ok);
        cp_async_commit();
        q_gp += step;
        q_b += dq;
        q_p += dr;
        if (q_p >= hw) {
            q_p -= hw;
            ++q_b;
        }
    };
#pragma unroll
    for (int k = 0; k < NSTAGE - 1; ++k) prefetch(k);

    for (int e = tid; e < 4 * 2 * 32; e += WARPS * 32) {
        const int ln = e & 31, j = (e >> 5) & 1, s = e >> 6, gg = ln >> 2, tt = ln & 3;
        f0b[e] = (EXACT ? with_lo_exact : with_lo)(__ldg(W0 + (8 * j + gg) * 32 + 8 * s + tt), __ldg(W0 + (8 * j + gg) * 32 + 8 * s + tt + 4));
    }
    for (int e = tid; e < 2 * 32; e += WARPS * 32) {
        const int ln = e & 31, s = e >> 5, gg = ln >> 2, tt = ln & 3;
        f1b[e] = (EXACT ? with_lo_exact : with_lo)(__ldg(W1 + gg * 16 + 8 * s + 2 * tt), __ldg(W1 + gg * 16 + 8 * s + 2 * tt + 1));
    }
    const float w2a = __ldg(W2 + 2 * t), w2b = __ldg(W2 + 2 * t + 1);
    __syncthreads();

    int slot = 0;
#pragma unroll 1
    for (int G = first; G < ngroups; G += stride) {
        cp_async_wait<NSTAGE - 2>();
        __syncwarp();
        {
            int pst = slot + NSTAGE - 1;
            if (pst >= NSTAGE) pst -= NSTAGE;
            prefetch(pst);
        }
        const void *xt = mytiles + slot * kTileFloats;
        if (++slot == NSTAGE) slot = 0;

        float a1[2][2][4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            FragA fa[2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) x_frag_a<BF>(xt, 8 * s + t, 16 * mt + g, fa[mt]);
            FragB fb[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) fb[j] = EXACT ? frag_b_exact(f0b[(s * 2 + j) * 32 + lane]) : frag_b(f0b[(s * 2 + j) * 32 + lane]);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    if (s == 0) mma_small<EXACT, true>(a1[mt][j], fa[mt], fb[j]);
                    else mma_small<EXACT, false>(a1[mt][j], fa[mt], fb[j]);
                }
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) mma2_main(a1[mt][j], fa[mt], fb[j]);
        }
        float a2[2][4];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const FragB fb = EXACT ? frag_b_exact(f1b[s * 32 + lane]) : frag_b(f1b[s * 32 + lane]);
            FragA fa[2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                make_a(elu_fast(a1[mt][s][0]), elu_fast(a1[mt][s][2]), elu_fast(a1[mt][s][1]), elu_fast(a1[mt][s][3]), fa[mt]);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                if (s == 0) mma_small<EXACT, true>(a2[mt], fa[mt], fb);
                else mma_small<EXACT, false>(a2[mt], fa[mt], fb);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) mma2_main(a2[mt], fa[mt], fb);
        }
        // z of pixels 8q + g (q = 2 mt + r), the same in the four t lanes; lane (g, t) stores pixel 8t + g: one 128-byte line per warp
        float z[4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                float zp = fmaf(elu_fast(a2[mt][2 * r + 1]), w2b, elu_fast(a2[mt][2 * r]) * w2a);
                zp += __shfl_xor_sync(0xffffffffu, zp, 1);
                zp += __shfl_xor_sync(0xffffffffu, zp, 2);
                z[2 * mt + r] = zp;
            }
        const float zm = t == 0 ? z[0] : (t == 1 ? z[1] : (t == 2 ? z[2] : z[3]));
        const unsigned gp = (unsigned)G * kGP + 8u * t + g;
        if (gp < npx) io_store(reinterpret_cast<TIO *>(a.out) + gp, sigmoidf(zm));
    }
    cp_async_wait<0>();
}

template <int WARPS, int NSTAGE, int MINB, bool BF>
int launch_cfg2(const MmaArgs &ma, int sms, cudaStream_t s)
{
    constexpr size_t smem_bytes = sizeof(float4) * (4 * 2 * 32 + 2 * 32 + 2 * 32 + 2 * 2 * 32 * 2) + sizeof(float) * kSumFloats +
                                  sizeof(float) * (size_t)WARPS * NSTAGE * kTileFloats;
    auto kern = head_bwd_mma_kernel<WARPS, NSTAGE, MINB, BF>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long ngroups = (ma.a.npx + kGP - 1) / kGP;
    long long grid = (ngroups + WARPS - 1) / WARPS;
    if (grid > (long long)sms * MINB) grid = (long long)sms * MINB;
    kern<<<(unsigned)grid, WARPS * 32, smem_bytes, s>>>(ma);
    return check_launch("head_bwd_mma");
}
template <int WARPS, int NSTAGE, int MINB>
int launch_cfg(const MmaArgs &ma, int sms, cudaStream_t s)
{
    return ma.a.bf16_io ? launch_cfg2<WARPS, NSTAGE, MINB, true>(ma, sms, s) : launch_cfg2<WARPS, NSTAGE, MINB, false>(ma, sms, s);
}

template <int WARPS, int NSTAGE, int MINB, bool EXACT, bool BF>
int launch_fwd_cfg2(const MmaArgs &ma, int sms, cudaStream_t s)
{
    constexpr size_t smem_bytes = sizeof(float4) * (4 * 2 * 32 + 2 * 32) + sizeof(float) * (size_t)WARPS * NSTAGE * kTileFloats;
    auto kern = head_fwd_mma_kernel<WARPS, NSTAGE, MINB, EXACT, BF>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long ngroups = (ma.a.npx + kGP - 1) / kGP;
    long long grid = (ngroups + WARPS - 1) / WARPS;
    if (grid > (long long)sms * MINB) grid = (long long)sms * MINB;
    kern<<<(unsigned)grid, WARPS * 32, smem_bytes, s>>>(ma);
    return check_launch("head_fwd_mma");
}
template <int WARPS, int NSTAGE, int MINB, bool EXACT = false>
int launch_fwd_cfg(const MmaArgs &ma, int sms, cudaStream_t s)
{
    return ma.a.bf16_io ? launch_fwd_cfg2<WARPS, NSTAGE, MINB, EXACT, true>(ma, sms, s) : launch_fwd_cfg2<WARPS, NSTAGE, MINB, EXACT, false>(ma, sms, s);
}

}  // namespace

// The head's forward on warp-level MMAs: fp32 tensors, 128-bit-copyable features, no fused depth / stash.
// BTS_HEAD_MMA=0 keeps the tcgen05 kernel (chain_tc_fwd.cu).
bool head_fwd_mma_wanted(const ChainParams &cp, const TcArgs &a)
{
    static const int env = getenv("BTS_HEAD_MMA") ? atoi(getenv("BTS_HEAD_MMA")) : 1;
    if (env == 0 || !a.x_vec16 || a.depth != nullptr || a.stash_act != nullptr || a.out == nullptr) return false;
    return cfg_matches<Cfg32_16>(cp) && cp.is_final && cp.dims[3] == 1;
}

int launch_head_fwd_mma(const ChainParams &cp, const TcArgs &a, int sms, cudaStream_t s)
{
    MmaArgs ma;
    ma.a = a;
    tc_set_weights(ma.a, cp);
    ma.atomic_w = 0;
    static const int variant = getenv("BTS_HEAD_MMA") ? atoi(getenv("BTS_HEAD_MMA")) : -1;   // unset: the default case
    static const bool exact = getenv("BTS_HEAD_EXACT") != nullptr && atoi(getenv("BTS_HEAD_EXACT")) != 0;
    if (exact) return launch_fwd_cfg<8, 2, 2, true>(ma, sms, s);
    switch (variant) {
    case 1: return launch_fwd_cfg<4, 3, 3>(ma, sms, s);
    case 3: return launch_fwd_cfg<4, 4, 3>(ma, sms, s);
    case 2: return launch_fwd_cfg<4, 2, 4>(ma, sms, s);
    case 5: return launch_fwd_cfg<4, 2, 5>(ma, sms, s);           // 20 warps per SM
    default: return launch_fwd_cfg<8, 2, 2>(ma, sms, s);          // 16 warps per SM in two CTAs (measured best)
    }
}

// The head's backward on warp-level MMAs: fp32 tensors, 128-bit-copyable features, atomic weight-gradient accumulation.
// BTS_HEAD_MMA=0 keeps the tcgen05 kernel (chain_bwd_fused.cu).
bool head_bwd_mma_wanted(const ChainParams &cp, const TcArgs &a, int mode)
{
    static const int env = getenv("BTS_HEAD_MMA") ? atoi(getenv("BTS_HEAD_MMA")) : 1;
    static const bool deterministic = getenv("BTS_DETERMINISTIC") != nullptr && atoi(getenv("BTS_DETERMINISTIC")) != 0;
    static const bool nwg_forced = getenv("BTS_HEAD_NWG") != nullptr;    // a tcgen05 variant was asked for by name
    if (env == 0 || deterministic || nwg_forced || mode != 0 || !a.x_vec16) return false;
    if (!cfg_matches<Cfg32_16>(cp) || !cp.is_final || cp.dims[3] != 1) return false;
    if (a.grad_x != nullptr && (reinterpret_cast<uintptr_t>(a.grad_x) & (a.bf16_io ? 3u : 7u)) != 0) return false;
    return true;
}

int launch_head_bwd_mma(const ChainParams &cp, const TcArgs &a, int sms, cudaStream_t s)
{
    MmaArgs ma;
    ma.a = a;
    tc_set_weights(ma.a, cp);
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) ma.a.gw[l] = l < cp.n_layers ? cp.gW[l] : nullptr;
    ma.atomic_w = cp.gW[0] != nullptr ? 1 : 0;
    static const int variant = getenv("BTS_HEAD_MMA") ? atoi(getenv("BTS_HEAD_MMA")) : -1;   // unset: the default case
    switch (variant) {                                            // (tuning variants; the default is the measured best)
    case 1: return launch_cfg<4, 3, 3>(ma, sms, s);               // 12 warps per SM, 168 registers, two tiles ahead
    case 3: return launch_cfg<4, 3, 2>(ma, sms, s);               // 8 warps per SM, 255 registers
    case 4: return launch_cfg<8, 2, 2>(ma, sms, s);               // 16 warps per SM in two CTAs
    case 5: return launch_cfg<16, 2, 1>(ma, sms, s);              // 16 warps per SM in one CTA
    default: return launch_cfg<4, 2, 4>(ma, sms, s);              // 16 warps per SM, 128 registers, one tile ahead
    }
}

}  // namespace bts

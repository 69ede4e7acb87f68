// upconv.cu -- the decoder's `upconv` block (pytorch/bts.py:83-94), forward:
//     y = ELU(conv3x3(interpolate(x, scale_factor=2, mode='nearest')), bias=False, padding=1)
// with the up-sampled tensor never materialised and 2.25x fewer multiply-adds than the reference's sequence.
//
// Folding. Output pixel (2i+a, 2j+b) of phase (a, b) reads rows {2i+a-1, 2i+a, 2i+a+1} of the up-sampled image =
// source rows {i-1, i, i} (a = 0) or {i, i, i+1} (a = 1), likewise for columns. So every phase is a 2x2 convolution of the
// LOW-RES input with weights that are sums of the 3x3 taps:
//     y[o][2i+a][2j+b] = sum_c sum_{dy,dx in {0,1}} Wf[a][b][dy][dx][o][c] * x[c][i-1+a+dy][j-1+b+dx]      (zero outside)
//     Wf[a][.][dy] sums the tap rows  a=0: {0} | {1,2},   a=1: {0,1} | {2}        (same for b / dx and the tap columns)
// 16 Cin multiply-adds per low-res pixel and output channel for the four phases, against 36 Cin.
//
// Kernel: implicit GEMM on warp-level tensor-core MMAs with fp32-accurate two-instruction products (chain_head_mma.cu:
// hi*hi as tf32 m16n8k8, lo*hi + hi*lo as one bf16 m16n8k16). Out channels on M (folded weights from a shared-memory
// chunk), low-res pixels on N (eight consecutive pixels of a row: a shifted window of the halo tile is just another
// shared-memory offset for register fragments -- no descriptor alignment to respect), input channels x taps on K. A CTA
// computes 32 out channels x (8 x 16 low-res pixels) x both column phases of row phase a = blockIdx.z; a thread ends up
// with output columns 4k..4k+3 of a row: 128-bit stores. cp.async double buffering, 16 input channels per stage.
#include <cstdlib>

#include "chain_tc.cuh"

namespace bts {
namespace {

constexpr int kUcKC = 16;                            // input channels per stage
// Low-res tile of a CTA: 128 pixels as 8 rows x 16 columns (WIDE) or 16 x 8 (TALL: small images whose width wastes less that
// way, e.g. the 15 x 20 input of upconv5). A warp owns four runs of eight pixels along a row: rows 2w, 2w+1 x both column
// halves (WIDE) or rows 4w .. 4w+3 (TALL).
template <bool TALL>
struct UcTile {
    static constexpr int TW = TALL ? 8 : 16, TH = TALL ? 16 : 8;
    static constexpr int XR = TH + 1, XC = TW + 2;   // halo tile: 9 x 18 | 17 x 10
    static constexpr int XS = TALL ? 200 : 168;      // floats per channel, = 8 mod 32: conflict-free B fragments (k = channel)
    static constexpr int WXS = TALL ? 196 : 164;     // same for the weight gradient, = 4 mod 32 (k = pixel)
    __device__ static constexpr int row(int warp, int n) { return TALL ? 4 * warp + n : 2 * warp + (n >> 1); }
    __device__ static constexpr int colb(int n) { return TALL ? 0 : 8 * (n & 1); }
};
constexpr int kUcMO = 32;                            // out channels per CTA
constexpr int kUcFrag = 256;                         // words of one A fragment in lane order: 32 lanes x {4 hi words, 4 bf16 pairs}

__device__ __forceinline__ void hmma_t(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void hmma_b(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t bf2(uint32_t lo, uint32_t hi)
{
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(__uint_as_float(hi)), "f"(__uint_as_float(lo)));
    return r;
}
__device__ __forceinline__ void split(float v, uint32_t &h, uint32_t &l)
{
    h = __float_as_uint(v) & 0xFFFFE000u;
    l = __float_as_uint(v - __uint_as_float(h));
}
// fragments of an fp32-accurate product (see chain_head_mma.cu): tf32 hi words + bf16 pairs {lo, lo} / {hi, hi}
struct UFragA { uint32_t h[4], s[4], l[4]; };
struct UFragB { uint32_t h0, h1, s0, s1, l0, l1; };
#ifndef BTS_UC_EXACT
#define BTS_UC_EXACT 0       // 1: plain 3xTF32 (three tf32 instructions per product)
#endif
// The tensor core's fp32 accumulator truncates what it adds (tests/cuda/probe_hmma.cu), a bias toward zero that grows with
// the K steps an accumulator sees: measured against the fp64 oracle 1.1e-5 of peak at cin = 512 and 4.6e-5 at cin = 2208
// (cuDNN's fp32 kernels show 3.8e-5 there). So the accumulators leave the tensor core every BTS_UC_DRAIN stages (32 steps)
// and continue as fp32 master sums with round-to-nearest adds: 1.4e-6 at every cin, for 64 more registers (two CTAs per SM
// instead of three: 5.3 instead of 4.95 ms for the five decoder upconvs at nyu_b8). 0 = no drains.
#ifndef BTS_UC_DRAIN
#define BTS_UC_DRAIN 2
#endif
__device__ __forceinline__ void make_a(float v0, float v1, float v2, float v3, UFragA &f)
{
    split(v0, f.h[0], f.l[0]);
    split(v1, f.h[1], f.l[1]);
    split(v2, f.h[2], f.l[2]);
    split(v3, f.h[3], f.l[3]);
    f.s[0] = bf2(f.l[0], f.l[2]);
    f.s[1] = bf2(f.l[1], f.l[3]);
    f.s[2] = bf2(f.h[0], f.h[2]);
    f.s[3] = bf2(f.h[1], f.h[3]);
}
__device__ __forceinline__ void make_b(float v0, float v1, UFragB &f)
{
    split(v0, f.h0, f.l0);
    split(v1, f.h1, f.l1);
    f.s0 = bf2(f.h0, f.h1);
    f.s1 = bf2(f.l0, f.l1);
}

struct UpconvArgs {
    const float *x;          // [B][Cin][h][w]
    const uint4 *wfrag;      // folded weights as A fragments (upconv_fold_frag_kernel)
    const uint4 *xsplit;     // x as B fragments (upconv_split_kernel)
    float *y;                // [B][Cout][2h][2w]
    int B, Cin, Cout, h, w, tiles_x, tiles_y;
};

// The forward's A operand, ready to use: the folded weights split (hi words, bf16 pairs of the remainders and of the hi
// parts) and laid out as the MMA fragments the lanes load, [a][out-channel block of 32][k-step of 8 c][product (b,dy,dx)]
// [m-tile][lane][8 words] -- a stage's chunk is one contiguous 128-bit-copyable run, a fragment two 128-bit loads, and no
// warp splits a weight again (the four warps of a CTA used to split the same fragments four times).
__global__ void upconv_fold_frag_kernel(const float *__restrict__ w, uint4 *__restrict__ wfrag, int cout, int cin)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)2 * (cout / 32) * (cin / 8) * 8 * 2 * 32;
    if (idx >= total) return;
    const int lane = (int)(idx & 31), mt = (int)((idx >> 5) & 1), p = (int)((idx >> 6) & 7);
    long long r = idx >> 9;
    const int ks = (int)(r % (cin / 8));
    r /= (cin / 8);
    const int ob = (int)(r % (cout / 32)), a = (int)(r / (cout / 32));
    const int b = p >> 2, dy = (p >> 1) & 1, dx = p & 1, g = lane >> 2, t = lane & 3;
    const int y0 = a == 0 ? (dy == 0 ? 0 : 1) : (dy == 0 ? 0 : 2), y1 = a == 0 ? (dy == 0 ? 0 : 2) : (dy == 0 ? 1 : 2);
    const int x0 = b == 0 ? (dx == 0 ? 0 : 1) : (dx == 0 ? 0 : 2), x1 = b == 0 ? (dx == 0 ? 0 : 2) : (dx == 0 ? 1 : 2);
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {                                    // (row g, k t), (row g+8, k t), (row g, k t+4), (row g+8, k t+4)
        const int o = 32 * ob + 16 * mt + g + 8 * (q & 1), c = 8 * ks + t + 4 * (q >> 1);
        const float *src = w + ((size_t)o * cin + c) * 9;
        float sum = 0.f;
        for (int ky = y0; ky <= y1; ++ky)
            for (int kx = x0; kx <= x1; ++kx) sum += __ldg(src + ky * 3 + kx);
        v[q] = sum;
    }
    UFragA f;
    make_a(v[0], v[1], v[2], v[3], f);
    wfrag[2 * idx] = make_uint4(f.h[0], f.h[1], f.h[2], f.h[3]);
    wfrag[2 * idx + 1] = make_uint4(f.s[0], f.s[1], f.s[2], f.s[3]);
}

// The forward's B operand, ready to use. A B fragment pairs channels (c, c+4) of ONE pixel, so the split does not depend on
// the tap window: xs[img][c / 8][t = c % 4][pixel] = {hi(c), hi(c+4), bf16 pair of the hi parts, bf16 pair of the remainders}
// (16 bytes = one fragment of lane (g, t) once the pixel is chosen). One elementwise pass (x read once, 2x its bytes
// written) replaces two loads and six ALU instructions per fragment in every warp, window and out-channel block that uses it
// -- the kernel was issue-bound on exactly that.
__global__ void upconv_split_kernel(const float *__restrict__ x, uint4 *__restrict__ xs, long long n, int hw, int cin)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // ((img * cin/8 + oct) * 4 + t) * hw + px
    if (idx >= n) return;
    const int px = (int)(idx % hw);
    const long long r = idx / hw;
    const int t = (int)(r & 3);
    const long long io = r >> 2;                                     // img * (cin / 8) + oct
    const long long img = io / (cin / 8);
    const int oct = (int)(io - img * (cin / 8));
    const float *src = x + ((size_t)img * cin + 8 * oct + t) * hw + px;
    UFragB f;
    make_b(__ldg(src), __ldg(src + (size_t)4 * hw), f);
    xs[idx] = make_uint4(f.h0, f.h1, f.s0, f.s1);
}

// MT: m-tiles (16 out channels each) per CTA; 2 by default (see launch_upconv_fwd for the measured alternative).
template <bool TALL, int MT>
__global__ void __launch_bounds__(128, MT == 1 ? 4 : 2) upconv_fwd_kernel(const __grid_constant__ UpconvArgs ua)
{
    using T = UcTile<TALL>;
    constexpr int NPX = T::XR * T::XC;                               // halo pixels; a plane (k-step, t) is NPX 16-byte fragments
    constexpr int XFloats = (kUcKC / 8) * 4 * NPX * 4;               // (NPX * 4 words = 8 mod 32: the eight lanes of a 128-bit load phase hit eight bank groups)
    constexpr int WFloats = (kUcKC / 8) * 8 * MT * kUcFrag, StageFloats = XFloats + WFloats;
    static_assert((NPX * 4) % 32 == 8, "conflict-free fragment loads");
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int a = blockIdx.z, oblk = blockIdx.y / (2 / MT), msel = blockIdx.y % (2 / MT), o0 = oblk * kUcMO + 16 * msel;
    int tile = blockIdx.x;
    const int tx = tile % ua.tiles_x;
    tile /= ua.tiles_x;
    const int ty = tile % ua.tiles_y, b_img = tile / ua.tiles_y;
    const int i0 = ty * T::TH, j0 = tx * T::TW;
    const int h = ua.h, w = ua.w, Cin = ua.Cin, Cout = ua.Cout;
    const uint4 *xb = ua.xsplit + (size_t)b_img * (Cin / 8) * 4 * h * w;
    // fragments of row phase a, out-channel block blockIdx.y: [k-step][8 products][2 m-tiles][32 lanes][8 words]
    const uint4 *wchunk = ua.wfrag + ((size_t)a * (Cout / kUcMO) + oblk) * (size_t)(Cin / 8) * (8 * 2 * kUcFrag / 4);

    auto load_stage = [&](int kc, int buf) {
        float *xs = smem + buf * StageFloats, *ws = xs + XFloats;
        // halo tile of ready B fragments, planes [k-step][t]: smem row r <-> source row i0 - 1 + a + r, column c <-> j0 - 1 + c
        for (int idx = tid; idx < (kUcKC / 8) * 4 * NPX; idx += 128) {
            const int pl = idx / NPX, rem = idx - pl * NPX;
            const int r = rem / T::XC, col = rem - r * T::XC;
            const int sr = i0 - 1 + a + r, sc = j0 - 1 + col;
            const bool ok = sr >= 0 && sr < h && sc >= 0 && sc < w;
            const uint4 *src = xb + ((size_t)(kc * (kUcKC / 8) * 4 + pl) * h + (ok ? sr : 0)) * w + (ok ? sc : 0);
            cp_async16(smem_u32(xs + 4 * idx), src, ok);
        }
        // weight chunk: the stage's 2 k-steps x 8 products x MT m-tiles of ready fragments (1 KB each; global order
        // [k-step][product][2 m-tiles])
        {
            const uint4 *src = wchunk + (size_t)kc * ((kUcKC / 8) * 8 * 2 * kUcFrag / 4);
            for (int idx = tid; idx < WFloats / 4; idx += 128) {
                const int fr = idx / (MT * 64), j = idx - fr * (MT * 64);             // fr = k-step * 8 + product
                cp_async16(smem_u32(ws + 4 * idx), src + (size_t)(fr * 2 + (MT == 1 ? msel : 0)) * 64 + j, true);
            }
        }
        cp_async_commit();
    };

    float acc[MT][4][2][4];                                           // [m-tile][n-tile][column phase b][fragment]
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[mt][n][b][q] = 0.f;
#if BTS_UC_DRAIN
    float master[MT][4][2][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int q = 0; q < 4; ++q) master[mt][n][b][q] = 0.f;
#endif

    const int nk = Cin / kUcKC;
    load_stage(0, 0);
    for (int kc = 0; kc < nk; ++kc) {
        const int buf = kc & 1;
        if (kc + 1 < nk) {
            load_stage(kc + 1, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const float *xs = smem + buf * StageFloats, *ws = xs + XFloats;
#pragma unroll
        for (int ks = 0; ks < kUcKC / 8; ++ks) {
            // six distinct windows of the halo tile: tap row dy x column shift sh = b + dx; each feeds the products (b, dx)
            // with b + dx = sh (one, two, one of them)
#pragma unroll
            for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                for (int sh = 0; sh < 3; ++sh) {
                    UFragB fb[4];
#pragma unroll
                    for (int n = 0; n < 4; ++n) {                    // n-tile: eight pixels of tile row T::row(warp, n); lane (g, t) takes pixel g's fragment t
                        const uint4 v = reinterpret_cast<const uint4 *>(xs)[(ks * 4 + t) * NPX + (T::row(warp, n) + dy) * T::XC + T::colb(n) + sh + g];
                        fb[n].h0 = v.x; fb[n].h1 = v.y; fb[n].s0 = v.z; fb[n].s1 = v.w;
                    }
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int dx = sh - b;
                        if (dx < 0 || dx > 1) continue;
                        const int p = b * 4 + dy * 2 + dx;
                        UFragA fa[MT];
#pragma unroll
                        for (int mt = 0; mt < MT; ++mt) {
                            const uint4 *q = reinterpret_cast<const uint4 *>(ws + ((ks * 8 + p) * MT + mt) * kUcFrag) + 2 * lane;
                            const uint4 vh = q[0], vs = q[1];
                            fa[mt].h[0] = vh.x; fa[mt].h[1] = vh.y; fa[mt].h[2] = vh.z; fa[mt].h[3] = vh.w;
                            fa[mt].s[0] = vs.x; fa[mt].s[1] = vs.y; fa[mt].s[2] = vs.z; fa[mt].s[3] = vs.w;
                        }
#pragma unroll
                        for (int n = 0; n < 4; ++n)
#pragma unroll
                            for (int mt = 0; mt < MT; ++mt) hmma_b(acc[mt][n][b], fa[mt].s, fb[n].s0, fb[n].s1);
#pragma unroll
                        for (int n = 0; n < 4; ++n)
#pragma unroll
                            for (int mt = 0; mt < MT; ++mt) hmma_t(acc[mt][n][b], fa[mt].h, fb[n].h0, fb[n].h1);
                    }
                }
        }
        __syncthreads();
#if BTS_UC_DRAIN
        if ((kc + 1) % (BTS_UC_DRAIN * 16 / kUcKC) == 0 || kc + 1 == nk) {   // every BTS_UC_DRAIN x 16 input channels
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int n = 0; n < 4; ++n)
#pragma unroll
                    for (int b = 0; b < 2; ++b)
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            master[mt][n][b][q] += acc[mt][n][b][q];
                            acc[mt][n][b][q] = 0.f;
                        }
        }
#endif
    }
#if BTS_UC_DRAIN
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[mt][n][b][q] = master[mt][n][b][q];
#endif

    // ---- ELU + stores: thread holds (o = 16 mt + g | + 8) x (low-res columns 2t, 2t+1 of the n-tile) x (b = 0, 1)
    //      = four consecutive output columns of output row 2 i + a
    const int H2 = 2 * h, W2 = 2 * w;
    const bool vec_ok = (W2 % 4) == 0 && (reinterpret_cast<uintptr_t>(ua.y) & 15u) == 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        const int i = i0 + T::row(warp, n), j = j0 + T::colb(n) + 2 * t;
        if (i >= h || j >= w) continue;
        const bool two = j + 1 < w;
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                const int o = o0 + 16 * mt + g + 8 * hf;
                float *dst = ua.y + (((size_t)b_img * Cout + o) * H2 + 2 * i + a) * W2 + 2 * j;
                const float v0 = elu(acc[mt][n][0][2 * hf]), v1 = elu(acc[mt][n][1][2 * hf]);
                const float v2 = elu(acc[mt][n][0][2 * hf + 1]), v3 = elu(acc[mt][n][1][2 * hf + 1]);
                if (two && vec_ok) {
                    *reinterpret_cast<float4 *>(dst) = make_float4(v0, v1, v2, v3);
                } else {
                    dst[0] = v0;
                    dst[1] = v1;
                    if (two) {
                        dst[2] = v2;
                        dst[3] = v3;
                    }
                }
            }
    }
}


// =====================================================================================================================
// backward.  e = grad_y . ELU'(y)  (ELU'(t) through y = ELU(t): y > 0 ? 1 : y + 1), split into its four phase planes
//     E[a][b][img][o][i][j] = e[img][o][2i+a][2j+b]          (low-resolution sized, so every plane is a dense image)
//   grad_x[c][i][j]          = sum_{a,b,dy,dx} sum_o Wf[a][b][dy][dx][o][c] * E[a][b][o][i+1-a-dy][j+1-b-dx]     (zero outside)
//   grad_Wf[a][b][dy][dx][o][c] = sum_{img,i,j} E[a][b][o][i][j] * x[c][i-1+a+dy][j-1+b+dx]
//   grad_W[o][c][ky][kx]     = sum of the grad_Wf entries whose tap sets contain (ky, kx)
// Both products are implicit GEMMs of the forward's kind: the input gradient has the phase planes as its four sources
// (K = out channels x 16 products), the weight gradient has the pixel axis as K.
// =====================================================================================================================
__global__ void upconv_eplanes_kernel(const float *__restrict__ gy, const float *__restrict__ y, float *__restrict__ e,
                                      long long nlow, int h, int w)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (img * Cout + o, i, j)
    if (idx >= nlow) return;
    const int j = (int)(idx % w);
    const long long r = idx / w;
    const int i = (int)(r % h);
    const long long plane = r / h;
    const size_t src = ((size_t)plane * 2 * h + 2 * i) * (2 * (size_t)w) + 2 * j;
    float g4[4], y4[4];
    g4[0] = __ldg(gy + src); g4[1] = __ldg(gy + src + 1); g4[2] = __ldg(gy + src + 2 * w); g4[3] = __ldg(gy + src + 2 * w + 1);
    y4[0] = __ldg(y + src); y4[1] = __ldg(y + src + 1); y4[2] = __ldg(y + src + 2 * w); y4[3] = __ldg(y + src + 2 * w + 1);
#pragma unroll
    for (int ab = 0; ab < 4; ++ab) e[(size_t)ab * nlow + idx] = g4[ab] * elu_grad_from_out(y4[ab]);
}

// The input gradient's A operand (A[m = c][k = o] = Wf[p][o][c]) as ready fragments, like the forward's:
// [input-channel block of 32][k-step of 8 out channels][16 products][2 m-tiles][lane][8 words].
__global__ void upconv_fold_frag_t_kernel(const float *__restrict__ w, uint4 *__restrict__ wfrag, int cout, int cin)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)(cin / 32) * (cout / 8) * 16 * 2 * 32;
    if (idx >= total) return;
    const int lane = (int)(idx & 31), mt = (int)((idx >> 5) & 1), p = (int)((idx >> 6) & 15);
    const long long r = idx >> 10;
    const int ks = (int)(r % (cout / 8)), cb = (int)(r / (cout / 8));
    const int a = p >> 3, b = (p >> 2) & 1, dy = (p >> 1) & 1, dx = p & 1, g = lane >> 2, t = lane & 3;
    const int y0 = a == 0 ? (dy == 0 ? 0 : 1) : (dy == 0 ? 0 : 2), y1 = a == 0 ? (dy == 0 ? 0 : 2) : (dy == 0 ? 1 : 2);
    const int x0 = b == 0 ? (dx == 0 ? 0 : 1) : (dx == 0 ? 0 : 2), x1 = b == 0 ? (dx == 0 ? 0 : 2) : (dx == 0 ? 1 : 2);
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {                                    // (row g, k t), (row g+8, k t), (row g, k t+4), (row g+8, k t+4)
        const int c = 32 * cb + 16 * mt + g + 8 * (q & 1), o = 8 * ks + t + 4 * (q >> 1);
        const float *src = w + ((size_t)o * cin + c) * 9;
        float sum = 0.f;
        for (int ky = y0; ky <= y1; ++ky)
            for (int kx = x0; kx <= x1; ++kx) sum += __ldg(src + ky * 3 + kx);
        v[q] = sum;
    }
    UFragA f;
    make_a(v[0], v[1], v[2], v[3], f);
    wfrag[2 * idx] = make_uint4(f.h[0], f.h[1], f.h[2], f.h[3]);
    wfrag[2 * idx + 1] = make_uint4(f.s[0], f.s[1], f.s[2], f.s[3]);
}

constexpr int kDgKC = 8;                              // out channels (K) per stage: one k-step
constexpr int kDgWFloats = 16 * 2 * kUcFrag;             // the stage's 16 products x 2 m-tiles of ready A fragments
constexpr int kDgEFloats = 4 * kDgKC * 168;
constexpr int kDgStageFloats = kDgEFloats + kDgWFloats;

struct UpconvBwdArgs {
    const float *x;          // [B][Cin][h][w]
    const uint4 *wfrag_t;    // folded weights as the input gradient's A fragments (upconv_fold_frag_t_kernel)
    const float *e;          // phase planes [4][B][Cout][h][w]
    float *gx;               // [B][Cin][h][w]
    float *gwf;              // [16 p][Cout][Cin], zero-filled before the weight-gradient kernel
    int B, Cin, Cout, h, w, tiles_x, tiles_y;
    int tiles_per_cta;       // weight gradient: pixel tiles per CTA (split K)
};

// grad_x: CTA = 32 input channels (M) x 8x16 pixels (N); K = out channels, 8 per stage, x 16 (plane, tap) products
template <bool TALL>
__global__ void __launch_bounds__(128, 2) upconv_dgrad_kernel(const __grid_constant__ UpconvBwdArgs ua)
{
    using T = UcTile<TALL>;
    // halo of a plane: TH + 1 rows x TW + 1 columns (153 pixels either way), channel stride 168 = 8 mod 32
    constexpr int DXR = T::TH + 1, DXC = T::TW + 1, DXS = 168;
    static_assert(DXR * DXC <= DXS, "halo fits its stride");
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int c0 = blockIdx.y * 32;
    int tile = blockIdx.x;
    const int tx = tile % ua.tiles_x;
    tile /= ua.tiles_x;
    const int ty = tile % ua.tiles_y, b_img = tile / ua.tiles_y;
    const int i0 = ty * T::TH, j0 = tx * T::TW;
    const int h = ua.h, w = ua.w, Cin = ua.Cin, Cout = ua.Cout;
    const size_t plane_sz = (size_t)ua.B * Cout * h * w;

    auto load_stage = [&](int kc, int buf) {
        float *es = smem + buf * kDgStageFloats, *ws = es + kDgEFloats;
        // plane (a, b): smem row r <-> source row i0 - a + r, smem column c <-> source column j0 - b + c
        for (int idx = tid; idx < 4 * kDgKC * DXR * DXC; idx += 128) {
            const int po = idx / (DXR * DXC), rem = idx - po * (DXR * DXC);   // po = plane * 8 + o
            const int r = rem / DXC, col = rem - r * DXC;
            const int ab = po >> 3, o = po & 7;
            const int sr = i0 - (ab >> 1) + r, sc = j0 - (ab & 1) + col;
            const bool ok = sr >= 0 && sr < h && sc >= 0 && sc < w;
            const float *src = ua.e + (size_t)ab * plane_sz + (((size_t)b_img * Cout + kc * kDgKC + o) * h + (ok ? sr : 0)) * w + (ok ? sc : 0);
            cp_async4(smem_u32(es + po * DXS + r * DXC + col), src, ok);
        }
        // weights: the stage's 16 products x 2 m-tiles of ready fragments, one contiguous run
        {
            const uint4 *src = ua.wfrag_t + ((size_t)blockIdx.y * (Cout / kDgKC) + kc) * (kDgWFloats / 4);
            for (int idx = tid; idx < kDgWFloats / 4; idx += 128) cp_async16(smem_u32(ws + 4 * idx), src + idx, true);
        }
        cp_async_commit();
    };

    float acc[2][4][4], master[2][4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[mt][n][q] = master[mt][n][q] = 0.f;

    const int nk = Cout / kDgKC;
    load_stage(0, 0);
    for (int kc = 0; kc < nk; ++kc) {
        const int buf = kc & 1;
        if (kc + 1 < nk) {
            load_stage(kc + 1, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const float *es = smem + buf * kDgStageFloats, *ws = es + kDgEFloats;
#pragma unroll
        for (int p = 0; p < 16; ++p) {                               // p = ((a*2 + b)*2 + dy)*2 + dx
            const int ab = p >> 2, dy = (p >> 1) & 1, dx = p & 1;
            UFragA fa[2];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {                         // A[m = c][k = o] = Wf[p][o][c], pre-split
                const uint4 *q = reinterpret_cast<const uint4 *>(ws + (p * 2 + mt) * kUcFrag) + 2 * lane;
                const uint4 vh = q[0], vs = q[1];
                fa[mt].h[0] = vh.x; fa[mt].h[1] = vh.y; fa[mt].h[2] = vh.z; fa[mt].h[3] = vh.w;
                fa[mt].s[0] = vs.x; fa[mt].s[1] = vs.y; fa[mt].s[2] = vs.z; fa[mt].s[3] = vs.w;
            }
#pragma unroll
            for (int n = 0; n < 4; ++n) {
                const float *q = es + (ab * kDgKC + t) * DXS + (T::row(warp, n) + 1 - dy) * DXC + T::colb(n) + 1 - dx + g;
                UFragB fb;
                make_b(q[0], q[4 * DXS], fb);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) hmma_b(acc[mt][n], fa[mt].s, fb.s0, fb.s1);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) hmma_t(acc[mt][n], fa[mt].h, fb.h0, fb.h1);
            }
        }
        __syncthreads();
        // 32 accumulate steps per stage: the sums leave the tensor core (truncating accumulator, see BTS_UC_DRAIN)
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int n = 0; n < 4; ++n)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    master[mt][n][q] += acc[mt][n][q];
                    acc[mt][n][q] = 0.f;
                }
    }

    const bool vec_ok = (w % 2) == 0 && (reinterpret_cast<uintptr_t>(ua.gx) & 7u) == 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        const int i = i0 + T::row(warp, n), j = j0 + T::colb(n) + 2 * t;
        if (i >= h || j >= w) continue;
        const bool two = j + 1 < w;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                float *dst = ua.gx + (((size_t)b_img * Cin + c0 + 16 * mt + g + 8 * hf) * h + i) * w + j;
                if (two && vec_ok) {
                    *reinterpret_cast<float2 *>(dst) = make_float2(master[mt][n][2 * hf], master[mt][n][2 * hf + 1]);
                } else {
                    dst[0] = master[mt][n][2 * hf];
                    if (two) dst[1] = master[mt][n][2 * hf + 1];
                }
            }
    }
}


// grad_Wf: the pixel axis is K. CTA = one phase plane (a, b) [its four (dy, dx) products share the plane's tile and one halo
// tile of x] x 16 out channels (M) x 32 input channels (N) x a run of 8x16 pixel tiles (split K; fp32 atomics at the end).
// A warp takes two rows of every tile = four k-steps of eight pixels along a row.
constexpr int kWgES = 132;                            // floats per out-channel row of the plane tile (128 px + 4: conflict-free A)
constexpr int kWgEFloats = 16 * kWgES;
template <bool TALL> constexpr int kWgStageFloats = kWgEFloats + 32 * UcTile<TALL>::WXS;   // + halo tile of x, 32 input channels

template <bool TALL>
__global__ void __launch_bounds__(128, 2) upconv_wgrad_kernel(const __grid_constant__ UpconvBwdArgs ua)
{
    using T = UcTile<TALL>;
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int h = ua.h, w = ua.w, Cin = ua.Cin, Cout = ua.Cout;
    const int nob = Cout / 16;
    const int ab = blockIdx.y / nob, o0 = (blockIdx.y % nob) * 16, c0 = blockIdx.z * 32;
    const int a = ab >> 1, b = ab & 1;
    const int ntiles = ua.tiles_x * ua.tiles_y * ua.B;
    const int tile_begin = blockIdx.x * ua.tiles_per_cta;
    const int tile_end = min(ntiles, tile_begin + ua.tiles_per_cta);
    const float *eplane = ua.e + (size_t)ab * ua.B * Cout * h * w;

    auto load_stage = [&](int tile, int buf) {
        float *es = smem + buf * kWgStageFloats<TALL>, *xs = es + kWgEFloats;
        const int tx = tile % ua.tiles_x;
        const int rest = tile / ua.tiles_x;
        const int ty = rest % ua.tiles_y, img = rest / ua.tiles_y;
        const int i0 = ty * T::TH, j0 = tx * T::TW;
        {   // plane tile [16 o][8 x 16 px]: thread = pixel
            const int r = tid / T::TW, col = tid % T::TW;
            const int sr = i0 + r, sc = j0 + col;
            const bool ok = sr < h && sc < w;
            const float *src = eplane + (((size_t)img * Cout + o0) * h + (ok ? sr : 0)) * w + (ok ? sc : 0);
            const uint32_t dst = smem_u32(es + tid);
#pragma unroll
            for (int o = 0; o < 16; ++o) cp_async4(dst + o * kWgES * 4, src + (size_t)o * h * w, ok);
        }
        // halo tile of x [32 c][9 x 18]: smem row r <-> source row i0 - 1 + a + r, column c <-> j0 - 1 + b + c
        for (int pos = tid; pos < T::XR * T::XC; pos += 128) {
            const int r = pos / T::XC, col = pos - r * T::XC;
            const int sr = i0 - 1 + a + r, sc = j0 - 1 + b + col;
            const bool ok = sr >= 0 && sr < h && sc >= 0 && sc < w;
            const float *src = ua.x + (((size_t)img * Cin + c0) * h + (ok ? sr : 0)) * w + (ok ? sc : 0);
            const uint32_t dst = smem_u32(xs + pos);
#pragma unroll 8
            for (int c = 0; c < 32; ++c) cp_async4(dst + c * T::WXS * 4, src + (size_t)c * h * w, ok);
        }
        cp_async_commit();
    };

    float acc[4][4][4], master[4][4][4];                             // [product (dy, dx)][n-tile (8 c)][fragment]
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[p][n][q] = master[p][n][q] = 0.f;

    if (tile_begin < tile_end) load_stage(tile_begin, 0);
    for (int tile = tile_begin; tile < tile_end; ++tile) {
        const int buf = (tile - tile_begin) & 1;
        if (tile + 1 < tile_end) {
            load_stage(tile + 1, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const float *es = smem + buf * kWgStageFloats<TALL>, *xs = es + kWgEFloats;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {                             // eight pixels: row 2 warp + ks/2, columns 8 (ks%2) ..+7
            const int rowl = T::row(warp, ks), colb = T::colb(ks);
            UFragA fa;                                               // A[m = o][k = px] = E[o][px]
            {
                const float *q = es + g * kWgES + rowl * T::TW + colb + t;
                make_a(q[0], q[8 * kWgES], q[4], q[8 * kWgES + 4], fa);
            }
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int dy = p >> 1, dx = p & 1;
#pragma unroll
                for (int n = 0; n < 4; ++n) {                        // B[k = px][n = c] = x[c][row + dy][col + dx]
                    const float *q = xs + (8 * n + g) * T::WXS + (rowl + dy) * T::XC + colb + dx + t;
                    UFragB fb;
                    make_b(q[0], q[4], fb);
                    hmma_b(acc[p][n], fa.s, fb.s0, fb.s1);
                    hmma_t(acc[p][n], fa.h, fb.h0, fb.h1);
                }
            }
        }
        __syncthreads();
        if (((tile - tile_begin) & 3) == 3 || tile + 1 == tile_end) {   // 32 accumulate steps: drain (see BTS_UC_DRAIN)
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int n = 0; n < 4; ++n)
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        master[p][n][q] += acc[p][n][q];
                        acc[p][n][q] = 0.f;
                    }
        }
    }
    // the four warps' sums -> shared memory -> one atomic per element and CTA
    float *red = smem;                                               // [4 p][16 o][32 c]
    for (int e = tid; e < 4 * 16 * 32; e += 128) red[e] = 0.f;
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int q = 0; q < 4; ++q) atomicAdd(red + (p * 16 + g + 8 * (q >> 1)) * 32 + 8 * n + 2 * t + (q & 1), master[p][n][q]);
    __syncthreads();
    for (int e = tid; e < 4 * 16 * 32; e += 128) {
        const int p = e >> 9, o = (e >> 5) & 15, c = e & 31;
        atomicAdd(ua.gwf + ((size_t)(ab * 4 + p) * Cout + o0 + o) * Cin + c0 + c, red[e]);
    }
}

// grad_W[o][c][ky][kx] = sum of the folded gradients whose tap sets contain (ky, kx)
__global__ void upconv_unfold_kernel(const float *__restrict__ gwf, float *__restrict__ gw, int cout, int cin)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= cout * cin) return;
    const size_t n = (size_t)cout * cin;
    float f[2][2][2][2];
#pragma unroll
    for (int p = 0; p < 16; ++p) f[p >> 3][(p >> 2) & 1][(p >> 1) & 1][p & 1] = __ldg(gwf + p * n + e);
    // tap k belongs to (phase, d): k=0: (0,0), (1,0);  k=1: (0,1), (1,0);  k=2: (0,1), (1,1)
    const int ph[3][2] = {{0, 1}, {0, 1}, {0, 1}}, dd[3][2] = {{0, 0}, {1, 0}, {1, 1}};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
            float s = 0.f;
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int v = 0; v < 2; ++v) s += f[ph[ky][u]][ph[kx][v]][dd[ky][u]][dd[kx][v]];
            gw[(size_t)e * 9 + ky * 3 + kx] = s;
        }
}

}  // namespace

// folded weights as A fragments (hi word + bf16 halves per weight) + x as B fragments (same, per element)
size_t upconv_workspace_bytes(int B, int cin, int cout, int h, int w)
{
    return ((size_t)16 * cin * cout + (size_t)B * cin * h * w) * 2 * sizeof(float);
}

bool upconv_supported(int cin, int cout) { return cin > 0 && cout > 0 && cin % kUcKC == 0 && cout % kUcMO == 0; }

// 16 x 8 tiles when they cover the image with less padding than 8 x 16 (BTS_UPCONV_TILE=wide|tall overrides)
static bool upconv_tall(int h, int w)
{
    static const char *env = getenv("BTS_UPCONV_TILE");
    if (env != nullptr && env[0] == 'w') return false;
    if (env != nullptr && env[0] == 't') return true;
    const long long wide = (long long)((h + 7) / 8) * ((w + 15) / 16), tall = (long long)((h + 15) / 16) * ((w + 7) / 8);
    return tall < wide;
}

template <bool TALL, int MT>
static int launch_upconv_fwd_t(UpconvArgs &ua, int B, cudaStream_t s)
{
    using T = UcTile<TALL>;
    ua.tiles_x = (ua.w + T::TW - 1) / T::TW;
    ua.tiles_y = (ua.h + T::TH - 1) / T::TH;
    const size_t smem_bytes = sizeof(float) * 2 * ((kUcKC / 8) * 4 * T::XR * T::XC * 4 + (kUcKC / 8) * 8 * MT * kUcFrag);
    cudaError_t e = cudaFuncSetAttribute(upconv_fwd_kernel<TALL, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long tiles = (long long)ua.tiles_x * ua.tiles_y * B;
    if (tiles > 0x7fffffffLL) return set_error(BTS_ERR_BAD_ARG, "too many tiles");
    dim3 grid((unsigned)tiles, (unsigned)(ua.Cout / kUcMO * (2 / MT)), 2);
    upconv_fwd_kernel<TALL, MT><<<grid, 128, smem_bytes, s>>>(ua);
    return check_launch("upconv_fwd");
}

int launch_upconv_fwd(const float *x, const float *w, float *y, int B, int cin, int cout, int h, int wd, void *ws, cudaStream_t s)
{
    uint4 *wfrag = static_cast<uint4 *>(ws);
    const long long nfrag = (long long)2 * (cout / 32) * (cin / 8) * 8 * 2 * 32;
    upconv_fold_frag_kernel<<<(unsigned)((nfrag + 255) / 256), 256, 0, s>>>(w, wfrag, cout, cin);
    if (int e = check_launch("upconv_fold_frag")) return e;
    uint4 *xsplit = wfrag + (size_t)16 * cin * cout / 2;              // (16 cin cout weights x 8 bytes)
    const long long nsplit = (long long)B * (cin / 2) * h * wd;
    upconv_split_kernel<<<(unsigned)((nsplit + 255) / 256), 256, 0, s>>>(x, xsplit, nsplit, h * wd, cin);
    if (int e = check_launch("upconv_split")) return e;
    UpconvArgs ua;
    ua.x = x; ua.wfrag = wfrag; ua.xsplit = xsplit; ua.y = y;
    ua.B = B; ua.Cin = cin; ua.Cout = cout; ua.h = h; ua.w = wd;
    // 32 out channels per CTA. BTS_UPCONV_MT=1 (tuning): 16 per CTA = half the accumulators, 16 instead of 8 warps per SM --
    // measured SLOWER (5.6 vs 4.5 ms for the five decoder upconvs at nyu_b8): every B fragment then serves half as many MMAs
    static const int mt_env = getenv("BTS_UPCONV_MT") ? atoi(getenv("BTS_UPCONV_MT")) : 2;
    if (mt_env == 1) return upconv_tall(h, wd) ? launch_upconv_fwd_t<true, 1>(ua, B, s) : launch_upconv_fwd_t<false, 1>(ua, B, s);
    return upconv_tall(h, wd) ? launch_upconv_fwd_t<true, 2>(ua, B, s) : launch_upconv_fwd_t<false, 2>(ua, B, s);
}

bool upconv_bwd_supported(int cin, int cout) { return cin > 0 && cout > 0 && cin % 32 == 0 && cout % 32 == 0; }

size_t upconv_bwd_workspace_bytes(int B, int cin, int cout, int h, int w)
{
    return sizeof(float) * ((size_t)48 * cin * cout + (size_t)4 * B * cout * h * w);
}

template <bool TALL>
static int launch_upconv_bwd_t(UpconvBwdArgs &ua, const float *w, uint4 *wfrag_t, float *gw, cudaStream_t s)
{
    using T = UcTile<TALL>;
    const int cin = ua.Cin, cout = ua.Cout;
    ua.tiles_x = (ua.w + T::TW - 1) / T::TW;
    ua.tiles_y = (ua.h + T::TH - 1) / T::TH;
    ua.tiles_per_cta = 1;
    const long long tiles = (long long)ua.tiles_x * ua.tiles_y * ua.B;
    if (tiles > 0x7fffffffLL) return set_error(BTS_ERR_BAD_ARG, "too many tiles");
    if (ua.gx != nullptr) {
        const long long nfrag = (long long)(cin / 32) * (cout / 8) * 16 * 2 * 32;
        upconv_fold_frag_t_kernel<<<(unsigned)((nfrag + 255) / 256), 256, 0, s>>>(w, wfrag_t, cout, cin);
        if (int er = check_launch("upconv_fold_frag_t")) return er;
        const size_t smem_bytes = sizeof(float) * 2 * kDgStageFloats;
        cudaError_t ce = cudaFuncSetAttribute(upconv_dgrad_kernel<TALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
        if (ce != cudaSuccess) return set_error((int)ce, "cudaFuncSetAttribute: %s", cudaGetErrorString(ce));
        upconv_dgrad_kernel<TALL><<<dim3((unsigned)tiles, (unsigned)(cin / 32), 1), 128, smem_bytes, s>>>(ua);
        if (int er = check_launch("upconv_dgrad")) return er;
    }
    if (gw != nullptr) {
        cudaError_t ce = cudaMemsetAsync(ua.gwf, 0, sizeof(float) * (size_t)16 * cin * cout, s);
        if (ce != cudaSuccess) return set_error((int)ce, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
        // split K so that the grid has a few CTAs per SM slot
        const long long blocks = (long long)4 * (cout / 16) * (cin / 32);
        long long ksplit = (8 * 148 * 2 + blocks - 1) / blocks;
        if (ksplit < 1) ksplit = 1;
        if (ksplit > tiles) ksplit = tiles;
        ua.tiles_per_cta = (int)((tiles + ksplit - 1) / ksplit);
        const unsigned gx_ = (unsigned)((tiles + ua.tiles_per_cta - 1) / ua.tiles_per_cta);
        if (4 * (cout / 16) > 65535 || cin / 32 > 65535) return set_error(BTS_ERR_UNSUPPORTED, "upconv: too many channels");
        const size_t smem_bytes = sizeof(float) * 2 * kWgStageFloats<TALL>;
        ce = cudaFuncSetAttribute(upconv_wgrad_kernel<TALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
        if (ce != cudaSuccess) return set_error((int)ce, "cudaFuncSetAttribute: %s", cudaGetErrorString(ce));
        upconv_wgrad_kernel<TALL><<<dim3(gx_, (unsigned)(4 * (cout / 16)), (unsigned)(cin / 32)), 128, smem_bytes, s>>>(ua);
        if (int er = check_launch("upconv_wgrad")) return er;
        upconv_unfold_kernel<<<(cin * cout + 255) / 256, 256, 0, s>>>(ua.gwf, gw, cout, cin);
        if (int er = check_launch("upconv_unfold")) return er;
    }
    return BTS_OK;
}

// ws = [folded weights as fragments, 16 Cout Cin x 8 bytes | folded weight gradients 16 Cout Cin | phase planes 4 B Cout h w]
int launch_upconv_bwd(const float *x, const float *w, const float *y, const float *gy, float *gx, float *gw, int B, int cin, int cout,
                      int h, int wd, void *ws, cudaStream_t s)
{
    uint4 *wfrag_t = static_cast<uint4 *>(ws);
    float *gwf = static_cast<float *>(ws) + (size_t)32 * cin * cout, *e = gwf + (size_t)16 * cin * cout;
    const long long nlow = (long long)B * cout * h * wd;
    upconv_eplanes_kernel<<<(unsigned)((nlow + 255) / 256), 256, 0, s>>>(gy, y, e, nlow, h, wd);
    if (int er = check_launch("upconv_eplanes")) return er;
    UpconvBwdArgs ua;
    ua.x = x; ua.wfrag_t = wfrag_t; ua.e = e; ua.gx = gx; ua.gwf = gwf;
    ua.B = B; ua.Cin = cin; ua.Cout = cout; ua.h = h; ua.w = wd;
    return upconv_tall(h, wd) ? launch_upconv_bwd_t<true>(ua, w, wfrag_t, gw, s) : launch_upconv_bwd_t<false>(ua, w, wfrag_t, gw, s);
}

}  // namespace bts

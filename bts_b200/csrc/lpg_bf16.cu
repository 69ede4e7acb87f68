// lpg_bf16.cu -- bf16 twin of the standalone local_planar_guidance kernels (pytorch/bts.py:149-173 and its autograd
// backward; north_star: "1e-2 for an optional bf16 path"). plane_eq, depth and their gradients are bf16 tensors; the
// arithmetic is the fp32 kernels' (lpg.cu), in the reference's operation order, rounded to bf16 once at the store.
// One thread per LOW-RES pixel: it owns an R x R block, R in {1, 2, 4, 8}, whose rows are R contiguous bf16 values
// (2 .. 16 bytes: one vector access per row; consecutive lanes own consecutive blocks, so a warp's row accesses are
// contiguous).  fwd bytes / full-res pixel: 8 / R^2 + 2, bwd: 2 + 16 / R^2.
#include "chain_tc.cuh"

namespace bts {

template <int R>
__device__ __forceinline__ void load_row_bf16(const unsigned short *p, float (&v)[R])
{
    if constexpr (R == 8) {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(p));
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = bf16_to_f32((unsigned short)(w[i] & 0xffffu));
            v[2 * i + 1] = bf16_to_f32((unsigned short)(w[i] >> 16));
        }
    } else if constexpr (R == 4) {
        const uint2 q = __ldg(reinterpret_cast<const uint2 *>(p));
        v[0] = bf16_to_f32((unsigned short)(q.x & 0xffffu)); v[1] = bf16_to_f32((unsigned short)(q.x >> 16));
        v[2] = bf16_to_f32((unsigned short)(q.y & 0xffffu)); v[3] = bf16_to_f32((unsigned short)(q.y >> 16));
    } else if constexpr (R == 2) {
        const uint32_t q = __ldg(reinterpret_cast<const uint32_t *>(p));
        v[0] = bf16_to_f32((unsigned short)(q & 0xffffu)); v[1] = bf16_to_f32((unsigned short)(q >> 16));
    } else {
        v[0] = bf16_to_f32(__ldg(p));
    }
}

template <int R>
__device__ __forceinline__ void store_row_bf16(unsigned short *p, const float (&v)[R])
{
    if constexpr (R == 8) {
        *reinterpret_cast<uint4 *>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                                   pack_bf16x2(v[6], v[7]));
    } else if constexpr (R == 4) {
        io_store4(p, v[0], v[1], v[2], v[3]);
    } else if constexpr (R == 2) {
        io_store2(p, v[0], v[1]);
    } else {
        *p = f32_to_bf16(v[0]);
    }
}

template <int R>
__global__ void __launch_bounds__(256)
lpg_fwd_bf16_kernel(const unsigned short *__restrict__ plane_eq, unsigned short *__restrict__ depth, float *abs_min, int B, int h,
                    int w, float eps)
{
    const size_t hw = (size_t)h * w;
    const long long total = (long long)B * hw;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int key = 0x7f800000;
    if (t < total) {
        const int b = (int)(t / hw);
        const int ij = (int)(t - (long long)b * hw);
        const int i = ij / w, j = ij - i * w;
        const unsigned short *pe = plane_eq + (size_t)b * 4 * hw + ij;
        const float n1 = bf16_to_f32(__ldg(pe)), n2 = bf16_to_f32(__ldg(pe + hw));
        const float n3 = bf16_to_f32(__ldg(pe + 2 * hw)), n4 = bf16_to_f32(__ldg(pe + 3 * hw));
        const int Wd = w * R;
        unsigned short *out = depth + (size_t)b * h * R * Wd + (size_t)i * R * Wd + (size_t)j * R;
        float n1u[R];
#pragma unroll
        for (int k = 0; k < R; ++k) n1u[k] = __fmul_rn(n1, lpg_offset(k, R));
#pragma unroll
        for (int p = 0; p < R; ++p) {
            const float n2v = __fmul_rn(n2, lpg_offset(p, R));
            float o[R];
#pragma unroll
            for (int k = 0; k < R; ++k) {
                const float den = lpg_den(n1u[k], n2v, n3);
                key = min(key, absmin_key(den));
                o[k] = __fdiv_rn(n4, lpg_clamp(den, eps));
            }
            store_row_bf16<R>(out + (size_t)p * Wd, o);
        }
    }
    block_absmin_commit(key, abs_min);
}

// TP / TO: storage of plane_eq and of its gradient. (bf16, bf16) is the standalone twin; (float, float) serves the bf16
// backward of the 128-channel chains, whose planes come from the fp32 stash and whose gradient feeds the fp32 dgrad
// kernel. `scale` = 1 / max_depth when the forward wrote depth / max_depth (bts.py:255), else 1.
template <int R, class TP, class TO>
__global__ void __launch_bounds__(256)
lpg_bwd_bf16_kernel(const unsigned short *__restrict__ grad_depth, const TP *__restrict__ plane_eq,
                    TO *__restrict__ grad_plane_eq, int B, int h, int w, float eps, float scale)
{
    const size_t hw = (size_t)h * w;
    const long long total = (long long)B * hw;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / hw);
    const int ij = (int)(t - (long long)b * hw);
    const int i = ij / w, j = ij - i * w;
    const size_t base = (size_t)b * 4 * hw + ij;
    const float n1 = io_load<TP>(plane_eq + base), n2 = io_load<TP>(plane_eq + base + hw);
    const float n3 = io_load<TP>(plane_eq + base + 2 * hw), n4 = io_load<TP>(plane_eq + base + 3 * hw);
    const int Wd = w * R;
    const unsigned short *gp = grad_depth + (size_t)b * h * R * Wd + (size_t)i * R * Wd + (size_t)j * R;
    LpgAcc a{0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int p = 0; p < R; ++p) {
        float g[R];
        load_row_bf16<R>(gp + (size_t)p * Wd, g);
        const float v = lpg_offset(p, R);
        const float n2v = __fmul_rn(n2, v);
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const float u = lpg_offset(k, R);
            lpg_bwd_accum(a, g[k], __fmul_rn(n1, u), n2v, n3, u, v, eps);
        }
    }
    io_store(grad_plane_eq + base, -n4 * a.tu * scale);
    io_store(grad_plane_eq + base + hw, -n4 * a.tv * scale);
    io_store(grad_plane_eq + base + 2 * hw, -n4 * a.t * scale);
    io_store(grad_plane_eq + base + 3 * hw, a.s4 * scale);
}

static bool row_aligned(const void *p, int w, int r)
{
    const int bytes = 2 * r;                                     // one block row; rows start at multiples of it when w*r*2 is
    return (reinterpret_cast<uintptr_t>(p) % bytes) == 0 && ((size_t)w * r * 2) % bytes == 0;
}

int launch_lpg_fwd_bf16(const void *plane_eq, void *depth, float *abs_min, int B, int h, int w, int r, float eps, cudaStream_t s)
{
    if (!row_aligned(depth, w, r)) return set_error(BTS_ERR_BAD_ARG, "bf16 depth must be aligned to one block row (%d bytes)", 2 * r);
    const long long total = (long long)B * h * w;
    const unsigned grid = (unsigned)((total + 255) / 256);
    const unsigned short *pe = static_cast<const unsigned short *>(plane_eq);
    unsigned short *d = static_cast<unsigned short *>(depth);
    switch (r) {
    case 8: lpg_fwd_bf16_kernel<8><<<grid, 256, 0, s>>>(pe, d, abs_min, B, h, w, eps); break;
    case 4: lpg_fwd_bf16_kernel<4><<<grid, 256, 0, s>>>(pe, d, abs_min, B, h, w, eps); break;
    case 2: lpg_fwd_bf16_kernel<2><<<grid, 256, 0, s>>>(pe, d, abs_min, B, h, w, eps); break;
    case 1: lpg_fwd_bf16_kernel<1><<<grid, 256, 0, s>>>(pe, d, abs_min, B, h, w, eps); break;
    default: return set_error(BTS_ERR_UNSUPPORTED, "the bf16 twin serves upratio 1, 2, 4, 8 (got %d)", r);
    }
    return check_launch("bts_lpg_fwd_bf16");
}

int launch_lpg_bwd_bf16(const void *grad_depth, const void *plane_eq, void *grad_plane_eq, int B, int h, int w, int r, float eps,
                        cudaStream_t s)
{
    if (!row_aligned(grad_depth, w, r))
        return set_error(BTS_ERR_BAD_ARG, "bf16 grad_depth must be aligned to one block row (%d bytes)", 2 * r);
    const long long total = (long long)B * h * w;
    const unsigned grid = (unsigned)((total + 255) / 256);
    const unsigned short *g = static_cast<const unsigned short *>(grad_depth), *pe = static_cast<const unsigned short *>(plane_eq);
    unsigned short *o = static_cast<unsigned short *>(grad_plane_eq);
    switch (r) {
    case 8: lpg_bwd_bf16_kernel<8><<<grid, 256, 0, s>>>(g, pe, o, B, h, w, eps, 1.f); break;
    case 4: lpg_bwd_bf16_kernel<4><<<grid, 256, 0, s>>>(g, pe, o, B, h, w, eps, 1.f); break;
    case 2: lpg_bwd_bf16_kernel<2><<<grid, 256, 0, s>>>(g, pe, o, B, h, w, eps, 1.f); break;
    case 1: lpg_bwd_bf16_kernel<1><<<grid, 256, 0, s>>>(g, pe, o, B, h, w, eps, 1.f); break;
    default: return set_error(BTS_ERR_UNSUPPORTED, "the bf16 twin serves upratio 1, 2, 4, 8 (got %d)", r);
    }
    return check_launch("bts_lpg_bwd_bf16");
}

// bf16 grad_depth, fp32 planes (the forward's stash) -> fp32 gradient of the plane parameters (the dgrad kernel's input)
int launch_lpg_bwd_bf16_to_f32(const void *grad_depth, const float *plane_eq, float *grad_plane_eq, int B, int h, int w, int r,
                               float eps, float depth_div, cudaStream_t s)
{
    if (!row_aligned(grad_depth, w, r))
        return set_error(BTS_ERR_BAD_ARG, "bf16 grad_depth must be aligned to one block row (%d bytes)", 2 * r);
    const long long total = (long long)B * h * w;
    const unsigned grid = (unsigned)((total + 255) / 256);
    const unsigned short *g = static_cast<const unsigned short *>(grad_depth);
    const float sc = depth_div != 1.f ? 1.f / depth_div : 1.f;
    switch (r) {
    case 8: lpg_bwd_bf16_kernel<8><<<grid, 256, 0, s>>>(g, plane_eq, grad_plane_eq, B, h, w, eps, sc); break;
    case 4: lpg_bwd_bf16_kernel<4><<<grid, 256, 0, s>>>(g, plane_eq, grad_plane_eq, B, h, w, eps, sc); break;
    case 2: lpg_bwd_bf16_kernel<2><<<grid, 256, 0, s>>>(g, plane_eq, grad_plane_eq, B, h, w, eps, sc); break;
    case 1: lpg_bwd_bf16_kernel<1><<<grid, 256, 0, s>>>(g, plane_eq, grad_plane_eq, B, h, w, eps, sc); break;
    default: return set_error(BTS_ERR_UNSUPPORTED, "the bf16 twin serves upratio 1, 2, 4, 8 (got %d)", r);
    }
    return check_launch("bts_lpg_bwd_bf16_to_f32");
}

__global__ void __launch_bounds__(256) cvt_bf16_to_f32_kernel(const unsigned short *__restrict__ in, float *__restrict__ out, long long n)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = bf16_to_f32(__ldg(in + i));
}

int launch_cvt_bf16_to_f32(const void *in, float *out, long long n, cudaStream_t s)
{
    long long grid = (n + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    cvt_bf16_to_f32_kernel<<<(unsigned)(grid < 1 ? 1 : grid), 256, 0, s>>>(static_cast<const unsigned short *>(in), out, n);
    return check_launch("cvt_bf16_to_f32");
}

}  // namespace bts

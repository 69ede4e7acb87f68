// lpg.cu -- standalone local_planar_guidance forward / backward for sm_100a.
//
// Replaces local_planar_guidance.forward (pytorch/bts.py:149-173: ~25 eager ATen kernels plus two
// host->device uploads of full-resolution u/v grids per call) and its autograd backward with one
// HBM-streaming kernel each. Prior art: tensorflow/custom_layer/local_planar_guidance.cu (one
// thread per output pixel, scalar loads/stores, read-modify-write gradient accumulation in global
// memory, host sync after launch) -- none of which is kept.
//
// Layout. One thread owns a 128-bit column group (4 consecutive output columns) of one low-res row
// and sweeps the r output rows of that low-res row, so the four plane coefficients are loaded once
// per r*4 outputs and every warp-level store is a contiguous 512 B run of one output row.
//   forward bytes : 16 B per low-res pixel read + 4 B per output pixel written
//   backward bytes: 4 B per output pixel read + 32 B per low-res pixel (plane_eq read, grad written)
#include "common.cuh"

namespace bts {

// ------------------------------------------------------------------------------------------------
// forward, vector path: requires (w * R) % 4 == 0
// ------------------------------------------------------------------------------------------------
template <int R>
__global__ void __launch_bounds__(256)
lpg_fwd_vec_kernel(const float *__restrict__ plane_eq, float *__restrict__ depth, float *abs_min,
                   int B, int h, int w, float eps, LpgGlue gl)
{
    const int Wd = w * R;              // output width
    const int W4 = Wd >> 2;            // float4 groups per output row
    const long long total = (long long)B * h * W4;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int key = 0x7f800000;
    if (t < total) {
        const int c4 = (int)(t % W4);
        const long long bi = t / W4;
        const int i = (int)(bi % h);
        const int b = (int)(bi / h);
        const size_t hw = (size_t)h * w;
        const float *pe = plane_eq + (size_t)b * 4 * hw + (size_t)i * w;
        const int x0 = c4 * 4;         // first output column of this thread

        // per-column coefficients (a thread's 4 columns touch 1 low-res pixel for R>=4, 2 for R=2, 4 for R=1)
        float n1u[4], n2[4], n3[4], n4[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int j = (x0 + k) / R, q = (x0 + k) % R;
            if (k == 0 || R < 4) {
                n1u[k] = __ldg(pe + j);
                n2[k] = __ldg(pe + hw + j);
                n3[k] = __ldg(pe + 2 * hw + j);
                n4[k] = __ldg(pe + 3 * hw + j);
            } else {
                n1u[k] = n1u[0]; n2[k] = n2[0]; n3[k] = n3[0]; n4[k] = n4[0];
            }
            (void)q;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) n1u[k] = __fmul_rn(n1u[k], lpg_offset((x0 + k) % R, R));
        if (gl.div != 1.f) {
#pragma unroll
            for (int k = 0; k < 4; ++k) n4[k] = __fdiv_rn(n4[k], gl.div);
        }
        const size_t bstride = gl.bstride ? (size_t)gl.bstride : (size_t)h * R * Wd;
        float *out = depth + (size_t)b * bstride + (size_t)i * R * Wd + x0;
#pragma unroll
        for (int p = 0; p < R; ++p) {
            const float v = lpg_offset(p, R);
            float4 o;
            float *op = reinterpret_cast<float *>(&o);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float den = lpg_den(n1u[k], __fmul_rn(n2[k], v), n3[k]);
                key = min(key, absmin_key(den));
                den = lpg_clamp(den, eps);
                op[k] = __fdiv_rn(n4[k], den);
            }
            *reinterpret_cast<float4 *>(out + (size_t)p * Wd) = o;
            if constexpr (R == 4 || R == 8) {                      // rows/cols 0 and R/2 of every block
                if (gl.ds != nullptr && p % (R / 2) == 0) {
                    float *d = gl.ds + (size_t)b * gl.ds_bstride + (size_t)(2 * i + p / (R / 2)) * (2 * w) + x0 / (R / 2);
                    d[0] = op[0];
                    if (R == 4) d[1] = op[2];
                }
            }
        }
    }
    block_absmin_commit(key, abs_min);
}

// forward, scalar path: any r >= 1, any w. One thread per output pixel.
__global__ void __launch_bounds__(256)
lpg_fwd_scalar_kernel(const float *__restrict__ plane_eq, float *__restrict__ depth, float *abs_min,
                      int B, int h, int w, int r, float eps, LpgGlue gl)
{
    const int Wd = w * r, Hd = h * r;
    const long long total = (long long)B * Hd * Wd;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int key = 0x7f800000;
    if (t < total) {
        const int xx = (int)(t % Wd);
        const long long by = t / Wd;
        const int y = (int)(by % Hd);
        const int b = (int)(by / Hd);
        const size_t hw = (size_t)h * w;
        const float *pe = plane_eq + (size_t)b * 4 * hw + (size_t)(y / r) * w + xx / r;
        const float n1u = __fmul_rn(__ldg(pe), lpg_offset(xx % r, r));
        const float n2v = __fmul_rn(__ldg(pe + hw), lpg_offset(y % r, r));
        float den = lpg_den(n1u, n2v, __ldg(pe + 2 * hw));
        key = absmin_key(den);
        den = lpg_clamp(den, eps);
        float n4 = __ldg(pe + 3 * hw);
        if (gl.div != 1.f) n4 = __fdiv_rn(n4, gl.div);
        const float o = __fdiv_rn(n4, den);
        const size_t bstride = gl.bstride ? (size_t)gl.bstride : (size_t)Hd * Wd;
        depth[(size_t)b * bstride + (size_t)y * Wd + xx] = o;
        if (gl.ds != nullptr && (r & 1) == 0 && y % (r / 2) == 0 && xx % (r / 2) == 0)
            gl.ds[(size_t)b * gl.ds_bstride + (size_t)(y / (r / 2)) * (2 * w) + xx / (r / 2)] = o;
    }
    block_absmin_commit(key, abs_min);
}

// ------------------------------------------------------------------------------------------------
// backward
//   G4 = sum g / den_c ;  unclamped pixels: G1 = -n4 sum g u / den^2, G2 = -n4 sum g v / den^2,
//   G3 = -n4 sum g / den^2 ; clamped pixels add nothing to G1..G3 (torch.where blocks the gradient).
// ------------------------------------------------------------------------------------------------
template <int R>
__global__ void __launch_bounds__(256)
lpg_bwd_vec_kernel(const float *__restrict__ grad_depth, const float *__restrict__ plane_eq,
                   float *__restrict__ grad_plane_eq, int B, int h, int w, float eps, LpgGlue gl)
{
    constexpr int NPX = (R >= 4) ? 1 : 4 / R;    // low-res pixels fully or partly covered by a thread
    pdl_trigger();                                // the dgrad kernel behind me may set up while I run
    const int Wd = w * R;
    const int W4 = Wd >> 2;
    const long long total = (long long)B * h * W4;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = t < total;
    LpgAcc acc[NPX];
#pragma unroll
    for (int n = 0; n < NPX; ++n) acc[n] = LpgAcc{0.f, 0.f, 0.f, 0.f};
    float n4[NPX];
    size_t base = 0;
    const size_t hw = (size_t)h * w;
    if (valid) {
        const int c4 = (int)(t % W4);
        const long long bi = t / W4;
        const int i = (int)(bi % h);
        const int b = (int)(bi / h);
        const int x0 = c4 * 4;
        const int j0 = x0 / R;
        base = (size_t)b * 4 * hw + (size_t)i * w + j0;
        float n1[NPX], n2[NPX], n3[NPX];
#pragma unroll
        for (int n = 0; n < NPX; ++n) {
            n1[n] = __ldg(plane_eq + base + n);
            n2[n] = __ldg(plane_eq + base + hw + n);
            n3[n] = __ldg(plane_eq + base + 2 * hw + n);
            n4[n] = __ldg(plane_eq + base + 3 * hw + n);
        }
        const size_t bstride = gl.bstride ? (size_t)gl.bstride : (size_t)h * R * Wd;
        const float *gp = grad_depth + (size_t)b * bstride + (size_t)i * R * Wd + x0;
        float4 g[R];
#pragma unroll
        for (int p = 0; p < R; ++p) g[p] = __ldcs(reinterpret_cast<const float4 *>(gp + (size_t)p * Wd));
        if constexpr (R == 4 || R == 8) {                          // gradient of the down-sampled map
            if (gl.gds != nullptr) {
#pragma unroll
                for (int p = 0; p < R; p += R / 2) {
                    const float *d = gl.gds + (size_t)b * gl.ds_bstride + (size_t)(2 * i + p / (R / 2)) * (2 * w) + x0 / (R / 2);
                    g[p].x += __ldcs(d);
                    if (R == 4) g[p].z += __ldcs(d + 1);
                }
            }
        }
#pragma unroll
        for (int p = 0; p < R; ++p) {
            const float v = lpg_offset(p, R);
            const float *gk = reinterpret_cast<const float *>(&g[p]);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int n = (R >= 4) ? 0 : k / R;
                const float u = lpg_offset((x0 + k) % R, R);
                lpg_bwd_accum(acc[n], gk[k], __fmul_rn(n1[n], u), __fmul_rn(n2[n], v), n3[n], u, v, eps);
            }
        }
    }
    // lanes sharing a low-res pixel (R/4 adjacent lanes when R > 4) combine; pairs never straddle a
    // warp because W4 is a multiple of R/4 and blockDim is a multiple of 32.
    if (R > 4) {
#pragma unroll
        for (int o = 1; o < R / 4; o <<= 1) {
            acc[0].s4 += __shfl_xor_sync(0xffffffffu, acc[0].s4, o);
            acc[0].t += __shfl_xor_sync(0xffffffffu, acc[0].t, o);
            acc[0].tu += __shfl_xor_sync(0xffffffffu, acc[0].tu, o);
            acc[0].tv += __shfl_xor_sync(0xffffffffu, acc[0].tv, o);
        }
    }
    if (valid && (R <= 4 || (threadIdx.x & (R / 4 - 1)) == 0)) {
        const float sc = gl.div != 1.f ? __fdiv_rn(1.f, gl.div) : 1.f;
#pragma unroll
        for (int n = 0; n < NPX; ++n) {
            grad_plane_eq[base + n] = -n4[n] * acc[n].tu * sc;
            grad_plane_eq[base + hw + n] = -n4[n] * acc[n].tv * sc;
            grad_plane_eq[base + 2 * hw + n] = -n4[n] * acc[n].t * sc;
            grad_plane_eq[base + 3 * hw + n] = acc[n].s4 * sc;
        }
    }
}

// backward, scalar path: one thread per low-res pixel, any r / w.
__global__ void __launch_bounds__(256)
lpg_bwd_scalar_kernel(const float *__restrict__ grad_depth, const float *__restrict__ plane_eq,
                      float *__restrict__ grad_plane_eq, int B, int h, int w, int r, float eps, LpgGlue gl)
{
    const size_t hw = (size_t)h * w;
    const long long total = (long long)B * hw;
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / hw);
    const int ij = (int)(t % hw);
    const int i = ij / w, j = ij % w;
    const size_t base = (size_t)b * 4 * hw + ij;
    const float n1 = __ldg(plane_eq + base), n2 = __ldg(plane_eq + base + hw);
    const float n3 = __ldg(plane_eq + base + 2 * hw), n4 = __ldg(plane_eq + base + 3 * hw);
    const int Wd = w * r;
    const size_t bstride = gl.bstride ? (size_t)gl.bstride : (size_t)h * r * Wd;
    const float *gp = grad_depth + (size_t)b * bstride + (size_t)i * r * Wd + (size_t)j * r;
    const bool ds = gl.gds != nullptr && (r & 1) == 0;
    LpgAcc a{0.f, 0.f, 0.f, 0.f};
    for (int p = 0; p < r; ++p) {
        const float v = lpg_offset(p, r);
        for (int q = 0; q < r; ++q) {
            const float u = lpg_offset(q, r);
            float g = __ldg(gp + (size_t)p * Wd + q);
            if (ds && p % (r / 2) == 0 && q % (r / 2) == 0)
                g += __ldg(gl.gds + (size_t)b * gl.ds_bstride + (size_t)(2 * i + p / (r / 2)) * (2 * w) + 2 * j + q / (r / 2));
            lpg_bwd_accum(a, g, __fmul_rn(n1, u), __fmul_rn(n2, v), n3, u, v, eps);
        }
    }
    const float sc = gl.div != 1.f ? __fdiv_rn(1.f, gl.div) : 1.f;
    grad_plane_eq[base] = -n4 * a.tu * sc;
    grad_plane_eq[base + hw] = -n4 * a.tv * sc;
    grad_plane_eq[base + 2 * hw] = -n4 * a.t * sc;
    grad_plane_eq[base + 3 * hw] = a.s4 * sc;
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int launch_lpg_fwd_glue(const float *plane_eq, float *depth, float *abs_min, int B, int h, int w, int r,
                        float eps, const LpgGlue &gl, cudaStream_t s);
int launch_lpg_bwd_glue(const float *grad_depth, const float *plane_eq, float *grad_plane_eq, int B, int h,
                        int w, int r, float eps, const LpgGlue &gl, cudaStream_t s);

int launch_lpg_fwd(const float *plane_eq, float *depth, float *abs_min, int B, int h, int w, int r,
                   float eps, cudaStream_t s)
{
    return launch_lpg_fwd_glue(plane_eq, depth, abs_min, B, h, w, r, eps, lpg_glue_none(), s);
}

int launch_lpg_bwd(const float *grad_depth, const float *plane_eq, float *grad_plane_eq, int B, int h,
                   int w, int r, float eps, cudaStream_t s)
{
    return launch_lpg_bwd_glue(grad_depth, plane_eq, grad_plane_eq, B, h, w, r, eps, lpg_glue_none(), s);
}

int launch_lpg_fwd_glue(const float *plane_eq, float *depth, float *abs_min, int B, int h, int w, int r,
                        float eps, const LpgGlue &gl, cudaStream_t s)
{
    const bool vec = ((w * r) % 4 == 0) && aligned16(depth) && gl.bstride % 4 == 0 && (r == 1 || r == 2 || r == 4 || r == 8);
    if (vec) {
        const long long total = (long long)B * h * (w * r / 4);
        const unsigned grid = (unsigned)((total + 255) / 256);
        switch (r) {
        case 8: lpg_fwd_vec_kernel<8><<<grid, 256, 0, s>>>(plane_eq, depth, abs_min, B, h, w, eps, gl); break;
        case 4: lpg_fwd_vec_kernel<4><<<grid, 256, 0, s>>>(plane_eq, depth, abs_min, B, h, w, eps, gl); break;
        case 2: lpg_fwd_vec_kernel<2><<<grid, 256, 0, s>>>(plane_eq, depth, abs_min, B, h, w, eps, gl); break;
        default: lpg_fwd_vec_kernel<1><<<grid, 256, 0, s>>>(plane_eq, depth, abs_min, B, h, w, eps, gl); break;
        }
    } else {
        const long long total = (long long)B * h * r * w * r;
        const unsigned grid = (unsigned)((total + 255) / 256);
        lpg_fwd_scalar_kernel<<<grid, 256, 0, s>>>(plane_eq, depth, abs_min, B, h, w, r, eps, gl);
    }
    return check_launch("bts_lpg_fwd");
}

int launch_lpg_bwd_glue(const float *grad_depth, const float *plane_eq, float *grad_plane_eq, int B, int h,
                        int w, int r, float eps, const LpgGlue &gl, cudaStream_t s)
{
    const bool vec = ((w * r) % 4 == 0) && aligned16(grad_depth) && gl.bstride % 4 == 0 && (r == 1 || r == 2 || r == 4 || r == 8);
    if (vec) {
        const long long total = (long long)B * h * (w * r / 4);
        const unsigned grid = (unsigned)((total + 255) / 256);
        switch (r) {
        case 8: lpg_bwd_vec_kernel<8><<<grid, 256, 0, s>>>(grad_depth, plane_eq, grad_plane_eq, B, h, w, eps, gl); break;
        case 4: lpg_bwd_vec_kernel<4><<<grid, 256, 0, s>>>(grad_depth, plane_eq, grad_plane_eq, B, h, w, eps, gl); break;
        case 2: lpg_bwd_vec_kernel<2><<<grid, 256, 0, s>>>(grad_depth, plane_eq, grad_plane_eq, B, h, w, eps, gl); break;
        default: lpg_bwd_vec_kernel<1><<<grid, 256, 0, s>>>(grad_depth, plane_eq, grad_plane_eq, B, h, w, eps, gl); break;
        }
    } else {
        const long long total = (long long)B * h * w;
        const unsigned grid = (unsigned)((total + 255) / 256);
        lpg_bwd_scalar_kernel<<<grid, 256, 0, s>>>(grad_depth, plane_eq, grad_plane_eq, B, h, w, r, eps, gl);
    }
    return check_launch("bts_lpg_bwd");
}

}  // namespace bts

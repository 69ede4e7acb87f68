// common.cuh -- shared device helpers for the BTS LPG decoder kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/bts_b200.h"

namespace bts {

constexpr int kWarp = 32;

// Chain description handed to kernels by value (kernel parameter space, no device allocation).
struct ChainParams {
    const float *W[BTS_MAX_LAYERS];
    float *gW[BTS_MAX_LAYERS];
    int dims[BTS_MAX_LAYERS + 1];
    int n_layers;
    int is_final;
    float max_depth;
};

// ELU(alpha=1) as ATen evaluates it: t > 0 ? t : exp(t) - 1        (pytorch/bts.py:116-119)
__device__ __forceinline__ float elu(float t) { return t > 0.f ? t : expf(t) - 1.f; }
// ELU'(t) expressed through a = ELU(t): 1 if a > 0 else exp(t) = a + 1
__device__ __forceinline__ float elu_grad_from_out(float a) { return a > 0.f ? 1.f : a + 1.f; }
__device__ __forceinline__ float sigmoidf(float t) { return 1.f / (1.f + expf(-t)); }

constexpr float kPi = 3.14159265358979323846f;

// Plane parameterisation of reduction_1x1.forward (pytorch/bts.py:127-134).
// z[3] logits -> (n1,n2,n3,n4). Python evaluates sigmoid(z)*math.pi/3 left to right in fp32.
struct PlaneFwd {
    float n1, n2, n3, n4;
    float s0, s1, s2;          // sigmoids
    float st, ct, sp, cp;      // sin/cos of theta, phi
};
__device__ __forceinline__ PlaneFwd plane_from_logits(float z0, float z1, float z2, float max_depth)
{
    PlaneFwd p;
    p.s0 = sigmoidf(z0); p.s1 = sigmoidf(z1); p.s2 = sigmoidf(z2);
    const float theta = __fdiv_rn(__fmul_rn(p.s0, kPi), 3.f);
    const float phi = __fmul_rn(__fmul_rn(p.s1, kPi), 2.f);
    sincosf(theta, &p.st, &p.ct);
    sincosf(phi, &p.sp, &p.cp);
    p.n1 = p.st * p.cp; p.n2 = p.st * p.sp; p.n3 = p.ct;
    p.n4 = p.s2 * max_depth;
    return p;
}
// Backward of the parameterisation: (g1..g4) -> dz[3]                
__device__ __forceinline__ void plane_bwd(const PlaneFwd &p, float g1, float g2, float g3, float g4,
                                          float max_depth, float dz[3])
{
    const float dtheta = g1 * p.ct * p.cp + g2 * p.ct * p.sp - g3 * p.st;
    const float dphi = -g1 * p.st * p.sp + g2 * p.st * p.cp;
    dz[0] = dtheta * (kPi / 3.f) * p.s0 * (1.f - p.s0);
    dz[1] = dphi * (kPi * 2.f) * p.s1 * (1.f - p.s1);
    dz[2] = g4 * max_depth * p.s2 * (1.f - p.s2);
}

// LPG sub-pixel offset: u_q = (q - (r-1)/2) / r  -- exact in fp32 for power-of-two r (bts.py:157-161)
__device__ __forceinline__ float lpg_offset(int q, int r)
{
    return __fdiv_rn(__fsub_rn((float)q, __fmul_rn((float)(r - 1), 0.5f)), (float)r);
}
// den = (n1*u + n2*v) + n3 with the reference's separately rounded products (bts.py:166)
__device__ __forceinline__ float lpg_den(float n1u, float n2v, float n3)
{
    return __fadd_rn(__fadd_rn(n1u, n2v), n3);
}
// clamp of bts.py:168-171; den == 0 stays 0
__device__ __forceinline__ float lpg_clamp(float den, float eps)
{
    den = (den > 0.f && den < eps) ? eps : den;
    den = (den < 0.f && den > -eps) ? -eps : den;
    return den;
}

// ---- LPG backward accumulation for one output pixel
// post-LPG glue of the decoder (pytorch/bts.py:255-256, 269-270, 283) carried by the standalone LPG kernels when
// they serve the fused stage on the generic route / the stash backward. Defaults = plain local_planar_guidance.
struct LpgGlue {
    float div;                 // depth is divided by this (max_depth); 1 = plain
    long long bstride;         // elements between images of the full-resolution map (0 = dense H*W)
    float *ds;                 // forward: nearest down-sampling by r/2, [B][2h][2w] (nullable; even r only)
    const float *gds;          // backward: gradient of that map (nullable)
    long long ds_bstride;
};
inline LpgGlue lpg_glue_none() { return LpgGlue{1.f, 0, nullptr, nullptr, 0}; }

struct LpgAcc {
    float s4, t, tu, tv;
};
__device__ __forceinline__ void lpg_bwd_accum(LpgAcc &a, float g, float n1u, float n2v, float n3,
                                              float u, float v, float eps)
{
    const float den = lpg_den(n1u, n2v, n3);
    const float denc = lpg_clamp(den, eps);
    const float inv = __fdiv_rn(1.f, denc);
    const float gi = g * inv;
    a.s4 += gi;
    const float gi2 = (denc == den) ? gi * inv : 0.f;   // clamped => no gradient to n1..n3
    a.t += gi2;
    a.tu = fmaf(gi2, u, a.tu);
    a.tv = fmaf(gi2, v, a.tv);
}

// ---- abs_min bookkeeping: key = bits(|den|) as signed int, NaN -> 0xFFC00000 (negative => wins the
// signed min and reads back as NaN, like torch.min).
__device__ __forceinline__ int absmin_key(float den)
{
    const float a = fabsf(den);
    return (a != a) ? (int)0xFFC00000 : __float_as_int(a);
}
__device__ __forceinline__ int warp_min_int(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// Every thread of the block must call. One atomic per block, skipped when it cannot lower the value.
__device__ __forceinline__ void block_absmin_commit(int key, float *abs_min)
{
    __shared__ int s_key[32];
    key = warp_min_int(key);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_key[wid] = key;
    __syncthreads();
    if (wid == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        key = lane < nw ? s_key[lane] : 0x7f800000;
        key = warp_min_int(key);
        if (lane == 0 && abs_min != nullptr) {
            int *addr = reinterpret_cast<int *>(abs_min);
            if (key < *reinterpret_cast<volatile int *>(addr)) atomicMin(addr, key);
        }
    }
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- programmatic dependent launch between kernels of ONE C-ABI call (lpg_bwd -> dgrad -> wgrad -> reduce).
// A kernel that calls pdl_trigger() lets the next kernel of the stream, if launched with launch_dependent(), become
// resident while this one still runs; the dependent kernel sets up (weight images, barriers, TMEM) and calls
// pdl_wait() before it touches anything a predecessor wrote. Both are no-ops for plainly launched kernels.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

}  // namespace bts

// ---- host-side error plumbing (defined in cabi.cu)
namespace bts {
int set_error(int code, const char *fmt, ...);
int check_launch(const char *what);

// launch `k` as a programmatic dependent of the preceding kernel of the stream (only ever used when that kernel is
// ours and was launched by the same C-ABI call). BTS_PDL=0 falls back to a plain launch.
bool pdl_enabled();
template <class... KArgs, class... Args>
inline void launch_dependent(void (*k)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, k, KArgs(args)...);
}
}  // namespace bts

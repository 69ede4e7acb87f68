// chain_generic.cu -- reduction_1x1 forward / backward for ARBITRARY chains (the chain is data).
//
// Replaces reduction_1x1.forward (pytorch/bts.py:124-136: one cuDNN 1x1 conv + one ELU kernel per
// layer, every activation round-tripping HBM, then ~20 pointwise kernels for the plane
// parameterisation) and its autograd backward by ONE kernel each: a warp owns 32 consecutive
// low-res pixels, keeps every layer's activations in shared memory ([channel][pixel], padded to 33
// so the transposed reads of the weight-gradient step are conflict-free) and never writes an
// intermediate to HBM. This is the correctness path for dims the specialised kernels
// (chain_head.cu, chain_tc.cu) do not cover, and the on-GPU cross-check for those kernels.
#include "common.cuh"

namespace bts {

constexpr int kGP = 32;   // pixels per block (one warp)
constexpr int kGS = 33;   // padded pixel stride in shared memory

__device__ __forceinline__ int sum_dims(const ChainParams &cp)
{
    int s = 0;
    for (int l = 0; l <= cp.n_layers; ++l) s += cp.dims[l];
    return s;
}

// forward of the whole chain for the warp's 32 pixels; act = [sum dims][33]. Returns offset (in
// channels) of the last layer's logits.
__device__ __forceinline__ int chain_forward_smem(const ChainParams &cp, float *act, int lane)
{
    int off = 0;
    for (int l = 0; l < cp.n_layers; ++l) {
        const int cin = cp.dims[l], cout = cp.dims[l + 1];
        const float *a = act + (size_t)off * kGS + lane;
        float *z = act + (size_t)(off + cin) * kGS + lane;
        const float *Wl = cp.W[l];
        const bool last = (l + 1 == cp.n_layers);
        for (int o = 0; o < cout; ++o) {
            const float *wr = Wl + (size_t)o * cin;
            float s0 = 0.f, s1 = 0.f;
            int c = 0;
            for (; c + 1 < cin; c += 2) {
                s0 = fmaf(__ldg(wr + c), a[(size_t)c * kGS], s0);
                s1 = fmaf(__ldg(wr + c + 1), a[(size_t)(c + 1) * kGS], s1);
            }
            if (c < cin) s0 = fmaf(__ldg(wr + c), a[(size_t)c * kGS], s0);
            const float s = s0 + s1;
            z[(size_t)o * kGS] = last ? s : elu(s);
        }
        off += cin;
    }
    return off;
}

__global__ void __launch_bounds__(kGP)
chain_generic_fwd_kernel(ChainParams cp, const float *__restrict__ x, float *__restrict__ out,
                         int B, int hw)
{
    extern __shared__ float act[];
    const int lane = threadIdx.x;
    const long long npx = (long long)B * hw;
    const long long gp = (long long)blockIdx.x * kGP + lane;
    const bool valid = gp < npx;
    const int b = valid ? (int)(gp / hw) : 0;
    const int p = valid ? (int)(gp % hw) : 0;
    const int c0 = cp.dims[0];
    const float *xp = x + ((size_t)b * c0) * hw + p;
    for (int c = 0; c < c0; ++c) act[(size_t)c * kGS + lane] = valid ? __ldg(xp + (size_t)c * hw) : 0.f;
    const int off = chain_forward_smem(cp, act, lane);
    if (!valid) return;
    const float *z = act + (size_t)off * kGS + lane;
    if (cp.is_final) {
        out[(size_t)b * hw + p] = sigmoidf(z[0]);
    } else {
        const PlaneFwd pl = plane_from_logits(z[0], z[kGS], z[2 * kGS], cp.max_depth);
        float *o = out + (size_t)b * 4 * hw + p;
        o[0] = pl.n1; o[(size_t)hw] = pl.n2; o[(size_t)2 * hw] = pl.n3; o[(size_t)3 * hw] = pl.n4;
    }
}

// backward: recompute, then walk the chain backwards. delta ping-pongs in dbuf = [2][cmax][33].
__global__ void __launch_bounds__(kGP)
chain_generic_bwd_kernel(ChainParams cp, const float *__restrict__ x, const float *__restrict__ grad_out,
                         float *__restrict__ grad_x, int B, int hw, int tot, int cmax)
{
    extern __shared__ float smem[];
    float *act = smem;
    float *dA = smem + (size_t)tot * kGS;
    float *dB = dA + (size_t)cmax * kGS;
    const int lane = threadIdx.x;
    const long long npx = (long long)B * hw;
    const long long gp = (long long)blockIdx.x * kGP + lane;
    const bool valid = gp < npx;
    const int b = valid ? (int)(gp / hw) : 0;
    const int p = valid ? (int)(gp % hw) : 0;
    const int c0 = cp.dims[0];
    const float *xp = x + ((size_t)b * c0) * hw + p;
    for (int c = 0; c < c0; ++c) act[(size_t)c * kGS + lane] = valid ? __ldg(xp + (size_t)c * hw) : 0.f;
    int off_out = chain_forward_smem(cp, act, lane);

    // upstream gradient -> delta at the last layer's logits (zero for padding lanes)
    {
        const float *z = act + (size_t)off_out * kGS + lane;
        float *d = dA + lane;
        if (cp.is_final) {
            const float s = sigmoidf(z[0]);
            d[0] = valid ? __ldg(grad_out + (size_t)b * hw + p) * s * (1.f - s) : 0.f;
        } else {
            const PlaneFwd pl = plane_from_logits(z[0], z[kGS], z[2 * kGS], cp.max_depth);
            float dz[3] = {0.f, 0.f, 0.f};
            if (valid) {
                const float *g = grad_out + (size_t)b * 4 * hw + p;
                plane_bwd(pl, __ldg(g), __ldg(g + hw), __ldg(g + (size_t)2 * hw), __ldg(g + (size_t)3 * hw),
                          cp.max_depth, dz);
            }
            d[0] = dz[0]; d[kGS] = dz[1]; d[2 * kGS] = dz[2];
        }
    }
    __syncwarp();
    float *delta = dA, *dprev = dB;
    for (int l = cp.n_layers - 1; l >= 0; --l) {
        const int cin = cp.dims[l], cout = cp.dims[l + 1];
        const int off_in = off_out - cin;
        const float *a_in = act + (size_t)off_in * kGS;
        // weight gradient: lane owns (o, c) pairs, sums over the 32 pixels (transposed smem reads)
        if (cp.gW[l] != nullptr) {
            float *gw = cp.gW[l];
            const int nw = cin * cout;
            for (int e = lane; e < nw; e += kGP) {
                const int o = e / cin, c = e - o * cin;
                const float *dr = delta + (size_t)o * kGS;
                const float *ar = a_in + (size_t)c * kGS;
                float s = 0.f;
#pragma unroll 8
                for (int q = 0; q < kGP; ++q) s = fmaf(dr[q], ar[q], s);
                atomicAdd(gw + e, s);
            }
        }
        // data gradient: lane = pixel again
        if (l > 0 || grad_x != nullptr) {
            const float *Wl = cp.W[l];
            for (int c = 0; c < cin; ++c) {
                float s0 = 0.f, s1 = 0.f;
                int o = 0;
                for (; o + 1 < cout; o += 2) {
                    s0 = fmaf(__ldg(Wl + (size_t)o * cin + c), delta[(size_t)o * kGS + lane], s0);
                    s1 = fmaf(__ldg(Wl + (size_t)(o + 1) * cin + c), delta[(size_t)(o + 1) * kGS + lane], s1);
                }
                if (o < cout) s0 = fmaf(__ldg(Wl + (size_t)o * cin + c), delta[(size_t)o * kGS + lane], s0);
                float s = s0 + s1;
                if (l > 0) s *= elu_grad_from_out(a_in[(size_t)c * kGS + lane]);
                dprev[(size_t)c * kGS + lane] = s;
            }
        }
        __syncwarp();
        float *t = delta; delta = dprev; dprev = t;
        off_out = off_in;
    }
    if (valid && grad_x != nullptr) {
        float *gx = grad_x + ((size_t)b * c0) * hw + p;
        for (int c = 0; c < c0; ++c) gx[(size_t)c * hw] = delta[(size_t)c * kGS + lane];
    }
}

static int chain_smem_floats(const ChainParams &cp, bool bwd, int *tot_out, int *cmax_out)
{
    int tot = 0, cmax = 0;
    for (int l = 0; l <= cp.n_layers; ++l) {
        tot += cp.dims[l];
        if (cp.dims[l] > cmax) cmax = cp.dims[l];
    }
    if (tot_out) *tot_out = tot;
    if (cmax_out) *cmax_out = cmax;
    return (tot + (bwd ? 2 * cmax : 0)) * kGS;
}

int launch_chain_generic_fwd(const ChainParams &cp, const float *x, float *out, int B, int hw, cudaStream_t s)
{
    const size_t smem = sizeof(float) * (size_t)chain_smem_floats(cp, false, nullptr, nullptr);
    if (smem > 227 * 1024) return set_error(BTS_ERR_UNSUPPORTED, "chain too wide for the generic kernel (%zu B smem)", smem);
    cudaError_t e = cudaFuncSetAttribute(chain_generic_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long npx = (long long)B * hw;
    const unsigned grid = (unsigned)((npx + kGP - 1) / kGP);
    chain_generic_fwd_kernel<<<grid, kGP, smem, s>>>(cp, x, out, B, hw);
    return check_launch("chain_generic_fwd");
}

int launch_chain_generic_bwd(const ChainParams &cp, const float *x, const float *grad_out, float *grad_x,
                             int B, int hw, cudaStream_t s)
{
    int tot, cmax;
    const size_t smem = sizeof(float) * (size_t)chain_smem_floats(cp, true, &tot, &cmax);
    if (smem > 227 * 1024) return set_error(BTS_ERR_UNSUPPORTED, "chain too wide for the generic kernel (%zu B smem)", smem);
    cudaError_t e = cudaFuncSetAttribute(chain_generic_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long long npx = (long long)B * hw;
    const unsigned grid = (unsigned)((npx + kGP - 1) / kGP);
    chain_generic_bwd_kernel<<<grid, kGP, smem, s>>>(cp, x, grad_out, grad_x, B, hw, tot, cmax);
    return check_launch("chain_generic_bwd");
}

}  // namespace bts

// silog.cu -- scale-invariant log loss of the BTS training step as one streaming reduction.
//
// Replaces silog_loss.forward (pytorch/bts.py:41-48; caller pytorch/bts_main.py:412, 565):
//     d = log(depth_est[mask]) - log(depth_gt[mask]);  loss = sqrt(mean(d^2) - vf * mean(d)^2) * 10
// which in eager PyTorch is two boolean-mask gathers (dynamic shape, host sync), two logs, a subtraction,
// a square, two means and a handful of scalar kernels -- and its autograd backward (index_put scatter).
// Here: one pass over (est, gt, mask) producing per-block partial sums in fp64, a one-block finish in a
// fixed order (deterministic), and a one-pass elementwise backward. No gather, no host sync.
//   forward bytes : 9 B per pixel read (est, gt, mask byte)
//   backward bytes: 9 B per pixel read + 4 B per pixel written
#include "common.cuh"

namespace bts {

constexpr int kSilogThreads = 256;

struct SilogState {        // written by the forward's finish kernel, read by the backward
    double sum_d, sum_d2, count, loss_over_10;
};

__device__ __forceinline__ void silog_acc(float e, float g, unsigned m, double &s1, double &s2, double &n)
{
    if (m) {
        const float d = logf(e) - logf(g);
        s1 += (double)d;
        s2 += (double)d * (double)d;
        n += 1.0;
    }
}

__global__ void __launch_bounds__(kSilogThreads)
silog_partial_kernel(const float *__restrict__ est, const float *__restrict__ gt, const unsigned char *__restrict__ mask,
                     long long n, int vec, double *__restrict__ partials)
{
    double s1 = 0.0, s2 = 0.0, cnt = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {                                   // 4 pixels per step: 2 x 128-bit + 1 x 32-bit loads
        const long long n4 = n >> 2;
        for (long long i = t0; i < n4; i += stride) {
            const float4 e = __ldcs(reinterpret_cast<const float4 *>(est) + i);
            const float4 g = __ldcs(reinterpret_cast<const float4 *>(gt) + i);
            const unsigned m = __ldcs(reinterpret_cast<const unsigned *>(mask) + i);
            silog_acc(e.x, g.x, m & 0xffu, s1, s2, cnt);
            silog_acc(e.y, g.y, m & 0xff00u, s1, s2, cnt);
            silog_acc(e.z, g.z, m & 0xff0000u, s1, s2, cnt);
            silog_acc(e.w, g.w, m & 0xff000000u, s1, s2, cnt);
        }
        for (long long i = (n4 << 2) + t0; i < n; i += stride) silog_acc(est[i], gt[i], mask[i], s1, s2, cnt);
    } else {
        for (long long i = t0; i < n; i += stride) silog_acc(est[i], gt[i], mask[i], s1, s2, cnt);
    }
    // block reduction in a fixed order: warp shuffles, then warp 0 over the per-warp sums
    __shared__ double red[3][kSilogThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { red[0][warp] = s1; red[1][warp] = s2; red[2][warp] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0, c = 0.0;
        for (int w = 0; w < kSilogThreads / 32; ++w) { a += red[0][w]; b += red[1][w]; c += red[2][w]; }
        partials[3 * blockIdx.x] = a; partials[3 * blockIdx.x + 1] = b; partials[3 * blockIdx.x + 2] = c;
    }
}

__global__ void __launch_bounds__(32) silog_finish_kernel(const double *__restrict__ partials, int nparts, float vf,
                                                          float *__restrict__ loss, SilogState *__restrict__ state)
{
    double a = 0.0, b = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < nparts; i += 32) { a += partials[3 * i]; b += partials[3 * i + 1]; c += partials[3 * i + 2]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    if (threadIdx.x == 0) {
        // empty mask: mean of nothing is NaN in the reference too
        const double m1 = a / c, m2 = b / c;
        const double s = sqrt(m2 - (double)vf * m1 * m1);
        *loss = (float)(s * 10.0);
        state->sum_d = a; state->sum_d2 = b; state->count = c; state->loss_over_10 = s;
    }
}

// d loss / d est_i = g * 10 / (s * n) * (d_i - vf * mean(d)) / est_i   for masked pixels, 0 elsewhere
__global__ void __launch_bounds__(kSilogThreads)
silog_bwd_kernel(const float *__restrict__ est, const float *__restrict__ gt, const unsigned char *__restrict__ mask,
                 long long n, int vec, float vf, const float *__restrict__ grad_loss, const SilogState *__restrict__ state,
                 float *__restrict__ grad_est)
{
    const double cnt = state->count, s = state->loss_over_10;
    const float k = (float)((double)__ldg(grad_loss) * 10.0 / (s * cnt));
    const float vm = (float)((double)vf * state->sum_d / cnt);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    auto one = [&](float e, float g, unsigned m) { return m ? k * ((logf(e) - logf(g)) - vm) / e : 0.f; };
    if (vec) {
        const long long n4 = n >> 2;
        for (long long i = t0; i < n4; i += stride) {
            const float4 e = __ldcs(reinterpret_cast<const float4 *>(est) + i);
            const float4 g = __ldcs(reinterpret_cast<const float4 *>(gt) + i);
            const unsigned m = __ldcs(reinterpret_cast<const unsigned *>(mask) + i);
            float4 o;
            o.x = one(e.x, g.x, m & 0xffu); o.y = one(e.y, g.y, m & 0xff00u);
            o.z = one(e.z, g.z, m & 0xff0000u); o.w = one(e.w, g.w, m & 0xff000000u);
            __stcs(reinterpret_cast<float4 *>(grad_est) + i, o);
        }
        for (long long i = (n4 << 2) + t0; i < n; i += stride) grad_est[i] = one(est[i], gt[i], mask[i]);
    } else {
        for (long long i = t0; i < n; i += stride) grad_est[i] = one(est[i], gt[i], mask[i]);
    }
}

static bool al16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static bool al4(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 3u) == 0; }

static int silog_grid(long long n, int sms)
{
    long long blocks = (n / 4 + kSilogThreads - 1) / kSilogThreads;
    const long long cap = (long long)sms * 8;
    if (blocks > cap) blocks = cap;
    return blocks < 1 ? 1 : (int)blocks;
}

size_t silog_workspace_bytes(int sms) { return sizeof(double) * 3 * (size_t)sms * 8; }

int launch_silog_fwd(const float *est, const float *gt, const unsigned char *mask, long long n, float vf, float *loss,
                     void *state, void *ws, int sms, cudaStream_t s)
{
    const int vec = al16(est) && al16(gt) && al4(mask);
    const int grid = silog_grid(n, sms);
    silog_partial_kernel<<<grid, kSilogThreads, 0, s>>>(est, gt, mask, n, vec, static_cast<double *>(ws));
    if (int e = check_launch("silog_partial")) return e;
    silog_finish_kernel<<<1, 32, 0, s>>>(static_cast<const double *>(ws), grid, vf, loss, static_cast<SilogState *>(state));
    return check_launch("silog_finish");
}

int launch_silog_bwd(const float *est, const float *gt, const unsigned char *mask, long long n, float vf,
                     const float *grad_loss, const void *state, float *grad_est, int sms, cudaStream_t s)
{
    const int vec = al16(est) && al16(gt) && al4(mask) && al16(grad_est);
    silog_bwd_kernel<<<silog_grid(n, sms), kSilogThreads, 0, s>>>(est, gt, mask, n, vec, vf, grad_loss,
                                                                  static_cast<const SilogState *>(state), grad_est);
    return check_launch("silog_bwd");
}

// ------------------------------------------------------------------------------------------------
// No-allocation variant of the forward (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops).
// The scratch of bts_silog_fwd -- per-block fp64 partial sums + the 64-byte state -- lives in a file-scope device
// array instead of a caller workspace. Every GPU that loads the library holds these 96 KB + 256 B until exit.
// All calls on a device share the array and it keeps the last call's contents: every word the finish kernel
// reads is written by the partial kernel of the SAME call (blocks 0..grid-1), so nothing has to start at zero;
// two calls that overlap on the GPU (different streams) race on it.
// ------------------------------------------------------------------------------------------------
constexpr int kSilogStaticMaxBlocks = 4096;                       // 3 doubles each = 96 KB
constexpr long long kSilogStaticMaxN = 1ll << 30;                 // elements per call (BTS_SILOG_STATIC_MAX_N)
static __device__ __align__(256) double g_silog_scratch[3 * kSilogStaticMaxBlocks + 32];   // partials | state (256 B)

constexpr int kMaxDevices = 64;
static void *g_silog_scratch_ptr[kMaxDevices];                    // device address of the array, per device

// In host code `g_silog_scratch` is a host-side stand-in, not a device address: look the address up once per
// device, with that device current (it is: every entry point runs on the caller's current device).
static int silog_static_scratch(double **out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_error((int)e, "cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return set_error(BTS_ERR_UNSUPPORTED, "device ordinal %d >= %d", dev, kMaxDevices);
    void *p = __atomic_load_n(&g_silog_scratch_ptr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        e = cudaGetSymbolAddress(&p, g_silog_scratch);
        if (e != cudaSuccess || p == nullptr)
            return set_error(e != cudaSuccess ? (int)e : BTS_ERR_UNSUPPORTED, "cudaGetSymbolAddress(silog scratch): %s", cudaGetErrorString(e));
        __atomic_store_n(&g_silog_scratch_ptr[dev], p, __ATOMIC_RELEASE);   // two threads racing here store the same value
    }
    *out = static_cast<double *>(p);
    return BTS_OK;
}

}  // namespace bts

extern "C" BTS_API int bts_silog_fwd_static(const float *depth_est, const float *depth_gt, const unsigned char *mask,
                                            long long n, float variance_focus, float *loss, void *state_out, void *stream)
{
    using namespace bts;
    if (n < 0) return set_error(BTS_ERR_BAD_ARG, "negative element count");
    if (n > kSilogStaticMaxN)
        return set_error(BTS_ERR_UNSUPPORTED, "bts_silog_fwd_static: n = %lld exceeds the static scratch's limit of %lld elements "
                         "(use bts_silog_fwd with a caller workspace)", n, kSilogStaticMaxN);
    if (!depth_est || !depth_gt || !mask || !loss) return set_error(BTS_ERR_BAD_ARG, "depth_est / depth_gt / mask / loss is NULL");
    if (state_out != nullptr && (reinterpret_cast<uintptr_t>(state_out) & 15u))
        return set_error(BTS_ERR_BAD_ARG, "state_out must be 16-byte aligned");
    double *scratch = nullptr;
    if (int e = silog_static_scratch(&scratch)) return e;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaStream_t s = (cudaStream_t)stream;
    const int vec = al16(depth_est) && al16(depth_gt) && al4(mask);
    int grid = silog_grid(n, sms);                                // same grid as bts_silog_fwd => same summation order
    if (grid > kSilogStaticMaxBlocks) grid = kSilogStaticMaxBlocks;
    SilogState *state = state_out ? static_cast<SilogState *>(state_out) : reinterpret_cast<SilogState *>(scratch + 3 * kSilogStaticMaxBlocks);
    silog_partial_kernel<<<grid, kSilogThreads, 0, s>>>(depth_est, depth_gt, mask, n, vec, scratch);
    if (int e = check_launch("silog_partial(static)")) return e;
    silog_finish_kernel<<<1, 32, 0, s>>>(scratch, grid, variance_focus, loss, state);
    return check_launch("silog_finish(static)");
}

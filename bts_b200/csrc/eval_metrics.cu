// eval_metrics.cu -- the nine depth-evaluation metrics of BTS's online_eval as one masked reduction per image.
//
// Replaces, per evaluated sample (pytorch/bts_main.py:218-254 and compute_errors :87-108): the device->host copy
// of prediction and ground truth, the numpy clamp of the prediction to [min_depth_eval, max_depth_eval] (inf ->
// max, nan -> min), the validity mask (min < gt < max, optional crop rectangle), the boolean-mask gathers and
// the nine numpy reductions. Here: one pass over (pred, gt) per image with fp64 partial sums in a fixed order.
//   out[b] = [silog, abs_rel, log10, rms, sq_rel, log_rms, d1, d2, d3]          (order of eval_metrics, :84)
//   bytes : 8 B per pixel read
#include "common.cuh"

namespace bts {

constexpr int kEvalThreads = 256;
constexpr int kEvalSums = 11;   // count, d1, d2, d3, sq, logsq, abs_rel, sq_rel, err, err2(=logsq), log10

struct EvalArgs {
    const float *pred, *gt;     // [B][H][W]
    int B, H, W;
    float min_depth, max_depth;
    int top, bottom, left, right;   // crop rectangle [top, bottom) x [left, right); the whole image when no crop
    double *partials;           // [B][gridDim.x][kEvalSums]
    float *out;                 // [B][9]
    int *count;                 // [B] valid pixels (nullable)
};

__global__ void __launch_bounds__(kEvalThreads) eval_partial_kernel(EvalArgs a)
{
    const int b = blockIdx.y;
    const long long hw = (long long)a.H * a.W;
    const float *pred = a.pred + (size_t)b * hw, *gt = a.gt + (size_t)b * hw;
    double s[kEvalSums];
#pragma unroll
    for (int k = 0; k < kEvalSums; ++k) s[k] = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (long long)gridDim.x * blockDim.x) {
        const int y = (int)(i / a.W), x = (int)(i - (long long)y * a.W);
        const float g = __ldcs(gt + i);
        float p = __ldcs(pred + i);
        // bts_main.py:233-236, in that order
        if (p < a.min_depth) p = a.min_depth;
        if (p > a.max_depth) p = a.max_depth;
        if (isinf(p)) p = a.max_depth;
        if (isnan(p)) p = a.min_depth;
        const bool valid = g > a.min_depth && g < a.max_depth && y >= a.top && y < a.bottom && x >= a.left && x < a.right;
        if (valid) {
            const float th = fmaxf(g / p, p / g);
            const float d = g - p;
            const float lg = logf(g), lp = logf(p);
            const float e = lp - lg;
            s[0] += 1.0;
            s[1] += th < 1.25f ? 1.0 : 0.0;
            s[2] += (double)th < 1.25 * 1.25 ? 1.0 : 0.0;
            s[3] += (double)th < 1.25 * 1.25 * 1.25 ? 1.0 : 0.0;
            s[4] += (double)(d * d);
            s[5] += (double)((lg - lp) * (lg - lp));
            s[6] += (double)(fabsf(d) / g);
            s[7] += (double)((d * d) / g);
            s[8] += (double)e;
            s[9] += (double)(e * e);
            s[10] += (double)fabsf(log10f(p) - log10f(g));
        }
    }
    __shared__ double red[kEvalSums][kEvalThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kEvalSums; ++k) {
        double v = s[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[k][warp] = v;
    }
    __syncthreads();
    if (threadIdx.x < kEvalSums) {
        double v = 0.0;
        for (int w = 0; w < kEvalThreads / 32; ++w) v += red[threadIdx.x][w];
        a.partials[((size_t)b * gridDim.x + blockIdx.x) * kEvalSums + threadIdx.x] = v;
    }
}

__global__ void __launch_bounds__(32) eval_finish_kernel(EvalArgs a, int nparts)
{
    const int b = blockIdx.x;
    __shared__ double s[kEvalSums];
    if (threadIdx.x < kEvalSums) {
        double v = 0.0;
        for (int i = 0; i < nparts; ++i) v += a.partials[((size_t)b * nparts + i) * kEvalSums + threadIdx.x];
        s[threadIdx.x] = v;
    }
    __syncwarp();
    if (threadIdx.x == 0) {
        const double n = s[0];
        float *o = a.out + (size_t)b * 9;
        const double me = s[8] / n;
        o[0] = (float)(sqrt(s[9] / n - me * me) * 100.0);   // silog
        o[1] = (float)(s[6] / n);                           // abs_rel
        o[2] = (float)(s[10] / n);                          // log10
        o[3] = (float)sqrt(s[4] / n);                       // rms
        o[4] = (float)(s[7] / n);                           // sq_rel
        o[5] = (float)sqrt(s[5] / n);                       // log_rms
        o[6] = (float)(s[1] / n);                           // d1
        o[7] = (float)(s[2] / n);                           // d2
        o[8] = (float)(s[3] / n);                           // d3
        if (a.count) a.count[b] = (int)n;
    }
}

static int eval_blocks(long long hw, int sms, int B)
{
    long long blocks = (hw + kEvalThreads * 4 - 1) / (kEvalThreads * 4);
    long long cap = ((long long)sms * 8 + B - 1) / B;
    if (cap < 1) cap = 1;
    if (blocks > cap) blocks = cap;
    return blocks < 1 ? 1 : (int)blocks;
}

size_t eval_workspace_bytes(int B, int H, int W, int sms)
{
    return sizeof(double) * kEvalSums * (size_t)B * eval_blocks((long long)H * W, sms, B);
}

int launch_eval_metrics(const float *pred, const float *gt, int B, int H, int W, float min_depth, float max_depth,
                        int top, int bottom, int left, int right, float *out, int *count, void *ws, int sms, cudaStream_t s)
{
    EvalArgs a{pred, gt, B, H, W, min_depth, max_depth, top, bottom, left, right, static_cast<double *>(ws), out, count};
    const int nb = eval_blocks((long long)H * W, sms, B);
    eval_partial_kernel<<<dim3(nb, B), kEvalThreads, 0, s>>>(a);
    if (int e = check_launch("eval_partial")) return e;
    eval_finish_kernel<<<B, 32, 0, s>>>(a, nb);
    return check_launch("eval_finish");
}

}  // namespace bts

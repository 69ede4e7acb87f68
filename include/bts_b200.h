/*
 * bts_b200.h -- C ABI of the B200-native BTS LPG decoder hot path (libbts_b200.so).
 *
 * The reference (minghanz/bts, pytorch/bts.py) implements this path as two nn.Modules on top of
 * ATen; its only native prior art is the TF custom op in tensorflow/custom_layer/. Each entry point
 * below names the reference code it replaces. The host side (bts_b200/modules.py) keeps the
 * reference's nn.Module constructor / forward signatures and calls these through ctypes.
 *
 * Conventions
 *  - All tensors are fp32, NCHW-contiguous DEVICE pointers owned by the caller. The library
 *    allocates nothing, frees nothing and keeps no pointer after return.
 *  - `stream` is a cudaStream_t passed as void*. Every call only enqueues work on it: no
 *    cudaDeviceSynchronize / cudaStreamSynchronize anywhere (unlike the TF op's d.synchronize(),
 *    tensorflow/custom_layer/local_planar_guidance.cu:91,170). Calls run on the caller's current
 *    device; there is no global mutable state, so one thread per device (nn.DataParallel,
 *    pytorch/bts_test.py:91) is safe.
 *  - Return value: 0 on success, otherwise a BTS_ERR_* or cudaError_t code; bts_last_error()
 *    returns a thread-local message. Nothing is printed, nothing exits.
 *  - The chain of 1x1 convolutions is DATA: `dims[0..n_layers]` are the channel counts
 *    (e.g. {128,128,64,32,16,8,3}), `W[l]` is layer l's weight, PyTorch layout [dims[l+1]][dims[l]]
 *    (= nn.Conv2d.weight [out,in,1,1]). `W` and `grad_W` are HOST arrays of DEVICE pointers.
 *  - `abs_min` points at ONE device float that the caller pre-sets to +inf; kernels lower it with
 *    an atomic min of |den| (pre-clamp, pytorch/bts.py:167). NaN denominators yield NaN, as
 *    torch.min would.
 */
#ifndef BTS_B200_H
#define BTS_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define BTS_API __attribute__((visibility("default")))
#else
#define BTS_API
#endif

#define BTS_ABI_VERSION 4
#define BTS_MAX_LAYERS 8      /* reduction_1x1 builds at most log2(num_out) layers */
#define BTS_MAX_CHANNELS 512  /* per layer */

enum {
    BTS_OK = 0,
    BTS_ERR_BAD_ARG = -1,      /* null pointer, non-positive size, bad dims, misaligned pointer */
    BTS_ERR_UNSUPPORTED = -2,  /* valid request this build cannot serve (e.g. chain too wide)   */
    BTS_ERR_WORKSPACE = -3     /* workspace smaller than bts_workspace_bytes() reports          */
};

/* Which implementation serves a chain call (returned by bts_reduc_path; for tests and profiles). */
enum {
    BTS_PATH_GENERIC = 0,      /* any dims: one thread per low-res pixel                          */
    BTS_PATH_SIMT_REG = 1,     /* narrow chains (<= 32 input channels): register-resident SIMT    */
    BTS_PATH_TCGEN05 = 2,      /* wide chains: 3xTF32 tcgen05 MMAs for the >= 64-channel layers   */
    BTS_PATH_TCGEN05_SPLIT = 3 /* force only: tcgen05 path with the multi-kernel (workspace) backward  */
};

BTS_API int bts_abi_version(void);
BTS_API const char *bts_last_error(void);

/* Debug / test hook: force a path for subsequent calls of the PROCESS (-1 = automatic). Process-wide because
 * PyTorch runs autograd backward nodes on its own worker threads. Not for production use. */
BTS_API void bts_force_path(int path);
BTS_API int bts_reduc_path(const int *dims, int n_layers, int is_final);

/* Number of kernels this library has launched in the calling process (all threads). Lets a
 * benchmark state, not guess, how many of OUR kernels ran inside a timed region. */
BTS_API unsigned long long bts_launch_count(void);

/*
 * local_planar_guidance.forward(plane_eq, focal)            pytorch/bts.py:149-173
 * (TF prior art: LocalPlanarGuidanceKernel<GPU>, tensorflow/custom_layer/local_planar_guidance.cu:33-93)
 *   plane_eq [B,4,h,w] -> depth [B,h*r,w*r];  depth = n4 / clamp(n1*u + n2*v + n3, +-eps)
 * `focal` is not a parameter: the reference accepts and ignores it (bts.py:149 vs 150-173).
 * The arithmetic order is the reference's, so depth is bit-identical to the reference's fp32 result.
 */
BTS_API int bts_lpg_fwd(const float *plane_eq, float *depth, float *abs_min,
                int B, int h, int w, int r, float eps, void *stream);

/*
 * autograd backward of the above                            (implicit in the reference)
 * (TF prior art: LocalPlanarGuidanceGradKernel<GPU>, local_planar_guidance.cu:95-172 -- which
 *  omits the n4 factor; this follows PyTorch autograd)
 *   grad_depth [B,h*r,w*r], plane_eq [B,4,h,w] -> grad_plane_eq [B,4,h,w]  (fully overwritten)
 */
BTS_API int bts_lpg_bwd(const float *grad_depth, const float *plane_eq, float *grad_plane_eq,
                int B, int h, int w, int r, float eps, void *stream);

/* Scratch the chain entry points need for this problem (0 is a valid answer). */
BTS_API size_t bts_workspace_bytes(const int *dims, int n_layers, int B, int h, int w, int r);

/*
 * reduction_1x1.forward(net)                                pytorch/bts.py:124-136
 *   x [B,dims[0],h,w] -> out [B,4,h,w] = (sin t cos p, sin t sin p, cos t, dist)   (is_final == 0)
 *                        out [B,1,h,w] = sigmoid(chain(x))                          (is_final != 0)
 */
BTS_API int bts_reduc_fwd(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                  float max_depth, int B, int h, int w, float *out,
                  void *ws, size_t ws_bytes, void *stream);

/*
 * autograd backward of reduction_1x1.forward. The chain is recomputed from x; nothing is saved.
 *   grad_out shaped like out -> grad_x shaped like x (overwritten; may be NULL to skip),
 *   grad_W[l] shaped like W[l] (ACCUMULATED: caller zero-initialises; entries may be NULL only
 *   if all are NULL).
 */
BTS_API int bts_reduc_bwd(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                  float max_depth, int B, int h, int w, const float *grad_out,
                  float *grad_x, float *const *grad_W,
                  void *ws, size_t ws_bytes, void *stream);

/*
 * Fused decoder stage  reducNxN -> normalise/cat -> lpgNxN   pytorch/bts.py:249-254 (and 263-268,
 * 277-282): one pass from features to full-resolution depth.
 *   x [B,dims[0],h,w] -> depth [B,h*r,w*r];  plane_eq_out [B,4,h,w] optional (NULL to skip).
 * The normalisation (bts.py:251) is the identity on (sin t cos p, sin t sin p, cos t) up to
 * rounding and is not re-applied.
 */
BTS_API int bts_reduc_lpg_fwd(const float *x, const float *const *W, const int *dims, int n_layers,
                      float max_depth, int B, int h, int w, int r, float eps,
                      float *plane_eq_out, float *depth, float *abs_min,
                      void *ws, size_t ws_bytes, void *stream);

/* Backward of the fused stage: grad_depth [B,h*r,w*r] -> grad_x, grad_W (as bts_reduc_bwd). */
BTS_API int bts_reduc_lpg_bwd(const float *grad_depth, const float *x, const float *const *W, const int *dims,
                      int n_layers, float max_depth, int B, int h, int w, int r, float eps,
                      float *grad_x, float *const *grad_W,
                      void *ws, size_t ws_bytes, void *stream);

/*
 * Activation stash (ABI 2). PyTorch autograd keeps every intermediate of reduction_1x1 alive for the
 * backward pass (pytorch/bts.py:124-136 under autograd); the entry points above keep nothing and
 * recompute. The *_stash variants make that a per-call choice: a forward call that is handed a
 * `stash` buffer of bts_stash_bytes() bytes leaves the chain's ELU outputs (and, for the fused stage,
 * plane_eq) there, and the backward call that is handed the SAME buffer, x, W and shape skips the
 * recompute. bts_stash_bytes() == 0 means the chain is always recomputed (narrow chains, whose
 * one-pass backward is cheaper than the extra traffic; chains outside the tensor-core path).
 * stash == NULL / too small / misaligned (16 B) -> exactly the behaviour of the plain entry points.
 * CONTRACT: the stash carries no validity tag (the backward cannot see, without a host synchronisation, whether the
 * forward filled it). The forward fills it only when it takes the tensor-core route: weights and `depth` 16-byte
 * aligned, a workspace of bts_workspace_bytes() bytes, the route not forced elsewhere (bts_force_path). Hand the backward
 * a stash only from a forward call made under those conditions with the same workspace size (bts_b200/functional.py
 * allocates all of them itself and does); otherwise pass stash = NULL and the backward recomputes.
 */
BTS_API size_t bts_stash_bytes(const int *dims, int n_layers, int B, int h, int w);

BTS_API int bts_reduc_fwd_stash(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                        float max_depth, int B, int h, int w, float *out,
                        void *ws, size_t ws_bytes, void *stash, size_t stash_bytes, void *stream);
BTS_API int bts_reduc_bwd_stash(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                        float max_depth, int B, int h, int w, const float *grad_out,
                        float *grad_x, float *const *grad_W,
                        void *ws, size_t ws_bytes, const void *stash, size_t stash_bytes, void *stream);
BTS_API int bts_reduc_lpg_fwd_stash(const float *x, const float *const *W, const int *dims, int n_layers,
                        float max_depth, int B, int h, int w, int r, float eps,
                        float *plane_eq_out, float *depth, float *abs_min,
                        void *ws, size_t ws_bytes, void *stash, size_t stash_bytes, void *stream);
BTS_API int bts_reduc_lpg_bwd_stash(const float *grad_depth, const float *x, const float *const *W, const int *dims,
                        int n_layers, float max_depth, int B, int h, int w, int r, float eps,
                        float *grad_x, float *const *grad_W,
                        void *ws, size_t ws_bytes, const void *stash, size_t stash_bytes, void *stream);

/*
 * Post-LPG glue folded into the fused stage (ABI 3)            pytorch/bts.py:255-256, 269-270, 283
 *   depth_NxN_scaled    = depth_NxN.unsqueeze(1) / max_depth
 *   depth_NxN_scaled_ds = F.interpolate(depth_NxN_scaled, scale_factor=2/N, mode='nearest')   (N = 8, 4)
 * The forward writes depth / depth_div (pass max_depth; 1 = plain depth) with `depth_bstride` elements between
 * images -- 0 = dense h*r*w*r, anything larger lets the map land in a channel slice of a concat buffer
 * (bts.py:260, 274, 287) -- and, when ds_out != NULL (r = 8 or 4 only), the nearest-neighbour down-sampling by r/2
 * (= rows / columns 0 and r/2 of every r x r block) as [B][2h][2w] with batch stride ds_bstride (0 = dense).
 * The backward takes the gradient of the scaled map (same stride convention) and, optionally, the gradient of
 * the down-sampled map; everything else as the *_stash entry points.
 */
BTS_API int bts_reduc_lpg_fwd_glue(const float *x, const float *const *W, const int *dims, int n_layers,
                        float max_depth, int B, int h, int w, int r, float eps, float depth_div,
                        float *plane_eq_out, float *depth, long long depth_bstride,
                        float *ds_out, long long ds_bstride, float *abs_min,
                        void *ws, size_t ws_bytes, void *stash, size_t stash_bytes, void *stream);
BTS_API int bts_reduc_lpg_bwd_glue(const float *grad_depth, long long grad_bstride,
                        const float *grad_ds, long long grad_ds_bstride, float depth_div,
                        const float *x, const float *const *W, const int *dims, int n_layers,
                        float max_depth, int B, int h, int w, int r, float eps,
                        float *grad_x, float *const *grad_W,
                        void *ws, size_t ws_bytes, const void *stash, size_t stash_bytes, void *stream);

/*
 * silog_loss.forward(depth_est, depth_gt, mask)                 pytorch/bts.py:41-48
 *   d = log(depth_est[mask]) - log(depth_gt[mask]);  loss = sqrt(mean(d^2) - variance_focus * mean(d)^2) * 10
 * as one streaming reduction (no boolean-mask gather, no host synchronisation). `mask` is one byte per element
 * (torch.bool), `n` the element count of the three equally shaped tensors, `loss` ONE device float.
 * `ws` (bts_silog_workspace_bytes(), 16-byte aligned) keeps the sums the backward needs: hand the same buffer to
 * bts_silog_bwd, which writes grad_est (zeros outside the mask) for the upstream gradient `grad_loss` (one float).
 * An empty mask yields NaN, as the reference's mean of an empty tensor does.
 */
#define BTS_SILOG_STATE_BYTES 64
BTS_API size_t bts_silog_workspace_bytes(void);
BTS_API int bts_silog_fwd(const float *depth_est, const float *depth_gt, const unsigned char *mask, long long n,
                  float variance_focus, float *loss, void *ws, size_t ws_bytes, void *stream);
BTS_API int bts_silog_bwd(const float *depth_est, const float *depth_gt, const unsigned char *mask, long long n,
                  float variance_focus, const float *grad_loss, const void *ws, float *grad_est, void *stream);

/*
 * bf16 twin of the path (ABI 4; BASELINE.json north_star: "1e-2 for an optional bf16 path", configs[4] "fp32 and bf16").
 * Each entry point is the bf16-storage twin of the fp32 one of the same name (and cites the same reference lines:
 * pytorch/bts.py:124-136 reduction_1x1.forward, :149-173 local_planar_guidance.forward, :249-254 the fused stage, and
 * their autograd backward): x, out / plane_eq, depth and the gradients of those are bf16 tensors (`void *`, 2-byte
 * elements, same NCHW shapes), weights and weight gradients stay fp32 (master weights), abs_min stays one fp32 word.
 * Arithmetic is the fp32 path's (3xTF32 tensor-core layers, fp32 SIMT tail) with ONE rounding to bf16 at each store; a
 * bf16 input is exact in tf32, so its remainder pass is skipped. Bytes per full-resolution pixel are half of the fp32
 * figures of SURVEY.md 8(d). Served: forward of every tensor-core chain (bts_bf16_supported(dims, n, final, 0) == 1).
 * Backward, bts_bf16_supported(.., 1): 1 = one-pass kernel, nothing saved -- head 32->16->8->1 and 64->32->16->8->3
 * (upratio 1 / 2), when h*w % 8 == 0 and x is 16-byte aligned; 2 = stash route -- the wider chains: hand the forward a
 * `stash` of bts_stash_bytes_bf16() bytes and the backward the same buffer (as the fp32 *_stash entry points; it also
 * keeps an fp32 copy of the features for the TMA weight-gradient kernel); 0 = not served. Anything not served returns
 * BTS_ERR_UNSUPPORTED and the caller converts to fp32 (bts_b200/functional.py does). depth_div: 1, or max_depth for
 * depth / max_depth (bts.py:255). Workspace: bts_workspace_bytes_bf16(). No strided destinations and no down-sampled map.
 */
BTS_API int bts_bf16_supported(const int *dims, int n_layers, int is_final, int backward);
BTS_API size_t bts_workspace_bytes_bf16(const int *dims, int n_layers, int B, int h, int w, int r);
BTS_API size_t bts_stash_bytes_bf16(const int *dims, int n_layers, int B, int h, int w);
BTS_API int bts_lpg_fwd_bf16(const void *plane_eq, void *depth, float *abs_min, int B, int h, int w, int r, float eps,
                             void *stream);
BTS_API int bts_lpg_bwd_bf16(const void *grad_depth, const void *plane_eq, void *grad_plane_eq, int B, int h, int w, int r,
                             float eps, void *stream);
BTS_API int bts_reduc_fwd_bf16(const void *x, const float *const *W, const int *dims, int n_layers, int is_final,
                               float max_depth, int B, int h, int w, void *out, void *ws, size_t ws_bytes,
                               void *stash, size_t stash_bytes, void *stream);
BTS_API int bts_reduc_bwd_bf16(const void *x, const float *const *W, const int *dims, int n_layers, int is_final,
                               float max_depth, int B, int h, int w, const void *grad_out, void *grad_x,
                               float *const *grad_W, void *ws, size_t ws_bytes, const void *stash, size_t stash_bytes,
                               void *stream);
BTS_API int bts_reduc_lpg_fwd_bf16(const void *x, const float *const *W, const int *dims, int n_layers, float max_depth,
                                   int B, int h, int w, int r, float eps, float depth_div, void *plane_eq_out, void *depth,
                                   float *abs_min, void *ws, size_t ws_bytes, void *stash, size_t stash_bytes, void *stream);
BTS_API int bts_reduc_lpg_bwd_bf16(const void *grad_depth, float depth_div, const void *x, const float *const *W,
                                   const int *dims, int n_layers, float max_depth, int B, int h, int w, int r, float eps,
                                   void *grad_x, float *const *grad_W, void *ws, size_t ws_bytes, const void *stash,
                                   size_t stash_bytes, void *stream);

/*
 * bts_silog_fwd without a workspace and without any allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it). The scratch (per-block fp64 partial sums) is a file-scope `static __device__` array
 * of 96 KB + 256 B that every GPU that loads the library holds until exit; the entry point calls no cudaMalloc /
 * cudaFree / allocating routine. Limit: n <= BTS_SILOG_STATIC_MAX_N elements per call (BTS_ERR_UNSUPPORTED beyond;
 * use bts_silog_fwd). `state_out`: NULL, or BTS_SILOG_STATE_BYTES of caller device memory (16-byte aligned) that
 * receives the sums bts_silog_bwd needs -- pass it there as `ws`. Result = bts_silog_fwd's, bit for bit (same grid,
 * same summation order). All calls on a device SHARE the array, which keeps the last call's contents (every word
 * read is rewritten by the same call, nothing needs clearing): calls that overlap on the GPU (two streams) race.
 */
#define BTS_SILOG_STATIC_MAX_N (1ll << 30)
BTS_API int bts_silog_fwd_static(const float *depth_est, const float *depth_gt, const unsigned char *mask, long long n,
                                 float variance_focus, float *loss, void *state_out, void *stream);

/*
 * upconv.forward                                         pytorch/bts.py:83-94 (callers :226, 231, 258, 272, 285)
 * (SURVEY.md 8(f)-3), FORWARD:  y = ELU(conv3x3(interpolate(x, scale_factor=2, mode='nearest'), padding=1, bias=False)).
 * x [B][cin][h][w], weight [cout][cin][3][3] (the module's `conv.weight`), y [B][cout][2h][2w], all fp32, contiguous.
 * The up-sampled tensor is never materialised: the nearest x2 is folded into the weights (four phases, each a 2x2
 * convolution of the low-resolution input with sums of the 3x3 taps: 16 cin multiply-adds per low-res pixel and output
 * channel instead of 36 cin), run as an implicit GEMM on tensor cores with fp32-accurate products, ELU in the epilogue.
 * `ws`: bts_upconv_workspace_bytes() bytes (the folded weights and the input, both pre-split into ready MMA fragments by
 * two small kernels of the same call), 16-byte aligned.
 * Served: cin % 16 == 0 and cout % 32 == 0 (every upconv of the BTS decoders); BTS_ERR_UNSUPPORTED otherwise.
 *
 * bts_upconv_bwd: its autograd backward. y = the forward's output (ELU' is taken through it: y > 0 ? 1 : y + 1),
 * grad_y like y; grad_x [B][cin][h][w] and grad_weight [cout][cin][3][3] are WRITTEN (either may be NULL = not wanted).
 * e = grad_y . ELU'(y) is split into its four phase planes, grad_x gathers them through the folded weights, the folded
 * weight gradient has the pixel axis as K (split over CTAs, fp32 atomics) and is un-folded into the 3x3 taps.
 * `ws`: bts_upconv_bwd_workspace_bytes() bytes (folded weights, folded weight gradients, phase planes), 16-byte aligned.
 * Served: cin % 32 == 0 and cout % 32 == 0.
 */
BTS_API size_t bts_upconv_workspace_bytes(int B, int cin, int cout, int h, int w);
BTS_API int bts_upconv_fwd(const float *x, const float *weight, int B, int cin, int cout, int h, int w, float *y,
                           void *ws, size_t ws_bytes, void *stream);
BTS_API size_t bts_upconv_bwd_workspace_bytes(int B, int cin, int cout, int h, int w);
BTS_API int bts_upconv_bwd(const float *x, const float *weight, const float *y, const float *grad_y, int B, int cin, int cout,
                           int h, int w, float *grad_x, float *grad_weight, void *ws, size_t ws_bytes, void *stream);

/*
 * online_eval's per-sample metrics                       pytorch/bts_main.py:218-254, compute_errors :87-108
 * (SURVEY.md 8(f)-5). For each of the B images: clamp the prediction to [min_depth, max_depth] (inf -> max,
 * nan -> min, :233-236), mask = min_depth < gt < max_depth inside the crop rectangle [crop_top, crop_bottom) x
 * [crop_left, crop_right) (:238-252; pass 0, H, 0, W for no crop), and reduce the nine metrics in ONE pass:
 *   out[b][0..8] = silog, abs_rel, log10, rms, sq_rel, log_rms, d1, d2, d3     (order of eval_metrics, :84)
 * `count[b]` (nullable) receives the number of valid pixels. No device->host copy, no boolean-mask gather.
 */
BTS_API size_t bts_eval_workspace_bytes(int B, int H, int W);
BTS_API int bts_eval_metrics(const float *pred, const float *gt, int B, int H, int W, float min_depth, float max_depth,
                     int crop_top, int crop_bottom, int crop_left, int crop_right, float *out, int *count,
                     void *ws, size_t ws_bytes, void *stream);

/*
 * Live-viewer point cloud, headless            pytorch/bts_live_3d.py:132-136 (edges), 414-429 (worldCoords,
 * posFromDepth). depth [B][H][W] -> points [B][H*W][3]:
 *   hide (z := hidden) where |sobel_rows| + |sobel_cols| > edge_thresh  (scipy.ndimage.sobel, reflect boundary),
 *   (x, y, z) = ((u - cx) / fx * z, (v - cy) / fy * z, z).   The viewer uses fx = fy = 518.8579, cx = W/2,
 *   cy = H/2, edge_thresh = 0.3, hidden = 1e6. The OpenGL window / VBO upload is unavailable on this box; the
 *   buffer written here is what it would upload.
 */
BTS_API int bts_backproject(const float *depth, float *points, int B, int H, int W, float fx, float fy, float cx, float cy,
                    float edge_thresh, float hidden, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* BTS_B200_H */

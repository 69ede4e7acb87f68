// cabi.cu -- extern "C" entry points of libbts_b200.so (declared in include/bts_b200.h).
// Argument validation, path selection and error plumbing only; kernels live in the other files.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "chain_tc.cuh"

namespace bts {

static thread_local char g_err[512] = "";
// Process-wide on purpose: PyTorch runs a Function's backward on its autograd worker thread, not on the thread
// that called bts_force_path(), so a thread-local hook would pin the forward only. Read with relaxed atomics.
static int g_force_path_v = -1;
#define g_force_path (__atomic_load_n(&g_force_path_v, __ATOMIC_RELAXED))

int set_error(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static unsigned long long g_launches = 0;

int check_launch(const char *what)
{
    __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error((int)e, "%s: launch failed: %s", what, cudaGetErrorString(e));
    return BTS_OK;
}

bool pdl_enabled()
{
    static const bool on = !(getenv("BTS_PDL") != nullptr && atoi(getenv("BTS_PDL")) == 0);
    return on;
}

// kernels / launchers implemented elsewhere
int launch_lpg_fwd(const float *, float *, float *, int, int, int, int, float, cudaStream_t);
int launch_lpg_bwd(const float *, const float *, float *, int, int, int, int, float, cudaStream_t);
int launch_lpg_fwd_glue(const float *, float *, float *, int, int, int, int, float, const LpgGlue &, cudaStream_t);
int launch_lpg_bwd_glue(const float *, const float *, float *, int, int, int, int, float, const LpgGlue &, cudaStream_t);
int launch_chain_generic_fwd(const ChainParams &, const float *, float *, int, int, cudaStream_t);
int launch_chain_generic_bwd(const ChainParams &, const float *, const float *, float *, int, int, cudaStream_t);
int launch_chain_tc_fwd(const ChainParams &, const TcArgs &, int r, int sms, cudaStream_t);
int launch_chain_tc_bwd(const ChainParams &, const TcArgs &, int mode, int r, void *ws, int sms, cudaStream_t);
int launch_chain_bwd_fused(const ChainParams &, const TcArgs &, int mode, int r, void *ws, int sms, cudaStream_t);
size_t upconv_workspace_bytes(int B, int cin, int cout, int h, int w);
bool upconv_supported(int cin, int cout);
int launch_upconv_fwd(const float *x, const float *w, float *y, int B, int cin, int cout, int h, int wd, void *ws, cudaStream_t s);
bool upconv_bwd_supported(int cin, int cout);
size_t upconv_bwd_workspace_bytes(int B, int cin, int cout, int h, int w);
int launch_upconv_bwd(const float *x, const float *w, const float *y, const float *gy, float *gx, float *gw, int B, int cin, int cout,
                      int h, int wd, void *ws, cudaStream_t s);
bool fused_bwd_supported(const ChainParams &);
bool tc_supported(const ChainParams &);
size_t tc_pack_bytes();
size_t tc_bwd_workspace_bytes(const ChainParams &, int B, int hw, int sms);
int launch_chain_tc_bwd_stash(const ChainParams &, const TcArgs &, int mode, int r, void *ws, int sms, cudaStream_t);
bool tc_stash_supported(const ChainParams &, int hw, const void *x);
size_t tc_stash_bytes(const ChainParams &, int B, int hw);
size_t tc_stash_bytes_bf16(const ChainParams &, int B, int hw);
int tc_stash_bind(const ChainParams &, TcArgs &, void *stash, bool clear, cudaStream_t);

size_t silog_workspace_bytes(int sms);
int launch_silog_fwd(const float *, const float *, const unsigned char *, long long, float, float *, void *, void *, int,
                     cudaStream_t);
int launch_silog_bwd(const float *, const float *, const unsigned char *, long long, float, const float *, const void *,
                     float *, int, cudaStream_t);

size_t eval_workspace_bytes(int B, int H, int W, int sms);
int launch_eval_metrics(const float *, const float *, int, int, int, float, float, int, int, int, int, float *, int *, void *,
                        int, cudaStream_t);

int launch_lpg_fwd_bf16(const void *, void *, float *, int, int, int, int, float, cudaStream_t);
int launch_lpg_bwd_bf16(const void *, const void *, void *, int, int, int, int, float, cudaStream_t);

int launch_backproject(const float *, float *, int, int, int, float, float, float, float, float, float, cudaStream_t);

static int sm_count()
{
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

static bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// tensor-core path usable for this call? (chain covered, not forced off, workspace + alignment ok)
static bool use_tc(const ChainParams &cp, const void *ws, size_t ws_bytes, const void *depth)
{
    if (g_force_path == BTS_PATH_GENERIC) return false;
    if (!tc_supported(cp)) return false;
    for (int l = 0; l < cp.n_layers; ++l)                         // the operand images are packed with 128-bit loads
        if (!aligned16(cp.W[l])) return false;
    if (ws == nullptr || ws_bytes < tc_pack_bytes() || !aligned16(ws)) return false;
    return depth == nullptr || aligned16(depth);
}

static bool use_tc_bwd(const ChainParams &cp, const void *ws, size_t ws_bytes, int B, int hw, const void *grad16)
{
    if (!use_tc(cp, ws, ws_bytes, grad16)) return false;
    return ws_bytes >= 2 * tc_pack_bytes() + tc_bwd_workspace_bytes(cp, B, hw, sm_count());
}

// the stash is honoured when the chain stashes at all, the buffer is big enough and the call takes the
// tensor-core route; forward and backward evaluate the same predicate on the same (chain, x, shape)
static bool use_stash(const ChainParams &cp, const float *x, int B, int hw, const void *stash, size_t stash_bytes)
{
    if (stash == nullptr || !aligned16(stash) || g_force_path == BTS_PATH_GENERIC) return false;
    return tc_stash_supported(cp, hw, x) && stash_bytes >= tc_stash_bytes(cp, B, hw);
}

static TcArgs make_tc_args(const float *x, int B, int h, int w, float max_depth, float eps)
{
    TcArgs a;
    memset(&a, 0, sizeof(a));
    a.x = x; a.B = B; a.h = h; a.w = w; a.hw = h * w; a.npx = (long long)B * h * w;
    a.max_depth = max_depth; a.eps = eps;
    a.x_vec16 = ((h * w) % 4 == 0 && aligned16(x)) ? 1 : 0;
    a.depth_div = 1.f;
    return a;
}

static int fill_chain(ChainParams &cp, const float *const *W, float *const *gW, const int *dims,
                      int n_layers, int is_final, float max_depth)
{
    if (dims == nullptr || W == nullptr) return set_error(BTS_ERR_BAD_ARG, "dims / W is NULL");
    if (n_layers < 1 || n_layers > BTS_MAX_LAYERS)
        return set_error(BTS_ERR_BAD_ARG, "n_layers=%d outside [1,%d]", n_layers, BTS_MAX_LAYERS);
    memset(&cp, 0, sizeof(cp));
    for (int l = 0; l <= n_layers; ++l) {
        if (dims[l] < 1 || dims[l] > BTS_MAX_CHANNELS)
            return set_error(BTS_ERR_BAD_ARG, "dims[%d]=%d outside [1,%d]", l, dims[l], BTS_MAX_CHANNELS);
        cp.dims[l] = dims[l];
    }
    const int last = dims[n_layers];
    if (is_final ? last != 1 : last != 3)
        return set_error(BTS_ERR_BAD_ARG, "last layer must have %d outputs, got %d", is_final ? 1 : 3, last);
    int n_null = 0;
    for (int l = 0; l < n_layers; ++l) {
        if (W[l] == nullptr) return set_error(BTS_ERR_BAD_ARG, "W[%d] is NULL", l);
        cp.W[l] = W[l];
        cp.gW[l] = gW ? gW[l] : nullptr;
        if (gW && gW[l] == nullptr) ++n_null;
    }
    if (gW && n_null != 0 && n_null != n_layers)
        return set_error(BTS_ERR_BAD_ARG, "grad_W entries must be all NULL or all non-NULL");
    cp.n_layers = n_layers;
    cp.is_final = is_final ? 1 : 0;
    cp.max_depth = max_depth;
    return BTS_OK;
}

static int check_shape(int B, int h, int w)
{
    if (B < 0 || h < 0 || w < 0) return set_error(BTS_ERR_BAD_ARG, "negative size B=%d h=%d w=%d", B, h, w);
    if ((long long)B * h * w > (1ll << 31) - 1) return set_error(BTS_ERR_UNSUPPORTED, "more than 2^31 low-res pixels");
    return BTS_OK;
}

}  // namespace bts

using namespace bts;

extern "C" {

int bts_abi_version(void) { return BTS_ABI_VERSION; }
const char *bts_last_error(void) { return g_err; }
void bts_force_path(int path) { __atomic_store_n(&g_force_path_v, path, __ATOMIC_RELAXED); }
unsigned long long bts_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

int bts_reduc_path(const int *dims, int n_layers, int is_final)
{
    ChainParams cp;
    const float *fake[BTS_MAX_LAYERS];
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) fake[l] = reinterpret_cast<const float *>(16);
    if (fill_chain(cp, fake, nullptr, dims, n_layers, is_final, 1.f)) return -1;
    return (g_force_path != BTS_PATH_GENERIC && tc_supported(cp)) ? BTS_PATH_TCGEN05 : BTS_PATH_GENERIC;
}

int bts_lpg_fwd(const float *plane_eq, float *depth, float *abs_min, int B, int h, int w, int r,
                float eps, void *stream)
{
    if (int e = check_shape(B, h, w)) return e;
    if (r < 1) return set_error(BTS_ERR_BAD_ARG, "upratio r=%d must be >= 1", r);
    if ((long long)B * h * w == 0) return BTS_OK;                     // empty input: nothing to do
    if (!plane_eq || !depth) return set_error(BTS_ERR_BAD_ARG, "plane_eq / depth is NULL");
    if ((long long)B * h * r * w * r > (1ll << 40)) return set_error(BTS_ERR_UNSUPPORTED, "output too large");
    return launch_lpg_fwd(plane_eq, depth, abs_min, B, h, w, r, eps, (cudaStream_t)stream);
}

int bts_lpg_bwd(const float *grad_depth, const float *plane_eq, float *grad_plane_eq, int B, int h,
                int w, int r, float eps, void *stream)
{
    if (int e = check_shape(B, h, w)) return e;
    if (r < 1) return set_error(BTS_ERR_BAD_ARG, "upratio r=%d must be >= 1", r);
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!grad_depth || !plane_eq || !grad_plane_eq)
        return set_error(BTS_ERR_BAD_ARG, "grad_depth / plane_eq / grad_plane_eq is NULL");
    return launch_lpg_bwd(grad_depth, plane_eq, grad_plane_eq, B, h, w, r, eps, (cudaStream_t)stream);
}

size_t bts_workspace_bytes(const int *dims, int n_layers, int B, int h, int w, int r)
{
    if (B <= 0 || h <= 0 || w <= 0) return 0;
    // layout: [packed weights | packed transposed weights | backward spill rows + wgrad partials]
    // (tensor-core path) or [.. | plane_eq, grad_plane_eq staging] (generic fused route, r > 0)
    const size_t npx = (size_t)B * h * w;
    size_t tail = r > 0 ? (size_t)2 * 4 * sizeof(float) * npx : 0;
    ChainParams cp;
    const float *fake[BTS_MAX_LAYERS];
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) fake[l] = reinterpret_cast<const float *>(16);
    const int last = (dims && n_layers >= 1 && n_layers <= BTS_MAX_LAYERS) ? dims[n_layers] : 0;
    if (last && fill_chain(cp, fake, nullptr, dims, n_layers, last == 1, 1.f) == BTS_OK && tc_supported(cp)) {
        const size_t t = tc_bwd_workspace_bytes(cp, B, h * w, sm_count());
        if (t > tail) tail = t;
    }
    return 2 * tc_pack_bytes() + tail;
}

size_t bts_stash_bytes(const int *dims, int n_layers, int B, int h, int w)
{
    if (B <= 0 || h <= 0 || w <= 0 || g_force_path == BTS_PATH_GENERIC) return 0;
    ChainParams cp;
    const float *fake[BTS_MAX_LAYERS];
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) fake[l] = reinterpret_cast<const float *>(16);
    const int last = (dims && n_layers >= 1 && n_layers <= BTS_MAX_LAYERS) ? dims[n_layers] : 0;
    if (!last || fill_chain(cp, fake, nullptr, dims, n_layers, last == 1, 1.f) != BTS_OK) return 0;
    if (!tc_stash_supported(cp, h * w, fake[0])) return 0;
    return tc_stash_bytes(cp, B, h * w);
}

int bts_reduc_fwd(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                  float max_depth, int B, int h, int w, float *out, void *ws, size_t ws_bytes, void *stream)
{
    return bts_reduc_fwd_stash(x, W, dims, n_layers, is_final, max_depth, B, h, w, out, ws, ws_bytes, nullptr, 0, stream);
}

int bts_reduc_fwd_stash(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                        float max_depth, int B, int h, int w, float *out, void *ws, size_t ws_bytes,
                        void *stash, size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, nullptr, dims, n_layers, is_final, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !out) return set_error(BTS_ERR_BAD_ARG, "x / out is NULL");
    if (use_tc(cp, ws, ws_bytes, nullptr)) {
        TcArgs a = make_tc_args(x, B, h, w, max_depth, 0.f);
        a.out = out;
        if (use_stash(cp, x, B, h * w, stash, stash_bytes))
            if (int e = tc_stash_bind(cp, a, stash, true, (cudaStream_t)stream)) return e;
        return launch_chain_tc_fwd(cp, a, 0, sm_count(), (cudaStream_t)stream);
    }
    return launch_chain_generic_fwd(cp, x, out, B, h * w, (cudaStream_t)stream);
}

int bts_reduc_bwd(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                  float max_depth, int B, int h, int w, const float *grad_out, float *grad_x,
                  float *const *grad_W, void *ws, size_t ws_bytes, void *stream)
{
    return bts_reduc_bwd_stash(x, W, dims, n_layers, is_final, max_depth, B, h, w, grad_out, grad_x, grad_W, ws,
                               ws_bytes, nullptr, 0, stream);
}

int bts_reduc_bwd_stash(const float *x, const float *const *W, const int *dims, int n_layers, int is_final,
                        float max_depth, int B, int h, int w, const float *grad_out, float *grad_x,
                        float *const *grad_W, void *ws, size_t ws_bytes, const void *stash, size_t stash_bytes,
                        void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, grad_W, dims, n_layers, is_final, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !grad_out) return set_error(BTS_ERR_BAD_ARG, "x / grad_out is NULL");
    if (use_tc_bwd(cp, ws, ws_bytes, B, h * w, nullptr)) {
        TcArgs a = make_tc_args(x, B, h, w, max_depth, 0.f);
        a.grad = grad_out; a.grad_x = grad_x;
        if (use_stash(cp, x, B, h * w, stash, stash_bytes)) {
            if (int e = tc_stash_bind(cp, a, const_cast<void *>(stash), false, (cudaStream_t)stream)) return e;
            return launch_chain_tc_bwd_stash(cp, a, is_final ? 0 : 1, 0, ws, sm_count(), (cudaStream_t)stream);
        }
        if (fused_bwd_supported(cp) && a.x_vec16 && g_force_path != BTS_PATH_TCGEN05_SPLIT)
            return launch_chain_bwd_fused(cp, a, is_final ? 0 : 1, 0, ws, sm_count(), (cudaStream_t)stream);
        return launch_chain_tc_bwd(cp, a, is_final ? 0 : 1, 0, ws, sm_count(), (cudaStream_t)stream);
    }
    return launch_chain_generic_bwd(cp, x, grad_out, grad_x, B, h * w, (cudaStream_t)stream);
}

int bts_reduc_lpg_fwd(const float *x, const float *const *W, const int *dims, int n_layers,
                      float max_depth, int B, int h, int w, int r, float eps, float *plane_eq_out,
                      float *depth, float *abs_min, void *ws, size_t ws_bytes, void *stream)
{
    return bts_reduc_lpg_fwd_stash(x, W, dims, n_layers, max_depth, B, h, w, r, eps, plane_eq_out, depth, abs_min, ws,
                                   ws_bytes, nullptr, 0, stream);
}

int bts_reduc_lpg_fwd_stash(const float *x, const float *const *W, const int *dims, int n_layers,
                            float max_depth, int B, int h, int w, int r, float eps, float *plane_eq_out,
                            float *depth, float *abs_min, void *ws, size_t ws_bytes, void *stash,
                            size_t stash_bytes, void *stream)
{
    return bts_reduc_lpg_fwd_glue(x, W, dims, n_layers, max_depth, B, h, w, r, eps, 1.f, plane_eq_out, depth, 0, nullptr,
                                  0, abs_min, ws, ws_bytes, stash, stash_bytes, stream);
}

static int check_glue(int r, float depth_div, long long bstride, const void *ds, long long ds_bstride, int h, int w)
{
    if (!(depth_div > 0.f)) return set_error(BTS_ERR_BAD_ARG, "depth_div must be positive");
    if (bstride != 0 && bstride < (long long)h * r * w * r)
        return set_error(BTS_ERR_BAD_ARG, "batch stride %lld smaller than one %dx%d map", bstride, h * r, w * r);
    if (ds != nullptr && r != 4 && r != 8)     // bts.py:256, 270: only lpg8x8 (x1/4) and lpg4x4 (x1/2) are down-sampled
        return set_error(BTS_ERR_BAD_ARG, "the down-sampled map exists for upratio 8 and 4 only, got %d", r);
    if (ds != nullptr && ds_bstride != 0 && ds_bstride < (long long)4 * h * w)
        return set_error(BTS_ERR_BAD_ARG, "down-sampled batch stride %lld smaller than one %dx%d map", ds_bstride, 2 * h, 2 * w);
    return BTS_OK;
}

int bts_reduc_lpg_fwd_glue(const float *x, const float *const *W, const int *dims, int n_layers,
                           float max_depth, int B, int h, int w, int r, float eps, float depth_div,
                           float *plane_eq_out, float *depth, long long depth_bstride, float *ds_out,
                           long long ds_bstride, float *abs_min, void *ws, size_t ws_bytes, void *stash,
                           size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, nullptr, dims, n_layers, 0, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if (r < 1) return set_error(BTS_ERR_BAD_ARG, "upratio r=%d must be >= 1", r);
    if (int e = check_glue(r, depth_div, depth_bstride, ds_out, ds_bstride, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !depth) return set_error(BTS_ERR_BAD_ARG, "x / depth is NULL");
    const long long dbs = depth_bstride ? depth_bstride : (long long)h * r * w * r;
    const long long sbs = ds_bstride ? ds_bstride : (long long)4 * h * w;
    if ((r == 1 || r == 2 || r == 4 || r == 8) && dbs % 4 == 0 && use_tc(cp, ws, ws_bytes, depth)) {
        TcArgs a = make_tc_args(x, B, h, w, max_depth, eps);
        a.out = plane_eq_out; a.depth = depth; a.abs_min = abs_min;
        a.depth_div = depth_div; a.depth_bstride = dbs; a.ds_out = ds_out; a.ds_bstride = sbs;
        if (use_stash(cp, x, B, h * w, stash, stash_bytes))
            if (int e = tc_stash_bind(cp, a, stash, true, (cudaStream_t)stream)) return e;
        return launch_chain_tc_fwd(cp, a, r, sm_count(), (cudaStream_t)stream);
    }
    float *pe = plane_eq_out;
    if (pe == nullptr) {
        const size_t need = tc_pack_bytes() + (size_t)4 * sizeof(float) * (size_t)B * h * w;
        if (ws == nullptr || ws_bytes < need)
            return set_error(BTS_ERR_WORKSPACE, "workspace %zu B < %zu B", ws_bytes, need);
        pe = reinterpret_cast<float *>(static_cast<char *>(ws) + tc_pack_bytes());
    }
    if (int e = launch_chain_generic_fwd(cp, x, pe, B, h * w, (cudaStream_t)stream)) return e;
    const LpgGlue gl{depth_div, dbs, ds_out, nullptr, sbs};
    return launch_lpg_fwd_glue(pe, depth, abs_min, B, h, w, r, eps, gl, (cudaStream_t)stream);
}

int bts_reduc_lpg_bwd(const float *grad_depth, const float *x, const float *const *W, const int *dims,
                      int n_layers, float max_depth, int B, int h, int w, int r, float eps,
                      float *grad_x, float *const *grad_W, void *ws, size_t ws_bytes, void *stream)
{
    return bts_reduc_lpg_bwd_stash(grad_depth, x, W, dims, n_layers, max_depth, B, h, w, r, eps, grad_x, grad_W, ws,
                                   ws_bytes, nullptr, 0, stream);
}

int bts_reduc_lpg_bwd_stash(const float *grad_depth, const float *x, const float *const *W, const int *dims,
                            int n_layers, float max_depth, int B, int h, int w, int r, float eps,
                            float *grad_x, float *const *grad_W, void *ws, size_t ws_bytes, const void *stash,
                            size_t stash_bytes, void *stream)
{
    return bts_reduc_lpg_bwd_glue(grad_depth, 0, nullptr, 0, 1.f, x, W, dims, n_layers, max_depth, B, h, w, r, eps, grad_x,
                                  grad_W, ws, ws_bytes, stash, stash_bytes, stream);
}

int bts_reduc_lpg_bwd_glue(const float *grad_depth, long long grad_bstride, const float *grad_ds,
                           long long grad_ds_bstride, float depth_div, const float *x, const float *const *W,
                           const int *dims, int n_layers, float max_depth, int B, int h, int w, int r, float eps,
                           float *grad_x, float *const *grad_W, void *ws, size_t ws_bytes, const void *stash,
                           size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, grad_W, dims, n_layers, 0, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if (r < 1) return set_error(BTS_ERR_BAD_ARG, "upratio r=%d must be >= 1", r);
    if (int e = check_glue(r, depth_div, grad_bstride, grad_ds, grad_ds_bstride, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !grad_depth) return set_error(BTS_ERR_BAD_ARG, "x / grad_depth is NULL");
    const long long gbs = grad_bstride ? grad_bstride : (long long)h * r * w * r;
    const long long sbs = grad_ds_bstride ? grad_ds_bstride : (long long)4 * h * w;
    if ((r == 1 || r == 2 || r == 4 || r == 8) && gbs % 4 == 0 && use_tc_bwd(cp, ws, ws_bytes, B, h * w, grad_depth)) {
        TcArgs a = make_tc_args(x, B, h, w, max_depth, eps);
        a.grad = grad_depth; a.grad_x = grad_x;
        a.depth_div = depth_div; a.grad_bstride = gbs; a.grad_ds = grad_ds; a.grad_ds_bstride = sbs;
        if (use_stash(cp, x, B, h * w, stash, stash_bytes)) {
            if (int e = tc_stash_bind(cp, a, const_cast<void *>(stash), false, (cudaStream_t)stream)) return e;
            return launch_chain_tc_bwd_stash(cp, a, 2, r, ws, sm_count(), (cudaStream_t)stream);
        }
        if (fused_bwd_supported(cp) && a.x_vec16 && g_force_path != BTS_PATH_TCGEN05_SPLIT)
            return launch_chain_bwd_fused(cp, a, 2, r, ws, sm_count(), (cudaStream_t)stream);
        return launch_chain_tc_bwd(cp, a, 2, r, ws, sm_count(), (cudaStream_t)stream);
    }
    const size_t one = (size_t)4 * sizeof(float) * (size_t)B * h * w;
    if (ws == nullptr || ws_bytes < 2 * tc_pack_bytes() + 2 * one)
        return set_error(BTS_ERR_WORKSPACE, "workspace %zu B < %zu B", ws_bytes, 2 * tc_pack_bytes() + 2 * one);
    float *pe = reinterpret_cast<float *>(static_cast<char *>(ws) + 2 * tc_pack_bytes());
    float *gpe = pe + (size_t)4 * B * h * w;
    cudaStream_t s = (cudaStream_t)stream;
    // generic route: recompute plane_eq, reduce grad_depth over each r x r block, then the chain.
    ChainParams cf = cp;
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) cf.gW[l] = nullptr;
    if (int e = launch_chain_generic_fwd(cf, x, pe, B, h * w, s)) return e;
    const LpgGlue gl{depth_div, gbs, nullptr, grad_ds, sbs};
    if (int e = launch_lpg_bwd_glue(grad_depth, pe, gpe, B, h, w, r, eps, gl, s)) return e;
    return launch_chain_generic_bwd(cp, x, gpe, grad_x, B, h * w, s);
}

size_t bts_silog_workspace_bytes(void) { return BTS_SILOG_STATE_BYTES + silog_workspace_bytes(sm_count()); }

int bts_silog_fwd(const float *depth_est, const float *depth_gt, const unsigned char *mask, long long n,
                  float variance_focus, float *loss, void *ws, size_t ws_bytes, void *stream)
{
    if (n < 0) return set_error(BTS_ERR_BAD_ARG, "negative element count");
    if (!depth_est || !depth_gt || !mask || !loss) return set_error(BTS_ERR_BAD_ARG, "depth_est / depth_gt / mask / loss is NULL");
    if (ws == nullptr || ws_bytes < bts_silog_workspace_bytes() || (reinterpret_cast<uintptr_t>(ws) & 15u))
        return set_error(BTS_ERR_WORKSPACE, "workspace %zu B < %zu B (or not 16-byte aligned)", ws_bytes, bts_silog_workspace_bytes());
    char *w = static_cast<char *>(ws);
    return launch_silog_fwd(depth_est, depth_gt, mask, n, variance_focus, loss, w, w + BTS_SILOG_STATE_BYTES, sm_count(),
                            (cudaStream_t)stream);
}

int bts_silog_bwd(const float *depth_est, const float *depth_gt, const unsigned char *mask, long long n,
                  float variance_focus, const float *grad_loss, const void *ws, float *grad_est, void *stream)
{
    if (n < 0) return set_error(BTS_ERR_BAD_ARG, "negative element count");
    if (!depth_est || !depth_gt || !mask || !grad_loss || !ws || !grad_est)
        return set_error(BTS_ERR_BAD_ARG, "NULL argument");
    if (n == 0) return BTS_OK;
    return launch_silog_bwd(depth_est, depth_gt, mask, n, variance_focus, grad_loss, ws, grad_est, sm_count(),
                            (cudaStream_t)stream);
}

size_t bts_eval_workspace_bytes(int B, int H, int W)
{
    if (B <= 0 || H <= 0 || W <= 0) return 0;
    return eval_workspace_bytes(B, H, W, sm_count());
}

int bts_eval_metrics(const float *pred, const float *gt, int B, int H, int W, float min_depth, float max_depth,
                     int crop_top, int crop_bottom, int crop_left, int crop_right, float *out, int *count,
                     void *ws, size_t ws_bytes, void *stream)
{
    if (B < 0 || H < 0 || W < 0) return set_error(BTS_ERR_BAD_ARG, "negative size B=%d H=%d W=%d", B, H, W);
    if ((long long)B * H * W == 0) return BTS_OK;
    if (!pred || !gt || !out) return set_error(BTS_ERR_BAD_ARG, "pred / gt / out is NULL");
    if (ws == nullptr || ws_bytes < bts_eval_workspace_bytes(B, H, W) || (reinterpret_cast<uintptr_t>(ws) & 7u))
        return set_error(BTS_ERR_WORKSPACE, "workspace %zu B < %zu B (or not 8-byte aligned)", ws_bytes,
                         bts_eval_workspace_bytes(B, H, W));
    if (B > 65535) return set_error(BTS_ERR_UNSUPPORTED, "more than 65535 images per call");
    return launch_eval_metrics(pred, gt, B, H, W, min_depth, max_depth, crop_top, crop_bottom, crop_left, crop_right, out,
                               count, ws, sm_count(), (cudaStream_t)stream);
}

int bts_backproject(const float *depth, float *points, int B, int H, int W, float fx, float fy, float cx, float cy,
                    float edge_thresh, float hidden, void *stream)
{
    if (B < 0 || H < 0 || W < 0) return set_error(BTS_ERR_BAD_ARG, "negative size B=%d H=%d W=%d", B, H, W);
    if ((long long)B * H * W == 0) return BTS_OK;
    if (!depth || !points) return set_error(BTS_ERR_BAD_ARG, "depth / points is NULL");
    if (fx == 0.f || fy == 0.f) return set_error(BTS_ERR_BAD_ARG, "zero focal length");
    if ((long long)B * H * W > (1ll << 31) - 1) return set_error(BTS_ERR_UNSUPPORTED, "more than 2^31 pixels");
    return launch_backproject(depth, points, B, H, W, fx, fy, cx, cy, edge_thresh, hidden, (cudaStream_t)stream);
}

size_t bts_upconv_workspace_bytes(int B, int cin, int cout, int h, int w)
{
    if (B < 0 || cin <= 0 || cout <= 0 || h < 0 || w < 0) return 0;
    return upconv_workspace_bytes(B, cin, cout, h, w);
}

int bts_upconv_fwd(const float *x, const float *weight, int B, int cin, int cout, int h, int w, float *y, void *ws,
                   size_t ws_bytes, void *stream)
{
    if (B < 0 || h < 0 || w < 0 || cin <= 0 || cout <= 0) return set_error(BTS_ERR_BAD_ARG, "bad size B=%d cin=%d cout=%d h=%d w=%d", B, cin, cout, h, w);
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !weight || !y) return set_error(BTS_ERR_BAD_ARG, "x / weight / y is NULL");
    if (!upconv_supported(cin, cout)) return set_error(BTS_ERR_UNSUPPORTED, "upconv: cin %% 16 == 0 and cout %% 32 == 0 required (cin=%d cout=%d)", cin, cout);
    if ((long long)B * cout * 4 * h * w > (1ll << 40) || (long long)4 * h * w > (1ll << 31) - 1) return set_error(BTS_ERR_UNSUPPORTED, "upconv: tensor too large");
    if (ws == nullptr || !aligned16(ws) || ws_bytes < upconv_workspace_bytes(B, cin, cout, h, w))
        return set_error(BTS_ERR_BAD_ARG, "upconv: workspace of %zu bytes (16-byte aligned) required", upconv_workspace_bytes(B, cin, cout, h, w));
    if (!aligned16(x) || !aligned16(weight)) return set_error(BTS_ERR_BAD_ARG, "upconv: x / weight must be 16-byte aligned");
    return launch_upconv_fwd(x, weight, y, B, cin, cout, h, w, ws, (cudaStream_t)stream);
}

size_t bts_upconv_bwd_workspace_bytes(int B, int cin, int cout, int h, int w)
{
    if (B <= 0 || cin <= 0 || cout <= 0 || h <= 0 || w <= 0) return 0;
    return upconv_bwd_workspace_bytes(B, cin, cout, h, w);
}

int bts_upconv_bwd(const float *x, const float *weight, const float *y, const float *grad_y, int B, int cin, int cout, int h, int w,
                   float *grad_x, float *grad_weight, void *ws, size_t ws_bytes, void *stream)
{
    if (B < 0 || h < 0 || w < 0 || cin <= 0 || cout <= 0) return set_error(BTS_ERR_BAD_ARG, "bad size B=%d cin=%d cout=%d h=%d w=%d", B, cin, cout, h, w);
    if (!upconv_bwd_supported(cin, cout)) return set_error(BTS_ERR_UNSUPPORTED, "upconv backward: cin %% 32 == 0 and cout %% 32 == 0 required (cin=%d cout=%d)", cin, cout);
    if ((long long)B * h * w == 0) {
        if (grad_weight != nullptr) cudaMemsetAsync(grad_weight, 0, sizeof(float) * (size_t)9 * cin * cout, (cudaStream_t)stream);
        return BTS_OK;
    }
    if (!x || !weight || !y || !grad_y) return set_error(BTS_ERR_BAD_ARG, "x / weight / y / grad_y is NULL");
    if ((long long)B * cout * 4 * h * w > (1ll << 40) || (long long)4 * h * w > (1ll << 31) - 1) return set_error(BTS_ERR_UNSUPPORTED, "upconv: tensor too large");
    if (ws == nullptr || !aligned16(ws) || ws_bytes < upconv_bwd_workspace_bytes(B, cin, cout, h, w))
        return set_error(BTS_ERR_BAD_ARG, "upconv backward: workspace of %zu bytes (16-byte aligned) required", upconv_bwd_workspace_bytes(B, cin, cout, h, w));
    if (!aligned16(x) || !aligned16(weight)) return set_error(BTS_ERR_BAD_ARG, "upconv: x / weight must be 16-byte aligned");
    return launch_upconv_bwd(x, weight, y, grad_y, grad_x, grad_weight, B, cin, cout, h, w, ws, (cudaStream_t)stream);
}

/* ------------------------------------------------------------------ bf16 twin (ABI 4) ------------------------------ */

static TcArgs make_tc_args_bf16(const void *x, int B, int h, int w, float max_depth, float eps)
{
    TcArgs a = make_tc_args(static_cast<const float *>(x), B, h, w, max_depth, eps);
    a.bf16_io = 1;
    a.x_vec16 = ((h * w) % 8 == 0 && aligned16(x)) ? 1 : 0;      // 16-byte copies move 8 bf16 pixels
    return a;
}

// bytes the bf16 entry points need: the one-pass backward only keeps per-CTA partial sums (deterministic route); the
// stash backward of the wider chains needs the delta rows and partials of the fp32 path
static size_t bf16_ws_bytes(const ChainParams &cp, int B, int hw)
{
    if (!fused_bwd_supported(cp)) return 2 * tc_pack_bytes() + tc_bwd_workspace_bytes(cp, B, hw, sm_count());
    size_t tw = 0;
    for (int l = 0; l < cp.n_layers; ++l) tw += (size_t)cp.dims[l] * cp.dims[l + 1];
    return 2 * tc_pack_bytes() + sizeof(float) * tw * (size_t)sm_count();
}

static bool fake_chain(ChainParams &cp, const int *dims, int n_layers, int is_final)
{
    static const float *fake[BTS_MAX_LAYERS];
    for (int l = 0; l < BTS_MAX_LAYERS; ++l) fake[l] = reinterpret_cast<const float *>(16);
    return dims && n_layers >= 1 && n_layers <= BTS_MAX_LAYERS && fill_chain(cp, fake, nullptr, dims, n_layers, is_final, 1.f) == BTS_OK;
}

size_t bts_workspace_bytes_bf16(const int *dims, int n_layers, int B, int h, int w, int r)
{
    (void)r;
    if (B <= 0 || h <= 0 || w <= 0) return 0;
    ChainParams cp;
    const int last = (dims && n_layers >= 1 && n_layers <= BTS_MAX_LAYERS) ? dims[n_layers] : 0;
    if (!last || !fake_chain(cp, dims, n_layers, last == 1)) return 0;
    return bf16_ws_bytes(cp, B, h * w);
}

size_t bts_stash_bytes_bf16(const int *dims, int n_layers, int B, int h, int w)
{
    if (B <= 0 || h <= 0 || w <= 0) return 0;
    ChainParams cp;
    const int last = (dims && n_layers >= 1 && n_layers <= BTS_MAX_LAYERS) ? dims[n_layers] : 0;
    if (!last || !fake_chain(cp, dims, n_layers, last == 1)) return 0;
    if (!tc_stash_supported(cp, h * w, reinterpret_cast<const void *>(16))) return 0;
    return tc_stash_bytes_bf16(cp, B, h * w);
}

// backward: 1 = one-pass kernel (no stash), 2 = stash route (hand the forward a bts_stash_bytes_bf16() buffer), 0 = none
int bts_bf16_supported(const int *dims, int n_layers, int is_final, int backward)
{
    ChainParams cp;
    if (!fake_chain(cp, dims, n_layers, is_final)) return 0;
    if (!tc_supported(cp)) return 0;
    if (!backward) return 1;
    if (fused_bwd_supported(cp)) return 1;
    return tc_stash_supported(cp, 32, reinterpret_cast<const void *>(16)) ? 2 : 0;
}

static bool use_stash_bf16(const ChainParams &cp, const void *x, int B, int hw, const void *stash, size_t stash_bytes)
{
    if (stash == nullptr || !aligned16(stash)) return false;
    return tc_stash_supported(cp, hw, x) && stash_bytes >= tc_stash_bytes_bf16(cp, B, hw);
}

int bts_lpg_fwd_bf16(const void *plane_eq, void *depth, float *abs_min, int B, int h, int w, int r, float eps, void *stream)
{
    if (int e = check_shape(B, h, w)) return e;
    if (r < 1) return set_error(BTS_ERR_BAD_ARG, "upratio r=%d must be >= 1", r);
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!plane_eq || !depth) return set_error(BTS_ERR_BAD_ARG, "plane_eq / depth is NULL");
    return launch_lpg_fwd_bf16(plane_eq, depth, abs_min, B, h, w, r, eps, (cudaStream_t)stream);
}

int bts_lpg_bwd_bf16(const void *grad_depth, const void *plane_eq, void *grad_plane_eq, int B, int h, int w, int r, float eps,
                     void *stream)
{
    if (int e = check_shape(B, h, w)) return e;
    if (r < 1) return set_error(BTS_ERR_BAD_ARG, "upratio r=%d must be >= 1", r);
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!grad_depth || !plane_eq || !grad_plane_eq) return set_error(BTS_ERR_BAD_ARG, "grad_depth / plane_eq / grad_plane_eq is NULL");
    return launch_lpg_bwd_bf16(grad_depth, plane_eq, grad_plane_eq, B, h, w, r, eps, (cudaStream_t)stream);
}

int bts_reduc_fwd_bf16(const void *x, const float *const *W, const int *dims, int n_layers, int is_final, float max_depth,
                       int B, int h, int w, void *out, void *ws, size_t ws_bytes, void *stash, size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, nullptr, dims, n_layers, is_final, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !out) return set_error(BTS_ERR_BAD_ARG, "x / out is NULL");
    if (!use_tc(cp, ws, ws_bytes, nullptr))
        return set_error(BTS_ERR_UNSUPPORTED, "the bf16 twin serves the chains of the tensor-core path (and needs a workspace)");
    TcArgs a = make_tc_args_bf16(x, B, h, w, max_depth, 0.f);
    a.out = static_cast<float *>(out);
    if (use_stash_bf16(cp, x, B, h * w, stash, stash_bytes))
        if (int e = tc_stash_bind(cp, a, stash, true, (cudaStream_t)stream)) return e;
    return launch_chain_tc_fwd(cp, a, 0, sm_count(), (cudaStream_t)stream);
}

int bts_reduc_bwd_bf16(const void *x, const float *const *W, const int *dims, int n_layers, int is_final, float max_depth,
                       int B, int h, int w, const void *grad_out, void *grad_x, float *const *grad_W, void *ws,
                       size_t ws_bytes, const void *stash, size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, grad_W, dims, n_layers, is_final, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !grad_out) return set_error(BTS_ERR_BAD_ARG, "x / grad_out is NULL");
    TcArgs a = make_tc_args_bf16(x, B, h, w, max_depth, 0.f);
    a.grad = static_cast<const float *>(grad_out);
    a.grad_x = static_cast<float *>(grad_x);
    if (!use_tc(cp, ws, ws_bytes, nullptr) || ws_bytes < bf16_ws_bytes(cp, B, h * w))
        return set_error(BTS_ERR_UNSUPPORTED, "bf16 backward: chain outside the tensor-core path, or workspace below bts_workspace_bytes_bf16()");
    if (fused_bwd_supported(cp) && a.x_vec16)
        return launch_chain_bwd_fused(cp, a, is_final ? 0 : 1, 0, ws, sm_count(), (cudaStream_t)stream);
    if (use_stash_bf16(cp, x, B, h * w, stash, stash_bytes)) {
        if (int e = tc_stash_bind(cp, a, const_cast<void *>(stash), false, (cudaStream_t)stream)) return e;
        return launch_chain_tc_bwd_stash(cp, a, is_final ? 0 : 1, 0, ws, sm_count(), (cudaStream_t)stream);
    }
    return set_error(BTS_ERR_UNSUPPORTED, "bf16 backward needs h*w %% 8 == 0 and a 16-byte aligned x (one-pass chains) or the stash the forward filled");
}

int bts_reduc_lpg_fwd_bf16(const void *x, const float *const *W, const int *dims, int n_layers, float max_depth, int B, int h,
                           int w, int r, float eps, float depth_div, void *plane_eq_out, void *depth, float *abs_min,
                           void *ws, size_t ws_bytes, void *stash, size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, nullptr, dims, n_layers, 0, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if (int e = check_glue(r < 1 ? 1 : r, depth_div, 0, nullptr, 0, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !depth) return set_error(BTS_ERR_BAD_ARG, "x / depth is NULL");
    if (!(r == 1 || r == 2 || r == 4 || r == 8) || !use_tc(cp, ws, ws_bytes, depth))
        return set_error(BTS_ERR_UNSUPPORTED, "the bf16 twin serves the tensor-core chains at upratio 1, 2, 4, 8 (16-byte aligned depth)");
    TcArgs a = make_tc_args_bf16(x, B, h, w, max_depth, eps);
    a.out = static_cast<float *>(plane_eq_out); a.depth = static_cast<float *>(depth); a.abs_min = abs_min;
    a.depth_div = depth_div; a.depth_bstride = (long long)h * r * w * r;
    if (use_stash_bf16(cp, x, B, h * w, stash, stash_bytes))
        if (int e = tc_stash_bind(cp, a, stash, true, (cudaStream_t)stream)) return e;
    return launch_chain_tc_fwd(cp, a, r, sm_count(), (cudaStream_t)stream);
}

int bts_reduc_lpg_bwd_bf16(const void *grad_depth, float depth_div, const void *x, const float *const *W, const int *dims,
                           int n_layers, float max_depth, int B, int h, int w, int r, float eps, void *grad_x,
                           float *const *grad_W, void *ws, size_t ws_bytes, const void *stash, size_t stash_bytes, void *stream)
{
    ChainParams cp;
    if (int e = fill_chain(cp, W, grad_W, dims, n_layers, 0, max_depth)) return e;
    if (int e = check_shape(B, h, w)) return e;
    if (int e = check_glue(r < 1 ? 1 : r, depth_div, 0, nullptr, 0, h, w)) return e;
    if ((long long)B * h * w == 0) return BTS_OK;
    if (!x || !grad_depth) return set_error(BTS_ERR_BAD_ARG, "x / grad_depth is NULL");
    TcArgs a = make_tc_args_bf16(x, B, h, w, max_depth, eps);
    a.grad = static_cast<const float *>(grad_depth); a.grad_x = static_cast<float *>(grad_x);
    a.depth_div = depth_div; a.grad_bstride = (long long)h * r * w * r;
    if (!(r == 1 || r == 2 || r == 4 || r == 8) || !use_tc(cp, ws, ws_bytes, grad_depth) || ws_bytes < bf16_ws_bytes(cp, B, h * w))
        return set_error(BTS_ERR_UNSUPPORTED, "bf16 fused backward: tensor-core chains at upratio 1, 2, 4, 8 with a bts_workspace_bytes_bf16() workspace");
    // the one-pass backward prefetches the r x r block of grad_depth per low-res pixel: upratio 1 and 2 only
    if (fused_bwd_supported(cp) && a.x_vec16 && r <= 2)
        return launch_chain_bwd_fused(cp, a, 2, r, ws, sm_count(), (cudaStream_t)stream);
    if (use_stash_bf16(cp, x, B, h * w, stash, stash_bytes)) {
        if (int e = tc_stash_bind(cp, a, const_cast<void *>(stash), false, (cudaStream_t)stream)) return e;
        return launch_chain_tc_bwd_stash(cp, a, 2, r, ws, sm_count(), (cudaStream_t)stream);
    }
    return set_error(BTS_ERR_UNSUPPORTED, "bf16 fused backward needs the one-pass chain (upratio <= 2, h*w %% 8 == 0) or the stash the forward filled");
}

}  // extern "C"

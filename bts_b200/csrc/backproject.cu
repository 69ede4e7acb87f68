// backproject.cu -- depth map -> 3-D point cloud of the live viewer, headless (no graphics API involved).
//
// Replaces posFromDepth / worldCoords / edges of pytorch/bts_live_3d.py:132-136, 414-429 (numpy + scipy on the
// host): Sobel edge magnitude |dx| + |dy| of the depth map (scipy.ndimage.sobel, 'reflect' boundary), depth
// hidden (set to `hidden`, 1e6 in the viewer) where the magnitude exceeds `edge_thresh` (0.3), and the pinhole
// back-projection  (x, y, z) = ((u - cx) / fx * z, (v - cy) / fy * z, z)  into a [H*W][3] device buffer -- the
// vertex buffer the viewer would upload. The GL window / VBO / swapchain part is unavailable on this box.
//   bytes: 4 B per pixel read (the 3x3 neighbourhood comes from L1/L2) + 12 B per pixel written
#include "common.cuh"

namespace bts {

__device__ __forceinline__ int reflect_idx(int i, int n)
{
    // scipy 'reflect': (d c b a | a b c d | d c b a)
    if (i < 0) i = -i - 1;
    if (i >= n) i = 2 * n - 1 - i;
    return i;
}

__global__ void __launch_bounds__(256)
backproject_kernel(const float *__restrict__ depth, float *__restrict__ points, int B, int H, int W, float fx, float fy,
                   float cx, float cy, float edge_thresh, float hidden)
{
    const long long hw = (long long)H * W;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * hw) return;
    const int b = (int)(i / hw);
    const int r = (int)(i - (long long)b * hw);
    const int y = r / W, x = r - y * W;
    const float *d = depth + (size_t)b * hw;
    float v[3][3];
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
        const int yy = reflect_idx(y + dy, H);
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) v[dy + 1][dx + 1] = __ldg(d + (size_t)yy * W + reflect_idx(x + dx, W));
    }
    // ndimage.sobel(d, axis): derivative [-1, 0, 1] along `axis`, smoothing [1, 2, 1] along the other one
    const float g0 = (v[2][0] + 2.f * v[2][1] + v[2][2]) - (v[0][0] + 2.f * v[0][1] + v[0][2]);   // axis 0 (rows)
    const float g1 = (v[0][2] + 2.f * v[1][2] + v[2][2]) - (v[0][0] + 2.f * v[1][0] + v[2][0]);   // axis 1 (columns)
    float z = v[1][1];
    if (fabsf(g0) + fabsf(g1) > edge_thresh) z = hidden;
    float *p = points + (size_t)i * 3;
    p[0] = ((float)x - cx) / fx * z;
    p[1] = ((float)y - cy) / fy * z;
    p[2] = z;
}

int launch_backproject(const float *depth, float *points, int B, int H, int W, float fx, float fy, float cx, float cy,
                       float edge_thresh, float hidden, cudaStream_t s)
{
    const long long n = (long long)B * H * W;
    backproject_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(depth, points, B, H, W, fx, fy, cx, cy, edge_thresh, hidden);
    return check_launch("backproject");
}

}  // namespace bts

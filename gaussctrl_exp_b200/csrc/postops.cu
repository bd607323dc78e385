// Depth post-ops behind the render path: what the reference does with a rendered depth map
// before the depth ControlNet.
//   * ray directions      /root/reference/gaussctrl/gc_pipeline.py:503-536 (an O(H*W) Python double loop there)
//   * depth -> world pts  gc_pipeline.py:286-314
//   * depth -> disparity  gc_pipeline.py:439-457
// One thread per pixel, no staging: each kernel streams the depth map once (HBM/L2 bound, 4-28 bytes per pixel; a
// 512x512 view is one wave). The arithmetic follows the reference's order of operations and precisions (float64 ray
// set-up rounded to float32, float32 norm and divisions, correctly rounded) so that the ray directions and the
// disparity are bit-identical to numpy / torch; the 4x4 product is within 1 ulp-class error of torch's matmul.
#include <algorithm>
#include <cmath>

#include "gs_common.cuh"

namespace {

// ray of pixel (x, y): numpy float64 arithmetic, stored as float32, then normalised in float32
// (np.linalg.norm = sqrt((x*x + y*y) + z*z), no contraction)
__device__ __forceinline__ float3 ray_direction(int x, int y, int width, int height, double vp_w, double vp_h,
                                                double near) {
    const double ndc_x = ((double)x + 0.5) / (double)width;
    const double ndc_y = ((double)y + 0.5) / (double)height;
    const float rx = (float)__dsub_rn(__dmul_rn(ndc_x, vp_w), vp_w / 2.0);      // no DFMA contraction: numpy has none
    const float ry = (float)__dsub_rn(__dmul_rn(ndc_y, vp_h), vp_h / 2.0);
    const float rz = (float)(-near);
    const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)), __fmul_rn(rz, rz));
    const float n = __fsqrt_rn(n2);
    return make_float3(__fdiv_rn(rx, n), __fdiv_rn(ry, n), __fdiv_rn(rz, n));
}

__global__ void __launch_bounds__(256)
ray_directions_kernel(int width, int height, double vp_w, double vp_h, double near, float *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)width * height) return;
    const int x = (int)(i % width), y = (int)(i / width);
    const float3 d = ray_direction(x, y, width, height, vp_w, vp_h, near);
    out[3 * i] = d.x; out[3 * i + 1] = d.y; out[3 * i + 2] = d.z;
}

struct Mat4 {
    float m[16];
};

// pts_world = inv(view) * [ray * depth, 1], divided by its w (gc_pipeline.py:300-313)
__global__ void __launch_bounds__(256)
depth_to_world_kernel(int width, int height, double vp_w, double vp_h, double near, const float *__restrict__ depth,
                      Mat4 inv, float *__restrict__ out) {
    gs_pdl_sync();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)width * height) return;
    const int x = (int)(i % width), y = (int)(i / width);
    const float3 d = ray_direction(x, y, width, height, vp_w, vp_h, near);
    const float z = depth[i];
    const float px = __fmul_rn(d.x, z), py = __fmul_rn(d.y, z), pz = __fmul_rn(d.z, z);
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
        r[k] = fmaf(inv.m[4 * k + 2], pz, fmaf(inv.m[4 * k + 1], py, fmaf(inv.m[4 * k], px, 0.f))) + inv.m[4 * k + 3];
    out[3 * i] = __fdiv_rn(r[0], r[3]);
    out[3 * i + 1] = __fdiv_rn(r[1], r[3]);
    out[3 * i + 2] = __fdiv_rn(r[2], r[3]);
}

// order-preserving map float -> unsigned, so that one atomicMax serves every sign
__device__ __forceinline__ unsigned ordered_bits(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_ordered_bits(unsigned o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__device__ __forceinline__ float disparity_of(float depth) { return __frcp_rn(__fadd_rn(depth, 1e-5f)); }

__global__ void __launch_bounds__(256)
disparity_max_kernel(int64_t n, const float *__restrict__ depth, unsigned *__restrict__ max_bits) {
    gs_pdl_sync();
    unsigned best = 0u;                                     // below every ordered value
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        best = max(best, ordered_bits(disparity_of(depth[i])));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    __shared__ unsigned s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 1; w < 8; ++w) best = max(best, s[w]);
        atomicMax(max_bits, best);
    }
}

__global__ void __launch_bounds__(256)
disparity_write_kernel(int64_t n, const float *__restrict__ depth, const unsigned *__restrict__ max_bits,
                       float *__restrict__ out) {
    gs_pdl_sync();
    const float top = from_ordered_bits(*max_bits);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = __fdiv_rn(disparity_of(depth[i]), top);
        out[i] = v; out[n + i] = v; out[2 * n + i] = v;      // the three identical channels of the ControlNet input
    }
}

// inverse of a 4x4 (row-major) in double, Gauss-Jordan with partial pivoting; false if singular
bool invert4(const float *a, float *out) {
    double m[4][8];
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) {
            m[r][c] = a[4 * r + c];
            m[r][4 + c] = r == c ? 1.0 : 0.0;
        }
    for (int c = 0; c < 4; ++c) {
        int p = c;
        for (int r = c + 1; r < 4; ++r)
            if (std::fabs(m[r][c]) > std::fabs(m[p][c])) p = r;
        if (!(std::fabs(m[p][c]) > 0.0)) return false;
        if (p != c)
            for (int k = 0; k < 8; ++k) std::swap(m[p][k], m[c][k]);
        const double inv = 1.0 / m[c][c];
        for (int k = 0; k < 8; ++k) m[c][k] *= inv;
        for (int r = 0; r < 4; ++r) {
            if (r == c) continue;
            const double f = m[r][c];
            if (f != 0.0)
                for (int k = 0; k < 8; ++k) m[r][k] -= f * m[c][k];
        }
    }
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) out[4 * r + c] = (float)m[r][4 + c];
    return true;
}

// the reference unpacks `width, height = resolution` (gc_pipeline.py:505) while its caller passes (h, w) (:299):
// the ABI keeps the reference's argument order, i.e. res0 is what the reference calls "width"
void viewport(int res0, int res1, double fov_y_rad, double near, double &vp_w, double &vp_h) {
    const double aspect = (double)res0 / (double)res1;
    vp_h = 2.0 * std::tan(fov_y_rad / 2.0) * near;
    vp_w = vp_h * aspect;
}

}  // namespace

extern "C" int gs_ray_directions(int res0, int res1, double fov_y_rad, double near, float *out, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (res0 <= 0 || res1 <= 0 || !out) return GS_ERR_BAD_ARG;
    double vp_w, vp_h;
    viewport(res0, res1, fov_y_rad, near, vp_w, vp_h);
    const int64_t n = (int64_t)res0 * res1;
    ray_directions_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(res0, res1, vp_w, vp_h, near, out);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_depth_to_world(int h, int w, const float *depth, const float *mat_view_host, double fov_y_rad,
                                 double near, float *pts_world, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (h <= 0 || w <= 0 || !depth || !mat_view_host || !pts_world) return GS_ERR_BAD_ARG;
    Mat4 inv;
    if (!invert4(mat_view_host, inv.m)) return GS_ERR_BAD_ARG;
    // the caller's (h, w) reaches the ray builder as (width, height): rows of `height = w` ... the depth map is then
    // read as (h, w) against rays of shape (w, h, 3); only square views broadcast, as in the reference
    if (h != w) return GS_ERR_BAD_ARG;
    double vp_w, vp_h;
    viewport(h, w, fov_y_rad, near, vp_w, vp_h);
    const int64_t n = (int64_t)h * w;
    GS_PDL_LAUNCH(depth_to_world_kernel, (unsigned)((n + 255) / 256), 256, 0, st, h, w, vp_w, vp_h, near, depth, inv,
                  pts_world);
    return 0;
}

extern "C" int gs_depth_to_disparity(int64_t n, const float *depth, float *disparity3, void *scratch4,
                                     void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n <= 0 || !depth || !disparity3 || !scratch4) return GS_ERR_BAD_ARG;
    unsigned *mx = (unsigned *)scratch4;
    GS_CUDA(cudaMemsetAsync(mx, 0, sizeof(unsigned), st));
    const unsigned grid = (unsigned)std::min<int64_t>((n + 255) / 256, GS_NUM_SMS * 8);
    disparity_max_kernel<<<grid, 256, 0, st>>>(n, depth, mx);
    GS_LAUNCH_CHECK();
    GS_PDL_LAUNCH(disparity_write_kernel, grid, 256, 0, st, n, depth, (const unsigned *)mx, disparity3);
    return 0;
}

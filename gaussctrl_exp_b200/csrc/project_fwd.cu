// Stage 1 — EWA projection forward (replaces gsplat's
// project_gaussians forward called at /root/reference/gaussctrl/gc_model.py:174-188).
//
// COMPILED WITH -fmad=false: radii / tiles_hit are integers derived from fp32
// arithmetic (ceil, float->int truncation) and must be bit-exact against the CPU
// oracle, so every multiply and add below rounds separately, in the written order.
//
// One Gaussian per thread. HBM-bound: 40 B in + 56 B out per Gaussian. The (N,3)
// inputs/outputs are AoS in the gsplat API; a warp's accesses cover one contiguous
// 384 B span so every fetched sector is fully used.
#include "gs_common.cuh"

namespace {

struct ProjConst {
    float V[12];
    float P[16];
};

__global__ void __launch_bounds__(256)
project_fwd_kernel(int N, const float *__restrict__ means, const float *__restrict__ scales, float glob,
                   const float4 *__restrict__ quats, const float *__restrict__ viewmat,
                   const float *__restrict__ projmat, float fx, float fy, float cx, float cy, int H, int W,
                   int tiles_x, int tiles_y, float clip, float *__restrict__ cov3d,
                   float2 *__restrict__ xys, float *__restrict__ depths, int32_t *__restrict__ radii,
                   float *__restrict__ conics, int32_t *__restrict__ tiles_hit) {
    __shared__ ProjConst k;
    if (threadIdx.x < 12) k.V[threadIdx.x] = viewmat[threadIdx.x];
    if (threadIdx.x >= 32 && threadIdx.x < 48) k.P[threadIdx.x - 32] = projmat[threadIdx.x - 32];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;

    const float p0 = means[3 * i], p1 = means[3 * i + 1], p2 = means[3 * i + 2];
    const float s0 = scales[3 * i], s1 = scales[3 * i + 1], s2 = scales[3 * i + 2];
    const float4 q = quats[i];
    const float *V = k.V, *P = k.P;

    float o_cov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float2 o_xy = make_float2(0.f, 0.f);
    float o_depth = 0.f, o_ca = 0.f, o_cb = 0.f, o_cc = 0.f;
    int o_radius = 0, o_hit = 0;

    // A.1.1 view transform + near clip
    const float t0 = V[0] * p0 + V[1] * p1 + V[2] * p2 + V[3];
    const float t1 = V[4] * p0 + V[5] * p1 + V[6] * p2 + V[7];
    const float tz = V[8] * p0 + V[9] * p1 + V[10] * p2 + V[11];
    if (tz > clip) {
        // A.1.2 cov3d = (R S)(R S)^T
        float w = q.x, x = q.y, y = q.z, z = q.w;
        const float nrm = sqrtf(w * w + x * x + y * y + z * z);
        const float inv = 1.0f / nrm;
        w = w * inv; x = x * inv; y = y * inv; z = z * inv;
        const float R0 = 1.0f - 2.0f * (y * y + z * z), R1 = 2.0f * (x * y - w * z), R2 = 2.0f * (x * z + w * y);
        const float R3 = 2.0f * (x * y + w * z), R4 = 1.0f - 2.0f * (x * x + z * z), R5 = 2.0f * (y * z - w * x);
        const float R6 = 2.0f * (x * z - w * y), R7 = 2.0f * (y * z + w * x), R8 = 1.0f - 2.0f * (x * x + y * y);
        const float sx = glob * s0, sy = glob * s1, sz = glob * s2;
        const float M0 = R0 * sx, M1 = R1 * sy, M2 = R2 * sz;
        const float M3 = R3 * sx, M4 = R4 * sy, M5 = R5 * sz;
        const float M6 = R6 * sx, M7 = R7 * sy, M8 = R8 * sz;
        const float S0 = M0 * M0 + M1 * M1 + M2 * M2;
        const float S1 = M0 * M3 + M1 * M4 + M2 * M5;
        const float S2 = M0 * M6 + M1 * M7 + M2 * M8;
        const float S3 = M3 * M3 + M4 * M4 + M5 * M5;
        const float S4 = M3 * M6 + M4 * M7 + M5 * M8;
        const float S5 = M6 * M6 + M7 * M7 + M8 * M8;
        o_cov[0] = S0; o_cov[1] = S1; o_cov[2] = S2; o_cov[3] = S3; o_cov[4] = S4; o_cov[5] = S5;

        // A.1.3 EWA
        const float limx = 1.3f * (0.5f * (float)W / fx), limy = 1.3f * (0.5f * (float)H / fy);
        const float rx = t0 / tz, ry = t1 / tz;
        const float tcx = tz * fminf(limx, fmaxf(-limx, rx));
        const float tcy = tz * fminf(limy, fmaxf(-limy, ry));
        const float rz = 1.0f / tz, rz2 = rz * rz;
        const float J00 = fx * rz, J02 = -fx * tcx * rz2;
        const float J11 = fy * rz, J12 = -fy * tcy * rz2;
        const float T0 = J00 * V[0] + J02 * V[8], T1 = J00 * V[1] + J02 * V[9], T2 = J00 * V[2] + J02 * V[10];
        const float T3 = J11 * V[4] + J12 * V[8], T4 = J11 * V[5] + J12 * V[9], T5 = J11 * V[6] + J12 * V[10];
        const float u0 = S0 * T0 + S1 * T1 + S2 * T2;
        const float u1 = S1 * T0 + S3 * T1 + S4 * T2;
        const float u2 = S2 * T0 + S4 * T1 + S5 * T2;
        const float w0 = S0 * T3 + S1 * T4 + S2 * T5;
        const float w1 = S1 * T3 + S3 * T4 + S4 * T5;
        const float w2 = S2 * T3 + S4 * T4 + S5 * T5;
        const float a = (T0 * u0 + T1 * u1 + T2 * u2) + GS_BLUR;
        const float b = T3 * u0 + T4 * u1 + T5 * u2;
        const float c = (T3 * w0 + T4 * w1 + T5 * w2) + GS_BLUR;

        // A.1.4 conic + radius
        const float det = a * c - b * b;
        if (det != 0.0f) {
            const float inv_det = 1.0f / det;
            const float mid = 0.5f * (a + c);
            const float disc = sqrtf(fmaxf(0.1f, mid * mid - det));
            const float v1 = mid + disc, v2 = mid - disc;
            const float radius = ceilf(3.0f * sqrtf(fmaxf(v1, v2)));
            // A.1.5 pixel centre
            const float hx = P[0] * p0 + P[1] * p1 + P[2] * p2 + P[3];
            const float hy = P[4] * p0 + P[5] * p1 + P[6] * p2 + P[7];
            const float hw = P[12] * p0 + P[13] * p1 + P[14] * p2 + P[15];
            const float rw = 1.0f / (hw + 1e-6f);
            const float px = 0.5f * (float)W * (hx * rw) + cx - 0.5f;
            const float py = 0.5f * (float)H * (hy * rw) + cy - 0.5f;
            // A.1.6 tile bbox
            int x0, y0, x1, y1;
            gs_tile_bbox(px, py, radius, tiles_x, tiles_y, x0, y0, x1, y1);
            const int area = (x1 - x0) * (y1 - y0);
            if (area > 0) {
                o_hit = area; o_depth = tz; o_radius = (int)radius;
                o_xy = make_float2(px, py);
                o_ca = c * inv_det; o_cb = -b * inv_det; o_cc = a * inv_det;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) cov3d[6 * i + j] = o_cov[j];
    xys[i] = o_xy;
    depths[i] = o_depth;
    radii[i] = o_radius;
    conics[3 * i] = o_ca; conics[3 * i + 1] = o_cb; conics[3 * i + 2] = o_cc;
    tiles_hit[i] = o_hit;
}

}  // namespace

extern "C" int gs_project_fwd(int N, const float *means, const float *scales, float glob_scale,
                              const float *quats, const float *viewmat, const float *projmat,
                              float fx, float fy, float cx, float cy, int H, int W,
                              int tiles_x, int tiles_y, float clip_thresh,
                              float *cov3d, float *xys, float *depths, int32_t *radii,
                              float *conics, int32_t *tiles_hit, void *stream) {
    if (N < 0 || H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!means || !scales || !quats || !viewmat || !projmat || !cov3d || !xys || !depths || !radii ||
        !conics || !tiles_hit)
        return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || (reinterpret_cast<uintptr_t>(xys) & 7u)) return GS_ERR_MISALIGNED;
    const int threads = 256, blocks = (N + threads - 1) / threads;
    project_fwd_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(
        N, means, scales, glob_scale, reinterpret_cast<const float4 *>(quats), viewmat, projmat, fx, fy,
        cx, cy, H, W, tiles_x, tiles_y, clip_thresh, cov3d, reinterpret_cast<float2 *>(xys), depths, radii,
        conics, tiles_hit);
    GS_LAUNCH_CHECK();
    return 0;
}

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
#include "project_math.cuh"

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

    ProjOut o;
    gs_project_one(k.V, k.P, means[3 * i], means[3 * i + 1], means[3 * i + 2], scales[3 * i], scales[3 * i + 1],
                   scales[3 * i + 2], glob, quats[i], fx, fy, cx, cy, H, W, tiles_x, tiles_y, clip, o);
#pragma unroll
    for (int j = 0; j < 6; ++j) cov3d[6 * i + j] = o.cov[j];
    xys[i] = o.xy;
    depths[i] = o.depth;
    radii[i] = o.radius;
    conics[3 * i] = o.ca; conics[3 * i + 1] = o.cb; conics[3 * i + 2] = o.cc;
    tiles_hit[i] = o.hit;
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

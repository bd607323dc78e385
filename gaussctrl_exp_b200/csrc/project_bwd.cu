// Stage 4 (tail) — projection backward: exact vjp of project_fwd.cu w.r.t. means,
// scales and quats (replaces gsplat's
// project_gaussians_backward reached from /root/reference/gaussctrl/gc_trainer.py:275).
// One Gaussian per thread; HBM-bound: 40 (inputs) + 4 (radii) + 24 (v_xy, v_depth,
// v_conic) in, 40 out per Gaussian. FMA contraction is allowed here (floats only).
#include "project_math.cuh"

namespace {

__global__ void __launch_bounds__(256)
project_bwd_kernel(int N, const float *__restrict__ means, const float *__restrict__ scales, float glob,
                   const float4 *__restrict__ quats, const float *__restrict__ viewmat,
                   const float *__restrict__ projmat, float fx, float fy, int H, int W,
                   const int32_t *__restrict__ radii, const float2 *__restrict__ v_xys,
                   const float *__restrict__ v_depths, const float *__restrict__ v_conics,
                   float *__restrict__ v_means, float *__restrict__ v_scales, float4 *__restrict__ v_quats) {
    __shared__ float V[12];
    __shared__ float P[16];
    if (threadIdx.x < 12) V[threadIdx.x] = viewmat[threadIdx.x];
    if (threadIdx.x >= 32 && threadIdx.x < 48) P[threadIdx.x - 32] = projmat[threadIdx.x - 32];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;

    ProjGrad g;
    g.vp[0] = g.vp[1] = g.vp[2] = 0.f;
    g.vs[0] = g.vs[1] = g.vs[2] = 0.f;
    g.vq = make_float4(0.f, 0.f, 0.f, 0.f);
    if (radii[i] > 0)
        gs_project_one_bwd(V, P, means[3 * i], means[3 * i + 1], means[3 * i + 2], scales[3 * i], scales[3 * i + 1],
                           scales[3 * i + 2], glob, quats[i], fx, fy, H, W, v_xys[i], v_depths[i], v_conics[3 * i],
                           v_conics[3 * i + 1], v_conics[3 * i + 2], g);
    v_means[3 * i] = g.vp[0]; v_means[3 * i + 1] = g.vp[1]; v_means[3 * i + 2] = g.vp[2];
    v_scales[3 * i] = g.vs[0]; v_scales[3 * i + 1] = g.vs[1]; v_scales[3 * i + 2] = g.vs[2];
    v_quats[i] = g.vq;
}

}  // namespace

extern "C" int gs_project_bwd(int N, const float *means, const float *scales, float glob_scale,
                              const float *quats, const float *viewmat, const float *projmat,
                              float fx, float fy, int H, int W, const int32_t *radii,
                              const float *v_xys, const float *v_depths, const float *v_conics,
                              float *v_means, float *v_scales, float *v_quats, void *stream) {
    if (N < 0 || H <= 0 || W <= 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!means || !scales || !quats || !viewmat || !projmat || !radii || !v_xys || !v_depths || !v_conics ||
        !v_means || !v_scales || !v_quats)
        return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(v_quats) || (reinterpret_cast<uintptr_t>(v_xys) & 7u))
        return GS_ERR_MISALIGNED;
    const int threads = 256, blocks = (N + threads - 1) / threads;
    project_bwd_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(
        N, means, scales, glob_scale, reinterpret_cast<const float4 *>(quats), viewmat, projmat, fx, fy, H, W,
        radii, reinterpret_cast<const float2 *>(v_xys), v_depths, v_conics, v_means, v_scales,
        reinterpret_cast<float4 *>(v_quats));
    GS_LAUNCH_CHECK();
    return 0;
}

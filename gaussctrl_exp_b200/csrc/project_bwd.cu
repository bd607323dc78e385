// Stage 4 (tail) — projection backward: exact vjp of project_fwd.cu w.r.t. means,
// scales and quats (replaces gsplat's
// project_gaussians_backward reached from /root/reference/gaussctrl/gc_trainer.py:275).
// One Gaussian per thread; HBM-bound: 40 (inputs) + 4 (radii) + 24 (v_xy, v_depth,
// v_conic) in, 40 out per Gaussian. FMA contraction is allowed here (floats only).
#include "gs_common.cuh"

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

    float vp0 = 0.f, vp1 = 0.f, vp2 = 0.f, vs0 = 0.f, vs1 = 0.f, vs2 = 0.f;
    float4 vq = make_float4(0.f, 0.f, 0.f, 0.f);
    if (radii[i] > 0) {
        const float p0 = means[3 * i], p1 = means[3 * i + 1], p2 = means[3 * i + 2];
        // ---- pixel centre ----
        const float hx = P[0] * p0 + P[1] * p1 + P[2] * p2 + P[3];
        const float hy = P[4] * p0 + P[5] * p1 + P[6] * p2 + P[7];
        const float hw = P[12] * p0 + P[13] * p1 + P[14] * p2 + P[15];
        const float rw = 1.0f / (hw + 1e-6f);
        const float2 vxy = v_xys[i];
        const float v_hx = 0.5f * (float)W * rw * vxy.x, v_hy = 0.5f * (float)H * rw * vxy.y;
        const float v_rw = 0.5f * (float)W * hx * vxy.x + 0.5f * (float)H * hy * vxy.y;
        const float v_hw = -rw * rw * v_rw;
        vp0 = P[0] * v_hx + P[4] * v_hy + P[12] * v_hw;
        vp1 = P[1] * v_hx + P[5] * v_hy + P[13] * v_hw;
        vp2 = P[2] * v_hx + P[6] * v_hy + P[14] * v_hw;

        // ---- recompute forward mid-state ----
        const float t0 = V[0] * p0 + V[1] * p1 + V[2] * p2 + V[3];
        const float t1 = V[4] * p0 + V[5] * p1 + V[6] * p2 + V[7];
        const float tz = V[8] * p0 + V[9] * p1 + V[10] * p2 + V[11];
        const float4 q = quats[i];
        const float inv = 1.0f / sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        const float w = q.x * inv, x = q.y * inv, y = q.z * inv, z = q.w * inv;
        const float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - w * z), 2.f * (x * z + w * y),
                            2.f * (x * y + w * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - w * x),
                            2.f * (x * z - w * y), 2.f * (y * z + w * x), 1.f - 2.f * (x * x + y * y)};
        const float s[3] = {glob * scales[3 * i], glob * scales[3 * i + 1], glob * scales[3 * i + 2]};
        float M[9], S[9];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) M[3 * r + c] = R[3 * r + c] * s[c];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                S[3 * r + c] = M[3 * r] * M[3 * c] + M[3 * r + 1] * M[3 * c + 1] + M[3 * r + 2] * M[3 * c + 2];
        const float rz = 1.0f / tz, rz2 = rz * rz, rz3 = rz2 * rz;
        const float limx = 1.3f * (0.5f * (float)W / fx), limy = 1.3f * (0.5f * (float)H / fy);
        // clamp decision exactly as the forward takes it (IEEE division)
        const float rx = __fdiv_rn(t0, tz), ry = __fdiv_rn(t1, tz);
        const int clx = rx > limx ? 1 : (rx < -limx ? -1 : 0);
        const int cly = ry > limy ? 1 : (ry < -limy ? -1 : 0);
        const float tcx = clx == 0 ? t0 : tz * ((float)clx * limx);
        const float tcy = cly == 0 ? t1 : tz * ((float)cly * limy);
        const float J00 = fx * rz, J02 = -fx * tcx * rz2, J11 = fy * rz, J12 = -fy * tcy * rz2;
        const float Tm[6] = {J00 * V[0] + J02 * V[8], J00 * V[1] + J02 * V[9], J00 * V[2] + J02 * V[10],
                             J11 * V[4] + J12 * V[8], J11 * V[5] + J12 * V[9], J11 * V[6] + J12 * V[10]};
        float ST0[3], ST1[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            ST0[r] = S[3 * r] * Tm[0] + S[3 * r + 1] * Tm[1] + S[3 * r + 2] * Tm[2];
            ST1[r] = S[3 * r] * Tm[3] + S[3 * r + 1] * Tm[4] + S[3 * r + 2] * Tm[5];
        }
        const float a = Tm[0] * ST0[0] + Tm[1] * ST0[1] + Tm[2] * ST0[2] + GS_BLUR;
        const float b = Tm[3] * ST0[0] + Tm[4] * ST0[1] + Tm[5] * ST0[2];
        const float c = Tm[3] * ST1[0] + Tm[4] * ST1[1] + Tm[5] * ST1[2] + GS_BLUR;
        const float det = a * c - b * b;
        const float id2 = 1.0f / (det * det);

        // ---- conic = (c,-b,a)/det ----
        const float vA = v_conics[3 * i], vB = v_conics[3 * i + 1], vC = v_conics[3 * i + 2];
        const float v_a = (-c * c * vA + b * c * vB - b * b * vC) * id2;
        const float v_b = (2.f * b * c * vA - (det + 2.f * b * b) * vB + 2.f * a * b * vC) * id2;
        const float v_c = (-b * b * vA + a * b * vB - a * a * vC) * id2;

        // ---- cov2d = Tm S Tm^T ----
        float vS[9], vT[6];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int k = 0; k < 3; ++k)
                vS[3 * r + k] = v_a * Tm[r] * Tm[k] + v_b * Tm[r] * Tm[3 + k] + v_c * Tm[3 + r] * Tm[3 + k];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            vT[k] = 2.f * v_a * ST0[k] + v_b * ST1[k];
            vT[3 + k] = 2.f * v_c * ST1[k] + v_b * ST0[k];
        }
        const float vJ00 = vT[0] * V[0] + vT[1] * V[1] + vT[2] * V[2];
        const float vJ02 = vT[0] * V[8] + vT[1] * V[9] + vT[2] * V[10];
        const float vJ11 = vT[3] * V[4] + vT[4] * V[5] + vT[5] * V[6];
        const float vJ12 = vT[3] * V[8] + vT[4] * V[9] + vT[5] * V[10];
        const float v_tcx = -fx * rz2 * vJ02, v_tcy = -fy * rz2 * vJ12;
        float v_t2 = -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.f * fx * tcx * rz3 * vJ02 + 2.f * fy * tcy * rz3 * vJ12;
        const float v_t0 = clx == 0 ? v_tcx : 0.f;
        const float v_t1 = cly == 0 ? v_tcy : 0.f;
        if (clx != 0) v_t2 += (float)clx * limx * v_tcx;
        if (cly != 0) v_t2 += (float)cly * limy * v_tcy;
        v_t2 += v_depths[i];
        vp0 += V[0] * v_t0 + V[4] * v_t1 + V[8] * v_t2;
        vp1 += V[1] * v_t0 + V[5] * v_t1 + V[9] * v_t2;
        vp2 += V[2] * v_t0 + V[6] * v_t1 + V[10] * v_t2;

        // ---- S = M M^T, M = R diag(s) ----
        float vR[9];
        float vsc[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float acc = 0.f;
#pragma unroll
                for (int j = 0; j < 3; ++j) acc += (vS[3 * r + j] + vS[3 * j + r]) * M[3 * j + k];
                vsc[k] += R[3 * r + k] * acc;
                vR[3 * r + k] = acc * s[k];
            }
        vs0 = vsc[0] * glob; vs1 = vsc[1] * glob; vs2 = vsc[2] * glob;
        // ---- R(qn), qn = q/|q| ----
        const float g0 = 2.f * (-z * vR[1] + y * vR[2] + z * vR[3] - x * vR[5] - y * vR[6] + x * vR[7]);
        const float g1 = 2.f * (y * vR[1] + z * vR[2] + y * vR[3] - 2.f * x * vR[4] - w * vR[5] + z * vR[6] + w * vR[7] - 2.f * x * vR[8]);
        const float g2 = 2.f * (-2.f * y * vR[0] + x * vR[1] + w * vR[2] + x * vR[3] + z * vR[5] - w * vR[6] + z * vR[7] - 2.f * y * vR[8]);
        const float g3 = 2.f * (-2.f * z * vR[0] - w * vR[1] + x * vR[2] + w * vR[3] - 2.f * z * vR[4] + y * vR[5] + x * vR[6] + y * vR[7]);
        const float dot = g0 * w + g1 * x + g2 * y + g3 * z;
        vq = make_float4((g0 - w * dot) * inv, (g1 - x * dot) * inv, (g2 - y * dot) * inv, (g3 - z * dot) * inv);
    }
    v_means[3 * i] = vp0; v_means[3 * i + 1] = vp1; v_means[3 * i + 2] = vp2;
    v_scales[3 * i] = vs0; v_scales[3 * i + 1] = vs1; v_scales[3 * i + 2] = vs2;
    v_quats[i] = vq;
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

// Fused per-view preprocess, backward: from the compositor's per-Gaussian gradients
// (v_xys, v_conics, v_rgbs, v_opac) straight to the six leaf gradients of
// /root/reference/gaussctrl/gc_model.py:157-170 in ONE pass — the chain the reference walks with
// ~20 autograd nodes after loss.backward() (gc_trainer.py:275): clamp / +0.5, SH backward,
// cat backward (split into features_dc / features_rest), sigmoid, exp, quaternion normalise,
// projection backward. 59 floats of gradient out per Gaussian (236 B), every byte written once.
//
// The (N,15,3) v_features_rest block of a 128-Gaussian CTA is built in shared memory (rows of
// 45 floats: odd pitch, conflict-free) and leaves with ONE TMA bulk store.
#include "project_math.cuh"
#include "sh_math.cuh"

namespace {

constexpr int PB_TPB = 128;
constexpr int PB_REST = 45;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(PB_TPB)
preprocess_bwd_kernel(int N, const float *__restrict__ means, const float *__restrict__ log_scales,
                      const float4 *__restrict__ quats, const float *__restrict__ opac_logits, int use_sh, int deg,
                      const __grid_constant__ GsCamera cam, const int32_t *__restrict__ radii,
                      const float *__restrict__ rgbs, const float2 *__restrict__ v_xys,
                      const float *__restrict__ v_conics, const float *__restrict__ v_rgbs,
                      const float *__restrict__ v_opac, float *__restrict__ v_means,
                      float *__restrict__ v_log_scales, float4 *__restrict__ v_quats,
                      float *__restrict__ v_opac_logits, float *__restrict__ v_dc, float *__restrict__ v_rest) {
    __shared__ __align__(128) float rest_s[PB_TPB * PB_REST];
    const int tid = threadIdx.x;
    const int g0 = blockIdx.x * PB_TPB;
    const int rows = min(PB_TPB, N - g0);
    const int i = g0 + tid;

    if (i < N) {
        float vm[3] = {0.f, 0.f, 0.f}, vls[3] = {0.f, 0.f, 0.f};
        float4 vq = make_float4(0.f, 0.f, 0.f, 0.f);
        float vol = 0.f, vd[3] = {0.f, 0.f, 0.f};
        float *row = rest_s + tid * PB_REST;
        const bool vis = radii[i] > 0;
        float Y[16];
        int nb = 0;
        float vc0 = 0.f, vc1 = 0.f, vc2 = 0.f;
        if (vis) {
            const float p0 = means[3 * i], p1 = means[3 * i + 1], p2 = means[3 * i + 2];
            const float s0 = expf(log_scales[3 * i]), s1 = expf(log_scales[3 * i + 1]), s2 = expf(log_scales[3 * i + 2]);
            const float4 qr = quats[i];
            const float qn = sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w);
            const float iq = 1.0f / qn;
            const float4 q = make_float4(qr.x * iq, qr.y * iq, qr.z * iq, qr.w * iq);
            ProjGrad g;
            gs_project_one_bwd(cam.viewmat, cam.projmat, p0, p1, p2, s0, s1, s2, cam.glob_scale, q, cam.fx, cam.fy,
                               cam.H, cam.W, v_xys[i], 0.0f, v_conics[3 * i], v_conics[3 * i + 1], v_conics[3 * i + 2], g);
            vm[0] = g.vp[0]; vm[1] = g.vp[1]; vm[2] = g.vp[2];
            vls[0] = g.vs[0] * s0; vls[1] = g.vs[1] * s1; vls[2] = g.vs[2] * s2;          // d exp
            const float dq = g.vq.x * q.x + g.vq.y * q.y + g.vq.z * q.z + g.vq.w * q.w;    // d (q/|q|)
            vq = make_float4((g.vq.x - q.x * dq) * iq, (g.vq.y - q.y * dq) * iq, (g.vq.z - q.z * dq) * iq,
                             (g.vq.w - q.w * dq) * iq);
            const float o = 1.0f / (1.0f + expf(-opac_logits[i]));
            vol = v_opac[i] * o * (1.0f - o);                                              // d sigmoid
            const float c0 = rgbs[3 * i], c1 = rgbs[3 * i + 1], c2 = rgbs[3 * i + 2];
            const float u0 = v_rgbs[3 * i], u1 = v_rgbs[3 * i + 1], u2 = v_rgbs[3 * i + 2];
            if (use_sh) {
                vc0 = c0 > 0.f ? u0 : 0.f; vc1 = c1 > 0.f ? u1 : 0.f; vc2 = c2 > 0.f ? u2 : 0.f;   // d clamp(min=0)
                float vx = p0 - cam.cam_pos[0], vy = p1 - cam.cam_pos[1], vz = p2 - cam.cam_pos[2];
                const float vn = sqrtf(vx * vx + vy * vy + vz * vz);
                vx = vx / vn; vy = vy / vn; vz = vz / vn;
                sh_basis<3>(deg, vx, vy, vz, Y);
                nb = sh_num_bases(deg);
                vd[0] = Y[0] * vc0; vd[1] = Y[0] * vc1; vd[2] = Y[0] * vc2;
            } else {
                vd[0] = u0 * c0 * (1.0f - c0); vd[1] = u1 * c1 * (1.0f - c1); vd[2] = u2 * c2 * (1.0f - c2);
            }
        }
#pragma unroll
        for (int k = 1; k < 16; ++k) {
            const float yk = (vis && k < nb) ? Y[k] : 0.f;
            row[3 * (k - 1)] = yk * vc0; row[3 * (k - 1) + 1] = yk * vc1; row[3 * (k - 1) + 2] = yk * vc2;
        }
        v_means[3 * i] = vm[0]; v_means[3 * i + 1] = vm[1]; v_means[3 * i + 2] = vm[2];
        v_log_scales[3 * i] = vls[0]; v_log_scales[3 * i + 1] = vls[1]; v_log_scales[3 * i + 2] = vls[2];
        v_quats[i] = vq;
        v_opac_logits[i] = vol;
        v_dc[3 * i] = vd[0]; v_dc[3 * i + 1] = vd[1]; v_dc[3 * i + 2] = vd[2];
    }
    // make the generic-proxy smem writes visible to the async proxy, then one bulk store
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int rows4 = rows & ~3;
    float *dst = v_rest + (size_t)g0 * PB_REST;
    if (tid == 0 && rows4 > 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(rest_s)),
                     "r"((unsigned)rows4 * PB_REST * 4u)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    for (int f = rows4 * PB_REST + tid; f < rows * PB_REST; f += PB_TPB) dst[f] = rest_s[f];
    if (tid == 0 && rows4 > 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

}  // namespace

extern "C" int gs_preprocess_bwd(int N, const float *means, const float *log_scales, const float *quats,
                                 const float *opac_logits, int use_sh, int sh_deg, const GsCamera *cam_host,
                                 const int32_t *radii, const float *rgbs, const float *v_xys, const float *v_conics,
                                 const float *v_rgbs, const float *v_opac, float *v_means, float *v_log_scales,
                                 float *v_quats, float *v_opac_logits, float *v_feat_dc, float *v_feat_rest,
                                 void *stream) {
    if (N < 0 || !cam_host) return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (N == 0) return 0;
    if (!means || !log_scales || !quats || !opac_logits || !radii || !rgbs || !v_xys || !v_conics || !v_rgbs ||
        !v_opac || !v_means || !v_log_scales || !v_quats || !v_opac_logits || !v_feat_dc || !v_feat_rest)
        return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(v_quats) || !gs_aligned16(v_feat_rest) ||
        (reinterpret_cast<uintptr_t>(v_xys) & 7u))
        return GS_ERR_MISALIGNED;
    const int blocks = (N + PB_TPB - 1) / PB_TPB;
    preprocess_bwd_kernel<<<blocks, PB_TPB, 0, (cudaStream_t)stream>>>(
        N, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg, *cam_host, radii,
        rgbs, reinterpret_cast<const float2 *>(v_xys), v_conics, v_rgbs, v_opac, v_means, v_log_scales,
        reinterpret_cast<float4 *>(v_quats), v_opac_logits, v_feat_dc, v_feat_rest);
    GS_LAUNCH_CHECK();
    return 0;
}

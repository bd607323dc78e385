// Fused per-view preprocess, backward: from the compositor's per-Gaussian gradients
// (v_xys, v_conics, v_rgbs, v_opac) straight to the six leaf gradients of
// /root/reference/gaussctrl/gc_model.py:157-170 in ONE pass — the chain the reference walks with
// ~20 autograd nodes after loss.backward() (gc_trainer.py:275): clamp / +0.5, SH backward,
// cat backward (split into features_dc / features_rest), sigmoid, exp, quaternion normalise,
// projection backward. 59 floats of gradient out per Gaussian (236 B), every byte written once.
//
// The (N,15,3) v_features_rest block of a 128-Gaussian CTA is built in shared memory (rows of
// 45 floats: odd pitch, conflict-free) and leaves with ONE TMA bulk store.
#define GS_PDL_GROUP 3
#include "project_math.cuh"
#include "sh_math.cuh"

namespace {

#ifndef GS_PB_TPB
#define GS_PB_TPB 128
#endif
constexpr int PB_TPB = GS_PB_TPB;
constexpr int PB_REST = 45;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(PB_TPB)
preprocess_bwd_kernel(int N, const float *__restrict__ means, const float *__restrict__ log_scales,
                      const float4 *__restrict__ quats, const float *__restrict__ opac_logits, int use_sh, int deg,
                      const __grid_constant__ GsCamera cam, const int32_t *__restrict__ radii,
                      const float *__restrict__ rgbs, int rgb_stride, const float2 *__restrict__ v_xys,
                      const float *__restrict__ v_conics, const float *__restrict__ v_rgbs,
                      const float *__restrict__ v_opac, float *__restrict__ v_means,
                      float *__restrict__ v_log_scales, float4 *__restrict__ v_quats,
                      float *__restrict__ v_opac_logits, float *__restrict__ v_dc, float *__restrict__ v_rest) {
    __shared__ __align__(128) float rest_s[PB_TPB * PB_REST];
    const int tid = threadIdx.x;
    const int g0 = blockIdx.x * PB_TPB;
    const int rows = min(PB_TPB, N - g0);
    const int i = g0 + tid;
    gs_pdl_sync();

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
            // the forward's correctly rounded exp(scales) (preprocess_fwd.cu)
            const float s0 = (float)exp((double)log_scales[3 * i]), s1 = (float)exp((double)log_scales[3 * i + 1]),
                        s2 = (float)exp((double)log_scales[3 * i + 2]);
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
            const float *crow = rgbs + (size_t)i * rgb_stride;   // (N,3) array, or the colour quad of the packed records
            const float c0 = crow[0], c1 = crow[1], c2 = crow[2];
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

// rgbs: post-activation colours, row i at rgbs + i * rgb_stride floats (3 = an (N,3) array; 12 with rgbs pointing at
// float 8 of the packed records)
int gs_launch_preprocess_bwd(int N, const float *means, const float *log_scales, const float *quats,
                             const float *opac_logits, int use_sh, int sh_deg, const GsCamera *cam_host,
                             const int32_t *radii, const float *rgbs, int rgb_stride, const float *v_xys,
                             const float *v_conics, const float *v_rgbs, const float *v_opac, float *v_means,
                             float *v_log_scales, float *v_quats, float *v_opac_logits, float *v_feat_dc,
                             float *v_feat_rest, cudaStream_t st) {
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
    GS_PDL_LAUNCH(preprocess_bwd_kernel, blocks, PB_TPB, 0, st, N, means, log_scales,
                  reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg, *cam_host, radii, rgbs, rgb_stride,
                  reinterpret_cast<const float2 *>(v_xys), v_conics, v_rgbs, v_opac, v_means, v_log_scales,
                  reinterpret_cast<float4 *>(v_quats), v_opac_logits, v_feat_dc, v_feat_rest);
    return 0;
}

extern "C" int gs_preprocess_bwd(int N, const float *means, const float *log_scales, const float *quats,
                                 const float *opac_logits, int use_sh, int sh_deg, const GsCamera *cam_host,
                                 const int32_t *radii, const float *rgbs, const float *v_xys, const float *v_conics,
                                 const float *v_rgbs, const float *v_opac, float *v_means, float *v_log_scales,
                                 float *v_quats, float *v_opac_logits, float *v_feat_dc, float *v_feat_rest,
                                 void *stream) {
    return gs_launch_preprocess_bwd(N, means, log_scales, quats, opac_logits, use_sh, sh_deg, cam_host, radii, rgbs, 3,
                                    v_xys, v_conics, v_rgbs, v_opac, v_means, v_log_scales, v_quats, v_opac_logits,
                                    v_feat_dc, v_feat_rest, (cudaStream_t)stream);
}

// ==========================================================================================
// Multi-view variant for the data-parallel fine-tune step.
// Every rank holds all parameters, so instead of all-reducing the 59-float leaf gradients
// (236 B/Gaussian) the ranks all-GATHER the 9 compositor-level gradients of their view
// (v_xy, v_conic, v_rgb [clamp already applied], v_opacity = 36 B/Gaussian, plus the camera)
// and each rank replays this kernel over all views: 6.5x fewer NVLink bytes at 2 GPUs, 1.6x
// at 8, and the leaf gradients are produced already summed.
// Block layout of one view in `views` (stride floats apart), Np = N rounded up to 4:
//   [0,2Np) v_xys | [2Np,5Np) v_conics | [5Np,8Np) v_rgbs | [8Np,9Np) v_opac | [9Np, 9Np+64) GsCamera
// A Gaussian whose 9 values are all zero in a view is skipped for that view (culled or unseen).
// ==========================================================================================
namespace {

constexpr int PB_MAX_VIEWS = 16;

struct ViewGrad {   // the 9 compositor-level gradients of one Gaussian in one view
    float2 xy;
    float a, b, c, u0, u1, u2, o;
};

// `views` != nullptr: V blocks `stride` floats apart in local memory (all-gathered by NCCL).
// `peer_ptrs` != nullptr: V base pointers (one per rank, symmetric-memory mapping of the PEERS' buffers) +
// `peer_off` floats: the blocks are read in place over NVLink — compute and collective in one kernel, no
// gathered staging copy. Loads of view v+1 are issued before view v is processed (NVLink latency ~2 us).
__global__ void __launch_bounds__(PB_TPB)
preprocess_bwd_multi_kernel(int N, int Np, int V, const float *__restrict__ views, size_t stride,
                            const unsigned long long *__restrict__ peer_ptrs, size_t peer_off, float scale,
                            const float *__restrict__ means, const float *__restrict__ log_scales,
                            const float4 *__restrict__ quats, const float *__restrict__ opac_logits, int use_sh,
                            int deg, float *__restrict__ v_means, float *__restrict__ v_log_scales,
                            float4 *__restrict__ v_quats, float *__restrict__ v_opac_logits,
                            float *__restrict__ v_dc, float *__restrict__ v_rest) {
    __shared__ __align__(128) float rest_s[PB_TPB * PB_REST];
    __shared__ GsCamera cam_s[PB_MAX_VIEWS];
    __shared__ const float *blk_s[PB_MAX_VIEWS];
    const int tid = threadIdx.x;
    const int g0 = blockIdx.x * PB_TPB;
    const int rows = min(PB_TPB, N - g0);
    const int i = g0 + tid;
    float *row = rest_s + tid * PB_REST;
#pragma unroll
    for (int k = 0; k < PB_REST; ++k) row[k] = 0.f;
    if (tid < V)
        blk_s[tid] = peer_ptrs ? reinterpret_cast<const float *>(peer_ptrs[tid]) + peer_off : views + (size_t)tid * stride;
    __syncthreads();
    constexpr int CAMF = (int)(sizeof(GsCamera) / 4);
    for (int k = tid; k < V * CAMF; k += PB_TPB)      // all cameras up front: one remote round trip, no per-view barrier
        reinterpret_cast<float *>(cam_s)[k] = blk_s[k / CAMF][(size_t)9 * Np + k % CAMF];

    float vm[3] = {0.f, 0.f, 0.f}, vls[3] = {0.f, 0.f, 0.f}, vd[3] = {0.f, 0.f, 0.f};
    float4 vq = make_float4(0.f, 0.f, 0.f, 0.f);
    float vop = 0.f;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f, s0 = 1.f, s1 = 1.f, s2 = 1.f, iq = 1.f, o = 0.f;
    float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
    if (i < N) {
        p0 = means[3 * i]; p1 = means[3 * i + 1]; p2 = means[3 * i + 2];
        s0 = gs_exp_scale(log_scales[3 * i]); s1 = gs_exp_scale(log_scales[3 * i + 1]); s2 = gs_exp_scale(log_scales[3 * i + 2]);
        const float4 qr = quats[i];
        iq = 1.0f / sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w);
        q = make_float4(qr.x * iq, qr.y * iq, qr.z * iq, qr.w * iq);
        o = 1.0f / (1.0f + expf(-opac_logits[i]));
    }
    auto load_view = [&](int v) {
        ViewGrad g;
        const float *blk = blk_s[v];
        g.xy = reinterpret_cast<const float2 *>(blk)[i];
        const float *vc = blk + (size_t)2 * Np + 3 * (size_t)i;
        const float *vr = blk + (size_t)5 * Np + 3 * (size_t)i;
        g.a = vc[0]; g.b = vc[1]; g.c = vc[2];
        g.u0 = vr[0]; g.u1 = vr[1]; g.u2 = vr[2];
        g.o = blk[(size_t)8 * Np + i];
        return g;
    };
    __syncthreads();   // cameras staged
    if (i < N) {
        ViewGrad cur = load_view(0);
        for (int v = 0; v < V; ++v) {
            ViewGrad nxt = cur;
            if (v + 1 < V) nxt = load_view(v + 1);
            const ViewGrad w = cur;
            cur = nxt;
            if (w.xy.x == 0.f && w.xy.y == 0.f && w.a == 0.f && w.b == 0.f && w.c == 0.f && w.u0 == 0.f && w.u1 == 0.f &&
                w.u2 == 0.f && w.o == 0.f)
                continue;
            const GsCamera &cam = cam_s[v];
            ProjGrad g;
            gs_project_one_bwd(cam.viewmat, cam.projmat, p0, p1, p2, s0, s1, s2, cam.glob_scale, q, cam.fx, cam.fy,
                               cam.H, cam.W, w.xy, 0.0f, w.a, w.b, w.c, g);
            vm[0] += g.vp[0]; vm[1] += g.vp[1]; vm[2] += g.vp[2];
            vls[0] += g.vs[0]; vls[1] += g.vs[1]; vls[2] += g.vs[2];
            vq.x += g.vq.x; vq.y += g.vq.y; vq.z += g.vq.z; vq.w += g.vq.w;
            vop += w.o;
            if (use_sh) {
                float dx = p0 - cam.cam_pos[0], dy = p1 - cam.cam_pos[1], dz = p2 - cam.cam_pos[2];
                const float vn = sqrtf(dx * dx + dy * dy + dz * dz);
                dx = dx / vn; dy = dy / vn; dz = dz / vn;
                float Y[16];
                sh_basis<3>(deg, dx, dy, dz, Y);
                const int nb = sh_num_bases(deg);
                vd[0] += Y[0] * w.u0; vd[1] += Y[0] * w.u1; vd[2] += Y[0] * w.u2;
#pragma unroll
                for (int k = 1; k < 16; ++k)
                    if (k < nb) {
                        row[3 * (k - 1)] += Y[k] * w.u0; row[3 * (k - 1) + 1] += Y[k] * w.u1;
                        row[3 * (k - 1) + 2] += Y[k] * w.u2;
                    }
            } else {
                vd[0] += w.u0; vd[1] += w.u1; vd[2] += w.u2;     // sender already applied d sigmoid
            }
        }
        v_means[3 * i] = vm[0] * scale; v_means[3 * i + 1] = vm[1] * scale; v_means[3 * i + 2] = vm[2] * scale;
        v_log_scales[3 * i] = vls[0] * s0 * scale; v_log_scales[3 * i + 1] = vls[1] * s1 * scale;
        v_log_scales[3 * i + 2] = vls[2] * s2 * scale;
        const float dq = vq.x * q.x + vq.y * q.y + vq.z * q.z + vq.w * q.w;
        v_quats[i] = make_float4((vq.x - q.x * dq) * iq * scale, (vq.y - q.y * dq) * iq * scale,
                                 (vq.z - q.z * dq) * iq * scale, (vq.w - q.w * dq) * iq * scale);
        v_opac_logits[i] = vop * o * (1.0f - o) * scale;
        v_dc[3 * i] = vd[0] * scale; v_dc[3 * i + 1] = vd[1] * scale; v_dc[3 * i + 2] = vd[2] * scale;
#pragma unroll
        for (int k = 0; k < PB_REST; ++k) row[k] *= scale;
    }
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

// sender side: apply the colour activation's derivative to v_rgbs in place and append the camera
__global__ void __launch_bounds__(256)
finish_view_block_kernel(int N, int use_sh, const float *__restrict__ rgbs, int rgb_stride, float *__restrict__ v_rgbs,
                         const __grid_constant__ GsCamera cam, float *__restrict__ cam_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    gs_pdl_sync();
    if (blockIdx.x == 0 && threadIdx.x < (int)(sizeof(GsCamera) / 4))
        cam_out[threadIdx.x] = reinterpret_cast<const float *>(&cam)[threadIdx.x];
    if (i >= 3 * N) return;
    const float c = rgbs[(size_t)(i / 3) * rgb_stride + (i % 3)];
    v_rgbs[i] = use_sh ? (c > 0.f ? v_rgbs[i] : 0.f) : v_rgbs[i] * c * (1.0f - c);
}

}  // namespace

extern "C" int gs_view_block_floats(int N) { return 9 * ((N + 3) / 4 * 4) + 64; }

int gs_launch_finish_view_block(int N, int use_sh, const GsCamera *cam_host, const float *rgbs, int rgb_stride,
                                float *block, cudaStream_t st) {
    if (N < 0 || !cam_host || !block || (N > 0 && !rgbs)) return GS_ERR_BAD_ARG;
    static_assert(sizeof(GsCamera) <= 64 * 4, "camera must fit the 64-float tail of a view block");
    const int Np = (N + 3) / 4 * 4;
    const int blocks = (3 * N + 255) / 256;
    GS_PDL_LAUNCH(finish_view_block_kernel, blocks > 0 ? blocks : 1, 256, 0, st, N, use_sh, rgbs, rgb_stride,
                  block + (size_t)5 * Np, *cam_host, block + (size_t)9 * Np);
    return 0;
}

extern "C" int gs_finish_view_block(int N, int use_sh, const GsCamera *cam_host, const float *rgbs, float *block,
                                    void *stream) {
    return gs_launch_finish_view_block(N, use_sh, cam_host, rgbs, 3, block, (cudaStream_t)stream);
}

static int launch_bwd_multi(int N, int V, const float *views, int64_t stride_floats,
                            const unsigned long long *peer_ptrs, int64_t peer_off, float scale, const float *means,
                            const float *log_scales, const float *quats, const float *opac_logits, int use_sh,
                            int sh_deg, float *v_means, float *v_log_scales, float *v_quats, float *v_opac_logits,
                            float *v_feat_dc, float *v_feat_rest, void *stream) {
    if (N < 0 || V <= 0 || V > PB_MAX_VIEWS || (!views && !peer_ptrs)) return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (N == 0) return 0;
    if (!means || !log_scales || !quats || !opac_logits || !v_means || !v_log_scales || !v_quats || !v_opac_logits ||
        !v_feat_dc || !v_feat_rest)
        return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(v_quats) || !gs_aligned16(v_feat_rest) || (views && !gs_aligned16(views)) ||
        (stride_floats & 3) || (peer_off & 3))
        return GS_ERR_MISALIGNED;
    const int Np = (N + 3) / 4 * 4;
    if (views && stride_floats < 9 * (int64_t)Np + 64) return GS_ERR_BAD_ARG;
    const int blocks = (N + PB_TPB - 1) / PB_TPB;
    preprocess_bwd_multi_kernel<<<blocks, PB_TPB, 0, (cudaStream_t)stream>>>(
        N, Np, V, views, (size_t)stride_floats, peer_ptrs, (size_t)peer_off, scale, means, log_scales,
        reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg, v_means, v_log_scales,
        reinterpret_cast<float4 *>(v_quats), v_opac_logits, v_feat_dc, v_feat_rest);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_preprocess_bwd_multi(int N, int V, const float *views, int64_t stride_floats, float scale,
                                       const float *means, const float *log_scales, const float *quats,
                                       const float *opac_logits, int use_sh, int sh_deg, float *v_means,
                                       float *v_log_scales, float *v_quats, float *v_opac_logits, float *v_feat_dc,
                                       float *v_feat_rest, void *stream) {
    return launch_bwd_multi(N, V, views, stride_floats, nullptr, 0, scale, means, log_scales, quats, opac_logits, use_sh,
                            sh_deg, v_means, v_log_scales, v_quats, v_opac_logits, v_feat_dc, v_feat_rest, stream);
}

extern "C" int gs_preprocess_bwd_multi_peer(int N, int V, const uint64_t *peer_base_ptrs_dev, int64_t offset_floats,
                                            float scale, const float *means, const float *log_scales,
                                            const float *quats, const float *opac_logits, int use_sh, int sh_deg,
                                            float *v_means, float *v_log_scales, float *v_quats,
                                            float *v_opac_logits, float *v_feat_dc, float *v_feat_rest, void *stream) {
    return launch_bwd_multi(N, V, nullptr, 0, reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev),
                            offset_floats, scale, means, log_scales, quats, opac_logits, use_sh, sh_deg, v_means,
                            v_log_scales, v_quats, v_opac_logits, v_feat_dc, v_feat_rest, stream);
}

// Fused per-view preprocess, forward: everything the reference's host code does per Gaussian
// between the leaf parameters and rasterize_gaussians, in ONE pass over HBM
// (/root/reference/gaussctrl/gc_model.py:172-203 + :215):
//   cat(features_dc, features_rest) | exp(scales) | quats/|quats| | project_gaussians |
//   normalise(means - cam_pos) | spherical_harmonics | clamp(+0.5, min 0) | sigmoid(opacities)
// The unfused path moves ~1.1 GB per 1 M Gaussians for this (the cat alone re-writes the 192 MB
// SH block); fused it is the compulsory 236 B in + 48 B out per Gaussian.
//
// The (N,15,3) features_rest block of a 128-Gaussian CTA is one dense 23 040 B span: it is
// fetched by ONE TMA bulk copy (cp.async.bulk, completion on an mbarrier) that overlaps the
// projection arithmetic; rows are 45 floats (odd) so per-thread row reads are bank-conflict free.
//
// COMPILED WITH -fmad=false (shares the bit-exact projection math with project_fwd.cu).
#define GS_PDL_GROUP 3
#include "project_math.cuh"
#include "sh_math.cuh"
#include "raster_packed.cuh"

namespace {

#ifndef GS_PP_TPB
#define GS_PP_TPB 128
#endif
constexpr int PP_TPB = GS_PP_TPB;
constexpr int PP_REST = 45;  // floats of features_rest per Gaussian (15 bases x 3 channels)

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// PART 0: everything. PART 1: geometry only (projection + opacity; the binning chain depends on nothing else).
// PART 2: colour only (reads the radii PART 1 wrote) — the host runs it on a second stream underneath the
// latency-bound binning kernels, which takes the 216 B/Gaussian SH traffic off the critical path.
// PACKED (PART 0 only): instead of the operator-shaped arrays (xys, conics, rgbs, opac) the kernel writes ONE 48-byte
// splat record per Gaussian (raster_packed.cuh: pixel centre + alpha extents | opacity + pre-scaled conic | colour +
// depth) — what the packed compositing kernels copy with cp.async — plus depths / radii / tiles_hit for the binning.
// subset != nullptr: element i is Gaussian subset[i] (the eval crop box of gc_model.py:108-113, :157-170, as an index
// list instead of six gathered copies); its SH rows are then fetched per thread, not by the bulk copy.
template <int PART, bool PACKED>
__global__ void __launch_bounds__(PP_TPB)
preprocess_fwd_kernel(int N, const float *__restrict__ means, const float *__restrict__ log_scales,
                      const float4 *__restrict__ quats, const float *__restrict__ opac_logits,
                      const float *__restrict__ feat_dc, const float *__restrict__ feat_rest, int use_sh, int deg,
                      const __grid_constant__ GsCamera cam, float2 *__restrict__ xys, float *__restrict__ depths,
                      int32_t *__restrict__ radii, float *__restrict__ conics, int32_t *__restrict__ tiles_hit,
                      float *__restrict__ rgbs, float *__restrict__ opac, const int32_t *__restrict__ subset,
                      float4 *__restrict__ rec) {
    gs_pdl_sync();   // lets the first binning kernel become resident while this one drains
    __shared__ __align__(128) float rest_s[PP_TPB * PP_REST];
    __shared__ __align__(8) unsigned long long mbar;
    const int tid = threadIdx.x;
    const int g0 = blockIdx.x * PP_TPB;
    const int rows = min(PP_TPB, N - g0);
    const bool need_rest = PART != 1 && use_sh && deg >= 1;
    const int src = (g0 + tid < N) ? (subset ? subset[g0 + tid] : g0 + tid) : 0;

    if (need_rest && subset) {
        if (g0 + tid < N)
            for (int f = 0; f < PP_REST; ++f) rest_s[tid * PP_REST + f] = feat_rest[(size_t)src * PP_REST + f];
    } else if (need_rest) {
        const int rows4 = rows & ~3;                      // 4 rows = 720 B: keeps the bulk size a multiple of 16
        const unsigned bytes = (unsigned)rows4 * PP_REST * 4u;
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&mbar)), "r"(bytes)
                         : "memory");
            if (bytes)
                asm volatile(
                    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                        smem_u32(rest_s)),
                    "l"(feat_rest + (size_t)g0 * PP_REST), "r"(bytes), "r"(smem_u32(&mbar))
                    : "memory");
        }
        for (int f = rows4 * PP_REST + tid; f < rows * PP_REST; f += PP_TPB)
            rest_s[f] = feat_rest[(size_t)g0 * PP_REST + f];
    }

    const int i = g0 + tid;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f;
    ProjOut o;
    o.radius = 0;
    float op = 0.f;
    if (PART == 2) {
        if (i < N) {
            p0 = means[3 * src]; p1 = means[3 * src + 1]; p2 = means[3 * src + 2];
            o.radius = radii[i];
        }
    } else if (i < N) {
        p0 = means[3 * src]; p1 = means[3 * src + 1]; p2 = means[3 * src + 2];
        // exp(scales) (gc_model.py:176) correctly rounded (evaluated in double, rounded once): the integer outputs
        // radius / num_tiles_hit depend on it through ceil(), and a 1-ulp difference between two libm's expf flips them
        // for ~1e-6 of the Gaussians — visible at 1 M. With a correctly rounded exp the fused path's integers are
        // reproducible on any platform (the CPU oracle does the same); 3 DP exp per Gaussian cost nothing here.
        const float s0 = (float)exp((double)log_scales[3 * src]), s1 = (float)exp((double)log_scales[3 * src + 1]),
                    s2 = (float)exp((double)log_scales[3 * src + 2]);
        float4 q = quats[src];
        const float qn = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);   // quats / quats.norm()
        q.x = q.x / qn; q.y = q.y / qn; q.z = q.z / qn; q.w = q.w / qn;
        gs_project_one(cam.viewmat, cam.projmat, p0, p1, p2, s0, s1, s2, cam.glob_scale, q, cam.fx, cam.fy, cam.cx,
                       cam.cy, cam.H, cam.W, cam.tiles_x, cam.tiles_y, cam.clip, o);
        op = 1.0f / (1.0f + expf(-opac_logits[src]));
        depths[i] = o.depth;
        radii[i] = o.radius;
        tiles_hit[i] = o.hit;
        if (!PACKED) {
            xys[i] = o.xy;
            conics[3 * i] = o.ca; conics[3 * i + 1] = o.cb; conics[3 * i + 2] = o.cc;
            opac[i] = op;
        }
    }

    if (need_rest && subset) {
        __syncthreads();
    } else if (need_rest) {
        __syncthreads();  // mbarrier init + remainder rows visible
        unsigned done = 0;
        while (!done)
            asm volatile(
                "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}"
                : "=r"(done)
                : "r"(smem_u32(&mbar))
                : "memory");
    }
    if (!PACKED && (i >= N || PART == 1)) return;

    float r = 0.f, g = 0.f, b = 0.f;
    if (i < N && o.radius > 0) {      // culled Gaussians never reach the compositor: colour left at 0
        const float d0 = feat_dc[3 * src], d1 = feat_dc[3 * src + 1], d2 = feat_dc[3 * src + 2];
        if (use_sh) {
            // viewdirs = normalise(means - cam_pos)  (gc_model.py:197-198); the SH routine normalises again
            float vx = p0 - cam.cam_pos[0], vy = p1 - cam.cam_pos[1], vz = p2 - cam.cam_pos[2];
            const float vn = sqrtf(vx * vx + vy * vy + vz * vz);
            vx = vx / vn; vy = vy / vn; vz = vz / vn;
            float Y[16];
            sh_basis<3>(deg, vx, vy, vz, Y);
            r = Y[0] * d0; g = Y[0] * d1; b = Y[0] * d2;
            const int nb = sh_num_bases(deg);
            const float *row = rest_s + tid * PP_REST;
#pragma unroll
            for (int k = 1; k < 16; ++k)
                if (k < nb) {
                    r += Y[k] * row[3 * (k - 1)];
                    g += Y[k] * row[3 * (k - 1) + 1];
                    b += Y[k] * row[3 * (k - 1) + 2];
                }
            r = fmaxf(r + 0.5f, 0.0f); g = fmaxf(g + 0.5f, 0.0f); b = fmaxf(b + 0.5f, 0.0f);   // :201
        } else {
            r = 1.0f / (1.0f + expf(-d0)); g = 1.0f / (1.0f + expf(-d1)); b = 1.0f / (1.0f + expf(-d2));   // :203
        }
    }
    if (!PACKED) {
        rgbs[3 * i] = r; rgbs[3 * i + 1] = g; rgbs[3 * i + 2] = b;
        return;
    }
    // PACKED: the CTA's 128 records are one dense 6 144-byte span; they are assembled in shared memory (the SH stage,
    // free once every thread has consumed its row) and leave with fully coalesced 16-byte stores
    float4 q0 = make_float4(0.f, 0.f, -1.f, -1.f), q1 = make_float4(0.f, 0.f, 0.f, 0.f), q2 = q1;   // culled: zero row
    if (i < N && o.radius > 0) gs_make_record(o.xy.x, o.xy.y, o.ca, o.cb, o.cc, op, r, g, b, o.depth, q0, q1, q2);
    __syncthreads();
    float4 *rs = reinterpret_cast<float4 *>(rest_s);
    rs[3 * tid] = q0; rs[3 * tid + 1] = q1; rs[3 * tid + 2] = q2;
    __syncthreads();
    for (int f = tid; f < 3 * rows; f += PP_TPB) rec[3 * (size_t)g0 + f] = rs[f];
}

}  // namespace

static int launch_preprocess_fwd(int part, int N, const float *means, const float *log_scales, const float *quats,
                                 const float *opac_logits, const float *feat_dc, const float *feat_rest, int use_sh,
                                 int sh_deg, const GsCamera *cam_host, float *xys, float *depths, int32_t *radii,
                                 float *conics, int32_t *tiles_hit, float *rgbs, float *opac, void *stream) {
    const int blocks = (N + PP_TPB - 1) / PP_TPB;
    cudaStream_t st = (cudaStream_t)stream;
#define GS_PP_ARGS                                                                                                       \
    N, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, feat_dc, feat_rest, use_sh, sh_deg,       \
        *cam_host, reinterpret_cast<float2 *>(xys), depths, radii, conics, tiles_hit, rgbs, opac,                        \
        (const int32_t *)nullptr, (float4 *)nullptr
    if (part == 0) preprocess_fwd_kernel<0, false><<<blocks, PP_TPB, 0, st>>>(GS_PP_ARGS);
    else if (part == 1) preprocess_fwd_kernel<1, false><<<blocks, PP_TPB, 0, st>>>(GS_PP_ARGS);
    else preprocess_fwd_kernel<2, false><<<blocks, PP_TPB, 0, st>>>(GS_PP_ARGS);
#undef GS_PP_ARGS
    GS_LAUNCH_CHECK();
    return 0;
}

// internal (view.cu): packed-record form, optional index list; N = number of elements (subset length when given)
int gs_launch_preprocess_fwd_packed(int N, const float *means, const float *log_scales, const float *quats,
                                    const float *opac_logits, const float *feat_dc, const float *feat_rest, int use_sh,
                                    int sh_deg, const GsCamera *cam_host, const int32_t *subset, float *rec, float *depths,
                                    int32_t *radii, int32_t *tiles_hit, cudaStream_t st) {
    if (N <= 0) return 0;
    const int blocks = (N + PP_TPB - 1) / PP_TPB;
    preprocess_fwd_kernel<0, true><<<blocks, PP_TPB, 0, st>>>(
        N, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, feat_dc, feat_rest, use_sh, sh_deg,
        *cam_host, (float2 *)nullptr, depths, radii, (float *)nullptr, tiles_hit, (float *)nullptr, (float *)nullptr,
        subset, reinterpret_cast<float4 *>(rec));
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_preprocess_fwd(int N, const float *means, const float *log_scales, const float *quats,
                                 const float *opac_logits, const float *feat_dc, const float *feat_rest,
                                 int use_sh, int sh_deg, const GsCamera *cam_host,
                                 float *xys, float *depths, int32_t *radii, float *conics, int32_t *tiles_hit,
                                 float *rgbs, float *opac, void *stream) {
    if (N < 0 || !cam_host) return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (cam_host->H <= 0 || cam_host->W <= 0 || cam_host->tiles_x <= 0 || cam_host->tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!means || !log_scales || !quats || !opac_logits || !feat_dc || !xys || !depths || !radii || !conics ||
        !tiles_hit || !rgbs || !opac)
        return GS_ERR_BAD_ARG;
    if (use_sh && sh_deg >= 1 && !feat_rest) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(feat_rest) || (reinterpret_cast<uintptr_t>(xys) & 7u))
        return GS_ERR_MISALIGNED;
    return launch_preprocess_fwd(0, N, means, log_scales, quats, opac_logits, feat_dc, feat_rest, use_sh, sh_deg, cam_host,
                                 xys, depths, radii, conics, tiles_hit, rgbs, opac, stream);
}

extern "C" int gs_preprocess_fwd_part(int part, int N, const float *means, const float *log_scales, const float *quats,
                                      const float *opac_logits, const float *feat_dc, const float *feat_rest,
                                      int use_sh, int sh_deg, const GsCamera *cam_host, float *xys, float *depths,
                                      int32_t *radii, float *conics, int32_t *tiles_hit, float *rgbs, float *opac,
                                      void *stream) {
    if (part != 1 && part != 2) return GS_ERR_BAD_ARG;
    if (N < 0 || !cam_host) return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (cam_host->H <= 0 || cam_host->W <= 0 || cam_host->tiles_x <= 0 || cam_host->tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (part == 1 && (!means || !log_scales || !quats || !opac_logits || !xys || !depths || !radii || !conics ||
                      !tiles_hit || !opac))
        return GS_ERR_BAD_ARG;
    if (part == 2 && (!means || !feat_dc || !radii || !rgbs || (use_sh && sh_deg >= 1 && !feat_rest)))
        return GS_ERR_BAD_ARG;
    if ((part == 1 && (!gs_aligned16(quats) || (reinterpret_cast<uintptr_t>(xys) & 7u))) ||
        (part == 2 && !gs_aligned16(feat_rest)))
        return GS_ERR_MISALIGNED;
    return launch_preprocess_fwd(part, N, means, log_scales, quats, opac_logits, feat_dc, feat_rest, use_sh, sh_deg,
                                 cam_host, xys, depths, radii, conics, tiles_hit, rgbs, opac, stream);
}

extern "C" int gs_preprocess_fwd_packed(int N, const float *means, const float *log_scales, const float *quats,
                                        const float *opac_logits, const float *feat_dc, const float *feat_rest,
                                        const int32_t *subset, int use_sh, int sh_deg, const GsCamera *cam_host,
                                        float *rec, float *depths, int32_t *radii, int32_t *tiles_hit, void *stream) {
    if (N < 0 || !cam_host) return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (cam_host->H <= 0 || cam_host->W <= 0 || cam_host->tiles_x <= 0 || cam_host->tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!means || !log_scales || !quats || !opac_logits || !feat_dc || !rec || !depths || !radii || !tiles_hit)
        return GS_ERR_BAD_ARG;
    if (use_sh && sh_deg >= 1 && !feat_rest) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(feat_rest) || !gs_aligned16(rec)) return GS_ERR_MISALIGNED;
    return gs_launch_preprocess_fwd_packed(N, means, log_scales, quats, opac_logits, feat_dc, feat_rest, use_sh, sh_deg,
                                           cam_host, subset, rec, depths, radii, tiles_hit, (cudaStream_t)stream);
}

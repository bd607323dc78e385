// Sparse NVLink exchange of per-view compositor gradients (data-parallel fine-tune step).
//
// Only ~22 % of the Gaussians receive a non-zero gradient from one view (the rest are culled or hidden
// behind saturated pixels; measured on the 1 M bench scene), so the view block a rank publishes to its
// peers is compacted first:
//     meta[W]     {bitmap word: 1 bit per Gaussian = any of its 9 compositor-level gradients non-zero,
//                  prefix: number of set bits before this 32-Gaussian word}
//     camera      64-float slot (GsCamera)
//     packed      v_xy | v_conic | v_rgb | v_opacity of the set Gaussians, in index order (SoA)
// Receivers (gs_preprocess_bwd_multi_peer_sparse) walk the Gaussians; a warp owns one bitmap word, so the
// word and its prefix are one broadcast load each and the packed values of the warp's set lanes are
// consecutive (coalesced peer loads). NVLink bytes per view drop from 36 N to ~0.25 N + 36 nnz.
#include "project_math.cuh"
#include "sh_math.cuh"

namespace {

constexpr int VX_TPB = 128;
constexpr int VX_REST = 45;
constexpr int VX_MAX_VIEWS = 16;

__host__ __device__ inline int vx_words(int N) { return (((N + 31) / 32) + 3) & ~3; }

__device__ __forceinline__ unsigned vx_smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// ---- sender 1: bitmap + per-word popcount ------------------------------------------------------
__global__ void __launch_bounds__(256)
vx_bitmap_kernel(int N, int Np, const float *__restrict__ blk, uint32_t *__restrict__ bitmap,
                 int32_t *__restrict__ wcount) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool nz = false;
    if (i < N) {
        const float2 xy = reinterpret_cast<const float2 *>(blk)[i];
        const float *vc = blk + (size_t)2 * Np + 3 * (size_t)i;
        const float *vr = blk + (size_t)5 * Np + 3 * (size_t)i;
        nz = xy.x != 0.f || xy.y != 0.f || vc[0] != 0.f || vc[1] != 0.f || vc[2] != 0.f || vr[0] != 0.f ||
             vr[1] != 0.f || vr[2] != 0.f || blk[(size_t)8 * Np + i] != 0.f;
    }
    const unsigned m = __ballot_sync(0xffffffffu, nz);
    if ((threadIdx.x & 31) == 0) {
        const int w = i >> 5;
        if (w < vx_words(N)) { bitmap[w] = m; wcount[w] = __popc(m); }
    }
}

// ---- sender 2: pack the set Gaussians ---------------------------------------------------------------
__global__ void __launch_bounds__(256)
vx_pack_kernel(int N, int Np, int Wp, const float *__restrict__ blk, const uint32_t *__restrict__ bitmap,
               const int32_t *__restrict__ prefix, uint2 *__restrict__ meta, float *__restrict__ packed) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if ((threadIdx.x & 31) == 0 && (i >> 5) < Wp)      // one 8-byte {bitmap word, prefix} record per 32 Gaussians
        meta[i >> 5] = make_uint2(bitmap[i >> 5], (uint32_t)prefix[i >> 5]);
    if (i >= N) return;
    const uint32_t word = bitmap[i >> 5];
    if (!((word >> (i & 31)) & 1u)) return;
    const int k = prefix[i >> 5] + __popc(word & gs_lanemask_lt());
    const float2 xy = reinterpret_cast<const float2 *>(blk)[i];
    reinterpret_cast<float2 *>(packed)[k] = xy;
    const float *vc = blk + (size_t)2 * Np + 3 * (size_t)i;
    const float *vr = blk + (size_t)5 * Np + 3 * (size_t)i;
    float *pc = packed + (size_t)2 * Np + 3 * (size_t)k;
    float *pr = packed + (size_t)5 * Np + 3 * (size_t)k;
    pc[0] = vc[0]; pc[1] = vc[1]; pc[2] = vc[2];
    pr[0] = vr[0]; pr[1] = vr[1]; pr[2] = vr[2];
    packed[(size_t)8 * Np + k] = blk[(size_t)8 * Np + i];
}

struct ViewGrad {
    float2 xy;
    float a, b, c, u0, u1, u2, o;
    bool set;
};

// ---- receiver: multi-view leaf backward over the peers' compacted blocks ---------------------------
__global__ void __launch_bounds__(VX_TPB)
vx_bwd_multi_sparse_kernel(int N, int Np, int Wp, int V, const unsigned long long *__restrict__ peer_ptrs,
                           size_t peer_off, float scale, const float *__restrict__ means,
                           const float *__restrict__ log_scales, const float4 *__restrict__ quats,
                           const float *__restrict__ opac_logits, int use_sh, int deg, float *__restrict__ v_means,
                           float *__restrict__ v_log_scales, float4 *__restrict__ v_quats,
                           float *__restrict__ v_opac_logits, float *__restrict__ v_dc, float *__restrict__ v_rest) {
    __shared__ __align__(128) float rest_s[VX_TPB * VX_REST];
    __shared__ GsCamera cam_s[VX_MAX_VIEWS];
    __shared__ const float *blk_s[VX_MAX_VIEWS];
    const int tid = threadIdx.x, lane = tid & 31;
    const int g0 = blockIdx.x * VX_TPB;
    const int rows = min(VX_TPB, N - g0);
    const int i = g0 + tid;
    const int w = i >> 5;    // this warp's bitmap word (g0 is a multiple of 128)
    float *row = rest_s + tid * VX_REST;
#pragma unroll
    for (int k = 0; k < VX_REST; ++k) row[k] = 0.f;
    if (tid < V) blk_s[tid] = reinterpret_cast<const float *>(peer_ptrs[tid]) + peer_off;
    __syncthreads();
    constexpr int CAMF = (int)(sizeof(GsCamera) / 4);
    for (int k = tid; k < V * CAMF; k += VX_TPB)
        reinterpret_cast<float *>(cam_s)[k] = blk_s[k / CAMF][(size_t)2 * Wp + k % CAMF];

    float vm[3] = {0.f, 0.f, 0.f}, vls[3] = {0.f, 0.f, 0.f}, vd[3] = {0.f, 0.f, 0.f};
    float4 vq = make_float4(0.f, 0.f, 0.f, 0.f);
    float vop = 0.f;
    float p0 = 0.f, p1 = 0.f, p2 = 0.f, s0 = 1.f, s1 = 1.f, s2 = 1.f, iq = 1.f, o = 0.f;
    float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
    if (i < N) {
        p0 = means[3 * i]; p1 = means[3 * i + 1]; p2 = means[3 * i + 2];
        s0 = expf(log_scales[3 * i]); s1 = expf(log_scales[3 * i + 1]); s2 = expf(log_scales[3 * i + 2]);
        const float4 qr = quats[i];
        iq = 1.0f / sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w);
        q = make_float4(qr.x * iq, qr.y * iq, qr.z * iq, qr.w * iq);
        o = 1.0f / (1.0f + expf(-opac_logits[i]));
    }
    // {word, prefix} of EVERY view first (one broadcast 8-byte peer load per view, all in flight together),
    // then the packed values, software-prefetched one view ahead: NVLink latency (~2 us) is paid ~twice per
    // CTA instead of three times per view
    uint2 meta[VX_MAX_VIEWS];
#pragma unroll
    for (int v = 0; v < VX_MAX_VIEWS; ++v)
        meta[v] = (v < V && i < N) ? reinterpret_cast<const uint2 *>(blk_s[v])[w] : make_uint2(0u, 0u);
    auto load_view = [&](int v) {
        ViewGrad g;
        const float *blk = blk_s[v];
        const uint32_t word = meta[v].x;
        g.set = (word >> lane) & 1u;
        if (g.set) {
            const int k = (int)meta[v].y + __popc(word & gs_lanemask_lt());
            const float *pk = blk + (size_t)2 * Wp + 64;
            g.xy = reinterpret_cast<const float2 *>(pk)[k];
            const float *vc = pk + (size_t)2 * Np + 3 * (size_t)k;
            const float *vr = pk + (size_t)5 * Np + 3 * (size_t)k;
            g.a = vc[0]; g.b = vc[1]; g.c = vc[2];
            g.u0 = vr[0]; g.u1 = vr[1]; g.u2 = vr[2];
            g.o = pk[(size_t)8 * Np + k];
        }
        return g;
    };
    __syncthreads();   // cameras staged
    if (i < N) {
        ViewGrad cur = load_view(0);
#pragma unroll
        for (int v = 0; v < VX_MAX_VIEWS; ++v) {
            if (v >= V) break;
            ViewGrad nxt = cur;
            if (v + 1 < V && v + 1 < VX_MAX_VIEWS) nxt = load_view(v + 1);
            const ViewGrad g9 = cur;
            cur = nxt;
            if (!g9.set) continue;
            const GsCamera &cam = cam_s[v];
            ProjGrad g;
            gs_project_one_bwd(cam.viewmat, cam.projmat, p0, p1, p2, s0, s1, s2, cam.glob_scale, q, cam.fx, cam.fy,
                               cam.H, cam.W, g9.xy, 0.0f, g9.a, g9.b, g9.c, g);
            vm[0] += g.vp[0]; vm[1] += g.vp[1]; vm[2] += g.vp[2];
            vls[0] += g.vs[0]; vls[1] += g.vs[1]; vls[2] += g.vs[2];
            vq.x += g.vq.x; vq.y += g.vq.y; vq.z += g.vq.z; vq.w += g.vq.w;
            vop += g9.o;
            if (use_sh) {
                float dx = p0 - cam.cam_pos[0], dy = p1 - cam.cam_pos[1], dz = p2 - cam.cam_pos[2];
                const float vn = sqrtf(dx * dx + dy * dy + dz * dz);
                dx = dx / vn; dy = dy / vn; dz = dz / vn;
                float Y[16];
                sh_basis<3>(deg, dx, dy, dz, Y);
                const int nb = sh_num_bases(deg);
                vd[0] += Y[0] * g9.u0; vd[1] += Y[0] * g9.u1; vd[2] += Y[0] * g9.u2;
#pragma unroll
                for (int k = 1; k < 16; ++k)
                    if (k < nb) {
                        row[3 * (k - 1)] += Y[k] * g9.u0; row[3 * (k - 1) + 1] += Y[k] * g9.u1;
                        row[3 * (k - 1) + 2] += Y[k] * g9.u2;
                    }
            } else {
                vd[0] += g9.u0; vd[1] += g9.u1; vd[2] += g9.u2;
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
        for (int k = 0; k < VX_REST; ++k) row[k] *= scale;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int rows4 = rows & ~3;
    float *dst = v_rest + (size_t)g0 * VX_REST;
    if (tid == 0 && rows4 > 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(vx_smem_u32(rest_s)),
                     "r"((unsigned)rows4 * VX_REST * 4u)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    for (int f = rows4 * VX_REST + tid; f < rows * VX_REST; f += VX_TPB) dst[f] = rest_s[f];
    if (tid == 0 && rows4 > 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

}  // namespace

extern "C" int gs_view_xbuf_floats(int N) {
    const int Np = (N + 3) / 4 * 4;
    return 2 * vx_words(N) + 64 + 9 * Np;
}

extern "C" size_t gs_view_pack_scratch_bytes(int N) {
    return gs_scan_scratch_bytes(vx_words(N)) + 3 * gs_align_up((size_t)vx_words(N) * 4, 256) + 256;
}

// block: finished dense view block (gs_finish_view_block already applied). xbuf: gs_view_xbuf_floats(N) floats.
extern "C" int gs_pack_view_block(int N, const float *block, float *xbuf, void *scratch, size_t scratch_bytes,
                                  void *stream) {
    if (N < 0 || !block || !xbuf || !scratch) return GS_ERR_BAD_ARG;
    if (scratch_bytes < gs_view_pack_scratch_bytes(N)) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(xbuf) || !gs_aligned16(scratch) || !gs_aligned16(block)) return GS_ERR_MISALIGNED;
    cudaStream_t st = (cudaStream_t)stream;
    const int Np = (N + 3) / 4 * 4, Wp = vx_words(N);
    uint2 *meta = reinterpret_cast<uint2 *>(xbuf);
    float *cam = xbuf + 2 * (size_t)Wp;
    float *packed = cam + 64;
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    const size_t w4 = gs_align_up((size_t)Wp * 4, 256);
    int32_t *wcount = reinterpret_cast<int32_t *>(p);
    uint32_t *bitmap = reinterpret_cast<uint32_t *>(p + w4);
    int32_t *prefix = reinterpret_cast<int32_t *>(p + 2 * w4);
    void *scan_scratch = p + 3 * w4;
    GS_CUDA(cudaMemcpyAsync(cam, block + (size_t)9 * Np, 64 * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (N == 0) return 0;
    const int threads_total = Wp * 32;
    vx_bitmap_kernel<<<(threads_total + 255) / 256, 256, 0, st>>>(N, Np, block, bitmap, wcount);
    GS_LAUNCH_CHECK();
    const int rc = gs_internal_exclusive_scan(Wp, wcount, prefix, nullptr, scan_scratch, scratch_bytes - 3 * w4, st);
    if (rc) return rc;
    vx_pack_kernel<<<(threads_total + 255) / 256, 256, 0, st>>>(N, Np, Wp, block, bitmap, prefix, meta, packed);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_preprocess_bwd_multi_peer_sparse(int N, int V, const uint64_t *peer_base_ptrs_dev,
                                                   int64_t offset_floats, float scale, const float *means,
                                                   const float *log_scales, const float *quats,
                                                   const float *opac_logits, int use_sh, int sh_deg,
                                                   float *v_means, float *v_log_scales, float *v_quats,
                                                   float *v_opac_logits, float *v_feat_dc, float *v_feat_rest,
                                                   void *stream) {
    if (N < 0 || V <= 0 || V > VX_MAX_VIEWS || !peer_base_ptrs_dev) return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (N == 0) return 0;
    if (!means || !log_scales || !quats || !opac_logits || !v_means || !v_log_scales || !v_quats || !v_opac_logits ||
        !v_feat_dc || !v_feat_rest)
        return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(v_quats) || !gs_aligned16(v_feat_rest) || (offset_floats & 3))
        return GS_ERR_MISALIGNED;
    const int Np = (N + 3) / 4 * 4;
    vx_bwd_multi_sparse_kernel<<<(N + VX_TPB - 1) / VX_TPB, VX_TPB, 0, (cudaStream_t)stream>>>(
        N, Np, vx_words(N), V, reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev), (size_t)offset_floats,
        scale, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg, v_means,
        v_log_scales, reinterpret_cast<float4 *>(v_quats), v_opac_logits, v_feat_dc, v_feat_rest);
    GS_LAUNCH_CHECK();
    return 0;
}

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
#include <algorithm>
#include <cstdlib>

#define GS_PDL_GROUP 4
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
    gs_pdl_sync();
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
               const int32_t *__restrict__ prefix, float *__restrict__ xbuf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    gs_pdl_sync();
    if (blockIdx.x == 0 && threadIdx.x < 64) xbuf[(size_t)2 * Wp + threadIdx.x] = blk[(size_t)9 * Np + threadIdx.x];
    if ((threadIdx.x & 31) == 0 && (i >> 5) < Wp)      // one 8-byte {bitmap word, prefix} record per 32 Gaussians
        reinterpret_cast<uint2 *>(xbuf)[i >> 5] = make_uint2(bitmap[i >> 5], (uint32_t)prefix[i >> 5]);
    if (i >= N) return;
    const uint32_t word = bitmap[i >> 5];
    if (!((word >> (i & 31)) & 1u)) return;
    const int k = prefix[i >> 5] + __popc(word & gs_lanemask_lt());
    float *packed = xbuf + (size_t)2 * Wp + 64;
    reinterpret_cast<float2 *>(packed)[k] = reinterpret_cast<const float2 *>(blk)[i];
    const float *vc = blk + (size_t)2 * Np + 3 * (size_t)i;
    const float *vr = blk + (size_t)5 * Np + 3 * (size_t)i;
    float *pc = packed + (size_t)2 * Np + 3 * (size_t)k;
    float *pr = packed + (size_t)5 * Np + 3 * (size_t)k;
    pc[0] = vc[0]; pc[1] = vc[1]; pc[2] = vc[2];
    pr[0] = vr[0]; pr[1] = vr[1]; pr[2] = vr[2];
    packed[(size_t)8 * Np + k] = blk[(size_t)8 * Np + i];
}

// ---- sender 3 (PUSH exchange): copy the compacted buffer into every peer's slot ----------------------------
// Posted NVLink stores, 16 B per lane and 512 contiguous bytes per warp instruction; only the used part of
// each packed array travels (the count is read from the last meta record). Scattered 4-byte peer stores
// straight from the pack kernel were ~3x slower than pack-locally-then-copy.
constexpr int VX_PUSH_TPB = 256;

__global__ void __launch_bounds__(VX_PUSH_TPB)
vx_push_kernel(int Np, int Wp, int D, int self, const float4 *__restrict__ src,
               const unsigned long long *__restrict__ peer_ptrs, size_t peer_off) {
    __shared__ float4 *dst_s[VX_MAX_VIEWS];
    gs_pdl_sync();
    if (threadIdx.x < D) dst_s[threadIdx.x] = reinterpret_cast<float4 *>(reinterpret_cast<float *>(peer_ptrs[threadIdx.x]) + peer_off);
    __syncthreads();
    const uint2 last = reinterpret_cast<const uint2 *>(src)[Wp - 1];
    const int nnz = (int)last.y + __popc(last.x);
    const int h4 = (2 * Wp + 64) / 4, q = Np / 4;
    const int n_xy = (2 * nnz + 3) / 4, n_c = (3 * nnz + 3) / 4, n_o = (nnz + 3) / 4;
    const int total = h4 + n_xy + 2 * n_c + n_o;
    for (int t = blockIdx.x * VX_PUSH_TPB + threadIdx.x; t < total; t += gridDim.x * VX_PUSH_TPB) {
        int idx = t;                                        // float4 index inside the exchange buffer
        if (t >= h4) {
            int r = t - h4;
            if (r < n_xy) idx = h4 + r;
            else if ((r -= n_xy) < n_c) idx = h4 + 2 * q + r;
            else if ((r -= n_c) < n_c) idx = h4 + 5 * q + r;
            else idx = h4 + 8 * q + (r - n_c);
        }
        const float4 val = src[idx];
        for (int d = 0; d < D; ++d)
            if (d != self) dst_s[d][idx] = val;
    }
}

// ---- sender 3b (SLICE exchange): peer d owns the Gaussians [d * S, (d + 1) * S) for the leaf backward, so it only needs
// that range of this view's compacted block: its meta words, the camera, and — the packed arrays being in index
// order — ONE contiguous run [k_lo, k_hi) of each of them (k = packed index from the meta prefix at the slice
// boundaries). ~1/p of the block per peer instead of all of it: the bytes a rank sends per step no longer grow with p.
// grid (blocks, D): blockIdx.y = destination. Runs are copied as float4 supersets (a few values of the neighbouring
// slices ride along; the destination never reads them).
__global__ void __launch_bounds__(VX_PUSH_TPB)
vx_push_slices_kernel(int Np, int Wp, int D, int self, int slice_words, const float4 *__restrict__ src,
                      const unsigned long long *__restrict__ peer_ptrs, size_t peer_off) {
    gs_pdl_sync();
    const int d = blockIdx.y;
    if (d == self) return;
    const int wlo = min(d * slice_words, Wp), whi = min(wlo + slice_words, Wp);
    if (wlo >= whi) return;
    float4 *dst = reinterpret_cast<float4 *>(reinterpret_cast<float *>(peer_ptrs[d]) + peer_off);
    const uint2 *meta = reinterpret_cast<const uint2 *>(src);
    const uint2 last = meta[Wp - 1];
    const int nnz = (int)last.y + __popc(last.x);
    const int k_lo = (int)meta[wlo].y, k_hi = whi < Wp ? (int)meta[whi].y : nnz;
    const int h4 = (2 * Wp + 64) / 4, q = Np / 4;
    // float4 runs: [begin, end) pairs inside the exchange buffer
    const int m0 = wlo / 2, m1 = (whi + 1) / 2;                               // meta (2 words per float4)
    const int c0 = (2 * Wp) / 4, c1 = h4;                                     // camera
    const int xy0 = h4 + (2 * k_lo) / 4, xy1 = h4 + (2 * k_hi + 3) / 4;
    const int cn0 = h4 + 2 * q + (3 * k_lo) / 4, cn1 = h4 + 2 * q + (3 * k_hi + 3) / 4;
    const int rg0 = h4 + 5 * q + (3 * k_lo) / 4, rg1 = h4 + 5 * q + (3 * k_hi + 3) / 4;
    const int op0 = h4 + 8 * q + k_lo / 4, op1 = h4 + 8 * q + (k_hi + 3) / 4;
    const int n0 = m1 - m0, n1 = n0 + (c1 - c0), n2 = n1 + (xy1 - xy0), n3 = n2 + (cn1 - cn0), n4 = n3 + (rg1 - rg0);
    const int total = n4 + (op1 - op0);
    for (int t = blockIdx.x * VX_PUSH_TPB + threadIdx.x; t < total; t += gridDim.x * VX_PUSH_TPB) {
        int idx;
        if (t < n0) idx = m0 + t;
        else if (t < n1) idx = c0 + (t - n0);
        else if (t < n2) idx = xy0 + (t - n1);
        else if (t < n3) idx = cn0 + (t - n2);
        else if (t < n4) idx = rg0 + (t - n3);
        else idx = op0 + (t - n4);
        dst[idx] = src[idx];
    }
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
        s0 = gs_exp_scale(log_scales[3 * i]); s1 = gs_exp_scale(log_scales[3 * i + 1]); s2 = gs_exp_scale(log_scales[3 * i + 2]);
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


// ---- receiver v2: item-compacted multi-view leaf backward ------------------------------------------------
// The loop-over-views kernel above runs the whole projection/SH backward for a view whenever ANY of a warp's
// 32 Gaussians is set in it — with ~22 % density that is every warp, every view, at ~7 useful lanes. Here a
// warp first enumerates its (Gaussian, view) ITEMS from the bitmap words, Gaussian-major, then processes them
// 32 at a time with all lanes busy: lane j computes item j into a 31-float slot (projection vjp 10, opacity 1,
// colour gradient 3, SH basis 16, constant 1). The slots are then folded into the 59 per-Gaussian sums with
// the lanes TRANSPOSED over the component index (lane e owns components e and e+32 of every Gaussian): the
// items of one Gaussian are consecutive, so each lane keeps a running sum in a register and stores it when the
// owner changes — no shared-memory read-modify-write chain. Every Gaussian sums its views in view order:
// deterministic, and the same on every rank.
constexpr int VI_TPB = 128;
constexpr int VI_HEAD = 14;      // vm3 vls3 vq4 vop1 vdc3
constexpr int VI_HSTRIDE = 15;
constexpr int VI_SLOT = 31;      // odd stride: per-lane slot rows stay off each other's banks; [30] holds 1.0f
constexpr int VI_PRM = 11;

struct __align__(16) ViWarp {
    float rest[32 * VX_REST];      // contiguous: bulk-stored to v_feat_rest
    float head[32 * VI_HSTRIDE];
    float slot[32 * VI_SLOT];
    float prm[32 * VI_PRM];
    uint32_t word[VX_MAX_VIEWS];
    uint32_t pref[VX_MAX_VIEWS];
};
static_assert(sizeof(ViWarp) % 16 == 0, "per-warp smem must keep 16 B alignment");

// dynamic smem: 4 ViWarp | block pointers | V cameras | 4 x 32 V item records (owner | view << 8, Gaussian-major).
// 56 KB at V = 8: four CTAs per SM.
__host__ __device__ inline size_t vi_smem_bytes(int V) {
    return 4 * sizeof(ViWarp) + VX_MAX_VIEWS * sizeof(const float *) + (size_t)V * sizeof(GsCamera) +
           4 * 32 * (size_t)V * sizeof(uint16_t);
}

// g_first (a multiple of VI_TPB): first Gaussian of the range this launch covers; N = its end. The six output
// pointers are those of ROW 0 (for a slice-shaped output the caller passes them pre-offset by -g_first rows).
__global__ void __launch_bounds__(VI_TPB)
vx_bwd_items_kernel(int g_first, int N, int Np, int Wp, int V, const unsigned long long *__restrict__ peer_ptrs, size_t peer_off,
                    float scale, const float *__restrict__ means, const float *__restrict__ log_scales,
                    const float4 *__restrict__ quats, const float *__restrict__ opac_logits, int use_sh, int deg,
                    float *__restrict__ v_means, float *__restrict__ v_log_scales, float4 *__restrict__ v_quats,
                    float *__restrict__ v_opac_logits, float *__restrict__ v_dc, float *__restrict__ v_rest) {
    extern __shared__ __align__(16) unsigned char vi_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    ViWarp &ws = reinterpret_cast<ViWarp *>(vi_smem)[warp];
    const float **blk_s = reinterpret_cast<const float **>(vi_smem + 4 * sizeof(ViWarp));
    GsCamera *cam_s = reinterpret_cast<GsCamera *>(blk_s + VX_MAX_VIEWS);
    uint16_t *items = reinterpret_cast<uint16_t *>(cam_s + V) + warp * 32 * V;
    const int i = g_first + blockIdx.x * VI_TPB + tid;
    const int w = i >> 5;
    if (tid < V) blk_s[tid] = reinterpret_cast<const float *>(peer_ptrs[tid]) + peer_off;
    __syncthreads();
    constexpr int CAMF = (int)(sizeof(GsCamera) / 4);
    for (int k = tid; k < V * CAMF; k += VI_TPB)
        reinterpret_cast<float *>(cam_s)[k] = blk_s[k / CAMF][(size_t)2 * Wp + k % CAMF];
    // this warp's bitmap word + packed-array prefix of every view (lane v fetches view v)
    uint2 m = make_uint2(0u, 0u);
    if (lane < V) m = reinterpret_cast<const uint2 *>(blk_s[lane])[w];
    if (lane < VX_MAX_VIEWS) { ws.word[lane] = m.x; ws.pref[lane] = m.y; }

    float s0 = 1.f, s1 = 1.f, s2 = 1.f, iq = 1.f, o = 0.f;
    float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
    if (i < N) {
        const float p0 = means[3 * i], p1 = means[3 * i + 1], p2 = means[3 * i + 2];
        s0 = gs_exp_scale(log_scales[3 * i]); s1 = gs_exp_scale(log_scales[3 * i + 1]); s2 = gs_exp_scale(log_scales[3 * i + 2]);
        const float4 qr = quats[i];
        iq = 1.0f / sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w);
        q = make_float4(qr.x * iq, qr.y * iq, qr.z * iq, qr.w * iq);
        o = 1.0f / (1.0f + expf(-opac_logits[i]));
        float *pr = ws.prm + lane * VI_PRM;
        pr[0] = p0; pr[1] = p1; pr[2] = p2; pr[3] = s0; pr[4] = s1; pr[5] = s2;
        pr[6] = q.x; pr[7] = q.y; pr[8] = q.z; pr[9] = q.w;
    }
    ws.slot[lane * VI_SLOT + 30] = 1.0f;
    __syncwarp();
    // item list: lane = Gaussian; its set views in ascending order, Gaussians in index order
    uint32_t vmask = 0;
#pragma unroll
    for (int v = 0; v < VX_MAX_VIEWS; ++v) vmask |= ((ws.word[v] >> lane) & 1u) << v;
    int incl = __popc(vmask);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const int T = __shfl_sync(0xffffffffu, incl, 31);
    {
        int pos = incl - __popc(vmask);
        for (uint32_t mk = vmask; mk; mk &= mk - 1) items[pos++] = (uint16_t)(lane | ((__ffs(mk) - 1) << 8));
    }
    // Gaussians without any item keep zero sums (the fold below only ever writes rows that have items)
    const uint32_t empty = __ballot_sync(0xffffffffu, vmask == 0);
    if (__popc(empty) > 10) {
        for (int e = lane; e < 32 * VX_REST; e += 32) ws.rest[e] = 0.f;
        for (int e = lane; e < 32 * VI_HSTRIDE; e += 32) ws.head[e] = 0.f;
    } else {
        for (uint32_t em = empty; em; em &= em - 1) {
            const int row = __ffs(em) - 1;
            ws.rest[row * VX_REST + lane] = 0.f;
            if (lane < VX_REST - 32) ws.rest[row * VX_REST + 32 + lane] = 0.f;
            if (lane < VI_HSTRIDE) ws.head[row * VI_HSTRIDE + lane] = 0.f;
        }
    }
    __syncwarp();
    __syncthreads();   // cameras + per-warp tables staged

    // fold: this lane owns components e1 = lane and e2 = lane + 32 (of 59) of every Gaussian.
    //   component e < 11: slot[e] * slot[30] (= 1);  e >= 11: slot[14 + (e-11)/3] * slot[11 + (e-11)%3]
    const int e1 = lane, e2 = lane + 32;
    const bool e2_ok = e2 < VI_HEAD + VX_REST;
    const int a1 = e1 < 11 ? e1 : 14 + (e1 - 11) / 3, b1 = e1 < 11 ? 30 : 11 + (e1 - 11) % 3;
    const int a2 = e2_ok ? 14 + (e2 - 11) / 3 : 30, b2 = e2_ok ? 11 + (e2 - 11) % 3 : 30;
    const int nb = use_sh ? sh_num_bases(deg) : 1;
    // running sums of the current owner; stored after EVERY item (last store of an owner = its full sum), which
    // keeps the loop free of branches: the items of one Gaussian are consecutive
    int cur = -1;
    float sum1 = 0.f, sum2 = 0.f;
    float *dst1 = e1 < VI_HEAD ? ws.head + e1 : ws.rest + (e1 - VI_HEAD);
    const int stride1 = e1 < VI_HEAD ? VI_HSTRIDE : VX_REST;
    float *dst2 = e2_ok ? ws.rest + (e2 - VI_HEAD) : ws.prm + lane * VI_PRM + 10;     // lanes 27..31: an unused word
    const int stride2 = e2_ok ? VX_REST : 0;

    for (int r0 = 0; r0 < T; r0 += 32) {
        const int j = r0 + lane;
        if (j < T) {
            const int it = items[j];
            const int ow = it & 31, v = it >> 8;
            const int k = (int)ws.pref[v] + __popc(ws.word[v] & ((1u << ow) - 1u));
            const float *pk = blk_s[v] + (size_t)2 * Wp + 64;
            float2 xy = reinterpret_cast<const float2 *>(pk)[k];
            const float *vc = pk + (size_t)2 * Np + 3 * (size_t)k;
            const float *vr = pk + (size_t)5 * Np + 3 * (size_t)k;
            const float ga = vc[0] * scale, gb = vc[1] * scale, gc = vc[2] * scale;
            const float u0 = vr[0] * scale, u1 = vr[1] * scale, u2 = vr[2] * scale;
            const float go = pk[(size_t)8 * Np + k] * scale;
            xy.x *= scale; xy.y *= scale;
            const float *pr = ws.prm + ow * VI_PRM;
            const float p0 = pr[0], p1 = pr[1], p2 = pr[2];
            const GsCamera &cam = cam_s[v];
            ProjGrad g;
            gs_project_one_bwd(cam.viewmat, cam.projmat, p0, p1, p2, pr[3], pr[4], pr[5], cam.glob_scale,
                               make_float4(pr[6], pr[7], pr[8], pr[9]), cam.fx, cam.fy, cam.H, cam.W, xy, 0.0f, ga, gb,
                               gc, g);
            float *sl = ws.slot + lane * VI_SLOT;
            sl[0] = g.vp[0]; sl[1] = g.vp[1]; sl[2] = g.vp[2];
            sl[3] = g.vs[0]; sl[4] = g.vs[1]; sl[5] = g.vs[2];
            sl[6] = g.vq.x; sl[7] = g.vq.y; sl[8] = g.vq.z; sl[9] = g.vq.w;
            sl[10] = go; sl[11] = u0; sl[12] = u1; sl[13] = u2;
            float Y[16];
#pragma unroll
            for (int b = 0; b < 16; ++b) Y[b] = 0.f;
            if (use_sh) {
                float dx = p0 - cam.cam_pos[0], dy = p1 - cam.cam_pos[1], dz = p2 - cam.cam_pos[2];
                const float vn = sqrtf(dx * dx + dy * dy + dz * dz);
                dx = dx / vn; dy = dy / vn; dz = dz / vn;
                sh_basis<3>(deg, dx, dy, dz, Y);
            } else {
                Y[0] = 1.0f;
            }
#pragma unroll
            for (int b = 0; b < 16; ++b) sl[14 + b] = b < nb ? Y[b] : 0.f;
        }
        __syncwarp();
        const int cnt = min(32, T - r0);
#pragma unroll 4
        for (int jj = 0; jj < cnt; ++jj) {
            const int ow = items[r0 + jj] & 31;
            const float *sl = ws.slot + jj * VI_SLOT;
            const float x1 = sl[a1], y1 = sl[b1], x2 = sl[a2], y2 = sl[b2];
            const bool same = ow == cur;
            cur = ow;
            sum1 = fmaf(x1, y1, same ? sum1 : 0.f);
            sum2 = fmaf(x2, y2, same ? sum2 : 0.f);
            dst1[ow * stride1] = sum1;
            dst2[ow * stride2] = sum2;
        }
        __syncwarp();
    }
    __syncwarp();

    const int w0 = w << 5;                    // first Gaussian of this warp
    const int rows = min(32, N - w0);
    if (i < N) {
        const float *a = ws.head + lane * VI_HSTRIDE;
        v_means[3 * i] = a[0]; v_means[3 * i + 1] = a[1]; v_means[3 * i + 2] = a[2];
        v_log_scales[3 * i] = a[3] * s0; v_log_scales[3 * i + 1] = a[4] * s1; v_log_scales[3 * i + 2] = a[5] * s2;
        const float4 vq = make_float4(a[6], a[7], a[8], a[9]);
        const float dq = vq.x * q.x + vq.y * q.y + vq.z * q.z + vq.w * q.w;
        v_quats[i] = make_float4((vq.x - q.x * dq) * iq, (vq.y - q.y * dq) * iq, (vq.z - q.z * dq) * iq,
                                 (vq.w - q.w * dq) * iq);
        v_opac_logits[i] = a[10] * o * (1.0f - o);
        v_dc[3 * i] = a[11]; v_dc[3 * i + 1] = a[12]; v_dc[3 * i + 2] = a[13];
    }
    if (rows == 32) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(v_rest + (size_t)w0 * VX_REST),
                         "r"(vx_smem_u32(ws.rest)), "r"((unsigned)(32 * VX_REST * 4))
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
    } else if (rows > 0) {
        for (int f = lane; f < rows * VX_REST; f += 32) v_rest[(size_t)w0 * VX_REST + f] = ws.rest[f];
    }
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
    if (N < 0 || !block || !scratch || !xbuf) return GS_ERR_BAD_ARG;
    if (scratch_bytes < gs_view_pack_scratch_bytes(N)) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(scratch) || !gs_aligned16(block) || !gs_aligned16(xbuf)) return GS_ERR_MISALIGNED;
    cudaStream_t st = (cudaStream_t)stream;
    const int Np = (N + 3) / 4 * 4, Wp = vx_words(N);
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    const size_t w4 = gs_align_up((size_t)Wp * 4, 256);
    int32_t *wcount = reinterpret_cast<int32_t *>(p);
    uint32_t *bitmap = reinterpret_cast<uint32_t *>(p + w4);
    int32_t *prefix = reinterpret_cast<int32_t *>(p + 2 * w4);
    void *scan_scratch = p + 3 * w4;
    const int threads_total = Wp * 32;
    vx_bitmap_kernel<<<(threads_total + 255) / 256, 256, 0, st>>>(N, Np, block, bitmap, wcount);
    GS_LAUNCH_CHECK();
    const int rc = gs_internal_exclusive_scan(Wp, wcount, prefix, nullptr, scan_scratch, scratch_bytes - 3 * w4, st);
    if (rc) return rc;
    GS_PDL_LAUNCH(vx_pack_kernel, (threads_total + 255) / 256, 256, 0, st, N, Np, Wp, block, (const uint32_t *)bitmap,
                  (const int32_t *)prefix, xbuf);
    return 0;
}

// PUSH: pack into the caller's own slot (xbuf_self == peer_base[self] + offset_floats, local memory), then copy
// the used part of it into the same slot of every other peer.
extern "C" int gs_pack_view_block_push(int N, const float *block, float *xbuf_self, int n_peers, int self,
                                       const uint64_t *peer_base_ptrs_dev, int64_t offset_floats, void *scratch,
                                       size_t scratch_bytes, void *stream) {
    if (!peer_base_ptrs_dev || offset_floats < 0 || n_peers <= 0 || n_peers > VX_MAX_VIEWS || self < 0 || self >= n_peers)
        return GS_ERR_BAD_ARG;
    if (offset_floats & 3) return GS_ERR_MISALIGNED;
    const int rc = gs_pack_view_block(N, block, xbuf_self, scratch, scratch_bytes, stream);
    if (rc || n_peers == 1) return rc;
    const int Np = (N + 3) / 4 * 4;
    GS_PDL_LAUNCH(vx_push_kernel, 148 * 4, VX_PUSH_TPB, 0, (cudaStream_t)stream, Np, vx_words(N), n_peers, self,
                  reinterpret_cast<const float4 *>(xbuf_self),
                  reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev), (size_t)offset_floats);
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
    static const bool loop_kernel = getenv("GS_VX_LOOP") != nullptr;    // A/B switch: the loop-over-views receiver
    if (!loop_kernel) {
        GS_CUDA(cudaFuncSetAttribute(vx_bwd_items_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)vi_smem_bytes(VX_MAX_VIEWS)));
        vx_bwd_items_kernel<<<(N + VI_TPB - 1) / VI_TPB, VI_TPB, vi_smem_bytes(V), (cudaStream_t)stream>>>(
            0, N, Np, vx_words(N), V, reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev), (size_t)offset_floats,
            scale, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg, v_means,
            v_log_scales, reinterpret_cast<float4 *>(v_quats), v_opac_logits, v_feat_dc, v_feat_rest);
        GS_LAUNCH_CHECK();
        return 0;
    }
    vx_bwd_multi_sparse_kernel<<<(N + VX_TPB - 1) / VX_TPB, VX_TPB, 0, (cudaStream_t)stream>>>(
        N, Np, vx_words(N), V, reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev), (size_t)offset_floats,
        scale, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg, v_means,
        v_log_scales, reinterpret_cast<float4 *>(v_quats), v_opac_logits, v_feat_dc, v_feat_rest);
    GS_LAUNCH_CHECK();
    return 0;
}

// ---- owner-sliced exchange (data-parallel step whose cost per rank does not grow with the number of ranks) ----------
// Slice size in Gaussians for `world` owners: ceil(N / world) rounded up to 128 (the leaf-backward CTA), so that slice
// boundaries are bitmap-word and float4 aligned. Owner d holds [d * S, min((d + 1) * S, N)).
extern "C" int gs_view_slice_size(int N, int world) {
    if (N < 0 || world <= 0) return GS_ERR_BAD_ARG;
    const int per = (N + world - 1) / world;
    return (per + VI_TPB - 1) / VI_TPB * VI_TPB;
}

// Like gs_pack_view_block_push, but every peer receives only the part of the compacted block that covers ITS slice.
extern "C" int gs_pack_view_block_push_slices(int N, const float *block, float *xbuf_self, int n_peers, int self,
                                              const uint64_t *peer_base_ptrs_dev, int64_t offset_floats, void *scratch,
                                              size_t scratch_bytes, void *stream) {
    if (!peer_base_ptrs_dev || offset_floats < 0 || n_peers <= 0 || n_peers > VX_MAX_VIEWS || self < 0 || self >= n_peers)
        return GS_ERR_BAD_ARG;
    if (offset_floats & 3) return GS_ERR_MISALIGNED;
    const int rc = gs_pack_view_block(N, block, xbuf_self, scratch, scratch_bytes, stream);
    if (rc || n_peers == 1 || N == 0) return rc;
    const int Np = (N + 3) / 4 * 4;
    const int S = gs_view_slice_size(N, n_peers);
    cudaLaunchConfig_t c = {};
    c.gridDim = dim3(148, n_peers);
    c.blockDim = dim3(VX_PUSH_TPB);
    c.stream = (cudaStream_t)stream;
    cudaLaunchAttribute a[1];
    a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    a[0].val.programmaticStreamSerializationAllowed = 1;
    c.attrs = a;
    c.numAttrs = gs_pdl_enabled(GS_PDL_GROUP) ? 1 : 0;
    GS_CUDA(cudaLaunchKernelEx(&c, vx_push_slices_kernel, Np, vx_words(N), n_peers, self, S / 32,
                               reinterpret_cast<const float4 *>(xbuf_self),
                               reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev), (size_t)offset_floats));
    ++g_gs_launches;
    return 0;
}

// Leaf backward of the Gaussians [g_lo, g_hi) over V compacted view blocks (g_lo a multiple of 128). Outputs are
// SLICE-shaped: row 0 of every output array is Gaussian g_lo. v_feat_rest must be 16-byte aligned.
extern "C" int gs_leaf_bwd_slice(int N, int g_lo, int g_hi, int V, const uint64_t *peer_base_ptrs_dev,
                                 int64_t offset_floats, float scale, const float *means, const float *log_scales,
                                 const float *quats, const float *opac_logits, int use_sh, int sh_deg, float *v_means,
                                 float *v_log_scales, float *v_quats, float *v_opac_logits, float *v_feat_dc,
                                 float *v_feat_rest, void *stream) {
    if (N < 0 || V <= 0 || V > VX_MAX_VIEWS || !peer_base_ptrs_dev || g_lo < 0 || g_hi > N || (g_lo % VI_TPB) != 0)
        return GS_ERR_BAD_ARG;
    if (sh_deg < 0 || sh_deg > 3) return GS_ERR_UNSUPPORTED;
    if (g_hi <= g_lo) return 0;
    if (!means || !log_scales || !quats || !opac_logits || !v_means || !v_log_scales || !v_quats || !v_opac_logits ||
        !v_feat_dc || !v_feat_rest)
        return GS_ERR_BAD_ARG;
    if (!gs_aligned16(quats) || !gs_aligned16(v_quats) || !gs_aligned16(v_feat_rest) || (offset_floats & 3))
        return GS_ERR_MISALIGNED;
    const int Np = (N + 3) / 4 * 4;
    GS_CUDA(cudaFuncSetAttribute(vx_bwd_items_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)vi_smem_bytes(VX_MAX_VIEWS)));
    const size_t o = (size_t)g_lo;
    vx_bwd_items_kernel<<<(g_hi - g_lo + VI_TPB - 1) / VI_TPB, VI_TPB, vi_smem_bytes(V), (cudaStream_t)stream>>>(
        g_lo, g_hi, Np, vx_words(N), V, reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev),
        (size_t)offset_floats, scale, means, log_scales, reinterpret_cast<const float4 *>(quats), opac_logits, use_sh, sh_deg,
        v_means - 3 * o, v_log_scales - 3 * o, reinterpret_cast<float4 *>(v_quats) - o, v_opac_logits - o, v_feat_dc - 3 * o,
        v_feat_rest - (size_t)VX_REST * o);
    GS_LAUNCH_CHECK();
    return 0;
}

// ---- optimiser step of the owner-sliced mode, fused with the parameter exchange ---------------------------------------
// Adam (torch.optim.Adam's update rule; what gc_trainer.py:276-298 steps after backward, with the per-group learning
// rates of gc_config.py:58-87) over `count` consecutive elements of ONE parameter tensor — this rank's rows — and, in the
// same kernel, the all-gather that a replicated-parameter data-parallel step needs afterwards: every updated value is
// stored into the same element of every peer's copy of the tensor (posted NVLink stores into symmetric memory), so the
// transfer rides under the kernel's own HBM traffic (28 B read+written per element locally, 4 B per peer sent).
namespace {
__device__ __forceinline__ void adam_one(float &p, float g, float &m, float &v, float step_size, float b1, float b2, float eps,
                                         float inv_sqrt_c2) {
    m = b1 * m + (1.0f - b1) * g;
    v = b2 * v + (1.0f - b2) * g * g;
    p -= step_size * m / (sqrtf(v) * inv_sqrt_c2 + eps);
}

__global__ void __launch_bounds__(256)
adam_slice_kernel(int64_t count, float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                  float *__restrict__ v, float step_size, float b1, float b2, float eps, float inv_sqrt_c2, int D, int self,
                  const unsigned long long *__restrict__ peer_ptrs, size_t peer_off) {
    __shared__ float *dst_s[VX_MAX_VIEWS];
    if (threadIdx.x < D) dst_s[threadIdx.x] = peer_ptrs ? reinterpret_cast<float *>(peer_ptrs[threadIdx.x]) + peer_off : nullptr;
    __syncthreads();
    const int64_t n4 = count / 4;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n4; t += (int64_t)gridDim.x * blockDim.x) {
        const float4 gg = reinterpret_cast<const float4 *>(g)[t];
        float4 mm = reinterpret_cast<float4 *>(m)[t], vv = reinterpret_cast<float4 *>(v)[t], pp = reinterpret_cast<float4 *>(p)[t];
        adam_one(pp.x, gg.x, mm.x, vv.x, step_size, b1, b2, eps, inv_sqrt_c2);
        adam_one(pp.y, gg.y, mm.y, vv.y, step_size, b1, b2, eps, inv_sqrt_c2);
        adam_one(pp.z, gg.z, mm.z, vv.z, step_size, b1, b2, eps, inv_sqrt_c2);
        adam_one(pp.w, gg.w, mm.w, vv.w, step_size, b1, b2, eps, inv_sqrt_c2);
        reinterpret_cast<float4 *>(m)[t] = mm;
        reinterpret_cast<float4 *>(v)[t] = vv;
        reinterpret_cast<float4 *>(p)[t] = pp;
        for (int d = 0; d < D; ++d)
            if (d != self && dst_s[d]) reinterpret_cast<float4 *>(dst_s[d])[t] = pp;
    }
    // tail (count not a multiple of 4)
    for (int64_t e = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (int64_t)gridDim.x * blockDim.x) {
        float pp = p[e], mm = m[e], vv = v[e];
        adam_one(pp, g[e], mm, vv, step_size, b1, b2, eps, inv_sqrt_c2);
        m[e] = mm; v[e] = vv; p[e] = pp;
        for (int d = 0; d < D; ++d)
            if (d != self && dst_s[d]) dst_s[d][e] = pp;
    }
}
}  // namespace

// p, g, m, v: `count` fp32 elements each, 16-byte aligned (p = first owned element of the parameter tensor).
// step_size = lr / (1 - b1^t), inv_sqrt_c2 = 1 / sqrt(1 - b2^t). n_peers > 1 with peer_base_ptrs_dev != NULL: p[e] is also
// stored at peer_base[d] + offset_floats + e of every other peer d (offset_floats = position of p[0] inside the
// symmetric buffer; a multiple of 4). The caller orders this after every peer's last read of the old parameters (the
// exchange barrier of the step) and runs a device barrier before the next forward.
extern "C" int gs_adam_slice(int64_t count, float *p, const float *g, float *m, float *v, float step_size, float b1, float b2,
                             float eps, float inv_sqrt_c2, int n_peers, int self, const uint64_t *peer_base_ptrs_dev,
                             int64_t offset_floats, void *stream) {
    if (count < 0 || n_peers < 0 || n_peers > VX_MAX_VIEWS || (n_peers > 1 && (self < 0 || self >= n_peers))) return GS_ERR_BAD_ARG;
    if (count == 0) return 0;
    if (!p || !g || !m || !v) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(p) || !gs_aligned16(g) || !gs_aligned16(m) || !gs_aligned16(v) || (offset_floats & 3)) return GS_ERR_MISALIGNED;
    const int64_t n4 = count / 4 + 1;
    const int blocks = (int)std::min<int64_t>((n4 + 255) / 256, 148 * 8);
    adam_slice_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
        count, p, g, m, v, step_size, b1, b2, eps, inv_sqrt_c2, peer_base_ptrs_dev ? n_peers : 0, self,
        reinterpret_cast<const unsigned long long *>(peer_base_ptrs_dev), (size_t)offset_floats);
    GS_LAUNCH_CHECK();
    return 0;
}

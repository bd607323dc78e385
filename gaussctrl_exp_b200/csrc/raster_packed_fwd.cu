// Stage 3 — forward alpha compositing over PACKED splat records (replaces gsplat's
// rasterize_forward reached from /root/reference/gaussctrl/gc_model.py:208-220 and :225-236).
//
// One CTA per 16x16 tile (of one of `views` cameras), one pixel per thread. Per 256-Gaussian batch of the tile's
// depth-sorted list:
//   * thread t copies record ids_sorted[k0 + t] (48 B) into the shared-memory stage with three 16-byte cp.async;
//     the copy of batch b+1 is issued right after the barrier that publishes batch b, so it runs underneath the
//     compositing of batch b (double-buffered stage, ONE barrier per batch; the id of batch b+2 is prefetched
//     into a register at the same time so that no cp.async ever waits for its address);
//   * the barrier is __syncthreads_count(done): the tile retires as soon as all 256 pixels have saturated;
//   * each warp owns an 8x4 pixel rect and walks the batch 32 records at a time: lane l tests record l's
//     alpha >= 1/255 box against the rect, the ballot is the list of survivors, and the warp composites exactly
//     the set bits front to back. Culled pairs are pairs the alpha test would skip, so no output bit changes.
// Bound: SM issue rate (FP32 + MUFU.EX2) and shared-memory broadcast reads, not HBM.
#define GS_PDL_GROUP 2
#include <cstdlib>
#include "raster_packed.cuh"

namespace {

constexpr int RP_BATCH = 256;

template <bool WITH_DEPTH, int RP_UNROLL>
__global__ void __launch_bounds__(256)
raster_fwd_packed_kernel(int H, int W, int tiles_x, int T_view, const int32_t *__restrict__ ids_sorted,
                         const int2 *__restrict__ tile_bins, const float4 *__restrict__ rec,
                         const float *__restrict__ bg, float *__restrict__ out_img, float *__restrict__ out_depth,
                         float *__restrict__ final_T, int32_t *__restrict__ final_idx,
                         const int32_t *__restrict__ tile_order, int32_t *__restrict__ tile_work,
                         float *__restrict__ out_clamped, float *__restrict__ out_alpha,
                         const int32_t *__restrict__ total_dev, int I_cap) {
    __shared__ float4 s0[2][RP_BATCH];   // x, y, ex, ey
    __shared__ float4 s1[2][RP_BATCH];   // opac, qa, qb, qc
    __shared__ float4 s2[2][RP_BATCH];   // r, g, b, depth
    __shared__ int s_work;
    gs_pdl_sync();
    // heaviest tiles first (LPT): CTAs are dispatched in blockIdx order
    const int gtile = tile_order ? tile_order[blockIdx.x] : blockIdx.x;
    const int view = gtile / T_view, tile = gtile - view * T_view;
    const int tile_x = tile % tiles_x, tile_y = tile / tiles_x;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) s_work = 0;
    int px, py;
    gs_pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < W && py < H;
    const float fx = (float)px + GS_PIX_CENTER, fy = (float)py + GS_PIX_CENTER;
    float rx0, ry0;
    gs_warp_rect(tile_x, tile_y, rx0, ry0);
    const float rx1 = rx0 + 7.0f, ry1 = ry0 + 3.0f;
    const int2 range = tile_bins[gtile];
    const int n = range.y - range.x;
    const int nbatch = (n + RP_BATCH - 1) / RP_BATCH;

    float T = 1.0f, r = 0.f, g = 0.f, b = 0.f, dacc = 0.f;
    int last = range.x;
    bool done = !inside;

    auto load_id = [&](int batch) -> int {
        const int k = range.x + batch * RP_BATCH + tid;
        return k < range.y ? ids_sorted[k] : -1;
    };
    auto issue = [&](int id, int stage) {
        if (id >= 0) {
            const float4 *src = rec + (size_t)GS_REC_QUADS * (size_t)id;
            gs_cp_async16(gs_smem_addr(&s0[stage][tid]), src);
            gs_cp_async16(gs_smem_addr(&s1[stage][tid]), src + 1);
            gs_cp_async16(gs_smem_addr(&s2[stage][tid]), src + 2);
        }
        gs_cp_async_commit();
    };
    int id_next = -1;
    if (nbatch > 0) {
        issue(load_id(0), 0);
        if (nbatch > 1) id_next = load_id(1);
    }
    for (int batch = 0; batch < nbatch; ++batch) {
        gs_cp_async_wait_all();
        if (__syncthreads_count(done) == 256) break;           // stage published; tile retires when saturated
        if (batch + 1 < nbatch) {
            issue(id_next, (batch + 1) & 1);                    // the other stage was last read in batch - 1
            if (batch + 2 < nbatch) id_next = load_id(batch + 2);
        }
        if (__all_sync(0xffffffffu, done)) continue;            // warp-uniform: this warp's pixels are saturated
        const int stage = batch & 1;
        const int cnt = min(RP_BATCH, n - batch * RP_BATCH);
        const int kbase = range.x + batch * RP_BATCH;
        const unsigned a0 = gs_smem_addr(s0[stage]), a1 = gs_smem_addr(s1[stage]), a2 = gs_smem_addr(s2[stage]);
        bool warp_done = false;
        for (int c32 = 0; c32 < cnt && !warp_done; c32 += 32) {
            const int j = c32 + lane;
            bool keep = false;
            if (j < cnt) keep = gs_rect_hit(gs_lds128(a0 + (unsigned)j * 16u), rx0, ry0, rx1, ry1);
            unsigned word = __ballot_sync(0xffffffffu, keep);
            // survivors RP_UNROLL at a time, all three quads of a survivor loaded up front (the colour quad used to be
            // loaded inside the branch that needs it: a dependent shared-memory round trip on the critical path; the
            // loop spent ~47 % of its stall samples on fixed-latency and shared-memory dependencies,
            // profiles/r02_raster_source_sass.csv). Measured in one process (tools/raster_ab.py): 125 us before,
            // RP_UNROLL = 1: 116.5 us (40 registers), 2 / 3 / 4: 124 us (48 / 54 / 60 registers: one CTA per SM fewer) ->
            // default 1. The updates are applied in list order; the saturation vote is taken once per group.
            while (word) {
                unsigned jj[RP_UNROLL];
                bool has[RP_UNROLL];
#pragma unroll
                for (int u = 0; u < RP_UNROLL; ++u) {
                    has[u] = word != 0;
                    jj[u] = has[u] ? (unsigned)c32 + (unsigned)(__ffs(word) - 1) : jj[0];
                    word &= word - 1;
                }
                float alpha[RP_UNROLL];
                float4 C[RP_UNROLL];
                bool ok[RP_UNROLL];
#pragma unroll
                for (int u = 0; u < RP_UNROLL; ++u) {
                    const float4 A = gs_lds128(a0 + jj[u] * 16u), B = gs_lds128(a1 + jj[u] * 16u);
                    C[u] = gs_lds128(a2 + jj[u] * 16u);
                    GsSplat sp = {A.x, A.y, B.x, B.y, B.z, B.w};
                    float dx, dy, vis;
                    ok[u] = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha[u]) && has[u];
                }
#pragma unroll
                for (int u = 0; u < RP_UNROLL; ++u) {
                    if (!done && ok[u]) {
                        const float Tn = T * (1.0f - alpha[u]);
                        if (Tn <= GS_T_EPS) {
                            done = true;
                        } else {
                            const float w = alpha[u] * T;
                            r += w * C[u].x; g += w * C[u].y; b += w * C[u].z;
                            if (WITH_DEPTH) dacc += w * C[u].w;
                            T = Tn;
                            last = kbase + (int)jj[u] + 1;
                        }
                    }
                }
                if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
            }
        }
    }
    gs_cp_async_wait_all();     // a copy issued for a batch the tile never reached must land before the CTA exits
    if (inside) {
        const size_t pix = (size_t)view * H * W + (size_t)py * W + px;
        float c0 = r + T * bg[0], c1 = g + T * bg[1], c2 = b + T * bg[2];
        // the tile lists did not fit their capacity (the binning wrote nothing): poison the image instead of
        // returning a background-only render that looks valid
        if (total_dev && *total_dev > I_cap) c0 = c1 = c2 = __int_as_float(0x7fc00000);
        out_img[3 * pix] = c0;
        out_img[3 * pix + 1] = c1;
        out_img[3 * pix + 2] = c2;
        if (out_clamped) {     // the caller's clamp(rgb, max=1) (gc_model.py:222), NaN-propagating like torch
            out_clamped[3 * pix] = c0 > 1.0f ? 1.0f : c0;
            out_clamped[3 * pix + 1] = c1 > 1.0f ? 1.0f : c1;
            out_clamped[3 * pix + 2] = c2 > 1.0f ? 1.0f : c2;
        }
        if (out_alpha) out_alpha[pix] = 1.0f - T;
        if (WITH_DEPTH) out_depth[pix] = dacc;
        final_T[pix] = T;
        final_idx[pix] = last;
    }
    if (tile_work) {   // depth of the deepest contributor: the backward's work for this tile
        atomicMax(&s_work, last - range.x);
        __syncthreads();
        if (threadIdx.x == 0) tile_work[gtile] = s_work;
    }
}

// Operator-API front end: (xys, conics, colors, opac[, depths]) -> records. One thread per Gaussian.
__global__ void __launch_bounds__(256)
pack_splats_kernel(int N, const float2 *__restrict__ xys, const float *__restrict__ conics,
                   const float *__restrict__ colors, const float *__restrict__ opac, const float *__restrict__ depths,
                   float4 *__restrict__ rec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const float2 xy = xys[i];
    float4 q0, q1, q2;
    gs_make_record(xy.x, xy.y, conics[3 * i], conics[3 * i + 1], conics[3 * i + 2], opac[i], colors[3 * i],
                   colors[3 * i + 1], colors[3 * i + 2], depths ? depths[i] : 0.f, q0, q1, q2);
    rec[3 * (size_t)i] = q0;
    rec[3 * (size_t)i + 1] = q1;
    rec[3 * (size_t)i + 2] = q2;
}

}  // namespace

namespace {
int &raster_fwd_unroll() {
    static int u = [] {
        const char *e = getenv("GS_RASTER_FWD_UNROLL");
        return (e && e[0] >= '1' && e[0] <= '4') ? e[0] - '0' : 1;
    }();
    return u;
}
}  // namespace

extern "C" int gs_set_raster_fwd_unroll(int u) {
    const int prev = raster_fwd_unroll();
    if (u >= 1 && u <= 4) raster_fwd_unroll() = u;
    return prev;
}

int gs_launch_raster_fwd_packed(int views, int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                const int32_t *tile_bins, const float *rec, const float *bg, float *out_img,
                                float *out_depth, float *final_T, int32_t *final_idx, const int32_t *tile_order,
                                int32_t *tile_work, float *out_clamped, float *out_alpha, const int32_t *total_dev,
                                int64_t I_cap, cudaStream_t st) {
    const int T_view = tiles_x * tiles_y, tiles = views * T_view;
    const int2 *bins = reinterpret_cast<const int2 *>(tile_bins);
    const float4 *rec4 = reinterpret_cast<const float4 *>(rec);
    const bool pdl = tiles <= GS_RASTER_PDL_MAX_TILES;   // raster_common.cuh: the early launch hurts on many-wave grids
    // survivors composited per loop iteration (default 1, see the kernel): gs_set_raster_fwd_unroll / environment
    // GS_RASTER_FWD_UNROLL=1..4 for A/B
    const int unroll = raster_fwd_unroll();
#define GS_RP_LAUNCH(D, U)                                                                                              \
    GS_PDL_LAUNCH_IF(pdl, (raster_fwd_packed_kernel<D, U>), tiles, 256, 0, st, H, W, tiles_x, T_view, ids_sorted, bins,  \
                     rec4, bg, out_img, out_depth, final_T, final_idx, tile_order, tile_work, out_clamped, out_alpha,   \
                     total_dev, (int)I_cap)
    if (out_depth) {
        if (unroll == 1) GS_RP_LAUNCH(true, 1); else if (unroll == 2) GS_RP_LAUNCH(true, 2);
        else if (unroll == 3) GS_RP_LAUNCH(true, 3); else GS_RP_LAUNCH(true, 4);
    } else {
        if (unroll == 1) GS_RP_LAUNCH(false, 1); else if (unroll == 2) GS_RP_LAUNCH(false, 2);
        else if (unroll == 3) GS_RP_LAUNCH(false, 3); else GS_RP_LAUNCH(false, 4);
    }
#undef GS_RP_LAUNCH
    return 0;
}

extern "C" int gs_pack_splats(int N, const float *xys, const float *conics, const float *colors, const float *opac,
                              const float *depths, float *rec, void *stream) {
    if (N < 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!xys || !conics || !colors || !opac || !rec) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(rec) || (reinterpret_cast<uintptr_t>(xys) & 7u)) return GS_ERR_MISALIGNED;
    pack_splats_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        N, reinterpret_cast<const float2 *>(xys), conics, colors, opac, depths, reinterpret_cast<float4 *>(rec));
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_raster_fwd_packed(int views, int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                    const int32_t *tile_bins, const float *rec, const float *bg, float *out_img,
                                    float *out_depth, float *final_T, int32_t *final_idx, const int32_t *tile_order,
                                    int32_t *tile_work, float *out_clamped, float *out_alpha, void *stream) {
    if (views <= 0 || H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !rec || !bg || !out_img || !final_T || !final_idx) return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(rec)) return GS_ERR_MISALIGNED;
    return gs_launch_raster_fwd_packed(views, H, W, tiles_x, tiles_y, ids_sorted, tile_bins, rec, bg, out_img, out_depth,
                                       final_T, final_idx, tile_order, tile_work, out_clamped, out_alpha, nullptr, 0,
                                       (cudaStream_t)stream);
}

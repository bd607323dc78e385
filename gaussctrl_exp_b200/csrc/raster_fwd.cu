// Stage 3 — forward alpha compositing (replaces gsplat's
// rasterize_forward / nd_rasterize_forward reached from
// /root/reference/gaussctrl/gc_model.py:208-220 and :225-236).
//
// One CTA per 16x16 tile, one pixel per thread. The tile's depth-sorted Gaussians are
// gathered in batches of 256 into a double-buffered shared-memory stage: the gather for
// batch b+1 is issued into registers before batch b is composited, so its latency hides
// behind the arithmetic; one barrier per batch (__syncthreads_count) both publishes the
// stage and retires the tile as soon as every pixel has saturated.
// Bound: SM issue rate (FP32 + MUFU.EX2) and shared-memory broadcast reads, not HBM.
#define GS_PDL_GROUP 2
#include "raster_common.cuh"

namespace {

constexpr int RF_BATCH = 256;

struct StageRGB {
    float4 a[RF_BATCH];  // x, y, opac, qa          (q* = pre-scaled conic, raster_common.cuh)
    float4 b[RF_BATCH];  // qb, qc, r, g
    float2 c[RF_BATCH];  // b, view-space depth (RGB-D variant)
    float2 e[RF_BATCH];  // alpha >= 1/255 half-extents (gs_alpha_extent)
};

// WITH_DEPTH: a 4th channel (the Gaussian's view-space depth, zero background) is composited in the
// same pass: the eval path of gc_model.py:208-238 needs RGB + alpha + depth and gsplat renders (and
// bins, and sorts) twice for it.
template <bool WITH_DEPTH>
__global__ void __launch_bounds__(256)
raster_fwd3_kernel(int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                   const int2 *__restrict__ tile_bins, const float2 *__restrict__ xys,
                   const float *__restrict__ conics, const float *__restrict__ colors,
                   const float *__restrict__ opac, const float *__restrict__ depths,
                   const float *__restrict__ bg, float *__restrict__ out_img, float *__restrict__ out_depth,
                   float *__restrict__ final_T, int32_t *__restrict__ final_idx,
                   const int32_t *__restrict__ tile_order, int32_t *__restrict__ tile_work,
                   float *__restrict__ out_clamped, float *__restrict__ out_alpha) {
    __shared__ StageRGB st[2];
    __shared__ unsigned char wlist[8][RF_BATCH];   // per-warp compacted survivors of the current batch
    __shared__ int s_work;
    gs_pdl_sync();
    // heaviest tiles first (LPT): CTAs are dispatched in blockIdx order, so a descending-work permutation
    // keeps the long tiles out of the tail of the launch
    const int tile = tile_order ? tile_order[blockIdx.x] : blockIdx.x;
    if (threadIdx.x == 0) s_work = 0;
    const int tile_x = tile % tiles_x, tile_y = tile / tiles_x;
    int px, py;
    gs_pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < W && py < H;
    const float fx = (float)px + GS_PIX_CENTER, fy = (float)py + GS_PIX_CENTER;
    float rx0, ry0;
    gs_warp_rect(tile_x, tile_y, rx0, ry0);
    const float rx1 = rx0 + 7.0f, ry1 = ry0 + 3.0f;
    const int2 range = tile_bins[tile];
    const int n = range.y - range.x;
    const int nbatch = (n + RF_BATCH - 1) / RF_BATCH;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned char *mylist = wlist[warp];

    float T = 1.0f, r = 0.f, g = 0.f, b = 0.f, dacc = 0.f;
    int last = range.x;
    bool done = !inside;

    // register prefetch of the next batch: the gather latency hides behind the compositing
    float4 pa = make_float4(0.f, 0.f, 0.f, 0.f), pb = pa;
    float2 pc = make_float2(0.f, 0.f);
    float2 pe = make_float2(-1.f, -1.f);
    auto gather = [&](int batch) {
        const int k = range.x + batch * RF_BATCH + tid;
        if (k < range.y) {
            const int id = ids_sorted[k];
            const float2 xy = xys[id];
            const float ca = conics[3 * id], cb = conics[3 * id + 1], cc = conics[3 * id + 2];
            const float o = opac[id];
            float qa, qb, qc;
            gs_prescale_conic(ca, cb, cc, qa, qb, qc);
            pa = make_float4(xy.x, xy.y, o, qa);
            pb = make_float4(qb, qc, colors[3 * id], colors[3 * id + 1]);
            pc = make_float2(colors[3 * id + 2], WITH_DEPTH ? depths[id] : 0.f);
            pe = gs_alpha_extent(o, ca, cb, cc);
        }
    };
    if (nbatch > 0) gather(0);
    for (int batch = 0; batch < nbatch; ++batch) {
        StageRGB &s = st[batch & 1];
        s.a[tid] = pa; s.b[tid] = pb; s.c[tid] = pc; s.e[tid] = pe;
        if (__syncthreads_count(done) == 256) break;
        if (batch + 1 < nbatch) gather(batch + 1);
        if (__all_sync(0xffffffffu, done)) continue;          // warp-uniform: this warp's pixels are saturated
        const int cnt = min(RF_BATCH, n - batch * RF_BATCH);
        const int kbase = range.x + batch * RF_BATCH;
        // ---- per-warp culling: keep the Gaussians whose alpha-extent box touches this warp's 8x4 rect ----
        int nl = 0;
        for (int j0 = 0; j0 < cnt; j0 += 32) {
            const int j = j0 + lane;
            bool keep = false;
            if (j < cnt) {
                const float4 A = s.a[j];
                const float2 E = s.e[j];
                keep = (A.x + E.x >= rx0) && (A.x - E.x <= rx1) && (A.y + E.y >= ry0) && (A.y - E.y <= ry1);
            }
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) mylist[nl + __popc(m & gs_lanemask_lt())] = (unsigned char)j;
            nl += __popc(m);
        }
        __syncwarp();
        // ---- front-to-back compositing over the survivors ----
        const unsigned sa_a = gs_smem_addr(s.a), sb_a = gs_smem_addr(s.b), sc_a = gs_smem_addr(s.c);
        const unsigned list_a = gs_smem_addr(mylist);
        for (int t = 0; t < nl; ++t) {
            const unsigned j = gs_lds8(list_a + t);
            const float4 A = gs_lds128(sa_a + j * 16u);
            const float4 B = gs_lds128(sb_a + j * 16u);
            GsSplat sp = {A.x, A.y, A.z, A.w, B.x, B.y};
            float dx, dy, vis, alpha;
            const bool ok = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha);
            if (!done && ok) {
                const float Tn = T * (1.0f - alpha);
                if (Tn <= GS_T_EPS) {
                    done = true;
                } else {
                    const float w = alpha * T;
                    const float2 Cc = gs_lds64(sc_a + j * 8u);
                    r += w * B.z; g += w * B.w; b += w * Cc.x;
                    if (WITH_DEPTH) dacc += w * Cc.y;
                    T = Tn;
                    last = kbase + (int)j + 1;
                }
            }
            if (__all_sync(0xffffffffu, done)) break;
        }
    }
    if (inside) {
        const size_t pix = (size_t)py * W + px;
        const float c0 = r + T * bg[0], c1 = g + T * bg[1], c2 = b + T * bg[2];
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
        if (threadIdx.x == 0) tile_work[tile] = s_work;
    }
}

// Generic channel count (C <= CPAD): geometry staged in shared memory, colours read
// through L1 from global for contributing pairs only.
struct StageGeo {
    float4 a[RF_BATCH];  // x, y, opac, conic.a
    float2 b[RF_BATCH];  // conic.b, conic.c
    int id[RF_BATCH];
};

template <int CPAD>
__global__ void __launch_bounds__(256)
raster_fwdN_kernel(int C, int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                   const int2 *__restrict__ tile_bins, const float2 *__restrict__ xys,
                   const float *__restrict__ conics, const float *__restrict__ colors,
                   const float *__restrict__ opac, const float *__restrict__ bg, float *__restrict__ out_img,
                   float *__restrict__ final_T, int32_t *__restrict__ final_idx) {
    __shared__ StageGeo s;
    const int tile = blockIdx.x;
    const int tile_x = tile % tiles_x, tile_y = tile / tiles_x;
    int px, py;
    gs_pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < W && py < H;
    const float fx = (float)px + GS_PIX_CENTER, fy = (float)py + GS_PIX_CENTER;
    const int2 range = tile_bins[tile];
    const int n = range.y - range.x;
    const int nbatch = (n + RF_BATCH - 1) / RF_BATCH;
    const int tid = threadIdx.x;
    float T = 1.0f;
    float acc[CPAD];
#pragma unroll
    for (int c = 0; c < CPAD; ++c) acc[c] = 0.f;
    int last = range.x;
    bool done = !inside;
    for (int batch = 0; batch < nbatch; ++batch) {
        if (__syncthreads_count(done) == 256) break;
        const int k = range.x + batch * RF_BATCH + tid;
        if (k < range.y) {
            const int id = ids_sorted[k];
            const float2 xy = xys[id];
            float qa, qb, qc;
            gs_prescale_conic(conics[3 * id], conics[3 * id + 1], conics[3 * id + 2], qa, qb, qc);
            s.a[tid] = make_float4(xy.x, xy.y, opac[id], qa);
            s.b[tid] = make_float2(qb, qc);
            s.id[tid] = id;
        }
        __syncthreads();
        if (!done) {
            const int cnt = min(RF_BATCH, n - batch * RF_BATCH);
            const int kbase = range.x + batch * RF_BATCH;
            for (int j = 0; j < cnt; ++j) {
                const float4 A = s.a[j];
                const float2 B = s.b[j];
                GsSplat sp = {A.x, A.y, A.z, A.w, B.x, B.y};
                float dx, dy, vis, alpha;
                if (!gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha)) continue;
                const float Tn = T * (1.0f - alpha);
                if (Tn <= GS_T_EPS) { done = true; break; }
                const float w = alpha * T;
                const float *col = colors + (size_t)s.id[j] * C;
#pragma unroll
                for (int c = 0; c < CPAD; ++c)
                    if (c < C) acc[c] += w * col[c];
                T = Tn;
                last = kbase + (int)j + 1;
            }
        }
    }
    if (inside) {
        const size_t pix = (size_t)py * W + px;
#pragma unroll
        for (int c = 0; c < CPAD; ++c)
            if (c < C) out_img[pix * C + c] = acc[c] + T * bg[c];
        final_T[pix] = T;
        final_idx[pix] = last;
    }
}

}  // namespace

extern "C" int gs_raster_fwd(int C, int H, int W, int tiles_x, int tiles_y,
                             const int32_t *ids_sorted, const int32_t *tile_bins,
                             const float *xys, const float *conics, const float *colors, const float *opac,
                             const float *bg, float *out_img, float *final_T, int32_t *final_idx,
                             void *stream) {
    if (C <= 0 || H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (C > 32) return GS_ERR_UNSUPPORTED;
    if (!tile_bins || !xys || !conics || !colors || !opac || !bg || !out_img || !final_T || !final_idx)
        return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const int tiles = tiles_x * tiles_y;
    const int2 *bins = reinterpret_cast<const int2 *>(tile_bins);
    const float2 *xy2 = reinterpret_cast<const float2 *>(xys);
#define GS_FWD_ARGS H, W, tiles_x, ids_sorted, bins, xy2, conics, colors, opac, bg, out_img, final_T, final_idx
    if (C == 3)
        raster_fwd3_kernel<false><<<tiles, 256, 0, st>>>(H, W, tiles_x, ids_sorted, bins, xy2, conics, colors, opac, nullptr,
                                                         bg, out_img, nullptr, final_T, final_idx, nullptr, nullptr, nullptr, nullptr);
    else if (C <= 4) raster_fwdN_kernel<4><<<tiles, 256, 0, st>>>(C, GS_FWD_ARGS);
    else if (C <= 8) raster_fwdN_kernel<8><<<tiles, 256, 0, st>>>(C, GS_FWD_ARGS);
    else if (C <= 16) raster_fwdN_kernel<16><<<tiles, 256, 0, st>>>(C, GS_FWD_ARGS);
    else raster_fwdN_kernel<32><<<tiles, 256, 0, st>>>(C, GS_FWD_ARGS);
#undef GS_FWD_ARGS
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_raster_fwd_rgbd(int H, int W, int tiles_x, int tiles_y,
                                  const int32_t *ids_sorted, const int32_t *tile_bins,
                                  const float *xys, const float *conics, const float *colors, const float *opac,
                                  const float *depths, const float *bg, float *out_img, float *out_depth,
                                  float *final_T, int32_t *final_idx, void *stream) {
    if (H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !xys || !conics || !colors || !opac || !depths || !bg || !out_img || !out_depth || !final_T ||
        !final_idx)
        return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
    raster_fwd3_kernel<true><<<tiles_x * tiles_y, 256, 0, (cudaStream_t)stream>>>(
        H, W, tiles_x, ids_sorted, reinterpret_cast<const int2 *>(tile_bins), reinterpret_cast<const float2 *>(xys),
        conics, colors, opac, depths, bg, out_img, out_depth, final_T, final_idx, nullptr, nullptr, nullptr, nullptr);
    GS_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Tile scheduling: rank tiles by descending work (ties by tile id). All-pairs counting in shared
// memory: T <= a few thousand, so this is a few microseconds.
namespace {
__global__ void __launch_bounds__(256)
tile_order_kernel(int T, const int32_t *__restrict__ work, int work_is_range, int32_t *__restrict__ order) {
    // one warp per tile: rank = number of tiles with more work (ties: lower tile id first)
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    gs_pdl_sync();
    if (t >= T) return;
    auto load = [&](int u) { return work_is_range ? work[2 * u + 1] - work[2 * u] : work[u]; };
    const int mine = load(t);
    int rank = 0;
    for (int u = lane; u < T; u += 32) {
        const int w = load(u);
        rank += (w > mine) || (w == mine && u < t);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rank += __shfl_xor_sync(0xffffffffu, rank, o);
    if (lane == 0) order[rank] = t;
}

// More than 2 048 tiles (1080p views: 8 160; batches of views): the all-pairs ranking above grows with T^2 (37 us at
// 8 192 tiles). A schedule does not need an exact sort: ONE CTA bins the tiles into 1 024 work levels relative to the
// heaviest tile (shared-memory histogram, scan, cursor per level) and emits them heaviest level first; the order inside
// a level is whatever the atomics give. `order` is a permutation of 0..T-1 by construction (counted and placed with
// the same level function).
constexpr int TO_LEVELS = 1024;
__global__ void __launch_bounds__(TO_LEVELS)
tile_order_levels_kernel(int T, const int32_t *__restrict__ work, int work_is_range, int32_t *__restrict__ order) {
    __shared__ int s_cnt[TO_LEVELS];
    __shared__ int s_red[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    gs_pdl_sync();
    auto load = [&](int u) { return max(work_is_range ? work[2 * u + 1] - work[2 * u] : work[u], 0); };
    s_cnt[tid] = 0;
    int mx = 0;
    for (int u = tid; u < T; u += TO_LEVELS) mx = max(mx, load(u));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
    mx = s_red[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const long long wmax = max(mx, 1);
    auto level = [&](int w) { return (TO_LEVELS - 1) - (int)(((long long)w * (TO_LEVELS - 1)) / wmax); };   // 0 = heaviest
    for (int u = tid; u < T; u += TO_LEVELS) atomicAdd(&s_cnt[level(load(u))], 1);
    __syncthreads();
    // exclusive scan of the level counts (one per thread)
    const int c = s_cnt[tid];
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    __syncthreads();                 // every thread has read s_red above and its own count
    if (lane == 31) s_red[warp] = inc;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; ++w) woff += s_red[w];
    s_cnt[tid] = woff + inc - c;     // cursor of the level
    __syncthreads();
    for (int u = tid; u < T; u += TO_LEVELS) order[atomicAdd(&s_cnt[level(load(u))], 1)] = u;
}
}  // namespace

extern "C" int gs_tile_order(int T, const int32_t *work, int work_is_range, int32_t *order, void *stream) {
    if (T <= 0 || !work || !order) return GS_ERR_BAD_ARG;
    if (T > GS_RASTER_PDL_MAX_TILES) {
        tile_order_levels_kernel<<<1, TO_LEVELS, 0, (cudaStream_t)stream>>>(T, work, work_is_range, order);
        GS_LAUNCH_CHECK();
        return 0;
    }
    GS_PDL_LAUNCH(tile_order_kernel, (T * 32 + 255) / 256, 256, 0, (cudaStream_t)stream, T, work, work_is_range, order);
    return 0;
}

extern "C" int gs_raster_fwd_ordered_post(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                     const int32_t *tile_bins, const float *xys, const float *conics,
                                     const float *colors, const float *opac, const float *depths, const float *bg,
                                     float *out_img, float *out_depth, float *final_T, int32_t *final_idx,
                                     const int32_t *tile_order, int32_t *tile_work, float *out_clamped, float *out_alpha,
                                     void *stream) {
    if (H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !xys || !conics || !colors || !opac || !bg || !out_img || !final_T || !final_idx)
        return GS_ERR_BAD_ARG;
    if ((depths == nullptr) != (out_depth == nullptr)) return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const int2 *bins = reinterpret_cast<const int2 *>(tile_bins);
    const float2 *xy2 = reinterpret_cast<const float2 *>(xys);
    const int tiles = tiles_x * tiles_y;
    const bool pdl = tiles <= GS_RASTER_PDL_MAX_TILES;   // raster_common.cuh: the early launch hurts on many-wave grids
    if (depths)
        GS_PDL_LAUNCH_IF(pdl, (raster_fwd3_kernel<true>), tiles, 256, 0, st, H, W, tiles_x, ids_sorted, bins, xy2, conics,
                         colors, opac, depths, bg, out_img, out_depth, final_T, final_idx, tile_order, tile_work,
                         out_clamped, out_alpha);
    else
        GS_PDL_LAUNCH_IF(pdl, (raster_fwd3_kernel<false>), tiles, 256, 0, st, H, W, tiles_x, ids_sorted, bins, xy2, conics,
                         colors, opac, (const float *)nullptr, bg, out_img, (float *)nullptr, final_T, final_idx,
                         tile_order, tile_work, out_clamped, out_alpha);
    return 0;
}

extern "C" int gs_raster_fwd_ordered(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                     const int32_t *tile_bins, const float *xys, const float *conics,
                                     const float *colors, const float *opac, const float *depths, const float *bg,
                                     float *out_img, float *out_depth, float *final_T, int32_t *final_idx,
                                     const int32_t *tile_order, int32_t *tile_work, void *stream) {
    return gs_raster_fwd_ordered_post(H, W, tiles_x, tiles_y, ids_sorted, tile_bins, xys, conics, colors, opac, depths,
                                      bg, out_img, out_depth, final_T, final_idx, tile_order, tile_work, nullptr, nullptr,
                                      stream);
}

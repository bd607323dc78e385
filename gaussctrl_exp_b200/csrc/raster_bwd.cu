// Stage 4 — backward alpha compositing (replaces gsplat's
// rasterize_backward / nd_rasterize_backward reached from
// /root/reference/gaussctrl/gc_trainer.py:275).
//
// Exact derivative of raster_fwd.cu. One CTA per tile, one pixel per thread, back-to-front
// replay from each pixel's final_idx. gsplat 0.1.2 issues >= 9 global float atomics per
// (pixel, Gaussian) hit; here the 9 partial gradients of a Gaussian are
//   (1) skipped outright when the splat's alpha >= 1/255 box misses the warp's 8x4 pixel rect
//       (per-warp compacted survivor lists) or no lane of the warp hit it (warp ballot),
//   (2) reduced across the warp with a halving butterfly (9 + 5 shuffles instead of 45),
//   (3) accumulated in shared memory — one row set per tile that the warps add into with shared-memory atomics
//       (default), or eight private per-warp row sets summed at flush time (earlier variant, kept for A/B),
//   (4) flushed with ONE red.global.add per (tile, Gaussian, component) after the batch.
// Bound: SM issue rate + shuffle throughput; global atomics drop from 9*P to <= 9*I.
#include <stdlib.h>

#define GS_PDL_GROUP 2
#include "raster_common.cuh"

namespace {

constexpr int RB_BATCH = 128;
constexpr int RB_NV = 9;  // v_x, v_y, v_ca, v_cb, v_cc, v_opac, v_r, v_g, v_b

// in: v[0..7] per lane. out: sum over the warp of v[(lane >> 2) & 7] (same value on the 4
// lanes that share lane>>2).
__device__ __forceinline__ float warp_reduce8(const float *v, int lane) {
    float w[4], u[2];
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float send = h16 ? v[i] : v[i + 4];
        const float keep = h16 ? v[i + 4] : v[i];
        w[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float send = h8 ? w[i] : w[i + 2];
        const float keep = h8 ? w[i + 2] : w[i];
        u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    const float send = h4 ? u[0] : u[1];
    const float keep = h4 ? u[1] : u[0];
    float x = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    x += __shfl_xor_sync(0xffffffffu, x, 2);
    x += __shfl_xor_sync(0xffffffffu, x, 1);
    return x;
}

__device__ __forceinline__ float warp_sum(float x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
}

__device__ __forceinline__ int warp_max(int x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x = max(x, __shfl_xor_sync(0xffffffffu, x, o));
    return x;
}

// RB_B: Gaussians staged per batch. ATOMIC_ACC: the per-(tile, Gaussian) moment sums live in ONE shared row set that
// the warps add into with shared-memory atomics (a CAS loop on this architecture: LDS + FADD + ATOMS.CAST) instead of
// eight private per-warp row sets — 8x less shared memory for the sums (so a batch can be 256 Gaussians: half the
// barriers per Gaussian) and a flush that reads 9 instead of 72 values per Gaussian.
template <int RB_B, bool ATOMIC_ACC>
__global__ void __launch_bounds__(256)
raster_bwd3_kernel(int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                   const int2 *__restrict__ tile_bins, const float2 *__restrict__ xys,
                   const float *__restrict__ conics, const float *__restrict__ colors,
                   const float *__restrict__ opac, const float *__restrict__ bg,
                   const float *__restrict__ final_T, const int32_t *__restrict__ final_idx,
                   const float *__restrict__ v_out, const float *__restrict__ v_out_alpha,
                   float *__restrict__ v_xys, float *__restrict__ v_conics, float *__restrict__ v_colors,
                   float *__restrict__ v_opac, const int32_t *__restrict__ tile_order,
                   const float *__restrict__ clamp_ref) {
    __shared__ float4 sa[RB_B];  // x, y, opac, qa   (q* = pre-scaled conic, raster_common.cuh)
    __shared__ float4 sb[RB_B];  // qb, qc, r, g
    __shared__ float sc[RB_B];   // b
    __shared__ float2 se[RB_B];  // alpha >= 1/255 half-extents
    __shared__ int sid[RB_B];
    __shared__ float acc[ATOMIC_ACC ? 1 : 8][RB_B * RB_NV];
    __shared__ unsigned char wlist[8][RB_B];
    __shared__ int s_wmax[8];

    gs_pdl_sync();
    const int tile = tile_order ? tile_order[blockIdx.x] : blockIdx.x;   // heaviest tiles first
    const int tile_x = tile % tiles_x, tile_y = tile / tiles_x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int px, py;
    gs_pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < W && py < H;
    const float fx = (float)px + GS_PIX_CENTER, fy = (float)py + GS_PIX_CENTER;
    float rx0, ry0;
    gs_warp_rect(tile_x, tile_y, rx0, ry0);
    const float rx1 = rx0 + 7.0f, ry1 = ry0 + 3.0f;
    const int2 range = tile_bins[tile];
    const size_t pix = inside ? (size_t)py * W + px : 0;

    const int my_end = inside ? final_idx[pix] : range.x;
    const float Tf = inside ? final_T[pix] : 0.f;
    float T = Tf;
    float vo0 = 0.f, vo1 = 0.f, vo2 = 0.f, voa = 0.f;
    if (inside) {
        vo0 = v_out[3 * pix]; vo1 = v_out[3 * pix + 1]; vo2 = v_out[3 * pix + 2];
        if (clamp_ref) {       // v_out is w.r.t. clamp(out, max=1): no gradient where the raw output exceeds 1
            if (clamp_ref[3 * pix] > 1.0f) vo0 = 0.f;
            if (clamp_ref[3 * pix + 1] > 1.0f) vo1 = 0.f;
            if (clamp_ref[3 * pix + 2] > 1.0f) vo2 = 0.f;
        }
        if (v_out_alpha) voa = v_out_alpha[pix];
    }
    // d out / d alpha_k through the background and the alpha output: Tf/(1-a) * (voa - bg.vo)
    const float tail = Tf * (voa - (bg[0] * vo0 + bg[1] * vo1 + bg[2] * vo2));
    float Svo = 0.f;

    const int wmax = warp_max(my_end);
    if (lane == 0) s_wmax[warp] = wmax;
    for (int i = tid; i < (ATOMIC_ACC ? 1 : 8) * RB_B * RB_NV; i += 256) (&acc[0][0])[i] = 0.f;
    __syncthreads();
    int bmax = range.x;
#pragma unroll
    for (int w = 0; w < 8; ++w) bmax = max(bmax, s_wmax[w]);

    unsigned char *mylist = wlist[warp];
    const unsigned sa_a = gs_smem_addr(sa), sb_a = gs_smem_addr(sb), sc_a = gs_smem_addr(sc);
    const unsigned list_a = gs_smem_addr(mylist);
    // this lane's accumulator column: lanes 0,4,..,28 own moments 0..7, lane 1 owns the ninth
    const int acc_col = lane == 1 ? 8 : (lane >> 2);
    const unsigned acc_a = gs_smem_addr(acc[ATOMIC_ACC ? 0 : warp]) + (unsigned)acc_col * 4u;
    const bool acc_lane = (lane & 3) == 0 || lane == 1;
    for (int hi = bmax; hi > range.x; hi -= RB_B) {
        const int lo = max(range.x, hi - RB_B);
        const int cnt = hi - lo;
        if (tid < cnt) {
            const int id = ids_sorted[lo + tid];
            const float2 xy = xys[id];
            const float ca = conics[3 * id], cb = conics[3 * id + 1], cc = conics[3 * id + 2];
            const float o = opac[id];
            float qa, qb, qc;
            gs_prescale_conic(ca, cb, cc, qa, qb, qc);
            sa[tid] = make_float4(xy.x, xy.y, o, qa);
            sb[tid] = make_float4(qb, qc, colors[3 * id], colors[3 * id + 1]);
            sc[tid] = colors[3 * id + 2];
            se[tid] = gs_alpha_extent(o, ca, cb, cc);
            sid[tid] = id;
        }
        __syncthreads();
        // ---- per-warp culling (same exact-skip argument as the forward) ----
        const int jend = min(cnt, wmax - lo);   // warp-uniform: nothing past this warp's last contributor
        int nl = 0;
        for (int j0 = 0; j0 < jend; j0 += 32) {
            const int j = j0 + lane;
            bool keep = false;
            if (j < jend) {
                const float4 A = sa[j];
                const float2 E = se[j];
                keep = (A.x + E.x >= rx0) && (A.x - E.x <= rx1) && (A.y + E.y >= ry0) && (A.y - E.y <= ry1);
            }
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) mylist[nl + __popc(m & gs_lanemask_lt())] = (unsigned char)j;
            nl += __popc(m);
        }
        __syncwarp();
        // ---- back-to-front replay over the survivors ----
        // Per lane only the raw moments of v_sigma are formed; the conic / xy gradients are linear in
        // them and are assembled once per (tile, Gaussian) in the flush. S.vo (colour accumulated behind,
        // dotted with the upstream gradient) is carried as ONE running scalar.
        for (int t = nl - 1; t >= 0; --t) {
            const unsigned j = gs_lds8(list_a + t);
            const float4 A = gs_lds128(sa_a + j * 16u);
            const float4 B = gs_lds128(sb_a + j * 16u);
            GsSplat sp = {A.x, A.y, A.z, A.w, B.x, B.y};
            float dx, dy, vis, alpha;
            bool hit = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha);
            hit = hit && (lo + (int)j < my_end);
            if (!__any_sync(0xffffffffu, hit)) continue;
            // branch-free: a lane that does not hit runs the same arithmetic with alpha = 0 (T, Svo unchanged,
            // all nine contributions exactly zero)
            const float al = hit ? alpha : 0.0f;
            const float ra = gs_rcp(1.0f - al);
            T *= ra;
            const float fac = al * T;
            const float cvo = B.z * vo0 + B.w * vo1 + gs_lds32(sc_a + j * 4u) * vo2;
            const float v_alpha = T * cvo - ra * (Svo - tail);
            Svo += fac * cvo;
            const bool live = hit && (A.z * vis <= GS_ALPHA_MAX);      // not clamped at 0.999: d alpha / d sigma != 0
            const float visl = live ? vis : 0.0f;
            const float v_sigma = -A.z * visl * v_alpha;
            const float mx = v_sigma * dx, my = v_sigma * dy;
            float v[8];
            v[0] = mx; v[1] = my;
            v[2] = mx * dx; v[3] = mx * dy; v[4] = my * dy;
            v[5] = visl * v_alpha;
            v[6] = fac * vo0; v[7] = fac * vo1;
            const float v8 = fac * vo2;
            const float r8 = warp_reduce8(v, lane);
            const float r9 = warp_sum(v8);
            if (acc_lane) {
                const float add = lane == 1 ? r9 : r8;
                if (ATOMIC_ACC) {
                    if (add != 0.0f) atomicAdd(&acc[0][j * RB_NV + acc_col], add);
                } else {
                    const unsigned a = acc_a + j * (RB_NV * 4u);
                    gs_sts32(a, gs_lds32(a) + add);
                }
            }
        }
        __syncthreads();
        // flush: one thread per Gaussian sums the 8 warp rows, turns the moments into gradients
        //   v_xy = (a mx + b my, b mx + c my),  v_conic = (mxx/2, mxy, myy/2)
        // and issues one global reduction per (tile, Gaussian, component)
        if (tid < cnt) {
            float m[RB_NV];
#pragma unroll
            for (int c = 0; c < RB_NV; ++c) {
                float sum = 0.f;
#pragma unroll
                for (int w = 0; w < (ATOMIC_ACC ? 1 : 8); ++w) {
                    sum += acc[w][tid * RB_NV + c];
                    acc[w][tid * RB_NV + c] = 0.f;
                }
                m[c] = sum;
            }
            const int id = sid[tid];
            float ca, cb, cc;
            gs_unscale_conic(sa[tid].w, sb[tid].x, sb[tid].y, ca, cb, cc);
            if (m[0] != 0.f || m[1] != 0.f) {
                atomicAdd(v_xys + 2 * id, ca * m[0] + cb * m[1]);
                atomicAdd(v_xys + 2 * id + 1, cb * m[0] + cc * m[1]);
            }
            if (m[2] != 0.f) atomicAdd(v_conics + 3 * id, 0.5f * m[2]);
            if (m[3] != 0.f) atomicAdd(v_conics + 3 * id + 1, m[3]);
            if (m[4] != 0.f) atomicAdd(v_conics + 3 * id + 2, 0.5f * m[4]);
            if (m[5] != 0.f) atomicAdd(v_opac + id, m[5]);
            if (m[6] != 0.f) atomicAdd(v_colors + 3 * id, m[6]);
            if (m[7] != 0.f) atomicAdd(v_colors + 3 * id + 1, m[7]);
            if (m[8] != 0.f) atomicAdd(v_colors + 3 * id + 2, m[8]);
        }
        __syncthreads();
    }
}

// Generic channel count: same replay, per-value warp reduction + one global atomic per warp.
__global__ void __launch_bounds__(256)
raster_bwdN_kernel(int C, int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                   const int2 *__restrict__ tile_bins, const float2 *__restrict__ xys,
                   const float *__restrict__ conics, const float *__restrict__ colors,
                   const float *__restrict__ opac, const float *__restrict__ bg,
                   const float *__restrict__ final_T, const int32_t *__restrict__ final_idx,
                   const float *__restrict__ v_out, const float *__restrict__ v_out_alpha,
                   float *__restrict__ v_xys, float *__restrict__ v_conics, float *__restrict__ v_colors,
                   float *__restrict__ v_opac) {
    constexpr int CMAX = 32;
    __shared__ float4 sa[RB_BATCH];
    __shared__ float2 sb[RB_BATCH];
    __shared__ int sid[RB_BATCH];
    __shared__ int s_wmax[8];
    const int tile = blockIdx.x;
    const int tile_x = tile % tiles_x, tile_y = tile / tiles_x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int px, py;
    gs_pixel_of_thread(tile_x, tile_y, px, py);
    const bool inside = px < W && py < H;
    const float fx = (float)px + GS_PIX_CENTER, fy = (float)py + GS_PIX_CENTER;
    const int2 range = tile_bins[tile];
    const size_t pix = inside ? (size_t)py * W + px : 0;
    const int my_end = inside ? final_idx[pix] : range.x;
    const float Tf = inside ? final_T[pix] : 0.f;
    float T = Tf;
    float S[CMAX];
#pragma unroll
    for (int c = 0; c < CMAX; ++c) S[c] = 0.f;
    float bgdot = 0.f;
    if (inside)
        for (int c = 0; c < C; ++c) bgdot += bg[c] * v_out[pix * C + c];
    const float voa = (inside && v_out_alpha) ? v_out_alpha[pix] : 0.f;
    const float tail = Tf * (voa - bgdot);
    const int wmax = warp_max(my_end);
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    int bmax = range.x;
    for (int w = 0; w < 8; ++w) bmax = max(bmax, s_wmax[w]);
    for (int hi = bmax; hi > range.x; hi -= RB_BATCH) {
        const int lo = max(range.x, hi - RB_BATCH);
        const int cnt = hi - lo;
        __syncthreads();
        if (tid < cnt) {
            const int id = ids_sorted[lo + tid];
            const float2 xy = xys[id];
            float qa, qb, qc;
            gs_prescale_conic(conics[3 * id], conics[3 * id + 1], conics[3 * id + 2], qa, qb, qc);
            sa[tid] = make_float4(xy.x, xy.y, opac[id], qa);
            sb[tid] = make_float2(qb, qc);
            sid[tid] = id;
        }
        __syncthreads();
        const int jtop = min(cnt, wmax - lo) - 1;
        for (int j = jtop; j >= 0; --j) {
            const float4 A = sa[j];
            const float2 B = sb[j];
            GsSplat sp = {A.x, A.y, A.z, A.w, B.x, B.y};
            float dx, dy, vis, alpha;
            bool hit = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha);
            hit = hit && (lo + (int)j < my_end);
            if (!__any_sync(0xffffffffu, hit)) continue;
            const int id = sid[j];
            const float *col = colors + (size_t)id * C;
            float ra = 0.f, fac = 0.f, v_alpha = 0.f;
            if (hit) {
                ra = 1.0f / (1.0f - alpha);
                T *= ra;
                fac = alpha * T;
                v_alpha = tail * ra;
            }
#pragma unroll
            for (int c = 0; c < CMAX; ++c) {
                if (c < C) {
                    float vc = 0.f;
                    if (hit) {
                        const float vo = v_out[pix * C + c];
                        vc = fac * vo;
                        v_alpha += (col[c] * T - S[c] * ra) * vo;
                        S[c] += col[c] * fac;
                    }
                    vc = warp_sum(vc);
                    if (lane == 0 && vc != 0.f) atomicAdd(v_colors + (size_t)id * C + c, vc);
                }
            }
            float g[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (hit && A.z * vis <= GS_ALPHA_MAX) {
                const float v_sigma = -A.z * vis * v_alpha;
                float ca, cb, cc;
                gs_unscale_conic(A.w, B.x, B.y, ca, cb, cc);
                g[0] = v_sigma * (ca * dx + cb * dy);
                g[1] = v_sigma * (cb * dx + cc * dy);
                g[2] = 0.5f * v_sigma * dx * dx;
                g[3] = v_sigma * dx * dy;
                g[4] = 0.5f * v_sigma * dy * dy;
                g[5] = vis * v_alpha;
            }
#pragma unroll
            for (int q = 0; q < 6; ++q) g[q] = warp_sum(g[q]);
            if (lane == 0) {
                if (g[0] != 0.f) atomicAdd(v_xys + 2 * id, g[0]);
                if (g[1] != 0.f) atomicAdd(v_xys + 2 * id + 1, g[1]);
                if (g[2] != 0.f) atomicAdd(v_conics + 3 * id, g[2]);
                if (g[3] != 0.f) atomicAdd(v_conics + 3 * id + 1, g[3]);
                if (g[4] != 0.f) atomicAdd(v_conics + 3 * id + 2, g[4]);
                if (g[5] != 0.f) atomicAdd(v_opac + id, g[5]);
            }
        }
    }
}

}  // namespace

extern "C" int gs_raster_bwd(int C, int H, int W, int tiles_x, int tiles_y,
                             const int32_t *ids_sorted, const int32_t *tile_bins,
                             const float *xys, const float *conics, const float *colors, const float *opac,
                             const float *bg, const float *final_T, const int32_t *final_idx,
                             const float *v_out, const float *v_out_alpha,
                             float *v_xys, float *v_conics, float *v_colors, float *v_opac, void *stream) {
    if (C <= 0 || H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (C > 32) return GS_ERR_UNSUPPORTED;
    if (!tile_bins || !xys || !conics || !colors || !opac || !bg || !final_T || !final_idx || !v_out ||
        !v_xys || !v_conics || !v_colors || !v_opac)
        return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const int tiles = tiles_x * tiles_y;
    const int2 *bins = reinterpret_cast<const int2 *>(tile_bins);
    const float2 *xy2 = reinterpret_cast<const float2 *>(xys);
    if (C == 3)
        raster_bwd3_kernel<256, true><<<tiles, 256, 0, st>>>(H, W, tiles_x, ids_sorted, bins, xy2, conics, colors, opac, bg,
                                                  final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics,
                                                  v_colors, v_opac, nullptr, nullptr);
    else
        raster_bwdN_kernel<<<tiles, 256, 0, st>>>(C, H, W, tiles_x, ids_sorted, bins, xy2, conics, colors, opac,
                                                  bg, final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics,
                                                  v_colors, v_opac);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_raster_bwd_ordered_post(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                     const int32_t *tile_bins, const float *xys, const float *conics,
                                     const float *colors, const float *opac, const float *bg, const float *final_T,
                                     const int32_t *final_idx, const float *v_out, const float *v_out_alpha,
                                     float *v_xys, float *v_conics, float *v_colors, float *v_opac,
                                     const int32_t *tile_order, const float *clamp_ref, void *stream) {
    if (H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !xys || !conics || !colors || !opac || !bg || !final_T || !final_idx || !v_out || !v_xys ||
        !v_conics || !v_colors || !v_opac)
        return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
#define GS_RB_LAUNCH(B, A)                                                                                              \
    GS_PDL_LAUNCH_IF(tiles_x * tiles_y <= GS_RASTER_PDL_MAX_TILES, (raster_bwd3_kernel<B, A>), tiles_x * tiles_y, 256, 0, (cudaStream_t)stream, H, W, tiles_x, ids_sorted,  \
                  reinterpret_cast<const int2 *>(tile_bins), reinterpret_cast<const float2 *>(xys), conics, colors, opac,  \
                  bg, final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics, v_colors, v_opac, tile_order, clamp_ref)
    // default: 256-Gaussian batches with shared-atomic sums (0.715 ms/step); GS_RB_VARIANT=0 selects the earlier
    // 128-Gaussian batches with per-warp rows (0.727), =1 128-Gaussian batches with atomic sums (0.721). Same gradients
    // up to the order of fp32 additions inside a tile.
    static const int variant = [] { const char *e = getenv("GS_RB_VARIANT"); return e ? atoi(e) : 2; }();
    if (variant == 0) GS_RB_LAUNCH(128, false);
    else if (variant == 1) GS_RB_LAUNCH(128, true);
    else GS_RB_LAUNCH(256, true);
#undef GS_RB_LAUNCH
    return 0;
}

extern "C" int gs_raster_bwd_ordered(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                     const int32_t *tile_bins, const float *xys, const float *conics,
                                     const float *colors, const float *opac, const float *bg, const float *final_T,
                                     const int32_t *final_idx, const float *v_out, const float *v_out_alpha,
                                     float *v_xys, float *v_conics, float *v_colors, float *v_opac,
                                     const int32_t *tile_order, void *stream) {
    return gs_raster_bwd_ordered_post(H, W, tiles_x, tiles_y, ids_sorted, tile_bins, xys, conics, colors, opac, bg,
                                      final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics, v_colors, v_opac,
                                      tile_order, nullptr, stream);
}

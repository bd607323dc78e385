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
//   (3) accumulated, conflict- and atomic-free, into per-warp shared-memory rows,
//   (4) summed over the 8 warps and flushed with ONE red.global.add per (tile, Gaussian,
//       component) after the batch.
// Bound: SM issue rate + shuffle throughput; global atomics drop from 9*P to <= 9*I.
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

__global__ void __launch_bounds__(256)
raster_bwd3_kernel(int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                   const int2 *__restrict__ tile_bins, const float2 *__restrict__ xys,
                   const float *__restrict__ conics, const float *__restrict__ colors,
                   const float *__restrict__ opac, const float *__restrict__ bg,
                   const float *__restrict__ final_T, const int32_t *__restrict__ final_idx,
                   const float *__restrict__ v_out, const float *__restrict__ v_out_alpha,
                   float *__restrict__ v_xys, float *__restrict__ v_conics, float *__restrict__ v_colors,
                   float *__restrict__ v_opac) {
    __shared__ float4 sa[RB_BATCH];  // x, y, opac, conic.a
    __shared__ float4 sb[RB_BATCH];  // conic.b, conic.c, r, g
    __shared__ float sc[RB_BATCH];   // b
    __shared__ float2 se[RB_BATCH];  // alpha >= 1/255 half-extents
    __shared__ int sid[RB_BATCH];
    __shared__ float acc[8][RB_BATCH * RB_NV];
    __shared__ unsigned char wlist[8][RB_BATCH];
    __shared__ int s_wmax[8];

    const int tile = blockIdx.x;
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
        if (v_out_alpha) voa = v_out_alpha[pix];
    }
    // d out / d alpha_k through the background and the alpha output: Tf/(1-a) * (voa - bg.vo)
    const float tail = Tf * (voa - (bg[0] * vo0 + bg[1] * vo1 + bg[2] * vo2));
    float S0 = 0.f, S1 = 0.f, S2 = 0.f;

    const int wmax = warp_max(my_end);
    if (lane == 0) s_wmax[warp] = wmax;
    for (int i = tid; i < 8 * RB_BATCH * RB_NV; i += 256) (&acc[0][0])[i] = 0.f;
    __syncthreads();
    int bmax = range.x;
#pragma unroll
    for (int w = 0; w < 8; ++w) bmax = max(bmax, s_wmax[w]);

    float *my_acc = acc[warp];
    unsigned char *mylist = wlist[warp];
    for (int hi = bmax; hi > range.x; hi -= RB_BATCH) {
        const int lo = max(range.x, hi - RB_BATCH);
        const int cnt = hi - lo;
        if (tid < cnt) {
            const int id = ids_sorted[lo + tid];
            const float2 xy = xys[id];
            const float4 A = make_float4(xy.x, xy.y, opac[id], conics[3 * id]);
            const float4 B = make_float4(conics[3 * id + 1], conics[3 * id + 2], colors[3 * id], colors[3 * id + 1]);
            sa[tid] = A; sb[tid] = B;
            sc[tid] = colors[3 * id + 2];
            se[tid] = gs_alpha_extent(A.z, A.w, B.x, B.y);
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
        for (int t = nl - 1; t >= 0; --t) {
            const int j = mylist[t];
            const float4 A = sa[j];
            const float4 B = sb[j];
            GsSplat sp = {A.x, A.y, A.z, A.w, B.x, B.y};
            float dx, dy, vis, alpha;
            bool hit = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha);
            hit = hit && (lo + j < my_end);
            if (!__any_sync(0xffffffffu, hit)) continue;
            float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            float v8 = 0.f;
            if (hit) {
                const float cb_ = sc[j];
                const float ra = 1.0f / (1.0f - alpha);
                T *= ra;
                const float fac = alpha * T;
                v[6] = fac * vo0; v[7] = fac * vo1; v8 = fac * vo2;
                float v_alpha = (B.z * T - S0 * ra) * vo0 + (B.w * T - S1 * ra) * vo1 + (cb_ * T - S2 * ra) * vo2;
                v_alpha += tail * ra;
                S0 += B.z * fac; S1 += B.w * fac; S2 += cb_ * fac;
                if (A.z * vis <= GS_ALPHA_MAX) {
                    const float v_sigma = -A.z * vis * v_alpha;
                    v[0] = v_sigma * (A.w * dx + B.x * dy);
                    v[1] = v_sigma * (B.x * dx + B.y * dy);
                    v[2] = 0.5f * v_sigma * dx * dx;
                    v[3] = v_sigma * dx * dy;
                    v[4] = 0.5f * v_sigma * dy * dy;
                    v[5] = vis * v_alpha;
                }
            }
            const float r8 = warp_reduce8(v, lane);
            const float r9 = warp_sum(v8);
            if ((lane & 3) == 0) my_acc[j * RB_NV + (lane >> 2)] += r8;
            if (lane == 1) my_acc[j * RB_NV + 8] += r9;
        }
        __syncthreads();
        // flush: sum the 8 warp rows, one global reduction per (Gaussian, component)
        for (int idx = tid; idx < cnt * RB_NV; idx += 256) {
            float sum = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                sum += acc[w][idx];
                acc[w][idx] = 0.f;
            }
            if (sum != 0.f) {
                const int j = idx / RB_NV, c = idx - j * RB_NV;
                const int id = sid[j];
                float *dst = c < 2 ? v_xys + 2 * id + c
                           : c < 5 ? v_conics + 3 * id + (c - 2)
                           : c == 5 ? v_opac + id
                                    : v_colors + 3 * id + (c - 6);
                atomicAdd(dst, sum);
            }
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
            sa[tid] = make_float4(xy.x, xy.y, opac[id], conics[3 * id]);
            sb[tid] = make_float2(conics[3 * id + 1], conics[3 * id + 2]);
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
            hit = hit && (lo + j < my_end);
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
                g[0] = v_sigma * (A.w * dx + B.x * dy);
                g[1] = v_sigma * (B.x * dx + B.y * dy);
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
        raster_bwd3_kernel<<<tiles, 256, 0, st>>>(H, W, tiles_x, ids_sorted, bins, xy2, conics, colors, opac, bg,
                                                  final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics,
                                                  v_colors, v_opac);
    else
        raster_bwdN_kernel<<<tiles, 256, 0, st>>>(C, H, W, tiles_x, ids_sorted, bins, xy2, conics, colors, opac,
                                                  bg, final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics,
                                                  v_colors, v_opac);
    GS_LAUNCH_CHECK();
    return 0;
}

// Stage 4 — backward alpha compositing over PACKED splat records (replaces gsplat's
// rasterize_backward reached from /root/reference/gaussctrl/gc_trainer.py:275). Exact derivative of
// raster_packed_fwd.cu; same arithmetic per (pixel, Gaussian) pair as raster_bwd.cu.
//
// What changed against raster_bwd.cu (gather -> barrier -> replay -> barrier -> flush -> barrier, no prefetch:
// ncu r01k: barrier = 31 % of the stall samples):
//   * the 256-Gaussian batch is copied into a DOUBLE-BUFFERED stage with three 16-byte cp.async per record, issued one
//     batch ahead (ids two batches ahead, in a register), so the gather latency sits under the replay;
//   * the nine moment sums per (tile, Gaussian) are double-buffered too: the flush of batch b (shared -> one
//     red.global per component) is done by each thread right AFTER the barrier that ends batch b, while other warps
//     already replay batch b+1 into the other accumulator set. One block barrier per batch instead of two, and the
//     global reductions overlap the replay;
//   * no per-warp survivor list in shared memory: lane l tests record l's alpha >= 1/255 box against the warp's 8x4
//     rect, the ballot word is the survivor set, and the warp replays its set bits back to front (one dependent
//     shared-memory load less on the list -> splat -> math chain).
// Per lane only the raw moments of v_sigma are formed; conic / xy gradients are linear in them and are assembled once
// per (tile, Gaussian) in the flush. Bound: SM issue rate + shuffle throughput; global atomics <= 9 * I.
#define GS_PDL_GROUP 2
#include <cstdlib>

#include "raster_packed.cuh"

namespace {

// CTAs per SM the direct kernel is compiled for: 5 = 48 registers, no spills. 6 (40 registers, two spilled) was measured
// slower, 248 vs 240 us, although its 27 KB of shared memory would allow eight.
#ifndef GS_BWD_MIN_CTAS
#define GS_BWD_MIN_CTAS 5
#endif
constexpr int RPB = 256;   // Gaussians per batch
constexpr int RPB_NV = 9;  // v_x, v_y, v_ca, v_cb, v_cc, v_opac, v_r, v_g, v_b (as moments)

// in: v[0..7] per lane. out: sum over the warp of v[(lane >> 2) & 7] (same value on the 4 lanes that share lane>>2).
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

// DET (debug switch, gs_view_bwd_det): bit-reproducible gradients. The two places where the order of float additions
// depends on scheduling are removed: (1) every warp accumulates its moments into its OWN shared-memory rows (plain
// adds) and the flush sums the eight rows in warp order instead of shared-memory atomics; (2) the flush writes the
// nine gradients of (tile, Gaussian) to row k of `isect` (k = position in ids_sorted, unique) instead of
// red.global; a segmented sum in list order (binning.cu: gs_internal_segment_reduce) then forms the per-Gaussian
// totals. 147 KB of dynamic shared memory (one CTA per SM): a debugging aid, not a fast path.
template <bool DET>
__global__ void __launch_bounds__(256)
raster_bwd_packed_kernel(int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                         const int2 *__restrict__ tile_bins, const float4 *__restrict__ rec,
                         const float *__restrict__ bg, const float *__restrict__ final_T,
                         const int32_t *__restrict__ final_idx, const float *__restrict__ v_out,
                         const float *__restrict__ v_out_alpha, float *__restrict__ v_xys,
                         float *__restrict__ v_conics, float *__restrict__ v_colors, float *__restrict__ v_opac,
                         const int32_t *__restrict__ tile_order, const float *__restrict__ clamp_ref,
                         float *__restrict__ isect) {
    __shared__ float4 s0[2][RPB];          // x, y, ex, ey
    __shared__ float4 s1[2][RPB];          // opac, qa, qb, qc
    __shared__ float4 s2[2][RPB];          // r, g, b, depth
    __shared__ float acc_s[DET ? 1 : 2 * RPB * RPB_NV];   // moment sums of the batch, one row of 9 per Gaussian
    extern __shared__ __align__(16) float acc_d[];        // DET: [2 stages][8 warps][RPB * 9]
    float *const acc_base = DET ? acc_d : acc_s;
    constexpr int ACC_STAGE = (DET ? 8 : 1) * RPB * RPB_NV;
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
    for (int i = tid; i < 2 * ACC_STAGE; i += 256) acc_base[i] = 0.f;
    __syncthreads();
    int bmax = range.x;
#pragma unroll
    for (int w = 0; w < 8; ++w) bmax = max(bmax, s_wmax[w]);
    const int nb = (bmax - range.x + RPB - 1) / RPB;       // batches, walked from the back of the list

    // batch b covers list entries [lo_b, hi_b), hi_b = bmax - b * RPB; thread t handles entry lo_b + t
    auto batch_lo = [&](int b) { return max(range.x, bmax - (b + 1) * RPB); };
    auto batch_cnt = [&](int b) { return (bmax - b * RPB) - batch_lo(b); };
    auto load_id = [&](int b) -> int { return tid < batch_cnt(b) ? ids_sorted[batch_lo(b) + tid] : -1; };
    auto issue = [&](int id, int stage) {
        if (id >= 0) {
            const float4 *src = rec + (size_t)GS_REC_QUADS * (size_t)id;
            gs_cp_async16(gs_smem_addr(&s0[stage][tid]), src);
            gs_cp_async16(gs_smem_addr(&s1[stage][tid]), src + 1);
            gs_cp_async16(gs_smem_addr(&s2[stage][tid]), src + 2);
        }
        gs_cp_async_commit();
    };
    // flush of one finished batch: thread t turns the 9 moment sums of ITS Gaussian into gradients
    //   v_xy = (a mx + b my, b mx + c my),  v_conic = (mxx/2, mxy, myy/2)
    // and issues one global reduction per (tile, Gaussian, component); the sums are cleared for re-use
    auto flush = [&](int id, int stage, int k) {      // k = position of this Gaussian's entry in ids_sorted
        if (id < 0) return;
        float m[RPB_NV];
        float *row = acc_base + stage * ACC_STAGE + tid * RPB_NV;
#pragma unroll
        for (int c = 0; c < RPB_NV; ++c) {
            if (DET) {
                float sum = 0.f;
#pragma unroll
                for (int w = 0; w < 8; ++w) {              // fixed order: warp 0 .. 7
                    sum += row[w * RPB * RPB_NV + c];
                    row[w * RPB * RPB_NV + c] = 0.f;
                }
                m[c] = sum;
            } else {
                m[c] = row[c];
                row[c] = 0.f;
            }
        }
        const float4 Q = s1[stage][tid];
        float ca, cb, cc;
        gs_unscale_conic(Q.y, Q.z, Q.w, ca, cb, cc);
        if (DET) {
            float *o = isect + (size_t)k * RPB_NV;
            o[0] = ca * m[0] + cb * m[1]; o[1] = cb * m[0] + cc * m[1];
            o[2] = 0.5f * m[2]; o[3] = m[3]; o[4] = 0.5f * m[4];
            o[5] = m[5]; o[6] = m[6]; o[7] = m[7]; o[8] = m[8];
            return;
        }
        if (m[0] != 0.f || m[1] != 0.f) {
            atomicAdd(v_xys + 2 * (size_t)id, ca * m[0] + cb * m[1]);
            atomicAdd(v_xys + 2 * (size_t)id + 1, cb * m[0] + cc * m[1]);
        }
        if (m[2] != 0.f) atomicAdd(v_conics + 3 * (size_t)id, 0.5f * m[2]);
        if (m[3] != 0.f) atomicAdd(v_conics + 3 * (size_t)id + 1, m[3]);
        if (m[4] != 0.f) atomicAdd(v_conics + 3 * (size_t)id + 2, 0.5f * m[4]);
        if (m[5] != 0.f) atomicAdd(v_opac + id, m[5]);
        if (m[6] != 0.f) atomicAdd(v_colors + 3 * (size_t)id, m[6]);
        if (m[7] != 0.f) atomicAdd(v_colors + 3 * (size_t)id + 1, m[7]);
        if (m[8] != 0.f) atomicAdd(v_colors + 3 * (size_t)id + 2, m[8]);
    };

    // this lane's accumulator column: lanes 0,4,..,28 own moments 0..7, lane 1 owns the ninth
    const int acc_col = lane == 1 ? 8 : (lane >> 2);
    const bool acc_lane = (lane & 3) == 0 || lane == 1;

    int id_prev = -1, id_cur = -1, id_next = -1;
    if (nb > 0) {
        id_cur = load_id(0);
        issue(id_cur, 0);
        if (nb > 1) id_next = load_id(1);
    }
    for (int b = 0; b < nb; ++b) {
        const int stage = b & 1;
        gs_cp_async_wait_all();
        __syncthreads();      // stage b published; every warp has finished replaying batch b-1
        if (b > 0) flush(id_prev, stage ^ 1, batch_lo(b - 1) + tid);
        if (b + 1 < nb) issue(id_next, stage ^ 1);     // my own slot of the stage I just flushed from
        id_prev = id_cur;
        id_cur = id_next;
        id_next = (b + 2 < nb) ? load_id(b + 2) : -1;

        const int lo = batch_lo(b);
        const int jend = min(batch_cnt(b), wmax - lo);   // warp-uniform: nothing past this warp's last contributor
        const unsigned a0 = gs_smem_addr(s0[stage]), a1 = gs_smem_addr(s1[stage]), a2 = gs_smem_addr(s2[stage]);
        float *accs = acc_base + stage * ACC_STAGE + (DET ? warp * RPB * RPB_NV : 0);
        for (int c32 = (jend - 1) & ~31; jend > 0 && c32 >= 0; c32 -= 32) {
            const int j = c32 + lane;
            bool keep = false;
            if (j < jend) keep = gs_rect_hit(gs_lds128(a0 + (unsigned)j * 16u), rx0, ry0, rx1, ry1);
            unsigned word = __ballot_sync(0xffffffffu, keep);
            // ---- back-to-front replay over the survivors ----
            while (word) {
                const int bit = 31 - __clz(word);
                word &= ~(1u << bit);
                const unsigned jj = (unsigned)(c32 + bit);
                const float4 A = gs_lds128(a0 + jj * 16u);
                const float4 B = gs_lds128(a1 + jj * 16u);
                const float4 Cc = gs_lds128(a2 + jj * 16u);      // up front: not a dependent load behind the vote
                GsSplat sp = {A.x, A.y, B.x, B.y, B.z, B.w};
                float dx, dy, vis, alpha;
                bool hit = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha);
                hit = hit && (lo + (int)jj < my_end);
                if (!__any_sync(0xffffffffu, hit)) continue;
                // branch-free: a lane that does not hit runs the same arithmetic with alpha = 0 (T, Svo unchanged,
                // all nine contributions exactly zero)
                const float al = hit ? alpha : 0.0f;
                const float ra = gs_rcp(1.0f - al);
                T *= ra;
                const float fac = al * T;
                const float cvo = Cc.x * vo0 + Cc.y * vo1 + Cc.z * vo2;
                const float v_alpha = T * cvo - ra * (Svo - tail);
                Svo += fac * cvo;
                const bool live = hit && (B.x * vis <= GS_ALPHA_MAX);      // not clamped at 0.999: d alpha / d sigma != 0
                const float visl = live ? vis : 0.0f;
                const float v_sigma = -B.x * visl * v_alpha;
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
                    if (DET) accs[jj * RPB_NV + acc_col] += add;        // this warp's own row: no other writer
                    else if (add != 0.0f) atomicAdd(&accs[jj * RPB_NV + acc_col], add);
                }
            }
        }
    }
    if (nb > 0) {
        __syncthreads();
        flush(id_prev, (nb - 1) & 1, batch_lo(nb - 1) + tid);
    }
}


// DIRECT variant (default): no shared-memory accumulators and no flush. Measured on the 1 M bench scene, a (tile, Gaussian)
// pair is hit by only ~1.35 of the tile's 8 warps (the alpha >= 1/255 box of a typical splat is smaller than a warp's
// 8x4 rect), so accumulating the warps' sums in shared memory first saved almost no global reductions (1.35 M hitting
// (warp, Gaussian) pairs against ~1.0 M non-empty (tile, Gaussian) pairs) while costing a compare-and-swap loop per
// warp sum, 18 KB of shared memory and a flush pass. Here the lanes that end the butterfly with a sum turn it into
// the gradient component they own (the xy pair needs one more shuffle: v_xy = Q (mx, my) with Q the conic) and issue
// ONE red.global each, straight into the compositor-gradient block.
__global__ void __launch_bounds__(256, GS_BWD_MIN_CTAS)
raster_bwd_direct_kernel(int H, int W, int tiles_x, const int32_t *__restrict__ ids_sorted,
                         const int2 *__restrict__ tile_bins, const float4 *__restrict__ rec,
                         const float *__restrict__ bg, const float *__restrict__ final_T,
                         const int32_t *__restrict__ final_idx, const float *__restrict__ v_out,
                         const float *__restrict__ v_out_alpha, float *__restrict__ v_xys,
                         float *__restrict__ v_conics, float *__restrict__ v_colors, float *__restrict__ v_opac,
                         const int32_t *__restrict__ tile_order, const float *__restrict__ clamp_ref) {
    __shared__ float4 s0[2][RPB];          // x, y, ex, ey
    __shared__ float4 s1[2][RPB];          // opac, qa, qb, qc
    __shared__ float4 s2[2][RPB];          // r, g, b, depth
    __shared__ int s_id[2][RPB];           // Gaussian of every staged record (destination of the reductions)
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
    const float tail = Tf * (voa - (bg[0] * vo0 + bg[1] * vo1 + bg[2] * vo2));
    float Svo = 0.f;

    int wmax = my_end;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    int bmax = range.x;
#pragma unroll
    for (int w = 0; w < 8; ++w) bmax = max(bmax, s_wmax[w]);
    const int nb = (bmax - range.x + RPB - 1) / RPB;       // batches, walked from the back of the list

    auto batch_lo = [&](int b) { return max(range.x, bmax - (b + 1) * RPB); };
    auto batch_cnt = [&](int b) { return (bmax - b * RPB) - batch_lo(b); };
    auto load_id = [&](int b) -> int { return tid < batch_cnt(b) ? ids_sorted[batch_lo(b) + tid] : -1; };
    auto issue = [&](int id, int stage) {
        if (id >= 0) {
            const float4 *src = rec + (size_t)GS_REC_QUADS * (size_t)id;
            gs_cp_async16(gs_smem_addr(&s0[stage][tid]), src);
            gs_cp_async16(gs_smem_addr(&s1[stage][tid]), src + 1);
            gs_cp_async16(gs_smem_addr(&s2[stage][tid]), src + 2);
            s_id[stage][tid] = id;
        }
        gs_cp_async_commit();
    };

    // what this lane does with the sum the butterfly leaves on it (value index lane >> 2; lane 1: the ninth value):
    //   0, 1: v_xy = (a mx + b my, b mx + c my) with (a, b, c) = conic = (K2 qa, K1 qb, K2 qc)   [needs the other sum]
    //   2, 3, 4: v_conic = (mxx / 2, mxy, myy / 2);  5: v_opacity;  6, 7, 8: v_rgb
    const int vi = lane >> 2;
    const bool acc_lane = (lane & 3) == 0 || lane == 1;
    const float K1 = -1.0f / GS_LOG2E, K2 = -2.0f / GS_LOG2E;
    float k2 = vi < 2 ? K2 : 0.f, k1 = vi < 2 ? K1 : 0.f;
    float mulc = vi < 2 ? 0.f : ((vi == 2 || vi == 4) ? 0.5f : 1.0f);
    float *gp;
    unsigned gstride;
    if (lane == 1) { gp = v_colors + 2; gstride = 3; }
    else if (vi < 2) { gp = v_xys + vi; gstride = 2; }
    else if (vi < 5) { gp = v_conics + (vi - 2); gstride = 3; }
    else if (vi == 5) { gp = v_opac; gstride = 1; }
    else { gp = v_colors + (vi - 6); gstride = 3; }
    const bool first = vi == 0;
    // opaque to the compiler: it otherwise re-derives these per-lane constants from the lane id inside the replay
    // loop (8 of ~110 issue slots per hitting (warp, Gaussian) in the first build of this kernel)
    asm volatile("" : "+f"(k1), "+f"(k2), "+f"(mulc), "+r"(gstride));

    int id_cur = -1, id_next = -1;
    if (nb > 0) {
        id_cur = load_id(0);
        issue(id_cur, 0);
        if (nb > 1) id_next = load_id(1);
    }
    for (int b = 0; b < nb; ++b) {
        const int stage = b & 1;
        gs_cp_async_wait_all();
        __syncthreads();      // stage b published; every warp has finished replaying batch b-1
        if (b + 1 < nb) issue(id_next, stage ^ 1);
        id_next = (b + 2 < nb) ? load_id(b + 2) : -1;

        const int lo = batch_lo(b);
        const int jend = min(batch_cnt(b), wmax - lo);   // warp-uniform: nothing past this warp's last contributor
        const unsigned a0 = gs_smem_addr(s0[stage]), a1 = gs_smem_addr(s1[stage]), a2 = gs_smem_addr(s2[stage]);
        const unsigned aid = gs_smem_addr(s_id[stage]);
        for (int c32 = (jend - 1) & ~31; jend > 0 && c32 >= 0; c32 -= 32) {
            const int j = c32 + lane;
            bool keep = false;
            if (j < jend) keep = gs_rect_hit(gs_lds128(a0 + (unsigned)j * 16u), rx0, ry0, rx1, ry1);
            unsigned word = __ballot_sync(0xffffffffu, keep);
            while (word) {
                const int bit = 31 - __clz(word);
                word &= ~(1u << bit);
                const unsigned jj = (unsigned)(c32 + bit);
                const float4 A = gs_lds128(a0 + jj * 16u);
                const float4 B = gs_lds128(a1 + jj * 16u);
                const float4 Cc = gs_lds128(a2 + jj * 16u);
                GsSplat sp = {A.x, A.y, B.x, B.y, B.z, B.w};
                float dx, dy, vis, alpha;
                bool hit = gs_eval_alpha(sp, fx, fy, dx, dy, vis, alpha);
                hit = hit && (lo + (int)jj < my_end);
                if (!__any_sync(0xffffffffu, hit)) continue;
                const int gid = __float_as_int(gs_lds32(aid + jj * 4u));
                const float al = hit ? alpha : 0.0f;
                const float ra = gs_rcp(1.0f - al);
                T *= ra;
                const float fac = al * T;
                const float cvo = Cc.x * vo0 + Cc.y * vo1 + Cc.z * vo2;
                const float v_alpha = T * cvo - ra * (Svo - tail);
                Svo += fac * cvo;
                const bool live = hit && (B.x * vis <= GS_ALPHA_MAX);
                const float visl = live ? vis : 0.0f;
                const float v_sigma = -B.x * visl * v_alpha;
                const float mx = v_sigma * dx, my = v_sigma * dy;
                float v[8];
                v[0] = mx; v[1] = my;
                v[2] = mx * dx; v[3] = mx * dy; v[4] = my * dy;
                v[5] = visl * v_alpha;
                v[6] = fac * vo0; v[7] = fac * vo1;
                const float v8 = fac * vo2;
                const float r8 = warp_reduce8(v, lane);
                const float r9 = warp_sum(v8);
                const float other = __shfl_xor_sync(0xffffffffu, r8, 4);      // lanes of sum 0 see sum 1 and vice versa
                const float coefA = fmaf(k2, first ? B.y : B.w, mulc);
                float add = fmaf(k1 * B.z, other, coefA * r8);
                if (lane == 1) add = r9;
                if (acc_lane && add != 0.0f) atomicAdd(gp + (unsigned)gid * gstride, add);   // < 2^32 floats per array
            }
        }
    }
    gs_cp_async_wait_all();
}

// 1 (default): reductions straight to global memory; 0: shared-memory accumulators + flush. GS_RASTER_BWD_DIRECT /
// gs_set_raster_bwd_direct: A/B switch
int g_bwd_direct = -1;
int raster_bwd_direct() {
    if (g_bwd_direct < 0) {
        const char *e = getenv("GS_RASTER_BWD_DIRECT");
        g_bwd_direct = e ? (atoi(e) != 0) : 1;
    }
    return g_bwd_direct;
}

}  // namespace

extern "C" int gs_set_raster_bwd_direct(int on) {
    const int prev = raster_bwd_direct();
    g_bwd_direct = on != 0;
    return prev;
}

int gs_launch_raster_bwd_packed(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                const int32_t *tile_bins, const float *rec, const float *bg, const float *final_T,
                                const int32_t *final_idx, const float *v_out, const float *v_out_alpha, float *v_xys,
                                float *v_conics, float *v_colors, float *v_opac, const int32_t *tile_order,
                                const float *clamp_ref, cudaStream_t st) {
    const int tiles = tiles_x * tiles_y;
    if (raster_bwd_direct()) {
        GS_PDL_LAUNCH_IF(tiles <= GS_RASTER_PDL_MAX_TILES, raster_bwd_direct_kernel, tiles, 256, 0, st, H, W, tiles_x,
                         ids_sorted, reinterpret_cast<const int2 *>(tile_bins), reinterpret_cast<const float4 *>(rec), bg,
                         final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics, v_colors, v_opac, tile_order, clamp_ref);
        return 0;
    }
    GS_PDL_LAUNCH_IF(tiles <= GS_RASTER_PDL_MAX_TILES, raster_bwd_packed_kernel<false>, tiles, 256, 0, st, H, W, tiles_x,
                     ids_sorted, reinterpret_cast<const int2 *>(tile_bins), reinterpret_cast<const float4 *>(rec), bg,
                     final_T, final_idx, v_out, v_out_alpha, v_xys, v_conics, v_colors, v_opac, tile_order, clamp_ref,
                     (float *)nullptr);
    return 0;
}

// deterministic variant: per-intersection gradient rows into isect (I x 9 floats, zero-filled by the caller)
int gs_launch_raster_bwd_packed_det(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                    const int32_t *tile_bins, const float *rec, const float *bg, const float *final_T,
                                    const int32_t *final_idx, const float *v_out, const float *v_out_alpha,
                                    const int32_t *tile_order, const float *clamp_ref, float *isect, cudaStream_t st) {
    const int tiles = tiles_x * tiles_y;
    const int smem = 2 * 8 * RPB * RPB_NV * (int)sizeof(float);
    GS_CUDA(cudaFuncSetAttribute(raster_bwd_packed_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    raster_bwd_packed_kernel<true><<<tiles, 256, smem, st>>>(
        H, W, tiles_x, ids_sorted, reinterpret_cast<const int2 *>(tile_bins), reinterpret_cast<const float4 *>(rec), bg, final_T,
        final_idx, v_out, v_out_alpha, nullptr, nullptr, nullptr, nullptr, tile_order, clamp_ref, isect);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_raster_bwd_packed(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                    const int32_t *tile_bins, const float *rec, const float *bg, const float *final_T,
                                    const int32_t *final_idx, const float *v_out, const float *v_out_alpha, float *v_xys,
                                    float *v_conics, float *v_colors, float *v_opac, const int32_t *tile_order,
                                    const float *clamp_ref, void *stream) {
    if (H <= 0 || W <= 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !rec || !bg || !final_T || !final_idx || !v_out || !v_xys || !v_conics || !v_colors || !v_opac)
        return GS_ERR_BAD_ARG;
    if (tiles_x * GS_TILE < W || tiles_y * GS_TILE < H) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(rec)) return GS_ERR_MISALIGNED;
    return gs_launch_raster_bwd_packed(H, W, tiles_x, tiles_y, ids_sorted, tile_bins, rec, bg, final_T, final_idx, v_out,
                                       v_out_alpha, v_xys, v_conics, v_colors, v_opac, tile_order, clamp_ref,
                                       (cudaStream_t)stream);
}

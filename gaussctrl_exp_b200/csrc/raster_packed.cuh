// Packed splat records and the cp.async staging shared by the packed compositing kernels
// (raster_packed_fwd.cu / raster_packed_bwd.cu).
//
// One record = everything the compositor needs about one projected Gaussian of one view, 48 bytes
// = three 16-byte quads, written ONCE per (view, Gaussian) by the fused preprocess (or by gs_pack_splats
// for the operator API):
//     q0 = (x, y, ex, ey)        pixel centre; half-extents of the alpha >= 1/255 region (gs_alpha_extent)
//     q1 = (opac, qa, qb, qc)    opacity in (0,1); conic pre-scaled for exp2 (gs_prescale_conic)
//     q2 = (r, g, b, depth)      post-activation colour; view-space depth (RGB-D pass)
// The un-packed kernels (raster_fwd.cu / raster_bwd.cu) gather nine scalars from five arrays per
// (tile, Gaussian) and recompute the pre-scaled conic and the extents there — once per INTERSECTION
// (~3 M per view) instead of once per Gaussian. Packed, a (tile, Gaussian) gather is three 16-byte
// cp.async (LDGSTS) straight into the shared-memory stage: two 32-byte sectors per record instead of
// five to six, no registers held across the batch, and the copy of batch b+1 runs under the compositing of
// batch b (north_star stage 3: "staged into shared memory with TMA/cp.async double-buffering"; a gather by id
// is not a dense box, so it is cp.async rather than a TMA tensor copy — SURVEY.md §7 "sm_100a specifics").
#pragma once
#include "raster_common.cuh"

#define GS_REC_FLOATS 12
#define GS_REC_QUADS 3

__device__ __forceinline__ void gs_cp_async16(unsigned smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void gs_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void gs_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// builds the three quads of one record (host-independent: used by the preprocess kernel and by gs_pack_splats)
__device__ __forceinline__ void gs_make_record(float x, float y, float ca, float cb, float cc, float opac, float r, float g,
                                               float b, float depth, float4 &q0, float4 &q1, float4 &q2) {
    float qa, qb, qc;
    gs_prescale_conic(ca, cb, cc, qa, qb, qc);
    const float2 e = gs_alpha_extent(opac, ca, cb, cc);
    q0 = make_float4(x, y, e.x, e.y);
    q1 = make_float4(opac, qa, qb, qc);
    q2 = make_float4(r, g, b, depth);
}

// does the splat's alpha >= 1/255 box touch the rect [rx0, rx1] x [ry0, ry1]?
__device__ __forceinline__ bool gs_rect_hit(const float4 &q0, float rx0, float ry0, float rx1, float ry1) {
    return (q0.x + q0.z >= rx0) && (q0.x - q0.z <= rx1) && (q0.y + q0.w >= ry0) && (q0.y - q0.w <= ry1);
}

// internal launchers (view.cu chains them; signatures mirror the extern "C" entry points)
int gs_launch_raster_fwd_packed(int views, int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                const int32_t *tile_bins, const float *rec, const float *bg, float *out_img,
                                float *out_depth, float *final_T, int32_t *final_idx, const int32_t *tile_order,
                                int32_t *tile_work, float *out_clamped, float *out_alpha, const int32_t *total_dev,
                                int64_t I_cap, cudaStream_t st);
int gs_launch_raster_bwd_packed(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                const int32_t *tile_bins, const float *rec, const float *bg, const float *final_T,
                                const int32_t *final_idx, const float *v_out, const float *v_out_alpha, float *v_xys,
                                float *v_conics, float *v_colors, float *v_opac, const int32_t *tile_order,
                                const float *clamp_ref, cudaStream_t st);
int gs_launch_raster_bwd_packed_det(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                                    const int32_t *tile_bins, const float *rec, const float *bg, const float *final_T,
                                    const int32_t *final_idx, const float *v_out, const float *v_out_alpha,
                                    const int32_t *tile_order, const float *clamp_ref, float *isect, cudaStream_t st);

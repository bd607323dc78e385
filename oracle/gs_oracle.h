/*
 * gs_oracle.h — CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the gsplat-0.1.2 operator semantics that
 * /root/reference/gaussctrl/gc_model.py:174-236 calls (project_gaussians,
 * spherical_harmonics, rasterize_gaussians) and of their exact derivatives.
 * The arithmetic itself lives in the third-party package gsplat==0.1.2.1
 * (reference README.md:58), which is NOT in /root/reference and not
 * installable here; the algorithm is restated from SURVEY.md Appendix A.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or renders for
 * this path. The only pieces pinned to reference code are
 * the caller-side matrices (tests/golden/, generated from gc_model.py:39-57
 * and ad_render.py:49-67).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 *
 * All arrays are fp32 / int32, row-major, caller-allocated. Every function is
 * deterministic; OpenMP is used only across independent Gaussians / tiles.
 */
#ifndef GS_ORACLE_H
#define GS_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* A.1 projection forward. Outputs must be zero-filled by the caller or not:
 * culled rows are written as zeros here. */
void orc_project_fwd(int N, const float *means, const float *scales, float glob_scale,
                     const float *quats, const float *viewmat /*>=12*/, const float *projmat /*16*/,
                     float fx, float fy, float cx, float cy, int H, int W,
                     int tiles_x, int tiles_y, float clip,
                     float *cov3d, float *xys, float *depths, int32_t *radii,
                     float *conics, int32_t *tiles_hit);

/* exact derivative of orc_project_fwd w.r.t. means, scales, quats */
void orc_project_bwd(int N, const float *means, const float *scales, float glob_scale,
                     const float *quats, const float *viewmat, const float *projmat,
                     float fx, float fy, float cx, float cy, int H, int W,
                     const int32_t *radii,
                     const float *v_xys, const float *v_depths, const float *v_conics,
                     float *v_means, float *v_scales, float *v_quats);

/* A.2 spherical harmonics. K = number of stored bases per Gaussian (>= (deg+1)^2) */
void orc_sh_fwd(int N, int K, int deg, const float *dirs, const float *coeffs, float *colors);
void orc_sh_bwd(int N, int K, int deg, const float *dirs, const float *v_colors, float *v_coeffs);

/* A.3 binning. cum[N] inclusive scan; returns I (total intersections). */
int64_t orc_scan(int N, const int32_t *tiles_hit, int32_t *cum);
void orc_emit(int N, const float *xys, const float *depths, const int32_t *radii,
              const int32_t *cum, int tiles_x, int tiles_y, int64_t *keys, int32_t *ids);
/* stable ascending sort of (keys, ids) */
void orc_sort(int64_t I, const int64_t *keys, const int32_t *ids, int64_t *keys_sorted,
              int32_t *ids_sorted);
void orc_tile_bins(int64_t I, int T, const int64_t *keys_sorted, int32_t *tile_bins /*T*2*/);

/* A.4 compositing forward, C channels. final_idx = 1 + global index (into
 * ids_sorted) of the last composited Gaussian, or the tile's range start when
 * none. Returns P = number of (pixel, Gaussian) pairs visited. */
int64_t orc_raster_fwd(int C, int H, int W, int tiles_x, int tiles_y,
                       const int32_t *ids_sorted, const int32_t *tile_bins,
                       const float *xys, const float *conics, const float *colors,
                       const float *opac, const float *bg,
                       float *out_img, float *final_T, int32_t *final_idx);

/* A.5 compositing backward (exact derivative of A.4). Outputs zero-filled here. */
void orc_raster_bwd(int C, int H, int W, int tiles_x, int tiles_y, int N,
                    const int32_t *ids_sorted, const int32_t *tile_bins,
                    const float *xys, const float *conics, const float *colors,
                    const float *opac, const float *bg,
                    const float *final_T, const int32_t *final_idx,
                    const float *v_out, const float *v_out_alpha /* may be NULL */,
                    float *v_xys, float *v_conics, float *v_colors, float *v_opac);

/* Test diagnostics (tests/test_parity_scale_gpu.py): per-pixel smallest relative distance of a forward
 * decision to its threshold (alpha vs 1/255, T(1-alpha) vs 1e-4, raw alpha vs 0.999), and the set of
 * Gaussians that can influence a masked pixel. */
void orc_raster_risk(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                     const int32_t *tile_bins, const float *xys, const float *conics,
                     const float *opac, float *risk_alpha, float *risk_T, float *risk_clamp);
void orc_mark_contributors(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                           const int32_t *tile_bins, const float *xys, const float *conics,
                           const float *opac, const uint8_t *mask, const int32_t *upto, uint8_t *touched);

/* Multi-view noise mask (ad_render.py:273-333 + shaders :384-479), drawn in order on the CPU. */
void orc_noise_mask(int P, const float *points, int V, const float *verts, int F, const int32_t *tris,
                    const float *view16, const float *proj16, int H, int W, const float *depth, float threshold,
                    uint8_t *out_rgba /*H*W*4*/);

int orc_num_threads(void);
void orc_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
#endif

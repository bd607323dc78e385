/*
 * gs_b200.h — C ABI of the B200-native 3D Gaussian splatting render path.
 *
 * Drop-in boundary for the gsplat-0.1.2 operators that the reference calls at
 *   /root/reference/gaussctrl/gc_model.py:174-188  project_gaussians
 *   /root/reference/gaussctrl/gc_model.py:199-200  spherical_harmonics
 *   /root/reference/gaussctrl/gc_model.py:208-220  rasterize_gaussians (RGB + alpha)
 *   /root/reference/gaussctrl/gc_model.py:225-236  rasterize_gaussians (depth)
 *   /root/reference/gaussctrl/gc_trainer.py:275    loss.backward() through all of the above
 * (imports at gc_model.py:32-36). In gsplat these are bound through a pybind11
 * torch extension (`gsplat.cuda._C`); here they are plain `extern "C"` entry
 * points taking raw DEVICE pointers, sizes and a `cudaStream_t` (passed as
 * void*), so that any host language can bind them (INTEGRATION.md shows the
 * ctypes binding the Python wrappers use).
 *
 * Conventions
 *   - all tensors fp32 / int32 / int64, row-major, contiguous, device memory;
 *     pointers must be 16-byte aligned (torch allocations are).
 *   - every function enqueues work on `stream` and returns immediately;
 *     nothing here synchronises the device except where stated.
 *   - return value: 0 = ok, >0 = cudaError_t, <0 = GS_ERR_* below.
 *   - the library never allocates persistent device memory: scratch is passed
 *     in by pointer + size (sizes from gs_*_scratch_bytes).
 */
#ifndef GS_B200_H
#define GS_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_ERR_BAD_ARG        (-1)
#define GS_ERR_SCRATCH_SMALL  (-2)
#define GS_ERR_UNSUPPORTED    (-3)
#define GS_ERR_MISALIGNED     (-4)

#define GS_TILE 16

/* ABI version (major*100 + minor) and the SM arch the kernels were built for. */
int gs_version(void);
int gs_built_arch(void);
/* count of kernel launches issued by this library since load (bench.py's gpu_launches). */
int64_t gs_launch_count(void);

/* One camera, by value, HOST memory (gc_model.py:117-155 builds exactly these numbers):
 * viewmat = rows 0-2 of the y/z-flipped world->camera matrix, projmat = P @ viewmat,
 * cam_pos = camera_to_worlds[:3,3]. */
typedef struct GsCamera {
    float viewmat[12];
    float projmat[16];
    float cam_pos[3];
    float fx, fy, cx, cy;
    float glob_scale;   /* 1 in the reference (gc_model.py:177) */
    float clip;         /* near clip, 0.01 */
    int32_t H, W, tiles_x, tiles_y;
} GsCamera;

/* ---- stage 1: projection (gc_model.py:174-188) ------------------------- */
/* viewmat: >=12 floats (3x4 row-major), projmat: 16 floats (P@V); both DEVICE pointers.
 * Outputs are fully written (culled rows = 0). */
int gs_project_fwd(int N, const float *means, const float *scales, float glob_scale,
                   const float *quats, const float *viewmat, const float *projmat,
                   float fx, float fy, float cx, float cy, int H, int W,
                   int tiles_x, int tiles_y, float clip_thresh,
                   float *cov3d, float *xys, float *depths, int32_t *radii,
                   float *conics, int32_t *tiles_hit, void *stream);

/* exact vjp of gs_project_fwd w.r.t. means, scales, quats. v_* inputs may not be NULL. */
int gs_project_bwd(int N, const float *means, const float *scales, float glob_scale,
                   const float *quats, const float *viewmat, const float *projmat,
                   float fx, float fy, int H, int W, const int32_t *radii,
                   const float *v_xys, const float *v_depths, const float *v_conics,
                   float *v_means, float *v_scales, float *v_quats, void *stream);

/* ---- stage 1b: spherical harmonics (gc_model.py:199-200; App. A.2) ------------------------- */
/* coeffs (N,K,3); deg <= 4 and (deg+1)^2 <= K. */
int gs_sh_fwd(int N, int K, int deg, const float *dirs, const float *coeffs, float *colors,
              void *stream);
/* v_coeffs (N,K,3) fully written (unused bases = 0). */
int gs_sh_bwd(int N, int K, int deg, const float *dirs, const float *v_colors, float *v_coeffs,
              void *stream);

/* ---- stage 1, fused: pre-activations + projection + SH in one pass over the leaf parameters -----
 * Replaces, for one view, gc_model.py:172-203 and :215: cat(features_dc, features_rest),
 * exp(scales), quats/|quats|, project_gaussians, normalise(means - cam_pos), spherical_harmonics,
 * clamp(+0.5, min 0) [use_sh != 0] or sigmoid(features_dc) [use_sh == 0], sigmoid(opacities).
 * Inputs are the RAW splatfacto leaves: log-scales, un-normalised quats (wxyz), opacity logits,
 * features_dc (N,3), features_rest (N,15,3). Outputs: xys (N,2), depths, radii, conics (N,3),
 * tiles_hit as gs_project_fwd; rgbs (N,3) post-activation colours; opac (N) in (0,1).
 * sh_deg <= 3 = degrees actually used (min(step // interval, sh_degree)). */
int gs_preprocess_fwd(int N, const float *means, const float *log_scales, const float *quats,
                      const float *opac_logits, const float *feat_dc, const float *feat_rest,
                      int use_sh, int sh_deg, const GsCamera *cam_host,
                      float *xys, float *depths, int32_t *radii, float *conics, int32_t *tiles_hit,
                      float *rgbs, float *opac, void *stream);
/* The same kernel in two parts: part 1 = geometry (projection + opacity: xys, depths, radii, conics, tiles_hit,
 * opac — all the binning needs), part 2 = colour (rgbs; reads the radii part 1 wrote). Results are bit-identical to
 * gs_preprocess_fwd. The host enqueues part 2 on a second stream so that its 216 B/Gaussian of SH traffic runs
 * underneath the latency-bound binning chain. Pointers a part does not use may be NULL. */
int gs_preprocess_fwd_part(int part, int N, const float *means, const float *log_scales, const float *quats,
                           const float *opac_logits, const float *feat_dc, const float *feat_rest, int use_sh,
                           int sh_deg, const GsCamera *cam_host, float *xys, float *depths, int32_t *radii,
                           float *conics, int32_t *tiles_hit, float *rgbs, float *opac, void *stream);
/* exact vjp of gs_preprocess_fwd: compositor gradients -> the six leaf gradients (all fully
 * written; culled Gaussians get zeros). rgbs = the forward's colour output (activation masks). */
int gs_preprocess_bwd(int N, const float *means, const float *log_scales, const float *quats,
                      const float *opac_logits, int use_sh, int sh_deg, const GsCamera *cam_host,
                      const int32_t *radii, const float *rgbs, const float *v_xys, const float *v_conics,
                      const float *v_rgbs, const float *v_opac, float *v_means, float *v_log_scales,
                      float *v_quats, float *v_opac_logits, float *v_feat_dc, float *v_feat_rest,
                      void *stream);

/* Data-parallel fine-tune step without a 236 B/Gaussian all-reduce: a "view block" holds
 * the 9 compositor-level gradients of one view (v_xys 2Np | v_conics 3Np | v_rgbs 3Np | v_opac Np, Np = N
 * rounded up to 4) followed by the 64-float camera slot = gs_view_block_floats(N) floats. It is the
 * buffer gs_raster_bwd accumulates into; gs_finish_view_block applies the colour activation's derivative
 * to v_rgbs in place and appends the camera. Ranks all-gather their blocks (36 B/Gaussian) and each runs
 * gs_preprocess_bwd_multi over all V blocks: leaf gradients = scale * sum over views. */
int gs_view_block_floats(int N);
int gs_finish_view_block(int N, int use_sh, const GsCamera *cam_host, const float *rgbs, float *block,
                         void *stream);
int gs_preprocess_bwd_multi(int N, int V, const float *views, int64_t stride_floats, float scale,
                            const float *means, const float *log_scales, const float *quats,
                            const float *opac_logits, int use_sh, int sh_deg, float *v_means,
                            float *v_log_scales, float *v_quats, float *v_opac_logits, float *v_feat_dc,
                            float *v_feat_rest, void *stream);

/* Same, with the collective fused into the kernel: peer_base_ptrs_dev = DEVICE array of V pointers, entry r =
 * this process's mapping of rank r's symmetric-memory buffer (NVLink peer memory); rank r's view block starts
 * offset_floats into it. The kernel reads the peers' blocks in place (software-prefetched peer loads), so no
 * gathered copy exists and the transfer overlaps the arithmetic. The caller provides the cross-rank barrier
 * between the writers (gs_raster_bwd + gs_finish_view_block) and this call. V <= 16. */
int gs_preprocess_bwd_multi_peer(int N, int V, const uint64_t *peer_base_ptrs_dev, int64_t offset_floats,
                                 float scale, const float *means, const float *log_scales,
                                 const float *quats, const float *opac_logits, int use_sh, int sh_deg,
                                 float *v_means, float *v_log_scales, float *v_quats,
                                 float *v_opac_logits, float *v_feat_dc, float *v_feat_rest, void *stream);

/* Sparse form of the same exchange: only ~1/5 of the Gaussians get a non-zero gradient from one view, so the
 * sender compacts its finished dense view block into an exchange buffer
 *   [{bitmap word, prefix} x W | camera 64 | packed v_xy 2Np | v_conic 3Np | v_rgb 3Np | v_opacity Np]   (floats/words)
 * (gs_view_xbuf_floats(N) floats, W = ceil(N/32) rounded up to 4) and the receivers index it with
 * bitmap + popcount. offset_floats = start of the rank's buffer inside its symmetric allocation. */
int gs_view_xbuf_floats(int N);
size_t gs_view_pack_scratch_bytes(int N);
int gs_pack_view_block(int N, const float *block, float *xbuf, void *scratch, size_t scratch_bytes, void *stream);
/* PUSH form: the block is compacted into the caller's own slot xbuf_self (== peer_base_ptrs[self] +
 * offset_floats, local memory) and the used part of it is then copied into the same slot of every other peer
 * (posted NVLink stores, 16 B per lane). After the device barrier each rank runs
 * gs_preprocess_bwd_multi_peer_sparse over its LOCAL slots (pointer array of local addresses). */
int gs_pack_view_block_push(int N, const float *block, float *xbuf_self, int n_peers, int self,
                            const uint64_t *peer_base_ptrs_dev, int64_t offset_floats, void *scratch,
                            size_t scratch_bytes, void *stream);
int gs_preprocess_bwd_multi_peer_sparse(int N, int V, const uint64_t *peer_base_ptrs_dev, int64_t offset_floats,
                                        float scale, const float *means, const float *log_scales,
                                        const float *quats, const float *opac_logits, int use_sh, int sh_deg,
                                        float *v_means, float *v_log_scales, float *v_quats,
                                        float *v_opac_logits, float *v_feat_dc, float *v_feat_rest, void *stream);

/* Owner-sliced form of the exchange: rank d owns the Gaussians [d * S, min((d + 1) * S, N)), S = gs_view_slice_size(N,
 * world) (ceil(N / world) rounded up to 128), for the leaf backward AND for the optimiser step that follows
 * (gc_trainer.py:276-298 steps the optimisers of gc_config.py:58-87 right after backward). Every rank pushes each peer
 * only the part of its compacted view block that covers the peer's slice (gs_pack_view_block_push_slices: ~1/world of
 * the block per peer), and after the device barrier runs gs_leaf_bwd_slice over ITS slice and all `V` views: slice-shaped
 * outputs (row 0 = Gaussian g_lo) holding scale * sum over the views. Work and NVLink bytes per rank are then
 * independent of the number of ranks. */
int gs_view_slice_size(int N, int world);
int gs_pack_view_block_push_slices(int N, const float *block, float *xbuf_self, int n_peers, int self,
                                   const uint64_t *peer_base_ptrs_dev, int64_t offset_floats, void *scratch,
                                   size_t scratch_bytes, void *stream);
int gs_leaf_bwd_slice(int N, int g_lo, int g_hi, int V, const uint64_t *peer_base_ptrs_dev, int64_t offset_floats,
                      float scale, const float *means, const float *log_scales, const float *quats,
                      const float *opac_logits, int use_sh, int sh_deg, float *v_means, float *v_log_scales,
                      float *v_quats, float *v_opac_logits, float *v_feat_dc, float *v_feat_rest, void *stream);

/* Optimiser step of the owner-sliced mode fused with the parameter exchange: Adam (torch.optim.Adam's rule; the step
 * gc_trainer.py:276-298 takes after backward with the learning rates of gc_config.py:58-87) over `count` consecutive fp32
 * elements of one parameter tensor (this rank's rows), each updated value also stored into every peer's copy of the
 * tensor over NVLink (symmetric memory; peer_base_ptrs_dev NULL or n_peers <= 1: local only). step_size = lr / (1 - b1^t),
 * inv_sqrt_c2 = 1 / sqrt(1 - b2^t); offset_floats = position of p[0] inside the symmetric buffer. */
int gs_adam_slice(int64_t count, float *p, const float *g, float *m, float *v, float step_size, float b1, float b2,
                  float eps, float inv_sqrt_c2, int n_peers, int self, const uint64_t *peer_base_ptrs_dev,
                  int64_t offset_floats, void *stream);

/* ---- stage 2: tile binning (inside rasterize_gaussians; App. A.3) -------------------------- */
size_t gs_scan_scratch_bytes(int N);
/* inclusive int32 scan of tiles_hit -> cum; the total is also written to *total_dev
 * (device int32) so the caller can copy it to pinned memory without reading cum[N-1]. */
int gs_bin_scan(int N, const int32_t *tiles_hit, int32_t *cum, int32_t *total_dev,
                void *scratch, size_t scratch_bytes, void *stream);
/* gs_bin_scan without a scratch argument — NO COUNTERPART IN THE REFERENCE (gsplat calls torch.cumsum, which allocates
 * through torch); it exists for hosts of this ABI that cannot give the library device memory of their own: plug-ins that
 * only ever receive their input and output buffers from the application (custom-op slots of an inference server,
 * cgo / JNI bindings whose runtime owns all device memory), and callers sitting at a point where cudaMalloc is not
 * allowed because it synchronises the device or would break a stream capture in progress. The look-back state lives
 * in an array inside the library image instead: GS_SCAN_STATIC_BYTES of device memory on every GPU the library is
 * loaded on, kept until the process exits. Limit: N <= GS_SCAN_STATIC_MAX_N, larger N returns GS_ERR_UNSUPPORTED
 * (use gs_bin_scan). The call allocates and frees nothing. Same outputs as gs_bin_scan, bit for bit.
 * The array is shared by every call on a device: each call clears the part it uses on `stream` first, so calls ordered
 * on the GPU are fine, calls that may overlap on the GPU (unordered streams, concurrent host threads) race. The first
 * call made with a given device current resolves the array's address on that device (cudaGetSymbolAddress) and caches
 * it; issue that first call outside any stream capture, with the device of `stream` current. */
#define GS_SCAN_STATIC_MAX_N (16 * 1024 * 1024)
#define GS_SCAN_STATIC_BYTES (16 + 8 * (GS_SCAN_STATIC_MAX_N / 2048 + 1))
int gs_bin_scan_static(int N, const int32_t *tiles_hit, int32_t *cum, int32_t *total_dev, void *stream);
/* 64-bit keys (tile_id << 32 | depth bits) + Gaussian ids, in emission order. */
int gs_bin_emit(int N, const float *xys, const float *depths, const int32_t *radii,
                const int32_t *cum, int tiles_x, int tiles_y, int64_t *keys, int32_t *ids,
                void *stream);
size_t gs_sort_scratch_bytes(int64_t I);
/* stable ascending radix sort over the live key bits [0, 32 + ceil(log2(T))). */
int gs_bin_sort(int64_t I, int T, const int64_t *keys, const int32_t *ids,
                int64_t *keys_sorted, int32_t *ids_sorted,
                void *scratch, size_t scratch_bytes, void *stream);
/* tile_bins (T,2) int32 = [start,end) per tile; fully written (empty tiles = 0,0). */
int gs_bin_edges(int64_t I, int T, const int64_t *keys_sorted, int32_t *tile_bins, void *stream);

/* Traffic-lean binning with the SAME outputs as scan+emit+sort+edges (ids_sorted, tile_bins)
 * but without materialising 64-bit keys: depth-sort the Gaussians once, emit in that order,
 * one stable partition by tile id. I_cap = capacity of ids_sorted; if the true I exceeds it,
 * nothing out of bounds is written and *total_dev (> I_cap) tells the caller to retry. */
size_t gs_bin_fast_scratch_bytes(int N, int64_t I_cap, int T);
int gs_bin_fast(int N, int64_t I_cap, const float *xys, const float *depths, const int32_t *radii,
                const int32_t *tiles_hit, int tiles_x, int tiles_y,
                int32_t *ids_sorted, int32_t *tile_bins, int32_t *total_dev,
                void *scratch, size_t scratch_bytes, void *stream);
/* The gs_set_* knobs below are PROCESS-GLOBAL and not thread-safe: they exist for A/B measurements (every setting
 * produces identical output), are meant to be set once before the first render, and must not be flipped while another
 * host thread is inside a gs_* call. A default run never needs them. */
/* Tuning knob of gs_bin_fast's depth sort: 4 (default; 8-bit digits, 12 launches) or 3 (11/11/10-bit digits, 9
 * launches; measured slower) stable LSD passes — identical output. Returns the previous setting; other values only query.
 * Initial value: environment variable GS_DEPTH_SORT_PASSES. */
int gs_set_depth_sort_passes(int passes);
/* CTA shape of gs_bin_fast's radix passes: which = 0 (depth sort over the N elements) or 1 (partition over the I pairs);
 * cfg = -1 (default: chosen from the problem size), 0 = 256 threads x 16 keys, 1 = 512 x 8, 2 = 384 x 8, 3 = 256 x 12.
 * Identical output. Returns the previous setting (-1 = automatic); other values only query. Initial value: environment
 * variables GS_RADIX_CFG_DEPTH / GS_RADIX_CFG_PART. */
int gs_set_radix_cfg(int which, int cfg);
/* Implementation behind gs_bin_fast / gs_view_fwd: 1 (default) = LSD pipeline — depth sort of the Gaussians, scan +
 * emission in depth order, stable partition by tile; 0 = bucket pipeline — per-tile counts, scan, unordered fill of
 * (depth bits, id) entries into each tile's segment, one CTA per tile merge-sorts its segment in shared memory (four
 * launches instead of sixteen, but measured slower: the per-tile sort is instruction-bound). Identical output. Returns
 * the previous setting; other values only query. Initial value: environment variable GS_BIN_IMPL=bucket. */
int gs_set_bin_impl(int impl);

/* ---- stage 3: forward compositing (App. A.4) ----------------------------------------------- */
/* colors (N,C); opac (N); bg (C) device; out_img (H,W,C); final_T (H,W); final_idx (H,W) =
 * 1 + index into ids_sorted of the last composited Gaussian (tile range start if none). */
int gs_raster_fwd(int C, int H, int W, int tiles_x, int tiles_y,
                  const int32_t *ids_sorted, const int32_t *tile_bins,
                  const float *xys, const float *conics, const float *colors, const float *opac,
                  const float *bg, float *out_img, float *final_T, int32_t *final_idx,
                  void *stream);

/* RGB + depth in ONE pass over the same tile lists (the eval path of gc_model.py:208-238 composites twice
 * and gsplat re-bins and re-sorts for the second call). out_depth (H,W) = sum_k alpha_k T_k depth_k with a
 * zero background, i.e. channel 0 of the reference's second rasterize call before its division by alpha. */
int gs_raster_fwd_rgbd(int H, int W, int tiles_x, int tiles_y,
                       const int32_t *ids_sorted, const int32_t *tile_bins,
                       const float *xys, const float *conics, const float *colors, const float *opac,
                       const float *depths, const float *bg, float *out_img, float *out_depth,
                       float *final_T, int32_t *final_idx, void *stream);

/* Tile scheduling (3-channel kernels): CTAs are dispatched in blockIdx order, so launching the tiles in
 * descending-work order (LPT) keeps the long tiles out of the tail. gs_tile_order ranks T tiles by
 * work[t] (work_is_range = 0) or by tile_bins[t].end - tile_bins[t].start (work_is_range = 1), ties by
 * tile id; above 2 048 tiles the ranking is by 1 024 work levels relative to the heaviest tile (heaviest level
 * first, order inside a level unspecified) — `order` is a permutation of 0..T-1 either way. The *_ordered variants take the permutation (NULL = natural order); the forward also writes
 * tile_work[t] = depth of the deepest contributor of tile t (NULL = skip), which is the backward's work.
 * depths/out_depth NULL = RGB only. Results are identical to the un-ordered entry points. */
int gs_tile_order(int T, const int32_t *work, int work_is_range, int32_t *order, void *stream);
int gs_raster_fwd_ordered(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                          const int32_t *tile_bins, const float *xys, const float *conics,
                          const float *colors, const float *opac, const float *depths, const float *bg,
                          float *out_img, float *out_depth, float *final_T, int32_t *final_idx,
                          const int32_t *tile_order, int32_t *tile_work, void *stream);
int gs_raster_bwd_ordered(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                          const int32_t *tile_bins, const float *xys, const float *conics,
                          const float *colors, const float *opac, const float *bg, const float *final_T,
                          const int32_t *final_idx, const float *v_out, const float *v_out_alpha,
                          float *v_xys, float *v_conics, float *v_colors, float *v_opac,
                          const int32_t *tile_order, void *stream);

/* The same two passes with the caller's per-pixel post-ops folded in (gc_model.py:221-222: alpha = 1 - T and
 * rgb = clamp(rgb, max=1)): out_clamped (H,W,3) and out_alpha (H,W) are extra outputs (either may be NULL;
 * out_img keeps the raw colour); clamp_ref = the raw out_img of the forward: v_out is then taken as the
 * gradient w.r.t. the CLAMPED colour, i.e. zeroed where clamp_ref > 1 (torch.clamp's backward). */
int gs_raster_fwd_ordered_post(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                               const int32_t *tile_bins, const float *xys, const float *conics,
                               const float *colors, const float *opac, const float *depths, const float *bg,
                               float *out_img, float *out_depth, float *final_T, int32_t *final_idx,
                               const int32_t *tile_order, int32_t *tile_work, float *out_clamped,
                               float *out_alpha, void *stream);
int gs_raster_bwd_ordered_post(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                               const int32_t *tile_bins, const float *xys, const float *conics,
                               const float *colors, const float *opac, const float *bg, const float *final_T,
                               const int32_t *final_idx, const float *v_out, const float *v_out_alpha,
                               float *v_xys, float *v_conics, float *v_colors, float *v_opac,
                               const int32_t *tile_order, const float *clamp_ref, void *stream);

/* ---- stage 4: backward compositing (App. A.5) ---------------------------------------------- */
/* v_out (H,W,C); v_out_alpha (H,W) or NULL. Outputs are ACCUMULATED INTO: caller zero-fills
 * v_xys (N,2), v_conics (N,3), v_colors (N,C), v_opac (N). */
int gs_raster_bwd(int C, int H, int W, int tiles_x, int tiles_y,
                  const int32_t *ids_sorted, const int32_t *tile_bins,
                  const float *xys, const float *conics, const float *colors, const float *opac,
                  const float *bg, const float *final_T, const int32_t *final_idx,
                  const float *v_out, const float *v_out_alpha,
                  float *v_xys, float *v_conics, float *v_colors, float *v_opac, void *stream);

/* ---- packed splat records: what the compositing kernels stage with cp.async ------------------------------------
 * One 48-byte record per projected Gaussian per view (12 floats):
 *   [0..3] x, y, ex, ey      pixel centre (xys), half-extents of the region where alpha can reach 1/255
 *   [4..7] opac, qa, qb, qc  opacity; conic scaled for exp2: qa = -a/2 log2e, qb = -b log2e, qc = -c/2 log2e
 *   [8..11] r, g, b, depth   colour handed to the compositor; view-space depth
 * gs_pack_splats builds them from the operator-API tensors (what rasterize_gaussians receives at
 * gc_model.py:208-220); depths may be NULL (depth field 0). rec: 12*N floats, 16-byte aligned. */
#define GS_REC_FLOATS_PER_SPLAT 12
int gs_pack_splats(int N, const float *xys, const float *conics, const float *colors, const float *opac,
                   const float *depths, float *rec, void *stream);
/* gs_preprocess_fwd writing records instead of the operator-shaped arrays: rec (N,12) + depths, radii, tiles_hit (what
 * the binning reads). subset != NULL: N = length of the index list, element i = Gaussian subset[i] (eval crop box,
 * gc_model.py:108-113 / :157-170). */
int gs_preprocess_fwd_packed(int N, const float *means, const float *log_scales, const float *quats,
                             const float *opac_logits, const float *feat_dc, const float *feat_rest,
                             const int32_t *subset, int use_sh, int sh_deg, const GsCamera *cam_host, float *rec,
                             float *depths, int32_t *radii, int32_t *tiles_hit, void *stream);
/* Forward / backward compositing over records (3 channels, optional depth channel = out_depth != NULL). `views`
 * images of H x W are composited in one launch: tile_bins has views*T rows (view-major), ids_sorted holds indices
 * into rec, outputs are (views,H,W,..). tile_order / tile_work / out_clamped / out_alpha as in the *_ordered_post
 * entry points (NULL = not wanted). Same results as gs_raster_fwd / gs_raster_bwd on the same inputs. */
int gs_raster_fwd_packed(int views, int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                         const int32_t *tile_bins, const float *rec, const float *bg, float *out_img,
                         float *out_depth, float *final_T, int32_t *final_idx, const int32_t *tile_order,
                         int32_t *tile_work, float *out_clamped, float *out_alpha, void *stream);
/* Tuning knob of the packed forward compositing kernel: survivors composited per loop iteration (1..4; default 1,
 * environment GS_RASTER_FWD_UNROLL). Identical output. Returns the previous setting; other values only query. */
int gs_set_raster_fwd_unroll(int u);
/* Tuning knob of the packed backward compositing kernel: 1 (default; environment GS_RASTER_BWD_DIRECT) = every warp
 * sum goes straight to the gradient block with one red.global; 0 = sums of a tile's warps are first accumulated in
 * shared memory and flushed once per (tile, Gaussian). Same gradients up to the order of float additions. Returns the
 * previous setting. */
int gs_set_raster_bwd_direct(int on);
int gs_raster_bwd_packed(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                         const int32_t *tile_bins, const float *rec, const float *bg, const float *final_T,
                         const int32_t *final_idx, const float *v_out, const float *v_out_alpha, float *v_xys,
                         float *v_conics, float *v_colors, float *v_opac, const int32_t *tile_order,
                         const float *clamp_ref, void *stream);

/* ---- one call per rendered view / batch of views (gc_model.py:172-238; gc_pipeline.py:222-228) ------------------
 * Workspace: ONE device buffer of gs_view_layout(...).bytes bytes holds everything a view keeps between forward
 * and backward; the layout gives the byte offset of each array so the host can alias them (records (views*n,12) f32,
 * depths / radii / tiles_hit (views*n), ids_sorted (I_cap) i32, tile_bins (views*T,2) i32, ...). */
typedef struct GsViewLayout {
    int64_t rec, depths, radii, tiles_hit, ids_sorted, tile_bins, tile_work, order_fwd, order_bwd, total, bytes;
} GsViewLayout;
int gs_view_layout(int n, int views, int64_t I_cap, int T_view, GsViewLayout *out);
size_t gs_view_scratch_bytes(int n, int views, int64_t I_cap, int T_view);
/* Forward of `views` cameras (cams_host[views], same image size) over the same n Gaussians (RAW splatfacto leaves
 * as gs_preprocess_fwd takes them; subset != NULL: n = length of the index list, element i = Gaussian subset[i]).
 * Outputs (views,H,W,3) out_img [+ out_clamped = clamp(out,max=1)] [+ out_alpha (views,H,W) = 1 - final_T]
 * [+ out_depth = sum alpha T depth], final_T, final_idx. I_cap = capacity of ids_sorted; the true intersection
 * count is left in the workspace (layout.total) and, if total_host != NULL (pinned host int32), copied there behind
 * the binning, after which count_event (a cudaEvent_t, may be NULL) is recorded — i.e. BEFORE the compositing
 * kernel is enqueued, so waiting on it does not wait for the compositing. count > I_cap: out_img is NaN-filled
 * (alpha 0) and the caller must redo the call with more room. */
int gs_view_fwd(int n, int views, const float *means, const float *log_scales, const float *quats,
                const float *opac_logits, const float *feat_dc, const float *feat_rest, const int32_t *subset,
                int use_sh, int sh_deg, const GsCamera *cams_host, const float *bg, int64_t I_cap, void *ws,
                size_t ws_bytes, void *scratch, size_t scratch_bytes, float *out_img, float *out_clamped,
                float *out_alpha, float *out_depth, float *final_T, int32_t *final_idx, int32_t *total_host,
                void *count_event, void *stream);
/* Backward of ONE view rendered by gs_view_fwd(views = 1, subset = NULL) from its workspace: gbuf (9*Np + 64 floats,
 * Np = n rounded up to 4; zero-filled here) receives the compositor-level gradients (view-block layout of
 * gs_view_block_floats); flat_leaf_grads (59*Np floats: v_means | v_log_scales at 3Np | v_quats at 6Np | v_opacity
 * logits at 10Np | v_features_dc at 11Np | v_features_rest at 14Np) receives the six leaf gradients, or, when NULL,
 * the view block is finished for the data-parallel exchange instead (gs_finish_view_block). clamp_ref = raw out_img
 * when v_out is w.r.t. the clamped image, else NULL. */
int gs_view_bwd(int n, const float *means, const float *log_scales, const float *quats, const float *opac_logits,
                int use_sh, int sh_deg, const GsCamera *cam_host, const float *bg, void *ws, size_t ws_bytes,
                int64_t I_cap, const float *final_T, const int32_t *final_idx, const float *v_out,
                const float *v_out_alpha, const float *clamp_ref, float *gbuf, float *flat_leaf_grads, void *stream);

/* Debug switch: the same backward with BIT-REPRODUCIBLE gradients (two runs on the same inputs agree bit for bit; the
 * default path adds floats with shared-memory atomics and red.global, whose order depends on scheduling). The
 * compositing kernel keeps per-warp moment rows summed in warp order and writes one 9-float row per (tile, Gaussian)
 * intersection; a stable sort of the list positions by Gaussian id and a per-Gaussian sum in list order replace the
 * global reductions. det_scratch: gs_view_bwd_det_scratch_bytes(n, I_cap) bytes, 16-byte aligned. Slower (one CTA per
 * SM in the compositing, 36 B per intersection of extra traffic): for comparing runs, not for training. */
size_t gs_view_bwd_det_scratch_bytes(int n, int64_t I_cap);
int gs_view_bwd_det(int n, const float *means, const float *log_scales, const float *quats, const float *opac_logits,
                    int use_sh, int sh_deg, const GsCamera *cam_host, const float *bg, void *ws, size_t ws_bytes,
                    int64_t I_cap, const float *final_T, const int32_t *final_idx, const float *v_out,
                    const float *v_out_alpha, const float *clamp_ref, float *gbuf, float *flat_leaf_grads,
                    void *det_scratch, size_t det_scratch_bytes, void *stream);

/* ---- depth post-ops between the render and the depth ControlNet ------------------------------
 * gs_ray_directions replaces compute_ray_directions (/root/reference/gaussctrl/gc_pipeline.py:503-536, a Python double
 * loop): out (res1, res0, 3) float device, unit ray through each pixel of a res0-wide, res1-high viewport at z = -near
 * (the reference unpacks `width, height = resolution`; res0/res1 keep that order). Bit-identical to the numpy result.
 * gs_depth_to_world replaces depth_map_screen_to_world (gc_pipeline.py:286-314): depth (h,w) device, mat_view = the
 * un-flipped world->camera matrix the model returns as outputs['mat_view'] (HOST pointer, 16 floats, row-major;
 * inverted here in double); pts_world (h,w,3) device = inverse(mat_view) * [ray * depth, 1] / w. As in the reference
 * the rays are built for resolution=(h, w), so only square views are defined: h != w returns GS_ERR_BAD_ARG (the
 * reference raises a broadcast error there).
 * gs_depth_to_disparity replaces depth2disparity / depth2disparity_torch (gc_pipeline.py:439-457): depth (n) device ->
 * disparity3 (3, n) device, three identical channels of (1 / (depth + 1e-5)) / max(1 / (depth + 1e-5));
 * bit-identical to torch. scratch4: 4 bytes of device memory, 4-byte aligned. */
int gs_ray_directions(int res0, int res1, double fov_y_rad, double near, float *out, void *stream);
int gs_depth_to_world(int h, int w, const float *depth, const float *mat_view_host, double fov_y_rad, double near,
                      float *pts_world, void *stream);
int gs_depth_to_disparity(int64_t n, const float *depth, float *disparity3, void *scratch4, void *stream);

/* ---- multi-view noise mask ------------------------------------------------------------------
 * Replaces MultiVeiwNoiseRenderer.render_noise_SS (/root/reference/gaussctrl/ad_render.py:273-333) and its GLSL
 * shaders (:384-479): P instances (points (P,3), device) of a triangle mesh (verts (V,3) float, tris (F,3) int32,
 * device; V <= 128) drawn with the row-major 4x4 `view` and `proj` matrices (HOST pointers, 16 floats each);
 * a fragment is kept iff 0 <= depth[r][c] - |view-space fragment position| <= threshold (depth (H,W) device, clipped
 * to [0, 1000]); no depth test, the last primitive in draw order wins; colour = (world position + 1) / 2.
 * out_rgba (H,W,4) uint8 device = the image the reference returns (after its fliplr/flipud); untouched pixels are
 * (0,0,0,0). scratch: gs_noise_mask_scratch_bytes(H, W) bytes, 8-byte aligned. Headless: no GL context or window. */
size_t gs_noise_mask_scratch_bytes(int H, int W);
int gs_noise_mask(int P, const float *points, int V, const float *verts, int F, const int32_t *tris,
                  const float *view16_host, const float *proj16_host, int H, int W, const float *depth,
                  float threshold, void *scratch, size_t scratch_bytes, uint8_t *out_rgba, void *stream);

#ifdef __cplusplus
}
#endif
#endif

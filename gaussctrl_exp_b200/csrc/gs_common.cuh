// Shared device/host helpers for the sm_100a kernels. Not part of the ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/gs_b200.h"

#define GS_ALPHA_MAX 0.999f
#define GS_ALPHA_MIN (1.0f / 255.0f)
#define GS_T_EPS 1e-4f
#define GS_BLUR 0.3f
// gsplat 0.1.2 samples pixel (i,j) at integer coordinates.
#define GS_PIX_CENTER 0.0f

#define GS_NUM_SMS 148  // B200; used only to size persistent grids

extern int64_t g_gs_launches;  // defined in abi.cu

#define GS_LAUNCH_CHECK()                                    \
    do {                                                     \
        ++g_gs_launches;                                     \
        cudaError_t _e = cudaPeekAtLastError();              \
        if (_e != cudaSuccess) return (int)_e;               \
    } while (0)

#define GS_CUDA(x)                                           \
    do {                                                     \
        cudaError_t _e = (x);                                \
        if (_e != cudaSuccess) return (int)_e;               \
    } while (0)

// Programmatic dependent launch (sm_90+): a kernel launched with GS_PDL_LAUNCH may be made resident while the
// previous kernel of the stream drains; it must call gs_pdl_sync() before touching ANY global memory (the wait
// returns once the previous grid has completed and flushed; it is a no-op for a normal launch). Used for the
// chains of short kernels (binning: 23 launches), where the ~1.2 us launch gaps added up to ~27 us per view.
__device__ __forceinline__ void gs_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void gs_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void gs_pdl_sync() {
    gs_pdl_wait();
    gs_pdl_trigger();
}
// abi.cu: false when GS_NO_PDL is set in the environment, or when GS_NO_PDL_GROUPS lists the calling file's group
// (1 = binning chain, 2 = compositing + tile order, 3 = preprocess, 4 = view exchange): A/B switches
bool gs_pdl_enabled(int group);
#ifndef GS_PDL_GROUP
#define GS_PDL_GROUP 0
#endif

#define GS_PDL_LAUNCH_IF(cond, kernel, grid, block, smem, st, ...)                            \
    do {                                                                                       \
        cudaLaunchConfig_t _c = {};                                                            \
        _c.gridDim = dim3(grid);                                                               \
        _c.blockDim = dim3(block);                                                             \
        _c.dynamicSmemBytes = (smem);                                                          \
        _c.stream = (st);                                                                      \
        cudaLaunchAttribute _a[1];                                                             \
        _a[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                         \
        _a[0].val.programmaticStreamSerializationAllowed = 1;                                  \
        _c.attrs = _a;                                                                         \
        _c.numAttrs = ((cond) && gs_pdl_enabled(GS_PDL_GROUP)) ? 1 : 0;                                               \
        cudaError_t _e = cudaLaunchKernelEx(&_c, kernel, __VA_ARGS__);                         \
        if (_e != cudaSuccess) return (int)_e;                                                 \
        ++g_gs_launches;                                                                       \
    } while (0)

#define GS_PDL_LAUNCH(kernel, grid, block, smem, st, ...) GS_PDL_LAUNCH_IF(true, kernel, grid, block, smem, st, __VA_ARGS__)

static inline bool gs_aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static inline int gs_ceil_log2(int v) { int b = 0; while ((1 << b) < v) ++b; return b; }
static inline size_t gs_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__device__ __forceinline__ int gs_clampi(int v, int lo, int hi) { return min(hi, max(lo, v)); }

// A.1.6: tile bbox of a splat (pixel centre, integer radius); C truncation toward zero.
__device__ __forceinline__ void gs_tile_bbox(float x, float y, float radius, int tiles_x, int tiles_y,
                                             int &x0, int &y0, int &x1, int &y1) {
    const float tcx = __fdiv_rn(x, 16.0f), tcy = __fdiv_rn(y, 16.0f), tr = __fdiv_rn(radius, 16.0f);
    x0 = gs_clampi((int)__fsub_rn(tcx, tr), 0, tiles_x);
    x1 = gs_clampi((int)__fadd_rn(__fadd_rn(tcx, tr), 1.0f), 0, tiles_x);
    y0 = gs_clampi((int)__fsub_rn(tcy, tr), 0, tiles_y);
    y1 = gs_clampi((int)__fadd_rn(__fadd_rn(tcy, tr), 1.0f), 0, tiles_y);
}

__device__ __forceinline__ unsigned gs_lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// streaming (read-once) loads that bypass L1 allocation
__device__ __forceinline__ float4 gs_ldg_stream4(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// internal helpers shared between translation units (not exported through the C ABI header)
int gs_internal_exclusive_scan(int n, const int32_t *in, int32_t *out, int32_t *total_dev, void *scratch,
                               size_t scratch_bytes, cudaStream_t st);
// binning.cu: gs_bin_fast over the elements of several views at once, xy read with a stride (packed records)
// deterministic per-Gaussian sums of per-intersection gradient rows (binning.cu; used by gs_view_bwd_det)
size_t gs_internal_segment_scratch_bytes(int n, int64_t I_cap);
int gs_internal_segment_reduce(int n, int64_t I_cap, const int32_t *total_dev, const int32_t *ids_sorted, const float *rows,
                               float *gbuf, void *scratch, size_t scratch_bytes, cudaStream_t st);
// bucket form of the binning (tilebucket.cu): same contract as gs_bin_fast_ex below
size_t gs_bin_bucket_scratch_bytes(int64_t I_cap, int T);
bool gs_bin_bucket_supported(int T_view, int64_t I_cap);
int gs_bin_bucket_ex(int N, int n_per_view, int64_t I_cap, const float *xys, int xy_stride, const float *depths,
                     const int32_t *radii, int tiles_x, int tiles_y, int32_t *ids_sorted, int32_t *tile_bins,
                     int32_t *total_dev, void *scratch, size_t scratch_bytes, void *stream);
int gs_bin_fast_ex(int N, int n_per_view, int64_t I_cap, const float *xys, int xy_stride, const float *depths,
                   const int32_t *radii, const int32_t *tiles_hit, int tiles_x, int tiles_y, int32_t *ids_sorted,
                   int32_t *tile_bins, int32_t *total_dev, void *scratch, size_t scratch_bytes, void *stream,
                   bool precleared);
int gs_bin_fast_clear(int N, int64_t I_cap, int T, int32_t *tile_bins, void *scratch, size_t scratch_bytes, void *stream);
// preprocess_fwd.cu / preprocess_bwd.cu: packed-record forms used by view.cu
int gs_launch_preprocess_fwd_packed(int N, const float *means, const float *log_scales, const float *quats,
                                    const float *opac_logits, const float *feat_dc, const float *feat_rest, int use_sh,
                                    int sh_deg, const GsCamera *cam_host, const int32_t *subset, float *rec, float *depths,
                                    int32_t *radii, int32_t *tiles_hit, cudaStream_t st);
int gs_launch_preprocess_bwd(int N, const float *means, const float *log_scales, const float *quats,
                             const float *opac_logits, int use_sh, int sh_deg, const GsCamera *cam_host,
                             const int32_t *radii, const float *rgbs, int rgb_stride, const float *v_xys,
                             const float *v_conics, const float *v_rgbs, const float *v_opac, float *v_means,
                             float *v_log_scales, float *v_quats, float *v_opac_logits, float *v_feat_dc,
                             float *v_feat_rest, cudaStream_t st);
int gs_launch_finish_view_block(int N, int use_sh, const GsCamera *cam_host, const float *rgbs, int rgb_stride,
                                float *block, cudaStream_t st);

// Shared between forward and backward compositing so both take identical skip / clamp
// decisions for every (pixel, Gaussian) pair.
#pragma once
#include "gs_common.cuh"

// The compositing kernels and the tile-order kernel are programmatic dependent launches only for images of up to this
// many tiles. Measured: at 512x512 (1 024 tiles, ~1 wave of CTAs) the early launch is worth 2 % of a training step
// (0.714 vs 0.729 ms); at 1920x1080 (8 160 tiles, 9 waves) it COSTS 15 % (5 M Gaussians: 272 vs 319 training views/s,
// 476 vs 560 eval views/s) — the dependent grid's CTAs take over SM slots while the many-wave grid is still running.
#define GS_RASTER_PDL_MAX_TILES 2048

// A 16x16 tile is covered by 8 warps of 8x4 pixels: compact warp footprints mean more
// warp-uniform early termination than the row-major 16x2 footprint.
__device__ __forceinline__ void gs_pixel_of_thread(int tile_x, int tile_y, int &px, int &py) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    px = tile_x * GS_TILE + (warp & 1) * 8 + (lane & 7);
    py = tile_y * GS_TILE + (warp >> 1) * 4 + (lane >> 3);
}

// Splat as staged in shared memory. The conic (a,b,c) is stored pre-scaled so that the exponent is one
// short FMA chain feeding MUFU.EX2 directly:  qa = -a/2 * log2(e), qb = -b * log2(e), qc = -c/2 * log2(e)
//   exp(-sigma) = exp2(qa dx^2 + qb dx dy + qc dy^2)
#define GS_LOG2E 1.4426950408889634f
struct GsSplat {
    float x, y, opac, qa, qb, qc;
};
__device__ __forceinline__ void gs_prescale_conic(float a, float b, float c, float &qa, float &qb, float &qc) {
    qa = -0.5f * GS_LOG2E * a; qb = -GS_LOG2E * b; qc = -0.5f * GS_LOG2E * c;
}
// true conic back from the staged form (backward flush)
__device__ __forceinline__ void gs_unscale_conic(float qa, float qb, float qc, float &a, float &b, float &c) {
    a = qa * (-2.0f / GS_LOG2E); b = qb * (-1.0f / GS_LOG2E); c = qc * (-2.0f / GS_LOG2E);
}
__device__ __forceinline__ float gs_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// returns false when the pair is skipped (sigma < 0 or alpha < 1/255)
__device__ __forceinline__ bool gs_eval_alpha(const GsSplat &g, float fx, float fy, float &dx, float &dy,
                                              float &vis, float &alpha) {
    dx = g.x - fx;
    dy = g.y - fy;
    const float p = dx * (g.qa * dx + g.qb * dy) + (g.qc * dy) * dy;   // = -sigma * log2(e)
    vis = gs_ex2(p);
    alpha = fminf(GS_ALPHA_MAX, g.opac * vis);
    return !(p > 0.0f) && (alpha >= GS_ALPHA_MIN);
}

// Conservative half-extents (in pixels) of the region where a splat can reach alpha >= 1/255:
//   alpha = opac * exp(-sigma) >= 1/255  <=>  sigma <= tau = ln(255 * opac)
//   {sigma <= tau} is the ellipse 1/2 d^T Q d <= tau (Q = conic), whose bounding box is
//   |dx| <= sqrt(2 tau c / det Q), |dy| <= sqrt(2 tau a / det Q).
// Pairs outside the box would be skipped by the alpha test anyway, so culling them changes no
// output bit; the 1e-4 inflation keeps the test on the safe side of fp32 / __expf rounding.
__device__ __forceinline__ float2 gs_alpha_extent(float opac, float ca, float cb, float cc) {
    const float det = ca * cc - cb * cb;
    if (!(det > 0.0f) || !(ca > 0.0f) || !(cc > 0.0f)) return make_float2(3.0e38f, 3.0e38f);  // not PD: never cull
    const float t = 255.0f * opac;
    if (!(t > 1.0f)) return make_float2(-1.0f, -1.0f);                                        // can never pass
    const float tau2 = 2.0f * __logf(t) * 1.001f + 1e-4f;
    const float inv = 1.0f / det;
    return make_float2(sqrtf(tau2 * cc * inv) * 1.0001f + 1e-4f, sqrtf(tau2 * ca * inv) * 1.0001f + 1e-4f);
}

// origin (pixel-centre coordinates) of the calling warp's 8x4 pixel rect
__device__ __forceinline__ void gs_warp_rect(int tile_x, int tile_y, float &x0, float &y0) {
    const int warp = threadIdx.x >> 5;
    x0 = (float)(tile_x * GS_TILE + (warp & 1) * 8) + GS_PIX_CENTER;
    y0 = (float)(tile_y * GS_TILE + (warp >> 1) * 4) + GS_PIX_CENTER;
}

// Shared-memory access through 32-bit shared-window addresses held in ordinary registers. With plain C++
// indexing the compiler re-derives the window base (S2UR SR_CgaCtaId / ULEA ...) inside the hot replay loops:
// ~9 of ~137 issue slots per (warp, Gaussian) iteration of the backward (ncu source view, profiles/r01h).
__device__ __forceinline__ unsigned gs_smem_addr(const void *p) {
    unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("" : "+r"(a));       // opaque: keeps the compiler from re-materialising the conversion per use
    return a;
}
__device__ __forceinline__ float2 gs_lds64(unsigned a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ float4 gs_lds128(unsigned a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ float gs_lds32(unsigned a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ unsigned gs_lds8(unsigned a) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void gs_sts32(unsigned a, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
}
__device__ __forceinline__ float gs_rcp(float x) {     // 1 ulp-class reciprocal, no range fix-up (x in [1e-3, 1])
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

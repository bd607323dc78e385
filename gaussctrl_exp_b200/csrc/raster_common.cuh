// Shared between forward and backward compositing so both take identical skip / clamp
// decisions for every (pixel, Gaussian) pair.
#pragma once
#include "gs_common.cuh"

// A 16x16 tile is covered by 8 warps of 8x4 pixels: compact warp footprints mean more
// warp-uniform early termination than the row-major 16x2 footprint.
__device__ __forceinline__ void gs_pixel_of_thread(int tile_x, int tile_y, int &px, int &py) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    px = tile_x * GS_TILE + (warp & 1) * 8 + (lane & 7);
    py = tile_y * GS_TILE + (warp >> 1) * 4 + (lane >> 3);
}

struct GsSplat {
    float x, y, opac, ca, cb, cc;
};

// returns false when the pair is skipped (sigma < 0 or alpha < 1/255)
__device__ __forceinline__ bool gs_eval_alpha(const GsSplat &g, float fx, float fy, float &dx, float &dy,
                                              float &vis, float &alpha) {
    dx = g.x - fx;
    dy = g.y - fy;
    const float sigma = 0.5f * (g.ca * dx * dx + g.cc * dy * dy) + g.cb * dx * dy;
    vis = __expf(-sigma);
    alpha = fminf(GS_ALPHA_MAX, g.opac * vis);
    return !(sigma < 0.0f || alpha < GS_ALPHA_MIN);
}

// Multi-view noise mask: the arithmetic of the reference's only OpenGL user,
// /root/reference/gaussctrl/ad_render.py:273-333 (render_noise_SS) with the shaders of :384-479, as a headless
// CUDA point-splat. The GL context, window, texture upload and glReadPixels are replaced by device buffers
// ("unavailable on this box"); what is restated is the pipeline's arithmetic:
//   vertex shader   : instanced sphere mesh, clip = P * (V * (aPos + instancePos)), colour = (world + 1) / 2
//   rasterisation   : pixel-centre coverage of every triangle (no face culling, no depth test: GL defaults, never
//                     changed by the reference), window coordinates snapped to 1/256 px, top-left tie rule,
//                     perspective-correct attribute interpolation
//   fragment shader : d_frag = |view-space position of the fragment|; keep iff 0 <= depth_tex - d_frag <= threshold
//                     (threshold 0.016 + 0.0001, :423); colour = interpolated (world + 1) / 2, alpha 1
//   framebuffer     : the LAST primitive in draw order (instance-major, triangle-minor) that keeps a fragment wins
//                     (no depth test, no blending) -> one 64-bit atomicMax per kept fragment on (order << 32 | rgba8)
//   read-back       : RGBA8, then np.fliplr / np.flipud of the PIL image (:326-327): the returned image pixel (r, c)
//                     is GL window pixel (W-1-c, H-1-r), and the depth texture (uploaded flipped the same way, :110-111)
//                     is read at depth[r][c] (sampled exactly at texel centres, so GL_LINEAR returns the texel).
// Coverage is decided in exact integer arithmetic, everything else in fp32 with a fixed operation order and no FMA
// contraction (this file is compiled with -fmad=false), so the CPU oracle reproduces the image bit for bit.
// One warp per instance: the lanes transform the mesh vertices into shared memory, then take the triangles in turn.
#include "gs_common.cuh"

namespace {

constexpr int NM_MAX_VERTS = 128;
constexpr int NM_WARPS = 8;
constexpr int NM_SUB = 256;                 // sub-pixel steps per pixel (8 fractional bits, like GL hardware)
constexpr float NM_GUARD = 1.0e5f;          // |window coordinate| beyond which a vertex is treated as unrasterisable

struct NmMats {
    float view[16];   // row-major world -> view
    float proj[16];   // row-major view -> clip
};

struct NmVert {
    int ix, iy;       // window coordinates, 1/256 px
    float invw;
    float vx, vy, vz; // view-space position
    float cr, cg, cb; // (world + 1) / 2
    int ok;           // w > 0 and inside the guard band
};

__device__ __forceinline__ unsigned nm_to_u8(float c) {
    c = fminf(fmaxf(c, 0.0f), 1.0f);
    return (unsigned)floorf(c * 255.0f + 0.5f);
}

__device__ __forceinline__ bool nm_edge_in(long long e, int dx, int dy) {
    // interior on the left of the directed edge (counter-clockwise, y up); ties go to left and top edges
    return e > 0 || (e == 0 && (dy < 0 || (dy == 0 && dx < 0)));
}

__global__ void __launch_bounds__(NM_WARPS * 32)
nm_raster_kernel(int P, const float *__restrict__ points, int V, const float *__restrict__ verts, int F,
                 const int32_t *__restrict__ tris, NmMats M, int H, int W, const float *__restrict__ depth,
                 float threshold, unsigned long long *__restrict__ winner) {
    __shared__ NmVert sv[NM_WARPS][NM_MAX_VERTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int inst = blockIdx.x * NM_WARPS + warp;
    if (inst >= P) return;
    const float ox = points[3 * inst], oy = points[3 * inst + 1], oz = points[3 * inst + 2];
    NmVert *vs = sv[warp];
    const float *Vm = M.view, *Pm = M.proj;
    for (int k = lane; k < V; k += 32) {
        const float wx = verts[3 * k] + ox, wy = verts[3 * k + 1] + oy, wz = verts[3 * k + 2] + oz;
        NmVert o;
        o.vx = Vm[0] * wx + Vm[1] * wy + Vm[2] * wz + Vm[3];
        o.vy = Vm[4] * wx + Vm[5] * wy + Vm[6] * wz + Vm[7];
        o.vz = Vm[8] * wx + Vm[9] * wy + Vm[10] * wz + Vm[11];
        const float cx = Pm[0] * o.vx + Pm[1] * o.vy + Pm[2] * o.vz + Pm[3];
        const float cy = Pm[4] * o.vx + Pm[5] * o.vy + Pm[6] * o.vz + Pm[7];
        const float cw = Pm[12] * o.vx + Pm[13] * o.vy + Pm[14] * o.vz + Pm[15];
        o.cr = (wx + 1.0f) / 2.0f; o.cg = (wy + 1.0f) / 2.0f; o.cb = (wz + 1.0f) / 2.0f;
        o.ok = 0; o.ix = o.iy = 0; o.invw = 0.f;
        if (cw > 0.0f) {
            o.invw = 1.0f / cw;
            const float sx = (cx * o.invw * 0.5f + 0.5f) * (float)W;
            const float sy = (cy * o.invw * 0.5f + 0.5f) * (float)H;
            if (fabsf(sx) < NM_GUARD && fabsf(sy) < NM_GUARD) {
                o.ix = (int)floorf(sx * (float)NM_SUB + 0.5f);
                o.iy = (int)floorf(sy * (float)NM_SUB + 0.5f);
                o.ok = 1;
            }
        }
        vs[k] = o;
    }
    __syncwarp();
    for (int f = lane; f < F; f += 32) {
        int ia = tris[3 * f], ib = tris[3 * f + 1], ic = tris[3 * f + 2];
        if ((unsigned)ia >= (unsigned)V || (unsigned)ib >= (unsigned)V || (unsigned)ic >= (unsigned)V) continue;
        if (!(vs[ia].ok && vs[ib].ok && vs[ic].ok)) continue;     // a vertex behind the eye: triangle dropped (no clipper)
        long long area2 = (long long)(vs[ib].ix - vs[ia].ix) * (vs[ic].iy - vs[ia].iy) -
                          (long long)(vs[ib].iy - vs[ia].iy) * (vs[ic].ix - vs[ia].ix);
        if (area2 == 0) continue;
        if (area2 < 0) { const int t = ib; ib = ic; ic = t; area2 = -area2; }
        const NmVert a = vs[ia], b = vs[ib], c = vs[ic];
        // pixels whose centre (256 x + 128, 256 y + 128) can lie inside
        const int minx = min(a.ix, min(b.ix, c.ix)), maxx = max(a.ix, max(b.ix, c.ix));
        const int miny = min(a.iy, min(b.iy, c.iy)), maxy = max(a.iy, max(b.iy, c.iy));
        int x0 = (minx - NM_SUB / 2 + NM_SUB - 1) >> 8, x1 = (maxx - NM_SUB / 2) >> 8;   // arithmetic shifts: floor
        int y0 = (miny - NM_SUB / 2 + NM_SUB - 1) >> 8, y1 = (maxy - NM_SUB / 2) >> 8;
        x0 = max(x0, 0); y0 = max(y0, 0); x1 = min(x1, W - 1); y1 = min(y1, H - 1);
        if (x0 > x1 || y0 > y1) continue;
        const float farea = (float)area2;
        const unsigned order = (unsigned)inst * (unsigned)F + (unsigned)f;
        for (int y = y0; y <= y1; ++y) {
            const int py = y * NM_SUB + NM_SUB / 2;
            for (int x = x0; x <= x1; ++x) {
                const int px = x * NM_SUB + NM_SUB / 2;
                const long long e_bc = (long long)(c.ix - b.ix) * (py - b.iy) - (long long)(c.iy - b.iy) * (px - b.ix);
                const long long e_ca = (long long)(a.ix - c.ix) * (py - c.iy) - (long long)(a.iy - c.iy) * (px - c.ix);
                const long long e_ab = (long long)(b.ix - a.ix) * (py - a.iy) - (long long)(b.iy - a.iy) * (px - a.ix);
                if (!nm_edge_in(e_bc, c.ix - b.ix, c.iy - b.iy) || !nm_edge_in(e_ca, a.ix - c.ix, a.iy - c.iy) ||
                    !nm_edge_in(e_ab, b.ix - a.ix, b.iy - a.iy))
                    continue;
                // perspective-correct weights
                const float la = ((float)e_bc / farea) * a.invw, lb = ((float)e_ca / farea) * b.invw,
                            lc = ((float)e_ab / farea) * c.invw;
                const float ls = la + lb + lc;
                const float wa = la / ls, wb = lb / ls, wc = lc / ls;
                const float qx = wa * a.vx + wb * b.vx + wc * c.vx;
                const float qy = wa * a.vy + wb * b.vy + wc * c.vy;
                const float qz = wa * a.vz + wb * b.vz + wc * c.vz;
                // depth clipping (-w <= z <= w) of the fragment
                const float zc = Pm[8] * qx + Pm[9] * qy + Pm[10] * qz + Pm[11];
                const float wc4 = Pm[12] * qx + Pm[13] * qy + Pm[14] * qz + Pm[15];
                if (zc < -wc4 || zc > wc4) continue;
                const float dfrag = sqrtf(qx * qx + qy * qy + qz * qz);
                // returned image pixel (r, c) <-> GL window pixel (W-1-c, H-1-r); depth texel = depth[r][c]
                const int r = H - 1 - y, cc = W - 1 - x;
                const float dtex = fminf(fmaxf(depth[(size_t)r * W + cc], 0.0f), 1000.0f);   // :108 np.clip
                const float diff = dtex - dfrag;
                if (diff < 0.0f || diff > threshold) continue;                               // :446-451 discard
                const unsigned rgba = nm_to_u8(wa * a.cr + wb * b.cr + wc * c.cr) |
                                      (nm_to_u8(wa * a.cg + wb * b.cg + wc * c.cg) << 8) |
                                      (nm_to_u8(wa * a.cb + wb * b.cb + wc * c.cb) << 16) | (255u << 24);
                // +1 so that primitive 0 beats the cleared framebuffer
                atomicMax(&winner[(size_t)r * W + cc], (((unsigned long long)order + 1ull) << 32) | rgba);
            }
        }
    }
}

__global__ void __launch_bounds__(256)
nm_resolve_kernel(int n, const unsigned long long *__restrict__ winner, uint32_t *__restrict__ out_rgba) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out_rgba[i] = (uint32_t)(winner[i] & 0xffffffffull);   // cleared framebuffer: (0,0,0,0)
}

}  // namespace

extern "C" size_t gs_noise_mask_scratch_bytes(int H, int W) {
    return (size_t)(H > 0 ? H : 0) * (size_t)(W > 0 ? W : 0) * sizeof(unsigned long long);
}

extern "C" int gs_noise_mask(int P, const float *points, int V, const float *verts, int F, const int32_t *tris,
                             const float *view16_host, const float *proj16_host, int H, int W, const float *depth,
                             float threshold, void *scratch, size_t scratch_bytes, uint8_t *out_rgba, void *stream) {
    if (P < 0 || V <= 0 || F <= 0 || H <= 0 || W <= 0) return GS_ERR_BAD_ARG;
    if (!verts || !tris || !view16_host || !proj16_host || !depth || !scratch || !out_rgba || (P > 0 && !points))
        return GS_ERR_BAD_ARG;
    if (V > NM_MAX_VERTS) return GS_ERR_UNSUPPORTED;
    if ((unsigned long long)P * (unsigned long long)F >= 0xffffffffull) return GS_ERR_UNSUPPORTED;   // 32-bit draw order
    if (scratch_bytes < gs_noise_mask_scratch_bytes(H, W)) return GS_ERR_SCRATCH_SMALL;
    if ((reinterpret_cast<uintptr_t>(scratch) & 7u) || (reinterpret_cast<uintptr_t>(out_rgba) & 3u)) return GS_ERR_MISALIGNED;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *winner = reinterpret_cast<unsigned long long *>(scratch);
    const int n = H * W;
    GS_CUDA(cudaMemsetAsync(winner, 0, (size_t)n * sizeof(unsigned long long), st));
    if (P > 0) {
        NmMats M;
        for (int i = 0; i < 16; ++i) { M.view[i] = view16_host[i]; M.proj[i] = proj16_host[i]; }
        nm_raster_kernel<<<(P + NM_WARPS - 1) / NM_WARPS, NM_WARPS * 32, 0, st>>>(P, points, V, verts, F, tris, M, H, W, depth,
                                                                                threshold, winner);
        GS_LAUNCH_CHECK();
    }
    nm_resolve_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, winner, reinterpret_cast<uint32_t *>(out_rgba));
    GS_LAUNCH_CHECK();
    return 0;
}

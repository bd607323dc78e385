// EWA projection math shared by project_fwd.cu / preprocess.cu (forward) and project_bwd.cu /
// preprocess.cu (backward). SURVEY.md App. A.1.
//
// The FORWARD function must only be instantiated in translation units compiled with
// -fmad=false: radii / tiles_hit are integers derived from this fp32 arithmetic and are
// compared bit-exactly with the CPU oracle, so every multiply and add rounds separately, in the
// written order.
#pragma once
#include "gs_common.cuh"

struct ProjOut {
    float cov[6];
    float2 xy;
    float depth, ca, cb, cc;
    int radius, hit;
};

// exp(log_scale) as every kernel of the path evaluates it: correctly rounded ((float) of the double result), so that the
// forward, the single-view backward and the multi-view / data-parallel leaf backwards all differentiate at the SAME
// scale (expf may differ by 1 ulp between builds and call sites; gc_model.py:172 torch.exp).
__device__ __forceinline__ float gs_exp_scale(float log_scale) { return (float)exp((double)log_scale); }

// V: 12 floats (3x4 row-major world->camera), P: 16 floats (P@V). q = (w,x,y,z) un-normalised.
__device__ __forceinline__ void gs_project_one(const float *V, const float *P, float p0, float p1, float p2,
                                               float s0, float s1, float s2, float glob, float4 q, float fx,
                                               float fy, float cx, float cy, int H, int W, int tiles_x,
                                               int tiles_y, float clip, ProjOut &o) {
#pragma unroll
    for (int j = 0; j < 6; ++j) o.cov[j] = 0.f;
    o.xy = make_float2(0.f, 0.f);
    o.depth = 0.f; o.ca = 0.f; o.cb = 0.f; o.cc = 0.f;
    o.radius = 0; o.hit = 0;
    // A.1.1 view transform + near clip
    const float t0 = V[0] * p0 + V[1] * p1 + V[2] * p2 + V[3];
    const float t1 = V[4] * p0 + V[5] * p1 + V[6] * p2 + V[7];
    const float tz = V[8] * p0 + V[9] * p1 + V[10] * p2 + V[11];
    if (!(tz > clip)) return;
    // A.1.2 cov3d = (R S)(R S)^T
    float w = q.x, x = q.y, y = q.z, z = q.w;
    const float nrm = sqrtf(w * w + x * x + y * y + z * z);
    const float inv = 1.0f / nrm;
    w = w * inv; x = x * inv; y = y * inv; z = z * inv;
    const float R0 = 1.0f - 2.0f * (y * y + z * z), R1 = 2.0f * (x * y - w * z), R2 = 2.0f * (x * z + w * y);
    const float R3 = 2.0f * (x * y + w * z), R4 = 1.0f - 2.0f * (x * x + z * z), R5 = 2.0f * (y * z - w * x);
    const float R6 = 2.0f * (x * z - w * y), R7 = 2.0f * (y * z + w * x), R8 = 1.0f - 2.0f * (x * x + y * y);
    const float sx = glob * s0, sy = glob * s1, sz = glob * s2;
    const float M0 = R0 * sx, M1 = R1 * sy, M2 = R2 * sz;
    const float M3 = R3 * sx, M4 = R4 * sy, M5 = R5 * sz;
    const float M6 = R6 * sx, M7 = R7 * sy, M8 = R8 * sz;
    const float S0 = M0 * M0 + M1 * M1 + M2 * M2;
    const float S1 = M0 * M3 + M1 * M4 + M2 * M5;
    const float S2 = M0 * M6 + M1 * M7 + M2 * M8;
    const float S3 = M3 * M3 + M4 * M4 + M5 * M5;
    const float S4 = M3 * M6 + M4 * M7 + M5 * M8;
    const float S5 = M6 * M6 + M7 * M7 + M8 * M8;
    o.cov[0] = S0; o.cov[1] = S1; o.cov[2] = S2; o.cov[3] = S3; o.cov[4] = S4; o.cov[5] = S5;
    // A.1.3 EWA
    const float limx = 1.3f * (0.5f * (float)W / fx), limy = 1.3f * (0.5f * (float)H / fy);
    const float rx = t0 / tz, ry = t1 / tz;
    const float tcx = tz * fminf(limx, fmaxf(-limx, rx));
    const float tcy = tz * fminf(limy, fmaxf(-limy, ry));
    const float rz = 1.0f / tz, rz2 = rz * rz;
    const float J00 = fx * rz, J02 = -fx * tcx * rz2;
    const float J11 = fy * rz, J12 = -fy * tcy * rz2;
    const float T0 = J00 * V[0] + J02 * V[8], T1 = J00 * V[1] + J02 * V[9], T2 = J00 * V[2] + J02 * V[10];
    const float T3 = J11 * V[4] + J12 * V[8], T4 = J11 * V[5] + J12 * V[9], T5 = J11 * V[6] + J12 * V[10];
    const float u0 = S0 * T0 + S1 * T1 + S2 * T2;
    const float u1 = S1 * T0 + S3 * T1 + S4 * T2;
    const float u2 = S2 * T0 + S4 * T1 + S5 * T2;
    const float w0 = S0 * T3 + S1 * T4 + S2 * T5;
    const float w1 = S1 * T3 + S3 * T4 + S4 * T5;
    const float w2 = S2 * T3 + S4 * T4 + S5 * T5;
    const float a = (T0 * u0 + T1 * u1 + T2 * u2) + GS_BLUR;
    const float b = T3 * u0 + T4 * u1 + T5 * u2;
    const float c = (T3 * w0 + T4 * w1 + T5 * w2) + GS_BLUR;
    // A.1.4 conic + radius
    const float det = a * c - b * b;
    if (det == 0.0f) return;
    const float inv_det = 1.0f / det;
    const float mid = 0.5f * (a + c);
    const float disc = sqrtf(fmaxf(0.1f, mid * mid - det));
    const float v1 = mid + disc, v2 = mid - disc;
    const float radius = ceilf(3.0f * sqrtf(fmaxf(v1, v2)));
    // A.1.5 pixel centre
    const float hx = P[0] * p0 + P[1] * p1 + P[2] * p2 + P[3];
    const float hy = P[4] * p0 + P[5] * p1 + P[6] * p2 + P[7];
    const float hw = P[12] * p0 + P[13] * p1 + P[14] * p2 + P[15];
    const float rw = 1.0f / (hw + 1e-6f);
    const float px = 0.5f * (float)W * (hx * rw) + cx - 0.5f;
    const float py = 0.5f * (float)H * (hy * rw) + cy - 0.5f;
    // A.1.6 tile bbox
    int x0, y0, x1, y1;
    gs_tile_bbox(px, py, radius, tiles_x, tiles_y, x0, y0, x1, y1);
    const int area = (x1 - x0) * (y1 - y0);
    if (area <= 0) return;
    o.hit = area; o.depth = tz; o.radius = (int)radius;
    o.xy = make_float2(px, py);
    o.ca = c * inv_det; o.cb = -b * inv_det; o.cc = a * inv_det;
}

struct ProjGrad {
    float vp[3], vs[3];
    float4 vq;
};

// Exact vjp of gs_project_one w.r.t. (p, s, q) for a visible Gaussian. vA,vB,vC = v_conic.
__device__ __forceinline__ void gs_project_one_bwd(const float *V, const float *P, float p0, float p1, float p2,
                                                   float sc0, float sc1, float sc2, float glob, float4 q,
                                                   float fx, float fy, int H, int W, float2 vxy, float v_depth,
                                                   float vA, float vB, float vC, ProjGrad &g) {
    // ---- pixel centre ----
    const float hx = P[0] * p0 + P[1] * p1 + P[2] * p2 + P[3];
    const float hy = P[4] * p0 + P[5] * p1 + P[6] * p2 + P[7];
    const float hw = P[12] * p0 + P[13] * p1 + P[14] * p2 + P[15];
    const float rw = 1.0f / (hw + 1e-6f);
    const float v_hx = 0.5f * (float)W * rw * vxy.x, v_hy = 0.5f * (float)H * rw * vxy.y;
    const float v_rw = 0.5f * (float)W * hx * vxy.x + 0.5f * (float)H * hy * vxy.y;
    const float v_hw = -rw * rw * v_rw;
    float vp0 = P[0] * v_hx + P[4] * v_hy + P[12] * v_hw;
    float vp1 = P[1] * v_hx + P[5] * v_hy + P[13] * v_hw;
    float vp2 = P[2] * v_hx + P[6] * v_hy + P[14] * v_hw;
    // ---- recompute forward mid-state ----
    const float t0 = V[0] * p0 + V[1] * p1 + V[2] * p2 + V[3];
    const float t1 = V[4] * p0 + V[5] * p1 + V[6] * p2 + V[7];
    const float tz = V[8] * p0 + V[9] * p1 + V[10] * p2 + V[11];
    const float inv = 1.0f / sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    const float w = q.x * inv, x = q.y * inv, y = q.z * inv, z = q.w * inv;
    const float R[9] = {1.f - 2.f * (y * y + z * z), 2.f * (x * y - w * z), 2.f * (x * z + w * y),
                        2.f * (x * y + w * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - w * x),
                        2.f * (x * z - w * y), 2.f * (y * z + w * x), 1.f - 2.f * (x * x + y * y)};
    const float s[3] = {glob * sc0, glob * sc1, glob * sc2};
    float M[9], S[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) M[3 * r + c] = R[3 * r + c] * s[c];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            S[3 * r + c] = M[3 * r] * M[3 * c] + M[3 * r + 1] * M[3 * c + 1] + M[3 * r + 2] * M[3 * c + 2];
    const float rz = 1.0f / tz, rz2 = rz * rz, rz3 = rz2 * rz;
    const float limx = 1.3f * (0.5f * (float)W / fx), limy = 1.3f * (0.5f * (float)H / fy);
    const float rx = __fdiv_rn(t0, tz), ry = __fdiv_rn(t1, tz);  // clamp decision as the forward takes it
    const int clx = rx > limx ? 1 : (rx < -limx ? -1 : 0);
    const int cly = ry > limy ? 1 : (ry < -limy ? -1 : 0);
    const float tcx = clx == 0 ? t0 : tz * ((float)clx * limx);
    const float tcy = cly == 0 ? t1 : tz * ((float)cly * limy);
    const float J00 = fx * rz, J02 = -fx * tcx * rz2, J11 = fy * rz, J12 = -fy * tcy * rz2;
    const float Tm[6] = {J00 * V[0] + J02 * V[8], J00 * V[1] + J02 * V[9], J00 * V[2] + J02 * V[10],
                         J11 * V[4] + J12 * V[8], J11 * V[5] + J12 * V[9], J11 * V[6] + J12 * V[10]};
    float ST0[3], ST1[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        ST0[r] = S[3 * r] * Tm[0] + S[3 * r + 1] * Tm[1] + S[3 * r + 2] * Tm[2];
        ST1[r] = S[3 * r] * Tm[3] + S[3 * r + 1] * Tm[4] + S[3 * r + 2] * Tm[5];
    }
    const float a = Tm[0] * ST0[0] + Tm[1] * ST0[1] + Tm[2] * ST0[2] + GS_BLUR;
    const float b = Tm[3] * ST0[0] + Tm[4] * ST0[1] + Tm[5] * ST0[2];
    const float c = Tm[3] * ST1[0] + Tm[4] * ST1[1] + Tm[5] * ST1[2] + GS_BLUR;
    const float det = a * c - b * b;
    const float id2 = 1.0f / (det * det);
    // ---- conic = (c,-b,a)/det ----
    const float v_a = (-c * c * vA + b * c * vB - b * b * vC) * id2;
    const float v_b = (2.f * b * c * vA - (det + 2.f * b * b) * vB + 2.f * a * b * vC) * id2;
    const float v_c = (-b * b * vA + a * b * vB - a * a * vC) * id2;
    // ---- cov2d = Tm S Tm^T ----
    float vS[9], vT[6];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int k = 0; k < 3; ++k)
            vS[3 * r + k] = v_a * Tm[r] * Tm[k] + v_b * Tm[r] * Tm[3 + k] + v_c * Tm[3 + r] * Tm[3 + k];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        vT[k] = 2.f * v_a * ST0[k] + v_b * ST1[k];
        vT[3 + k] = 2.f * v_c * ST1[k] + v_b * ST0[k];
    }
    const float vJ00 = vT[0] * V[0] + vT[1] * V[1] + vT[2] * V[2];
    const float vJ02 = vT[0] * V[8] + vT[1] * V[9] + vT[2] * V[10];
    const float vJ11 = vT[3] * V[4] + vT[4] * V[5] + vT[5] * V[6];
    const float vJ12 = vT[3] * V[8] + vT[4] * V[9] + vT[5] * V[10];
    const float v_tcx = -fx * rz2 * vJ02, v_tcy = -fy * rz2 * vJ12;
    float v_t2 = -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2.f * fx * tcx * rz3 * vJ02 + 2.f * fy * tcy * rz3 * vJ12;
    const float v_t0 = clx == 0 ? v_tcx : 0.f;
    const float v_t1 = cly == 0 ? v_tcy : 0.f;
    if (clx != 0) v_t2 += (float)clx * limx * v_tcx;
    if (cly != 0) v_t2 += (float)cly * limy * v_tcy;
    v_t2 += v_depth;
    vp0 += V[0] * v_t0 + V[4] * v_t1 + V[8] * v_t2;
    vp1 += V[1] * v_t0 + V[5] * v_t1 + V[9] * v_t2;
    vp2 += V[2] * v_t0 + V[6] * v_t1 + V[10] * v_t2;
    // ---- S = M M^T, M = R diag(s) ----
    float vR[9];
    float vsc[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 3; ++j) acc += (vS[3 * r + j] + vS[3 * j + r]) * M[3 * j + k];
            vsc[k] += R[3 * r + k] * acc;
            vR[3 * r + k] = acc * s[k];
        }
    // ---- R(qn), qn = q/|q| ----
    const float g0 = 2.f * (-z * vR[1] + y * vR[2] + z * vR[3] - x * vR[5] - y * vR[6] + x * vR[7]);
    const float g1 = 2.f * (y * vR[1] + z * vR[2] + y * vR[3] - 2.f * x * vR[4] - w * vR[5] + z * vR[6] + w * vR[7] - 2.f * x * vR[8]);
    const float g2 = 2.f * (-2.f * y * vR[0] + x * vR[1] + w * vR[2] + x * vR[3] + z * vR[5] - w * vR[6] + z * vR[7] - 2.f * y * vR[8]);
    const float g3 = 2.f * (-2.f * z * vR[0] - w * vR[1] + x * vR[2] + w * vR[3] - 2.f * z * vR[4] + y * vR[5] + x * vR[6] + y * vR[7]);
    const float dot = g0 * w + g1 * x + g2 * y + g3 * z;
    g.vq = make_float4((g0 - w * dot) * inv, (g1 - x * dot) * inv, (g2 - y * dot) * inv, (g3 - z * dot) * inv);
    g.vp[0] = vp0; g.vp[1] = vp1; g.vp[2] = vp2;
    g.vs[0] = vsc[0] * glob; g.vs[1] = vsc[1] * glob; g.vs[2] = vsc[2] * glob;
}

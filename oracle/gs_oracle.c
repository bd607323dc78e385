/*
 * gs_oracle.c — CPU ORACLE (test infrastructure, NOT product code). See gs_oracle.h.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC
 * (-ffp-contract=off matters: the integer outputs radii / tiles_hit / keys are
 *  compared BIT-EXACTLY with the CUDA path, which compiles its projection
 *  kernel with -fmad=false and uses the same fp32 operation order.)
 *
 * Follows SURVEY.md Appendix A (gsplat 0.1.2 semantics, recalled) for the
 * calls made at /root/reference/gaussctrl/gc_model.py:174-188 (A.1),
 * :199-200 (A.2), :208-236 (A.3-A.5). PARITY UNPINNED (no reference goldens).
 */
#include "gs_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16
#define ALPHA_MAX 0.999f
#define ALPHA_MIN (1.0f / 255.0f)
#define T_EPS 1e-4f
#define BLUR 0.3f

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

static inline float clampf(float v, float lo, float hi) { return fminf(hi, fmaxf(lo, v)); }
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* A.1.6 tile bbox from pixel centre + radius; C truncation toward zero. */
static inline void tile_bbox(float x, float y, float radius, int tiles_x, int tiles_y,
                             int *x0, int *y0, int *x1, int *y1) {
    float tcx = x / (float)TILE, tcy = y / (float)TILE, tr = radius / (float)TILE;
    *x0 = clampi((int)(tcx - tr), 0, tiles_x);
    *x1 = clampi((int)(tcx + tr + 1.0f), 0, tiles_x);
    *y0 = clampi((int)(tcy - tr), 0, tiles_y);
    *y1 = clampi((int)(tcy + tr + 1.0f), 0, tiles_y);
}

/* normalised quaternion (w,x,y,z) -> row-major rotation matrix */
static inline void quat_to_R(const float *q, float *R, float *qn /*4, may be NULL*/, float *inv_norm) {
    float w = q[0], x = q[1], y = q[2], z = q[3];
    float nrm = sqrtf(w * w + x * x + y * y + z * z);
    float inv = 1.0f / nrm;
    w = w * inv; x = x * inv; y = y * inv; z = z * inv;
    if (qn) { qn[0] = w; qn[1] = x; qn[2] = y; qn[3] = z; }
    if (inv_norm) *inv_norm = inv;
    R[0] = 1.0f - 2.0f * (y * y + z * z);
    R[1] = 2.0f * (x * y - w * z);
    R[2] = 2.0f * (x * z + w * y);
    R[3] = 2.0f * (x * y + w * z);
    R[4] = 1.0f - 2.0f * (x * x + z * z);
    R[5] = 2.0f * (y * z - w * x);
    R[6] = 2.0f * (x * z - w * y);
    R[7] = 2.0f * (y * z + w * x);
    R[8] = 1.0f - 2.0f * (x * x + y * y);
}

/* Shared by fwd and bwd: everything up to (a,b,c) of the blurred 2D covariance. */
typedef struct {
    float t[3];        /* view-space point (unclamped) */
    float tc[2];       /* clamped x,y used by the Jacobian */
    int clamped[2];    /* -1/0/+1: which side the fov clamp was active */
    float lim[2];
    float R[9], M[9], S[6]; /* rotation, R*diag(s), cov3d upper triangle */
    float Tm[6];       /* 2x3 = J * Wrot */
    float a, b, c;
} proj_mid;

static inline void proj_front(const float *p, const float *s, float glob, const float *q,
                              const float *V, float fx, float fy, int H, int W, proj_mid *m) {
    m->t[0] = V[0] * p[0] + V[1] * p[1] + V[2] * p[2] + V[3];
    m->t[1] = V[4] * p[0] + V[5] * p[1] + V[6] * p[2] + V[7];
    m->t[2] = V[8] * p[0] + V[9] * p[1] + V[10] * p[2] + V[11];
    quat_to_R(q, m->R, NULL, NULL);
    float sx = glob * s[0], sy = glob * s[1], sz = glob * s[2];
    float *M = m->M, *R = m->R;
    M[0] = R[0] * sx; M[1] = R[1] * sy; M[2] = R[2] * sz;
    M[3] = R[3] * sx; M[4] = R[4] * sy; M[5] = R[5] * sz;
    M[6] = R[6] * sx; M[7] = R[7] * sy; M[8] = R[8] * sz;
    float *S = m->S;
    S[0] = M[0] * M[0] + M[1] * M[1] + M[2] * M[2];
    S[1] = M[0] * M[3] + M[1] * M[4] + M[2] * M[5];
    S[2] = M[0] * M[6] + M[1] * M[7] + M[2] * M[8];
    S[3] = M[3] * M[3] + M[4] * M[4] + M[5] * M[5];
    S[4] = M[3] * M[6] + M[4] * M[7] + M[5] * M[8];
    S[5] = M[6] * M[6] + M[7] * M[7] + M[8] * M[8];
}

static inline void proj_ewa(const float *V, float fx, float fy, int H, int W, proj_mid *m) {
    float tz = m->t[2];
    m->lim[0] = 1.3f * (0.5f * (float)W / fx);
    m->lim[1] = 1.3f * (0.5f * (float)H / fy);
    float rx = m->t[0] / tz, ry = m->t[1] / tz;
    m->clamped[0] = rx > m->lim[0] ? 1 : (rx < -m->lim[0] ? -1 : 0);
    m->clamped[1] = ry > m->lim[1] ? 1 : (ry < -m->lim[1] ? -1 : 0);
    m->tc[0] = tz * clampf(rx, -m->lim[0], m->lim[0]);
    m->tc[1] = tz * clampf(ry, -m->lim[1], m->lim[1]);
    float rz = 1.0f / tz, rz2 = rz * rz;
    float J00 = fx * rz, J02 = -fx * m->tc[0] * rz2;
    float J11 = fy * rz, J12 = -fy * m->tc[1] * rz2;
    float *Tm = m->Tm;
    Tm[0] = J00 * V[0] + J02 * V[8];
    Tm[1] = J00 * V[1] + J02 * V[9];
    Tm[2] = J00 * V[2] + J02 * V[10];
    Tm[3] = J11 * V[4] + J12 * V[8];
    Tm[4] = J11 * V[5] + J12 * V[9];
    Tm[5] = J11 * V[6] + J12 * V[10];
    const float *S = m->S;
    float u0 = S[0] * Tm[0] + S[1] * Tm[1] + S[2] * Tm[2];
    float u1 = S[1] * Tm[0] + S[3] * Tm[1] + S[4] * Tm[2];
    float u2 = S[2] * Tm[0] + S[4] * Tm[1] + S[5] * Tm[2];
    float w0 = S[0] * Tm[3] + S[1] * Tm[4] + S[2] * Tm[5];
    float w1 = S[1] * Tm[3] + S[3] * Tm[4] + S[4] * Tm[5];
    float w2 = S[2] * Tm[3] + S[4] * Tm[4] + S[5] * Tm[5];
    m->a = (Tm[0] * u0 + Tm[1] * u1 + Tm[2] * u2) + BLUR;
    m->b = Tm[3] * u0 + Tm[4] * u1 + Tm[5] * u2;
    m->c = (Tm[3] * w0 + Tm[4] * w1 + Tm[5] * w2) + BLUR;
}

void orc_project_fwd(int N, const float *means, const float *scales, float glob_scale,
                     const float *quats, const float *V, const float *P,
                     float fx, float fy, float cx, float cy, int H, int W,
                     int tiles_x, int tiles_y, float clip,
                     float *cov3d, float *xys, float *depths, int32_t *radii,
                     float *conics, int32_t *tiles_hit) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        const float *p = means + 3 * i;
        memset(cov3d + 6 * i, 0, 6 * sizeof(float));
        xys[2 * i] = xys[2 * i + 1] = 0.0f;
        depths[i] = 0.0f; radii[i] = 0; tiles_hit[i] = 0;
        conics[3 * i] = conics[3 * i + 1] = conics[3 * i + 2] = 0.0f;

        proj_mid m;
        proj_front(p, scales + 3 * i, glob_scale, quats + 4 * i, V, fx, fy, H, W, &m);
        if (m.t[2] <= clip) continue;                      /* A.1.1 near clip */
        memcpy(cov3d + 6 * i, m.S, 6 * sizeof(float));     /* A.1.2 */
        proj_ewa(V, fx, fy, H, W, &m);                     /* A.1.3 */
        float a = m.a, b = m.b, c = m.c;
        float det = a * c - b * b;                         /* A.1.4 */
        if (det == 0.0f) continue;
        float inv_det = 1.0f / det;
        float mid = 0.5f * (a + c);
        float disc = sqrtf(fmaxf(0.1f, mid * mid - det));
        float v1 = mid + disc, v2 = mid - disc;
        float radius = ceilf(3.0f * sqrtf(fmaxf(v1, v2)));
        /* A.1.5 pixel centre through the full projection matrix */
        float hx = P[0] * p[0] + P[1] * p[1] + P[2] * p[2] + P[3];
        float hy = P[4] * p[0] + P[5] * p[1] + P[6] * p[2] + P[7];
        float hw = P[12] * p[0] + P[13] * p[1] + P[14] * p[2] + P[15];
        float rw = 1.0f / (hw + 1e-6f);
        float x = 0.5f * (float)W * (hx * rw) + cx - 0.5f;
        float y = 0.5f * (float)H * (hy * rw) + cy - 0.5f;
        int x0, y0, x1, y1;                                /* A.1.6 */
        tile_bbox(x, y, radius, tiles_x, tiles_y, &x0, &y0, &x1, &y1);
        int area = (x1 - x0) * (y1 - y0);
        if (area <= 0) continue;
        tiles_hit[i] = area;                               /* A.1.7 */
        depths[i] = m.t[2];
        radii[i] = (int32_t)radius;
        xys[2 * i] = x; xys[2 * i + 1] = y;
        conics[3 * i] = c * inv_det;
        conics[3 * i + 1] = -b * inv_det;
        conics[3 * i + 2] = a * inv_det;
    }
}

/* ------------------------------------------------------------------------- */
/* Projection backward: exact derivative of the forward above, evaluated in   */
/* double from the fp32 inputs (the oracle's best estimate of the true vjp).  */
/* SURVEY.md App. A.6 lists where the real gsplat-0.1.2 binary is believed to */
/* deviate from this; those deviations are documented, not reproduced.        */
/* ------------------------------------------------------------------------- */
void orc_project_bwd(int N, const float *means, const float *scales, float glob_scale,
                     const float *quats, const float *Vf, const float *Pf,
                     float fxf, float fyf, float cx, float cy, int H, int W,
                     const int32_t *radii,
                     const float *v_xys, const float *v_depths, const float *v_conics,
                     float *v_means, float *v_scales, float *v_quats) {
    (void)cx; (void)cy;
    double V[12], P[16];
    for (int k = 0; k < 12; ++k) V[k] = Vf[k];
    for (int k = 0; k < 16; ++k) P[k] = Pf[k];
    const double fx = fxf, fy = fyf, glob = glob_scale;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float *vm = v_means + 3 * i, *vs = v_scales + 3 * i, *vq = v_quats + 4 * i;
        vm[0] = vm[1] = vm[2] = 0.f; vs[0] = vs[1] = vs[2] = 0.f;
        vq[0] = vq[1] = vq[2] = vq[3] = 0.f;
        if (radii[i] <= 0) continue;
        const double p[3] = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
        double vp[3] = {0, 0, 0};

        /* --- pixel centre: xy = 0.5*W*hx*rw + cx - 0.5 --- */
        double hx = P[0] * p[0] + P[1] * p[1] + P[2] * p[2] + P[3];
        double hy = P[4] * p[0] + P[5] * p[1] + P[6] * p[2] + P[7];
        double hw = P[12] * p[0] + P[13] * p[1] + P[14] * p[2] + P[15];
        double rw = 1.0 / (hw + 1e-6);
        double vx = v_xys[2 * i], vy = v_xys[2 * i + 1];
        double v_hx = 0.5 * W * rw * vx, v_hy = 0.5 * H * rw * vy;
        double v_rw = 0.5 * W * hx * vx + 0.5 * H * hy * vy;
        double v_hw = -rw * rw * v_rw;
        for (int j = 0; j < 3; ++j) vp[j] += P[j] * v_hx + P[4 + j] * v_hy + P[12 + j] * v_hw;

        /* --- recompute forward mid-state in double --- */
        double t[3];
        for (int r = 0; r < 3; ++r) t[r] = V[4 * r] * p[0] + V[4 * r + 1] * p[1] + V[4 * r + 2] * p[2] + V[4 * r + 3];
        double qw = quats[4 * i], qx = quats[4 * i + 1], qy = quats[4 * i + 2], qz = quats[4 * i + 3];
        double inv = 1.0 / sqrt(qw * qw + qx * qx + qy * qy + qz * qz);
        double w = qw * inv, x = qx * inv, y = qy * inv, z = qz * inv;
        double R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                       2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                       2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
        double s[3] = {glob * scales[3 * i], glob * scales[3 * i + 1], glob * scales[3 * i + 2]};
        double M[9], S[9];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) M[3 * r + c] = R[3 * r + c] * s[c];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c)
            S[3 * r + c] = M[3 * r] * M[3 * c] + M[3 * r + 1] * M[3 * c + 1] + M[3 * r + 2] * M[3 * c + 2];
        double tz = t[2], rz = 1.0 / tz, rz2 = rz * rz;
        double lim[2] = {1.3 * (0.5 * W / fx), 1.3 * (0.5 * H / fy)};
        double tcl[2]; int cl[2];
        for (int k = 0; k < 2; ++k) {
            /* clamp decision taken in fp32, as the forward does */
            float rf = (float)t[k] / (float)tz, lf = (float)lim[k];
            cl[k] = rf > lf ? 1 : (rf < -lf ? -1 : 0);
            tcl[k] = cl[k] == 0 ? t[k] : tz * (cl[k] * lim[k]);
        }
        double J00 = fx * rz, J02 = -fx * tcl[0] * rz2, J11 = fy * rz, J12 = -fy * tcl[1] * rz2;
        double Tm[6] = {J00 * V[0] + J02 * V[8], J00 * V[1] + J02 * V[9], J00 * V[2] + J02 * V[10],
                        J11 * V[4] + J12 * V[8], J11 * V[5] + J12 * V[9], J11 * V[6] + J12 * V[10]};
        double ST0[3], ST1[3];
        for (int r = 0; r < 3; ++r) {
            ST0[r] = S[3 * r] * Tm[0] + S[3 * r + 1] * Tm[1] + S[3 * r + 2] * Tm[2];
            ST1[r] = S[3 * r] * Tm[3] + S[3 * r + 1] * Tm[4] + S[3 * r + 2] * Tm[5];
        }
        double a = Tm[0] * ST0[0] + Tm[1] * ST0[1] + Tm[2] * ST0[2] + (double)BLUR;
        double b = Tm[3] * ST0[0] + Tm[4] * ST0[1] + Tm[5] * ST0[2];
        double c = Tm[3] * ST1[0] + Tm[4] * ST1[1] + Tm[5] * ST1[2] + (double)BLUR;
        double det = a * c - b * b, id2 = 1.0 / (det * det);

        /* --- conic = (c, -b, a)/det  ->  (v_a, v_b, v_c) --- */
        double vA = v_conics[3 * i], vB = v_conics[3 * i + 1], vC = v_conics[3 * i + 2];
        double v_a = (-c * c * vA + b * c * vB - b * b * vC) * id2;
        double v_b = (2 * b * c * vA - (det + 2 * b * b) * vB + 2 * a * b * vC) * id2;
        double v_c = (-b * b * vA + a * b * vB - a * a * vC) * id2;

        /* --- cov2d = Tm S Tm^T --- */
        double vS[9], vT[6];
        for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q)
            vS[3 * r + q] = v_a * Tm[r] * Tm[q] + v_b * Tm[r] * Tm[3 + q] + v_c * Tm[3 + r] * Tm[3 + q];
        for (int k = 0; k < 3; ++k) {
            vT[k] = 2 * v_a * ST0[k] + v_b * ST1[k];
            vT[3 + k] = 2 * v_c * ST1[k] + v_b * ST0[k];
        }
        /* Tm = J Wrot -> v_J = vT Wrot^T (only 4 live entries) */
        double vJ00 = vT[0] * V[0] + vT[1] * V[1] + vT[2] * V[2];
        double vJ02 = vT[0] * V[8] + vT[1] * V[9] + vT[2] * V[10];
        double vJ11 = vT[3] * V[4] + vT[4] * V[5] + vT[5] * V[6];
        double vJ12 = vT[3] * V[8] + vT[4] * V[9] + vT[5] * V[10];
        double rz3 = rz2 * rz;
        double v_tcx = -fx * rz2 * vJ02, v_tcy = -fy * rz2 * vJ12;
        double v_t[3];
        v_t[2] = -fx * rz2 * vJ00 - fy * rz2 * vJ11 + 2 * fx * tcl[0] * rz3 * vJ02 + 2 * fy * tcl[1] * rz3 * vJ12;
        v_t[0] = cl[0] == 0 ? v_tcx : 0.0;
        v_t[1] = cl[1] == 0 ? v_tcy : 0.0;
        if (cl[0] != 0) v_t[2] += cl[0] * lim[0] * v_tcx;
        if (cl[1] != 0) v_t[2] += cl[1] * lim[1] * v_tcy;
        v_t[2] += v_depths[i];
        for (int j = 0; j < 3; ++j) vp[j] += V[j] * v_t[0] + V[4 + j] * v_t[1] + V[8 + j] * v_t[2];

        /* --- S = M M^T, M = R diag(s) --- */
        double vM[9], vR[9], vsc[3] = {0, 0, 0};
        for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) {
            double acc = 0;
            for (int k = 0; k < 3; ++k) acc += (vS[3 * r + k] + vS[3 * k + r]) * M[3 * k + q];
            vM[3 * r + q] = acc;
        }
        for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) {
            vsc[q] += R[3 * r + q] * vM[3 * r + q];
            vR[3 * r + q] = vM[3 * r + q] * s[q];
        }
        /* --- R(qn) --- */
        double g[4];
        g[0] = 2 * (-z * vR[1] + y * vR[2] + z * vR[3] - x * vR[5] - y * vR[6] + x * vR[7]);
        g[1] = 2 * (y * vR[1] + z * vR[2] + y * vR[3] - 2 * x * vR[4] - w * vR[5] + z * vR[6] + w * vR[7] - 2 * x * vR[8]);
        g[2] = 2 * (-2 * y * vR[0] + x * vR[1] + w * vR[2] + x * vR[3] + z * vR[5] - w * vR[6] + z * vR[7] - 2 * y * vR[8]);
        g[3] = 2 * (-2 * z * vR[0] - w * vR[1] + x * vR[2] + w * vR[3] - 2 * z * vR[4] + y * vR[5] + x * vR[6] + y * vR[7]);
        /* qn = q/|q| */
        double dot = g[0] * w + g[1] * x + g[2] * y + g[3] * z;
        vq[0] = (float)((g[0] - w * dot) * inv);
        vq[1] = (float)((g[1] - x * dot) * inv);
        vq[2] = (float)((g[2] - y * dot) * inv);
        vq[3] = (float)((g[3] - z * dot) * inv);
        for (int j = 0; j < 3; ++j) { vm[j] = (float)vp[j]; vs[j] = (float)(vsc[j] * glob); }
    }
}

/* ------------------------------------------------------------------------- */
/* A.2 spherical harmonics (real SH, Inria / svox2 constants), deg <= 4.      */
/* ------------------------------------------------------------------------- */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};
static const float SH_C4[9] = {2.5033429417967046f, -1.7701307697799304f, 0.9461746957575601f,
                               -0.6690465435572892f, 0.10578554691520431f, -0.6690465435572892f,
                               0.47308734787878004f, -1.7701307697799304f, 0.6258357354491761f};

/* basis values Y[0..(deg+1)^2) for direction d (normalised here, as the kernel does) */
static inline int sh_basis(int deg, const float *d, float *Y) {
    float nrm = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    float inv = 1.0f / nrm;
    float x = d[0] * inv, y = d[1] * inv, z = d[2] * inv;
    Y[0] = SH_C0;
    if (deg < 1) return 1;
    Y[1] = -SH_C1 * y; Y[2] = SH_C1 * z; Y[3] = -SH_C1 * x;
    if (deg < 2) return 4;
    float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    Y[4] = SH_C2[0] * xy; Y[5] = SH_C2[1] * yz; Y[6] = SH_C2[2] * (2.0f * zz - xx - yy);
    Y[7] = SH_C2[3] * xz; Y[8] = SH_C2[4] * (xx - yy);
    if (deg < 3) return 9;
    Y[9] = SH_C3[0] * y * (3.0f * xx - yy);
    Y[10] = SH_C3[1] * xy * z;
    Y[11] = SH_C3[2] * y * (4.0f * zz - xx - yy);
    Y[12] = SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
    Y[13] = SH_C3[4] * x * (4.0f * zz - xx - yy);
    Y[14] = SH_C3[5] * z * (xx - yy);
    Y[15] = SH_C3[6] * x * (xx - 3.0f * yy);
    if (deg < 4) return 16;
    Y[16] = SH_C4[0] * xy * (xx - yy);
    Y[17] = SH_C4[1] * yz * (3.0f * xx - yy);
    Y[18] = SH_C4[2] * xy * (7.0f * zz - 1.0f);
    Y[19] = SH_C4[3] * yz * (7.0f * zz - 3.0f);
    Y[20] = SH_C4[4] * (zz * (35.0f * zz - 30.0f) + 3.0f);
    Y[21] = SH_C4[5] * xz * (7.0f * zz - 3.0f);
    Y[22] = SH_C4[6] * (xx - yy) * (7.0f * zz - 1.0f);
    Y[23] = SH_C4[7] * xz * (xx - 3.0f * yy);
    Y[24] = SH_C4[8] * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy));
    return 25;
}

void orc_sh_fwd(int N, int K, int deg, const float *dirs, const float *coeffs, float *colors) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float Y[25];
        int nb = sh_basis(deg, dirs + 3 * i, Y);
        const float *c = coeffs + (size_t)i * K * 3;
        double acc[3] = {0, 0, 0};
        for (int k = 0; k < nb; ++k)
            for (int ch = 0; ch < 3; ++ch) acc[ch] += (double)Y[k] * (double)c[3 * k + ch];
        for (int ch = 0; ch < 3; ++ch) colors[3 * i + ch] = (float)acc[ch];
    }
}

void orc_sh_bwd(int N, int K, int deg, const float *dirs, const float *v_colors, float *v_coeffs) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float Y[25];
        int nb = sh_basis(deg, dirs + 3 * i, Y);
        float *vc = v_coeffs + (size_t)i * K * 3;
        for (int k = 0; k < K; ++k)
            for (int ch = 0; ch < 3; ++ch) vc[3 * k + ch] = k < nb ? Y[k] * v_colors[3 * i + ch] : 0.0f;
    }
}

/* ------------------------------------------------------------------------- */
/* A.3 binning: scan, key emission, stable sort, tile ranges. Integer-exact.  */
/* ------------------------------------------------------------------------- */
int64_t orc_scan(int N, const int32_t *tiles_hit, int32_t *cum) {
    int64_t run = 0;
    for (int i = 0; i < N; ++i) { run += tiles_hit[i]; cum[i] = (int32_t)run; }
    return run;
}

void orc_emit(int N, const float *xys, const float *depths, const int32_t *radii,
              const int32_t *cum, int tiles_x, int tiles_y, int64_t *keys, int32_t *ids) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        if (radii[i] <= 0) continue;
        int x0, y0, x1, y1;
        tile_bbox(xys[2 * i], xys[2 * i + 1], (float)radii[i], tiles_x, tiles_y, &x0, &y0, &x1, &y1);
        int64_t at = i == 0 ? 0 : cum[i - 1];
        uint32_t dbits;
        memcpy(&dbits, depths + i, 4);
        for (int ty = y0; ty < y1; ++ty)
            for (int tx = x0; tx < x1; ++tx) {
                int64_t tile = (int64_t)ty * tiles_x + tx;
                keys[at] = (tile << 32) | (int64_t)dbits;
                ids[at] = i;
                ++at;
            }
    }
}

void orc_sort(int64_t I, const int64_t *keys, const int32_t *ids, int64_t *ko, int32_t *io) {
    /* stable LSD radix sort, 16-bit digits over bits [0,64) (keys are non-negative) */
    if (I <= 0) return;
    int64_t *ka = (int64_t *)malloc(sizeof(int64_t) * I), *kb = (int64_t *)malloc(sizeof(int64_t) * I);
    int32_t *ia = (int32_t *)malloc(sizeof(int32_t) * I), *ib = (int32_t *)malloc(sizeof(int32_t) * I);
    int64_t *cnt = (int64_t *)malloc(sizeof(int64_t) * 65537);
    memcpy(ka, keys, sizeof(int64_t) * I);
    memcpy(ia, ids, sizeof(int32_t) * I);
    for (int pass = 0; pass < 4; ++pass) {
        int sh = 16 * pass;
        memset(cnt, 0, sizeof(int64_t) * 65537);
        for (int64_t k = 0; k < I; ++k) cnt[((ka[k] >> sh) & 0xFFFF) + 1]++;
        for (int d = 0; d < 65536; ++d) cnt[d + 1] += cnt[d];
        for (int64_t k = 0; k < I; ++k) {
            int64_t at = cnt[(ka[k] >> sh) & 0xFFFF]++;
            kb[at] = ka[k]; ib[at] = ia[k];
        }
        int64_t *tk = ka; ka = kb; kb = tk;
        int32_t *ti = ia; ia = ib; ib = ti;
    }
    memcpy(ko, ka, sizeof(int64_t) * I);
    memcpy(io, ia, sizeof(int32_t) * I);
    free(ka); free(kb); free(ia); free(ib); free(cnt);
}

void orc_tile_bins(int64_t I, int T, const int64_t *ks, int32_t *bins) {
    memset(bins, 0, sizeof(int32_t) * 2 * (size_t)T);
    for (int64_t k = 0; k < I; ++k) {
        int32_t t = (int32_t)(ks[k] >> 32);
        if (k == 0 || (int32_t)(ks[k - 1] >> 32) != t) bins[2 * t] = (int32_t)k;
        if (k == I - 1 || (int32_t)(ks[k + 1] >> 32) != t) bins[2 * t + 1] = (int32_t)(k + 1);
    }
}

/* ------------------------------------------------------------------------- */
/* A.4 compositing forward. Pixel centres are the integer coordinates (j,i)   */
/* (SURVEY.md App. A.4, U1). One OpenMP task per tile; inside, the literal    */
/* per-pixel front-to-back loop.                                              */
/* ------------------------------------------------------------------------- */
#define PIX_CENTER 0.0f

int64_t orc_raster_fwd(int C, int H, int W, int tiles_x, int tiles_y,
                       const int32_t *ids_sorted, const int32_t *tile_bins,
                       const float *xys, const float *conics, const float *colors,
                       const float *opac, const float *bg,
                       float *out_img, float *final_T, int32_t *final_idx) {
    int64_t pairs = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : pairs)
    for (int tile = 0; tile < tiles_x * tiles_y; ++tile) {
        int tx = tile % tiles_x, ty = tile / tiles_x;
        int s = tile_bins[2 * tile], e = tile_bins[2 * tile + 1];
        float *acc = (float *)malloc(sizeof(float) * (size_t)C);
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int i = ty * TILE + ly, j = tx * TILE + lx;
                if (i >= H || j >= W) continue;
                float px = (float)j + PIX_CENTER, py = (float)i + PIX_CENTER;
                float T = 1.0f;
                int last = s;
                for (int ch = 0; ch < C; ++ch) acc[ch] = 0.0f;
                for (int k = s; k < e; ++k) {
                    ++pairs;
                    int g = ids_sorted[k];
                    float dx = xys[2 * g] - px, dy = xys[2 * g + 1] - py;
                    float ca = conics[3 * g], cb = conics[3 * g + 1], cc = conics[3 * g + 2];
                    float sigma = 0.5f * (ca * dx * dx + cc * dy * dy) + cb * dx * dy;
                    if (sigma < 0.0f) continue;
                    float alpha = fminf(ALPHA_MAX, opac[g] * expf(-sigma));
                    if (alpha < ALPHA_MIN) continue;
                    float Tn = T * (1.0f - alpha);
                    if (Tn <= T_EPS) break;
                    float vis = alpha * T;
                    for (int ch = 0; ch < C; ++ch) acc[ch] += vis * colors[(size_t)g * C + ch];
                    T = Tn;
                    last = k + 1;
                }
                size_t pix = (size_t)i * W + j;
                for (int ch = 0; ch < C; ++ch) out_img[pix * C + ch] = acc[ch] + T * bg[ch];
                final_T[pix] = T;
                final_idx[pix] = last;
            }
        free(acc);
    }
    return pairs;
}

/* ------------------------------------------------------------------------- */
/* A.5 compositing backward: back-to-front replay from final_idx, exact       */
/* derivative of the forward (alpha clamp has zero slope where active).       */
/* Per-Gaussian sums are accumulated in double.                               */
/* ------------------------------------------------------------------------- */
void orc_raster_bwd(int C, int H, int W, int tiles_x, int tiles_y, int N,
                    const int32_t *ids_sorted, const int32_t *tile_bins,
                    const float *xys, const float *conics, const float *colors,
                    const float *opac, const float *bg,
                    const float *final_T, const int32_t *final_idx,
                    const float *v_out, const float *v_out_alpha,
                    float *v_xys, float *v_conics, float *v_colors, float *v_opac) {
    size_t per = (size_t)(6 + C);
    double *G = (double *)calloc((size_t)N * per, sizeof(double));
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < tiles_x * tiles_y; ++tile) {
        int tx = tile % tiles_x, ty = tile / tiles_x;
        int s = tile_bins[2 * tile], e = tile_bins[2 * tile + 1];
        if (e <= s) continue;
        int cnt = e - s;
        double *L = (double *)calloc((size_t)cnt * per, sizeof(double));
        double *S = (double *)malloc(sizeof(double) * (size_t)C);
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int i = ty * TILE + ly, j = tx * TILE + lx;
                if (i >= H || j >= W) continue;
                size_t pix = (size_t)i * W + j;
                float px = (float)j + PIX_CENTER, py = (float)i + PIX_CENTER;
                double Tf = final_T[pix], T = Tf;
                const float *vo = v_out + pix * C;
                double voa = v_out_alpha ? (double)v_out_alpha[pix] : 0.0;
                double bgdot = 0.0;
                for (int ch = 0; ch < C; ++ch) { S[ch] = 0.0; bgdot += (double)bg[ch] * vo[ch]; }
                for (int k = final_idx[pix] - 1; k >= s; --k) {
                    int g = ids_sorted[k];
                    float dxf = xys[2 * g] - px, dyf = xys[2 * g + 1] - py;
                    float ca = conics[3 * g], cb = conics[3 * g + 1], cc = conics[3 * g + 2];
                    float sigma = 0.5f * (ca * dxf * dxf + cc * dyf * dyf) + cb * dxf * dyf;
                    if (sigma < 0.0f) continue;
                    float visf = expf(-sigma);
                    float alphaf = fminf(ALPHA_MAX, opac[g] * visf);
                    if (alphaf < ALPHA_MIN) continue;
                    double alpha = alphaf, ra = 1.0 / (1.0 - alpha);
                    T *= ra;                                   /* T_k */
                    double fac = alpha * T, v_alpha = 0.0;
                    double *Lg = L + (size_t)(k - s) * per;
                    const float *col = colors + (size_t)g * C;
                    for (int ch = 0; ch < C; ++ch) {
                        Lg[6 + ch] += fac * vo[ch];
                        v_alpha += ((double)col[ch] * T - S[ch] * ra) * vo[ch];
                        S[ch] += (double)col[ch] * fac;
                    }
                    v_alpha += Tf * ra * voa;
                    v_alpha += -Tf * ra * bgdot;
                    if (opac[g] * visf <= ALPHA_MAX) {
                        double vis = visf, dx = dxf, dy = dyf;
                        double v_sigma = -(double)opac[g] * vis * v_alpha;
                        Lg[0] += v_sigma * ((double)ca * dx + (double)cb * dy);
                        Lg[1] += v_sigma * ((double)cb * dx + (double)cc * dy);
                        Lg[2] += 0.5 * v_sigma * dx * dx;
                        Lg[3] += v_sigma * dx * dy;
                        Lg[4] += 0.5 * v_sigma * dy * dy;
                        Lg[5] += vis * v_alpha;
                    }
                }
            }
        for (int k = 0; k < cnt; ++k) {
            int g = ids_sorted[s + k];
            for (size_t c = 0; c < per; ++c) {
                double v = L[(size_t)k * per + c];
                if (v != 0.0) {
#pragma omp atomic
                    G[(size_t)g * per + c] += v;
                }
            }
        }
        free(L); free(S);
    }
    for (int g = 0; g < N; ++g) {
        const double *Gg = G + (size_t)g * per;
        v_xys[2 * g] = (float)Gg[0]; v_xys[2 * g + 1] = (float)Gg[1];
        v_conics[3 * g] = (float)Gg[2]; v_conics[3 * g + 1] = (float)Gg[3]; v_conics[3 * g + 2] = (float)Gg[4];
        v_opac[g] = (float)Gg[5];
        for (int ch = 0; ch < C; ++ch) v_colors[(size_t)g * C + ch] = (float)Gg[6 + ch];
    }
    free(G);
}

/* ------------------------------------------------------------------------- */
/* Test diagnostics for the parity tests at full size (tests/test_parity_scale_gpu.py).          */
/* The forward of A.4 takes three discontinuous decisions per (pixel, Gaussian) pair:             */
/*   alpha >= 1/255 (skip), T(1-alpha) <= 1e-4 (stop), opac*exp(-sigma) vs 0.999 (clamp: only the */
/*   derivative is discontinuous). Two correct fp32 implementations (expf here, ex2.approx on the */
/*   GPU, gsplat's __expf) can take a decision differently only where the quantity sits within    */
/*   rounding distance of its threshold. orc_raster_risk returns, per pixel, the smallest RELATIVE */
/*   distance |q/threshold - 1| over all decisions this oracle took at that pixel, so a test can   */
/*   assert "every pixel that deviates is one where a decision was within eps of its threshold".   */
/* ------------------------------------------------------------------------- */
void orc_raster_risk(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                     const int32_t *tile_bins, const float *xys, const float *conics,
                     const float *opac, float *risk_alpha, float *risk_T, float *risk_clamp) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < tiles_x * tiles_y; ++tile) {
        int tx = tile % tiles_x, ty = tile / tiles_x;
        int s = tile_bins[2 * tile], e = tile_bins[2 * tile + 1];
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int i = ty * TILE + ly, j = tx * TILE + lx;
                if (i >= H || j >= W) continue;
                float px = (float)j + PIX_CENTER, py = (float)i + PIX_CENTER;
                float T = 1.0f;
                double ra = 1e30, rt = 1e30, rc = 1e30;
                for (int k = s; k < e; ++k) {
                    int g = ids_sorted[k];
                    float dx = xys[2 * g] - px, dy = xys[2 * g + 1] - py;
                    float ca = conics[3 * g], cb = conics[3 * g + 1], cc = conics[3 * g + 2];
                    float sigma = 0.5f * (ca * dx * dx + cc * dy * dy) + cb * dx * dy;
                    if (sigma < 0.0f) continue;
                    double raw = (double)opac[g] * exp(-(double)sigma);
                    double da = fabs(raw / (double)ALPHA_MIN - 1.0);
                    if (da < ra) ra = da;
                    float alpha = fminf(ALPHA_MAX, opac[g] * expf(-sigma));
                    if (alpha < ALPHA_MIN) continue;
                    double dc = fabs(raw / (double)ALPHA_MAX - 1.0);
                    if (dc < rc) rc = dc;
                    float Tn = T * (1.0f - alpha);
                    double dt = fabs((double)Tn / (double)T_EPS - 1.0);
                    if (dt < rt) rt = dt;
                    if (Tn <= T_EPS) break;
                    T = Tn;
                }
                size_t pix = (size_t)i * W + j;
                risk_alpha[pix] = (float)ra; risk_T[pix] = (float)rt; risk_clamp[pix] = (float)rc;
            }
    }
}

/* touched[g] = 1 for every Gaussian that can influence (forward or backward) a pixel with mask != 0:
 * all list entries k in [range start, upto[pix]) whose raw alpha at that pixel is at least half the
 * 1/255 threshold. upto = per-pixel exclusive end index (max of the two implementations' final_idx). */
void orc_mark_contributors(int H, int W, int tiles_x, int tiles_y, const int32_t *ids_sorted,
                           const int32_t *tile_bins, const float *xys, const float *conics,
                           const float *opac, const uint8_t *mask, const int32_t *upto, uint8_t *touched) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int tile = 0; tile < tiles_x * tiles_y; ++tile) {
        int tx = tile % tiles_x, ty = tile / tiles_x;
        int s = tile_bins[2 * tile], e = tile_bins[2 * tile + 1];
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int i = ty * TILE + ly, j = tx * TILE + lx;
                if (i >= H || j >= W) continue;
                size_t pix = (size_t)i * W + j;
                if (!mask[pix]) continue;
                float px = (float)j + PIX_CENTER, py = (float)i + PIX_CENTER;
                int hi = upto[pix] < e ? upto[pix] : e;
                for (int k = s; k < hi; ++k) {
                    int g = ids_sorted[k];
                    float dx = xys[2 * g] - px, dy = xys[2 * g + 1] - py;
                    float ca = conics[3 * g], cb = conics[3 * g + 1], cc = conics[3 * g + 2];
                    float sigma = 0.5f * (ca * dx * dx + cc * dy * dy) + cb * dx * dy;
                    if (sigma < 0.0f) continue;
                    if (opac[g] * expf(-sigma) >= 0.5f * ALPHA_MIN) touched[g] = 1;   /* benign race: all writers store 1 */
                }
            }
    }
}

/* ------------------------------------------------------------------------------------------
 * Multi-view noise mask — CPU restatement of /root/reference/gaussctrl/ad_render.py:273-333
 * (render_noise_SS) and the GLSL shaders at :384-479, drawn in order on the CPU: instances in
 * buffer order, triangles in index order, later fragments overwrite earlier ones (the reference
 * never enables GL_DEPTH_TEST or blending). Coverage: pixel centres, window coordinates snapped
 * to 1/256 px, exact integer edge functions, left/top tie rule. Everything else fp32 in the
 * order written here (the build uses -ffp-contract=off). PARITY UNPINNED against a real GL
 * implementation (no graphics stack on this box); see gs_oracle.h.
 * view / proj: row-major 4x4. out_rgba (H,W,4) = the image render_noise_SS returns (after its
 * fliplr / flipud); depth (H,W) is the image handed to depth_np2texture (:104-123).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int ix, iy, ok;
    float invw, vx, vy, vz, cr, cg, cb;
} orc_nm_vert;

static unsigned orc_nm_u8(float c) {
    c = fminf(fmaxf(c, 0.0f), 1.0f);
    return (unsigned)floorf(c * 255.0f + 0.5f);
}

static int orc_nm_edge_in(long long e, int dx, int dy) {
    return e > 0 || (e == 0 && (dy < 0 || (dy == 0 && dx < 0)));
}

static int orc_floor_div256(int v) { return v >= 0 ? v / 256 : -((-v + 255) / 256); }

void orc_noise_mask(int P, const float *points, int V, const float *verts, int F, const int32_t *tris,
                    const float *Vm, const float *Pm, int H, int W, const float *depth, float threshold,
                    uint8_t *out_rgba) {
    memset(out_rgba, 0, (size_t)H * W * 4);
    orc_nm_vert *vs = (orc_nm_vert *)malloc(sizeof(orc_nm_vert) * (size_t)V);
    for (int inst = 0; inst < P; ++inst) {
        const float ox = points[3 * inst], oy = points[3 * inst + 1], oz = points[3 * inst + 2];
        for (int k = 0; k < V; ++k) {
            const float wx = verts[3 * k] + ox, wy = verts[3 * k + 1] + oy, wz = verts[3 * k + 2] + oz;
            orc_nm_vert o;
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
                if (fabsf(sx) < 1.0e5f && fabsf(sy) < 1.0e5f) {
                    o.ix = (int)floorf(sx * 256.0f + 0.5f);
                    o.iy = (int)floorf(sy * 256.0f + 0.5f);
                    o.ok = 1;
                }
            }
            vs[k] = o;
        }
        for (int f = 0; f < F; ++f) {
            int ia = tris[3 * f], ib = tris[3 * f + 1], ic = tris[3 * f + 2];
            if (ia < 0 || ia >= V || ib < 0 || ib >= V || ic < 0 || ic >= V) continue;
            if (!(vs[ia].ok && vs[ib].ok && vs[ic].ok)) continue;
            long long area2 = (long long)(vs[ib].ix - vs[ia].ix) * (vs[ic].iy - vs[ia].iy) -
                              (long long)(vs[ib].iy - vs[ia].iy) * (vs[ic].ix - vs[ia].ix);
            if (area2 == 0) continue;
            if (area2 < 0) { int t = ib; ib = ic; ic = t; area2 = -area2; }
            const orc_nm_vert a = vs[ia], b = vs[ib], c = vs[ic];
            int minx = a.ix < b.ix ? a.ix : b.ix, maxx = a.ix > b.ix ? a.ix : b.ix;
            int miny = a.iy < b.iy ? a.iy : b.iy, maxy = a.iy > b.iy ? a.iy : b.iy;
            if (c.ix < minx) minx = c.ix;
            if (c.ix > maxx) maxx = c.ix;
            if (c.iy < miny) miny = c.iy;
            if (c.iy > maxy) maxy = c.iy;
            int x0 = orc_floor_div256(minx - 128 + 255), x1 = orc_floor_div256(maxx - 128);
            int y0 = orc_floor_div256(miny - 128 + 255), y1 = orc_floor_div256(maxy - 128);
            if (x0 < 0) x0 = 0;
            if (y0 < 0) y0 = 0;
            if (x1 > W - 1) x1 = W - 1;
            if (y1 > H - 1) y1 = H - 1;
            const float farea = (float)area2;
            for (int y = y0; y <= y1; ++y) {
                const int py = y * 256 + 128;
                for (int x = x0; x <= x1; ++x) {
                    const int px = x * 256 + 128;
                    const long long e_bc = (long long)(c.ix - b.ix) * (py - b.iy) - (long long)(c.iy - b.iy) * (px - b.ix);
                    const long long e_ca = (long long)(a.ix - c.ix) * (py - c.iy) - (long long)(a.iy - c.iy) * (px - c.ix);
                    const long long e_ab = (long long)(b.ix - a.ix) * (py - a.iy) - (long long)(b.iy - a.iy) * (px - a.ix);
                    if (!orc_nm_edge_in(e_bc, c.ix - b.ix, c.iy - b.iy) || !orc_nm_edge_in(e_ca, a.ix - c.ix, a.iy - c.iy) ||
                        !orc_nm_edge_in(e_ab, b.ix - a.ix, b.iy - a.iy))
                        continue;
                    const float la = ((float)e_bc / farea) * a.invw, lb = ((float)e_ca / farea) * b.invw,
                                lc = ((float)e_ab / farea) * c.invw;
                    const float ls = la + lb + lc;
                    const float wa = la / ls, wb = lb / ls, wc = lc / ls;
                    const float qx = wa * a.vx + wb * b.vx + wc * c.vx;
                    const float qy = wa * a.vy + wb * b.vy + wc * c.vy;
                    const float qz = wa * a.vz + wb * b.vz + wc * c.vz;
                    const float zc = Pm[8] * qx + Pm[9] * qy + Pm[10] * qz + Pm[11];
                    const float wc4 = Pm[12] * qx + Pm[13] * qy + Pm[14] * qz + Pm[15];
                    if (zc < -wc4 || zc > wc4) continue;
                    const float dfrag = sqrtf(qx * qx + qy * qy + qz * qz);
                    const int r = H - 1 - y, cc = W - 1 - x;
                    const float dtex = fminf(fmaxf(depth[(size_t)r * W + cc], 0.0f), 1000.0f);
                    const float diff = dtex - dfrag;
                    if (diff < 0.0f || diff > threshold) continue;
                    uint8_t *o = out_rgba + ((size_t)r * W + cc) * 4;
                    o[0] = (uint8_t)orc_nm_u8(wa * a.cr + wb * b.cr + wc * c.cr);
                    o[1] = (uint8_t)orc_nm_u8(wa * a.cg + wb * b.cg + wc * c.cg);
                    o[2] = (uint8_t)orc_nm_u8(wa * a.cb + wb * b.cb + wc * c.cb);
                    o[3] = 255;
                }
            }
        }
    }
    free(vs);
}

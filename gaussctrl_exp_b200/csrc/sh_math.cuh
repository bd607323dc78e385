// Real spherical-harmonics basis shared by sh.cu and the fused preprocess kernels
// (Inria / svox2 constants).
#pragma once
#include "gs_common.cuh"

__device__ __forceinline__ int sh_num_bases(int deg) { return (deg + 1) * (deg + 1); }

// Real SH basis (Inria / svox2 constants). Y must hold >= 25 floats when deg == 4.
template <int MAXDEG>
__device__ __forceinline__ void sh_basis(int deg, float dx, float dy, float dz, float *Y) {
    const float inv = 1.0f / sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx * inv, y = dy * inv, z = dz * inv;
    Y[0] = 0.28209479177387814f;
    if (MAXDEG < 1 || deg < 1) return;
    Y[1] = -0.4886025119029199f * y;
    Y[2] = 0.4886025119029199f * z;
    Y[3] = -0.4886025119029199f * x;
    if (MAXDEG < 2 || deg < 2) return;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    Y[4] = 1.0925484305920792f * xy;
    Y[5] = -1.0925484305920792f * yz;
    Y[6] = 0.31539156525252005f * (2.0f * zz - xx - yy);
    Y[7] = -1.0925484305920792f * xz;
    Y[8] = 0.5462742152960396f * (xx - yy);
    if (MAXDEG < 3 || deg < 3) return;
    Y[9] = -0.5900435899266435f * y * (3.0f * xx - yy);
    Y[10] = 2.890611442640554f * xy * z;
    Y[11] = -0.4570457994644658f * y * (4.0f * zz - xx - yy);
    Y[12] = 0.3731763325901154f * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
    Y[13] = -0.4570457994644658f * x * (4.0f * zz - xx - yy);
    Y[14] = 1.445305721320277f * z * (xx - yy);
    Y[15] = -0.5900435899266435f * x * (xx - 3.0f * yy);
    if (MAXDEG < 4 || deg < 4) return;
    Y[16] = 2.5033429417967046f * xy * (xx - yy);
    Y[17] = -1.7701307697799304f * yz * (3.0f * xx - yy);
    Y[18] = 0.9461746957575601f * xy * (7.0f * zz - 1.0f);
    Y[19] = -0.6690465435572892f * yz * (7.0f * zz - 3.0f);
    Y[20] = 0.10578554691520431f * (zz * (35.0f * zz - 30.0f) + 3.0f);
    Y[21] = -0.6690465435572892f * xz * (7.0f * zz - 3.0f);
    Y[22] = 0.47308734787878004f * (xx - yy) * (7.0f * zz - 1.0f);
    Y[23] = -1.7701307697799304f * xz * (xx - 3.0f * yy);
    Y[24] = 0.6258357354491761f * (xx * (xx - 3.0f * yy) - yy * (3.0f * xx - yy));
}


// Stage 1b — spherical harmonics forward / backward (replaces
// gsplat's compute_sh_forward/backward called at
// /root/reference/gaussctrl/gc_model.py:199-200 and from gc_trainer.py:275).
//
// HBM-bound: deg 3 reads 192 B of coefficients per Gaussian. The (N,16,3) block of a
// 128-Gaussian CTA is one contiguous 24 KB span: it is moved with 16 B cp.async
// (LDGSTS, no register staging) into shared memory rows padded to 52 floats, which
// makes both the cooperative store and the per-thread LDS.128 row reads conflict-free.
// The backward does the mirror image: rows built in shared memory, streamed out as
// coalesced float4.
#include "sh_math.cuh"

namespace {

constexpr int SH_TPB = 128;     // Gaussians per CTA
constexpr int SH_ROW4 = 13;     // padded row pitch in float4 (12 payload + 1)

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// K == 16 fast path.
__global__ void __launch_bounds__(SH_TPB)
sh_fwd16_kernel(int N, int deg, const float *__restrict__ dirs, const float4 *__restrict__ coeffs,
                float *__restrict__ colors) {
    __shared__ float4 tile[SH_TPB * SH_ROW4];
    const int g0 = blockIdx.x * SH_TPB;
    const int rows = min(SH_TPB, N - g0);
    const float4 *src = coeffs + (size_t)g0 * 12;
    const int n4 = rows * 12;
    for (int f = threadIdx.x; f < n4; f += SH_TPB) cp_async16(&tile[(f / 12) * SH_ROW4 + (f % 12)], src + f);
    const int i = g0 + threadIdx.x;
    float Y[16];
    if (i < N) sh_basis<3>(deg, dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], Y);
    cp_async_commit_wait_all();
    __syncthreads();
    if (i >= N) return;
    const int nb = sh_num_bases(deg);
    float c[48];
    const float4 *row = &tile[threadIdx.x * SH_ROW4];
#pragma unroll
    for (int j = 0; j < 12; ++j) {
        const float4 v = row[j];
        c[4 * j] = v.x; c[4 * j + 1] = v.y; c[4 * j + 2] = v.z; c[4 * j + 3] = v.w;
    }
    float r = 0.f, g = 0.f, b = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k)
        if (k < nb) { r += Y[k] * c[3 * k]; g += Y[k] * c[3 * k + 1]; b += Y[k] * c[3 * k + 2]; }
    colors[3 * i] = r; colors[3 * i + 1] = g; colors[3 * i + 2] = b;
}

__global__ void __launch_bounds__(SH_TPB)
sh_bwd16_kernel(int N, int deg, const float *__restrict__ dirs, const float *__restrict__ v_colors,
                float4 *__restrict__ v_coeffs) {
    __shared__ float4 tile[SH_TPB * SH_ROW4];
    const int g0 = blockIdx.x * SH_TPB;
    const int rows = min(SH_TPB, N - g0);
    const int i = g0 + threadIdx.x;
    if (i < N) {
        float Y[16];
        sh_basis<3>(deg, dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], Y);
        const int nb = sh_num_bases(deg);
        const float vr = v_colors[3 * i], vg = v_colors[3 * i + 1], vb = v_colors[3 * i + 2];
        float c[48];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const float yk = k < nb ? Y[k] : 0.f;
            c[3 * k] = yk * vr; c[3 * k + 1] = yk * vg; c[3 * k + 2] = yk * vb;
        }
        float4 *row = &tile[threadIdx.x * SH_ROW4];
#pragma unroll
        for (int j = 0; j < 12; ++j) row[j] = make_float4(c[4 * j], c[4 * j + 1], c[4 * j + 2], c[4 * j + 3]);
    }
    __syncthreads();
    float4 *dst = v_coeffs + (size_t)g0 * 12;
    const int n4 = rows * 12;
    for (int f = threadIdx.x; f < n4; f += SH_TPB) __stcs(dst + f, tile[(f / 12) * SH_ROW4 + (f % 12)]);
}

// Generic path: any K, deg <= 4. One Gaussian per thread, strided accesses.
__global__ void __launch_bounds__(128)
sh_fwd_generic_kernel(int N, int K, int deg, const float *__restrict__ dirs, const float *__restrict__ coeffs,
                      float *__restrict__ colors) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float Y[25];
    sh_basis<4>(deg, dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], Y);
    const int nb = sh_num_bases(deg);
    const float *c = coeffs + (size_t)i * K * 3;
    float r = 0.f, g = 0.f, b = 0.f;
    for (int k = 0; k < nb; ++k) { r += Y[k] * c[3 * k]; g += Y[k] * c[3 * k + 1]; b += Y[k] * c[3 * k + 2]; }
    colors[3 * i] = r; colors[3 * i + 1] = g; colors[3 * i + 2] = b;
}

__global__ void __launch_bounds__(128)
sh_bwd_generic_kernel(int N, int K, int deg, const float *__restrict__ dirs, const float *__restrict__ v_colors,
                      float *__restrict__ v_coeffs) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float Y[25];
    sh_basis<4>(deg, dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], Y);
    const int nb = sh_num_bases(deg);
    const float vr = v_colors[3 * i], vg = v_colors[3 * i + 1], vb = v_colors[3 * i + 2];
    float *c = v_coeffs + (size_t)i * K * 3;
    for (int k = 0; k < K; ++k) {
        const float yk = k < nb ? Y[k] : 0.f;
        c[3 * k] = yk * vr; c[3 * k + 1] = yk * vg; c[3 * k + 2] = yk * vb;
    }
}

}  // namespace

extern "C" int gs_sh_fwd(int N, int K, int deg, const float *dirs, const float *coeffs, float *colors,
                         void *stream) {
    if (N < 0 || deg < 0 || deg > 4 || K < (deg + 1) * (deg + 1)) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!dirs || !coeffs || !colors) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (K == 16 && gs_aligned16(coeffs)) {
        sh_fwd16_kernel<<<(N + SH_TPB - 1) / SH_TPB, SH_TPB, 0, st>>>(
            N, deg, dirs, reinterpret_cast<const float4 *>(coeffs), colors);
    } else {
        sh_fwd_generic_kernel<<<(N + 127) / 128, 128, 0, st>>>(N, K, deg, dirs, coeffs, colors);
    }
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_sh_bwd(int N, int K, int deg, const float *dirs, const float *v_colors, float *v_coeffs,
                         void *stream) {
    if (N < 0 || deg < 0 || deg > 4 || K < (deg + 1) * (deg + 1)) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!dirs || !v_colors || !v_coeffs) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (K == 16 && gs_aligned16(v_coeffs)) {
        sh_bwd16_kernel<<<(N + SH_TPB - 1) / SH_TPB, SH_TPB, 0, st>>>(
            N, deg, dirs, v_colors, reinterpret_cast<float4 *>(v_coeffs));
    } else {
        sh_bwd_generic_kernel<<<(N + 127) / 128, 128, 0, st>>>(N, K, deg, dirs, v_colors, v_coeffs);
    }
    GS_LAUNCH_CHECK();
    return 0;
}

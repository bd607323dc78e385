// Stage 2 — tile binning: the work gsplat's rasterize_gaussians does
// before compositing (torch.cumsum + map_gaussian_to_intersects + torch.sort +
// get_tile_bin_edges), reached from /root/reference/gaussctrl/gc_model.py:208 and :225.
// Integer work, bit-exact against the oracle. HBM/L2-bound.
#include "radix.cuh"

namespace {
namespace cg = cooperative_groups;

// ------------------------------------------------------------------------------------------
// Inclusive int32 scan, single pass, decoupled look-back (warp-parallel window of 32 tiles).
// ------------------------------------------------------------------------------------------
constexpr int SCAN_TPB = 256;
constexpr int SCAN_IPT = 8;
constexpr int SCAN_TILE = SCAN_TPB * SCAN_IPT;
#define FLAG_AGG (1ull << 32)
#define FLAG_INC (2ull << 32)

struct ScanCtl {
    unsigned int ticket;
    unsigned int pad[3];
};

__global__ void __launch_bounds__(SCAN_TPB)
scan_kernel(int N, const int32_t *__restrict__ in, int32_t *__restrict__ out, int32_t *__restrict__ total,
            ScanCtl *ctl, unsigned long long *status) {
    __shared__ unsigned int s_tile;
    __shared__ int s_warp[SCAN_TPB / 32];
    __shared__ int s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&ctl->ticket, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int base = tile * SCAN_TILE + tid * SCAN_IPT;
    int v[SCAN_IPT];
    if (base + SCAN_IPT <= N) {
        const int4 a = *reinterpret_cast<const int4 *>(in + base);
        const int4 b = *reinterpret_cast<const int4 *>(in + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_IPT; ++j) v[j] = base + j < N ? in[base + j] : 0;
    }
#pragma unroll
    for (int j = 1; j < SCAN_IPT; ++j) v[j] += v[j - 1];
    int inc = v[SCAN_IPT - 1];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < SCAN_TPB / 32; ++w) {
        if (w < warp) warp_off += s_warp[w];
        block_total += s_warp[w];
    }
    const int thread_off = warp_off + inc - v[SCAN_IPT - 1];

    if (warp == 0) {
        if (lane == 0)
            atomicExch(&status[tile], (tile == 0 ? FLAG_INC : FLAG_AGG) | (unsigned int)block_total);
        int excl = 0;
        int look = tile - 1;
        while (look >= 0) {
            const int idx = look - lane;
            unsigned long long w;
            do {
                w = idx >= 0 ? *reinterpret_cast<volatile unsigned long long *>(&status[idx]) : FLAG_INC;
            } while (__any_sync(0xffffffffu, (w >> 32) == 0));
            const unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 32) == 2);
            const int first = inc_mask ? __ffs(inc_mask) - 1 : 31;
            int part = lane <= first ? (int)(unsigned int)w : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            excl += part;
            if (inc_mask) break;
            look -= 32;
        }
        if (lane == 0) {
            if (tile > 0) atomicExch(&status[tile], FLAG_INC | (unsigned int)(excl + block_total));
            s_prefix = excl;
        }
    }
    __syncthreads();
    const int off = s_prefix + thread_off;
    if (base + SCAN_IPT <= N) {
        *reinterpret_cast<int4 *>(out + base) = make_int4(v[0] + off, v[1] + off, v[2] + off, v[3] + off);
        *reinterpret_cast<int4 *>(out + base + 4) = make_int4(v[4] + off, v[5] + off, v[6] + off, v[7] + off);
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_IPT; ++j)
            if (base + j < N) out[base + j] = v[j] + off;
    }
    if (total && tile == (N - 1) / SCAN_TILE && tid == SCAN_TPB - 1) *total = s_prefix + block_total;
}

// ------------------------------------------------------------------------------------------
// Key emission: key = tile_id << 32 | bits(depth), value = Gaussian id, row-major over the
// bbox starting at cum[i-1]. Small splats are written by their own thread; splats covering
// more than 8 tiles are written by the whole warp so one big splat does not serialise a lane.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
emit_kernel(int N, const float2 *__restrict__ xys, const float *__restrict__ depths,
            const int32_t *__restrict__ radii, const int32_t *__restrict__ cum, int tiles_x, int tiles_y,
            int64_t *__restrict__ keys, int32_t *__restrict__ ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0, start = 0;
    unsigned dbits = 0;
    if (i < N && radii[i] > 0) {
        const float2 c = xys[i];
        gs_tile_bbox(c.x, c.y, (float)radii[i], tiles_x, tiles_y, x0, y0, x1, y1);
        start = i == 0 ? 0 : cum[i - 1];
        dbits = __float_as_uint(depths[i]);
    }
    const int bw = x1 - x0, area = bw * (y1 - y0);
    if (area > 0 && area <= 8) {
        int at = start;
        for (int ty = y0; ty < y1; ++ty)
            for (int tx = x0; tx < x1; ++tx) {
                keys[at] = ((int64_t)(ty * tiles_x + tx) << 32) | (int64_t)dbits;
                ids[at] = i;
                ++at;
            }
    }
    unsigned big = __ballot_sync(0xffffffffu, area > 8);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const int sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
        const int sbw = __shfl_sync(0xffffffffu, bw, src), sarea = __shfl_sync(0xffffffffu, area, src);
        const int sstart = __shfl_sync(0xffffffffu, start, src);
        const unsigned sd = __shfl_sync(0xffffffffu, dbits, src);
        const int sid = __shfl_sync(0xffffffffu, i, src);
        for (int k = lane; k < sarea; k += 32) {
            const int ty = sy0 + k / sbw, tx = sx0 + k % sbw;
            keys[sstart + k] = ((int64_t)(ty * tiles_x + tx) << 32) | (int64_t)sd;
            ids[sstart + k] = sid;
        }
    }
}

// ------------------------------------------------------------------------------------------
// 64-bit (tile, depth) key sort: cooperative persistent LSD radix over the live bits only.
// ------------------------------------------------------------------------------------------
struct SortPlan {
    int npass;
    int shift[8];
    int bits[8];
};

__global__ void __launch_bounds__(gsradix::TPB)
sort64_kernel(int64_t n, const uint64_t *k0, const uint32_t *v0, uint64_t *ka, uint32_t *va, uint64_t *kb,
              uint32_t *vb, SortPlan plan, uint32_t *mat) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    auto &s = *reinterpret_cast<gsradix::Smem<uint64_t, 8> *>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const uint64_t *kin = k0;
    const uint32_t *vin = v0;
    // the final pass must land in (ka, va) = the caller's sorted outputs
    bool to_a = (plan.npass & 1) != 0;
    for (int p = 0; p < plan.npass; ++p) {
        uint64_t *ko = to_a ? ka : kb;
        uint32_t *vo = to_a ? va : vb;
        gsradix::radix_pass<uint64_t, 8>(grid, s, n, kin, vin, ko, vo, plan.shift[p], plan.bits[p], mat, nullptr);
        kin = ko; vin = vo;
        to_a = !to_a;
    }
}

// tile_bins[t] = [first index with tile t, last+1); zero-filled beforehand.
__global__ void __launch_bounds__(256)
edges_kernel(int64_t I, const int64_t *__restrict__ ks, int32_t *__restrict__ bins) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= I) return;
    const int32_t t = (int32_t)(ks[k] >> 32);
    if (k == 0 || (int32_t)(ks[k - 1] >> 32) != t) bins[2 * t] = (int32_t)k;
    if (k == I - 1 || (int32_t)(ks[k + 1] >> 32) != t) bins[2 * t + 1] = (int32_t)(k + 1);
}

int g_coop_blocks_sort64 = 0;

int coop_grid(const void *kernel, size_t smem, int *cached) {
    if (*cached > 0) return *cached;
    int dev = 0, sms = 0, occ = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    if (smem > 48 * 1024)
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, gsradix::TPB, smem) != cudaSuccess) return 0;
    if (occ > 4) occ = 4;
    *cached = occ * sms;
    return *cached;
}

}  // namespace

// ------------------------------------------------------------------------------------------
extern "C" size_t gs_scan_scratch_bytes(int N) {
    const size_t tiles = (size_t)(N + SCAN_TILE - 1) / SCAN_TILE + 1;
    return sizeof(ScanCtl) + tiles * sizeof(unsigned long long);
}

extern "C" int gs_bin_scan(int N, const int32_t *tiles_hit, int32_t *cum, int32_t *total_dev,
                           void *scratch, size_t scratch_bytes, void *stream) {
    if (N < 0) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) {
        if (total_dev) GS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), st));
        return 0;
    }
    if (!tiles_hit || !cum || !scratch) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(tiles_hit) || !gs_aligned16(cum) || !gs_aligned16(scratch)) return GS_ERR_MISALIGNED;
    const size_t need = gs_scan_scratch_bytes(N);
    if (scratch_bytes < need) return GS_ERR_SCRATCH_SMALL;
    GS_CUDA(cudaMemsetAsync(scratch, 0, need, st));
    ScanCtl *ctl = reinterpret_cast<ScanCtl *>(scratch);
    unsigned long long *status = reinterpret_cast<unsigned long long *>(ctl + 1);
    const int blocks = (N + SCAN_TILE - 1) / SCAN_TILE;
    scan_kernel<<<blocks, SCAN_TPB, 0, st>>>(N, tiles_hit, cum, total_dev, ctl, status);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_bin_emit(int N, const float *xys, const float *depths, const int32_t *radii,
                           const int32_t *cum, int tiles_x, int tiles_y, int64_t *keys, int32_t *ids,
                           void *stream) {
    if (N < 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (N == 0) return 0;
    if (!xys || !depths || !radii || !cum || !keys || !ids) return GS_ERR_BAD_ARG;
    const int blocks = (N + 255) / 256;
    emit_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(N, reinterpret_cast<const float2 *>(xys), depths, radii,
                                                          cum, tiles_x, tiles_y, keys, ids);
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" size_t gs_sort_scratch_bytes(int64_t I) {
    // ping-pong buffers + CTA x digit matrix (sized for the largest grid we ever launch)
    const size_t keys = gs_align_up((size_t)I * 8, 256), vals = gs_align_up((size_t)I * 4, 256);
    return keys + vals + (size_t)(1024 + 1) * 256 * sizeof(uint32_t) + 256;
}

extern "C" int gs_bin_sort(int64_t I, int T, const int64_t *keys, const int32_t *ids,
                           int64_t *keys_sorted, int32_t *ids_sorted,
                           void *scratch, size_t scratch_bytes, void *stream) {
    if (I < 0 || T <= 0) return GS_ERR_BAD_ARG;
    if (I == 0) return 0;
    if (I >= (1ll << 31)) return GS_ERR_UNSUPPORTED;
    if (!keys || !ids || !keys_sorted || !ids_sorted || !scratch) return GS_ERR_BAD_ARG;
    if (scratch_bytes < gs_sort_scratch_bytes(I)) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(scratch)) return GS_ERR_MISALIGNED;
    const size_t smem = sizeof(gsradix::Smem<uint64_t, 8>);
    const int blocks = coop_grid((const void *)sort64_kernel, smem, &g_coop_blocks_sort64);
    if (blocks <= 0 || blocks > 1024) return GS_ERR_UNSUPPORTED;
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    uint64_t *kb = reinterpret_cast<uint64_t *>(p); p += gs_align_up((size_t)I * 8, 256);
    uint32_t *vb = reinterpret_cast<uint32_t *>(p); p += gs_align_up((size_t)I * 4, 256);
    uint32_t *mat = reinterpret_cast<uint32_t *>(p);
    // live bits: 32 depth bits + ceil(log2 T) tile bits, split evenly into <= 8-bit digits
    const int total_bits = 32 + (T > 1 ? gs_ceil_log2(T) : 1);
    SortPlan plan;
    plan.npass = (total_bits + 7) / 8;
    int at = 0;
    for (int q = 0; q < plan.npass; ++q) {
        const int left = total_bits - at, passes_left = plan.npass - q;
        const int b = (left + passes_left - 1) / passes_left;
        plan.shift[q] = at; plan.bits[q] = b; at += b;
    }
    int64_t n = I;
    const uint64_t *k0 = reinterpret_cast<const uint64_t *>(keys);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(ids);
    uint64_t *ka = reinterpret_cast<uint64_t *>(keys_sorted);
    uint32_t *va = reinterpret_cast<uint32_t *>(ids_sorted);
    void *args[] = {&n, &k0, &v0, &ka, &va, &kb, &vb, &plan, &mat};
    GS_CUDA(cudaLaunchCooperativeKernel((const void *)sort64_kernel, dim3(blocks), dim3(gsradix::TPB), args, smem,
                                        (cudaStream_t)stream));
    GS_LAUNCH_CHECK();
    return 0;
}

extern "C" int gs_bin_edges(int64_t I, int T, const int64_t *keys_sorted, int32_t *tile_bins, void *stream) {
    if (I < 0 || T <= 0 || !tile_bins) return GS_ERR_BAD_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    GS_CUDA(cudaMemsetAsync(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)T, st));
    if (I == 0) return 0;
    if (!keys_sorted) return GS_ERR_BAD_ARG;
    const int blocks = (int)((I + 255) / 256);
    edges_kernel<<<blocks, 256, 0, st>>>(I, keys_sorted, tile_bins);
    GS_LAUNCH_CHECK();
    return 0;
}

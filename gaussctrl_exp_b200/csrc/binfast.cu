// Stage 2, traffic-lean variant (SURVEY.md §7 "sort bandwidth"): same ids_sorted / tile_bins
// as scan + emit + 64-bit sort + edges, produced without ever materialising 64-bit keys:
//   1. stable radix sort of the N Gaussians by depth bits (32-bit keys, 4 passes over N, not I)
//   2. scan of tiles_hit in that order
//   3. emission of (tile id, Gaussian id) in that order
//   4. ONE stable partition by tile id (<= 2048 tiles: a single 11-bit pass over I)
//   5. tile ranges from the partitioned tile ids
// sort-by-depth then stable-partition-by-tile == sort by (tile, depth) with ties in id order,
// i.e. exactly the order of a stable sort of the 64-bit keys. All in one cooperative launch.
#include "radix.cuh"

namespace {
namespace cg = cooperative_groups;
using namespace gsradix;

struct FastArgs {
    int N;
    int64_t I_cap;
    const float2 *xys;
    const float *depths;
    const int32_t *radii;
    const int32_t *tiles_hit;
    int tiles_x, tiles_y, T, tile_bits;
    int32_t *ids_sorted;
    int32_t *tile_bins;
    int32_t *total_dev;
    uint32_t *keyA, *valA, *keyB, *valB;  // N each
    int32_t *start;                       // N: exclusive scan of tiles_hit in depth order
    uint32_t *tkey, *tval;                // I_cap each: emitted (tile, gid)
    uint32_t *tkey2, *tval2;              // I_cap each: partition ping-pong / sorted tile ids
    uint32_t *mat;
    uint32_t *cta_sum;                    // gridDim.x + 1
};

union FastSmem {
    Smem<uint32_t, 8> s8;
    Smem<uint32_t, 11> s11;
};

__global__ void __launch_bounds__(TPB)
binfast_kernel(FastArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    FastSmem &sm = *reinterpret_cast<FastSmem *>(smem_raw);
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x, lane = tid & 31;
    const int nb = gridDim.x, b = blockIdx.x;
    const int64_t gtid = (int64_t)b * TPB + tid, gsize = (int64_t)nb * TPB;
    const int N = a.N;

    // ---- 0. depth keys (culled Gaussians sort to the end), zero tile_bins ----
    for (int64_t i = gtid; i < N; i += gsize)
        a.keyA[i] = a.radii[i] > 0 ? __float_as_uint(a.depths[i]) : 0xFFFFFFFFu;
    for (int64_t i = gtid; i < 2 * (int64_t)a.T; i += gsize) a.tile_bins[i] = 0;
    grid.sync();

    // ---- 1. depth sort: A -> B -> A -> B -> A ----
    radix_pass<uint32_t, 8>(grid, sm.s8, N, a.keyA, nullptr, a.keyB, a.valB, 0, 8, a.mat, nullptr);
    radix_pass<uint32_t, 8>(grid, sm.s8, N, a.keyB, a.valB, a.keyA, a.valA, 8, 8, a.mat, nullptr);
    radix_pass<uint32_t, 8>(grid, sm.s8, N, a.keyA, a.valA, a.keyB, a.valB, 16, 8, a.mat, nullptr);
    radix_pass<uint32_t, 8>(grid, sm.s8, N, a.keyB, a.valB, a.keyA, a.valA, 24, 8, a.mat, nullptr);
    const uint32_t *order = a.valA;

    // ---- 2. scan of tiles_hit in depth order (chunk per CTA) ----
    const int64_t nsub = ((int64_t)N + SUB - 1) / SUB;
    const int64_t sub_per = (nsub + nb - 1) / nb;
    const int64_t c0 = min((int64_t)N, (int64_t)b * sub_per * SUB), c1 = min((int64_t)N, (int64_t)(b + 1) * sub_per * SUB);
    {
        uint32_t part = 0;
        for (int64_t j = c0 + tid; j < c1; j += TPB) part += (uint32_t)a.tiles_hit[ld_cg(order + j)];
        uint32_t tot;
        block_excl_scan(part, sm.s8.scan_tmp, &tot);
        if (tid == 0) a.cta_sum[b] = tot;
    }
    grid.sync();
    __shared__ uint32_t s_base, s_total;
    if (tid < 32) {
        uint32_t below = 0, all = 0;
        for (int cb = lane; cb < nb; cb += 32) {
            const uint32_t c = __ldcg(&a.cta_sum[cb]);
            all += c;
            if (cb < b) below += c;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            below += __shfl_xor_sync(0xffffffffu, below, o);
            all += __shfl_xor_sync(0xffffffffu, all, o);
        }
        if (lane == 0) { s_base = below; s_total = all; }
    }
    __syncthreads();
    const int64_t I = (int64_t)s_total;
    if (b == 0 && tid == 0) *a.total_dev = (int32_t)min(I, (int64_t)0x7fffffff);
    if (I > a.I_cap || I == 0) return;  // grid-uniform: every CTA sees the same total
    {
        uint32_t run = s_base;
        for (int64_t base = c0; base < c1; base += SUB) {
            // blocked arrangement: thread owns IPT consecutive sorted positions
            uint32_t h[IPT], sum = 0;
#pragma unroll
            for (int k = 0; k < IPT; ++k) {
                const int64_t j = base + (int64_t)tid * IPT + k;
                h[k] = j < c1 ? (uint32_t)a.tiles_hit[ld_cg(order + j)] : 0u;
                sum += h[k];
            }
            uint32_t tot;
            uint32_t ex = run + block_excl_scan(sum, sm.s8.scan_tmp, &tot);
#pragma unroll
            for (int k = 0; k < IPT; ++k) {
                const int64_t j = base + (int64_t)tid * IPT + k;
                if (j < c1) a.start[j] = (int32_t)ex;
                ex += h[k];
            }
            run += tot;
        }
    }
    grid.sync();

    // ---- 3. emission in depth order ----
    for (int64_t j0 = (int64_t)b * TPB; j0 < N; j0 += gsize) {
        const int64_t j = j0 + tid;
        int x0 = 0, y0 = 0, x1 = 0, y1 = 0, st = 0, g = 0;
        if (j < N) {
            g = (int)ld_cg(order + j);
            const int r = a.radii[g];
            if (r > 0) {
                const float2 c = a.xys[g];
                gs_tile_bbox(c.x, c.y, (float)r, a.tiles_x, a.tiles_y, x0, y0, x1, y1);
                st = __ldcg(&a.start[j]);
            }
        }
        const int bw = x1 - x0, area = bw * (y1 - y0);
        if (area > 0 && area <= 8) {
            int at = st;
            for (int ty = y0; ty < y1; ++ty)
                for (int tx = x0; tx < x1; ++tx) {
                    a.tkey[at] = (uint32_t)(ty * a.tiles_x + tx);
                    a.tval[at] = (uint32_t)g;
                    ++at;
                }
        }
        unsigned big = __ballot_sync(0xffffffffu, area > 8);
        while (big) {
            const int src = __ffs(big) - 1;
            big &= big - 1;
            const int sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
            const int sbw = __shfl_sync(0xffffffffu, bw, src), sarea = __shfl_sync(0xffffffffu, area, src);
            const int sst = __shfl_sync(0xffffffffu, st, src), sg = __shfl_sync(0xffffffffu, g, src);
            for (int k = lane; k < sarea; k += 32) {
                a.tkey[sst + k] = (uint32_t)((sy0 + k / sbw) * a.tiles_x + sx0 + k % sbw);
                a.tval[sst + k] = (uint32_t)sg;
            }
        }
    }
    grid.sync();

    // ---- 4. stable partition by tile id ----
    uint32_t *ids_out = reinterpret_cast<uint32_t *>(a.ids_sorted);
    const uint32_t *tk_sorted;
    if (a.tile_bits <= 11) {
        radix_pass<uint32_t, 11>(grid, sm.s11, I, a.tkey, a.tval, a.tkey2, ids_out, 0, a.tile_bits, a.mat, nullptr);
        tk_sorted = a.tkey2;
    } else {
        const int lo_bits = (a.tile_bits + 1) / 2, hi_bits = a.tile_bits - lo_bits;
        radix_pass<uint32_t, 11>(grid, sm.s11, I, a.tkey, a.tval, a.tkey2, a.tval2, 0, lo_bits, a.mat, nullptr);
        radix_pass<uint32_t, 11>(grid, sm.s11, I, a.tkey2, a.tval2, a.tkey, ids_out, lo_bits, hi_bits, a.mat, nullptr);
        tk_sorted = a.tkey;
    }

    // ---- 5. tile ranges ----
    for (int64_t k = gtid; k < I; k += gsize) {
        const uint32_t t = ld_cg(tk_sorted + k);
        if (k == 0 || ld_cg(tk_sorted + k - 1) != t) a.tile_bins[2 * t] = (int32_t)k;
        if (k == I - 1 || ld_cg(tk_sorted + k + 1) != t) a.tile_bins[2 * t + 1] = (int32_t)(k + 1);
    }
}

int g_coop_blocks_fast = 0;
int g_sms = 0;

struct FastLayout {
    size_t keyA, valA, keyB, valB, start, tkey, tval, tkey2, tval2, mat, cta_sum, total;
};

FastLayout fast_layout(int N, int64_t I_cap) {
    FastLayout L;
    size_t at = 0;
    auto take = [&](size_t bytes) { size_t o = at; at += gs_align_up(bytes, 256); return o; };
    const size_t n4 = (size_t)(N > 0 ? N : 1) * 4, i4 = (size_t)(I_cap > 0 ? I_cap : 1) * 4;
    L.keyA = take(n4); L.valA = take(n4); L.keyB = take(n4); L.valB = take(n4); L.start = take(n4);
    L.tkey = take(i4); L.tval = take(i4); L.tkey2 = take(i4); L.tval2 = take(i4);
    L.mat = take((size_t)(1024 + 1) * 2048 * 4);
    L.cta_sum = take(2048 * 4);
    L.total = at;
    return L;
}

}  // namespace

extern "C" size_t gs_bin_fast_scratch_bytes(int N, int64_t I_cap, int T) {
    (void)T;
    return fast_layout(N, I_cap).total;
}

extern "C" int gs_bin_fast(int N, int64_t I_cap, const float *xys, const float *depths, const int32_t *radii,
                           const int32_t *tiles_hit, int tiles_x, int tiles_y,
                           int32_t *ids_sorted, int32_t *tile_bins, int32_t *total_dev,
                           void *scratch, size_t scratch_bytes, void *stream) {
    if (N < 0 || I_cap < 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !total_dev || !scratch) return GS_ERR_BAD_ARG;
    if (I_cap >= (1ll << 31)) return GS_ERR_UNSUPPORTED;
    const int T = tiles_x * tiles_y;
    const int tile_bits = T > 1 ? gs_ceil_log2(T) : 1;
    if (tile_bits > 22) return GS_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) {
        GS_CUDA(cudaMemsetAsync(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)T, st));
        GS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), st));
        return 0;
    }
    if (!xys || !depths || !radii || !tiles_hit || (!ids_sorted && I_cap > 0)) return GS_ERR_BAD_ARG;
    const FastLayout L = fast_layout(N, I_cap);
    if (scratch_bytes < L.total) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(scratch)) return GS_ERR_MISALIGNED;
    const size_t smem = sizeof(FastSmem);
    if (g_coop_blocks_fast <= 0) {
        int dev = 0, occ = 0;
        GS_CUDA(cudaGetDevice(&dev));
        GS_CUDA(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
        GS_CUDA(cudaFuncSetAttribute((const void *)binfast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        GS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void *)binfast_kernel, TPB, smem));
        if (occ <= 0) return GS_ERR_UNSUPPORTED;
        if (occ > 3) occ = 3;
        g_coop_blocks_fast = occ * g_sms;
        if (g_coop_blocks_fast > 1024) g_coop_blocks_fast = 1024;
    }
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    FastArgs a;
    a.N = N; a.I_cap = I_cap;
    a.xys = reinterpret_cast<const float2 *>(xys); a.depths = depths; a.radii = radii; a.tiles_hit = tiles_hit;
    a.tiles_x = tiles_x; a.tiles_y = tiles_y; a.T = T; a.tile_bits = tile_bits;
    a.ids_sorted = ids_sorted; a.tile_bins = tile_bins; a.total_dev = total_dev;
    a.keyA = (uint32_t *)(p + L.keyA); a.valA = (uint32_t *)(p + L.valA);
    a.keyB = (uint32_t *)(p + L.keyB); a.valB = (uint32_t *)(p + L.valB);
    a.start = (int32_t *)(p + L.start);
    a.tkey = (uint32_t *)(p + L.tkey); a.tval = (uint32_t *)(p + L.tval);
    a.tkey2 = (uint32_t *)(p + L.tkey2); a.tval2 = (uint32_t *)(p + L.tval2);
    a.mat = (uint32_t *)(p + L.mat); a.cta_sum = (uint32_t *)(p + L.cta_sum);
    void *args[] = {&a};
    GS_CUDA(cudaLaunchCooperativeKernel((const void *)binfast_kernel, dim3(g_coop_blocks_fast), dim3(TPB), args,
                                        smem, st));
    GS_LAUNCH_CHECK();
    return 0;
}

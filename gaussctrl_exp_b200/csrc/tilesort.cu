// Stage 2, tile-local variant (gs_bin_fast): same ids_sorted / tile_bins as a stable sort of the 64-bit
// (tile, depth) keys, built from five launches instead of a multi-pass global radix sort:
//   1. count      : per Gaussian, one RED.ADD per covered tile                        -> counts[T]
//   2. tile scan  : exclusive scan of counts (one CTA) -> tile starts, tile_bins, total (stays on device)
//   3. scatter    : (depth bits, Gaussian id) into its tile's segment at an atomic cursor (unordered)
//   4. tile sort  : ONE CTA PER TILE loads its segment into shared memory and LSD-radix-sorts it by the
//                   32 depth bits (digits that are uniform inside the tile — usually the exponent byte —
//                   are skipped), orders equal depths by Gaussian id, and writes ids_sorted.
// Why: at this path's sizes (I ~ 3 M pairs, ~2.9 K per tile) the global sort was bound by the latency of
// ~24 dependent micro-kernels (ncu: 216 us at 11 % of the HBM roofline, profiles/r01c); per-tile segments fit
// the 227 KB of shared memory, so the sort never leaves the SM. Result order = (tile, depth, id) exactly
// as the oracle's stable sort (ties in depth are broken by id explicitly).
// Segments larger than the shared-memory capacity of the kernel variant fall back to the same CTA-local
// sort running on global ping-pong buffers (correct, slower; only reached in the multi-million-Gaussian
// stress configurations).
#include "gs_common.cuh"

namespace {


// Global per-tile counters are padded to one per 128 B line: L2 serialises atomics per sector, and 1024
// adjacent 4-byte counters share only 32 lines (ncu: 3 M RED.ADD on them took 326 us, profiles/r01d).
constexpr int TS_PAD = 32;   // uint32 stride between per-tile counters

// bbox of Gaussian i (empty when culled)
__device__ __forceinline__ void ts_bbox(int i, int N, const float2 *__restrict__ xys, const int32_t *__restrict__ radii,
                                        int tiles_x, int tiles_y, int &x0, int &y0, int &bw, int &area) {
    x0 = y0 = bw = area = 0;
    if (i < N) {
        const int r = radii[i];
        if (r > 0) {
            int x1, y1;
            const float2 c = xys[i];
            gs_tile_bbox(c.x, c.y, (float)r, tiles_x, tiles_y, x0, y0, x1, y1);
            bw = x1 - x0;
            area = bw * (y1 - y0);
        }
    }
}

// Visit every (Gaussian, tile) pair of the calling CTA's chunk [g_lo, g_hi). Small splats are handled by their own
// thread, splats covering more than 8 tiles by the whole warp. f(tile, gaussian id) is called once per pair.
template <typename F>
__device__ __forceinline__ void ts_for_each_pair(int g_lo, int g_hi, int N, const float2 *__restrict__ xys,
                                                 const int32_t *__restrict__ radii, int tiles_x, int tiles_y, F f) {
    const int lane = threadIdx.x & 31;
    for (int base = g_lo; base < g_hi; base += blockDim.x) {      // block-uniform trip count
        const int i = base + threadIdx.x;
        int x0, y0, bw, area;
        ts_bbox(i < g_hi ? i : N, N, xys, radii, tiles_x, tiles_y, x0, y0, bw, area);
        if (area > 0 && area <= 8)
            for (int k = 0; k < area; ++k) f((y0 + k / bw) * tiles_x + x0 + k % bw, i);
        unsigned big = __ballot_sync(0xffffffffu, area > 8);
        while (big) {
            const int src = __ffs(big) - 1;
            big &= big - 1;
            const int sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
            const int sbw = __shfl_sync(0xffffffffu, bw, src), sarea = __shfl_sync(0xffffffffu, area, src);
            const int sid = __shfl_sync(0xffffffffu, i, src);
            for (int k = lane; k < sarea; k += 32) f((sy0 + k / sbw) * tiles_x + sx0 + k % sbw, sid);
        }
    }
}

__device__ __forceinline__ void ts_chunk(int N, int &g_lo, int &g_hi) {
    const int per = ((N + gridDim.x - 1) / gridDim.x + 31) & ~31;
    g_lo = min(N, (int)blockIdx.x * per);
    g_hi = min(N, g_lo + per);
}

// ---- 1. count: persistent CTAs, shared-memory histogram of the chunk, one global RED per (CTA, touched tile) ----
__global__ void __launch_bounds__(256)
ts_count_kernel(int N, int T, const float2 *__restrict__ xys, const int32_t *__restrict__ radii, int tiles_x,
                int tiles_y, uint32_t *__restrict__ counts) {
    extern __shared__ uint32_t s_hist[];
    for (int t = threadIdx.x; t < T; t += blockDim.x) s_hist[t] = 0;
    __syncthreads();
    int g_lo, g_hi;
    ts_chunk(N, g_lo, g_hi);
    ts_for_each_pair(g_lo, g_hi, N, xys, radii, tiles_x, tiles_y, [&](int t, int) { atomicAdd(&s_hist[t], 1u); });
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const uint32_t c = s_hist[t];
        if (c) atomicAdd(&counts[(size_t)t * TS_PAD], c);
    }
}

// ---- 2. scan over tiles (single CTA) -----------------------------------------------------------
__global__ void __launch_bounds__(1024)
ts_scan_kernel(int T, int64_t I_cap, const uint32_t *__restrict__ counts, uint32_t *__restrict__ starts,
               int32_t *__restrict__ tile_bins, int32_t *__restrict__ total_dev) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < T; base += 1024) {
        const int t = base + tid;
        const uint32_t c = t < T ? counts[(size_t)t * TS_PAD] : 0u;
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) warp_sum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_sum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += n;
            }
            warp_sum[lane] = w;   // inclusive
        }
        __syncthreads();
        const uint32_t start = s_carry + (warp ? warp_sum[warp - 1] : 0u) + inc - c;
        if (t < T) {
            starts[t] = start;
            tile_bins[2 * t] = c ? (int32_t)start : 0;
            tile_bins[2 * t + 1] = c ? (int32_t)(start + c) : 0;
        }
        __syncthreads();
        if (tid == 1023) s_carry = start + c;
        __syncthreads();
    }
    if (tid == 0) *total_dev = (int32_t)min(s_carry, 0x7fffffffu);
    // capacity too small: ids_sorted will not be written, so publish EMPTY tile ranges (the compositor
    // then renders the background and touches no list) and let the host retry with a larger buffer
    if ((int64_t)s_carry > I_cap)
        for (int t = tid; t < T; t += 1024) { tile_bins[2 * t] = 0; tile_bins[2 * t + 1] = 0; }
}

// ---- 3. unordered scatter: the CTA re-counts its chunk, reserves one contiguous range per touched tile with a
//         single global atomic, then places its pairs with shared-memory cursors ------------------------------
__global__ void __launch_bounds__(256)
ts_scatter_kernel(int N, int T, int64_t I_cap, const int32_t *__restrict__ total_dev, const float2 *__restrict__ xys,
                  const float *__restrict__ depths, const int32_t *__restrict__ radii, int tiles_x, int tiles_y,
                  const uint32_t *__restrict__ starts, uint32_t *__restrict__ cursors, uint2 *__restrict__ pairs) {
    extern __shared__ uint32_t s_hist[];
    if ((int64_t)*total_dev > I_cap) return;   // capacity too small: write nothing, the host retries
    for (int t = threadIdx.x; t < T; t += blockDim.x) s_hist[t] = 0;
    __syncthreads();
    int g_lo, g_hi;
    ts_chunk(N, g_lo, g_hi);
    ts_for_each_pair(g_lo, g_hi, N, xys, radii, tiles_x, tiles_y, [&](int t, int) { atomicAdd(&s_hist[t], 1u); });
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const uint32_t c = s_hist[t];
        s_hist[t] = c ? starts[t] + atomicAdd(&cursors[(size_t)t * TS_PAD], c) : 0u;   // -> this CTA's write cursor
    }
    __syncthreads();
    ts_for_each_pair(g_lo, g_hi, N, xys, radii, tiles_x, tiles_y, [&](int t, int g) {
        pairs[atomicAdd(&s_hist[t], 1u)] = make_uint2(__float_as_uint(depths[g]), (unsigned)g);
    });
}

// ---- 4. per-tile sort ----------------------------------------------------------------------------
template <int TPB>
struct TsSmem {
    static constexpr int WARPS = TPB / 32;
    uint32_t hist[256];
    uint32_t cursor[256];
    uint32_t dbase[256];
    uint16_t wc[WARPS][256];
    uint32_t tmp[64];
    uint32_t diff;
};

template <int TPB>
__device__ __forceinline__ uint32_t ts_scan256(uint32_t v, uint32_t *tmp) {
    // exclusive scan over threads 0..255 (the digit owners); all TPB threads call it
    constexpr int WARPS = TPB / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) tmp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < WARPS ? tmp[lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += n;
        }
        tmp[32 + lane] = w;
    }
    __syncthreads();
    const uint32_t r = (warp ? tmp[32 + warp - 1] : 0u) + inc - v;
    __syncthreads();
    return r;
}

// CTA-local stable LSD radix sort of n (depth, id) pairs by depth; A/B may live in shared or global
// memory (generic pointers). Only the bit range in which the keys of THIS tile differ is sorted (inside a
// tile the exponent byte is almost always constant). Returns the buffer that holds the result.
template <int TPB>
__device__ uint2 *ts_sort_segment(uint2 *A, uint2 *B, int n, TsSmem<TPB> &s) {
    constexpr int WARPS = TPB / 32;
    constexpr int IPT = 8;
    constexpr int CHUNK = TPB * IPT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- which bits vary? ----
    if (tid == 0) s.diff = 0;
    __syncthreads();
    {
        const uint32_t k0 = A[0].x;
        uint32_t d = 0;
        for (int i = tid; i < n; i += TPB) d |= A[i].x ^ k0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) d |= __shfl_xor_sync(0xffffffffu, d, o);
        if (lane == 0 && d) atomicOr(&s.diff, d);
    }
    __syncthreads();
    const uint32_t diff = s.diff;
    if (diff != 0) {
        const int lo = __ffs(diff) - 1, hi = 31 - __clz(diff);
        const int nbits = hi - lo + 1, npass = (nbits + 7) / 8, width = (nbits + npass - 1) / npass;
        for (int pass = 0; pass < npass; ++pass) {
            const int shift = lo + pass * width;
            const uint32_t mask = (1u << min(width, hi + 1 - shift)) - 1u;
            if (tid < 256) s.hist[tid] = 0;
            __syncthreads();
            // warp-aggregated histogram: one shared-memory atomic per distinct digit per warp round
            for (int i0 = 0; i0 < n; i0 += TPB) {
                const int i = i0 + tid;
                const uint32_t dg = i < n ? ((A[i].x >> shift) & mask) : (256u + (unsigned)lane);
                const unsigned peers = __match_any_sync(0xffffffffu, dg);
                if (dg < 256u && lane == __ffs(peers) - 1) atomicAdd(&s.hist[dg], (uint32_t)__popc(peers));
            }
            __syncthreads();
            const uint32_t c = tid < 256 ? s.hist[tid] : 0u;
            const uint32_t ex = ts_scan256<TPB>(c, s.tmp);
            if (tid < 256) s.cursor[tid] = ex;
            __syncthreads();
            for (int base = 0; base < n; base += CHUNK) {
                for (int k = tid; k < WARPS * 256; k += TPB) (&s.wc[0][0])[k] = 0;
                __syncthreads();
                uint2 key[IPT];
                uint32_t dig[IPT];
                unsigned peers[IPT];
                uint16_t rank[IPT];
#pragma unroll
                for (int i = 0; i < IPT; ++i) {
                    const int idx = base + warp * (32 * IPT) + i * 32 + lane;
                    const bool valid = idx < n;
                    key[i] = valid ? A[idx] : make_uint2(0u, 0u);
                    dig[i] = valid ? ((key[i].x >> shift) & mask) : (256u + (unsigned)lane);   // invalid: unique, uncounted
                    peers[i] = __match_any_sync(0xffffffffu, dig[i]);
                }
#pragma unroll
                for (int i = 0; i < IPT; ++i) {
                    const int leader = __ffs(peers[i]) - 1;
                    uint32_t pre = 0;
                    if (lane == leader && dig[i] < 256u) {
                        pre = s.wc[warp][dig[i]];
                        s.wc[warp][dig[i]] = (uint16_t)(pre + __popc(peers[i]));
                    }
                    pre = __shfl_sync(0xffffffffu, pre, leader);
                    rank[i] = (uint16_t)(pre + __popc(peers[i] & gs_lanemask_lt()));
                    __syncwarp();
                }
                __syncthreads();
                if (tid < 256) {   // thread d owns digit d: exclusive over warps, advance the running cursor
                    uint32_t r = 0;
#pragma unroll
                    for (int w = 0; w < WARPS; ++w) {
                        const uint32_t cc = s.wc[w][tid];
                        s.wc[w][tid] = (uint16_t)r;
                        r += cc;
                    }
                    s.dbase[tid] = s.cursor[tid];
                    s.cursor[tid] += r;
                }
                __syncthreads();
#pragma unroll
                for (int i = 0; i < IPT; ++i)
                    if (dig[i] < 256u) B[s.dbase[dig[i]] + s.wc[warp][dig[i]] + rank[i]] = key[i];
                __syncthreads();
            }
            uint2 *t = A; A = B; B = t;
        }
    }
    // equal depths: order by Gaussian id (what a stable sort of the id-ordered emission produces)
    for (int i = tid; i < n - 1; i += TPB) {
        const uint32_t d = A[i].x;
        if (A[i + 1].x == d && (i == 0 || A[i - 1].x != d)) {
            int e = i + 1;
            while (e + 1 < n && A[e + 1].x == d) ++e;
            for (int a = i + 1; a <= e; ++a) {          // insertion sort of the (normally tiny) run [i, e]
                const uint2 v = A[a];
                int b = a - 1;
                while (b >= i && A[b].y > v.y) { A[b + 1] = A[b]; --b; }
                A[b + 1] = v;
            }
        }
    }
    __syncthreads();
    return A;
}

// Handles the tiles whose segment length n satisfies lo < n <= hi. n <= cap: in shared memory.
template <int TPB>
__global__ void __launch_bounds__(TPB)
ts_tile_sort_kernel(int T, int64_t I_cap, const int32_t *__restrict__ total_dev, const int2 *__restrict__ tile_bins,
                    uint2 *pairs, uint2 *pairs_tmp, int32_t *__restrict__ ids_sorted, int cap, int lo, int hi) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TsSmem<TPB> &s = *reinterpret_cast<TsSmem<TPB> *>(smem_raw);
    uint2 *bufA = reinterpret_cast<uint2 *>(smem_raw + ((sizeof(TsSmem<TPB>) + 15) & ~size_t(15)));
    uint2 *bufB = bufA + cap;
    if ((int64_t)*total_dev > I_cap) return;
    const int tile = blockIdx.x;
    const int2 range = tile_bins[tile];
    const int n = range.y - range.x;
    if (n <= lo || n > hi) return;
    const int tid = threadIdx.x;
    uint2 *seg = pairs + range.x;
    uint2 *res;
    if (n <= cap) {
        for (int i = tid; i < n; i += TPB) bufA[i] = seg[i];
        __syncthreads();
        res = ts_sort_segment<TPB>(bufA, bufB, n, s);
    } else {
        res = ts_sort_segment<TPB>(seg, pairs_tmp + range.x, n, s);
    }
    for (int i = tid; i < n; i += TPB) ids_sorted[range.x + i] = (int32_t)res[i].y;
}

constexpr int TS_CAP_SMALL = 6144;    // 2 x 48 KB ping-pong + 7 KB: two 256-thread CTAs per SM
constexpr int TS_CAP_LARGE = 13056;   // 2 x 102 KB + 19 KB: one 1024-thread CTA per SM
constexpr int TS_TPB_SMALL = 256, TS_TPB_LARGE = 1024;

template <int TPB>
size_t ts_smem_bytes(int cap) { return ((sizeof(TsSmem<TPB>) + 15) & ~size_t(15)) + (size_t)cap * 16; }

struct TsLayout {
    size_t counts, starts, cursors, pairs, pairs_tmp, total;
};
TsLayout ts_layout(int T, int64_t I_cap) {
    TsLayout L;
    size_t at = 0;
    auto take = [&](size_t bytes) { size_t o = at; at += gs_align_up(bytes, 256); return o; };
    const size_t t4 = (size_t)(T > 0 ? T : 1) * 4, i8 = (size_t)(I_cap > 0 ? I_cap : 1) * 8;
    L.counts = take(t4 * TS_PAD); L.cursors = take(t4 * TS_PAD); L.starts = take(t4);
    L.pairs = take(i8); L.pairs_tmp = take(i8);
    L.total = at;
    return L;
}
bool g_ts_attr_set = false;
int g_ts_sms = 0;

}  // namespace

extern "C" size_t gs_bin_tile_scratch_bytes(int N, int64_t I_cap, int T) {
    (void)N;
    return ts_layout(T, I_cap).total;
}

extern "C" int gs_bin_tile(int N, int64_t I_cap, const float *xys, const float *depths, const int32_t *radii,
                           int tiles_x, int tiles_y, int32_t *ids_sorted, int32_t *tile_bins, int32_t *total_dev,
                           void *scratch, size_t scratch_bytes, void *stream) {
    if (N < 0 || I_cap < 0 || tiles_x <= 0 || tiles_y <= 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !total_dev || !scratch) return GS_ERR_BAD_ARG;
    if (I_cap >= (1ll << 31)) return GS_ERR_UNSUPPORTED;
    const int T = tiles_x * tiles_y;
    if ((size_t)T * 4 > 200 * 1024) return GS_ERR_UNSUPPORTED;   // per-CTA tile histogram must fit shared memory
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) {
        GS_CUDA(cudaMemsetAsync(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)T, st));
        GS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), st));
        return 0;
    }
    if (!xys || !depths || !radii || (!ids_sorted && I_cap > 0)) return GS_ERR_BAD_ARG;
    const TsLayout L = ts_layout(T, I_cap);
    if (scratch_bytes < L.total) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(scratch)) return GS_ERR_MISALIGNED;
    if (!g_ts_attr_set) {
        int dev = 0;
        GS_CUDA(cudaGetDevice(&dev));
        GS_CUDA(cudaDeviceGetAttribute(&g_ts_sms, cudaDevAttrMultiProcessorCount, dev));
        GS_CUDA(cudaFuncSetAttribute((const void *)ts_tile_sort_kernel<TS_TPB_SMALL>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)ts_smem_bytes<TS_TPB_SMALL>(TS_CAP_SMALL)));
        GS_CUDA(cudaFuncSetAttribute((const void *)ts_tile_sort_kernel<TS_TPB_LARGE>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)ts_smem_bytes<TS_TPB_LARGE>(TS_CAP_LARGE)));
        GS_CUDA(cudaFuncSetAttribute((const void *)ts_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        GS_CUDA(cudaFuncSetAttribute((const void *)ts_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        g_ts_attr_set = true;
    }
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    uint32_t *counts = (uint32_t *)(p + L.counts), *cursors = (uint32_t *)(p + L.cursors);
    uint32_t *starts = (uint32_t *)(p + L.starts);
    uint2 *pairs = (uint2 *)(p + L.pairs), *pairs_tmp = (uint2 *)(p + L.pairs_tmp);
    const float2 *xy2 = reinterpret_cast<const float2 *>(xys);
    const size_t hist_smem = (size_t)T * 4;
    // persistent CTAs: enough of them to fill the machine, few enough that each (CTA, tile) pair costs
    // only one global atomic
    int grid = g_ts_sms * (hist_smem <= 48 * 1024 ? 4 : 1);
    grid = grid < (N + 255) / 256 ? grid : (N + 255) / 256;
    GS_CUDA(cudaMemsetAsync(counts, 0, L.starts - L.counts, st));   // counts + cursors are adjacent
    ts_count_kernel<<<grid, 256, hist_smem, st>>>(N, T, xy2, radii, tiles_x, tiles_y, counts);
    GS_LAUNCH_CHECK();
    ts_scan_kernel<<<1, 1024, 0, st>>>(T, I_cap, counts, starts, tile_bins, total_dev);
    GS_LAUNCH_CHECK();
    if (I_cap == 0) return 0;
    ts_scatter_kernel<<<grid, 256, hist_smem, st>>>(N, T, I_cap, total_dev, xy2, depths, radii, tiles_x, tiles_y, starts,
                                                   cursors, pairs);
    GS_LAUNCH_CHECK();
    const int2 *bins2 = reinterpret_cast<const int2 *>(tile_bins);
    // the few long segments first (one 1024-thread CTA per SM), then everything else (two CTAs per SM)
    ts_tile_sort_kernel<TS_TPB_LARGE><<<T, TS_TPB_LARGE, ts_smem_bytes<TS_TPB_LARGE>(TS_CAP_LARGE), st>>>(
        T, I_cap, total_dev, bins2, pairs, pairs_tmp, ids_sorted, TS_CAP_LARGE, TS_CAP_SMALL, 0x7fffffff);
    GS_LAUNCH_CHECK();
    ts_tile_sort_kernel<TS_TPB_SMALL><<<T, TS_TPB_SMALL, ts_smem_bytes<TS_TPB_SMALL>(TS_CAP_SMALL), st>>>(
        T, I_cap, total_dev, bins2, pairs, pairs_tmp, ids_sorted, TS_CAP_SMALL, 0, TS_CAP_SMALL);
    GS_LAUNCH_CHECK();
    return 0;
}

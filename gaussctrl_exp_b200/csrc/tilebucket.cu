// Stage 2 — tile binning, bucket form (the SECOND, opt-in implementation behind gs_bin_fast / gs_view_fwd, selected with
// gs_set_bin_impl(0) / GS_BIN_IMPL=bucket — the default is the LSD pipeline of binning.cu, which measured faster; SURVEY.md App. A.3:
// torch.cumsum + map_gaussian_to_intersects + torch.sort + get_tile_bin_edges inside gsplat's rasterize_gaussians,
// reached from /root/reference/gaussctrl/gc_model.py:208 and :225). Same ids_sorted / tile_bins as a stable sort of
// the 64-bit (tile << 32 | depth bits) keys, bit for bit, from FOUR launches:
//   1. count : chunk of 4 096 Gaussians per CTA, shared-memory histogram of the tiles its splats touch, one global
//              RED per (CTA, touched tile) on 128-byte-padded counters                        -> counts[tile]
//   2. scan  : exclusive scan of the tile counts (decoupled look-back)                        -> starts[tile], total
//   3. fill  : the same chunks again: histogram, ONE atomic per (CTA, touched tile) reserves the chunk's slots in
//              the tile's segment, then every (tile, Gaussian) pair is written as (depth bits << 32 | id) at
//              segment start + reserved base + shared-memory rank. Order inside a segment is arbitrary.
//   4. sort  : one CTA per tile: segment -> shared memory, merge sort of the 64-bit (depth, id) keys (thread-local
//              sorting network, then log2 merge rounds by merge-path partition), ids out; tile_bins = segment.
// The (depth, id) key is a total order, so the arbitrary fill order cannot show in the output: equal depths come out
// in Gaussian-id order exactly as the stable key sort of the reference leaves them.
//
// Why (B200): the LSD pipeline (binning.cu: depth sort of N, scan, emission, partition over I) is 16 dependent
// latency-bound kernels over an L2-resident working set, 165 us at 1 M Gaussians / 3 M intersections
// (profiles/README.md); its traffic is nowhere near the HBM roofline, its length is. Here every pair is touched twice
// in global memory (8-byte write, 8-byte read) plus a 4-byte id write, the sort never leaves the SM, and global
// atomics are per (chunk, tile), not per pair (round 1 measured 3 M per-pair atomics on 1 024 hot counters at
// 326 us; per-chunk reservation makes them ~250 K on padded counters).
// Segments longer than the shared-memory capacity (8 192 entries) are sorted in runs and merged through global
// ping-pong buffers by the same CTA (correct for any length; slower — only met in pathological tiles).
#include <stdlib.h>
#define GS_PDL_GROUP 1
#include "gs_common.cuh"

namespace {

constexpr int TB_PAD = 32;            // uint32 stride of the per-tile global counters (one per 128-byte line)
constexpr int TB_TPB = 256;
constexpr int TB_EPT = 16;            // Gaussians per thread
constexpr int TB_CHUNK = TB_TPB * TB_EPT;

// ---------------------------------------------------------------------------------------------------------------
// pairs of a chunk. Each thread keeps the tile boxes of its TB_EPT Gaussians in registers; f(tile in view, element,
// depth bits) is called once per (tile, Gaussian) pair — by the owning thread for boxes of <= 8 tiles, by the whole
// warp for larger ones.
// ---------------------------------------------------------------------------------------------------------------
struct ChunkBoxes {
    uint32_t xy0[TB_EPT];     // x0 | y0 << 16
    uint32_t wh[TB_EPT];      // width | height << 16 (0 = culled)
};

__device__ __forceinline__ void load_boxes(ChunkBoxes &cb, int g0, int g_end, const float *__restrict__ xy_base, int xy_stride,
                                           const int32_t *__restrict__ radii, int tiles_x, int tiles_y) {
    int r[TB_EPT];
#pragma unroll
    for (int k = 0; k < TB_EPT; ++k) {
        const int g = g0 + k * TB_TPB + threadIdx.x;
        r[k] = g < g_end ? radii[g] : 0;
    }
#pragma unroll
    for (int k = 0; k < TB_EPT; ++k) {
        const int g = g0 + k * TB_TPB + threadIdx.x;
        cb.xy0[k] = cb.wh[k] = 0;
        if (r[k] > 0) {
            const float2 c = *reinterpret_cast<const float2 *>(xy_base + (size_t)g * xy_stride);
            int x0, y0, x1, y1;
            gs_tile_bbox(c.x, c.y, (float)r[k], tiles_x, tiles_y, x0, y0, x1, y1);
            if (x1 > x0 && y1 > y0) {
                cb.xy0[k] = (uint32_t)x0 | ((uint32_t)y0 << 16);
                cb.wh[k] = (uint32_t)(x1 - x0) | ((uint32_t)(y1 - y0) << 16);
            }
        }
    }
}

template <typename F>
__device__ __forceinline__ void for_each_pair(const ChunkBoxes &cb, int g0, int tiles_x, F f) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < TB_EPT; ++k) {
        const int g = g0 + k * TB_TPB + threadIdx.x;
        const int x0 = cb.xy0[k] & 0xffff, y0 = cb.xy0[k] >> 16, bw = cb.wh[k] & 0xffff, bh = cb.wh[k] >> 16;
        const int area = bw * bh;
        if (area > 0 && area <= 8) {
            for (int ty = y0; ty < y0 + bh; ++ty)
                for (int tx = x0; tx < x0 + bw; ++tx) f(ty * tiles_x + tx, g, k);
        }
        unsigned big = __ballot_sync(0xffffffffu, area > 8);
        while (big) {
            const int src = __ffs(big) - 1;
            big &= big - 1;
            const int sx0 = __shfl_sync(0xffffffffu, x0, src), sy0 = __shfl_sync(0xffffffffu, y0, src);
            const int sbw = __shfl_sync(0xffffffffu, bw, src), sarea = __shfl_sync(0xffffffffu, area, src);
            const int sg = __shfl_sync(0xffffffffu, g, src);
            for (int q = lane; q < sarea; q += 32) f((sy0 + q / sbw) * tiles_x + sx0 + q % sbw, sg, -1 - src);
        }
    }
}

// CTA -> (view, chunk of that view): a chunk never straddles two views, so its histogram covers one view's tiles
__device__ __forceinline__ void chunk_of(int n_per_view, int chunks_per_view, int &view, int &g0, int &g_end) {
    view = blockIdx.x / chunks_per_view;
    const int c = blockIdx.x - view * chunks_per_view;
    g0 = view * n_per_view + c * TB_CHUNK;
    g_end = min(g0 + TB_CHUNK, (view + 1) * n_per_view);
}

// ---- 1. count ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TB_TPB)
tb_count_kernel(int n_per_view, int chunks_per_view, int T_view, const float *__restrict__ xy_base, int xy_stride,
                const int32_t *__restrict__ radii, int tiles_x, int tiles_y, uint32_t *__restrict__ counts) {
    extern __shared__ uint32_t s_hist[];
    gs_pdl_sync();
    for (int t = threadIdx.x; t < T_view; t += TB_TPB) s_hist[t] = 0;
    int view, g0, g_end;
    chunk_of(n_per_view, chunks_per_view, view, g0, g_end);
    ChunkBoxes cb;
    load_boxes(cb, g0, g_end, xy_base, xy_stride, radii, tiles_x, tiles_y);
    __syncthreads();
    for_each_pair(cb, g0, tiles_x, [&](int t, int, int) { atomicAdd(&s_hist[t], 1u); });
    __syncthreads();
    uint32_t *cv = counts + (size_t)view * T_view * TB_PAD;
    for (int t = threadIdx.x; t < T_view; t += TB_TPB) {
        const uint32_t c = s_hist[t];
        if (c) atomicAdd(&cv[(size_t)t * TB_PAD], c);
    }
}

// ---- 2. scan of the tile counts: single pass, decoupled look-back (warp-wide window) ------------------------------
constexpr int TS_TPB = 256;
constexpr int TS_IPT = 4;
constexpr int TS_TILE = TS_TPB * TS_IPT;
#define TB_FLAG_AGG (1ull << 32)
#define TB_FLAG_INC (2ull << 32)

__global__ void __launch_bounds__(TS_TPB)
tb_scan_kernel(int T, const uint32_t *__restrict__ counts, uint32_t *__restrict__ starts, int32_t *__restrict__ total,
               unsigned int *ticket, unsigned long long *status) {
    __shared__ unsigned int s_tile;
    __shared__ uint32_t s_warp[TS_TPB / 32];
    __shared__ uint32_t s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    gs_pdl_sync();
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int base = tile * TS_TILE + tid * TS_IPT;
    uint32_t v[TS_IPT], sum = 0;
#pragma unroll
    for (int j = 0; j < TS_IPT; ++j) v[j] = base + j < T ? counts[(size_t)(base + j) * TB_PAD] : 0u;
#pragma unroll
    for (int j = 0; j < TS_IPT; ++j) sum += v[j];
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < TS_TPB / 32; ++w) {
        if (w < warp) warp_off += s_warp[w];
        block_total += s_warp[w];
    }
    if (warp == 0) {
        if (lane == 0) atomicExch(&status[tile], (tile == 0 ? TB_FLAG_INC : TB_FLAG_AGG) | block_total);
        uint32_t excl = 0;
        int look = tile - 1;
        while (look >= 0) {
            const int idx = look - lane;
            unsigned long long w;
            do {
                w = idx >= 0 ? *reinterpret_cast<volatile unsigned long long *>(&status[idx]) : TB_FLAG_INC;
            } while (__any_sync(0xffffffffu, (w >> 32) == 0));
            const unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 32) == 2);
            const int first = inc_mask ? __ffs(inc_mask) - 1 : 31;
            uint32_t part = lane <= first ? (uint32_t)w : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            excl += part;
            if (inc_mask) break;
            look -= 32;
        }
        if (lane == 0) {
            if (tile > 0) atomicExch(&status[tile], TB_FLAG_INC | (unsigned long long)(excl + block_total));
            s_prefix = excl;
        }
    }
    __syncthreads();
    uint32_t run = s_prefix + warp_off + inc - sum;
#pragma unroll
    for (int j = 0; j < TS_IPT; ++j) {
        if (base + j < T) starts[base + j] = run;
        run += v[j];
    }
    if (tile == (T - 1) / TS_TILE && tid == TS_TPB - 1) {
        starts[T] = s_prefix + block_total;
        *total = (int32_t)(s_prefix + block_total);
    }
}

// ---- 3. fill ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TB_TPB)
tb_fill_kernel(int n_per_view, int chunks_per_view, int T_view, int64_t I_cap, const int32_t *__restrict__ total_dev,
               const float *__restrict__ xy_base, int xy_stride, const float *__restrict__ depths,
               const int32_t *__restrict__ radii, int tiles_x, int tiles_y, const uint32_t *__restrict__ starts,
               uint32_t *__restrict__ cursors, unsigned long long *__restrict__ entries) {
    extern __shared__ uint32_t s_mem[];
    uint32_t *s_rank = s_mem, *s_base = s_mem + T_view;
    gs_pdl_sync();
    if ((int64_t)*total_dev > I_cap) return;      // capacity too small: write nothing, the host redoes the view
    for (int t = threadIdx.x; t < T_view; t += TB_TPB) s_rank[t] = 0;
    int view, g0, g_end;
    chunk_of(n_per_view, chunks_per_view, view, g0, g_end);
    ChunkBoxes cb;
    load_boxes(cb, g0, g_end, xy_base, xy_stride, radii, tiles_x, tiles_y);
    uint32_t dbits[TB_EPT];
#pragma unroll
    for (int k = 0; k < TB_EPT; ++k) {
        const int g = g0 + k * TB_TPB + threadIdx.x;
        dbits[k] = (g < g_end && cb.wh[k]) ? __float_as_uint(depths[g]) : 0u;
    }
    __syncthreads();
    for_each_pair(cb, g0, tiles_x, [&](int t, int, int) { atomicAdd(&s_rank[t], 1u); });
    __syncthreads();
    const size_t tb = (size_t)view * T_view;
    for (int t = threadIdx.x; t < T_view; t += TB_TPB) {
        const uint32_t c = s_rank[t];
        if (c) s_base[t] = starts[tb + t] + atomicAdd(&cursors[(tb + t) * TB_PAD], c);
        s_rank[t] = 0;
    }
    __syncthreads();
    for_each_pair(cb, g0, tiles_x, [&](int t, int g, int k) {
        // k >= 0: own element k; k < 0: warp-cooperative box of lane (-1 - k), whose depth sits in that lane
        uint32_t d = 0;
        if (k >= 0) {
#pragma unroll
            for (int q = 0; q < TB_EPT; ++q) d = q == k ? dbits[q] : d;
        } else {
            d = __float_as_uint(depths[g]);
        }
        const uint32_t at = s_base[t] + atomicAdd(&s_rank[t], 1u);
        entries[at] = ((unsigned long long)d << 32) | (unsigned long long)(uint32_t)g;
    });
}

// ---- 4. per-tile sort -----------------------------------------------------------------------------------------------
constexpr int SORT_TPB = 512;
constexpr int SORT_MAX_ITEMS = 16;
constexpr int SORT_CAP = SORT_TPB * SORT_MAX_ITEMS;           // 8 192 entries per shared-memory run
typedef unsigned long long u64;
constexpr u64 KEY_MAX = ~0ull;

// element i of the staged run: one padding slot per 16 keeps thread-blocked 64-bit accesses (stride ITEMS) off the
// same banks
__device__ __forceinline__ int sp(int i) { return i + (i >> 4); }
constexpr int SORT_SMEM_BYTES = (SORT_CAP + (SORT_CAP >> 4) + 16) * 8;

template <int ITEMS>
__device__ __forceinline__ void sort_registers(u64 (&k)[ITEMS]) {
    // bitonic sorting network on compile-time indices
#pragma unroll
    for (int size = 2; size <= ITEMS; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const int j = i ^ stride;
                if (j > i) {
                    const bool up = (i & size) == 0;
                    const u64 a = k[i], b = k[j];
                    const bool sw = up ? (a > b) : (a < b);
                    k[i] = sw ? b : a;
                    k[j] = sw ? a : b;
                }
            }
        }
    }
}

// merge of the two sorted runs s[a_off .. a_off+la) and s[b_off .. b_off+lb): the calling thread produces outputs
// [diag, diag + ITEMS) of the merged sequence into k (KEY_MAX past the end)
template <int ITEMS>
__device__ __forceinline__ void merge_path(const u64 *s, int a_off, int la, int b_off, int lb, int diag, u64 (&k)[ITEMS]) {
    int lo = max(0, diag - lb), hi = min(diag, la);
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s[sp(a_off + mid)] <= s[sp(b_off + diag - 1 - mid)]) lo = mid + 1; else hi = mid;
    }
    int a = lo, b = diag - lo;
    u64 ka = a < la ? s[sp(a_off + a)] : KEY_MAX, kb = b < lb ? s[sp(b_off + b)] : KEY_MAX;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        if (kb < ka) {
            k[i] = kb;
            ++b;
            kb = b < lb ? s[sp(b_off + b)] : KEY_MAX;
        } else {
            k[i] = ka;
            ++a;
            ka = a < la ? s[sp(a_off + a)] : KEY_MAX;
        }
    }
}

// sorts s[0 .. n) in place (n <= SORT_TPB * ITEMS; s[n .. padded) must hold KEY_MAX, padded = ITEMS << rounds)
template <int ITEMS>
__device__ __forceinline__ void sort_run(u64 *s, int n) {
    const int tid = threadIdx.x;
    const int runs = (n + ITEMS - 1) / ITEMS;
    int rounds = 0;
    while ((1 << rounds) < runs) ++rounds;
    const int padded = ITEMS << rounds;
    const int out0 = tid * ITEMS;
    const bool active = out0 < padded;
    u64 k[ITEMS];
    if (active) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) k[i] = s[sp(out0 + i)];
        sort_registers<ITEMS>(k);
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) s[sp(out0 + i)] = k[i];
    }
    __syncthreads();
    for (int L = ITEMS; L < padded; L <<= 1) {
        if (active) {
            const int pair = out0 & ~(2 * L - 1);
            merge_path<ITEMS>(s, pair, L, pair + L, L, out0 - pair, k);
        }
        __syncthreads();
        if (active) {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) s[sp(out0 + i)] = k[i];
        }
        __syncthreads();
    }
}

__device__ __forceinline__ void sort_run_any(u64 *s, int n) {
    if (n <= SORT_TPB * 2) sort_run<2>(s, n);
    else if (n <= SORT_TPB * 4) sort_run<4>(s, n);
    else if (n <= SORT_TPB * 8) sort_run<8>(s, n);
    else sort_run<16>(s, n);
}

// stages src[0 .. n) into s and pads to the size sort_run expects
__device__ __forceinline__ void stage_run(u64 *s, const u64 *__restrict__ src, int n) {
    const int items = n <= SORT_TPB * 2 ? 2 : (n <= SORT_TPB * 4 ? 4 : (n <= SORT_TPB * 8 ? 8 : 16));
    const int runs = (n + items - 1) / items;
    int rounds = 0;
    while ((1 << rounds) < runs) ++rounds;
    const int padded = items << rounds;
    for (int j = threadIdx.x; j < padded; j += SORT_TPB) s[sp(j)] = j < n ? src[j] : KEY_MAX;
    __syncthreads();
}

// first index a in [max(0, d - lb), min(d, la)] of the merge path of two sorted GLOBAL runs at diagonal d
__device__ __forceinline__ int global_split(const u64 *A, int la, const u64 *B, int lb, int d) {
    int lo = max(0, d - lb), hi = min(d, la);
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (A[mid] <= B[d - 1 - mid]) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(SORT_TPB, 2)
tb_sort_kernel(int T, int64_t I_cap, const int32_t *__restrict__ total_dev, const uint32_t *__restrict__ starts,
               u64 *entries, u64 *entries2, int32_t *__restrict__ ids_sorted, int32_t *__restrict__ tile_bins) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    u64 *s = reinterpret_cast<u64 *>(s_raw);
    __shared__ int s_split[2];
    gs_pdl_sync();
    if ((int64_t)*total_dev > I_cap) return;
    const int t = blockIdx.x, tid = threadIdx.x;
    const uint32_t start = starts[t];
    const int n = (int)(starts[t + 1] - start);
    if (tid == 0) {
        tile_bins[2 * t] = n ? (int32_t)start : 0;
        tile_bins[2 * t + 1] = n ? (int32_t)(start + n) : 0;
    }
    if (n == 0) return;
    u64 *seg = entries + start;
    int32_t *out = ids_sorted + start;
    if (n <= SORT_CAP) {
        stage_run(s, seg, n);
        sort_run_any(s, n);
        for (int j = tid; j < n; j += SORT_TPB) out[j] = (int32_t)(uint32_t)s[sp(j)];
        return;
    }
    // ---- long segment: sorted runs of SORT_CAP, then merge levels through global memory (seg <-> seg2) ----
    for (int r0 = 0; r0 < n; r0 += SORT_CAP) {
        const int m = min(SORT_CAP, n - r0);
        stage_run(s, seg + r0, m);
        sort_run_any(s, m);
        for (int j = tid; j < m; j += SORT_TPB) seg[r0 + j] = s[sp(j)];
        __syncthreads();
    }
    u64 *src = seg, *dst = entries2 + start;
    for (int L = SORT_CAP; L < n; L <<= 1) {
        __threadfence_block();
        for (long long p = 0; p < n; p += 2ll * L) {
            const int la = (int)min((long long)L, n - p), lb = (int)max(0ll, min((long long)L, n - p - L));
            const u64 *A = src + p, *B = src + p + la;
            for (int d0 = 0; d0 < la + lb; d0 += SORT_CAP) {
                const int d1 = min(d0 + SORT_CAP, la + lb);
                if (tid < 2) s_split[tid] = global_split(A, la, B, lb, tid == 0 ? d0 : d1);
                __syncthreads();
                const int a0 = s_split[0], a1 = s_split[1], b0 = d0 - a0, b1 = d1 - a1;
                const int ca = a1 - a0, cb = b1 - b0;
                for (int j = tid; j < ca; j += SORT_TPB) s[sp(j)] = A[a0 + j];
                for (int j = tid; j < cb; j += SORT_TPB) s[sp(ca + j)] = B[b0 + j];
                __syncthreads();
                u64 k[SORT_MAX_ITEMS];
                const int o0 = tid * SORT_MAX_ITEMS;
                if (o0 < ca + cb) merge_path<SORT_MAX_ITEMS>(s, 0, ca, ca, cb, o0, k);
                __syncthreads();
                if (o0 < ca + cb) {
#pragma unroll
                    for (int i = 0; i < SORT_MAX_ITEMS; ++i) s[sp(o0 + i)] = k[i];
                }
                __syncthreads();
                for (int j = tid; j < ca + cb; j += SORT_TPB) dst[p + d0 + j] = s[sp(j)];
                __syncthreads();
            }
        }
        __syncthreads();
        u64 *tmp = src; src = dst; dst = tmp;
    }
    __threadfence_block();
    for (int j = tid; j < n; j += SORT_TPB) out[j] = (int32_t)(uint32_t)src[j];
}

struct BucketLayout {
    size_t counts, cursors, starts, ctl, entries, entries2, total;
    size_t clear_bytes;     // counts + cursors + ctl, contiguous from offset 0: one memset
};

BucketLayout bucket_layout(int64_t I_cap, int T) {
    BucketLayout L;
    size_t at = 0;
    auto take = [&](size_t bytes) { size_t o = at; at += gs_align_up(bytes, 256); return o; };
    const size_t t = (size_t)(T > 0 ? T : 1);
    L.counts = take(t * TB_PAD * 4);
    L.cursors = take(t * TB_PAD * 4);
    L.ctl = take(256 + ((t + TS_TILE - 1) / TS_TILE + 1) * 8);
    L.clear_bytes = at;
    L.starts = take((t + 1) * 4);
    const size_t i8 = (size_t)(I_cap > 0 ? I_cap : 1) * 8;
    L.entries = take(i8);
    L.entries2 = take(i8);
    L.total = at;
    return L;
}

}  // namespace

size_t gs_bin_bucket_scratch_bytes(int64_t I_cap, int T) { return bucket_layout(I_cap, T).total; }

// can this view shape use the bucket pipeline? (the chunk histograms of count / fill live in shared memory)
bool gs_bin_bucket_supported(int T_view, int64_t I_cap) {
    return (size_t)T_view * 8 <= 160 * 1024 && I_cap < (1ll << 31);
}

// Same contract as gs_bin_fast_ex (binning.cu).
int gs_bin_bucket_ex(int N, int n_per_view, int64_t I_cap, const float *xys, int xy_stride, const float *depths,
                     const int32_t *radii, int tiles_x, int tiles_y, int32_t *ids_sorted, int32_t *tile_bins,
                     int32_t *total_dev, void *scratch, size_t scratch_bytes, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int views = N / n_per_view, T_view = tiles_x * tiles_y, T = views * T_view;
    const BucketLayout L = bucket_layout(I_cap, T);
    if (scratch_bytes < L.total) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(scratch)) return GS_ERR_MISALIGNED;
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    uint32_t *counts = (uint32_t *)(p + L.counts), *cursors = (uint32_t *)(p + L.cursors);
    uint32_t *starts = (uint32_t *)(p + L.starts);
    unsigned int *ticket = (unsigned int *)(p + L.ctl);
    unsigned long long *status = (unsigned long long *)(p + L.ctl + 256);
    u64 *entries = (u64 *)(p + L.entries), *entries2 = (u64 *)(p + L.entries2);

    GS_CUDA(cudaMemsetAsync(p, 0, L.clear_bytes, st));
    const int chunks_per_view = (n_per_view + TB_CHUNK - 1) / TB_CHUNK;
    const int grid = views * chunks_per_view;
    const size_t smem_count = (size_t)T_view * 4, smem_fill = (size_t)T_view * 8;
    if (smem_count > 48 * 1024)
        GS_CUDA(cudaFuncSetAttribute((const void *)tb_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_count));
    if (smem_fill > 48 * 1024)
        GS_CUDA(cudaFuncSetAttribute((const void *)tb_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_fill));
    GS_CUDA(cudaFuncSetAttribute((const void *)tb_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SORT_SMEM_BYTES));
    GS_PDL_LAUNCH(tb_count_kernel, grid, TB_TPB, smem_count, st, n_per_view, chunks_per_view, T_view, xys, xy_stride, radii,
                  tiles_x, tiles_y, counts);
    GS_PDL_LAUNCH(tb_scan_kernel, (T + TS_TILE - 1) / TS_TILE, TS_TPB, 0, st, T, (const uint32_t *)counts, starts, total_dev,
                  ticket, status);
    if (I_cap == 0) return 0;
    GS_PDL_LAUNCH(tb_fill_kernel, grid, TB_TPB, smem_fill, st, n_per_view, chunks_per_view, T_view, I_cap,
                  (const int32_t *)total_dev, xys, xy_stride, depths, radii, tiles_x, tiles_y, (const uint32_t *)starts, cursors,
                  entries);
    GS_PDL_LAUNCH(tb_sort_kernel, T, SORT_TPB, SORT_SMEM_BYTES, st, T, I_cap, (const int32_t *)total_dev,
                  (const uint32_t *)starts, entries, entries2, ids_sorted, tile_bins);
    return 0;
}

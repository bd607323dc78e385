// Stage 2 — tile binning: the work gsplat's rasterize_gaussians does
// before compositing (torch.cumsum + map_gaussian_to_intersects + torch.sort +
// get_tile_bin_edges), reached from /root/reference/gaussctrl/gc_model.py:208 and :225.
// Integer work, bit-exact against the oracle. HBM/L2-bound.
#include <stdlib.h>
#include <mutex>
#define GS_PDL_GROUP 1
#include "radix.cuh"

namespace {

// ------------------------------------------------------------------------------------------
// Inclusive int32 scan, single pass, decoupled look-back (warp-parallel window of 32 tiles).
// ------------------------------------------------------------------------------------------
constexpr int SCAN_TPB = 256;
constexpr int SCAN_IPT = 8;
constexpr int SCAN_TILE = SCAN_TPB * SCAN_IPT;
#define FLAG_AGG (1ull << 32)
#define FLAG_INC (2ull << 32)

// Prefixes beyond int32 (more than 2^31 - 1 tile intersections: huge radii at 1080p, many-view batches) SATURATE at
// INT32_MAX instead of wrapping: a wrapped negative prefix would pass the `offset + at < I_cap` guards of the emission
// and a wrapped total would pass `total > I_cap`. A saturated total is always above any capacity the entry points accept
// (I_cap < INT32_MAX), so every consumer sees "does not fit" and the host reports the count as unsupported. Operands
// are <= INT32_MAX, so the unsigned sum cannot wrap.
__device__ __forceinline__ unsigned sat_add31(unsigned a, unsigned b) { return min(a + b, 0x7fffffffu); }

struct ScanCtl {
    unsigned int ticket;
    unsigned int pad[3];
};

__global__ void __launch_bounds__(SCAN_TPB)
scan_kernel(int N, const int32_t *__restrict__ in, const uint32_t *__restrict__ gather, int exclusive,
            int32_t *__restrict__ out, int32_t *__restrict__ total, ScanCtl *ctl, unsigned long long *status) {
    __shared__ unsigned int s_tile;
    __shared__ int s_warp[SCAN_TPB / 32];
    __shared__ int s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    gs_pdl_sync();
    if (tid == 0) s_tile = atomicAdd(&ctl->ticket, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int base = tile * SCAN_TILE + tid * SCAN_IPT;
    int v[SCAN_IPT], raw[SCAN_IPT];
    if (gather) {   // in[gather[i]]: scan of tiles_hit in depth order
#pragma unroll
        for (int j = 0; j < SCAN_IPT; ++j) v[j] = base + j < N ? in[gather[base + j]] : 0;
    } else if (base + SCAN_IPT <= N) {
        const int4 a = *reinterpret_cast<const int4 *>(in + base);
        const int4 b = *reinterpret_cast<const int4 *>(in + base + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_IPT; ++j) v[j] = base + j < N ? in[base + j] : 0;
    }
#pragma unroll
    for (int j = 0; j < SCAN_IPT; ++j) raw[j] = v[j];
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
            unsigned part = lane <= first ? (unsigned int)w : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part = sat_add31(part, __shfl_xor_sync(0xffffffffu, part, o));
            excl = (int)sat_add31((unsigned)excl, part);
            if (inc_mask) break;
            look -= 32;
        }
        if (lane == 0) {
            if (tile > 0) atomicExch(&status[tile], FLAG_INC | sat_add31((unsigned)excl, (unsigned)block_total));
            s_prefix = excl;
        }
    }
    __syncthreads();
    const int off = s_prefix + thread_off;
    if (exclusive) {
#pragma unroll
        for (int j = 0; j < SCAN_IPT; ++j) v[j] -= raw[j];
    }
    if (base + SCAN_IPT <= N) {
        *reinterpret_cast<int4 *>(out + base) = make_int4(v[0] + off, v[1] + off, v[2] + off, v[3] + off);
        *reinterpret_cast<int4 *>(out + base + 4) = make_int4(v[4] + off, v[5] + off, v[6] + off, v[7] + off);
    } else {
#pragma unroll
        for (int j = 0; j < SCAN_IPT; ++j)
            if (base + j < N) out[base + j] = v[j] + off;
    }
    if (total && tile == (N - 1) / SCAN_TILE && tid == SCAN_TPB - 1)
        *total = (int)sat_add31((unsigned)s_prefix, (unsigned)block_total);
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

// tile_bins[t] = [first index with tile t, last+1); zero-filled beforehand.
__global__ void __launch_bounds__(256)
edges_kernel(int64_t I, const int64_t *__restrict__ ks, int32_t *__restrict__ bins) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= I) return;
    const int32_t t = (int32_t)(ks[k] >> 32);
    if (k == 0 || (int32_t)(ks[k - 1] >> 32) != t) bins[2 * t] = (int32_t)k;
    if (k == I - 1 || (int32_t)(ks[k + 1] >> 32) != t) bins[2 * t + 1] = (int32_t)(k + 1);
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
    scan_kernel<<<blocks, SCAN_TPB, 0, st>>>(N, tiles_hit, nullptr, 0, cum, total_dev, ctl, status);
    GS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------
// gs_bin_scan_static — gs_bin_scan for hosts that have no way to hand the library device memory of their own.
//
// Who needs it: gs_bin_scan wants a caller-provided look-back buffer (a ticket + one status word per 2 048-element
// tile). A torch caller just slices one out of its caching allocator, but not every host of this ABI has an
// allocator behind it: a plug-in that is only ever handed its input and output tensors by the application (an
// inference server's custom-op slot, a DCC / renderer plug-in, a cgo or JNI binding whose runtime owns all device
// memory and exposes no allocation call to native code), or code that runs at a point where cudaMalloc is off
// limits because it synchronises the whole device or would fail the surrounding stream capture. Such a caller can
// still pass tiles_hit / cum / total_dev, but has nowhere to get the scratch from — so here the scratch lives inside
// the library image, as a file-scope __device__ array.
//
// What that costs: GS_SCAN_STATIC_BYTES (64 KiB + 16 B) of device memory on every GPU this library is loaded on,
// held until the process exits, and a hard limit of GS_SCAN_STATIC_MAX_N elements per call (larger N is refused with
// GS_ERR_UNSUPPORTED; use gs_bin_scan with your own scratch for those). No cudaMalloc / cudaFree / allocating
// library call is made, on the first call or later.
//
// Sharing: there is ONE array per device and every call on that device uses it. Each call first clears the part it
// needs (the ticket and status words of the previous call are still in there) with a memset on `stream`, so calls
// that are ordered on the GPU — same stream, or ordered by events — are fine; two calls that can overlap on the GPU
// (different streams without ordering, or two host threads) race on the ticket and produce wrong sums.
//
// First call on a device: `g_scan_static` below is, in host code, only a handle; the kernel needs the array's device
// address, which cudaGetSymbolAddress returns for the CURRENT device. The first call made with a given device
// current looks it up (and thereby loads this module on that device) and remembers it; make that first call outside
// any stream capture. The device current at the call must be the one `stream` belongs to, as for every entry point.
// ------------------------------------------------------------------------------------------
namespace {
constexpr int GS_SCAN_STATIC_MAX_N_ = 16 * 1024 * 1024;                                  // = GS_SCAN_STATIC_MAX_N
constexpr int SCAN_STATIC_TILES = GS_SCAN_STATIC_MAX_N_ / SCAN_TILE + 1;                  // 8 193 status words
constexpr size_t SCAN_STATIC_WORDS = sizeof(ScanCtl) / sizeof(unsigned long long) + SCAN_STATIC_TILES;
static __device__ __align__(256) unsigned long long g_scan_static[SCAN_STATIC_WORDS];    // ticket | status[tiles]

constexpr int SCAN_STATIC_MAX_DEVICES = 64;
std::mutex g_scan_static_mu;
void *g_scan_static_ptr[SCAN_STATIC_MAX_DEVICES] = {};   // per device: address of g_scan_static there
}  // namespace

extern "C" int gs_bin_scan_static(int N, const int32_t *tiles_hit, int32_t *cum, int32_t *total_dev, void *stream) {
    static_assert(sizeof(ScanCtl) % sizeof(unsigned long long) == 0, "ticket block is a whole number of status words");
    static_assert(GS_SCAN_STATIC_MAX_N_ == GS_SCAN_STATIC_MAX_N && sizeof(g_scan_static) == GS_SCAN_STATIC_BYTES,
                  "header constants out of date");
    if (N < 0) return GS_ERR_BAD_ARG;
    if (N > GS_SCAN_STATIC_MAX_N) return GS_ERR_UNSUPPORTED;   // the built-in scratch has no room for more tiles
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) {
        if (total_dev) GS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), st));
        return 0;
    }
    if (!tiles_hit || !cum) return GS_ERR_BAD_ARG;
    if (!gs_aligned16(tiles_hit) || !gs_aligned16(cum)) return GS_ERR_MISALIGNED;

    // device address of the array on the current device, looked up once per device
    int dev = -1;
    GS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= SCAN_STATIC_MAX_DEVICES) return GS_ERR_UNSUPPORTED;
    void *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_scan_static_mu);
        if (!g_scan_static_ptr[dev]) {
            void *p = nullptr;
            GS_CUDA(cudaGetSymbolAddress(&p, g_scan_static));
            if (!p) return GS_ERR_UNSUPPORTED;
            g_scan_static_ptr[dev] = p;
        }
        scratch = g_scan_static_ptr[dev];
    }

    const size_t need = gs_scan_scratch_bytes(N);            // <= sizeof(g_scan_static) by the N check above
    GS_CUDA(cudaMemsetAsync(scratch, 0, need, st));          // the last call's ticket and flags are still in there
    ScanCtl *ctl = reinterpret_cast<ScanCtl *>(scratch);
    unsigned long long *status = reinterpret_cast<unsigned long long *>(ctl + 1);
    const int blocks = (N + SCAN_TILE - 1) / SCAN_TILE;
    scan_kernel<<<blocks, SCAN_TPB, 0, st>>>(N, tiles_hit, nullptr, 0, cum, total_dev, ctl, status);
    GS_LAUNCH_CHECK();
    return 0;
}

// internal (not part of the ABI): exclusive int32 scan used by view_exchange.cu
int gs_internal_exclusive_scan(int n, const int32_t *in, int32_t *out, int32_t *total_dev, void *scratch,
                               size_t scratch_bytes, cudaStream_t st) {
    if (n <= 0) return 0;
    const size_t need = gs_scan_scratch_bytes(n);
    if (scratch_bytes < need) return GS_ERR_SCRATCH_SMALL;
    GS_CUDA(cudaMemsetAsync(scratch, 0, need, st));
    ScanCtl *ctl = reinterpret_cast<ScanCtl *>(scratch);
    unsigned long long *status = reinterpret_cast<unsigned long long *>(ctl + 1);
    scan_kernel<<<(n + SCAN_TILE - 1) / SCAN_TILE, SCAN_TPB, 0, st>>>(n, in, nullptr, 1, out, total_dev, ctl, status);
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
    // ping-pong buffers + CTA x digit matrix
    const size_t keys = gs_align_up((size_t)I * 8, 256), vals = gs_align_up((size_t)I * 4, 256);
    return keys + vals + gs_align_up(gsradix::mat_entries(I, 256) * sizeof(uint32_t), 256) + 256;
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
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    uint64_t *kb = reinterpret_cast<uint64_t *>(p); p += gs_align_up((size_t)I * 8, 256);
    uint32_t *vb = reinterpret_cast<uint32_t *>(p); p += gs_align_up((size_t)I * 4, 256);
    uint32_t *mat = reinterpret_cast<uint32_t *>(p);
    // live bits: 32 depth bits + ceil(log2 T) tile bits, split evenly into <= 8-bit digits
    const int total_bits = 32 + (T > 1 ? gs_ceil_log2(T) : 1);
    const int npass = (total_bits + 7) / 8;
    const uint64_t *kin = reinterpret_cast<const uint64_t *>(keys);
    const uint32_t *vin = reinterpret_cast<const uint32_t *>(ids);
    uint64_t *ka = reinterpret_cast<uint64_t *>(keys_sorted);
    uint32_t *va = reinterpret_cast<uint32_t *>(ids_sorted);
    bool to_a = (npass & 1) != 0;   // the final pass must land in the caller's sorted outputs
    int at = 0;
    for (int q = 0; q < npass; ++q) {
        const int left = total_bits - at, passes_left = npass - q;
        const int b = (left + passes_left - 1) / passes_left;
        uint64_t *ko = to_a ? ka : kb;
        uint32_t *vo = to_a ? va : vb;
        const int rc = gsradix::launch_pass<uint64_t, 8>(I, nullptr, kin, vin, ko, vo, at, b, mat, (cudaStream_t)stream);
        if (rc) return rc;
        kin = ko; vin = vo; to_a = !to_a; at += b;
    }
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

// ==========================================================================================
// Traffic-lean binning (gs_bin_fast): same ids_sorted / tile_bins as scan + emit + 64-bit sort +
// edges, without ever materialising 64-bit keys (SURVEY.md §7 "sort bandwidth"):
//   1. stable radix sort of the N Gaussians by depth bits (32-bit keys, three passes over N, not I)
//   2. exclusive scan of tiles_hit in that order (single-pass look-back, gathered input)
//   3. emission of (tile id, Gaussian id) in that order
//   4. ONE stable partition by tile id (<= 2048 tiles: a single pass over I)
//   5. tile ranges from the partitioned tile ids
// sort-by-depth then stable-partition-by-tile == sort by (tile, depth) with ties in id order,
// i.e. exactly the order of a stable sort of the 64-bit keys. The intersection count never
// leaves the device: steps 3-5 read it from total_dev and are launched for the capacity.
// ==========================================================================================
namespace {

__global__ void __launch_bounds__(256)
emit_sorted_kernel(int N, int64_t I_cap, const int32_t *__restrict__ total_dev, const uint32_t *__restrict__ order,
                   const int32_t *__restrict__ start, const float *__restrict__ xy_base, int xy_stride,
                   const int32_t *__restrict__ radii, int tiles_x, int tiles_y, int n_per_view,
                   uint32_t *__restrict__ tkey, uint32_t *__restrict__ tval) {
    gs_pdl_sync();
    if ((int64_t)*total_dev > I_cap) return;   // capacity too small: write nothing, the host retries
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    // several views in one call (gs_view_fwd, views > 1): element g belongs to view g / n_per_view and its tiles
    // live in that view's slice [view * T, (view + 1) * T) of the tile-id space
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0, st = 0, g = 0, tbase = 0;
    if (j < N) {
        g = (int)order[j];
        const int r = radii[g];
        if (r > 0) {
            const float2 c = *reinterpret_cast<const float2 *>(xy_base + (size_t)g * xy_stride);
            gs_tile_bbox(c.x, c.y, (float)r, tiles_x, tiles_y, x0, y0, x1, y1);
            st = start[j];
            tbase = (g / n_per_view) * (tiles_x * tiles_y);
        }
    }
    const int bw = x1 - x0, area = bw * (y1 - y0);
    if (area > 0 && area <= 8) {
        int at = st;
        for (int ty = y0; ty < y1; ++ty)
            for (int tx = x0; tx < x1; ++tx) {
                tkey[at] = (uint32_t)(tbase + ty * tiles_x + tx);
                tval[at] = (uint32_t)g;
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
        const int stb = __shfl_sync(0xffffffffu, tbase, src);
        for (int k = lane; k < sarea; k += 32) {
            tkey[sst + k] = (uint32_t)(stb + (sy0 + k / sbw) * tiles_x + sx0 + k % sbw);
            tval[sst + k] = (uint32_t)sg;
        }
    }
}

// Scan + emission in ONE kernel: what scan_kernel (exclusive scan of tiles_hit in depth order) followed by
// emit_sorted_kernel did in two, without the `start` array and without reading tiles_hit at all — a Gaussian's tile
// count IS the area of the bbox that the emission recomputes from (xy, radius) anyway (A.1.6), so the scan's 1 M random
// 4-byte gathers (its whole cost: profiles/README.md item 7) disappear. Each CTA takes 1 024 consecutive elements of the
// depth order (4 per thread), scans their areas, gets its global offset by decoupled look-back (ticket order, 32-wide
// window) and writes its (tile, id) pairs. The total is published by the last tile; writes are bounded by the capacity
// (if the total exceeds it the consumers see that in *total_dev and do nothing).
// CTA shape of scan_emit_kernel: 128 threads x 4 elements with a 2 048-pair stage (measured against 256 / 4 096, the
// round-2 form: 145.6 -> 142.7 us for the chain at 1 M / 512^2, 282.9 -> 271.8 us at 1080p; 64 / 1 024 the same)
#ifndef GS_SE_TPB
#define GS_SE_TPB 128
#endif
#ifndef GS_SE_STAGE
#define GS_SE_STAGE 2048
#endif
constexpr int SE_TPB = GS_SE_TPB;
constexpr int SE_IPT = 4;
constexpr int SE_TILE = SE_TPB * SE_IPT;
constexpr int SE_STAGE = GS_SE_STAGE;
static_assert(SE_TPB >= 64 && (SE_TPB & (SE_TPB - 1)) == 0, "scan_emit_kernel: power-of-two CTA sizes only (192 x 4 miscounts)");       // pairs of one CTA staged in shared memory (a CTA's pairs are CONTIGUOUS in the output)

__global__ void __launch_bounds__(SE_TPB)
scan_emit_kernel(int N, int64_t I_cap, const uint32_t *__restrict__ order, const float *__restrict__ xy_base, int xy_stride,
                 const int32_t *__restrict__ radii, int tiles_x, int tiles_y, int n_per_view, uint32_t *__restrict__ tkey,
                 uint32_t *__restrict__ tval, int32_t *__restrict__ total, ScanCtl *ctl, unsigned long long *status) {
    __shared__ unsigned int s_tile;
    __shared__ int s_warp[SE_TPB / 32];
    __shared__ int s_prefix;
    // the pairs of this CTA form one contiguous run of the output: they are assembled here and stored with coalesced
    // 128-byte rows. Written straight from the threads, every lane of a store lands in a different sector (a thread
    // owns ~12 consecutive pairs) and the kernel is bound by the LSU's sector rate (ncu: mio_throttle, r02).
    __shared__ uint32_t s_key[SE_STAGE], s_val[SE_STAGE];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    gs_pdl_sync();
    if (tid == 0) s_tile = atomicAdd(&ctl->ticket, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int base = tile * SE_TILE + tid * SE_IPT;
    int g[SE_IPT], x0[SE_IPT], y0[SE_IPT], bw[SE_IPT], area[SE_IPT], tb[SE_IPT];
    if (base + SE_IPT <= N) {
        const uint4 o = *reinterpret_cast<const uint4 *>(order + base);
        g[0] = (int)o.x; g[1] = (int)o.y; g[2] = (int)o.z; g[3] = (int)o.w;
    } else {
#pragma unroll
        for (int k = 0; k < SE_IPT; ++k) g[k] = base + k < N ? (int)order[base + k] : -1;
    }
    int rr[SE_IPT];
#pragma unroll
    for (int k = 0; k < SE_IPT; ++k) rr[k] = g[k] >= 0 ? radii[g[k]] : 0;       // 4 independent gathers in flight
#pragma unroll
    for (int k = 0; k < SE_IPT; ++k) {
        x0[k] = y0[k] = bw[k] = area[k] = tb[k] = 0;
        if (rr[k] > 0) {
            const float2 c = *reinterpret_cast<const float2 *>(xy_base + (size_t)g[k] * xy_stride);
            int x1, y1;
            gs_tile_bbox(c.x, c.y, (float)rr[k], tiles_x, tiles_y, x0[k], y0[k], x1, y1);
            bw[k] = x1 - x0[k];
            area[k] = bw[k] * (y1 - y0[k]);
            tb[k] = (g[k] / n_per_view) * (tiles_x * tiles_y);
        }
    }
    int excl[SE_IPT], sum = 0;
#pragma unroll
    for (int k = 0; k < SE_IPT; ++k) { excl[k] = sum; sum += area[k]; }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int warp_off = 0, block_total = 0;
#pragma unroll
    for (int w = 0; w < SE_TPB / 32; ++w) {
        if (w < warp) warp_off += s_warp[w];
        block_total += s_warp[w];
    }
    const int thread_off = warp_off + inc - sum;
    if (warp == 0) {
        if (lane == 0)
            atomicExch(&status[tile], (tile == 0 ? FLAG_INC : FLAG_AGG) | (unsigned int)block_total);
        int ex = 0;
        int look = tile - 1;
        while (look >= 0) {
            const int idx = look - lane;
            unsigned long long w;
            do {
                w = idx >= 0 ? *reinterpret_cast<volatile unsigned long long *>(&status[idx]) : FLAG_INC;
            } while (__any_sync(0xffffffffu, (w >> 32) == 0));
            const unsigned inc_mask = __ballot_sync(0xffffffffu, (w >> 32) == 2);
            const int first = inc_mask ? __ffs(inc_mask) - 1 : 31;
            unsigned part = lane <= first ? (unsigned int)w : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part = sat_add31(part, __shfl_xor_sync(0xffffffffu, part, o));
            ex = (int)sat_add31((unsigned)ex, part);
            if (inc_mask) break;
            look -= 32;
        }
        if (lane == 0) {
            if (tile > 0) atomicExch(&status[tile], FLAG_INC | sat_add31((unsigned)ex, (unsigned)block_total));
            s_prefix = ex;
        }
    }
    __syncthreads();
    const int cta_off = s_prefix;
    if (tile == (N - 1) / SE_TILE && tid == SE_TPB - 1) *total = (int)sat_add31((unsigned)s_prefix, (unsigned)block_total);
    const bool staged = block_total <= SE_STAGE;
    // pair `at` (index inside this CTA's run): shared-memory stage, or straight to the output when the run is too long
    auto put = [&](int at, uint32_t t, uint32_t id) {
        if (staged) {
            s_key[at] = t;
            s_val[at] = id;
        } else if ((int64_t)cta_off + at < I_cap) {
            tkey[cta_off + at] = t;
            tval[cta_off + at] = id;
        }
    };
    // ---- emission: row-major over the bbox (outer y, inner x), small splats by their own thread, big ones by the warp
#pragma unroll
    for (int k = 0; k < SE_IPT; ++k) {
        const int st = thread_off + excl[k];
        if (area[k] > 0 && area[k] <= 8) {
            int at = st;
            for (int ty = y0[k]; ty < y0[k] + area[k] / bw[k]; ++ty)
                for (int tx = x0[k]; tx < x0[k] + bw[k]; ++tx) put(at++, (uint32_t)(tb[k] + ty * tiles_x + tx), (uint32_t)g[k]);
        }
        unsigned big = __ballot_sync(0xffffffffu, area[k] > 8);
        while (big) {
            const int src = __ffs(big) - 1;
            big &= big - 1;
            const int sx0 = __shfl_sync(0xffffffffu, x0[k], src), sy0 = __shfl_sync(0xffffffffu, y0[k], src);
            const int sbw = __shfl_sync(0xffffffffu, bw[k], src), sarea = __shfl_sync(0xffffffffu, area[k], src);
            const int sst = __shfl_sync(0xffffffffu, st, src), sg = __shfl_sync(0xffffffffu, g[k], src);
            const int stb = __shfl_sync(0xffffffffu, tb[k], src);
            for (int q = lane; q < sarea; q += 32)
                put(sst + q, (uint32_t)(stb + (sy0 + q / sbw) * tiles_x + sx0 + q % sbw), (uint32_t)sg);
        }
    }
    if (staged) {
        __syncthreads();
        for (int at = tid; at < block_total; at += SE_TPB)
            if ((int64_t)cta_off + at < I_cap) {
                tkey[cta_off + at] = s_key[at];
                tval[cta_off + at] = s_val[at];
            }
    }
}

// tile ranges from the sorted tile ids (two-pass partition only). Four consecutive entries per thread (one 16-byte
// load), the neighbours across threads by shuffle: the first version read tk[k-1], tk[k], tk[k+1] per entry and cost
// 146 us for the 23 M pairs of an 8-view batch — as much as a whole radix pass.
__global__ void __launch_bounds__(256)
edges32_kernel(int64_t I_cap, const int32_t *__restrict__ total_dev, const uint32_t *__restrict__ tk,
               int32_t *__restrict__ bins) {
    gs_pdl_sync();
    const int64_t I = *total_dev;
    if (I > I_cap) return;
    const int lane = threadIdx.x & 31;
    const int64_t k0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int64_t w0 = k0 - 4 * lane;                         // first entry of the warp
    if (w0 >= I) return;                                      // warp-uniform
    uint32_t t[4] = {0u, 0u, 0u, 0u};
    if (k0 + 4 <= I) {
        const uint4 v = *reinterpret_cast<const uint4 *>(tk + k0);
        t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (k0 + j < I) t[j] = tk[k0 + j];
    }
    uint32_t prev = __shfl_up_sync(0xffffffffu, t[3], 1), next = __shfl_down_sync(0xffffffffu, t[0], 1);
    if (lane == 0 && k0 > 0) prev = tk[k0 - 1];
    if (lane == 31 && k0 + 4 < I) next = tk[k0 + 4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t k = k0 + j;
        if (k >= I) break;
        const uint32_t before = j ? t[j - 1] : prev, after = j < 3 ? t[j + 1] : next;
        if (k == 0 || before != t[j]) bins[2 * t[j]] = (int32_t)k;
        if (k == I - 1 || after != t[j]) bins[2 * t[j] + 1] = (int32_t)(k + 1);
    }
}

struct FastLayout {
    size_t keyA, valA, keyB, valB, start, tkey, tval, tkey2, tval2, mat, ctl, total;
};

FastLayout fast_layout(int N, int64_t I_cap) {
    FastLayout L;
    size_t at = 0;
    auto take = [&](size_t bytes) { size_t o = at; at += gs_align_up(bytes, 256); return o; };
    const size_t n4 = (size_t)(N > 0 ? N : 1) * 4, i4 = (size_t)(I_cap > 0 ? I_cap : 1) * 4;
    L.keyA = take(n4); L.valA = take(n4); L.keyB = take(n4); L.valB = take(n4); L.start = take(n4);
    L.tkey = take(i4); L.tval = take(i4); L.tkey2 = take(i4); L.tval2 = take(i4);
    const size_t m = gsradix::mat_entries(N, 2048), mi = gsradix::mat_entries(I_cap, 2048);
    L.mat = take((m > mi ? m : mi) * 4);
    L.ctl = take((SCAN_TILE / SE_TILE) * gs_scan_scratch_bytes(N));      // look-back state of scan_emit_kernel (SE_TILE-element tiles)
    L.total = at;
    return L;
}

// number of LSD passes of the depth sort (4 = 8-bit digits, default; 3 = 11/11/10-bit digits, measured slower:
// 191 vs 169 us for the chain at 1 M Gaussians — 80 KB of shared memory per scatter CTA); initial value from the
// environment (GS_DEPTH_SORT_PASSES), changed at run time by gs_set_depth_sort_passes (tests compare both)
int &depth_sort_passes() {
    static int passes = [] {
        const char *e = getenv("GS_DEPTH_SORT_PASSES");
        return (e && e[0] == '3') ? 3 : 4;
    }();
    return passes;
}

// implementation behind gs_bin_fast: 1 = LSD pipeline (this file, default), 0 = bucket pipeline (tilebucket.cu)
int &bin_impl() {
    static int impl = [] {
        const char *e = getenv("GS_BIN_IMPL");
        return (e && (e[0] == 'b' || e[0] == '0')) ? 0 : 1;       // GS_BIN_IMPL=bucket
    }();
    return impl;
}

// CTA shape of the radix passes of gs_bin_fast (radix.cuh: launch_pass_cfg), [0] = depth sort over N elements,
// [1] = partition over I pairs. -1 = chosen from the problem size; GS_RADIX_CFG_DEPTH / GS_RADIX_CFG_PART or
// gs_set_radix_cfg pin one (A/B)
int g_radix_cfg[2] = {-2, -2};
int radix_cfg(int which, int64_t n) {
    if (g_radix_cfg[which] == -2) {
        const char *e = getenv(which == 0 ? "GS_RADIX_CFG_DEPTH" : "GS_RADIX_CFG_PART");
        g_radix_cfg[which] = (e && e[0] >= '0' && e[0] <= '3') ? e[0] - '0' : -1;
    }
    if (g_radix_cfg[which] >= 0) return g_radix_cfg[which];
    if (which == 1) return n <= (8ll << 20) ? 1 : 2;       // n = capacity of the pair list (~1.5x the pairs)
    // depth sort: a pass is one to a few waves of CTAs and as long as its slowest SM — take the chunk size that
    // leaves the fewest keys on the busiest SM (measured: 384 x 8 at 200 K / 2 M / 5 M elements, 512 x 8 at 0.5 / 1 M)
    const int64_t c4 = (n + 4095) / 4096, c3 = (n + 3071) / 3072;
    const int64_t w4 = (c4 + GS_NUM_SMS - 1) / GS_NUM_SMS * 4096, w3 = (c3 + GS_NUM_SMS - 1) / GS_NUM_SMS * 3072;
    return w4 < w3 ? 1 : 2;
}

template <int BITS>
int partition_pass(int64_t I_cap, const int32_t *total_dev, const uint32_t *kin, const uint32_t *vin, uint32_t *kout,
                   uint32_t *vout, int shift, int bits, uint32_t *mat, cudaStream_t st, int32_t *bins = nullptr,
                   int T = 0) {
    return gsradix::launch_pass_cfg<uint32_t, BITS>(radix_cfg(1, I_cap), I_cap, total_dev, kin, vin, kout, vout, shift, bits,
                                                    mat, st, bins, T);
}

}  // namespace

extern "C" int gs_set_bin_impl(int impl) {
    const int prev = bin_impl();
    if (impl == 0 || impl == 1) bin_impl() = impl;
    return prev;
}

extern "C" int gs_set_radix_cfg(int which, int cfg) {
    if (which < 0 || which > 1) return GS_ERR_BAD_ARG;
    radix_cfg(which, 0);
    const int prev = g_radix_cfg[which];
    if (cfg >= -1 && cfg <= 3) g_radix_cfg[which] = cfg;
    return prev;
}

extern "C" int gs_set_depth_sort_passes(int passes) {
    const int prev = depth_sort_passes();
    if (passes == 3 || passes == 4) depth_sort_passes() = passes;
    return prev;
}

extern "C" size_t gs_bin_fast_scratch_bytes(int N, int64_t I_cap, int T) {
    const size_t lsd = fast_layout(N, I_cap).total, bucket = gs_bin_bucket_scratch_bytes(I_cap, T > 0 ? T : 1);
    return lsd > bucket ? lsd : bucket;       // either implementation can run in the same scratch buffer
}

// N = elements of ALL views (view v owns [v * n_per_view, (v + 1) * n_per_view)); xy of element g at
// xy_base + g * xy_stride floats (2 = an (N,2) array, 12 = the packed records); tile_bins has views * T rows.
// Clears of gs_bin_fast_ex (tile ranges, look-back state) issued AHEAD of the kernels that feed it: gs_view_fwd calls this
// before its preprocess launches so that no memset node sits between the preprocess kernel and the first radix pass
// (every kernel of the chain is a programmatic dependent launch; a memset in between breaks the overlap). Returns 1 when
// the clears were done (pass precleared = true to gs_bin_fast_ex), 0 when the call should do them itself, < 0 on error.
int gs_bin_fast_clear(int N, int64_t I_cap, int T, int32_t *tile_bins, void *scratch, size_t scratch_bytes, void *stream) {
    if (N <= 0 || T <= 0 || !tile_bins || !scratch || bin_impl() != 1) return 0;
    const FastLayout L = fast_layout(N, I_cap);
    if (scratch_bytes < L.total) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)T, st) != cudaSuccess) return -1;
    if (cudaMemsetAsync(reinterpret_cast<unsigned char *>(scratch) + L.ctl, 0, (SCAN_TILE / SE_TILE) * gs_scan_scratch_bytes(N), st) != cudaSuccess)
        return -1;
    return 1;
}

int gs_bin_fast_ex(int N, int n_per_view, int64_t I_cap, const float *xys, int xy_stride, const float *depths,
                   const int32_t *radii, const int32_t *tiles_hit, int tiles_x, int tiles_y, int32_t *ids_sorted,
                   int32_t *tile_bins, int32_t *total_dev, void *scratch, size_t scratch_bytes, void *stream,
                   bool precleared) {
    if (N < 0 || I_cap < 0 || tiles_x <= 0 || tiles_y <= 0 || n_per_view <= 0 || (N % n_per_view) != 0) return GS_ERR_BAD_ARG;
    if (!tile_bins || !total_dev || !scratch) return GS_ERR_BAD_ARG;
    if (I_cap >= 0x7fffffffll) return GS_ERR_UNSUPPORTED;   // INT32_MAX is the saturated "too many pairs" total
    const int views = N > 0 ? N / n_per_view : 1;
    if ((int64_t)views * tiles_x * tiles_y >= (1ll << 22)) return GS_ERR_UNSUPPORTED;
    const int T = views * tiles_x * tiles_y;
    const int tile_bits = T > 1 ? gs_ceil_log2(T) : 1;
    if (tile_bits > 22) return GS_ERR_UNSUPPORTED;
    cudaStream_t st = (cudaStream_t)stream;
    if (!precleared) GS_CUDA(cudaMemsetAsync(tile_bins, 0, sizeof(int32_t) * 2 * (size_t)T, st));
    if (N == 0) {
        GS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), st));
        return 0;
    }
    if (!xys || !depths || !radii || !tiles_hit || (!ids_sorted && I_cap > 0)) return GS_ERR_BAD_ARG;
    if (bin_impl() == 0 && gs_bin_bucket_supported(tiles_x * tiles_y, I_cap))
        return gs_bin_bucket_ex(N, n_per_view, I_cap, xys, xy_stride, depths, radii, tiles_x, tiles_y, ids_sorted, tile_bins,
                                total_dev, scratch, scratch_bytes, stream);
    const FastLayout L = fast_layout(N, I_cap);
    if (scratch_bytes < L.total) return GS_ERR_SCRATCH_SMALL;
    if (!gs_aligned16(scratch)) return GS_ERR_MISALIGNED;
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    uint32_t *keyA = (uint32_t *)(p + L.keyA), *valA = (uint32_t *)(p + L.valA);
    uint32_t *keyB = (uint32_t *)(p + L.keyB), *valB = (uint32_t *)(p + L.valB);
    int32_t *start = (int32_t *)(p + L.start);
    uint32_t *tkey = (uint32_t *)(p + L.tkey), *tval = (uint32_t *)(p + L.tval);
    uint32_t *tkey2 = (uint32_t *)(p + L.tkey2), *tval2 = (uint32_t *)(p + L.tval2);
    uint32_t *mat = (uint32_t *)(p + L.mat);
    ScanCtl *ctl = reinterpret_cast<ScanCtl *>(p + L.ctl);
    unsigned long long *status = reinterpret_cast<unsigned long long *>(ctl + 1);
    const int nblk = (N + 255) / 256;
    int rc;

    // the scan's look-back state is cleared up front so that the kernel chain below is uninterrupted (every launch
    // is a programmatic dependent launch: the next kernel is resident before the previous one drains)
    if (!precleared) GS_CUDA(cudaMemsetAsync(ctl, 0, (SCAN_TILE / SE_TILE) * gs_scan_scratch_bytes(N), st));
    // 1. depth sort (values start as the element index; the first pass forms its keys from depths/radii on the fly,
    //    so there is no key-building kernel). Default: four passes of 8-bit digits; gs_set_depth_sort_passes(3) /
    //    GS_DEPTH_SORT_PASSES=3 selects three passes of 11/11/10-bit digits (fewer launches, but measured slower).
    //    Any split of the 32 bits into stable LSD passes yields the same order.
    const uint32_t *dkeys = reinterpret_cast<const uint32_t *>(depths);
    const uint32_t *order;
    if (depth_sort_passes() == 4) {
        const int cd = radix_cfg(0, N);
        if ((rc = gsradix::launch_pass_cfg<uint32_t, 8, 1>(cd, N, nullptr, dkeys, nullptr, keyB, valB, 0, 8, mat, st, nullptr, 0,
                                                          radii))) return rc;
        if ((rc = gsradix::launch_pass_cfg<uint32_t, 8>(cd, N, nullptr, keyB, valB, keyA, valA, 8, 8, mat, st))) return rc;
        if ((rc = gsradix::launch_pass_cfg<uint32_t, 8>(cd, N, nullptr, keyA, valA, keyB, valB, 16, 8, mat, st))) return rc;
        if ((rc = gsradix::launch_pass_cfg<uint32_t, 8>(cd, N, nullptr, keyB, valB, (uint32_t *)nullptr, valA, 24, 8, mat, st)))
            return rc;
        order = valA;
    } else {
        if ((rc = gsradix::launch_pass<uint32_t, 11, true>(N, nullptr, dkeys, nullptr, keyB, valB, 0, 11, mat, st, nullptr,
                                                          0, radii))) return rc;
        if ((rc = gsradix::launch_pass<uint32_t, 11>(N, nullptr, keyB, valB, keyA, valA, 11, 11, mat, st))) return rc;
        if ((rc = gsradix::launch_pass<uint32_t, 11>(N, nullptr, keyA, valA, (uint32_t *)nullptr, valB, 22, 10, mat, st)))
            return rc;
        order = valB;
    }

    // 2 + 3. scan of the tile counts and emission, both in depth order, in one kernel (GS_BIN_SPLIT_SCAN=1: the two
    //        separate kernels of round 1, kept for A/B). An element's count is recomputed from (xy, radius), so
    //        tiles_hit is not read at all.
    static const bool split_scan = getenv("GS_BIN_SPLIT_SCAN") != nullptr;
    if (!split_scan) {
        GS_PDL_LAUNCH(scan_emit_kernel, (N + SE_TILE - 1) / SE_TILE, SE_TPB, 0, st, N, I_cap, order, xys, xy_stride, radii,
                      tiles_x, tiles_y, n_per_view, tkey, tval, total_dev, ctl, status);
        if (I_cap == 0) return 0;
    } else {
        GS_PDL_LAUNCH(scan_kernel, (N + SCAN_TILE - 1) / SCAN_TILE, SCAN_TPB, 0, st, N, tiles_hit, order, 1, start, total_dev,
                      ctl, status);
        if (I_cap == 0) return 0;
        GS_PDL_LAUNCH(emit_sorted_kernel, nblk, 256, 0, st, N, I_cap, (const int32_t *)total_dev, order,
                      (const int32_t *)start, xys, xy_stride, radii, tiles_x, tiles_y, n_per_view, tkey, tval);
    }

    // 4. stable partition by tile id
    uint32_t *ids_out = reinterpret_cast<uint32_t *>(ids_sorted);
    // single pass: the digit is the tile id, so the tile ranges are the exclusive scan of the digit totals
    // (written by the pass itself) and the sorted tile ids need not be materialised at all
    const uint32_t *tk_sorted = nullptr;
    if (tile_bits <= 8) {
        rc = partition_pass<8>(I_cap, total_dev, tkey, tval, nullptr, ids_out, 0, tile_bits, mat, st, tile_bins, T);
    } else if (tile_bits <= 10) {
        rc = partition_pass<10>(I_cap, total_dev, tkey, tval, nullptr, ids_out, 0, tile_bits, mat, st, tile_bins, T);
    } else if (tile_bits <= 11) {
        rc = partition_pass<11>(I_cap, total_dev, tkey, tval, nullptr, ids_out, 0, tile_bits, mat, st, tile_bins, T);
    } else {
        const int lo_bits = (tile_bits + 1) / 2, hi_bits = tile_bits - lo_bits;
        if (lo_bits <= 8) {
            rc = partition_pass<8>(I_cap, total_dev, tkey, tval, tkey2, tval2, 0, lo_bits, mat, st);
            if (!rc) rc = partition_pass<8>(I_cap, total_dev, tkey2, tval2, tkey, ids_out, lo_bits, hi_bits, mat, st);
        } else {
            rc = partition_pass<11>(I_cap, total_dev, tkey, tval, tkey2, tval2, 0, lo_bits, mat, st);
            if (!rc) rc = partition_pass<11>(I_cap, total_dev, tkey2, tval2, tkey, ids_out, lo_bits, hi_bits, mat, st);
        }
        tk_sorted = tkey;
    }
    if (rc) return rc;

    // 5. tile ranges (two-pass partition only)
    if (tk_sorted) {
        edges32_kernel<<<(int)((I_cap + 1023) / 1024), 256, 0, st>>>(I_cap, total_dev, tk_sorted, tile_bins);
        GS_LAUNCH_CHECK();
    }
    return 0;
}

extern "C" int gs_bin_fast(int N, int64_t I_cap, const float *xys, const float *depths, const int32_t *radii,
                           const int32_t *tiles_hit, int tiles_x, int tiles_y,
                           int32_t *ids_sorted, int32_t *tile_bins, int32_t *total_dev,
                           void *scratch, size_t scratch_bytes, void *stream) {
    return gs_bin_fast_ex(N, N > 0 ? N : 1, I_cap, xys, 2, depths, radii, tiles_hit, tiles_x, tiles_y, ids_sorted, tile_bins,
                          total_dev, scratch, scratch_bytes, stream, false);
}

// ==========================================================================================
// Deterministic per-Gaussian totals of per-intersection rows (debug switch gs_view_bwd_det): stable radix sort of the
// list positions k by Gaussian id (ids_sorted[k]), segment bounds from the sorted ids, then ONE thread per Gaussian
// adds its rows in list order — a fixed order of float additions, so two runs agree bit for bit.
// ==========================================================================================
namespace {

__global__ void __launch_bounds__(256)
segment_bounds_kernel(int64_t I_cap, const int32_t *__restrict__ total_dev, const uint32_t *__restrict__ ids_by_id,
                      int32_t *__restrict__ seg) {
    const int64_t I = *total_dev;
    if (I > I_cap) return;
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= I) return;
    const uint32_t g = ids_by_id[j];
    if (j == 0 || ids_by_id[j - 1] != g) seg[2 * g] = (int32_t)j;
    if (j == I - 1 || ids_by_id[j + 1] != g) seg[2 * g + 1] = (int32_t)(j + 1);
}

__global__ void __launch_bounds__(256)
segment_sum_kernel(int n, size_t Np, const int32_t *__restrict__ seg, const uint32_t *__restrict__ pos_by_id,
                   const float *__restrict__ rows, float *__restrict__ gbuf) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const int lo = seg[2 * g], hi = seg[2 * g + 1];
    if (hi <= lo) return;                                   // no intersection: the zero-filled block stays
    float a[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) a[c] = 0.f;
    for (int j = lo; j < hi; ++j) {
        const float *r = rows + (size_t)pos_by_id[j] * 9;
#pragma unroll
        for (int c = 0; c < 9; ++c) a[c] += r[c];
    }
    gbuf[2 * (size_t)g] = a[0]; gbuf[2 * (size_t)g + 1] = a[1];
    float *vc = gbuf + 2 * Np + 3 * (size_t)g, *vr = gbuf + 5 * Np + 3 * (size_t)g;
    vc[0] = a[2]; vc[1] = a[3]; vc[2] = a[4];
    gbuf[8 * Np + g] = a[5];
    vr[0] = a[6]; vr[1] = a[7]; vr[2] = a[8];
}

struct SegLayout {
    size_t keyA, valA, keyB, valB, mat, seg, total;
};
SegLayout seg_layout(int n, int64_t I_cap) {
    SegLayout L;
    size_t at = 0;
    auto take = [&](size_t bytes) { size_t o = at; at += gs_align_up(bytes, 256); return o; };
    const size_t i4 = (size_t)(I_cap > 0 ? I_cap : 1) * 4;
    L.keyA = take(i4); L.valA = take(i4); L.keyB = take(i4); L.valB = take(i4);
    L.mat = take(gsradix::mat_entries(I_cap, 256) * 4);
    L.seg = take((size_t)(n > 0 ? n : 1) * 8);
    L.total = at;
    return L;
}

}  // namespace

size_t gs_internal_segment_scratch_bytes(int n, int64_t I_cap) { return seg_layout(n, I_cap).total; }

// rows: (I, 9) per-intersection gradients in ids_sorted order. gbuf: zero-filled compositor-gradient block
// (v_xy | v_conic | v_rgb | v_opacity, strides of Np). total_dev: device intersection count.
int gs_internal_segment_reduce(int n, int64_t I_cap, const int32_t *total_dev, const int32_t *ids_sorted, const float *rows,
                               float *gbuf, void *scratch, size_t scratch_bytes, cudaStream_t st) {
    const SegLayout L = seg_layout(n, I_cap);
    if (scratch_bytes < L.total) return GS_ERR_SCRATCH_SMALL;
    if (I_cap <= 0 || n <= 0) return 0;
    unsigned char *p = reinterpret_cast<unsigned char *>(scratch);
    uint32_t *keyA = (uint32_t *)(p + L.keyA), *valA = (uint32_t *)(p + L.valA);
    uint32_t *keyB = (uint32_t *)(p + L.keyB), *valB = (uint32_t *)(p + L.valB);
    uint32_t *mat = (uint32_t *)(p + L.mat);
    int32_t *seg = (int32_t *)(p + L.seg);
    GS_CUDA(cudaMemsetAsync(seg, 0, (size_t)n * 8, st));
    const int bits = n > 1 ? gs_ceil_log2(n) : 1;
    const int npass = (bits + 7) / 8;
    const uint32_t *kin = reinterpret_cast<const uint32_t *>(ids_sorted), *vin = nullptr;      // value = list position
    uint32_t *ko = keyA, *vo = valA;
    for (int q = 0; q < npass; ++q) {
        const int shift = 8 * q, b = bits - shift < 8 ? bits - shift : 8;
        const int rc = gsradix::launch_pass<uint32_t, 8>(I_cap, total_dev, kin, vin, ko, vo, shift, b, mat, st);
        if (rc) return rc;
        kin = ko; vin = vo;
        ko = ko == keyA ? keyB : keyA;
        vo = vo == valA ? valB : valA;
    }
    segment_bounds_kernel<<<(int)((I_cap + 255) / 256), 256, 0, st>>>(I_cap, total_dev, kin, seg);
    GS_LAUNCH_CHECK();
    const size_t Np = (size_t)(n + 3) / 4 * 4;
    segment_sum_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, Np, seg, vin, rows, gbuf);
    GS_LAUNCH_CHECK();
    return 0;
}

// Persistent, cooperative LSD radix-sort building blocks (keys + 32-bit values).
//
// Design for B200: at the sizes of this path (I ~ 3 M intersections, 36 B each) the whole
// sort working set sits in the 126 MB L2, so a pass is bound by L2 latency and launch
// count, not HBM. Instead of a decoupled look-back over hundreds of co-resident tiles,
// one cooperative grid (occupancy x #SMs CTAs) owns a contiguous chunk per CTA:
//     phase H: CTA histograms its chunk            -> mat[cta][digit]
//     grid.sync
//     phase X: one warp per digit scans its column of `mat` over the CTAs (exclusive)
//     grid.sync
//     phase S: CTA reads its row of `mat` + the digit totals -> per-digit write cursors,
//              then walks its chunk in 2048-key sub-tiles: warp match-any
//              ranking (stable), shared-memory staging, coalesced scatter.
//     grid.sync
// Stability: chunk order == CTA order, sub-tiles in order, warp-striped items in order.
#pragma once
#include <cooperative_groups.h>
#include "gs_common.cuh"

namespace gsradix {
namespace cg = cooperative_groups;

constexpr int TPB = 256;
constexpr int WARPS = TPB / 32;
constexpr int IPT = 8;
constexpr int SUB = TPB * IPT;  // keys per sub-tile

// Data written by other CTAs earlier in the same kernel must be read through L2 (.cg).
template <typename T> __device__ __forceinline__ T ld_cg(const T *p) { return __ldcg(p); }
template <> __device__ __forceinline__ uint64_t ld_cg<uint64_t>(const uint64_t *p) {
    return (uint64_t)__ldcg(reinterpret_cast<const unsigned long long *>(p));
}

template <typename KeyT, int BITS>
struct Smem {
    static constexpr int RADIX = 1 << BITS;
    KeyT stage_k[SUB];
    uint32_t stage_v[SUB];
    uint32_t hist[RADIX];           // phase H histogram / phase S per-sub-tile counts
    uint32_t cursor[RADIX];         // running global write cursor per digit (this CTA)
    int32_t dbase[RADIX];           // cursor - local start, per sub-tile
    uint16_t wc[WARPS][RADIX];      // per-warp digit counters
    uint32_t scan_tmp[TPB];
};

// block-wide exclusive scan of one value per thread (TPB threads); returns exclusive prefix,
// *total = block sum. Uses s.scan_tmp.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *tmp, uint32_t *total) {
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
        for (int o = 1; o < WARPS; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += n;
        }
        if (lane < WARPS) tmp[32 + lane] = w;  // inclusive warp totals
    }
    __syncthreads();
    const uint32_t warp_off = warp == 0 ? 0u : tmp[32 + warp - 1];
    if (total) *total = tmp[32 + WARPS - 1];
    const uint32_t r = warp_off + inc - v;
    __syncthreads();
    return r;
}

// One LSD pass over n (key,value) pairs. Must be called by every thread of a cooperative
// grid. kin/kout: keys; vin may be nullptr (value = element index); kout may be nullptr
// (keys dropped, e.g. final partition). `mat` holds (gridDim.x + 1) * RADIX uint32.
// If digit_excl_out != nullptr, CTA 0 writes the exclusive digit prefix (RADIX+1 entries)
// there: for a tile-id partition this is exactly the per-tile [start,end) table.
template <typename KeyT, int BITS>
__device__ void radix_pass(cg::grid_group &grid, Smem<KeyT, BITS> &s, int64_t n, const KeyT *kin,
                           const uint32_t *vin, KeyT *kout, uint32_t *vout, int shift, int bits,
                           uint32_t *mat, uint32_t *digit_excl_out) {
    constexpr int RADIX = 1 << BITS;
    constexpr int DPT = (RADIX + TPB - 1) / TPB;  // digits per thread (contiguous)
    const uint32_t mask = (1u << bits) - 1u;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = gridDim.x, b = blockIdx.x;
    // chunk of this CTA, aligned to SUB so sub-tiles never straddle CTAs
    const int64_t nsub = (n + SUB - 1) / SUB;
    const int64_t sub_per = (nsub + nb - 1) / nb;
    const int64_t c0 = min(n, (int64_t)b * sub_per * SUB), c1 = min(n, (int64_t)(b + 1) * sub_per * SUB);

    // ---------------- phase H ----------------
    for (int d = tid; d < RADIX; d += TPB) s.hist[d] = 0;
    __syncthreads();
    for (int64_t i = c0 + tid; i < c1; i += TPB) {
        const uint32_t d = (uint32_t)(ld_cg(kin + i) >> shift) & mask;
        atomicAdd(&s.hist[d], 1u);
    }
    __syncthreads();
    for (int d = tid; d < RADIX; d += TPB) mat[(size_t)b * RADIX + d] = s.hist[d];
    grid.sync();

    // ---------------- phase X: distributed column scan ----------------
    // One warp per digit (round-robin over the whole grid): exclusive prefix over CTAs of
    // mat[.][d], written back in place; the column total goes to totals[d].
    uint32_t *totals = mat + (size_t)nb * RADIX;
    {
        const int gw = b * WARPS + warp, nw = nb * WARPS;
        for (int d = gw; d < RADIX; d += nw) {
            uint32_t carry = 0;
            for (int cb = 0; cb < nb; cb += 32) {
                const int bb = cb + lane;
                const uint32_t c = bb < nb ? __ldcg(&mat[(size_t)bb * RADIX + d]) : 0u;
                uint32_t inc = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t nn = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += nn;
                }
                if (bb < nb) mat[(size_t)bb * RADIX + d] = carry + inc - c;
                carry += __shfl_sync(0xffffffffu, inc, 31);
            }
            if (lane == 0) totals[d] = carry;
        }
    }
    grid.sync();

    // ---------------- phase S: cursors ----------------
    uint32_t tot[DPT], below[DPT];
    uint32_t tsum = 0;
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int d = tid * DPT + j;
        tot[j] = d < RADIX ? __ldcg(&totals[d]) : 0u;
        below[j] = d < RADIX ? __ldcg(&mat[(size_t)b * RADIX + d]) : 0u;
        tsum += tot[j];
    }
    uint32_t run = block_excl_scan(tsum, s.scan_tmp, nullptr);
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int d = tid * DPT + j;
        if (d < RADIX) {
            s.cursor[d] = run + below[j];
            if (digit_excl_out && b == 0) {
                digit_excl_out[d] = run;
                if (d == RADIX - 1) digit_excl_out[RADIX] = run + tot[j];
            }
        }
        run += tot[j];
    }
    __syncthreads();

    // ---------------- phase S: sub-tiles ----------------
    for (int64_t base = c0; base < c1; base += SUB) {
        const int valid = (int)min((int64_t)SUB, c1 - base);
        for (int i = tid; i < WARPS * RADIX; i += TPB) (&s.wc[0][0])[i] = 0;
        __syncthreads();
        KeyT key[IPT];
        uint16_t rank[IPT];
        const int woff = warp * 32 * IPT;
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const int li = woff + i * 32 + lane;
            key[i] = li < valid ? ld_cg(kin + base + li) : (KeyT)~(KeyT)0;
        }
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const int li = woff + i * 32 + lane;
            // out-of-range items get digit RADIX-1 and sort behind every valid key of the tile
            const uint32_t d = li < valid ? ((uint32_t)(key[i] >> shift) & mask) : (uint32_t)(RADIX - 1);
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            const int leader = __ffs(peers) - 1;
            uint32_t pre = 0;
            if (lane == leader) {
                pre = s.wc[warp][d];
                s.wc[warp][d] = (uint16_t)(pre + __popc(peers));
            }
            pre = __shfl_sync(0xffffffffu, pre, leader);
            rank[i] = (uint16_t)(pre + __popc(peers & gs_lanemask_lt()));
            __syncwarp();
        }
        __syncthreads();
        // per digit: exclusive over warps, sub-tile count
        uint32_t cnt[DPT];
        uint32_t csum = 0;
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            const int d = tid * DPT + j;
            cnt[j] = 0;
            if (d < RADIX) {
                uint32_t r = 0;
#pragma unroll
                for (int w = 0; w < WARPS; ++w) {
                    const uint32_t c = s.wc[w][d];
                    s.wc[w][d] = (uint16_t)r;
                    r += c;
                }
                cnt[j] = r;
            }
            csum += cnt[j];
        }
        uint32_t dstart = block_excl_scan(csum, s.scan_tmp, nullptr);
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            const int d = tid * DPT + j;
            if (d < RADIX) {
                s.hist[d] = dstart;                                    // local start of digit d
                s.dbase[d] = (int32_t)(s.cursor[d] - dstart);          // dst = dbase + local pos
                // padding items (digit RADIX-1) must not advance the cursor
                uint32_t real = cnt[j];
                if (d == RADIX - 1) real -= (uint32_t)(SUB - valid);
                s.cursor[d] += real;
            }
            dstart += cnt[j];
        }
        __syncthreads();
        uint16_t pos[IPT];
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const int li = woff + i * 32 + lane;
            const uint32_t d = li < valid ? ((uint32_t)(key[i] >> shift) & mask) : (uint32_t)(RADIX - 1);
            pos[i] = (uint16_t)(s.hist[d] + s.wc[warp][d] + rank[i]);
            s.stage_k[pos[i]] = key[i];
        }
        __syncthreads();
        int64_t dst[IPT];
#pragma unroll
        for (int k = 0; k < IPT; ++k) {
            const int p = tid + k * TPB;
            dst[k] = -1;
            if (p < valid) {
                const KeyT kk = s.stage_k[p];
                const uint32_t d = (uint32_t)(kk >> shift) & mask;
                dst[k] = (int64_t)(uint32_t)(s.dbase[d] + p);
                if (kout) kout[dst[k]] = kk;
            }
        }
        // values
#pragma unroll
        for (int i = 0; i < IPT; ++i) {
            const int li = woff + i * 32 + lane;
            uint32_t v = 0;
            if (li < valid) v = vin ? ld_cg(vin + base + li) : (uint32_t)(base + li);
            s.stage_v[pos[i]] = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < IPT; ++k)
            if (dst[k] >= 0) vout[dst[k]] = s.stage_v[tid + k * TPB];
        __syncthreads();
    }
    grid.sync();
}

}  // namespace gsradix

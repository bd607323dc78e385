// Stable LSD radix-sort building blocks (keys + 32-bit values), one pass = three small kernels:
//     hist    : CTA b histograms its chunk (3 072 - 4 096 keys, Cfg)     -> mat[digit][b]
//     colscan : one warp per digit scans its row over the CTAs       -> mat (exclusive), totals[digit]
//     scatter : CTA b scans the digit totals itself (digit base), ranks its chunk (ballot matching,
//               stable), stages it in shared memory in digit order and writes coalesced runs at
//               digit_base + mat[digit][b] + local rank
// MODE 1 passes read their keys straight from (depths, radii): key = bits(depth), culled -> 0xFFFFFFFF.
//
// Design for B200: at the sizes of this path (N ~ 1 M Gaussians, I ~ 3 M intersections) the whole
// working set sits in the 126 MB L2, so a pass is bound by latency and launch count, not HBM.
// A first version ran everything as ONE cooperative persistent kernel with grid.sync between
// phases; ncu showed 10 % issue-active (19 grid syncs at 16 warps/SM, profiles/r01). Splitting into
// ordinary kernels lets every phase run at full occupancy with stream order as the barrier, and
// every kernel takes its element count from DEVICE memory (n_dev), so no host read is needed
// between the scan that produces the count and the passes that consume it.
// Stability: chunk order == CTA order, warp-striped items ranked in order.
#pragma once
#include <cstdlib>

#include "gs_common.cuh"

namespace gsradix {

// CTA shape of a pass: TPB threads x IPT keys per thread = SUB keys per CTA. The passes over N ~ 1 M keys are one-wave,
// latency-bound launches and want many warps per SM (512 x 8: 16 warps per CTA); the passes over I >= 6 M pairs are
// throughput-bound and want smaller chunks (384 x 8 / 256 x 12) — measured per size (tools/radix_cfg_sweep.py); the
// choice is made by the caller (binning.cu: radix_cfg) and handed to launch_pass_cfg below.
template <int TPB_, int IPT_>
struct Cfg {
    static constexpr int TPB = TPB_, IPT = IPT_, WARPS = TPB_ / 32, SUB = TPB_ * IPT_;
};
using CfgDefault = Cfg<256, 16>;
using Cfg512x8 = Cfg<512, 8>;
using Cfg384x8 = Cfg<384, 8>;
using Cfg256x12 = Cfg<256, 12>;
constexpr int MIN_SUB = 2048;   // smallest chunk of any configuration: sizes the count matrix

// element count: host value, or (n_dev != nullptr) the device value clamped by the capacity;
// returns -1 when the device value exceeds the capacity (caller retries with a larger buffer).
__device__ __forceinline__ int64_t resolve_n(int64_t n_cap, const int32_t *n_dev) {
    if (!n_dev) return n_cap;
    const int64_t v = *n_dev;
    return v > n_cap ? -1 : v;
}

// key of element `at`: the key array, or (MODE 1) the depth bits of a visible Gaussian / 0xFFFFFFFF for a
// culled one (sorts behind every visible Gaussian), so that no separate key-building kernel is needed
template <typename KeyT, int MODE>
__device__ __forceinline__ KeyT load_key(const KeyT *__restrict__ kin, const int32_t *__restrict__ radii, int64_t at) {
    const KeyT k = kin[at];
    if (MODE == 1) {   // both loads unconditional (a load behind the radius test would serialise two round trips per key)
        const int32_t r = radii[at];
        return r > 0 ? k : (KeyT) ~(KeyT)0;
    }
    return k;
}

// block-wide exclusive scan of one value per thread (TPB threads). tmp: >= 64 uint32.
template <int WARPS>
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
        for (int o = 1; o < 32; o <<= 1) {
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

// ---------------------------------------------------------------------------------------------
template <typename KeyT, int BITS, int MODE, class C>
__global__ void __launch_bounds__(C::TPB)
hist_kernel(int64_t n_cap, const int32_t *__restrict__ n_dev, const KeyT *__restrict__ kin,
            const int32_t *__restrict__ radii, int shift, int bits, uint32_t *__restrict__ mat, int pitch) {
    constexpr int RADIX = 1 << BITS;
    constexpr int TPB = C::TPB, IPT = C::IPT, SUB = C::SUB;
    __shared__ uint32_t hist[RADIX];
    gs_pdl_sync();
    const int64_t n = resolve_n(n_cap, n_dev);
    const int64_t base = (int64_t)blockIdx.x * SUB;
    if (n < 0 || base >= n) return;
    const uint32_t mask = (1u << bits) - 1u;
    for (int d = threadIdx.x; d < RADIX; d += TPB) hist[d] = 0;
    __syncthreads();
    KeyT key[IPT];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const int64_t at = base + i * TPB + threadIdx.x;
        key[i] = at < n ? load_key<KeyT, MODE>(kin, radii, at) : (KeyT)0;
    }
#pragma unroll
    for (int i = 0; i < IPT; ++i)
        if (base + i * TPB + threadIdx.x < n) atomicAdd(&hist[(uint32_t)(key[i] >> shift) & mask], 1u);
    __syncthreads();
    for (int d = threadIdx.x; d < RADIX; d += TPB) mat[(size_t)d * pitch + blockIdx.x] = hist[d];
}

// One warp per digit: exclusive prefix over the CTA rows of mat[.][d], in place; totals[d] = sum.
template <int BITS, int MODE, int SUB>
__global__ void __launch_bounds__(256)
colscan_kernel(int64_t n_cap, const int32_t *__restrict__ n_dev, uint32_t *__restrict__ mat, int pitch,
               uint32_t *__restrict__ totals) {
    constexpr int RADIX = 1 << BITS;
    gs_pdl_sync();
    const int64_t n = resolve_n(n_cap, n_dev);
    if (n < 0) return;
    const int nb = (int)((n + SUB - 1) / SUB);
    const int lane = threadIdx.x & 31;
    const int d = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (d >= RADIX) return;
    uint32_t *row = mat + (size_t)d * pitch;   // mat[digit][cta]: the scan walks contiguous memory
    uint32_t carry = 0;
    for (int cb = 0; cb < nb; cb += 32 * 4) {
        uint32_t c[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {   // 4 independent loads in flight per lane
            const int bb = cb + u * 32 + lane;
            c[u] = bb < nb ? row[bb] : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int bb = cb + u * 32 + lane;
            uint32_t inc = c[u];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t nn = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += nn;
            }
            if (bb < nb) row[bb] = carry + inc - c[u];
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    if (lane == 0) totals[d] = carry;
}

// Segmented column scan for passes with many chunks (5 M elements, 20-160 M pairs: 1 600 - 80 000 chunks). The kernel
// above gives one warp a whole row, i.e. 32 CTAs walk 80 000 entries each — 253 us per pass on an 8-view 1080p batch,
// as much as the histogram. Here a row is cut into segments of COLSEG chunks: (1) every (digit, segment) warp scans
// its segment in place and leaves the segment total, (2) every warp adds the totals of the segments before its own;
// the warps of segment 0 also write the digit totals.
constexpr int COLSEG = 1024;

template <int BITS, int SUB>
__global__ void __launch_bounds__(256)
colscan_seg_kernel(int64_t n_cap, const int32_t *__restrict__ n_dev, uint32_t *__restrict__ mat, int pitch,
                   uint32_t *__restrict__ segtot, int nseg) {
    constexpr int RADIX = 1 << BITS;
    gs_pdl_sync();
    const int64_t n = resolve_n(n_cap, n_dev);
    if (n < 0) return;
    const int nb = (int)((n + SUB - 1) / SUB);
    const int lane = threadIdx.x & 31;
    const int d = blockIdx.x * 8 + (threadIdx.x >> 5), seg = blockIdx.y;
    if (d >= RADIX) return;
    const int lo = seg * COLSEG, hi = min(nb, lo + COLSEG);
    uint32_t *row = mat + (size_t)d * pitch;
    uint32_t carry = 0;
    for (int cb = lo; cb < hi; cb += 32 * 4) {
        uint32_t c[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int bb = cb + u * 32 + lane;
            c[u] = bb < hi ? row[bb] : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int bb = cb + u * 32 + lane;
            uint32_t inc = c[u];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t nn = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += nn;
            }
            if (bb < hi) row[bb] = carry + inc - c[u];
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    if (lane == 0) segtot[(size_t)d * nseg + seg] = carry;      // 0 for a segment behind the last chunk
}

template <int BITS, int SUB>
__global__ void __launch_bounds__(256)
colscan_fix_kernel(int64_t n_cap, const int32_t *__restrict__ n_dev, uint32_t *__restrict__ mat, int pitch,
                   const uint32_t *__restrict__ segtot, int nseg, uint32_t *__restrict__ totals) {
    constexpr int RADIX = 1 << BITS;
    gs_pdl_sync();
    const int64_t n = resolve_n(n_cap, n_dev);
    if (n < 0) return;
    const int nb = (int)((n + SUB - 1) / SUB);
    const int lane = threadIdx.x & 31;
    const int d = blockIdx.x * 8 + (threadIdx.x >> 5), seg = blockIdx.y;
    if (d >= RADIX) return;
    const uint32_t *st = segtot + (size_t)d * nseg;
    uint32_t before = 0, all = 0;
    for (int s = lane; s < nseg; s += 32) {
        const uint32_t v = st[s];
        all += v;
        if (s < seg) before += v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        before += __shfl_xor_sync(0xffffffffu, before, o);
        all += __shfl_xor_sync(0xffffffffu, all, o);
    }
    if (seg == 0 && lane == 0) totals[d] = all;
    if (before == 0) return;
    const int lo = seg * COLSEG, hi = min(nb, lo + COLSEG);
    uint32_t *row = mat + (size_t)d * pitch;
    for (int bb = lo + lane; bb < hi; bb += 32) row[bb] += before;
}

template <typename KeyT, int BITS, class C>
struct ScatterSmem {
    static constexpr int RADIX = 1 << BITS;
    static constexpr int SUB = C::SUB, WARPS = C::WARPS;
    KeyT stage_k[SUB];
    uint32_t stage_v[SUB];
    uint32_t dstart[RADIX];         // local start of each digit inside the CTA's chunk
    int32_t dbase[RADIX];           // global cursor - local start
    uint16_t wc[WARPS][RADIX];      // per-warp digit counters
    uint32_t scan_tmp[64];
};

// vin == nullptr: value = element index. kout == nullptr: keys dropped.
// Every CTA scans the RADIX digit totals itself (<= 8 KB out of L2 + one block scan) instead of waiting for a
// one-CTA digit-base kernel between colscan and scatter: one launch less per pass. When the digit IS the tile id
// (single-pass partition, bins_out != nullptr) the tile ranges fall out of the same scan: CTA 0 writes
// bins[t] = [base, base + total) (0,0 for empty tiles).
template <typename KeyT, int BITS, int MODE, class C>
__global__ void __launch_bounds__(C::TPB)
scatter_kernel(int64_t n_cap, const int32_t *__restrict__ n_dev, const KeyT *__restrict__ kin,
               const int32_t *__restrict__ radii, const uint32_t *__restrict__ vin, KeyT *__restrict__ kout,
               uint32_t *__restrict__ vout, int shift, int bits, const uint32_t *__restrict__ mat, int pitch,
               const uint32_t *__restrict__ totals, int32_t *__restrict__ bins_out, int bins_T) {
    constexpr int RADIX = 1 << BITS;
    constexpr int TPB = C::TPB, IPT = C::IPT, SUB = C::SUB, WARPS = C::WARPS;
    constexpr int DPT = (RADIX + TPB - 1) / TPB;  // digits per thread (contiguous)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ScatterSmem<KeyT, BITS, C> &s = *reinterpret_cast<ScatterSmem<KeyT, BITS, C> *>(smem_raw);
    gs_pdl_sync();
    const int64_t n = resolve_n(n_cap, n_dev);
    const int64_t base = (int64_t)blockIdx.x * SUB;
    if (n < 0 || base >= n) return;
    const int valid = (int)min((int64_t)SUB, n - base);
    const uint32_t mask = (1u << bits) - 1u;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int i = tid; i < WARPS * RADIX; i += TPB) (&s.wc[0][0])[i] = 0;
    // keys (warp-striped: item i of lane l sits at warp*256 + i*32 + l) and values, loaded up front
    KeyT key[IPT];
    uint32_t val[IPT];
    const int woff = warp * 32 * IPT;
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const int li = woff + i * 32 + lane;
        key[i] = li < valid ? load_key<KeyT, MODE>(kin, radii, base + li) : (KeyT) ~(KeyT)0;
        val[i] = li < valid ? (vin ? vin[base + li] : (uint32_t)(base + li)) : 0u;
    }
    // global cursors of this CTA: digit base (exclusive scan of the digit totals) + keys of the same digit in the
    // CTAs below (mat is exclusive)
    uint32_t cursor[DPT];
    {
        uint32_t tot[DPT], sum = 0;
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            const int d = tid * DPT + j;
            tot[j] = d < RADIX ? totals[d] : 0u;
            cursor[j] = d < RADIX ? mat[(size_t)d * pitch + blockIdx.x] : 0u;
            sum += tot[j];
        }
        uint32_t run = block_excl_scan<WARPS>(sum, s.scan_tmp, nullptr);   // ends with a barrier: wc is zeroed after it
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            const int d = tid * DPT + j;
            cursor[j] += run;
            if (bins_out && blockIdx.x == 0 && d < bins_T) {
                bins_out[2 * d] = tot[j] ? (int32_t)run : 0;
                bins_out[2 * d + 1] = tot[j] ? (int32_t)(run + tot[j]) : 0;
            }
            run += tot[j];
        }
    }

    // stable ranking inside each warp: the 8 match.any are independent and issue back to back; only the
    // counter read-modify-write chain is sequential
    uint16_t rank[IPT];
    unsigned peers[IPT];
    uint32_t dig[IPT];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const int li = woff + i * 32 + lane;
        // out-of-range items get digit RADIX-1 and sort behind every valid key of the chunk
        dig[i] = li < valid ? ((uint32_t)(key[i] >> shift) & mask) : (uint32_t)(RADIX - 1);
        // match-any by ballots: `bits` ballot+select steps whatever the digit mix (the hardware match.any
        // serialises over the DISTINCT values in the warp: ~2x slower on uniformly random digits)
        unsigned pm = 0xffffffffu;
        for (int b = 0; b < bits; ++b) {
            const unsigned bal = __ballot_sync(0xffffffffu, (dig[i] >> b) & 1u);
            pm &= ((dig[i] >> b) & 1u) ? bal : ~bal;
        }
        if (bits < BITS) {   // the padding digit RADIX-1 lies outside the mask: keep it apart from the valid digit `mask`
            const unsigned bal = __ballot_sync(0xffffffffu, li < valid);
            pm &= (li < valid) ? bal : ~bal;
        }
        peers[i] = pm;
    }
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const int leader = __ffs(peers[i]) - 1;
        uint32_t pre = 0;
        if (lane == leader) {
            pre = s.wc[warp][dig[i]];
            s.wc[warp][dig[i]] = (uint16_t)(pre + __popc(peers[i]));
        }
        pre = __shfl_sync(0xffffffffu, pre, leader);
        rank[i] = (uint16_t)(pre + __popc(peers[i] & gs_lanemask_lt()));
        __syncwarp();
    }
    __syncthreads();
    // per digit: exclusive over warps, chunk count
    uint32_t cnt[DPT], csum = 0;
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
    uint32_t dstart = block_excl_scan<WARPS>(csum, s.scan_tmp, nullptr);
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int d = tid * DPT + j;
        if (d < RADIX) {
            s.dstart[d] = dstart;
            s.dbase[d] = (int32_t)(cursor[j] - dstart);   // dst = dbase + local position
        }
        dstart += cnt[j];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
        const uint32_t d = dig[i];
        const uint32_t pos = s.dstart[d] + s.wc[warp][d] + rank[i];
        s.stage_k[pos] = key[i];
        s.stage_v[pos] = val[i];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
        const int p = tid + k * TPB;
        if (p < valid) {   // padding items sit behind every valid one
            const KeyT kk = s.stage_k[p];
            const uint32_t d = (uint32_t)(kk >> shift) & mask;
            const uint32_t dst = (uint32_t)(s.dbase[d] + p);
            if (kout) kout[dst] = kk;
            vout[dst] = s.stage_v[p];
        }
    }
}

// host-side launcher of one pass. mat: RADIX * pitch + RADIX uint32 (pitch = ceil(n_cap/SUB); the tail holds
// totals[RADIX]). bins_out != nullptr: the digit is the tile id -> write tile ranges.
// radii != nullptr (uint32 keys only): kin is the float depth array and the keys are formed on the fly.
template <typename KeyT, int BITS, int MODE = 0, class C = CfgDefault>
inline int launch_pass(int64_t n_cap, const int32_t *n_dev, const KeyT *kin, const uint32_t *vin, KeyT *kout,
                       uint32_t *vout, int shift, int bits, uint32_t *mat, cudaStream_t st,
                       int32_t *bins_out = nullptr, int bins_T = 0, const int32_t *radii = nullptr) {
    constexpr int RADIX = 1 << BITS;
    if (n_cap <= 0) return 0;
    if (bits > BITS || (MODE != 0 && !radii)) return GS_ERR_BAD_ARG;
    constexpr int TPB = C::TPB, SUB = C::SUB;
    const int nb = (int)((n_cap + SUB - 1) / SUB);
    uint32_t *totals = mat + (size_t)nb * RADIX;
    const size_t smem = sizeof(ScatterSmem<KeyT, BITS, C>);
    // the attribute is per DEVICE: set it on every launch (a few hundred ns) rather than once per process — a second
    // GPU in the same process (tests, tools, set_device after first use) would otherwise launch without the opt-in
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute((const void *)scatter_kernel<KeyT, BITS, MODE, C>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    GS_PDL_LAUNCH((hist_kernel<KeyT, BITS, MODE, C>), nb, TPB, 0, st, n_cap, n_dev, kin, radii, shift, bits, mat, nb);
    static const int seg_min = [] {          // chunks above which the column scan is segmented (A/B: GS_COLSCAN_SEG_MIN)
        const char *e = getenv("GS_COLSCAN_SEG_MIN");
        return e ? atoi(e) : 4 * COLSEG;
    }();
    // measured on the 5 M / 1080p stress scene (tools/stress_ab.py with GS_COLSCAN_SEG_MIN): segmenting pays for the pair
    // partitions (10 240 chunks: training 354.6 -> 358.1 views/s; 79 531 chunks of an 8-view batch: eval 672 -> 699)
    // and costs on the depth passes of the batch (13 021 chunks: 699 -> 639) although the kernels themselves are
    // faster under ncu (15 vs 57 us) — so only passes whose count comes from the device (the partitions) take it
    if (nb > seg_min && n_dev != nullptr) {
        const int nseg = (nb + COLSEG - 1) / COLSEG;
        uint32_t *segtot = totals + RADIX;                       // behind the totals (mat_entries leaves room)
        GS_PDL_LAUNCH((colscan_seg_kernel<BITS, SUB>), dim3((RADIX + 7) / 8, nseg), 256, 0, st, n_cap, n_dev, mat, nb, segtot, nseg);
        GS_PDL_LAUNCH((colscan_fix_kernel<BITS, SUB>), dim3((RADIX + 7) / 8, nseg), 256, 0, st, n_cap, n_dev, mat, nb,
                      (const uint32_t *)segtot, nseg, totals);
    } else {
        GS_PDL_LAUNCH((colscan_kernel<BITS, MODE, SUB>), (RADIX + 7) / 8, 256, 0, st, n_cap, n_dev, mat, nb, totals);
    }
    GS_PDL_LAUNCH((scatter_kernel<KeyT, BITS, MODE, C>), nb, TPB, smem, st, n_cap, n_dev, kin, radii, vin, kout, vout,
                  shift, bits, (const uint32_t *)mat, nb, (const uint32_t *)totals, bins_out, bins_T);
    return 0;
}

inline size_t mat_entries(int64_t n_cap, int radix) {
    const size_t chunks = (size_t)((n_cap + MIN_SUB - 1) / MIN_SUB);
    return (chunks + 2 + (chunks + COLSEG - 1) / COLSEG + 1) * (size_t)radix;     // counts | totals | segment totals
}

// launch_pass with the CTA shape chosen at run time: 0 = 256 x 16 (round 1), 1 = 512 x 8, 2 = 384 x 8, 3 = 256 x 12
template <typename KeyT, int BITS, int MODE = 0>
inline int launch_pass_cfg(int cfg, int64_t n_cap, const int32_t *n_dev, const KeyT *kin, const uint32_t *vin, KeyT *kout,
                           uint32_t *vout, int shift, int bits, uint32_t *mat, cudaStream_t st, int32_t *bins_out = nullptr,
                           int bins_T = 0, const int32_t *radii = nullptr) {
#define GS_RADIX_GO(CFG) \
    return launch_pass<KeyT, BITS, MODE, CFG>(n_cap, n_dev, kin, vin, kout, vout, shift, bits, mat, st, bins_out, bins_T, radii)
    switch (cfg) {
        case 1: GS_RADIX_GO(Cfg512x8);
        case 2: GS_RADIX_GO(Cfg384x8);
        case 3: GS_RADIX_GO(Cfg256x12);
        default: GS_RADIX_GO(CfgDefault);
    }
#undef GS_RADIX_GO
}

}  // namespace gsradix

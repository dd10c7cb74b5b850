// finalize.cuh - per-cell mean / max / percentiles from the pane accumulators and run lists.
//
// mean follows gpu_util_stats.py:77-80 (sum/len, 0.0 for an empty cell) with an exact integer
// sum and one correctly rounded fp64 divide.  Percentiles (no reference counterpart; rule R5 in
// oracle/window_stats.py: Hyndman-Fan type 7, v[lo]*(1-w) + v[hi]*w, no FMA) are an exact
// selection: for the u8 metrics a count per byte value is a complete one-digit radix select.
// One THREAD owns one window cell and counts into a PRIVATE histogram in shared memory whose
// layout keeps thread t on bank t%32 (conflict free, no atomics).  Three kernels:
//   k_finalize_cells every cell k_finalize_groups (finalize_groups.cuh) has not finished: count/sum/max/mean;
//                    percentiles when the cell has few samples (<= thread_cell_max) and max < 104 (u16 bins,
//                    208 B per thread, 128 threads per CTA, 8 CTAs/SM); other non-empty cells go on the rare list;
//   k_finalize_rare  one launch for the two parts of the rare list: values up to 255 with few samples per
//                    cell (64 threads per CTA, u16 bins), and cells with many samples (a warp per cell, u32 bins).
#pragma once
#include "common.cuh"

#ifndef KGS_MAX_Q
#define KGS_MAX_Q 8
#endif

namespace kgs {

constexpr uint32_t kSmallMax = 65535;  // largest cell the u16-bin histogram can count

struct FinKind {
    uint32_t n_keys;
    uint32_t cell_begin;                 // first global cell index of this kind
    const uint32_t* cnt;                 // pane accumulators
    const unsigned long long* sum;
    const uint32_t* mx;
    const uint32_t* head;
    uint32_t* o_count;                   // window results
    unsigned long long* o_sum;
    uint32_t* o_max;
    double* o_mean;
    double* o_pct;                       // [n_q][n_keys * n_windows]
    // packed exchange tables (optional, nullptr = off): what another shard needs to rebuild the cell
    void* p_sum;                         // u32 or u64 per cell (FinParams::packed_sum64); compact: the low 16 bits
    uint8_t* p_sum_hi;                   // compact only: bits 16..23 of the sum
    void* p_count;                       // u32 per cell; compact: u16, 0xFFFF = see the escape list
    void* p_max;                         // u16 per cell; compact: u8
    void* p_rank;                        // per percentile and cell: the two selected values, u8|u8<<8 (u16|u16<<16 for fb_mib)
};

struct FinParams {
    FinKind kind[3];
    uint32_t n_kinds;                    // kinds actually present, packed at the front
    int twin_pod, twin_gpu;              // kind slots of the pod and gpu tables when both exist and W == 1, else -1
    uint32_t total_cells;
    uint32_t n_windows, n_panes, W;
    uint32_t n_q;
    double q[8];
    const uint8_t* vals;
    const uint4* desc;
    uint32_t* rare_list;                 // pairs (kind slot, cell) left to k_finalize_rare: the wide cells fill it from
                                         // the front, the big cells from the back; rare_cap = number of window cells,
                                         // so it cannot overflow
    uint32_t* ctl;                       // ctl[2] = big cells listed, ctl[3] = wide cells listed
    uint32_t rare_wide_ctas;             // k_finalize_rare: CTAs working on the wide cells (the rest: big cells)
    uint32_t rare_cap;
    uint32_t thread_cell_max;            // cells above this many samples are counted by a whole warp (big part)
    // k_finalize_groups (finalize_groups.cuh)
    uint8_t* done;                       // per global window cell: == done_gen when k_finalize_groups finished the cell
    uint32_t done_gen;
    int g_gpu, g_pod, g_node;            // kind slots of the gpu / pod / node tables, -1 = not part of the job
    uint32_t g_gs_log2;                  // lanes per (node, window) group = 1 << g_gs_log2 >= gpus per node
    uint32_t g_gpn, g_nodes;             // gpus per node; nodes covered by the groups
    uint32_t packed_sum64;               // packed sums are u64 (a cell's sum may not fit 32 bits)
    uint32_t packed_compact;             // compact packed tables: count u16, sum 24 bits, max u8 (8-bit metrics)
    uint32_t* esc_count;                 // compact: device counter of escape-list entries (in the job's control words)
    uint32_t* esc_hdr;                   // compact: [0] = cells on the escape list (written once the finalize kernels are done), list follows at esc
    uint4* esc;                          //   {global cell, count, sum low, sum high} of a cell above 65 534 samples
    uint32_t esc_cap;
    uint32_t slide_chunk;                // windows per thread of k_finalize_slide
    uint32_t u16;                        // 16-bit metric: the value log holds 2 B/sample, cells go to k_finalize_u16
};

__device__ __forceinline__ double nan_f64() { return __longlong_as_double(0x7FF8000000000000ll); }

// rule R5 ranks for a cell of n > 0 values
__device__ __forceinline__ void pct_ranks(double q, uint32_t n, uint32_t& ilo, uint32_t& ihi, double& w) {
    const double nm1 = (double)(n - 1);
    const double rank = __dmul_rn(q, nm1);
    const double lo = floor(rank);
    const double hi = fmin(nm1, __dadd_rn(lo, 1.0));
    w = __dsub_rn(rank, lo);
    ilo = (uint32_t)lo;
    ihi = (uint32_t)hi;
}

__device__ __forceinline__ double pct_lerp(uint32_t vlo, uint32_t vhi, double w) {
    return __dadd_rn(__dmul_rn((double)vlo, __dsub_rn(1.0, w)), __dmul_rn((double)vhi, w));
}

// One percentile of one cell: the interpolated value into the result table and, when the packed
// exchange tables are on, the two selected values (the receiver repeats pct_ranks + pct_lerp on
// them, which gives the same bits).
__device__ __forceinline__ void store_pct(const FinParams& P, const FinKind& K, uint32_t c, uint32_t iq,
                                          uint32_t vlo, uint32_t vhi, double w) {
    const size_t at = (size_t)iq * (K.n_keys * P.n_windows) + c;
    K.o_pct[at] = pct_lerp(vlo, vhi, w);
    if (K.p_rank) {
        if (P.u16) static_cast<uint32_t*>(K.p_rank)[at] = vlo | (vhi << 16);
        else static_cast<uint16_t*>(K.p_rank)[at] = (uint16_t)(vlo | (vhi << 8));
    }
}

// count / sum / max / mean of one cell (and NaN percentiles when it is empty).
__device__ __forceinline__ void store_basics(const FinParams& P, const FinKind& K, uint32_t c, unsigned long long count,
                                             unsigned long long total, uint32_t vmax) {
    K.o_count[c] = (uint32_t)count;
    K.o_sum[c] = total;
    K.o_max[c] = vmax;
    K.o_mean[c] = count ? __ddiv_rn((double)total, (double)count) : 0.0;
    if (K.p_count && P.packed_compact) {
        // a cell of up to 65 534 samples of an 8-bit metric: its sum is below 2^24
        const bool esc = count > 0xFFFEull;
        static_cast<uint16_t*>(K.p_count)[c] = esc ? (uint16_t)0xFFFFu : (uint16_t)count;
        static_cast<uint16_t*>(K.p_sum)[c] = (uint16_t)total;
        K.p_sum_hi[c] = (uint8_t)(total >> 16);
        static_cast<uint8_t*>(K.p_max)[c] = (uint8_t)vmax;
        if (esc) {   // rare: the full numbers go on a list the receiver patches in (overflow is reported by the host)
            const uint32_t at = atomicAdd(P.esc_count, 1u);
            if (at < P.esc_cap) P.esc[at] = make_uint4(K.cell_begin + c, (uint32_t)count, (uint32_t)total, (uint32_t)(total >> 32));
        }
    } else if (K.p_count) {
        static_cast<uint32_t*>(K.p_count)[c] = (uint32_t)count;
        static_cast<uint16_t*>(K.p_max)[c] = (uint16_t)vmax;
        if (P.packed_sum64) static_cast<unsigned long long*>(K.p_sum)[c] = total;
        else static_cast<uint32_t*>(K.p_sum)[c] = (uint32_t)total;
    }
    if (count == 0) {
        const size_t ncell = (size_t)K.n_keys * P.n_windows;
        for (uint32_t iq = 0; iq < P.n_q; ++iq) {
            K.o_pct[iq * ncell + c] = nan_f64();
            if (K.p_rank) {
                if (P.u16) static_cast<uint32_t*>(K.p_rank)[iq * ncell + c] = 0;
                else static_cast<uint16_t*>(K.p_rank)[iq * ncell + c] = 0;
            }
        }
    }
}

// ---- private histogram layouts -----------------------------------------------------------------
// u16 bins, two per 32-bit word; word index = (bin >> 1) * NT + tid, so every access of thread
// tid falls on bank tid % 32 whatever the bin.  All accesses go through explicit shared-space
// PTX with 32-bit addresses; `base` = shared address of the thread's word 0.
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

template <int NT_, int BINS_>
struct Layout16 {
    static constexpr int NT = NT_, BINS = BINS_, WORDS = BINS_ / 2;
    static constexpr int ROW = NT_ * 4;             // bytes between consecutive words of one thread
    static __device__ __forceinline__ uint32_t addr(uint32_t base, uint32_t bin) {
        return base + (bin >> 1) * ROW + ((bin & 1u) << 1);
    }
    // Two increments at once.  The loads are issued together so the two read-modify-write chains
    // overlap; when both bytes hit the same bin the second store carries both increments.
    static __device__ __forceinline__ void inc2(uint32_t base, uint32_t bin_a, uint32_t bin_b, uint32_t da, uint32_t db) {
        const uint32_t aa = addr(base, bin_a), ab = addr(base, bin_b);
        uint32_t va = lds_u16(aa), vb = lds_u16(ab);
        va += da;
        vb = (aa == ab) ? va : vb;
        vb += db;
        sts_u16(aa, va);
        sts_u16(ab, vb);
    }
};
using CellLayout = Layout16<128, 104>;   // values 0..103 (a utilisation is 0..100): 208 B per thread, 26 KB per CTA of 128 threads
using WideLayout = Layout16<64, 256>;    // values 0..255: 512 B per thread, 32 KB per CTA of 64 threads

// delta = 1 counts a value in, delta = 0xFFFF (-1 modulo 2^16, the bins are u16) takes it out again
// (sliding windows drop the pane that left the window).
template <class L>
__device__ __forceinline__ void add4(uint32_t base, uint32_t w, uint32_t delta = 1u) {
    L::inc2(base, w & 0xFFu, (w >> 8) & 0xFFu, delta, delta);
    L::inc2(base, (w >> 16) & 0xFFu, w >> 24, delta, delta);
}

template <class L>
__device__ __forceinline__ void add1(uint32_t base, uint32_t b, uint32_t delta = 1u) {
    const uint32_t a = L::addr(base, b);
    sts_u16(a, lds_u16(a) + delta);
}

// Count (delta = 1) or un-count (delta = 0xFFFF) every run of the list that starts at descriptor
// `at` (1-based, 0 = none) in the lane's histogram.  WARP-SYNCHRONOUS: all 32 lanes call it together
// (idle lanes pass at = 0) and walk their runs in lockstep.  A run is covered by 16-byte ALIGNED
// blocks of the value log; kChunk of them are loaded at once (one round trip to L2 per chunk, not
// per block) and then counted word by word: a word that lies completely inside the run takes the
// branch-free 4-byte body, the at most two words cut by the run's ends take a byte loop that the
// warp only enters when some lane needs it.
template <class L>
__device__ __forceinline__ void add_list_sync(uint32_t base, const FinParams& P, uint32_t at, uint32_t delta = 1u) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    constexpr int kChunk = 5;   // 80 bytes: a 60-sample run at any alignment
    while (__any_sync(kAll, at != 0)) {
        uint4 r = make_uint4(0, 0, 0, 0);
        if (at) r = __ldg(P.desc + (at - 1));
        const uint32_t start = r.x, end = r.x + r.y;
        uint32_t blk = start & ~15u;
        while (__any_sync(kAll, blk < end)) {
            uint4 x[kChunk];
#pragma unroll
            for (int b = 0; b < kChunk; ++b) {
                x[b] = make_uint4(0, 0, 0, 0);
                if (blk + 16 * b < end) x[b] = __ldg(reinterpret_cast<const uint4*>(P.vals + blk + 16 * b));
            }
#pragma unroll
            for (int b = 0; b < kChunk; ++b) {
                const uint32_t a = blk + 16 * b;
                if (!__any_sync(kAll, a < end)) break;
                const int lo = (int)(start - a), hi = (int)(end - a);   // run = bytes [lo, hi) of this block
                const uint32_t w[4] = {x[b].x, x[b].y, x[b].z, x[b].w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const bool whole = lo <= 4 * k && hi >= 4 * k + 4;
                    const bool cut = !whole && lo < 4 * k + 4 && hi > 4 * k;
                    if (whole) add4<L>(base, w[k], delta);
                    if (__any_sync(kAll, cut)) {
                        if (cut) {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (lo <= 4 * k + j && 4 * k + j < hi) add1<L>(base, (w[k] >> (8 * j)) & 0xFFu, delta);
                        }
                    }
                }
            }
            blk += 16 * kChunk;
        }
        at = r.z;
    }
}

// Count the W pane lists of this thread's window cell (lanes without a cell pass valid = false).
template <class L>
__device__ __forceinline__ void add_cell(uint32_t base, const FinParams& P, const FinKind& K, uint32_t pc0,
                                         bool valid) {
    for (uint32_t d = 0; d < P.W; ++d) add_list_sync<L>(base, P, valid ? K.head[pc0 + d] : 0u);
}

// The value of rank `ilo` and of rank `ihi` (ihi = ilo or ilo + 1) among the `count` values of the
// thread's histogram, whose highest occupied bin is vmax.  Ranks in the lower half are searched
// upwards from the first occupied bin, ranks in the upper half downwards from vmax.
template <class L>
__device__ __forceinline__ void select2(uint32_t base, uint32_t count, uint32_t vmax, uint32_t ilo, uint32_t ihi,
                                        uint32_t& vlo, uint32_t& vhi) {
    const uint32_t top = base + (uint32_t)(L::WORDS - 1) * L::ROW;
    uint32_t a, word, le;   // le = number of values <= vlo
    if (ilo < (count >> 1)) {
        a = base;
        while ((word = lds_u32(a)) == 0 && a < top) a += L::ROW;
        uint32_t cum = 0, ws = (word & 0xFFFFu) + (word >> 16);
        while (cum + ws <= ilo && a < top) {
            cum += ws;
            a += L::ROW;
            word = lds_u32(a);
            ws = (word & 0xFFFFu) + (word >> 16);
        }
        const bool in_lo = ilo - cum < (word & 0xFFFFu);
        vlo = 2 * ((a - base) / L::ROW) + (in_lo ? 0u : 1u);
        le = in_lo ? cum + (word & 0xFFFFu) : cum + ws;
    } else {
        const uint32_t u = count - 1 - ilo;   // rank counted from the top
        a = base + (vmax >> 1) * L::ROW;
        word = lds_u32(a);
        uint32_t cum = 0, ws = (word & 0xFFFFu) + (word >> 16);
        while (cum + ws <= u && a > base) {
            cum += ws;
            a -= L::ROW;
            word = lds_u32(a);
            ws = (word & 0xFFFFu) + (word >> 16);
        }
        const bool in_hi = u - cum < (word >> 16);
        vlo = 2 * ((a - base) / L::ROW) + (in_hi ? 1u : 0u);
        le = count - (in_hi ? cum : cum + (word >> 16));
    }
    vhi = vlo;
    if (ihi >= le) {   // the next occupied bin above vlo
        if (!(vlo & 1u) && (word >> 16)) {
            vhi = vlo + 1;
        } else {
            uint32_t word2;
            do {
                a += L::ROW;
                word2 = lds_u32(a);
            } while (!word2 && a < top);
            vhi = 2 * ((a - base) / L::ROW) + ((word2 & 0xFFFFu) ? 0u : 1u);
        }
    }
}

// Rule R5 for every requested percentile of one cell.
template <class L>
__device__ __forceinline__ void select_store(uint32_t base, const FinParams& P, const FinKind& K, uint32_t c,
                                             uint32_t count, uint32_t vmax, uint32_t twin_cell = kInvalid) {
    for (uint32_t iq = 0; iq < P.n_q; ++iq) {
        uint32_t ilo, ihi, vlo, vhi;
        double wt;
        pct_ranks(P.q[iq], count, ilo, ihi, wt);
        select2<L>(base, count, vmax, ilo, ihi, vlo, vhi);
        store_pct(P, K, c, iq, vlo, vhi, wt);
        if (twin_cell != kInvalid) store_pct(P, P.kind[P.twin_gpu], twin_cell, iq, vlo, vhi, wt);
    }
}

__device__ __forceinline__ const FinKind& kind_of_cell(const FinParams& P, uint32_t gc, uint32_t& ks) {
    ks = 0;
    if (P.n_kinds > 1 && gc >= P.kind[1].cell_begin) ks = 1;
    if (P.n_kinds > 2 && gc >= P.kind[2].cell_begin) ks = 2;
    return P.kind[ks];
}

// Cells the private u16 histograms should not take go on the rare list (one atomic per warp and
// part): `big` cells (many samples: u32 bins, a whole warp per cell) are appended from the back,
// `wide` cells (a value >= 104, or the 16-bit metric) from the front.  A cell is listed at most once
// and the list has one entry per window cell, so the two parts cannot meet.  Both parts are worked
// off by ONE launch (k_finalize_rare).
__device__ __forceinline__ void defer_cell(const FinParams& P, bool wide, bool big, uint32_t ks, uint32_t c) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    const uint32_t wide_mask = __ballot_sync(kAll, wide), big_mask = __ballot_sync(kAll, big);
    const uint32_t below = (1u << lane_id()) - 1u;
    if (wide_mask) {
        uint32_t at = 0;
        if (lane_id() == 0) at = atomicAdd(P.ctl + 3, (uint32_t)__popc(wide_mask));
        at = __shfl_sync(kAll, at, 0) + __popc(wide_mask & below);
        if (wide) {
            P.rare_list[2 * (size_t)at] = ks;
            P.rare_list[2 * (size_t)at + 1] = c;
        }
    }
    if (big_mask) {
        uint32_t at = 0;
        if (lane_id() == 0) at = atomicAdd(P.ctl + 2, (uint32_t)__popc(big_mask));
        at = P.rare_cap - 1u - (__shfl_sync(kAll, at, 0) + __popc(big_mask & below));
        if (big) {
            P.rare_list[2 * (size_t)at] = ks;
            P.rare_list[2 * (size_t)at + 1] = c;
        }
    }
}

// Compact exchange tables: the length of the escape list belongs in the header of the exchanged region, and it is
// final only when the last store_basics of the job has run.  Every CTA of the last kernel that stores basics calls
// this once, all its threads together, on its way out; the CTA that arrives last writes the header.
__device__ __forceinline__ void publish_escapes(const FinParams& P) {
    if (!P.esc_hdr || threadIdx.x != 0) return;
    __threadfence();
    if (atomicAdd(P.ctl + 6, 1u) == gridDim.x - 1) {
        P.esc_hdr[0] = *reinterpret_cast<volatile uint32_t*>(P.esc_count);
        P.esc_hdr[1] = 0; P.esc_hdr[2] = 0; P.esc_hdr[3] = 0;
    }
}

// One window cell by a whole warp: u32 bins in shared memory (hist = 256 words of the warp's own), every lane
// reads the same run descriptor and the warp counts the run's bytes together; the ranks are then located with a
// warp scan over the 256 bins (lane l owns bins 8l .. 8l+7).  For cells with many samples, where one thread
// walking the values is the long pole, and for cells with values >= 104, which the 104-bin thread histograms
// cannot take.  count = the cell's sample count (a percentile job holds fewer than 2^32 samples).
__device__ __forceinline__ void warp_cell_select(const FinParams& P, const FinKind& K, uint32_t c, uint32_t pc0,
                                                 uint32_t count, uint32_t* hist) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    const uint32_t lane = lane_id();
    for (uint32_t b = lane; b < 256; b += 32) hist[b] = 0;
    __syncwarp();
    for (uint32_t d = 0; d < P.W; ++d) {
        uint32_t at = K.head[pc0 + d];
        while (at) {
            const uint4 r = __ldg(P.desc + (at - 1));
            for (uint32_t o = lane; o < r.y; o += 32) atomicAdd(hist + P.vals[(uint64_t)r.x + o], 1u);
            at = r.z;
        }
    }
    __syncwarp();
    uint32_t mine[8], incl = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        mine[b] = hist[8 * lane + b];
        incl += mine[b];
    }
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(kAll, incl, d);
        if (lane >= (uint32_t)d) incl += o;
    }
    const uint32_t excl = incl - (mine[0] + mine[1] + mine[2] + mine[3] + mine[4] + mine[5] + mine[6] + mine[7]);
    auto value_of_rank = [&](uint32_t r) {   // every lane gets the answer
        uint32_t v = 0;
        if (excl <= r && r < incl) {
            uint32_t cum = excl;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                if (r >= cum && r - cum < mine[b]) v = 8 * lane + b;
                cum += mine[b];
            }
        }
        const uint32_t owner = __ballot_sync(kAll, excl <= r && r < incl);
        return __shfl_sync(kAll, v, owner ? __ffs(owner) - 1 : 0);
    };
    for (uint32_t iq = 0; iq < P.n_q; ++iq) {
        uint32_t ilo, ihi;
        double w;
        pct_ranks(P.q[iq], count, ilo, ihi, w);
        const uint32_t vlo = value_of_rank(ilo), vhi = value_of_rank(ihi);
        if (lane == 0) store_pct(P, K, c, iq, vlo, vhi, w);
    }
    __syncwarp();
}

// Every window cell that k_finalize_groups has not finished: count / sum / max / mean, and the
// percentiles of cells with up to thread_cell_max samples whose values stay below 104 by a thread each
// (private histogram); cells with more samples or larger values by their whole warp, one after the
// other, in place (warp_cell_select) - no list, no second kernel.  Cells are visited from the END of
// the index space (node cells, the heaviest, are spread first).  Cells of the 16-bit metric go on
// the list k_finalize_u16 works off.
__global__ void __launch_bounds__(CellLayout::NT, 8) k_finalize_cells(const __grid_constant__ FinParams P) {
    using L = CellLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    __shared__ uint32_t hist32[L::NT / 32][256];     // warp-per-cell path (warp_cell_select)
    cudaGridDependencySynchronize();                 // a programmatic dependent of the kernel before it (no-op otherwise)
    cudaTriggerProgrammaticLaunchCompletion();
    const uint32_t tid = threadIdx.x;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(hist) + tid * 4;
    const uint32_t n_blocks = (P.total_cells + L::NT - 1) / L::NT;
    // a CTA walks blocks of 128 cells, gridDim.x apart (node cells, the heaviest, are spread over all CTAs)
    if (P.done) {   // after k_finalize_groups the usual CTA has nothing left to do: one look at its cells' flags
        bool all = true;
        for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
            const uint32_t i = blk * L::NT + tid;
            if (i < P.total_cells) all = all && P.done[P.total_cells - 1 - i] == (uint8_t)P.done_gen;
        }
        if (__syncthreads_and(all)) {
            publish_escapes(P);
            return;
        }
    }
    for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        const uint32_t idx = blk * L::NT + tid;
        bool live = idx < P.total_cells;
        const uint32_t gc = live ? P.total_cells - 1 - idx : 0;
        if (live && P.done && P.done[gc] == (uint8_t)P.done_gen) live = false;   // finished by k_finalize_groups
        uint32_t ks = 0, c = 0, pc0 = 0, vmax = 0;
        unsigned long long count = 0;
        if (live) {
            const FinKind& K = kind_of_cell(P, gc, ks);
            c = gc - K.cell_begin;
            const uint32_t key = c / P.n_windows;
            const uint32_t j = c - key * P.n_windows;
            pc0 = key * P.n_panes + j;
            unsigned long long total = 0;
            for (uint32_t d = 0; d < P.W; ++d) {
                count += K.cnt[pc0 + d];
                total += K.sum[pc0 + d];
                vmax = max(vmax, K.mx[pc0 + d]);
            }
            store_basics(P, K, c, count, total, vmax);
        }
        // a run log that overflowed holds incomplete lists: the host reports the error, skip selection
        if (P.n_q == 0 || P.ctl[1]) continue;
        const FinKind& K = P.kind[ks];
        // a cell with many samples, or with a value the 104-bin thread histogram cannot take, is counted by its
        // whole warp right here (8-bit metrics); cells of the 16-bit metric go on the list k_finalize_u16 works off
        const bool big = live && !P.u16 && count > 0 && (count > P.thread_cell_max || vmax >= (uint32_t)L::BINS);
        const bool fast = live && count > 0 && !big && !P.u16;
        const bool wide = live && count > 0 && P.u16;
        defer_cell(P, wide, false, ks, c);
        {
            const uint32_t big_mask = __ballot_sync(0xFFFFFFFFu, big);
            if (big_mask) {
                if (lane_id() == 0) atomicAdd(P.ctl + 2, (uint32_t)__popc(big_mask));   // job counter
                for (uint32_t m = big_mask; m; m &= m - 1) {
                    const int src = __ffs(m) - 1;
                    const uint32_t ks_b = __shfl_sync(0xFFFFFFFFu, ks, src), c_b = __shfl_sync(0xFFFFFFFFu, c, src);
                    const uint32_t pc_b = __shfl_sync(0xFFFFFFFFu, pc0, src), n_b = __shfl_sync(0xFFFFFFFFu, (uint32_t)count, src);
                    warp_cell_select(P, P.kind[ks_b], c_b, pc_b, n_b, hist32[tid >> 5]);
                }
            }
        }
        // Twin cells.  With one process per GPU a (gpu, window) cell and its (pod, window) cell usually
        // consist of the very same single run, and k_accum wrote that run's descriptors side by side
        // (pod at index i, gpu at i + 1).  The pod cell's thread then selects once and stores the
        // percentiles in both tables; the gpu cell's thread, seeing the mirror-image condition on the
        // same immutable lists, does nothing.  Both sides test "my list is exactly this run" AND "the
        // sibling's list is exactly the sibling descriptor", so they always agree.  (A pair that
        // k_finalize_groups has finished never gets here: it marks both cells or neither.)
        bool work = fast;
        uint32_t twin_cell = kInvalid;
        if (fast && P.twin_pod >= 0) {
            const uint32_t h = K.head[pc0];
            if ((int)ks == P.twin_pod) {
                const uint4 mine = __ldg(P.desc + (h - 1));
                const uint4 sib = __ldg(P.desc + h);              // the gpu descriptor of the same run
                if (mine.z == 0 && sib.z == 0 && sib.w != kInvalid && sib.x == mine.x && sib.y == mine.y &&
                    P.kind[P.twin_gpu].head[sib.w] == h + 1)
                    twin_cell = sib.w;
            } else if ((int)ks == P.twin_gpu && h >= 2) {
                const uint4 mine = __ldg(P.desc + (h - 1));
                const uint4 sib = __ldg(P.desc + (h - 2));        // the pod descriptor of the same run
                if (mine.z == 0 && sib.z == 0 && sib.w != kInvalid && sib.x == mine.x && sib.y == mine.y &&
                    P.kind[P.twin_pod].head[sib.w] == h - 1)
                    work = false;                                  // the pod cell's thread stores my percentiles
            }
        }
        // (this barrier also separates the previous block's selection from the clearing below)
        if (!__syncthreads_or(work)) continue;   // nothing to count in this block (the usual case after k_finalize_groups)
        {   // clear the CTA's histograms together (STS.128)
            uint4* z = reinterpret_cast<uint4*>(hist);
#pragma unroll
            for (int i = 0; i < L::WORDS / 4; ++i) z[i * L::NT + tid] = make_uint4(0, 0, 0, 0);
            __syncthreads();
        }
        add_cell<L>(base, P, K, pc0, work);   // warp-synchronous: every lane takes part
        if (work) select_store<L>(base, P, K, c, (uint32_t)count, vmax, twin_cell);
    }
    __syncthreads();   // every store_basics of this CTA (and its escape-list entries) has been issued
    publish_escapes(P);
}

// The wide list: values up to 255 (u16 bins, 256 values), at most 65 535 samples per cell.
// CTAs [0, P.rare_wide_ctas) of k_finalize_rare.
__device__ __forceinline__ void finalize_wide_part(const FinParams& P, uint32_t cta, uint32_t n_ctas) {
    using L = WideLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    const uint32_t tid = threadIdx.x;
    const uint32_t n_wide = P.ctl[1] ? 0u : min(P.ctl[3], P.rare_cap);
    if (cta * L::NT >= n_wide) return;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(hist) + tid * 4;
    for (uint32_t a = base; a < base + L::WORDS * L::ROW; a += L::ROW) sts_u32(a, 0);
    for (uint32_t it0 = cta * L::NT; it0 < n_wide; it0 += n_ctas * L::NT) {   // CTA-uniform trips
        const uint32_t it = it0 + tid;
        const bool valid = it < n_wide;
        uint32_t ks = 0, c = 0, vmax = 0, pc0 = 0;
        unsigned long long count = 0;
        if (valid) {
            ks = P.rare_list[2 * (size_t)it];
            c = P.rare_list[2 * (size_t)it + 1];
            count = P.kind[ks].o_count[c];
        }
        const FinKind& K = P.kind[ks];
        if (valid) {
            vmax = K.o_max[c];
            const uint32_t key = c / P.n_windows;
            pc0 = key * P.n_panes + (c - key * P.n_windows);
        }
        add_cell<L>(base, P, K, pc0, valid);
        if (valid) select_store<L>(base, P, K, c, (uint32_t)count, vmax);
        // leave the histogram clean for the next cell (idle lanes only ever add 0 to bin 0)
        for (uint32_t a = base; a <= base + (vmax >> 1) * L::ROW; a += L::ROW) sts_u32(a, 0);
    }
}

// Cells with many samples (above thread_cell_max): one warp per cell, u32 bins in shared memory.
// Every lane reads the same descriptor, the whole warp counts each run's bytes; the ranks are then
// located with a warp scan over the 256 bins (lane l owns bins 8l .. 8l+7).  The CTAs behind the wide part.
__device__ __forceinline__ void finalize_big_part(const FinParams& P, uint32_t cta, uint32_t n_ctas) {
    constexpr uint32_t kWarps = WideLayout::NT / 32;
    __shared__ uint32_t hist_all[kWarps][256];
    const uint32_t wib = threadIdx.x >> 5;
    uint32_t* hist = hist_all[wib];
    const uint32_t n_big = P.ctl[1] ? 0u : min(P.ctl[2], P.rare_cap);
    const uint32_t warps = n_ctas * kWarps;
    for (uint32_t it = cta * kWarps + wib; it < n_big; it += warps) {
        const size_t e = (size_t)(P.rare_cap - 1u - it);
        const FinKind& K = P.kind[P.rare_list[2 * e]];
        const uint32_t c = P.rare_list[2 * e + 1];
        const uint32_t key = c / P.n_windows, j = c - key * P.n_windows;
        const uint32_t pc0 = key * P.n_panes + j;
        warp_cell_select(P, K, c, pc0, K.o_count[c], hist);
    }
}

// Both rare lists in one launch (each is almost always empty; two launches cost 6 % of a step).
__global__ void __launch_bounds__(WideLayout::NT) k_finalize_rare(const __grid_constant__ FinParams P) {
    cudaGridDependencySynchronize();                 // a programmatic dependent of k_finalize_cells / k_finalize_slide
    if (P.esc_hdr && blockIdx.x == 0 && threadIdx.x < 4)   // every store_basics has run: publish the escape list's length
        P.esc_hdr[threadIdx.x] = threadIdx.x == 0 ? *P.esc_count : 0u;
    if (blockIdx.x < P.rare_wide_ctas) finalize_wide_part(P, blockIdx.x, P.rare_wide_ctas);
    else finalize_big_part(P, blockIdx.x - P.rare_wide_ctas, gridDim.x - P.rare_wide_ctas);
}

// 16-bit metrics (fb_mib): a two-digit radix select, one warp per cell.  Pass 1 counts the HIGH
// byte of every value of the cell (256 u32 bins in shared memory) and finds, for each wanted rank,
// the high byte that holds it; pass 2 counts the LOW byte of the values with that high byte and
// finishes the selection.  Ranks that share a high byte share the second pass.  Consumes the wide
// list (k_finalize_cells sends every non-empty cell there when P.u16 is set).
__global__ void __launch_bounds__(128) k_finalize_u16(const __grid_constant__ FinParams P) {
    __shared__ uint32_t hist_all[4][256];
    const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
    uint32_t* hist = hist_all[wib];
    const uint16_t* vals = reinterpret_cast<const uint16_t*>(P.vals);
    const uint32_t n_cells = P.ctl[1] ? 0u : min(P.ctl[3], P.rare_cap);
    const uint32_t warps = gridDim.x * 4;
    for (uint32_t it = blockIdx.x * 4 + wib; it < n_cells; it += warps) {
        const FinKind& K = P.kind[P.rare_list[2 * (size_t)it]];
        const uint32_t c = P.rare_list[2 * (size_t)it + 1];
        const uint32_t key = c / P.n_windows, pc0 = key * P.n_panes + (c - key * P.n_windows);
        const uint32_t count = K.o_count[c];
        // digit pass over the cell's runs: `want` < 256 restricts to values with that high byte
        auto digit_pass = [&](uint32_t want) {
            for (uint32_t b = lane; b < 256; b += 32) hist[b] = 0;
            __syncwarp();
            for (uint32_t d = 0; d < P.W; ++d) {
                uint32_t at = K.head[pc0 + d];
                while (at) {   // warp-uniform: every lane reads the same descriptor
                    const uint4 r = __ldg(P.desc + (at - 1));
                    for (uint32_t o = lane; o < r.y; o += 32) {
                        const uint32_t v = vals[(uint64_t)r.x + o];
                        if (want > 255u) atomicAdd(hist + (v >> 8), 1u);
                        else if ((v >> 8) == want) atomicAdd(hist + (v & 0xFFu), 1u);
                    }
                    at = r.z;
                }
            }
            __syncwarp();
        };
        // rank -> (digit, rank inside that digit's bin); every lane computes the same answer
        auto locate = [&](uint32_t rank, uint32_t& digit, uint32_t& inside) {
            uint32_t cum = 0;
            digit = 255; inside = 0;
            for (uint32_t b = 0; b < 256; ++b) {
                const uint32_t cb = hist[b];
                if (rank - cum < cb) { digit = b; inside = rank - cum; break; }
                cum += cb;
            }
        };
        uint32_t hi_digit[2 * KGS_MAX_Q], hi_inside[2 * KGS_MAX_Q];
        double wq[KGS_MAX_Q];
        digit_pass(0xFFFFu);
        for (uint32_t iq = 0; iq < P.n_q; ++iq) {
            uint32_t ilo, ihi;
            pct_ranks(P.q[iq], count, ilo, ihi, wq[iq]);
            locate(ilo, hi_digit[2 * iq], hi_inside[2 * iq]);
            locate(ihi, hi_digit[2 * iq + 1], hi_inside[2 * iq + 1]);
        }
        uint32_t have = 0xFFFFu;   // high byte the second-pass histogram currently holds
        for (uint32_t iq = 0; iq < P.n_q; ++iq) {
            uint32_t val[2];
            for (int e = 0; e < 2; ++e) {
                const uint32_t hd = hi_digit[2 * iq + e];
                if (hd != have) {
                    digit_pass(hd);
                    have = hd;
                }
                uint32_t ld, unused;
                locate(hi_inside[2 * iq + e], ld, unused);
                val[e] = (hd << 8) | ld;
            }
            if (lane == 0) store_pct(P, K, c, iq, val[0], val[1], wq[iq]);
        }
        __syncwarp();
    }
}

}  // namespace kgs

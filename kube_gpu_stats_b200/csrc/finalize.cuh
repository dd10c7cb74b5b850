// finalize.cuh - per-cell mean / max / percentiles from the pane accumulators and run lists.
//
// mean follows gpu_util_stats.py:77-80 (sum/len, 0.0 for an empty cell) with an exact integer
// sum and one correctly rounded fp64 divide.  Percentiles (no reference counterpart; rule R5 in
// oracle/window_stats.py: Hyndman-Fan type 7, v[lo]*(1-w) + v[hi]*w, no FMA) are an exact
// selection: for the u8 metrics a count per byte value is a complete one-digit radix select.
// One THREAD owns one window cell and counts into a PRIVATE histogram in shared memory whose
// layout keeps thread t on bank t%32 (conflict free, no atomics).  Three kernels:
//   k_finalize_cells every cell: count/sum/max/mean; percentiles when the cell has <= 65 535
//                    samples and max < 128 (u16 bins, 256 B per thread, 128 threads per CTA);
//                    other non-empty cells are appended to the wide list;
//   k_finalize_wide  values up to 255, cells up to 65 535 samples (64 threads per CTA);
//   k_finalize_big   anything larger, one warp per cell with u32 bins.
#pragma once
#include "common.cuh"

namespace kgs {

constexpr uint32_t kSmallMax = 65535;  // largest cell the u16-bin histogram can count

struct FinKind {
    uint32_t n_keys;
    uint32_t cell_begin;                 // first global cell index of this kind
    const uint32_t* cnt;                 // pane accumulators
    const unsigned long long* sum;
    const uint32_t* mx;
    const uint32_t* head;
    unsigned long long* o_count;         // window results
    unsigned long long* o_sum;
    uint32_t* o_max;
    double* o_mean;
    double* o_pct;                       // [n_q][n_keys * n_windows]
};

struct FinParams {
    FinKind kind[3];
    uint32_t n_kinds;                    // kinds actually present, packed at the front
    uint32_t total_cells;
    uint32_t n_windows, n_panes, W;
    uint32_t n_q;
    double q[8];
    const uint8_t* vals;
    const uint4* desc;
    uint32_t q_row[8];                   // q[] is ascending; q_row[i] = the caller's index of q[i]
    uint32_t* wide_list;                 // pairs (kind slot, cell) left to k_finalize_wide
    uint32_t* big_list;                  // pairs left to k_finalize_big
    uint32_t* ctl;                       // ctl[2] = big_list length, ctl[3] = wide_list length
    uint32_t wide_cap, big_cap;
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

// ---- private histogram layouts -----------------------------------------------------------------
// u16 bins, two per 32-bit word; word index = (bin >> 1) * NT + tid, so every access of thread
// tid falls on bank tid % 32 whatever the bin.
template <int NT_, int BINS_>
struct Layout16 {
    static constexpr int NT = NT_, BINS = BINS_, WORDS = BINS_ / 2;
    static __device__ __forceinline__ void inc(unsigned char* hb, uint32_t tid4, uint32_t bin) {
        *reinterpret_cast<unsigned short*>(hb + ((bin >> 1) * (NT * 4)) + tid4 + ((bin & 1u) << 1)) += 1;
    }
};
using CellLayout = Layout16<128, 128>;   // values 0..127: 256 B per thread, 32 KB per CTA of 128 threads
using WideLayout = Layout16<64, 256>;    // values 0..255: 512 B per thread, 32 KB per CTA of 64 threads

// Count the bytes of x selected by `mask` (bit k = byte k) into the thread's histogram.
template <class L, bool MASKED>
__device__ __forceinline__ void add16(unsigned char* hb, uint32_t tid4, const uint4& x, uint32_t mask) {
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        if (!MASKED || (mask & (1u << k))) L::inc(hb, tid4, (w[k >> 2] >> (8 * (k & 3))) & 0xFFu);
    }
}

// Count the bytes vals[start, start+len) into the thread's histogram, 16 aligned bytes per load.
template <class L>
__device__ __forceinline__ void add_run(unsigned char* hb, uint32_t tid4, const uint8_t* __restrict__ vals,
                                        uint32_t start, uint32_t len) {
    const uint32_t end = start + len;
    uint32_t blk = start & ~15u;
    uint4 x = __ldg(reinterpret_cast<const uint4*>(vals + blk));
    while (true) {
        const uint32_t lo = max(start, blk) - blk, hi = min(end, blk + 16) - blk;
        const bool more = blk + 16 < end;
        uint4 nx = x;
        if (more) nx = __ldg(reinterpret_cast<const uint4*>(vals + blk + 16));
        if (hi - lo == 16) add16<L, false>(hb, tid4, x, 0xFFFFu);
        else add16<L, true>(hb, tid4, x, ((1u << hi) - 1u) & ~((1u << lo) - 1u));
        if (!more) break;
        x = nx;
        blk += 16;
    }
}

// Walk the W pane lists of one window cell.
template <class L>
__device__ __forceinline__ void add_cell(unsigned char* hb, uint32_t tid4, const FinParams& P, const FinKind& K,
                                         uint32_t pc0) {
    for (uint32_t d = 0; d < P.W; ++d) {
        uint32_t at = K.head[pc0 + d];
        while (at) {
            const uint4 r = __ldg(P.desc + (at - 1));
            add_run<L>(hb, tid4, P.vals, r.x, r.y);
            at = r.z;
        }
    }
}

// Select the percentiles from the thread's histogram and store them.  P.q is ascending (the host
// sorts it; P.q_row maps back to the caller's order), so one forward pass over the words serves
// every rank: `cum` counts the values below word w.
template <class L>
__device__ __forceinline__ void select_store(const uint32_t* hist, uint32_t tid, const FinParams& P,
                                             const FinKind& K, uint32_t c, uint32_t count) {
    const uint32_t ncell = K.n_keys * P.n_windows;
    uint32_t w = 0, cum = 0;
    uint32_t word = hist[tid];
    uint32_t ws = (word & 0xFFFFu) + (word >> 16);
    for (uint32_t iq = 0; iq < P.n_q; ++iq) {
        uint32_t ilo, ihi;
        double wt;
        pct_ranks(P.q[iq], count, ilo, ihi, wt);
        while (cum + ws <= ilo && w < (uint32_t)L::WORDS - 1) {   // count > ilo ends this inside the table
            cum += ws;
            ++w;
            word = hist[w * L::NT + tid];
            ws = (word & 0xFFFFu) + (word >> 16);
        }
        const uint32_t c0 = word & 0xFFFFu;
        const bool in_lo = ilo - cum < c0;                 // cum <= ilo < cum + c0
        const uint32_t vlo = 2 * w + (in_lo ? 0u : 1u);
        const uint32_t bin_end = in_lo ? cum + c0 : cum + ws;   // ranks below bin_end share vlo's value
        uint32_t vhi = vlo;
        if (ihi >= bin_end) {                              // the next occupied bin
            if (in_lo && (word >> 16)) {
                vhi = vlo + 1;
            } else {
                uint32_t w2 = w;
                uint32_t word2;
                do {
                    ++w2;
                    word2 = hist[w2 * L::NT + tid];
                } while (!word2 && w2 < (uint32_t)L::WORDS - 1);
                vhi = 2 * w2 + ((word2 & 0xFFFFu) ? 0u : 1u);
            }
        }
        K.o_pct[(size_t)P.q_row[iq] * ncell + c] = pct_lerp(vlo, vhi, wt);
    }
}

__device__ __forceinline__ const FinKind& kind_of_cell(const FinParams& P, uint32_t gc, uint32_t& ks) {
    ks = 0;
    if (P.n_kinds > 1 && gc >= P.kind[1].cell_begin) ks = 1;
    if (P.n_kinds > 2 && gc >= P.kind[2].cell_begin) ks = 2;
    return P.kind[ks];
}

// Every window cell: count / sum / max / mean, and the percentiles of cells with up to 65 535
// samples whose values stay below 128.  Cells are visited from the END of the index space
// (node cells, the heaviest, are scheduled first).  Other non-empty cells go to the wide list.
__global__ void __launch_bounds__(CellLayout::NT) k_finalize_cells(const __grid_constant__ FinParams P) {
    using L = CellLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    const uint32_t tid = threadIdx.x;
    const uint32_t idx = blockIdx.x * L::NT + tid;
    const bool live = idx < P.total_cells;
    const uint32_t gc = live ? P.total_cells - 1 - idx : 0;
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
        K.o_count[c] = count;
        K.o_sum[c] = total;
        K.o_max[c] = vmax;
        K.o_mean[c] = count ? __ddiv_rn((double)total, (double)count) : 0.0;
    }
    // a run log that overflowed holds incomplete lists: the host reports the error, skip selection
    if (P.n_q == 0 || P.ctl[1]) return;
    const FinKind& K = P.kind[ks];
    const bool fast = live && count > 0 && count <= kSmallMax && vmax < (uint32_t)L::BINS;
    const bool later = live && count > 0 && !fast;
    const uint32_t later_mask = __ballot_sync(0xFFFFFFFFu, later);   // one atomic per warp
    if (later_mask) {
        uint32_t base = 0;
        if (lane_id() == 0) base = atomicAdd(P.ctl + 3, (uint32_t)__popc(later_mask));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (later) {
            const uint32_t at = base + __popc(later_mask & ((1u << lane_id()) - 1u));
            if (at < P.wide_cap) {
                P.wide_list[2 * at] = ks;
                P.wide_list[2 * at + 1] = c;
            }
        }
    }
    if (live && count == 0) {
        const uint32_t ncell = K.n_keys * P.n_windows;
        for (uint32_t iq = 0; iq < P.n_q; ++iq) K.o_pct[(size_t)iq * ncell + c] = nan_f64();
    }
    if (!fast) return;
    const uint32_t nw = (vmax >> 1) + 1;
    for (uint32_t w = 0; w < nw; ++w) hist[w * L::NT + tid] = 0;
    add_cell<L>(reinterpret_cast<unsigned char*>(hist), tid * 4, P, K, pc0);
    select_store<L>(hist, tid, P, K, c, (uint32_t)count);
}

// The wide list: values up to 255 (u16 bins, 256 values); cells above 65 535 samples move on to
// the big list.
__global__ void __launch_bounds__(WideLayout::NT) k_finalize_wide(const __grid_constant__ FinParams P) {
    using L = WideLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    const uint32_t tid = threadIdx.x;
    const uint32_t n_wide = P.ctl[1] ? 0u : min(P.ctl[3], P.wide_cap);
    if (blockIdx.x * L::NT >= n_wide) return;
    for (uint32_t w = 0; w < L::WORDS; ++w) hist[w * L::NT + tid] = 0;
    for (uint32_t it = blockIdx.x * L::NT + tid; it < n_wide; it += gridDim.x * L::NT) {
        const uint32_t ks = P.wide_list[2 * it];
        const uint32_t c = P.wide_list[2 * it + 1];
        const FinKind& K = P.kind[ks];
        const unsigned long long count = K.o_count[c];
        if (count > kSmallMax) {
            const uint32_t at = atomicAdd(P.ctl + 2, 1u);
            if (at < P.big_cap) {
                P.big_list[2 * at] = ks;
                P.big_list[2 * at + 1] = c;
            }
            continue;
        }
        const uint32_t vmax = K.o_max[c];
        const uint32_t key = c / P.n_windows;
        const uint32_t pc0 = key * P.n_panes + (c - key * P.n_windows);
        add_cell<L>(reinterpret_cast<unsigned char*>(hist), tid * 4, P, K, pc0);
        select_store<L>(hist, tid, P, K, c, (uint32_t)count);
        const uint32_t nw = (vmax >> 1) + 1;   // leave the histogram clean for the next cell
        for (uint32_t w = 0; w < nw; ++w) hist[w * L::NT + tid] = 0;
    }
}

// Cells above kSmallMax samples: one warp per cell, u32 bins in shared memory.  Every lane reads
// the same descriptor, the whole warp counts each run's bytes.
__global__ void __launch_bounds__(128) k_finalize_big(const __grid_constant__ FinParams P) {
    __shared__ uint32_t hist_all[4][256];
    const uint32_t lane = lane_id(), wib = threadIdx.x >> 5;
    uint32_t* hist = hist_all[wib];
    const uint32_t n_big = P.ctl[1] ? 0u : min(P.ctl[2], P.big_cap);
    const uint32_t warps = gridDim.x * 4;
    for (uint32_t it = blockIdx.x * 4 + wib; it < n_big; it += warps) {
        const FinKind& K = P.kind[P.big_list[2 * it]];
        const uint32_t c = P.big_list[2 * it + 1];
        const uint32_t key = c / P.n_windows, j = c - key * P.n_windows;
        const uint32_t pc0 = key * P.n_panes + j;
        const uint32_t ncell = K.n_keys * P.n_windows;
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
        const unsigned long long count = K.o_count[c];
        if (lane == 0) {
            for (uint32_t iq = 0; iq < P.n_q; ++iq) {
                // a percentile job holds fewer than 2^32 samples, so u32 ranks suffice
                uint32_t ilo, ihi;
                double w;
                pct_ranks(P.q[iq], (uint32_t)count, ilo, ihi, w);
                uint32_t cum = 0, vlo = 0, vhi = 0;
                for (uint32_t b = 0; b < 256; ++b) {
                    const uint32_t cb = hist[b];
                    if (ilo - cum < cb) vlo = b;
                    if (ihi - cum < cb) vhi = b;
                    cum += cb;
                }
                K.o_pct[(size_t)P.q_row[iq] * ncell + c] = pct_lerp(vlo, vhi, w);
            }
        }
        __syncwarp();
    }
}

}  // namespace kgs

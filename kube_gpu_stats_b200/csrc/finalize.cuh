// finalize.cuh - per-cell mean / max / percentiles from the pane accumulators and run lists.
//
// mean follows gpu_util_stats.py:77-80 (sum/len, 0.0 for an empty cell) with an exact integer
// sum and one correctly rounded fp64 divide.  Percentiles (no reference counterpart; rule R5 in
// oracle/window_stats.py: Hyndman-Fan type 7, v[lo]*(1-w) + v[hi]*w, no FMA) are an exact
// selection: for the u8 metrics a count per byte value is a complete one-digit radix select.
// One THREAD owns one window cell and counts into a PRIVATE histogram in shared memory whose
// layout keeps thread t on bank t%32 (conflict free, no atomics).  Three size classes:
//   k_finalize_tiny  every cell: count/sum/max/mean; percentiles when the cell has <= 255
//                    samples and max < 128 (u8 bins, 128 B per thread, 128 threads per CTA);
//                    other non-empty cells are appended to the mid list;
//   k_finalize_mid   cells up to 65 535 samples (u16 bins, 256 values, 64 threads per CTA);
//   k_finalize_big   anything larger, one warp per cell with u32 bins.
#pragma once
#include "common.cuh"

namespace kgs {

constexpr uint32_t kTinyMax = 255;     // largest cell the u8-bin histogram can count
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
    uint32_t* mid_list;                  // pairs (kind slot, cell) left to k_finalize_mid
    uint32_t* big_list;                  // pairs left to k_finalize_big
    uint32_t* ctl;                       // ctl[2] = big_list length, ctl[3] = mid_list length
    uint32_t mid_cap, big_cap;
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
// Bytes of one 32-bit word hold consecutive bins; word index = (bin / bins_per_word) * NT + tid,
// so every access of thread tid falls on bank tid % 32.
struct TinyLayout {   // u8 bins, values 0..127
    static constexpr int NT = 128, WORDS = 32, PER_WORD = 4, BITS = 8;
    static __device__ __forceinline__ void inc(unsigned char* hb, uint32_t tid4, uint32_t bin) {
        hb[((bin & 0xFCu) << 7) + tid4 + (bin & 3u)] += 1;
    }
};
struct MidLayout {    // u16 bins, values 0..255
    static constexpr int NT = 64, WORDS = 128, PER_WORD = 2, BITS = 16;
    static __device__ __forceinline__ void inc(unsigned char* hb, uint32_t tid4, uint32_t bin) {
        *reinterpret_cast<unsigned short*>(hb + ((bin & 0xFEu) << 7) + tid4 + ((bin & 1u) << 1)) += 1;
    }
};

template <class L, bool MASKED>
__device__ __forceinline__ void add16(unsigned char* hb, uint32_t tid4, const uint4& x, uint32_t lo, uint32_t hi) {
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        if (!MASKED || ((uint32_t)k >= lo && (uint32_t)k < hi)) L::inc(hb, tid4, (w[k >> 2] >> (8 * (k & 3))) & 0xFFu);
    }
}

// Count the bytes vals[start, start+len) into the thread's histogram, 16 aligned bytes per load.
template <class L>
__device__ __forceinline__ void add_run(unsigned char* hb, uint32_t tid4, const uint8_t* __restrict__ vals,
                                        uint32_t start, uint32_t len) {
    const uint32_t end = start + len;
    uint32_t blk = start & ~15u;
    uint4 x = __ldg(reinterpret_cast<const uint4*>(vals + blk));
    {
        const uint32_t lo = start - blk, hi = min(16u, end - blk);
        uint4 nx = x;
        if (blk + 16 < end) nx = __ldg(reinterpret_cast<const uint4*>(vals + blk + 16));
        if (lo == 0 && hi == 16) add16<L, false>(hb, tid4, x, 0, 16);
        else add16<L, true>(hb, tid4, x, lo, hi);
        x = nx;
        blk += 16;
    }
    while (blk + 16 <= end) {
        uint4 nx = x;
        if (blk + 16 < end) nx = __ldg(reinterpret_cast<const uint4*>(vals + blk + 16));
        add16<L, false>(hb, tid4, x, 0, 16);
        x = nx;
        blk += 16;
    }
    if (blk < end) add16<L, true>(hb, tid4, x, 0, end - blk);
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

// Select the n_q percentiles from the thread's histogram (bins 0..vmax) and store them.
template <class L>
__device__ __forceinline__ void select_store(const uint32_t* hist, uint32_t tid, const FinParams& P,
                                             const FinKind& K, uint32_t c, uint32_t count, uint32_t vmax) {
    uint32_t ilo[KGS_MAX_Q_], ihi[KGS_MAX_Q_], vlo[KGS_MAX_Q_], vhi[KGS_MAX_Q_];
    double wq[KGS_MAX_Q_];
#pragma unroll
    for (int iq = 0; iq < KGS_MAX_Q_; ++iq) {
        ilo[iq] = ihi[iq] = vlo[iq] = vhi[iq] = 0;
        wq[iq] = 0.0;
        if (iq < (int)P.n_q) pct_ranks(P.q[iq], count, ilo[iq], ihi[iq], wq[iq]);
    }
    uint32_t cum = 0;
    const uint32_t nw = vmax / L::PER_WORD + 1;
    for (uint32_t w = 0; w < nw; ++w) {
        const uint32_t word = hist[w * L::NT + tid];
        if (!word) continue;
#pragma unroll
        for (int k = 0; k < L::PER_WORD; ++k) {
            const uint32_t cb = (word >> (L::BITS * k)) & ((1u << L::BITS) - 1u);
            if (!cb) continue;
            const uint32_t bin = L::PER_WORD * w + k;
#pragma unroll
            for (int iq = 0; iq < KGS_MAX_Q_; ++iq) {
                if (iq < (int)P.n_q) {       // unsigned wrap: true iff cum <= rank < cum + cb
                    if (ilo[iq] - cum < cb) vlo[iq] = bin;
                    if (ihi[iq] - cum < cb) vhi[iq] = bin;
                }
            }
            cum += cb;
        }
    }
    const uint32_t ncell = K.n_keys * P.n_windows;
#pragma unroll
    for (int iq = 0; iq < KGS_MAX_Q_; ++iq)
        if (iq < (int)P.n_q) K.o_pct[(size_t)iq * ncell + c] = pct_lerp(vlo[iq], vhi[iq], wq[iq]);
}

__device__ __forceinline__ const FinKind& kind_of_cell(const FinParams& P, uint32_t gc, uint32_t& ks) {
    ks = 0;
    if (P.n_kinds > 1 && gc >= P.kind[1].cell_begin) ks = 1;
    if (P.n_kinds > 2 && gc >= P.kind[2].cell_begin) ks = 2;
    return P.kind[ks];
}

__global__ void __launch_bounds__(TinyLayout::NT) k_finalize_tiny(const __grid_constant__ FinParams P) {
    using L = TinyLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    const uint32_t tid = threadIdx.x;
    const uint32_t gc = blockIdx.x * L::NT + tid;
    const bool live = gc < P.total_cells;
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
    if (P.n_q == 0) return;
    const FinKind& K = P.kind[ks];
    const bool tiny = live && count > 0 && count <= kTinyMax && vmax < 128;
    const bool later = live && count > 0 && !tiny;
    // cells this kernel cannot count go to the mid list: one atomic per warp
    const uint32_t later_mask = __ballot_sync(0xFFFFFFFFu, later);
    if (later_mask) {
        uint32_t base = 0;
        if (lane_id() == 0) base = atomicAdd(P.ctl + 3, (uint32_t)__popc(later_mask));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (later) {
            const uint32_t at = base + __popc(later_mask & ((1u << lane_id()) - 1u));
            if (at < P.mid_cap) {
                P.mid_list[2 * at] = ks;
                P.mid_list[2 * at + 1] = c;
            }
        }
    }
    if (live && count == 0) {
        const uint32_t ncell = K.n_keys * P.n_windows;
        for (uint32_t iq = 0; iq < P.n_q; ++iq) K.o_pct[(size_t)iq * ncell + c] = nan_f64();
    }
    if (!tiny) return;
    const uint32_t nw = (vmax >> 2) + 1;
    for (uint32_t w = 0; w < nw; ++w) hist[w * L::NT + tid] = 0;
    add_cell<L>(reinterpret_cast<unsigned char*>(hist), tid * 4, P, K, pc0);
    select_store<L>(hist, tid, P, K, c, (uint32_t)count, vmax);
}

__global__ void __launch_bounds__(MidLayout::NT) k_finalize_mid(const __grid_constant__ FinParams P) {
    using L = MidLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    const uint32_t tid = threadIdx.x;
    const uint32_t n_mid = min(P.ctl[3], P.mid_cap);
    for (uint32_t w = 0; w < L::WORDS; ++w) hist[w * L::NT + tid] = 0;
    for (uint32_t it = blockIdx.x * L::NT + tid; it < n_mid; it += gridDim.x * L::NT) {
        const uint32_t ks = P.mid_list[2 * it];
        const uint32_t c = P.mid_list[2 * it + 1];
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
        select_store<L>(hist, tid, P, K, c, (uint32_t)count, vmax);
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
    const uint32_t n_big = min(P.ctl[2], P.big_cap);
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
                K.o_pct[(size_t)iq * ncell + c] = pct_lerp(vlo, vhi, w);
            }
        }
        __syncwarp();
    }
}

}  // namespace kgs

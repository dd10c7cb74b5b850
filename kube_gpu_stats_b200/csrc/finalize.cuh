// finalize.cuh - per-cell mean / max / percentiles from the pane accumulators and run lists.
//
// mean follows gpu_util_stats.py:77-80 (sum/len, 0.0 for an empty cell) with an exact integer
// sum and one correctly rounded fp64 divide.  Percentiles (no reference counterpart; rule R5 in
// oracle/window_stats.py: Hyndman-Fan type 7, v[lo]*(1-w) + v[hi]*w, no FMA) are an exact
// selection: for the u8 metrics a 256-bin count of the cell's values is a complete one-digit
// radix select.  One thread owns one window cell and counts into a PRIVATE histogram in
// shared memory, laid out so that thread t only ever touches bank t%32 (conflict free, no
// atomics).  Cells with more than 65 535 samples go to a warp-per-cell kernel with u32 bins.
#pragma once
#include "common.cuh"

namespace kgs {

constexpr int kFinThreads = 64;       // threads per CTA of k_finalize; 256 u16 bins each = 32 KB
constexpr uint32_t kSmallMax = 65535; // largest cell a u16-bin private histogram can count

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
    uint32_t* big_list;                  // pairs (kind slot, cell) of cells left to k_finalize_big
    uint32_t* ctl;                       // ctl[2] = big_list length
    uint32_t big_cap;
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

__global__ void __launch_bounds__(kFinThreads) k_finalize(const __grid_constant__ FinParams P) {
    __shared__ uint32_t hist[128 * kFinThreads];   // word (bin>>1)*kFinThreads + tid, u16 halves
    const uint32_t tid = threadIdx.x;
    const uint32_t gc = blockIdx.x * kFinThreads + tid;
    if (gc >= P.total_cells) return;
    uint32_t ks = 0;
    if (P.n_kinds > 1 && gc >= P.kind[1].cell_begin) ks = 1;
    if (P.n_kinds > 2 && gc >= P.kind[2].cell_begin) ks = 2;
    const FinKind& K = P.kind[ks];
    const uint32_t c = gc - K.cell_begin;
    const uint32_t key = c / P.n_windows;
    const uint32_t j = c - key * P.n_windows;
    const uint32_t pc0 = key * P.n_panes + j;
    unsigned long long count = 0, total = 0;
    uint32_t vmax = 0;
    for (uint32_t d = 0; d < P.W; ++d) {
        count += K.cnt[pc0 + d];
        total += K.sum[pc0 + d];
        vmax = max(vmax, K.mx[pc0 + d]);
    }
    K.o_count[c] = count;
    K.o_sum[c] = total;
    K.o_max[c] = vmax;
    K.o_mean[c] = count ? __ddiv_rn((double)total, (double)count) : 0.0;
    if (P.n_q == 0) return;
    const uint32_t ncell = K.n_keys * P.n_windows;
    if (count == 0) {
        for (uint32_t iq = 0; iq < P.n_q; ++iq) K.o_pct[(size_t)iq * ncell + c] = nan_f64();
        return;
    }
    if (count > kSmallMax) {
        const uint32_t at = atomicAdd(P.ctl + 2, 1u);
        if (at < P.big_cap) {
            P.big_list[2 * at] = ks;
            P.big_list[2 * at + 1] = c;
        }
        return;
    }
    // private histogram over bins 0..vmax
    const uint32_t nw = (vmax >> 1) + 1;
    for (uint32_t w = 0; w < nw; ++w) hist[w * kFinThreads + tid] = 0;
    unsigned short* h16 = reinterpret_cast<unsigned short*>(hist);
    for (uint32_t d = 0; d < P.W; ++d) {
        uint32_t at = K.head[pc0 + d];
        while (at) {
            const uint4 r = __ldg(P.desc + (at - 1));
            const uint64_t a = r.x, b = a + r.y;
            for (uint64_t blk = a & ~15ull; blk < b; blk += 16) {   // 16-byte aligned slices of the run
                const uint4 x = __ldg(reinterpret_cast<const uint4*>(P.vals + blk));
                const uint32_t wds[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const uint64_t i = blk + k;
                    if (i >= a && i < b) {
                        const uint32_t bin = (wds[k >> 2] >> (8 * (k & 3))) & 0xFFu;
                        h16[(((bin >> 1) * kFinThreads + tid) << 1) + (bin & 1)] += 1;
                    }
                }
            }
            at = r.z;
        }
    }
    uint32_t ilo[KGS_MAX_Q_], ihi[KGS_MAX_Q_], vlo[KGS_MAX_Q_], vhi[KGS_MAX_Q_];
    double wq[KGS_MAX_Q_];
#pragma unroll
    for (int iq = 0; iq < KGS_MAX_Q_; ++iq) {
        ilo[iq] = ihi[iq] = vlo[iq] = vhi[iq] = 0;
        wq[iq] = 0.0;
        if (iq < (int)P.n_q) pct_ranks(P.q[iq], (uint32_t)count, ilo[iq], ihi[iq], wq[iq]);
    }
    uint32_t cum = 0;
    for (uint32_t w = 0; w < nw; ++w) {
        const uint32_t word = hist[w * kFinThreads + tid];
        if (!word) continue;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const uint32_t cb = half ? (word >> 16) : (word & 0xFFFFu);
            if (!cb) continue;
            const uint32_t bin = 2 * w + half, nc = cum + cb;
#pragma unroll
            for (int iq = 0; iq < KGS_MAX_Q_; ++iq) {
                if (iq < (int)P.n_q) {
                    if (ilo[iq] >= cum && ilo[iq] < nc) vlo[iq] = bin;
                    if (ihi[iq] >= cum && ihi[iq] < nc) vhi[iq] = bin;
                }
            }
            cum = nc;
        }
    }
#pragma unroll
    for (int iq = 0; iq < KGS_MAX_Q_; ++iq)
        if (iq < (int)P.n_q) K.o_pct[(size_t)iq * ncell + c] = pct_lerp(vlo[iq], vhi[iq], wq[iq]);
}

// Cells above kSmallMax samples: one warp per cell, u32 bins in shared memory.  Lane 0 walks the
// run list, the whole warp counts each run's bytes.
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
            while (at) {   // warp-uniform: every lane reads the same descriptor
                const uint4 r = __ldg(P.desc + (at - 1));
                for (uint32_t o = lane; o < r.y; o += 32) atomicAdd(hist + P.vals[(uint64_t)r.x + o], 1u);
                at = r.z;
            }
        }
        __syncwarp();
        const unsigned long long count = K.o_count[c];
        if (lane == 0) {
            for (uint32_t iq = 0; iq < P.n_q; ++iq) {
                // counts can exceed 2^32 only beyond the 4 Gi-sample log limit, so u32 ranks suffice
                uint32_t ilo, ihi;
                double w;
                pct_ranks(P.q[iq], (uint32_t)count, ilo, ihi, w);
                uint32_t cum = 0, vlo = 0, vhi = 0;
                for (uint32_t b = 0; b < 256; ++b) {
                    const uint32_t nc = cum + hist[b];
                    if (ilo >= cum && ilo < nc) vlo = b;
                    if (ihi >= cum && ihi < nc) vhi = b;
                    cum = nc;
                }
                K.o_pct[(size_t)iq * ncell + c] = pct_lerp(vlo, vhi, w);
            }
        }
        __syncwarp();
    }
}

}  // namespace kgs

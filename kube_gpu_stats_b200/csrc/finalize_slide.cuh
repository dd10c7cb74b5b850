// finalize_slide.cuh - percentiles of SLIDING windows (width = W panes, W > 1) without re-counting.
//
// k_finalize_cells counts all W panes of every window cell: with BASELINE configs[4] (5-min windows
// every 10 s, W = 30) each sample is counted 30 times per key kind.  Here one thread owns a CHUNK of
// consecutive windows of one key and keeps the window's histogram alive while it slides: entering
// window j+1 it takes pane j out (delta 0xFFFF = -1 in the u16 bins) and counts pane j+W in, then
// selects.  Only the first window of a chunk counts W panes.  Same exact selection (rule R5).
//   * count / sum / max / mean of every window cell come from k_finalize_basic (no shared memory);
//   * a chunk that meets a value >= 104 or a window above 65 535 samples hands its windows to the
//     wide or big list (k_finalize_rare counts them the slow way).
#pragma once
#include "finalize.cuh"

namespace kgs {

// count / sum / max / mean of every window cell, NaN percentiles for empty cells.
__global__ void __launch_bounds__(256) k_finalize_basic(const __grid_constant__ FinParams P) {
    const uint32_t gc = blockIdx.x * 256 + threadIdx.x;
    if (gc >= P.total_cells) return;
    uint32_t ks;
    const FinKind& K = kind_of_cell(P, gc, ks);
    const uint32_t c = gc - K.cell_begin;
    const uint32_t key = c / P.n_windows;
    const uint32_t pc0 = key * P.n_panes + (c - key * P.n_windows);
    unsigned long long count = 0, total = 0;
    uint32_t vmax = 0;
    for (uint32_t d = 0; d < P.W; ++d) {
        count += K.cnt[pc0 + d];
        total += K.sum[pc0 + d];
        vmax = max(vmax, K.mx[pc0 + d]);
    }
    store_basics(P, K, c, count, total, vmax);
}

// Four consecutive u32 of an array held in registers; reloaded (one 16-byte load) when the index
// leaves them.  A lane's panes lie 4*chunk bytes from its neighbour's, so a scalar load per step costs
// a 32-byte sector from L2 for 4 useful bytes (the L1 left beside the histograms cannot keep them):
// this turns four such loads into one.  Arrays are 256-byte aligned and padded inside the arena.
struct Quad {
    uint4 v;
    uint32_t base;
    __device__ __forceinline__ uint32_t at(const uint32_t* a, uint32_t e) {
        const uint32_t b = e & ~3u;
        if (b != base) {
            v = __ldg(reinterpret_cast<const uint4*>(a + b));
            base = b;
        }
        const uint32_t k = e & 3u;
        return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
    }
};

// One thread per (key kind, key, chunk of P.slide_chunk windows); a warp's lanes are consecutive
// chunks of (mostly) the same key, so they step through their windows in lockstep.  Count and max of
// every window come from k_finalize_basic (run before this kernel); of the panes only the list heads
// are read here.
__global__ void __launch_bounds__(CellLayout::NT) k_finalize_slide(const __grid_constant__ FinParams P) {
    using L = CellLayout;
    __shared__ uint32_t hist[L::WORDS * L::NT];
    if (P.ctl[1]) return;   // overflowed run log: the host reports the error
    const uint32_t tid = threadIdx.x;
    {   // clear the CTA's histograms together
        uint4* z = reinterpret_cast<uint4*>(hist);
#pragma unroll
        for (int i = 0; i < L::WORDS / 4; ++i) z[i * L::NT + tid] = make_uint4(0, 0, 0, 0);
        __syncthreads();
    }
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(hist) + tid * 4;
    const uint32_t C = P.slide_chunk;
    const uint32_t chunks_per_key = (P.n_windows + C - 1) / C;
    // which (kind, key, chunk): kinds are laid out one after the other
    uint32_t idx = blockIdx.x * L::NT + tid, ks = 0;
    bool live = false;
    for (uint32_t k = 0; k < P.n_kinds; ++k) {
        const uint32_t span = P.kind[k].n_keys * chunks_per_key;
        if (idx < span) { ks = k; live = true; break; }
        idx -= span;
    }
    const FinKind& K = P.kind[ks];
    const uint32_t key = live ? idx / chunks_per_key : 0;
    const uint32_t j0 = live ? (idx - key * chunks_per_key) * C : 0;
    const uint32_t j1 = live ? min(P.n_windows, j0 + C) : 0;
    const uint32_t prow = key * P.n_panes;          // first pane cell of the key
    const uint32_t crow = key * P.n_windows;        // first window cell of the key
    Quad q_in{make_uint4(0, 0, 0, 0), 0xFFFFFFFFu}, q_out = q_in, q_cnt = q_in, q_max = q_in;
    // A window whose largest value does not fit the 104-bin histogram goes to the wide list, and so do
    // the windows after it in this chunk (the histogram stops being maintained).
    bool bad = false;
    {   // first window of the chunk: count its W panes
        const bool on = live && j0 < j1;
        if (on) bad = q_max.at(K.o_max, crow + j0) >= (uint32_t)L::BINS;
        for (uint32_t d = 0; d < P.W; ++d) {
            uint32_t at = 0;
            if (on && !bad) at = q_in.at(K.head, prow + j0 + d);
            add_list_sync<L>(base, P, at);
        }
    }
    for (uint32_t s = 0; s < C; ++s) {              // warp-uniform trip count
        const uint32_t j = j0 + s;
        const bool on = live && j < j1;
        const uint32_t c = crow + j;
        uint32_t count = 0, vmax = 0;
        if (on) {
            count = q_cnt.at(K.o_count, c);
            vmax = q_max.at(K.o_max, c);
        }
        if (s > 0) {                                // slide: pane j-1 leaves, pane j+W-1 enters
            uint32_t out = 0, in = 0;
            if (on) {
                const bool was_bad = bad;
                bad |= vmax >= (uint32_t)L::BINS;
                if (!was_bad) out = q_out.at(K.head, prow + j - 1);     // un-count only what was counted
                if (!bad) in = q_in.at(K.head, prow + j + P.W - 1);
            }
            add_list_sync<L>(base, P, out, 0xFFFFu);
            add_list_sync<L>(base, P, in);
        }
        const bool later = on && count > 0 && (bad || count > kSmallMax);
        defer_cell(P, later && count <= kSmallMax, later && count > kSmallMax, ks, c);   // the slow kernel counts these windows pane by pane
        if (on && count > 0 && !later) {
            for (uint32_t iq = 0; iq < P.n_q; ++iq) {
                uint32_t ilo, ihi, vlo, vhi;
                double wt;
                pct_ranks(P.q[iq], count, ilo, ihi, wt);
                select2<L>(base, count, vmax, ilo, ihi, vlo, vhi);
                store_pct(P, K, c, iq, vlo, vhi, wt);
            }
        }
    }
}

}  // namespace kgs

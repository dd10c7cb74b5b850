// finalize_groups.cuh - the common case of a tumbling-window job finished without a histogram.
//
// In collector order (sorted by GPU, then time) the samples of one (gpu, window) cell are ONE
// contiguous piece of the value log - a single run, or the few runs of a window in which the
// tenant changed - and at 1 Hz with windows of up to a minute it has at most 64 values.  Such a
// cell fits in 17 registers (four values per word).  k_finalize_groups gives every (gpu, window)
// cell one thread which loads the cell's bytes ONCE and uses them three times:
//   * its own percentiles,
//   * the percentiles of the (pod, window) cell that consists of the very same run (one process
//     per GPU: the usual case; k_accum wrote the pod descriptor right before the gpu descriptor) -
//     and, in a window where the tenant changed, of the two pods' cells, each a sub-range of the
//     bytes the lane already holds,
//   * together with the other GPUs of its node - the threads of one (node, window) GROUP are
//     neighbouring lanes - the percentiles of the (node, window) cell, by summing the lanes' counts
//     with warp shuffles.
// Selection is an exact rank search by bisection on the value (rule R5 needs the values of rank lo
// and hi): "how many of my values are > t" is answered for four values at a time - add 127 - t to
// every byte (values are < 128, so no carry crosses a byte) and bit 7 of each byte is the answer;
// dp4a on the SIGNED bytes sums those flags for a whole word in one instruction:
//     sum_signed(x + c) = sum(x) + n_bytes * c - 256 * #{x > t}.
// Seven steps of 2 instructions per word, the same for every lane (no divergence, no shared
// memory, no dependent load-modify-store chain), instead of a 104-bin private histogram.
// Takes over, for these cells, what k_finalize_cells does (gpu_util_stats.py:77-80 for the mean;
// max and percentiles have no reference counterpart, rule R5 of oracle/window_stats.py).
//
// A cell that does not fit (more than 64 values, a value >= 128, runs that are not contiguous in
// the log, the 16-bit metric, sliding windows) is left alone: the kernel marks what it finished in
// P.done[] and k_finalize_cells, launched after it, skips exactly those cells.
#pragma once
#include "finalize.cuh"

namespace kgs {

constexpr int kGroupWords = 17;          // 64 values at any byte alignment
constexpr uint32_t kGroupMaxCount = 64;
constexpr int kGroupThreads = 128;

__device__ __forceinline__ int dp4a_s8u8(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Number of the lane's values that are > cand - 1, i.e. >= cand (1 <= cand <= 127).  x holds the values four per
// word, unused bytes are 0; byte_total = sum of all bytes of x.
__device__ __forceinline__ uint32_t count_ge(const uint32_t (&x)[kGroupWords], uint32_t byte_total, uint32_t cand) {
    const uint32_t c = 128u - cand;                  // 127 - (cand - 1), 1..127: x + c <= 254 per byte
    const uint32_t c4 = c * 0x01010101u;
    int a0 = 0, a1 = 0;
#pragma unroll
    for (int k = 0; k + 1 < kGroupWords; k += 2) {
        a0 = dp4a_s8u8(x[k] + c4, 0x01010101u, a0);
        a1 = dp4a_s8u8(x[k + 1] + c4, 0x01010101u, a1);
    }
    a0 = dp4a_s8u8(x[kGroupWords - 1] + c4, 0x01010101u, a0);
    return (byte_total + 4u * kGroupWords * c - (uint32_t)(a0 + a1)) >> 8;
}

// Sum over the lanes of a group (GS aligned lanes, a power of two); every lane gets the total.
template <int GS>
__device__ __forceinline__ uint32_t group_sum(uint32_t v) {
#pragma unroll
    for (int d = 1; d < GS; d <<= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
    return v;
}

// The values of rank r_lo and r_hi (0-based, ascending; r_hi = r_lo or r_lo + 1) among n values: the
// lane's own (GS = 1) or those of all GS lanes of its group together (n and the ranks are then the
// group's, equal in all its lanes).  Both ranks are found by bisection on the value, most significant
// bit first: v[r] >= cand  <=>  #{v >= cand} >= n - r.  The two searches ask the same question until
// one step has exactly n - r_hi values at or above its candidate; neighbouring order statistics differ
// in their low bits if at all, so the second count is needed in the last steps only (and only if some
// lane of the WARP has got there).
template <int GS>
__device__ __forceinline__ void rank_select2(const uint32_t (&x)[kGroupWords], uint32_t byte_total, uint32_t n,
                                             uint32_t r_lo, uint32_t r_hi, uint32_t& v_lo, uint32_t& v_hi) {
    const uint32_t need_hi = n - r_hi, need_lo = n - r_lo;
    uint32_t a_hi = 0, a_lo = 0;
#pragma unroll
    for (int b = 6; b >= 0; --b) {
        const uint32_t c_hi = a_hi | (1u << b), c_lo = a_lo | (1u << b);
        uint32_t ge_hi = count_ge(x, byte_total, c_hi);
        if (GS > 1) ge_hi = group_sum<GS>(ge_hi);
        uint32_t ge_lo = ge_hi;
        if (__any_sync(0xFFFFFFFFu, a_lo != a_hi)) {
            ge_lo = count_ge(x, byte_total, c_lo);
            if (GS > 1) ge_lo = group_sum<GS>(ge_lo);
        }
        if (ge_hi >= need_hi) a_hi = c_hi;
        if (ge_lo >= need_lo) a_lo = c_lo;
    }
    v_lo = a_lo;
    v_hi = a_hi;
}

// The lane's bytes [b0, b1) of the 4 * kGroupWords bytes that start at the word holding log position `start`.
__device__ __forceinline__ void load_values(const FinParams& P, uint32_t start, uint32_t b0, uint32_t b1,
                                            uint32_t (&x)[kGroupWords]) {
    const uint32_t w0 = start & ~3u;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(P.vals + w0);
#pragma unroll
    for (int k = 0; k < kGroupWords; ++k) {
        const int lo = (int)b0 - 4 * k, hi = (int)b1 - 4 * k;          // valid bytes of this word: [lo, hi)
        uint32_t w = 0;
        if (hi > 0 && lo < 4) w = __ldg(src + k);
        // bytes below hi, minus bytes below lo (both clamped to 0..4): two 64-bit shifts of all-ones
        const uint32_t below_hi = (uint32_t)(~(0xFFFFFFFFFFFFFFFFull << (8 * min(max(hi, 0), 4))));
        const uint32_t below_lo = (uint32_t)(~(0xFFFFFFFFFFFFFFFFull << (8 * min(max(lo, 0), 4))));
        x[k] = w & below_hi & ~below_lo;
    }
}

template <int GS>
__global__ void __launch_bounds__(kGroupThreads, 8) k_finalize_groups(const __grid_constant__ FinParams P) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    cudaGridDependencySynchronize();                 // launched as a programmatic dependent of k_accum (no-op otherwise)
    cudaTriggerProgrammaticLaunchCompletion();       // k_finalize_cells may be scheduled as this grid drains
    if (P.ctl[1]) return;                            // overflowed run log: lists are incomplete, the host reports it
    constexpr uint32_t gs = GS;                      // lanes per (node, window) group = 1 << P.g_gs_log2
    const uint32_t lane = lane_id();
    const uint64_t t = (uint64_t)blockIdx.x * kGroupThreads + threadIdx.x;
    const uint32_t l = (uint32_t)t & (gs - 1);
    const uint32_t grp = (uint32_t)(t >> P.g_gs_log2);
    const uint32_t node = grp / P.n_windows, j = grp - node * P.n_windows;
    const FinKind& G = P.kind[P.g_gpu];
    const uint32_t g = node * P.g_gpn + l;
    const bool have = node < P.g_nodes && l < P.g_gpn && g < G.n_keys;
    const uint32_t c = g * P.n_windows + j;          // W == 1: pane cell == window cell
    uint32_t count = 0, vmax = 0, h = 0;
    unsigned long long total = 0;
    if (have) {
        count = G.cnt[c];
        total = G.sum[c];
        vmax = G.mx[c];
        h = G.head[c];
    }
    // the cell's runs: at most four, together one contiguous piece of the value log; with the pod table
    // in the job, the pod descriptor of a run sits right before its gpu descriptor (k_accum, emit_run)
    bool ok = true;
    uint32_t start = 0, runs = 0;
    uint32_t run_start[4] = {0, 0, 0, 0}, run_len[4] = {0, 0, 0, 0}, run_pod[4] = {kInvalid, kInvalid, kInvalid, kInvalid};
    if (count) {
        ok = count <= kGroupMaxCount && vmax < 128u;
        if (ok) {
            uint32_t at = h, lo = 0xFFFFFFFFu, hi = 0, tot = 0;
#pragma unroll
            for (int r = 0; r < 4; ++r) {            // constant indices: the arrays stay in registers
                if (at) {
                    const uint4 d = __ldg(P.desc + (at - 1));
                    run_start[r] = d.x;
                    run_len[r] = d.y;
                    if (P.g_pod >= 0 && at >= 2) {
                        const uint4 sib = __ldg(P.desc + (at - 2));
                        if (sib.x == d.x && sib.y == d.y) run_pod[r] = sib.w;     // pod cell of this run (kInvalid: no pod)
                    }
                    lo = min(lo, d.x);
                    hi = max(hi, d.x + d.y);
                    tot += d.y;
                    at = d.z;
                    ++runs;
                }
            }
            ok = at == 0 && tot == count && hi - lo == count;
            start = lo;
        }
    }
    const uint32_t n = (ok && count) ? count : 0;    // values this lane holds
    uint32_t x[kGroupWords];
    load_values(P, start, start & 3u, (start & 3u) + n, x);
    const uint32_t byte_total = n ? (uint32_t)total : 0u;
    // ---- this GPU's cell, and the pod cell that holds exactly the same samples ----------------------------
    // A pod cell whose count equals the length of the runs of mine that carry its id consists of those runs
    // and nothing else.  Usually that is the whole cell (one process per GPU; a run cut in two where one
    // warp's share of the input ended is still one pod): the pod cell gets my numbers.
    const bool mine = have && (count == 0 || ok);    // this thread finishes its gpu cell
    uint32_t twin = kInvalid;
    bool split = false;                              // my runs belong to different pods: the tenant changed in this window
    if (mine && n && P.g_pod >= 0) {
        const bool same = (runs < 2 || run_pod[1] == run_pod[0]) && (runs < 3 || run_pod[2] == run_pod[0]) &&
                          (runs < 4 || run_pod[3] == run_pod[0]);
        if (same) {
            if (run_pod[0] != kInvalid && P.kind[P.g_pod].cnt[run_pod[0]] == count) twin = run_pod[0];
        } else {
            split = true;
        }
    }
    if (mine) {
        store_basics(P, G, c, count, total, vmax);
        P.done[G.cell_begin + c] = (uint8_t)P.done_gen;
        if (twin != kInvalid) {
            const FinKind& K = P.kind[P.g_pod];
            store_basics(P, K, twin, count, total, vmax);
            P.done[K.cell_begin + twin] = (uint8_t)P.done_gen;
        }
    }
    if (__any_sync(kAll, n != 0)) {
        for (uint32_t iq = 0; iq < P.n_q; ++iq) {
            uint32_t ilo = 0, ihi = 0;
            double w = 0.0;
            if (n) pct_ranks(P.q[iq], n, ilo, ihi, w);
            const uint32_t nn = n ? n : 1u;
            uint32_t vlo, vhi;
            rank_select2<1>(x, byte_total, nn, ilo, ihi, vlo, vhi);
            if (mine && n) {
                store_pct(P, G, c, iq, vlo, vhi, w);
                if (twin != kInvalid) store_pct(P, P.kind[P.g_pod], twin, iq, vlo, vhi, w);
            }
        }
    }
    // ---- the node's cell: the lanes of a group hold its values between them ----------------------------
    if (P.g_node >= 0) {                             // warp-uniform
        const FinKind& N = P.kind[P.g_node];
        const uint32_t group_lanes = (gs == 32 ? kAll : ((1u << gs) - 1u)) << (lane & ~(gs - 1));
        const bool lanes_ok = (__ballot_sync(kAll, ok) & group_lanes) == group_lanes;
        const uint32_t held = group_sum<GS>(n);
        const bool node_here = l == 0 && node < P.g_nodes && node < N.n_keys;
        const uint32_t nc = node * P.n_windows + j;
        uint32_t ncount = 0, nmax = 0;
        unsigned long long ntotal = 0;
        if (node_here) {
            ncount = N.cnt[nc];
            ntotal = N.sum[nc];
            nmax = N.mx[nc];
        }
        // every sample of the node cell must be in some lane's registers (a GPU outside the gpu table's
        // domain, or a cell some lane could not take, leaves the node cell to k_finalize_cells)
        const bool node_fast = node_here && lanes_ok && ncount == held;
        if (node_fast) {
            store_basics(P, N, nc, ncount, ntotal, nmax);
            P.done[N.cell_begin + nc] = (uint8_t)P.done_gen;
        }
        const uint32_t base_lane = lane & ~(gs - 1);
        const bool group_fast = __shfl_sync(kAll, (uint32_t)(node_fast && ncount > 0), base_lane) != 0;
        if (__any_sync(kAll, group_fast)) {
            const uint32_t gn = __shfl_sync(kAll, ncount, base_lane);
            for (uint32_t iq = 0; iq < P.n_q; ++iq) {
                uint32_t ilo = 0, ihi = 0;
                double w = 0.0;
                if (group_fast) pct_ranks(P.q[iq], gn, ilo, ihi, w);   // same n in every lane of the group: same ranks
                const uint32_t nn = group_fast ? gn : 1u;
                uint32_t vlo, vhi;
                rank_select2<GS>(x, byte_total, nn, ilo, ihi, vlo, vhi);
                if (node_fast && ncount) store_pct(P, N, nc, iq, vlo, vhi, w);
            }
        }
    }
    // ---- (pod, window) cells without a sample (a pod that did not exist yet, or no more): no thread above
    // ever gets to them; finishing them here leaves k_finalize_cells nothing to do in the usual job -------------
    if (P.g_pod >= 0) {
        const FinKind& K = P.kind[P.g_pod];
        const uint64_t n_pod_cells = (uint64_t)K.n_keys * P.n_windows, n_threads = (uint64_t)gridDim.x * kGroupThreads;
        for (uint64_t pc = t; pc < n_pod_cells; pc += n_threads) {
            if (K.cnt[pc] == 0) {
                store_basics(P, K, (uint32_t)pc, 0, 0, 0);
                P.done[K.cell_begin + (uint32_t)pc] = (uint8_t)P.done_gen;
            }
        }
    }
    // ---- a window in which the tenant changed: every pod's cell is a sub-range (or two) of the bytes this
    // lane has just used; they are read again (L1) with the other runs masked out ----------------------------
    if (P.g_pod < 0 || !__any_sync(kAll, split)) return;
    const FinKind& K = P.kind[P.g_pod];
#pragma unroll 1
    for (uint32_t r = 0; r < 4; ++r) {
        const uint32_t pod = r == 0 ? run_pod[0] : r == 1 ? run_pod[1] : r == 2 ? run_pod[2] : run_pod[3];
        // the first run with this pod id speaks for all runs of mine that carry it
        bool first = split && r < runs && pod != kInvalid;
        uint32_t len = 0, b0 = 0xFFFFFFFFu, b1 = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if ((uint32_t)q < runs && run_pod[q] == pod) {
                if ((uint32_t)q < r) first = false;
                len += run_len[q];
                b0 = min(b0, run_start[q]);
                b1 = max(b1, run_start[q] + run_len[q]);
            }
        }
        // its runs must be neighbours in the log (one byte range) and be all the pod cell holds
        uint32_t pn = 0, pmax = 0;
        unsigned long long ptotal = 0;
        bool here = first && b1 - b0 == len;
        if (here) {
            pn = K.cnt[pod];
            here = pn == len;
        }
        if (!__any_sync(kAll, here)) continue;
        if (here) {
            ptotal = K.sum[pod];
            pmax = K.mx[pod];
            store_basics(P, K, pod, pn, ptotal, pmax);
            P.done[K.cell_begin + pod] = (uint8_t)P.done_gen;
        }
        const uint32_t off = here ? b0 - (start & ~3u) : 0u;
        load_values(P, start, off, off + (here ? len : 0u), x);
        for (uint32_t iq = 0; iq < P.n_q; ++iq) {
            uint32_t ilo = 0, ihi = 0;
            double w = 0.0;
            if (here) pct_ranks(P.q[iq], pn, ilo, ihi, w);
            const uint32_t nn = here ? pn : 1u;
            uint32_t vlo, vhi;
            rank_select2<1>(x, (uint32_t)ptotal, nn, ilo, ihi, vlo, vhi);
            if (here) store_pct(P, K, pod, iq, vlo, vhi, w);
        }
    }
}

}  // namespace kgs

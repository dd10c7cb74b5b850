// accum.cuh - the hot kernel: one streaming pass over the 16 B/sample columns that does the
// (node,pid)->pod join and folds every sample into its (key, pane) cells for up to three key
// kinds at once.
//
// Takes over the per-sample loop of the reference (gpu_util_stats.py:72-78: decode, append,
// sum/len) and the dict probes of its join (gpu_util_stats.py:86-91).  Design:
//   * a warp owns a contiguous chunk of 256-sample tiles; each lane reads EIGHT consecutive
//     samples per column with one 256-bit load (LDG.E.256, sm_100), so a warp instruction moves
//     1 KB of one column (algorithmic traffic: 16 B/sample, read once); the next tile's loads are
//     issued before the current tile is processed;
//   * consecutive samples that share (gpu_gid, pid, pane) form a RUN; in collector order
//     (sorted by gpu, then time) runs are tens to thousands of samples long.  A lane walks its
//     eight samples sequentially; the usual lane is one piece of one run.  Lanes are chained:
//     a chain starts at a lane's LAST piece, runs over the following single-piece lanes with the
//     same key and ends by absorbing the FIRST piece of the next lane that has a boundary inside.
//     One segmented warp-shuffle reduction per tile (not per sample) sums each chain;
//   * chains become run records in a per-warp queue in shared memory; a run that continues over
//     a tile boundary is merged into the queued record;
//   * when 32 records are queued the warp retires them one per LANE: hash probe, one RED per
//     statistic and key kind, one run-log descriptor per key kind.  The join and all atomics
//     therefore cost once per run, executed 32 wide, and the run-log cursor (a single hot
//     address) is bumped once per 32 runs;
//   * sums are integers (u8/u16 metrics): RED.ADD.U64 / RED.MAX are order independent, so the
//     result is bit-exact whatever the scheduling;
//   * for percentiles the metric byte of every sample is appended to a value log (1 B/sample,
//     same order as the input) and each run leaves a 16 B descriptor {start, len, next, cell}
//     linked into its cell's list; the finalize kernels select from those.
// Any input order is handled correctly; run length only affects speed.
#pragma once
#include "common.cuh"

namespace kgs {

constexpr int kAccumWarps = 8;          // warps per CTA
constexpr int kLaneSamples = 8;         // consecutive samples per lane and tile
constexpr int kTileSamples = 32 * kLaneSamples;
constexpr int kQueueCap = 32 + 64;      // <= 32 kept + two records per lane
constexpr uint32_t kMergeMax = 32768;   // longest merged run (keeps the u32 sum exact for u16 metrics)

struct AccumParams {
    const uint32_t* ts;
    const uint32_t* gid;
    const uint32_t* pid;
    const uint32_t* util;
    uint32_t n;              // samples in this batch (< 2^32)
    uint32_t tiles_per_warp;
    uint64_t vals_base;      // offset of this batch in the value log (multiple of 16)
    const JoinSlot* slots;   // join table (nullptr when the pod kind is off)
    uint32_t tmask;
    uint32_t t0;
    FastDiv dstep;
    uint32_t n_panes;
    uint32_t step, win_len;  // win_len < step: sampled windows, only the first win_len seconds of a pane count
    FastDiv dgpn;
    uint32_t mshift, mmask;  // metric = (util >> mshift) & mmask
    uint32_t key_mask;
    uint32_t n_keys[3];
    uint32_t* cnt[3];              // [n_keys * n_panes]
    unsigned long long* sum[3];
    uint32_t* mx[3];
    uint32_t* head[3];             // run-list heads (1 + descriptor index, 0 = empty)
    uint8_t* vals;                 // value log
    uint4* desc;                   // run log
    uint32_t desc_cap;
    uint32_t log_base;             // this launch's private run-log slab: warp w owns
    uint32_t log_per_warp;         //   [log_base + w*log_per_warp, +log_per_warp)
    uint32_t pool_end;             // shared overflow pool = entries [0, pool_end), cursor ctl[0]
    uint32_t* ctl;                 // ctl[0] pool cursor, ctl[1] overflow flag, ctl[4] CTAs that used the accumulator cache
    uint32_t agg_on;               // every cell index < 2^30: the per-CTA accumulator cache may be used
    uint32_t dense_ok;             // step >= 8 and no sampled windows: the dense-tile shortcut applies
    uint32_t dense16_ok;           // step >= 16, an 8-bit metric: two tiles at once, 16 consecutive samples per lane
};

__device__ __forceinline__ uint64_t policy_evict_last() {
    uint64_t pol;
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));   // no side effect: the compiler may hoist it out of the tile loop
    return pol;
}
__device__ __forceinline__ void st_keep_v2(void* p, uint32_t a, uint32_t b, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v2.u32 [%0], {%1,%2}, %3;" ::"l"(p), "r"(a), "r"(b), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_keep_v4(void* p, uint4 v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol)
                 : "memory");
}

struct LogCursor {   // warp-uniform position inside the warp's private slab
    uint32_t at, left;
};

struct RunQueue {   // structure of arrays: lane l touching record l is conflict free
    uint32_t gid[kQueueCap], pid[kQueueCap], pane[kQueueCap], cnt[kQueueCap], sum[kQueueCap], mx[kQueueCap],
        start[kQueueCap];
};

// Per-CTA accumulator cache for input that is far from collector order (the flat random stream of
// BASELINE configs[2]): there a few hot cells (Zipf-sized pods) receive a large share of all
// samples and their global atomics serialise on one L2 address each.  Runs are first folded into
// this shared-memory table (open addressing, 4 probes, tag = cell*4 + kind + 1); the table is
// written out with one global RED per occupied slot when the CTA ends.  A run whose cell finds no
// slot goes to global memory directly, as always.
constexpr int kAggSlots = 1024;
struct AggTable {
    unsigned long long sum[kAggSlots];
    uint32_t tag[kAggSlots], cnt[kAggSlots], mx[kAggSlots];
};

__device__ __forceinline__ bool agg_add(AggTable& T, uint32_t tag, uint32_t c, uint32_t s, uint32_t m) {
    const uint32_t h = (tag * 0x9E3779B1u) >> 22;
#pragma unroll 1
    for (uint32_t probe = 0; probe < 4; ++probe) {
        const uint32_t slot = (h + probe) & (kAggSlots - 1);
        uint32_t old = *reinterpret_cast<volatile uint32_t*>(&T.tag[slot]);
        if (old == 0) old = atomicCAS(&T.tag[slot], 0u, tag);
        if (old == 0 || old == tag) {
            atomicAdd(&T.cnt[slot], c);
            atomicAdd(&T.sum[slot], (unsigned long long)s);
            atomicMax(&T.mx[slot], m);
            return true;
        }
    }
    return false;
}

// Retire one run: probe the join table once, then one RED per statistic and key kind and one
// run-log descriptor per key kind (slots di .. di+nk-1).
template <bool PCT>
__device__ __forceinline__ void emit_run(const AccumParams& P, AggTable* agg, uint32_t g, uint32_t pid, uint32_t pane,
                                         uint32_t c, uint32_t s, uint32_t m, uint32_t start, uint32_t di,
                                         bool log_ok) {
    const uint32_t node = fdiv(g, P.dgpn);
#pragma unroll
    for (int kind = 0; kind < 3; ++kind) {
        if (!(P.key_mask & (1u << kind))) continue;
        uint32_t key;
        if (kind == 0) key = join_lookup(P.slots, P.tmask, node, pid);
        else if (kind == 1) key = g;
        else key = node;
        const bool valid = key < P.n_keys[kind];
        uint32_t cell = kInvalid;
        if (valid) {
            cell = key * P.n_panes + pane;
            if (!(agg && agg_add(*agg, cell * 4 + kind + 1, c, s, m))) {
                atomicAdd(P.cnt[kind] + cell, c);
                atomicAdd(P.sum[kind] + cell, (unsigned long long)s);
                atomicMax(P.mx[kind] + cell, m);
            }
        }
        if (PCT && log_ok) {
            uint32_t next = 0;
            if (valid) next = atomicExch(P.head[kind] + cell, di + 1);
            st_keep_v4(P.desc + di, make_uint4(start, c, next, cell), policy_evict_last());
            ++di;
        }
    }
}

// Reserve `count` run-log entries for the warp; returns the first index.  Entries come from the
// warp's private slab (no atomics, nothing to wait for); a warp that exhausts its slab falls back
// to the shared pool, whose cursor is a single hot address.
__device__ __forceinline__ uint32_t reserve_log(const AccumParams& P, LogCursor& lc, uint32_t count, bool& log_ok) {
    if (count <= lc.left) {
        const uint32_t base = lc.at;
        lc.at += count;
        lc.left -= count;
        log_ok = true;
        return base;
    }
    uint32_t base = 0;
    if (lane_id() == 0) base = atomicAdd(P.ctl, count);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    log_ok = (uint64_t)base + count <= P.pool_end;
    if (!log_ok && lane_id() == 0) P.ctl[1] = 1;
    return base;
}

// Retire records [0, n_flush) of the warp's queue, one record per lane.
template <bool PCT>
__device__ __forceinline__ void flush_queue(const AccumParams& P, AggTable* agg, LogCursor& lc, const RunQueue& q,
                                            uint32_t n_flush, uint32_t nk) {
    const uint32_t lane = lane_id();
    for (uint32_t base = 0; base < n_flush; base += 32) {
        const uint32_t here = min(32u, n_flush - base);
        bool log_ok = false;
        uint32_t di = 0;
        if (PCT) di = reserve_log(P, lc, here * nk, log_ok) + lane * nk;
        const uint32_t r = base + lane;
        if (lane < here)
            emit_run<PCT>(P, agg, q.gid[r], q.pid[r], q.pane[r], q.cnt[r], q.sum[r], q.mx[r], q.start[r], di, log_ok);
    }
}

struct Oct {   // one lane's eight consecutive samples
    uint32_t t[8], g[8], p[8], u[8];
};

__device__ __forceinline__ void ld_stream_v8(const uint32_t* p, uint32_t* r) {
    // read once: keep it out of L1 and first in line for eviction from L2, so the value log, run log
    // and accumulators this kernel writes are still L2-resident when the finalize kernels read them
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}

__device__ __forceinline__ void load_oct(const AccumParams& P, uint32_t i0, bool full, Oct& o) {
    if (full) {
        ld_stream_v8(P.ts + i0, o.t);
        ld_stream_v8(P.gid + i0, o.g);
        ld_stream_v8(P.pid + i0, o.p);
        ld_stream_v8(P.util + i0, o.u);
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const bool in = i0 + j < P.n;   // the last tile only; i0 + j stays below 2^32 (n < 2^32 - 256)
            o.t[j] = in ? P.ts[i0 + j] : 0u;
            o.g[j] = in ? P.gid[i0 + j] : 0u;
            o.p[j] = in ? P.pid[i0 + j] : 0u;
            o.u[j] = in ? P.util[i0 + j] : 0u;
        }
    }
}

// What the chain machinery needs to know about one lane's consecutive samples.
struct LaneRuns {
    uint32_t g0, p0, pane0;        // key of the lane's first sample
    uint32_t g1, p1, pane1;        // key of its last sample
    uint32_t fc, fs, fm, fstart;   // forward piece (the whole lane when single, else its last piece): count, sum, max, offset
    uint32_t ac, as, am;           // first piece of a lane with a boundary inside
    bool single;                   // no boundary inside the lane
};

// Chains: a chain starts at a lane's LAST piece, runs over the following single-piece lanes with the same key and
// absorbs the first piece of the lane that ends it.  One segmented shuffle reduction sums every chain of the tile;
// chains (and stand-alone first pieces) become records of the warp's queue, a record that continues the queued tail
// is merged into it.  LS = samples per lane (8: the general tile; 16: two tiles at once, dense input only).
template <int LS>
__device__ __forceinline__ void append_chains(RunQueue& Q, uint32_t& queued, const LaneRuns& L, uint32_t start0, uint32_t lane) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    const bool single = L.single;
    // does my first piece continue the previous lane's last piece?
    const uint32_t pg = __shfl_up_sync(kAll, L.g1, 1);
    const uint32_t pp = __shfl_up_sync(kAll, L.p1, 1);
    const uint32_t ppn = __shfl_up_sync(kAll, L.pane1, 1);
    const bool cont = lane > 0 && pg == L.g0 && pp == L.p0 && ppn == L.pane0;
    // a chain starts at every lane with a boundary inside (its last piece) and at every single
    // lane that does not continue its predecessor
    const bool head = !single || !cont;
    const uint32_t heads = __ballot_sync(kAll, head);
    const uint32_t rest = (lane == 31) ? 0u : (heads & (0xFFFFFFFEu << lane));
    const uint32_t seg_end = rest ? (uint32_t)(__ffs(rest) - 2) : 31u;
    uint32_t fs = L.fs, fm = L.fm;
    // chain totals: sum and max by a segmented shuffle reduction, count is implied
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t s2 = __shfl_down_sync(kAll, fs, d);
        const uint32_t m2 = __shfl_down_sync(kAll, fm, d);
        if (lane + d <= seg_end) {
            fs += s2;
            fm = max(fm, m2);
        }
    }
    uint32_t cc = L.fc + (uint32_t)LS * (seg_end - lane);
    {   // absorb the first piece of the split lane that ends the chain, if it has the chain's key
        const uint32_t j = min(seg_end + 1, 31u);
        const bool joins = !single && cont;
        const uint32_t jj = __shfl_sync(kAll, (uint32_t)joins, j);
        const uint32_t jc = __shfl_sync(kAll, L.ac, j);
        const uint32_t js = __shfl_sync(kAll, L.as, j);
        const uint32_t jm = __shfl_sync(kAll, L.am, j);
        if (seg_end < 31 && jj) {
            cc += jc;
            fs += js;
            fm = max(fm, jm);
        }
    }
    // records this lane appends: its chain (head lanes, valid key) and a stand-alone first piece
    const bool rec_chain = head && L.pane1 != kInvalid;
    const bool rec_first = !single && !cont && L.pane0 != kInvalid;
    // the record that starts exactly at the tile's first sample may continue the queued tail
    bool merged = false;
    if (lane == 0 && queued && (single ? rec_chain : rec_first)) {
        const uint32_t r = queued - 1;
        const uint32_t mc = single ? cc : L.ac, ms = single ? fs : L.as, mm = single ? fm : L.am;
        if (Q.gid[r] == L.g0 && Q.pid[r] == L.p0 && Q.pane[r] == L.pane0 && Q.start[r] + Q.cnt[r] == start0 &&
            Q.cnt[r] + mc <= kMergeMax) {
            Q.cnt[r] += mc;
            Q.sum[r] += ms;
            Q.mx[r] = max(Q.mx[r], mm);
            merged = true;
        }
    }
    merged = __shfl_sync(kAll, (uint32_t)merged, 0) != 0;
    const uint32_t m_first = __ballot_sync(kAll, rec_first);
    const uint32_t m_chain = __ballot_sync(kAll, rec_chain);
    const uint32_t below = (1u << lane) - 1u;
    // slot order = sample order: a lane's first piece precedes its chain
    uint32_t slot = queued + __popc(m_first & below) + __popc(m_chain & below) - ((merged && lane > 0) ? 1u : 0u);
    if (rec_first) {
        if (!(lane == 0 && merged)) {
            Q.gid[slot] = L.g0; Q.pid[slot] = L.p0; Q.pane[slot] = L.pane0;
            Q.cnt[slot] = L.ac; Q.sum[slot] = L.as; Q.mx[slot] = L.am; Q.start[slot] = start0;
        }
        slot += (lane == 0 && merged) ? 0u : 1u;
    }
    if (rec_chain && !(lane == 0 && merged && single)) {
        Q.gid[slot] = L.g1; Q.pid[slot] = L.p1; Q.pane[slot] = L.pane1;
        Q.cnt[slot] = cc; Q.sum[slot] = fs; Q.mx[slot] = fm; Q.start[slot] = start0 + L.fstart;
    }
    queued += __popc(m_first) + __popc(m_chain) - (merged ? 1u : 0u);
}

// Two tiles at once for dense input.  Each tile is loaded the usual way (a lane reads 8 consecutive samples per
// column with one 256-bit load, a warp instruction covers 1 KB without gaps) and reduced to a lane summary: the
// 8 values packed into two words (they go to the value log as they are), the first sample's key and time, and
// whether the lane is 8 consecutive seconds of one (gpu, pid).  If that holds for all lanes of both tiles and every
// odd lane continues its even neighbour, the summaries of lanes (2l, 2l+1) are brought together by shuffles: lane l
// then stands for SIXTEEN consecutive samples, its pane boundary, if any, follows from one division, the piece sums
// come from dp4a on the masked words, and ONE pass of the chain machinery serves 512 samples instead of 256.
// Returns false (nothing done, nothing written) for any other pair of tiles; the caller then takes the general
// 8-samples-per-lane path for them.  Needs step >= 16 (at most one boundary in 16 seconds).
struct LaneSummary {
    uint32_t w0, w1, t0, g0, p0;
    bool ok;                     // dense lane, and (odd lanes) the continuation of the lane before
};

__device__ __forceinline__ LaneSummary summarize_lane(const AccumParams& P, uint32_t i0, uint32_t lane) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    Oct o;
    load_oct(P, i0, true, o);
    LaneSummary S;
    S.t0 = o.t[0]; S.g0 = o.g[0]; S.p0 = o.p[0];
    uint32_t acc = 0;
#pragma unroll
    for (int j = 1; j < 8; ++j) acc |= ((o.t[j] - S.t0) ^ (uint32_t)j) | (o.g[j] ^ S.g0) | (o.p[j] ^ S.p0);
    uint32_t v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (o.u[j] >> P.mshift) & 0xFFu;
    S.w0 = v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
    S.w1 = v[4] | (v[5] << 8) | (v[6] << 16) | (v[7] << 24);
    // an odd lane must continue its even neighbour: 8 seconds later, same gpu and process
    const uint32_t pt = __shfl_up_sync(kAll, S.t0, 1), pg = __shfl_up_sync(kAll, S.g0, 1), pp = __shfl_up_sync(kAll, S.p0, 1);
    const bool pair = !(lane & 1u) || (S.t0 == pt + 8u && S.g0 == pg && S.p0 == pp);
    S.ok = acc == 0 && pair && S.t0 >= P.t0 && S.t0 < 0xFFFFFF00u;
    return S;
}

template <bool PCT>
__device__ __forceinline__ bool dense16_tiles(const AccumParams& P, RunQueue& Q, uint32_t& queued, uint32_t tile, uint32_t lane) {
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    const uint32_t base = tile * kTileSamples;
    const LaneSummary A = summarize_lane(P, base + 8 * lane, lane);
    if (!__all_sync(kAll, A.ok)) return false;
    const LaneSummary B = summarize_lane(P, base + kTileSamples + 8 * lane, lane);
    if (!__all_sync(kAll, B.ok)) return false;
    if (PCT) {
        st_keep_v2(P.vals + P.vals_base + base + 8 * lane, A.w0, A.w1, policy_evict_last());
        st_keep_v2(P.vals + P.vals_base + base + kTileSamples + 8 * lane, B.w0, B.w1, policy_evict_last());
    }
    // lane l now stands for samples [16 l, 16 l + 16) of the pair: summaries 2l, 2l+1 of tile A (l < 16) or B
    const uint32_t even = (2u * lane) & 31u, odd = even + 1u;
    const bool from_a = lane < 16;
    uint32_t w[4];
    {
        const uint32_t a0 = __shfl_sync(kAll, A.w0, even), b0 = __shfl_sync(kAll, B.w0, even);
        const uint32_t a1 = __shfl_sync(kAll, A.w1, even), b1 = __shfl_sync(kAll, B.w1, even);
        const uint32_t a2 = __shfl_sync(kAll, A.w0, odd), b2 = __shfl_sync(kAll, B.w0, odd);
        const uint32_t a3 = __shfl_sync(kAll, A.w1, odd), b3 = __shfl_sync(kAll, B.w1, odd);
        w[0] = from_a ? a0 : b0; w[1] = from_a ? a1 : b1; w[2] = from_a ? a2 : b2; w[3] = from_a ? a3 : b3;
    }
    const uint32_t ta = __shfl_sync(kAll, A.t0, even), tb = __shfl_sync(kAll, B.t0, even);
    const uint32_t ga = __shfl_sync(kAll, A.g0, even), gb = __shfl_sync(kAll, B.g0, even);
    const uint32_t pa = __shfl_sync(kAll, A.p0, even), pb = __shfl_sync(kAll, B.p0, even);
    const uint32_t t0 = from_a ? ta : tb, g0 = from_a ? ga : gb, p0 = from_a ? pa : pb;
    const uint32_t i0 = base + 16 * lane;
    const uint32_t x0 = t0 - P.t0;
    const uint32_t pn0 = fdiv(x0, P.dstep);
    const uint32_t left = P.step - (x0 - pn0 * P.step);          // samples from t0 to the end of pane pn0
    const uint32_t p_a = pn0 < P.n_panes ? pn0 : kInvalid;
    const uint32_t p_b = pn0 + 1 < P.n_panes ? pn0 + 1 : kInvalid;
    const bool split = left < 16 && p_a != p_b;
    const uint32_t k = split ? left : 0u;                        // first piece = samples [0, k), forward piece = [k, 16)
    uint32_t tot = 0, as = 0, am = 0, fm = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int below = min(max((int)k - 4 * i, 0), 4);       // bytes of word i that belong to the first piece
        const uint32_t m = (uint32_t)(~(0xFFFFFFFFFFFFFFFFull << (8 * below)));
        const uint32_t aw = w[i] & m, fw = w[i] & ~m;
        tot = __dp4a(w[i], 0x01010101u, tot);
        as = __dp4a(aw, 0x01010101u, as);
        am = max(am, max(max(aw & 0xFFu, (aw >> 8) & 0xFFu), max((aw >> 16) & 0xFFu, aw >> 24)));
        fm = max(fm, max(max(fw & 0xFFu, (fw >> 8) & 0xFFu), max((fw >> 16) & 0xFFu, fw >> 24)));
    }
    LaneRuns L;
    L.single = !split;
    L.g0 = g0; L.p0 = p0; L.pane0 = p_a;
    L.g1 = g0; L.p1 = p0; L.pane1 = (left < 16) ? p_b : p_a;
    L.fc = 16 - k; L.fs = tot - as; L.fm = fm; L.fstart = k;
    L.ac = k; L.as = as; L.am = am;
    append_chains<16>(Q, queued, L, (uint32_t)P.vals_base + i0, lane);
    return true;
}

// ---- TMA staging (TMA = true) ----------------------------------------------------------------------
// Each warp owns kStages tile buffers in shared memory (4 columns x 1 KB) and one mbarrier
// per buffer.  Lane 0 arms the barrier with the 4 KB it expects and issues four 1-D bulk copies
// (cp.async.bulk -> UBLKCP): the copy engine moves the next tiles while the warp works on the current
// one, with no registers held for data in flight and no load instructions to issue.  The warp waits
// on the barrier's phase bit, reads its 8 samples per column with two LDS.128, and re-arms the
// buffer with the tile kStages ahead.  Warps never wait for each other.
constexpr int kStages = 1;   // one buffer per warp: it is refilled as soon as the warp has its tile in registers (32 KB per CTA: 4 CTAs per SM)
constexpr uint32_t kTileBytes = 4 * kTileSamples * 4;   // four u32 columns
constexpr uint32_t kTmaSmemBytes = kAccumWarps * kStages * (kTileBytes + 8);

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a) : "memory");
    return r;
}

template <bool PCT, bool TMA>
__global__ void __launch_bounds__(kAccumWarps * 32, 4) k_accum(const __grid_constant__ AccumParams P) {
    // dynamic shared memory: with TMA the warps' tile buffers and their barriers, else the accumulator cache
    // (the two are never needed together: the cache is for input far from collector order, the staging for dense input)
    extern __shared__ __align__(128) unsigned char tma_smem[];
    cudaTriggerProgrammaticLaunchCompletion();   // the finalize kernel behind this one may be scheduled while the last CTAs drain
    __shared__ RunQueue queues[kAccumWarps];
    AggTable& table = *reinterpret_cast<AggTable*>(tma_smem);
    constexpr uint32_t kAll = 0xFFFFFFFFu;
    if (!TMA && P.agg_on) {
        for (uint32_t i = threadIdx.x; i < kAggSlots; i += blockDim.x) {
            table.tag[i] = 0; table.cnt[i] = 0; table.mx[i] = 0; table.sum[i] = 0;
        }
        __syncthreads();
    }
    AggTable* agg = nullptr;   // switched on (per warp) by the first tile that is mostly many-piece lanes
    const uint32_t lane = lane_id();
    RunQueue& Q = queues[threadIdx.x >> 5];
    const uint32_t nk = __popc(P.key_mask);
    const uint32_t n_tiles = (uint32_t)(((uint64_t)P.n + kTileSamples - 1) / kTileSamples);
    const uint32_t warp = blockIdx.x * kAccumWarps + (threadIdx.x >> 5);
    const uint64_t tb = (uint64_t)warp * P.tiles_per_warp;
    const uint32_t tile_begin = (uint32_t)min((uint64_t)n_tiles, tb);
    const uint32_t tile_end = (uint32_t)min((uint64_t)n_tiles, tb + P.tiles_per_warp);
    const uint32_t full_tiles = P.n / kTileSamples;   // tiles below this index are complete
    uint32_t queued = 0;                               // warp-uniform
    LogCursor lc{P.log_base + warp * P.log_per_warp, P.log_per_warp};

    Oct cur;
    uint32_t ring = 0, bars = 0;     // shared-space addresses of this warp's tile buffers and barriers
    uint64_t pol_stream = 0;
    auto tma_issue = [&](uint32_t stage, uint32_t tile) {   // lane 0 only: arm the barrier, start the four copies
        const uint32_t dst = ring + stage * kTileBytes, bar = bars + stage * 8;
        const size_t at = (size_t)tile * kTileSamples;
        mbar_expect_tx(bar, kTileBytes);
        bulk_g2s(dst, P.ts + at, kTileSamples * 4, bar, pol_stream);
        bulk_g2s(dst + kTileSamples * 4, P.gid + at, kTileSamples * 4, bar, pol_stream);
        bulk_g2s(dst + 2 * kTileSamples * 4, P.pid + at, kTileSamples * 4, bar, pol_stream);
        bulk_g2s(dst + 3 * kTileSamples * 4, P.util + at, kTileSamples * 4, bar, pol_stream);
    };
    if (TMA) {
        const uint32_t wib = threadIdx.x >> 5;
        ring = (uint32_t)__cvta_generic_to_shared(tma_smem) + wib * kStages * kTileBytes;
        bars = (uint32_t)__cvta_generic_to_shared(tma_smem) + kAccumWarps * kStages * kTileBytes + wib * kStages * 8;
        if (lane == 0) {
            pol_stream = policy_evict_first();
#pragma unroll
            for (int st = 0; st < kStages; ++st) mbar_init(bars + st * 8, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
        if (lane == 0) {
#pragma unroll
            for (int st = 0; st < kStages; ++st)
                if (tile_begin + st < tile_end && tile_begin + st < full_tiles) tma_issue(st, tile_begin + st);
        }
    }
    uint32_t skip16 = 0, penalty16 = 8;   // after a pair of tiles that was not dense: no further attempt for a while
    auto drain = [&]() {   // retire all but the newest record (it may still grow)
        __syncwarp();
        if (queued > 32) {
            flush_queue<PCT>(P, agg, lc, Q, queued - 1, nk);
            __syncwarp();
            if (lane == 0) {
                const uint32_t r = queued - 1;
                Q.gid[0] = Q.gid[r]; Q.pid[0] = Q.pid[r]; Q.pane[0] = Q.pane[r];
                Q.cnt[0] = Q.cnt[r]; Q.sum[0] = Q.sum[r]; Q.mx[0] = Q.mx[r]; Q.start[0] = Q.start[r];
            }
            queued = 1;
            __syncwarp();
        }
    };
    for (uint32_t tile = tile_begin; tile < tile_end; ++tile) {
        if (!TMA && P.dense16_ok && tile + 1 < tile_end && tile + 1 < full_tiles) {   // warp-uniform
            if (skip16) {
                --skip16;
            } else if (dense16_tiles<PCT>(P, Q, queued, tile, lane)) {
                penalty16 = 8;
                drain();
                ++tile;
                continue;
            } else {
                skip16 = penalty16;
                penalty16 = min(penalty16 * 2, 1024u);
            }
        }
        const uint32_t i0 = tile * kTileSamples + 8 * lane;
        const bool full = tile < full_tiles;
        if (TMA && full) {
            const uint32_t k = tile - tile_begin, stage = k % kStages, parity = (k / kStages) & 1u;
            uint32_t spins = 0;
            while (!mbar_try_wait(bars + stage * 8, parity)) {
                if (++spins > (1u << 24)) { P.ctl[1] = 2; break; }   // never observed; a stuck copy must not hang the GPU
            }
            const uint32_t src = ring + stage * kTileBytes + lane * 32;
            uint4 a = lds_v4(src), b = lds_v4(src + 16);
            cur.t[0] = a.x; cur.t[1] = a.y; cur.t[2] = a.z; cur.t[3] = a.w; cur.t[4] = b.x; cur.t[5] = b.y; cur.t[6] = b.z; cur.t[7] = b.w;
            a = lds_v4(src + kTileSamples * 4); b = lds_v4(src + kTileSamples * 4 + 16);
            cur.g[0] = a.x; cur.g[1] = a.y; cur.g[2] = a.z; cur.g[3] = a.w; cur.g[4] = b.x; cur.g[5] = b.y; cur.g[6] = b.z; cur.g[7] = b.w;
            a = lds_v4(src + 2 * kTileSamples * 4); b = lds_v4(src + 2 * kTileSamples * 4 + 16);
            cur.p[0] = a.x; cur.p[1] = a.y; cur.p[2] = a.z; cur.p[3] = a.w; cur.p[4] = b.x; cur.p[5] = b.y; cur.p[6] = b.z; cur.p[7] = b.w;
            a = lds_v4(src + 3 * kTileSamples * 4); b = lds_v4(src + 3 * kTileSamples * 4 + 16);
            cur.u[0] = a.x; cur.u[1] = a.y; cur.u[2] = a.z; cur.u[3] = a.w; cur.u[4] = b.x; cur.u[5] = b.y; cur.u[6] = b.z; cur.u[7] = b.w;
            // The tile is in registers: refill the buffer with the next tile NOW, so that the copy runs while this
            // tile is processed.  Every lane's eight shared-memory loads must have returned before the copy engine
            // may overwrite the buffer: the vote below cannot complete before each lane has used one word of each.
            const uint32_t touch = cur.t[0] ^ cur.t[4] ^ cur.g[0] ^ cur.g[4] ^ cur.p[0] ^ cur.p[4] ^ cur.u[0] ^ cur.u[4];
            const uint32_t voted = __ballot_sync(kAll, touch == 0x5A5A5A5Au);   // only its completion matters
            asm volatile("" ::"r"(voted) : "memory");
            const uint32_t next = tile + 1;
            if (lane == 0 && next < tile_end && next < full_tiles) tma_issue(0u, next);
        } else {
            load_oct(P, i0, full, cur);
        }
        uint32_t pane[8], v[8], bm = 0;   // bm: bit j set when sample j starts a new piece inside the lane
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = (cur.u[j] >> P.mshift) & P.mmask;
        // Dense tile: every lane holds 8 consecutive seconds of one (gpu, pid).  Then the lane's panes
        // follow from ONE division: at most one pane boundary falls inside it (step >= 8), at a
        // position known from the remainder.  Collector-ordered 1 Hz data takes this path except
        // where a GPU stream or a tenancy ends inside the tile.
        bool dense = false;
        if (full && P.dense_ok) {
            uint32_t acc = 0;
#pragma unroll
            for (int j = 1; j < 8; ++j)
                acc |= ((cur.t[j] - cur.t[0]) ^ (uint32_t)j) | (cur.g[j] ^ cur.g[0]) | (cur.p[j] ^ cur.p[0]);
            dense = __all_sync(kAll, acc == 0 && cur.t[0] >= P.t0 && cur.t[0] < 0xFFFFFF00u);
        }
        if (dense) {
            const uint32_t x0 = cur.t[0] - P.t0;
            const uint32_t pn0 = fdiv(x0, P.dstep);
            const uint32_t left = P.step - (x0 - pn0 * P.step);          // samples from t[0] to the end of pane pn0
            const uint32_t p_a = pn0 < P.n_panes ? pn0 : kInvalid;
            const uint32_t p_b = pn0 + 1 < P.n_panes ? pn0 + 1 : kInvalid;
            // only the lane's first and last pane are used below (the per-sample panes matter to lanes with
            // two or more boundaries, which a dense tile cannot have)
#pragma unroll
            for (int j = 0; j < 7; ++j) pane[j] = p_a;
            pane[7] = left < 8 ? p_b : p_a;
            if (left < 8 && p_a != p_b) bm = 1u << left;
        } else {
            const uint32_t lim = full ? 8u : (P.n > i0 ? P.n - i0 : 0u);   // samples of this lane inside the batch
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t pn = fdiv(cur.t[j] - P.t0, P.dstep);
                bool ok = (cur.t[j] >= P.t0) & (pn < P.n_panes) & ((uint32_t)j < lim);
                if (P.win_len < P.step) ok &= (cur.t[j] - P.t0) - pn * P.step < P.win_len;   // warp-uniform branch
                pane[j] = ok ? pn : kInvalid;
            }
#pragma unroll
            for (int j = 1; j < 8; ++j) {
                const uint32_t diff = (cur.g[j] ^ cur.g[j - 1]) | (cur.p[j] ^ cur.p[j - 1]) | (pane[j] ^ pane[j - 1]);
                bm += min(diff, 1u) << j;   // distinct bits: + is |
            }
        }
        if (PCT && P.mmask > 0xFFu) {   // value log of a 16-bit metric (fb_mib): 2 B/sample (warp-uniform branch)
            uint16_t* dst = reinterpret_cast<uint16_t*>(P.vals) + P.vals_base + i0;
            if (full) {
                st_keep_v4(dst, make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16)),
                           policy_evict_last());
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (i0 + j < P.n) dst[j] = (uint16_t)v[j];
            }
        } else if (PCT) {   // value log, same order as the input (u8 metrics): 1 B/sample
            uint8_t* dst = P.vals + P.vals_base + i0;
            if (full) {
                st_keep_v2(dst, v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24),
                           v[4] | (v[5] << 8) | (v[6] << 16) | (v[7] << 24), policy_evict_last());
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (i0 + j < P.n) dst[j] = (uint8_t)v[j];
            }
        }
        LaneRuns L;
        L.single = bm == 0;
        L.g0 = cur.g[0]; L.p0 = cur.p[0]; L.pane0 = pane[0];
        L.g1 = cur.g[7]; L.p1 = cur.p[7]; L.pane1 = pane[7];
        L.fc = 8; L.fstart = 0; L.ac = 0; L.as = 0; L.am = 0;
        if (L.single) {
            L.fs = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
            L.fm = max(max(max(v[0], v[1]), max(v[2], v[3])), max(max(v[4], v[5]), max(v[6], v[7])));
        } else {
            const uint32_t first_end = __ffs(bm) - 1;             // first piece = samples [0, first_end)
            const uint32_t last_begin = 31 - __clz(bm);           // last piece  = samples [last_begin, 8)
            L.fs = 0; L.fm = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if ((uint32_t)j < first_end) { L.as += v[j]; L.am = max(L.am, v[j]); }
                if ((uint32_t)j >= last_begin) { L.fs += v[j]; L.fm = max(L.fm, v[j]); }
            }
            L.ac = first_end;
            L.fc = 8 - last_begin;
            L.fstart = last_begin;
        }
        const uint32_t start0 = (uint32_t)P.vals_base + i0;
        append_chains<kLaneSamples>(Q, queued, L, start0, lane);
        // lanes with two or more boundaries: the pieces between the first and the last are retired
        // directly (rare in collector order; every lane in shuffled input, then 32 wide)
        const bool many = (bm & (bm - 1)) != 0;
        uint32_t many_mask = 0;
        if (!dense) {                                  // warp-uniform
            many_mask = __ballot_sync(kAll, many);
            if (!TMA && P.agg_on && __popc(many_mask) >= 16) agg = &table;
        }
        if (many_mask) {
            uint32_t nr = 0;
            if (many) {
#pragma unroll
                for (int j = 1; j < 8; ++j) {      // middle piece starting at j: a boundary at j, another one later
                    const bool starts = (bm >> j) & 1u;
                    const bool last = (bm >> (j + 1)) == 0;
                    nr += (starts && !last && pane[j] != kInvalid) ? 1u : 0u;
                }
            }
            uint32_t di = 0;
            bool log_ok = false;
            if (PCT) {
                uint32_t incl = nr;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(kAll, incl, d);
                    if (lane >= (uint32_t)d) incl += o;
                }
                const uint32_t total = __shfl_sync(kAll, incl, 31) * nk;
                di = reserve_log(P, lc, total, log_ok) + (incl - nr) * nk;
            }
            if (nr) {
                uint32_t rs = 0, rm = 0, rc = 0, rstart = 0;
                bool open = false;
#pragma unroll
                for (int j = 1; j < 8; ++j) {
                    const bool starts = (bm >> j) & 1u;
                    const bool last = (bm >> (j + 1)) == 0;
                    if (starts) {
                        if (open && pane[j - 1] != kInvalid) {
                            emit_run<PCT>(P, agg, cur.g[j - 1], cur.p[j - 1], pane[j - 1], rc, rs, rm, start0 + rstart, di, log_ok);
                            di += nk;
                        }
                        open = !last;
                        rs = 0; rm = 0; rc = 0; rstart = j;
                    }
                    if (open) {
                        rs += v[j];
                        rm = max(rm, v[j]);
                        rc += 1;
                    }
                }
            }
        }
        drain();
    }
    flush_queue<PCT>(P, agg, lc, Q, queued, nk);
    if (!TMA && P.agg_on) {   // write the CTA's accumulator cache out: one RED per occupied slot and statistic
        __syncthreads();
        bool used = false;
        for (uint32_t i = threadIdx.x; i < kAggSlots; i += blockDim.x) {
            const uint32_t tag = table.tag[i];
            if (tag) {
                const uint32_t kind = (tag - 1) & 3u, cell = (tag - 1) >> 2;
                atomicAdd(P.cnt[kind] + cell, table.cnt[i]);
                atomicAdd(P.sum[kind] + cell, table.sum[i]);
                atomicMax(P.mx[kind] + cell, table.mx[i]);
                used = true;
            }
        }
        if (__syncthreads_or(used) && threadIdx.x == 0) atomicAdd(P.ctl + 4, 1u);   // job counter: CTAs that used the cache
    }
}

}  // namespace kgs

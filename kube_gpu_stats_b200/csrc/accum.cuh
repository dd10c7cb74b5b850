// accum.cuh - the hot kernel: one streaming pass over the 16 B/sample columns that does the
// (node,pid)->pod join and folds every sample into its (key, pane) cells for up to three key
// kinds at once.
//
// Takes over the per-sample loop of the reference (gpu_util_stats.py:72-78: decode, append,
// sum/len) and the dict probes of its join (gpu_util_stats.py:86-91).  Design:
//   * a warp owns a contiguous chunk of 128-sample tiles and reads each tile with one LDG.128
//     per column and lane (algorithmic traffic: 16 B/sample, read once); the next tile's loads
//     are issued before the current tile is processed;
//   * consecutive samples that share (gpu_gid, pid, pane) form a RUN; in collector order
//     (sorted by gpu, then time) runs are tens to thousands of samples long.  Lanes whose four
//     samples are one run are combined with a segmented warp-shuffle reduction into one run
//     record per segment; records go to a per-warp queue in shared memory, a run that
//     continues across a tile boundary is merged into the queued record;
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
constexpr int kQueueCap = 64;           // run records per warp queue (<= 31 kept + 32 new + 1)
constexpr uint32_t kMergeMax = 32768;   // longest merged run (keeps the u32 sum exact for u16 metrics)

struct AccumParams {
    const uint32_t* ts;
    const uint32_t* gid;
    const uint32_t* pid;
    const uint32_t* util;
    uint64_t n;              // samples in this batch
    uint64_t vals_base;      // offset of this batch in the value log (multiple of 16)
    uint32_t tiles_per_warp;
    const JoinSlot* slots;   // join table (nullptr when the pod kind is off)
    uint32_t tmask;
    uint32_t t0;
    FastDiv dstep;
    uint32_t n_panes;
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
    uint32_t* ctl;                 // ctl[0] run-log cursor, ctl[1] overflow flag
};

struct RunQueue {   // structure of arrays: lane l touching record l is conflict free
    uint32_t gid[kQueueCap], pid[kQueueCap], pane[kQueueCap], cnt[kQueueCap], sum[kQueueCap], mx[kQueueCap],
        start[kQueueCap];
};

// Retire one run: probe the join table once, then one RED per statistic and key kind and one
// run-log descriptor per key kind (slots di .. di+nk-1).
template <bool PCT>
__device__ __forceinline__ void emit_run(const AccumParams& P, uint32_t g, uint32_t pid, uint32_t pane,
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
            atomicAdd(P.cnt[kind] + cell, c);
            atomicAdd(P.sum[kind] + cell, (unsigned long long)s);
            atomicMax(P.mx[kind] + cell, m);
        }
        if (PCT && log_ok) {
            uint32_t next = 0;
            if (valid) next = atomicExch(P.head[kind] + cell, di + 1);
            P.desc[di] = make_uint4(start, c, next, cell);
            ++di;
        }
    }
}

// Reserve `count` run-log entries for the warp with one atomic; returns the first index.
__device__ __forceinline__ uint32_t reserve_log(const AccumParams& P, uint32_t count, bool& log_ok) {
    uint32_t base = 0;
    if (lane_id() == 0 && count) base = atomicAdd(P.ctl, count);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    log_ok = (uint64_t)base + count <= P.desc_cap;
    if (!log_ok && lane_id() == 0 && count) P.ctl[1] = 1;
    return base;
}

// Retire records [0, n_flush) of the warp's queue, one record per lane.
template <bool PCT>
__device__ __forceinline__ void flush_queue(const AccumParams& P, const RunQueue& q, uint32_t n_flush, uint32_t nk) {
    const uint32_t lane = lane_id();
    for (uint32_t base = 0; base < n_flush; base += 32) {
        const uint32_t here = min(32u, n_flush - base);
        bool log_ok = false;
        uint32_t di = 0;
        if (PCT) di = reserve_log(P, here * nk, log_ok) + lane * nk;
        const uint32_t r = base + lane;
        if (lane < here)
            emit_run<PCT>(P, q.gid[r], q.pid[r], q.pane[r], q.cnt[r], q.sum[r], q.mx[r], q.start[r], di, log_ok);
    }
}

struct Quad {   // one lane's four consecutive samples
    uint32_t t[4], g[4], p[4], u[4];
};

__device__ __forceinline__ void load_quad(const AccumParams& P, uint64_t i0, Quad& q) {
    if (i0 + 3 < P.n) {
        const uint4 a = ld_stream_v4(P.ts + i0), b = ld_stream_v4(P.gid + i0);
        const uint4 c = ld_stream_v4(P.pid + i0), d = ld_stream_v4(P.util + i0);
        q.t[0] = a.x; q.t[1] = a.y; q.t[2] = a.z; q.t[3] = a.w;
        q.g[0] = b.x; q.g[1] = b.y; q.g[2] = b.z; q.g[3] = b.w;
        q.p[0] = c.x; q.p[1] = c.y; q.p[2] = c.z; q.p[3] = c.w;
        q.u[0] = d.x; q.u[1] = d.y; q.u[2] = d.z; q.u[3] = d.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool in = i0 + j < P.n;
            q.t[j] = in ? P.ts[i0 + j] : 0u;
            q.g[j] = in ? P.gid[i0 + j] : 0u;
            q.p[j] = in ? P.pid[i0 + j] : 0u;
            q.u[j] = in ? P.util[i0 + j] : 0u;
        }
    }
}

template <bool PCT>
__global__ void __launch_bounds__(kAccumWarps * 32) k_accum(const __grid_constant__ AccumParams P) {
    __shared__ RunQueue queues[kAccumWarps];
    const uint32_t lane = lane_id();
    RunQueue& Q = queues[threadIdx.x >> 5];
    const uint32_t nk = __popc(P.key_mask);
    const uint64_t n_tiles = (P.n + 127) >> 7;
    const uint64_t warp = (uint64_t)blockIdx.x * kAccumWarps + (threadIdx.x >> 5);
    const uint64_t tile_begin = warp * P.tiles_per_warp;
    if (tile_begin >= n_tiles) return;
    const uint64_t tile_end = min(n_tiles, tile_begin + P.tiles_per_warp);
    uint32_t queued = 0;   // warp-uniform

    Quad cur, nxt;
    load_quad(P, (tile_begin << 7) + 4 * lane, cur);
    for (uint64_t tile = tile_begin; tile < tile_end; ++tile) {
        const uint64_t i0 = (tile << 7) + 4 * lane;
        if (tile + 1 < tile_end) load_quad(P, i0 + 128, nxt);   // in flight while this tile is processed
        const bool full = i0 + 3 < P.n;
        uint32_t pane[4], v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t x = cur.t[j] - P.t0;
            const uint32_t pn = fdiv(x, P.dstep);
            const bool ok = (cur.t[j] >= P.t0) && (pn < P.n_panes) && (full || i0 + j < P.n);
            pane[j] = ok ? pn : kInvalid;
            v[j] = (cur.u[j] >> P.mshift) & P.mmask;
        }
        if (PCT) {   // value log, same order as the input (u8 metrics)
            uint8_t* dst = P.vals + P.vals_base + i0;
            if (full) {
                *reinterpret_cast<uint32_t*>(dst) = v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (i0 + j < P.n) dst[j] = (uint8_t)v[j];
            }
        }
        const uint32_t* g = cur.g;
        const uint32_t* p = cur.p;
        // is this lane's quad one valid run?
        const bool uni = (pane[0] != kInvalid) && g[0] == g[1] && g[1] == g[2] && g[2] == g[3] &&
                         p[0] == p[1] && p[1] == p[2] && p[2] == p[3] && pane[0] == pane[1] &&
                         pane[1] == pane[2] && pane[2] == pane[3];
        // segment heads among uniform lanes
        const uint32_t pg = __shfl_up_sync(0xFFFFFFFFu, g[0], 1);
        const uint32_t pp = __shfl_up_sync(0xFFFFFFFFu, p[0], 1);
        const uint32_t ppn = __shfl_up_sync(0xFFFFFFFFu, pane[0], 1);
        const uint32_t unis = __ballot_sync(0xFFFFFFFFu, uni);
        const bool prev_uni = lane > 0 && ((unis >> (lane - 1)) & 1u);
        const bool head = uni && (!prev_uni || pg != g[0] || pp != p[0] || ppn != pane[0]);
        const uint32_t heads = __ballot_sync(0xFFFFFFFFu, head);
        const uint32_t breakers = heads | ~unis;
        const uint32_t rest = (lane == 31) ? 0u : (breakers & (0xFFFFFFFEu << lane));
        const uint32_t seg_end = rest ? (uint32_t)(__ffs(rest) - 2) : 31u;
        uint32_t s = 0, m = 0;
        if (uni) {
            s = v[0] + v[1] + v[2] + v[3];
            m = max(max(v[0], v[1]), max(v[2], v[3]));
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t s2 = __shfl_down_sync(0xFFFFFFFFu, s, d);
            const uint32_t m2 = __shfl_down_sync(0xFFFFFFFFu, m, d);
            if (lane + d <= seg_end) {
                s += s2;
                m = max(m, m2);
            }
        }
        const uint32_t c = 4 * (seg_end - lane + 1);
        const uint32_t start = (uint32_t)(P.vals_base + i0);
        // does the run at the head of this tile continue the last queued record?
        bool merged = false;
        if (lane == 0 && head && queued) {
            const uint32_t r = queued - 1;
            if (Q.gid[r] == g[0] && Q.pid[r] == p[0] && Q.pane[r] == pane[0] && Q.start[r] + Q.cnt[r] == start &&
                Q.cnt[r] + c <= kMergeMax) {
                Q.cnt[r] += c;
                Q.sum[r] += s;
                Q.mx[r] = max(Q.mx[r], m);
                merged = true;
            }
        }
        merged = __shfl_sync(0xFFFFFFFFu, merged, 0);
        if (head && !(lane == 0 && merged)) {
            const uint32_t r = queued + __popc(heads & ((1u << lane) - 1u)) - (merged ? 1u : 0u);
            Q.gid[r] = g[0]; Q.pid[r] = p[0]; Q.pane[r] = pane[0];
            Q.cnt[r] = c; Q.sum[r] = s; Q.mx[r] = m; Q.start[r] = start;
        }
        queued += __popc(heads) - (merged ? 1u : 0u);
        // ragged quads (a run boundary or an out-of-range sample inside the quad): each lane
        // retires its own pieces directly; rare in collector order
        if (unis != 0xFFFFFFFFu) {
            uint32_t nr = 0;
            if (!uni) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const bool first = (j == 0) || g[j] != g[j - 1] || p[j] != p[j - 1] || pane[j] != pane[j - 1];
                    nr += (first && pane[j] != kInvalid) ? 1u : 0u;
                }
            }
            if (__any_sync(0xFFFFFFFFu, nr != 0)) {
                uint32_t di = 0;
                bool log_ok = false;
                if (PCT) {
                    uint32_t incl = nr;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                        if (lane >= (uint32_t)d) incl += o;
                    }
                    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31) * nk;
                    di = reserve_log(P, total, log_ok) + (incl - nr) * nk;
                }
                if (nr) {
                    uint32_t rs = 0, rm = 0, rc = 0;
                    int rstart = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const bool first = (j == 0) || g[j] != g[j - 1] || p[j] != p[j - 1] || pane[j] != pane[j - 1];
                        if (first) {
                            rs = 0; rm = 0; rc = 0; rstart = j;
                        }
                        rs += v[j];
                        rm = max(rm, v[j]);
                        rc += 1;
                        const bool last = (j == 3) || g[j + 1] != g[j] || p[j + 1] != p[j] || pane[j + 1] != pane[j];
                        if (last && pane[j] != kInvalid) {
                            emit_run<PCT>(P, g[j], p[j], pane[j], rc, rs, rm, start + rstart, di, log_ok);
                            di += nk;
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (queued > 32) {   // retire all but the newest record (it may still grow)
            flush_queue<PCT>(P, Q, queued - 1, nk);
            __syncwarp();
            if (lane == 0) {
                const uint32_t r = queued - 1;
                Q.gid[0] = Q.gid[r]; Q.pid[0] = Q.pid[r]; Q.pane[0] = Q.pane[r];
                Q.cnt[0] = Q.cnt[r]; Q.sum[0] = Q.sum[r]; Q.mx[0] = Q.mx[r]; Q.start[0] = Q.start[r];
            }
            queued = 1;
            __syncwarp();
        }
        cur = nxt;
    }
    flush_queue<PCT>(P, Q, queued, nk);
}

}  // namespace kgs

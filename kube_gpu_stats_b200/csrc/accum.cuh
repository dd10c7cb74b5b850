// accum.cuh - the hot kernel: one streaming pass over the 16 B/sample columns that does the
// (node,pid)->pod join and folds every sample into its (key, pane) cells for up to three key
// kinds at once.
//
// Takes over the per-sample loop of the reference (gpu_util_stats.py:72-78: decode, append,
// sum/len) and the dict probes of its join (gpu_util_stats.py:86-91).  Design:
//   * a warp reads a tile of 128 consecutive samples, four per lane, with one LDG.128 per
//     column (algorithmic traffic: 16 B/sample, read once);
//   * consecutive samples that share (gpu_gid, pid, pane) form a RUN; in collector order
//     (sorted by gpu, then time) runs are tens to thousands of samples long.  Lanes whose four
//     samples are one run are combined with a segmented warp-shuffle reduction, so a run costs
//     one hash probe and one RED per statistic per key kind, not one per sample;
//   * sums are integers (u8/u16 metrics): RED.ADD.U64 / RED.MAX are order independent, so the
//     result is bit-exact whatever the scheduling;
//   * for percentiles the metric byte of every sample is appended to a value log (1 B/sample,
//     same order as the input) and each run leaves a 16 B descriptor {start, len, next, cell}
//     linked into its cell's list; the finalize kernel selects from those.
// Any input order is handled correctly; run length only affects speed.
#pragma once
#include "common.cuh"

namespace kgs {

struct AccumParams {
    const uint32_t* ts;
    const uint32_t* gid;
    const uint32_t* pid;
    const uint32_t* util;
    uint64_t n;              // samples in this batch
    uint64_t vals_base;      // offset of this batch in the value log (multiple of 16)
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

template <bool PCT>
__device__ __forceinline__ void emit_run(const AccumParams& P, uint32_t g, uint32_t pid, uint32_t pane,
                                         uint32_t c, uint32_t s, uint32_t m, uint64_t start, uint32_t di,
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
            P.desc[di] = make_uint4((uint32_t)start, c, next, valid ? (cell | 0u) : kInvalid);
            // high bits of start (logs beyond 4 Gi samples are rejected on the host)
            ++di;
        }
    }
}

template <bool PCT>
__global__ void __launch_bounds__(256) k_accum(const __grid_constant__ AccumParams P) {
    const uint32_t lane = lane_id();
    const uint32_t nk = __popc(P.key_mask);
    const uint64_t n_tiles = (P.n + 127) >> 7;
    const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
    for (uint64_t tile = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); tile < n_tiles;
         tile += warps) {
        const uint64_t i0 = (tile << 7) + 4 * lane;
        uint32_t t[4], g[4], p[4], u[4];
        const bool full = i0 + 3 < P.n;
        if (full) {
            const uint4 a = ld_stream_v4(P.ts + i0), b = ld_stream_v4(P.gid + i0);
            const uint4 c = ld_stream_v4(P.pid + i0), d = ld_stream_v4(P.util + i0);
            t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w;
            g[0] = b.x; g[1] = b.y; g[2] = b.z; g[3] = b.w;
            p[0] = c.x; p[1] = c.y; p[2] = c.z; p[3] = c.w;
            u[0] = d.x; u[1] = d.y; u[2] = d.z; u[3] = d.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool in = i0 + j < P.n;
                t[j] = in ? P.ts[i0 + j] : 0u;
                g[j] = in ? P.gid[i0 + j] : 0u;
                p[j] = in ? P.pid[i0 + j] : 0u;
                u[j] = in ? P.util[i0 + j] : 0u;
            }
        }
        uint32_t pane[4], v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t x = t[j] - P.t0;
            const uint32_t pn = fdiv(x, P.dstep);
            const bool ok = (t[j] >= P.t0) && (pn < P.n_panes) && (full || i0 + j < P.n);
            pane[j] = ok ? pn : kInvalid;
            v[j] = (u[j] >> P.mshift) & P.mmask;
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
        // is this lane's quad one valid run?
        const bool uni = (pane[0] != kInvalid) && g[0] == g[1] && g[1] == g[2] && g[2] == g[3] &&
                         p[0] == p[1] && p[1] == p[2] && p[2] == p[3] && pane[0] == pane[1] &&
                         pane[1] == pane[2] && pane[2] == pane[3];
        // runs this lane will emit by itself (non-uniform quad): count the valid local runs
        uint32_t nr = 0;
        if (!uni) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool first = (j == 0) || g[j] != g[j - 1] || p[j] != p[j - 1] || pane[j] != pane[j - 1];
                nr += (first && pane[j] != kInvalid) ? 1u : 0u;
            }
        }
        // segment heads among uniform lanes
        const uint32_t pg = __shfl_up_sync(0xFFFFFFFFu, g[0], 1);
        const uint32_t pp = __shfl_up_sync(0xFFFFFFFFu, p[0], 1);
        const uint32_t ppn = __shfl_up_sync(0xFFFFFFFFu, pane[0], 1);
        const uint32_t unis = __ballot_sync(0xFFFFFFFFu, uni);
        const bool prev_uni = lane > 0 && ((unis >> (lane - 1)) & 1u);
        const bool head = uni && (!prev_uni || pg != g[0] || pp != p[0] || ppn != pane[0]);
        const uint32_t heads = __ballot_sync(0xFFFFFFFFu, head);
        if (unis == 0 && __all_sync(0xFFFFFFFFu, nr == 0)) continue;   // nothing in range in this tile
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
        if (head) nr = 1;
        // reserve run-log entries for the whole warp with one atomic
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
            uint32_t base = 0;
            if (lane == 0 && total) base = atomicAdd(P.ctl, total);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            log_ok = (uint64_t)base + total <= P.desc_cap;
            if (!log_ok && lane == 0 && total) P.ctl[1] = 1;
            di = base + (incl - nr) * nk;
        }
        if (head) {
            emit_run<PCT>(P, g[0], p[0], pane[0], 4 * (seg_end - lane + 1), s, m, P.vals_base + i0, di, log_ok);
        } else if (nr) {
            // ragged quad: emit each local run on its own (rare in collector order)
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
                    emit_run<PCT>(P, g[j], p[j], pane[j], rc, rs, rm, P.vals_base + i0 + rstart, di, log_ok);
                    di += nk;
                }
            }
        }
    }
}

}  // namespace kgs

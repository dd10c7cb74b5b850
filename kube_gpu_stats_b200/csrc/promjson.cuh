// promjson.cuh - the reference's value decode (gpu_util_stats.py:25, :47-48, :73-74) on the GPU:
// from the BYTES of a Prometheus /query_range answer
//   {"status":"success","data":{"resultType":"matrix","result":[{"metric":{...},"values":[[ts,"v"],...]},...]}}
// to the CSR the statistics kernels reduce (fp64 values + per-series offsets), without building a
// Python object per sample.  SURVEY.md section 8 row a1 / (f)1: this is where the reference's time goes.
//
// A thread owns kMxChunk consecutive bytes.  What a chunk means depends on what came before it
// (inside a string or not, how deep in brackets), so the work is three walks over the text:
//   k_mx_classify  per chunk: the TRANSITION (quote parity flip, bracket-depth change under both
//                  "starts outside a string" and "starts inside"), combined per CTA;
//   k_mx_top       one CTA: exclusive scan of the CTA transitions (the operator is associative);
//   k_mx_count     per chunk: its real start state, then the number of samples and series that
//                  open in it; CTA sums;
//   k_mx_top2      one CTA: exclusive scan of the CTA counts, totals;
//   k_mx_emit      per chunk: byte position of every sample, per series its first sample index and
//                  the byte spans the host needs to decode the (small) label part itself;
//   k_mx_parse     per sample: ts and value text -> fp64 (strtod.cuh, bit-identical to float()).
// A '"' or bracket preceded by an odd run of backslashes is escaped; label values may hold any text.
// Depths (counting '{' and '['): a series object opens at depth 3, its "values" array at 4, a
// [ts,"v"] pair at 5; for an instant-query "vector" the pair itself opens at depth 4.
#pragma once
#include "common.cuh"
#include "strtod.cuh"
#include "promjson_bits.cuh"

namespace kgs {

constexpr int kMxThreads = 256;
constexpr uint32_t kMxChunk = 64;
constexpr uint32_t kMxTile = kMxThreads * kMxChunk;   // 16 KB of text per CTA
constexpr uint32_t kMxNoSpan = 0xFFFFFFFFu;

struct MxTrans {   // effect of a stretch of text on (in_string, depth)
    uint32_t flip; // quote parity
    int32_t d0;    // depth change if the stretch starts outside a string
    int32_t d1;    // ... inside a string
};

__device__ __forceinline__ MxTrans mx_then(const MxTrans& a, const MxTrans& b) {
    MxTrans r;
    r.flip = a.flip ^ b.flip;
    r.d0 = a.d0 + (a.flip ? b.d1 : b.d0);
    r.d1 = a.d1 + (a.flip ? b.d0 : b.d1);
    return r;
}

__device__ __forceinline__ MxTrans mx_shfl_up(const MxTrans& v, int d) {
    MxTrans r;
    r.flip = __shfl_up_sync(0xFFFFFFFFu, v.flip, d);
    r.d0 = __shfl_up_sync(0xFFFFFFFFu, v.d0, d);
    r.d1 = __shfl_up_sync(0xFFFFFFFFu, v.d1, d);
    return r;
}

// In-order scan over the CTA's threads: exclusive prefix for each thread, total for all.
__device__ __forceinline__ MxTrans mx_block_scan(MxTrans mine, MxTrans* warp_tot, MxTrans& total) {
    const uint32_t lane = lane_id(), w = threadIdx.x >> 5;
    MxTrans incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const MxTrans o = mx_shfl_up(incl, d);
        if (lane >= (uint32_t)d) incl = mx_then(o, incl);
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    MxTrans pre{0, 0, 0};
    total = pre;
    for (int k = 0; k < kMxThreads / 32; ++k) {
        if ((uint32_t)k == w) pre = total;
        total = mx_then(total, warp_tot[k]);
    }
    MxTrans before = mx_shfl_up(incl, 1);
    if (lane == 0) before = MxTrans{0, 0, 0};
    __syncthreads();
    return mx_then(pre, before);
}

struct MxParams {
    const char* text;          // padded with zero bytes to a whole number of tiles plus one chunk
    unsigned long long n;      // bytes of text
    uint32_t n_ctas;
    uint32_t cta_first, cta_count;   // the piece of the text this launch works on (pieces arrive one by one)
    uint32_t last_piece;
    uint32_t k_first, k_end;   // samples of the piece (k_mx_parse)
    uint32_t pair_depth;       // 5: matrix, 4: vector
    MxTrans* carry;            // transition of all text before the piece
    MxTrans* cta_trans;        // [n_ctas]
    uint2* cta_counts;         // [n_ctas] (samples, series)
    uint32_t* chunk_state;     // [chunks] depth << 1 | in_string at the chunk's first byte
    uint2* chunk_counts;       // [chunks]
    uint32_t* sample_pos;      // [n_samples] byte index of the '[' of each pair
    unsigned long long* offsets;   // [n_series + 1]
    unsigned long long* spans;     // [n_series][4]: '{' of the series, '[' and ']' of its values array, '}'
    double* ts;
    double* val;
    unsigned long long* ctl;   // [0] samples [1] series [2] first bad byte (~0: none) [3] long series [4] end state
};

// Is the byte at `pos` escaped (odd run of backslashes right before it)?
__device__ __forceinline__ bool mx_escaped_at(const char* text, unsigned long long pos) {
    uint32_t run = 0;
    while (pos > 0 && text[pos - 1] == '\\') {
        ++run;
        --pos;
    }
    return run & 1u;
}

__device__ __forceinline__ void mx_load(const MxParams& P, unsigned long long base, uint4 (&t)[4]) {
    const uint4* p = reinterpret_cast<const uint4*>(P.text + base);
#pragma unroll
    for (int v = 0; v < 4; ++v) t[v] = __ldg(p + v);
}

// The chunk as masks (promjson_bits.cuh): 16 words -> where the quotes, escapes and brackets are.
__device__ __forceinline__ MxLex mx_lex_chunk(const uint4 (&t)[4], bool first_escaped) {
    const uint32_t w[16] = {t[0].x, t[0].y, t[0].z, t[0].w, t[1].x, t[1].y, t[1].z, t[1].w,
                            t[2].x, t[2].y, t[2].z, t[2].w, t[3].x, t[3].y, t[3].z, t[3].w};
    return mx_lex(mx_masks(w), first_escaped);
}

__device__ __forceinline__ MxTrans mx_chunk_trans(const MxLex& x) {
    int d0, d1;
    mx_depth_change(x, d0, d1);
    return MxTrans{x.flip, d0, d1};
}

__global__ void __launch_bounds__(kMxThreads) k_mx_classify(const __grid_constant__ MxParams P) {
    __shared__ MxTrans warp_tot[kMxThreads / 32];
    const uint32_t cta = P.cta_first + blockIdx.x;
    const unsigned long long base = ((unsigned long long)cta * kMxThreads + threadIdx.x) * kMxChunk;
    uint4 t[4];
    mx_load(P, base, t);
    const MxTrans mine = mx_chunk_trans(mx_lex_chunk(t, mx_escaped_at(P.text, base)));
    MxTrans total;
    mx_block_scan(mine, warp_tot, total);
    if (threadIdx.x == 0) P.cta_trans[cta] = total;
}

// Exclusive scan of the piece's CTA transitions behind the carry of the text before it, one CTA;
// records the state at the end of the text seen so far.
__global__ void __launch_bounds__(kMxThreads) k_mx_top(const __grid_constant__ MxParams P) {
    __shared__ MxTrans warp_tot[kMxThreads / 32];
    const uint32_t end = P.cta_first + P.cta_count;
    const uint32_t per = (P.cta_count + kMxThreads - 1) / kMxThreads;
    const uint32_t a = min(end, P.cta_first + threadIdx.x * per), b = min(end, a + per);
    MxTrans seg{0, 0, 0};
    for (uint32_t i = a; i < b; ++i) seg = mx_then(seg, P.cta_trans[i]);
    const MxTrans carry = *P.carry;
    MxTrans total;
    MxTrans run = mx_then(carry, mx_block_scan(seg, warp_tot, total));
    for (uint32_t i = a; i < b; ++i) {
        const MxTrans x = P.cta_trans[i];
        P.cta_trans[i] = run;
        run = mx_then(run, x);
    }
    __syncthreads();   // every thread has read the carry
    if (threadIdx.x == 0) {
        total = mx_then(carry, total);
        *P.carry = total;
        P.ctl[4] = ((unsigned long long)(uint32_t)total.d0 << 1) | total.flip;   // 0 for balanced text
    }
}

__global__ void __launch_bounds__(kMxThreads) k_mx_count(const __grid_constant__ MxParams P) {
    __shared__ MxTrans warp_tot[kMxThreads / 32];
    __shared__ uint32_t wsum[2][kMxThreads / 32];
    const uint32_t cta = P.cta_first + blockIdx.x;
    const uint32_t chunk = cta * kMxThreads + threadIdx.x;
    const unsigned long long base = (unsigned long long)chunk * kMxChunk;
    uint4 t[4];
    mx_load(P, base, t);
    const MxLex lex = mx_lex_chunk(t, mx_escaped_at(P.text, base));
    MxTrans total;
    const MxTrans before = mx_then(P.cta_trans[cta], mx_block_scan(mx_chunk_trans(lex), warp_tot, total));
    const uint32_t in_string = before.flip;              // the text starts outside a string at depth 0
    const int32_t depth = before.d0;
    uint32_t n_samples = 0, n_series = 0;
    mx_brackets(lex, in_string, depth, (int)P.pair_depth, [&](int kind, uint32_t) {
        n_samples += kind == 0;
        n_series += kind == 1;
    });
    P.chunk_state[chunk] = ((uint32_t)depth << 1) | in_string;
    P.chunk_counts[chunk] = make_uint2(n_samples, n_series);
    const uint32_t s0 = __reduce_add_sync(0xFFFFFFFFu, n_samples), s1 = __reduce_add_sync(0xFFFFFFFFu, n_series);
    if (lane_id() == 0) {
        wsum[0][threadIdx.x >> 5] = s0;
        wsum[1][threadIdx.x >> 5] = s1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0, b = 0;
        for (int k = 0; k < kMxThreads / 32; ++k) {
            a += wsum[0][k];
            b += wsum[1][k];
        }
        P.cta_counts[cta] = make_uint2(a, b);
    }
}

__global__ void __launch_bounds__(kMxThreads) k_mx_top2(const __grid_constant__ MxParams P) {
    __shared__ unsigned long long wtot[2][kMxThreads / 32];
    const uint32_t end = P.cta_first + P.cta_count;
    const uint32_t per = (P.cta_count + kMxThreads - 1) / kMxThreads;
    const uint32_t a = min(end, P.cta_first + threadIdx.x * per), b = min(end, a + per);
    const unsigned long long c0 = P.ctl[0], c1 = P.ctl[1];   // samples and series before this piece
    unsigned long long s0 = 0, s1 = 0;
    for (uint32_t i = a; i < b; ++i) {
        const uint2 c = P.cta_counts[i];
        s0 += c.x;
        s1 += c.y;
    }
    unsigned long long i0 = s0, i1 = s1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o0 = __shfl_up_sync(0xFFFFFFFFu, i0, d), o1 = __shfl_up_sync(0xFFFFFFFFu, i1, d);
        if (lane_id() >= (uint32_t)d) {
            i0 += o0;
            i1 += o1;
        }
    }
    const uint32_t w = threadIdx.x >> 5;
    if (lane_id() == 31) {
        wtot[0][w] = i0;
        wtot[1][w] = i1;
    }
    __syncthreads();
    unsigned long long p0 = 0, p1 = 0, t0 = 0, t1 = 0;
    for (uint32_t k = 0; k < (uint32_t)kMxThreads / 32; ++k) {
        if (k == w) {
            p0 = t0;
            p1 = t1;
        }
        t0 += wtot[0][k];
        t1 += wtot[1][k];
    }
    unsigned long long r0 = c0 + p0 + i0 - s0, r1 = c1 + p1 + i1 - s1;
    // sample indices are 32-bit in the per-CTA table: the host refuses answers past 2^32 - 1 samples
    for (uint32_t i = a; i < b; ++i) {
        const uint2 c = P.cta_counts[i];
        P.cta_counts[i] = make_uint2((uint32_t)r0, (uint32_t)r1);
        r0 += c.x;
        r1 += c.y;
    }
    __syncthreads();   // every thread has read the running totals
    if (threadIdx.x == 0) {
        P.ctl[0] = c0 + t0;
        P.ctl[1] = c1 + t1;
    }
}

__global__ void __launch_bounds__(kMxThreads) k_mx_emit(const __grid_constant__ MxParams P) {
    __shared__ uint32_t wtot[2][kMxThreads / 32];
    const uint32_t cta_id = P.cta_first + blockIdx.x;
    const uint32_t chunk = cta_id * kMxThreads + threadIdx.x;
    const unsigned long long base = (unsigned long long)chunk * kMxChunk;
    const uint2 mine = P.chunk_counts[chunk];
    uint32_t i0 = mine.x, i1 = mine.y;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o0 = __shfl_up_sync(0xFFFFFFFFu, i0, d), o1 = __shfl_up_sync(0xFFFFFFFFu, i1, d);
        if (lane_id() >= (uint32_t)d) {
            i0 += o0;
            i1 += o1;
        }
    }
    const uint32_t w = threadIdx.x >> 5;
    if (lane_id() == 31) {
        wtot[0][w] = i0;
        wtot[1][w] = i1;
    }
    __syncthreads();
    const uint2 cta = P.cta_counts[cta_id];
    uint32_t k = cta.x + i0 - mine.x, s = cta.y + i1 - mine.y;
    for (uint32_t q = 0; q < w; ++q) {
        k += wtot[0][q];
        s += wtot[1][q];
    }
    if (P.last_piece && blockIdx.x == 0 && threadIdx.x == 0) P.offsets[P.ctl[1]] = P.ctl[0];   // closing entry of the CSR
    if ((mine.x | mine.y) == 0) {
        // nothing opens here, but a values array or a series may close here
        const uint32_t st = P.chunk_state[chunk];
        if ((st >> 1) < 4 || (st >> 1) > 6) return;
    }
    uint4 t[4];
    mx_load(P, base, t);
    const uint32_t st = P.chunk_state[chunk];
    const MxLex lex = mx_lex_chunk(t, mx_escaped_at(P.text, base));
    mx_brackets(lex, st & 1u, (int)(st >> 1), (int)P.pair_depth, [&](int kind, uint32_t i) {
        if (kind == 0) {
            P.sample_pos[k++] = (uint32_t)(base + i);
        } else if (kind == 1) {
            P.offsets[s] = k;
            P.spans[4ull * s] = base + i;
            ++s;
        } else if (s) {                                  // 2: '[' of values, 3: its ']', 4: '}' of the series
            P.spans[4ull * (s - 1) + (uint32_t)(kind - 1)] = base + i;
        }
    });
}

__device__ __forceinline__ bool mx_blank(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

// One thread per sample: '[' blanks number blanks ',' blanks ("text" | number) blanks ']'.
__global__ void __launch_bounds__(128) k_mx_parse(const __grid_constant__ MxParams P) {
    const unsigned long long k = (unsigned long long)P.k_first + (unsigned long long)blockIdx.x * 128 + threadIdx.x;
    if (k >= P.k_end) return;
    const char* s = P.text;
    const unsigned long long open = P.sample_pos[k], n = P.n;
    double ts = 0.0, v = 0.0;
    {   // the everyday form [digits[.digits],"digits[.digits]"] with short numbers: no grammar, one division each
        // (the text is padded with zero bytes, so the digit loops stop inside the buffer)
        const char* p = s + open + 1;
        const uint32_t a = parse_short_decimal(p, &ts);
        if (a && p[a] == ',' && p[a + 1] == '"' && json_number_ok(p, a)) {   // the timestamp is a bare JSON number
            const char* q = p + a + 2;
            const uint32_t b = parse_short_decimal(q, &v);
            if (b && q[b] == '"' && q[b + 1] == ']') {
                P.ts[k] = ts;
                P.val[k] = v;
                return;
            }
        }
    }
    unsigned long long i = open + 1;
    bool ok = true;
    while (i < n && mx_blank(s[i])) ++i;
    unsigned long long a = i;
    while (i < n && s[i] != ',' && s[i] != ']' && !mx_blank(s[i])) ++i;
    ok = ok && i > a && i - a < 4096 && json_number_ok(s + a, (uint32_t)(i - a)) && parse_float(s + a, (uint32_t)(i - a), &ts);
    while (i < n && mx_blank(s[i])) ++i;
    ok = ok && i < n && s[i] == ',';
    ++i;
    while (i < n && mx_blank(s[i])) ++i;
    if (i < n && s[i] == '"') {
        a = ++i;
        while (i < n && s[i] != '"' && s[i] != '\\') ++i;     // an escape inside a sample value is refused
        ok = ok && i < n && s[i] == '"' && i - a < (1u << 20) && parse_float(s + a, (uint32_t)(i - a), &v);
        ++i;
    } else {
        a = i;
        while (i < n && s[i] != ',' && s[i] != ']' && !mx_blank(s[i])) ++i;
        ok = ok && i > a && i - a < 4096 && json_number_ok(s + a, (uint32_t)(i - a)) && parse_float(s + a, (uint32_t)(i - a), &v);
    }
    while (i < n && mx_blank(s[i])) ++i;
    ok = ok && i < n && s[i] == ']';
    P.ts[k] = ts;
    P.val[k] = v;
    if (!ok) atomicMin(P.ctl + 2, open);
}

// Series longer than kSeqMax go to the warp-per-series mean kernel: list them.
__global__ void k_mx_long_series(const unsigned long long* __restrict__ offsets, uint32_t n_series, uint32_t limit,
                                 uint32_t* long_ids, unsigned long long* ctl) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_series) return;
    if (offsets[s + 1] - offsets[s] > limit) long_ids[atomicAdd(ctl + 3, 1ull)] = s;
}

}  // namespace kgs

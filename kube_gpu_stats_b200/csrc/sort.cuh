// sort.cuh - kgs_reorder_batch: bring a resident batch into series-major collector order on the GPU.
//
// The accumulate kernel is exact for any sample order but fast when consecutive samples share
// (gpu, pid): a time-major live feed, or several processes per GPU delivered second by second,
// interleave the series and every sample becomes its own run (25x slower end to end).
// This is the ingestion step for such feeds: a STABLE least-significant-digit radix sort of the
// four columns on the key (gpu_gid, pid), 8 bits per pass over just the bits the batch uses.
// Stability keeps each series in its input (time) order, so the result is (gpu, pid, time).
// No reference counterpart (the reference never sees per-process samples).
//
// One pass = k_sort_hist (per-CTA digit counts) -> exclusive scan of the [digit][cta] count matrix
// (k_scan_reduce / k_scan_top / k_scan_apply) -> k_sort_scatter.  A CTA owns kSortBlock consecutive
// elements, a warp a contiguous eighth of them, so the order "cta, warp, round, lane" IS the input
// order and a rank built from warp-level __match_any_sync groups is stable.
#pragma once
#include "common.cuh"

namespace kgs {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortRounds = 32;                                   // rounds of 32 elements per warp
constexpr uint32_t kSortWarpSpan = 32 * kSortRounds;              // 1024 elements per warp
constexpr uint32_t kSortBlock = kSortWarps * kSortWarpSpan;       // 8192 elements per CTA
constexpr uint32_t kScanBlock = 4096;                             // elements per CTA of the scan kernels

struct SortParams {
    const uint32_t* in[4];     // ts, gid, pid, util
    uint32_t* out[4];
    uint32_t n, n_ctas;
    uint32_t pid_bits, shift;  // key = gid << pid_bits | pid; this pass sorts on bits [shift, shift + 8)
    uint32_t* counts;          // [256][n_ctas]; after the scan: first output index of (digit, cta)
};

__device__ __forceinline__ uint32_t sort_digit(const SortParams& P, uint32_t i) {
    const unsigned long long key = ((unsigned long long)P.in[1][i] << P.pid_bits) | P.in[2][i];
    return (uint32_t)(key >> P.shift) & 0xFFu;
}

// max gpu_gid and max pid of the batch (decides how many bits the key has)
__global__ void k_sort_maxima(const uint32_t* __restrict__ gid, const uint32_t* __restrict__ pid, uint32_t n,
                              uint32_t* out) {
    uint32_t mg = 0, mp = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        mg = max(mg, gid[i]);
        mp = max(mp, pid[i]);
    }
    mg = __reduce_max_sync(0xFFFFFFFFu, mg);
    mp = __reduce_max_sync(0xFFFFFFFFu, mp);
    if (lane_id() == 0) {
        atomicMax(out, mg);
        atomicMax(out + 1, mp);
    }
}

__global__ void __launch_bounds__(kSortThreads) k_sort_hist(const __grid_constant__ SortParams P) {
    __shared__ uint32_t hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t begin = blockIdx.x * kSortBlock, end = min(P.n, begin + kSortBlock);
    for (uint32_t i = begin + threadIdx.x; i < end; i += kSortThreads) atomicAdd(&hist[sort_digit(P, i)], 1u);
    __syncthreads();
    P.counts[threadIdx.x * P.n_ctas + blockIdx.x] = hist[threadIdx.x];
}

// ---- exclusive scan of a u32 array (three small kernels) ------------------------------------------
__global__ void __launch_bounds__(256) k_scan_reduce(const uint32_t* __restrict__ a, uint32_t n, uint32_t* partial) {
    __shared__ uint32_t warp_sum[8];
    const uint32_t begin = blockIdx.x * kScanBlock;
    uint32_t s = 0;
    for (uint32_t i = begin + threadIdx.x; i < min(n, begin + kScanBlock); i += 256) s += a[i];
    s = __reduce_add_sync(0xFFFFFFFFu, s);
    if (lane_id() == 0) warp_sum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; ++w) t += warp_sum[w];
        partial[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024) k_scan_top(uint32_t* partial, uint32_t m) {   // m <= 64 Ki partials, one CTA
    __shared__ uint32_t carry;
    __shared__ uint32_t wsum[32];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < m; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < m ? partial[i] : 0;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane_id() >= (uint32_t)d) incl += o;
        }
        if (lane_id() == 31) wsum[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = wsum[threadIdx.x], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, wi, d);
                if (lane_id() >= (uint32_t)d) wi += o;
            }
            wsum[threadIdx.x] = wi - w;   // exclusive prefix of the warp sums
        }
        __syncthreads();
        const uint32_t excl = carry + wsum[threadIdx.x >> 5] + incl - v;
        if (i < m) partial[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_scan_apply(uint32_t* a, uint32_t n, const uint32_t* __restrict__ partial) {
    // the CTA's kScanBlock elements are scanned by warp 0 in 32-element steps (the matrix is small)
    const uint32_t begin = blockIdx.x * kScanBlock, end = min(n, begin + kScanBlock);
    if (threadIdx.x >= 32) return;
    uint32_t run = partial[blockIdx.x];
    for (uint32_t base = begin; base < end; base += 32) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < end ? a[i] : 0;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (threadIdx.x >= (uint32_t)d) incl += o;
        }
        if (i < end) a[i] = run + incl - v;
        run += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
}

// ---- stable scatter ---------------------------------------------------------------------------------
// The CTA first sorts its 8192 elements by digit in shared memory (stable: warps in order, rounds in
// order, lanes in order inside a match group), then writes them out bucket by bucket, so that the
// ~32 elements a CTA has for one digit leave as one contiguous 128-byte piece per column instead of
// 32 separate 4-byte stores (the input of a time-major feed has 32 different digits per warp load).
struct SortSmem {
    uint32_t slot[kSortWarps][256];   // per warp and digit: count, then next position inside the CTA's sorted order
    uint32_t gdelta[256];             // global index of the digit's first element of this CTA minus its local start
    uint32_t wsum[kSortWarps];
    uint32_t dst[kSortBlock];         // global destination of the element at each sorted position
    uint32_t val[kSortBlock];         // one column in sorted order
};
constexpr size_t kSortSmemBytes = sizeof(SortSmem);

__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const __grid_constant__ SortParams P) {
    extern __shared__ __align__(16) unsigned char sort_smem_raw[];
    SortSmem& S = *reinterpret_cast<SortSmem*>(sort_smem_raw);
    const uint32_t lane = lane_id(), w = threadIdx.x >> 5, t = threadIdx.x;
    for (int k = 0; k < kSortWarps; ++k) S.slot[k][t] = 0;
    __syncthreads();
    const uint32_t cbegin = blockIdx.x * kSortBlock;
    const uint32_t n_cta = min(kSortBlock, P.n - cbegin);
    const uint32_t wbegin = cbegin + w * kSortWarpSpan;
    // A: digit counts of each warp's span
    for (uint32_t r = 0; r < (uint32_t)kSortRounds; ++r) {
        const uint32_t i = wbegin + r * 32 + lane;
        const uint32_t d = i < P.n ? sort_digit(P, i) : 256u + lane;   // out-of-range lanes match nobody
        const uint32_t m = __match_any_sync(0xFFFFFFFFu, d);
        if (d < 256u && lane == (uint32_t)(__ffs(m) - 1)) S.slot[w][d] += __popc(m);
    }
    __syncthreads();
    // B: thread t owns digit t: where its elements start inside the CTA's sorted order (exclusive scan
    // over the digits), the warps following one another inside a digit, and the global base
    {
        uint32_t cnt = 0;
        for (int k = 0; k < kSortWarps; ++k) cnt += S.slot[k][t];
        uint32_t incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= (uint32_t)d) incl += o;
        }
        if (lane == 31) S.wsum[w] = incl;
        __syncthreads();
        uint32_t lstart = incl - cnt;
        for (uint32_t k = 0; k < w; ++k) lstart += S.wsum[k];
        S.gdelta[t] = P.counts[t * P.n_ctas + blockIdx.x] - lstart;
        uint32_t run = lstart;
        for (int k = 0; k < kSortWarps; ++k) {
            const uint32_t c = S.slot[k][t];
            S.slot[k][t] = run;
            run += c;
        }
    }
    __syncthreads();
    // C: position of every element in the sorted order (kept in registers), destination per position
    uint32_t lp[kSortRounds];
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
        const uint32_t i = wbegin + r * 32 + lane;
        const bool in = i < P.n;
        const uint32_t d = in ? sort_digit(P, i) : 256u + lane;
        const uint32_t m = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t leader = __ffs(m) - 1;
        uint32_t base = 0;
        if (in && lane == leader) {
            base = S.slot[w][d];
            S.slot[w][d] = base + __popc(m);
        }
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        lp[r] = base + __popc(m & ((1u << lane) - 1u));
        if (in) S.dst[lp[r]] = S.gdelta[d] + lp[r];
        __syncwarp();
    }
    // D: one column at a time through shared memory, written out in sorted order
    for (int c = 0; c < 4; ++c) {
        __syncthreads();   // dst complete (c == 0) / the previous column has left val
#pragma unroll
        for (int r = 0; r < kSortRounds; ++r) {
            const uint32_t i = wbegin + r * 32 + lane;
            if (i < P.n) S.val[lp[r]] = P.in[c][i];
        }
        __syncthreads();
        for (uint32_t k = t; k < n_cta; k += kSortThreads) P.out[c][S.dst[k]] = S.val[k];
    }
}

}  // namespace kgs

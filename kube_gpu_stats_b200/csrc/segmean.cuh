// segmean.cuh - the reference's own statistic on a CSR layout of fp64 series values.
//
// gpu_util_stats.py:77-80: mean = sum(val_list)/len(val_list), 0.0 for an empty series;
// gpu_util_stats.py:51: sum(val_list) * step_sec / time_range_sec.  Python 2's sum() adds left
// to right in fp64, so short series (the reference queries <= 25 points per series,
// gpu_util_stats.py:155-159) are summed by ONE thread in index order: bit-identical to the
// reference.  Long series (1 Hz replays) use a warp: strided partial sums, then a shuffle tree;
// that differs from the naive sum by rounding only (tests hold it to 1e-12 relative).
#pragma once
#include "common.cuh"

namespace kgs {

constexpr uint32_t kSeqMax = 64;   // longest series summed sequentially

// div == 0: the mean of :77-80.  Otherwise (sum * mul) / div, the evaluation order of :51.
__device__ __forceinline__ double seg_result(double s, unsigned long long len, double mul, double div) {
    if (div != 0.0) return __ddiv_rn(__dmul_rn(s, mul), div);
    return len ? __ddiv_rn(s, (double)len) : 0.0;
}

__global__ void k_segmean_short(const double* __restrict__ vals, const unsigned long long* __restrict__ off,
                                uint32_t n_series, double mul, double div, double* __restrict__ out) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_series) return;
    const unsigned long long a = off[s], b = off[s + 1];
    if (b - a > kSeqMax) return;   // left to k_segmean_long
    double acc = 0.0;
    for (unsigned long long i = a; i < b; ++i) acc = __dadd_rn(acc, vals[i]);
    out[s] = seg_result(acc, b - a, mul, div);
}

__global__ void k_segmean_long(const double* __restrict__ vals, const unsigned long long* __restrict__ off,
                               const uint32_t* __restrict__ long_ids, uint32_t n_long, double mul, double div,
                               double* __restrict__ out) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_long) return;
    const uint32_t s = long_ids[w];
    const unsigned long long a = off[s], b = off[s + 1];
    double acc = 0.0;
    for (unsigned long long i = a + lane_id(); i < b; i += 32) acc = __dadd_rn(acc, vals[i]);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc = __dadd_rn(acc, __shfl_down_sync(0xFFFFFFFFu, acc, d));
    if (lane_id() == 0) out[s] = seg_result(acc, b - a, mul, div);
}

}  // namespace kgs

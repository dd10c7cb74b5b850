// common.cuh - small device/host helpers shared by the libkgs kernels (sm_100a).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace kgs {

constexpr uint32_t kPodNone = 0xFFFFFFFFu;
constexpr uint32_t kInvalid = 0xFFFFFFFFu;
constexpr unsigned long long kEmptyKey = 0xFFFFFFFFFFFFFFFFull;

// Exact unsigned 32-bit division by a run-time constant (round-up magic number, two shifts):
//   q = (t + ((x - t) >> s1)) >> s2,  t = umulhi(m, x).
// d == 1 -> (0,0,0); d == 2^k -> (0,1,k-1); else m = floor(2^32 (2^L - d) / d) + 1, L = ceil(log2 d).
struct FastDiv {
    uint32_t m, s1, s2;
};

inline FastDiv make_fastdiv(uint32_t d) {
    FastDiv f{0, 0, 0};
    if (d <= 1) return f;
    uint32_t L = 0;
    while ((1ull << L) < d) ++L;
    if ((d & (d - 1)) == 0) {
        f.s1 = 1;
        f.s2 = L - 1;
        return f;
    }
    f.m = (uint32_t)((((1ull << L) - d) << 32) / d + 1);
    f.s1 = 1;
    f.s2 = L - 1;
    return f;
}

__device__ __forceinline__ uint32_t fdiv(uint32_t x, const FastDiv& f) {
    const uint32_t t = __umulhi(f.m, x);
    return (t + ((x - t) >> f.s1)) >> f.s2;
}

// 32-bit mix of the (node, pid) pair; the table index is hash & mask.
__host__ __device__ __forceinline__ uint32_t hash_node_pid(uint32_t node, uint32_t pid) {
    uint32_t h = node * 0x9E3779B1u ^ pid * 0x85EBCA6Bu;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 13;
    return h;
}

// Streaming 128-bit load: read-only path, do not keep the line in L1 (each sample column is
// read exactly once per pass).
__device__ __forceinline__ uint4 ld_stream_v4(const uint32_t* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// One slot of the open-addressing join table (16 B, one LDG.128 per probe step).
struct __align__(16) JoinSlot {
    unsigned long long key;   // node << 32 | pid, kEmptyKey when free
    uint32_t pod;             // value
    uint32_t src;             // 1 + index of the build pair that owns the slot (last pair wins)
};

__device__ __forceinline__ uint32_t join_lookup(const JoinSlot* __restrict__ slots, uint32_t mask,
                                                uint32_t node, uint32_t pid) {
    const unsigned long long key = ((unsigned long long)node << 32) | pid;
    uint32_t s = hash_node_pid(node, pid) & mask;
    while (true) {
        const uint4 e = __ldg(reinterpret_cast<const uint4*>(slots + s));
        const unsigned long long k = ((unsigned long long)e.y << 32) | e.x;
        if (k == key) return e.z;
        if (k == kEmptyKey) return kPodNone;
        s = (s + 1) & mask;
    }
}

}  // namespace kgs

// join.cuh - (node,pid) -> pod open-addressing hash join: build, resolve, probe.
//
// Takes over the dict build/probe of the reference's join (gpu_util_stats.py:139-148 build,
// :86-91 probe) with the key re-expressed as node << 32 | pid.  Duplicate build keys keep the
// LAST pair (dict assignment semantics, gpu_util_stats.py:78/:148): every inserter of a key
// does atomicMax on the slot's `src` with 1 + its pair index, and k_join_resolve then copies
// the winner's pod id into the slot.  The result is independent of thread scheduling.
#pragma once
#include "common.cuh"

namespace kgs {

__global__ void k_join_clear(JoinSlot* slots, uint32_t n_slots) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += gridDim.x * blockDim.x) {
        JoinSlot e;
        e.key = kEmptyKey;
        e.pod = kPodNone;
        e.src = 0;
        slots[i] = e;
    }
}

__global__ void k_join_build(JoinSlot* slots, uint32_t mask, const unsigned long long* __restrict__ keys,
                             uint32_t m) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        const unsigned long long key = keys[i];
        if (key == kEmptyKey) continue;   // reserved as the free marker; rejected on the host too
        uint32_t s = hash_node_pid((uint32_t)(key >> 32), (uint32_t)key) & mask;
        while (true) {
            const unsigned long long prev = atomicCAS(&slots[s].key, kEmptyKey, key);
            if (prev == kEmptyKey || prev == key) {
                atomicMax(&slots[s].src, i + 1);
                break;
            }
            s = (s + 1) & mask;
        }
    }
}

__global__ void k_join_resolve(JoinSlot* slots, uint32_t n_slots, const uint32_t* __restrict__ pod_ids) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += gridDim.x * blockDim.x) {
        const uint32_t src = slots[i].src;
        if (src) slots[i].pod = pod_ids[src - 1];
    }
}

// out[i] = pod of sample i.  Four samples per thread, 128-bit loads of the two key columns.
__global__ void __launch_bounds__(256)
k_join_probe(const JoinSlot* __restrict__ slots, uint32_t mask, const uint32_t* __restrict__ gid,
             const uint32_t* __restrict__ pid, uint64_t n, FastDiv dgpn, uint32_t* __restrict__ out) {
    const uint64_t n4 = n >> 2;
    for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += (uint64_t)gridDim.x * blockDim.x) {
        const uint4 g = ld_stream_v4(gid + 4 * q);
        const uint4 p = ld_stream_v4(pid + 4 * q);
        uint4 r;
        r.x = join_lookup(slots, mask, fdiv(g.x, dgpn), p.x);
        r.y = (g.y == g.x && p.y == p.x) ? r.x : join_lookup(slots, mask, fdiv(g.y, dgpn), p.y);
        r.z = (g.z == g.y && p.z == p.y) ? r.y : join_lookup(slots, mask, fdiv(g.z, dgpn), p.z);
        r.w = (g.w == g.z && p.w == p.z) ? r.z : join_lookup(slots, mask, fdiv(g.w, dgpn), p.w);
        *reinterpret_cast<uint4*>(out + 4 * q) = r;
    }
    // ragged tail (n not a multiple of 4)
    const uint64_t tail = n4 << 2;
    if (blockIdx.x == 0 && threadIdx.x < (n - tail)) {
        const uint64_t i = tail + threadIdx.x;
        out[i] = join_lookup(slots, mask, fdiv(gid[i], dgpn), pid[i]);
    }
}

}  // namespace kgs

// noalloc.cuh - kgs_join_probe_noalloc: the (node,pid)->pod join for callers that must not
// allocate GPU memory.
//
// A telemetry agent usually runs beside the jobs it watches, on the same GPUs.  There a
// cudaMalloc is unwelcome for two reasons: it can fail (or push a tenant out of memory) when
// the training job has sized itself to the card, and it synchronises the whole device, which
// shows up as a stall in the very workload being measured.  Hosts that give the agent a fixed
// arena up front, or that run it inside a process whose allocator owns all device memory, have
// the same constraint.  This entry point therefore takes every input and output as a device
// pointer from the caller and keeps its only scratch - the open-addressing hash table - in a
// fixed array that is part of the loaded module.
//
// Limits and sharp edges (also in include/kgs.h):
//   * at most KGS_NOALLOC_MAX_PAIRS (65 536) join pairs: the table has 131 072 slots of 16 B =
//     2 MiB of device memory per GPU that loads libkgs.so, held until the process exits;
//   * the first call on each device looks the table's device address up; make that first call
//     outside any stream capture;
//   * every call on a device shares the one table and rebuilds it from scratch (it is cleared
//     first, nothing of the previous call survives); two calls whose kernels overlap on the
//     same GPU - different streams, different threads - race on it.  Serialise them.
#pragma once
#include <mutex>

#include "join.cuh"

namespace kgs {

constexpr uint32_t kNoallocSlots = 1u << 17;
constexpr uint64_t kNoallocMaxPairs = kNoallocSlots / 2;
constexpr int kNoallocMaxDevices = 64;

// The table.  In host code this name is only a handle for cudaGetSymbolAddress - it is NOT a
// device address and must never be passed to a kernel as one.
static __device__ __align__(256) JoinSlot g_noalloc_slots[kNoallocSlots];

// Device address of g_noalloc_slots on `device` (which must be current), looked up once.
inline cudaError_t noalloc_table(int device, JoinSlot** out) {
    static std::mutex mu;
    static JoinSlot* cached[kNoallocMaxDevices] = {nullptr};
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_slots);
        if (e != cudaSuccess) return e;
        cached[device] = static_cast<JoinSlot*>(p);
    }
    *out = cached[device];
    return cudaSuccess;
}

}  // namespace kgs

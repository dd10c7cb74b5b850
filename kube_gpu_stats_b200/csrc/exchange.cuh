// exchange.cuh - the one exchange step of the path: every shard's finished tables to every GPU of
// the box (north_star: "per-shard partial aggregates are merged with NCCL allgather over NVLink").
//
// The reference has one global reduction only (who_use_gpu.py:18-20, a few integers); its tables are
// keyed by (server, pod) (gpu_util_stats.py:70-71), so with nodes partitioned over the GPUs every cell
// lives on one shard and the exchange is a pure all-gather of bytes, no arithmetic.
//
// Two transports behind the C ABI (include/kgs.h, kgs_exchange_*):
//   * peer pushes: the gathered buffers are peer-mapped by the caller (cudaIpc*, cuMem*, or a
//     symmetric-memory allocator); each rank copies its region into its slot of every rank's buffer
//     with cudaMemcpyAsync - NVLink through the copy engines, no SM - and then runs k_peer_barrier:
//     one CTA that tells every peer "my copies into your buffer are complete" and waits for the same
//     word from every peer.  Two buffer halves and two flag rows alternate between consecutive pushes;
//   * NCCL: one ncclAllGather of the region on the exchange stream.  libnccl is looked up at run time
//     (dlopen), so the library has no link-time dependency on it and, inside a process that already
//     loaded an NCCL (torch), uses that very copy - which is what makes a caller's ncclComm_t usable.
#pragma once
#include <dlfcn.h>

#include "common.cuh"

namespace kgs {

constexpr int kMaxRanks = 16;

struct PeerFlags {
    uint32_t* row[kMaxRanks];   // row[r]: rank r's flag words (2 halves x kMaxRanks u32) as mapped on this device
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Thread t < n: announce to rank t, then wait for rank t.  The copies this rank queued before the
// kernel (same stream) have completed when it starts; the release store orders them before the flag.
// A peer that never arrives must not hang the GPU: after timeout_ns the kernel gives up and raises
// *timed_out (the host reports it).
__global__ void k_peer_barrier(PeerFlags pf, int n, int me, int half, uint32_t seq, unsigned long long timeout_ns,
                               uint32_t* timed_out) {
    const int t = threadIdx.x;
    if (t >= n) return;
    __threadfence_system();
    st_release_sys(pf.row[t] + half * kMaxRanks + me, seq);
    const uint32_t* mine = pf.row[me] + half * kMaxRanks + t;
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(mine) - seq) < 0) {
        __nanosleep(200);
        if (global_ns() - t0 > timeout_ns) {
            atomicExch(timed_out, 1u);
            break;
        }
    }
}

// ---- NCCL by name ---------------------------------------------------------------------------------
struct NcclApi {
    typedef int (*all_gather_t)(const void*, void*, size_t, int, void*, cudaStream_t);
    typedef int (*get_unique_id_t)(void*);
    struct UniqueId { char bytes[128]; };
    typedef int (*comm_init_rank_t)(void**, int, UniqueId, int);
    typedef int (*comm_destroy_t)(void*);
    typedef const char* (*get_error_string_t)(int);
    all_gather_t all_gather = nullptr;
    get_unique_id_t get_unique_id = nullptr;
    comm_init_rank_t comm_init_rank = nullptr;
    comm_destroy_t comm_destroy = nullptr;
    get_error_string_t get_error_string = nullptr;
    const char* load_error = nullptr;
    bool ok = false;
};

inline const NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        // RTLD_NOLOAD first: the copy the process already has (torch ships its own), else the system's
        void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) {
            a.load_error = "libnccl.so.2 not found (dlopen)";
            return a;
        }
        a.all_gather = reinterpret_cast<NcclApi::all_gather_t>(dlsym(lib, "ncclAllGather"));
        a.get_unique_id = reinterpret_cast<NcclApi::get_unique_id_t>(dlsym(lib, "ncclGetUniqueId"));
        a.comm_init_rank = reinterpret_cast<NcclApi::comm_init_rank_t>(dlsym(lib, "ncclCommInitRank"));
        a.comm_destroy = reinterpret_cast<NcclApi::comm_destroy_t>(dlsym(lib, "ncclCommDestroy"));
        a.get_error_string = reinterpret_cast<NcclApi::get_error_string_t>(dlsym(lib, "ncclGetErrorString"));
        a.ok = a.all_gather && a.get_unique_id && a.comm_init_rank && a.comm_destroy && a.get_error_string;
        if (!a.ok) a.load_error = "libnccl.so.2 lacks ncclAllGather / ncclCommInitRank";
        return a;
    }();
    return api;
}

constexpr int kNcclUint8 = 1;   // ncclUint8 in nccl.h (ncclInt8 = 0, ncclUint8 = 1): stable since NCCL 2.0

}  // namespace kgs

// kgs.cu - host side of libkgs: the C ABI declared in include/kgs.h over the sm_100a kernels in
// join.cuh / accum.cuh / finalize.cuh / segmean.cuh.  CUDA runtime only; no torch, no CPU path.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kgs.h"

#include "accum.cuh"
#include "common.cuh"
#include "finalize.cuh"
#include "finalize_slide.cuh"
#include "finalize_groups.cuh"
#include "exchange.cuh"
#include "join.cuh"
#include "noalloc.cuh"
#include "segmean.cuh"
#include "sort.cuh"
#include "promjson.cuh"

using namespace kgs;

namespace {

thread_local std::string g_create_error;

struct DevBuf {   // grow-only device buffer
    void* p = nullptr;
    size_t cap = 0;
};

struct Batch {
    DevBuf col[4];
    const uint32_t* ptr[4] = {nullptr, nullptr, nullptr, nullptr};
    uint64_t n = 0;
    cudaEvent_t ready = nullptr, consumed = nullptr;
    int order_class = 0;        // what the host columns looked like at ingest (input_order_class): 0 long runs, 1 disordered, 2 interleaved series
};

struct KindState {
    uint32_t n_keys = 0;
    uint32_t* cnt = nullptr;
    unsigned long long* sum = nullptr;
    uint32_t* mx = nullptr;
    uint32_t* head = nullptr;
    uint32_t* o_count = nullptr;
    unsigned long long* o_sum = nullptr;
    uint32_t* o_max = nullptr;
    double* o_mean = nullptr;
    double* o_pct = nullptr;
    void* p_sum = nullptr;
    uint8_t* p_sum_hi = nullptr;
    void* p_count = nullptr;
    void* p_max = nullptr;
    void* p_rank = nullptr;
};

struct TimedLaunch {
    int kid;
    cudaEvent_t e0, e1;
};

}  // namespace

struct kgs_ctx {
    int device = 0;
    cudaStream_t copy = nullptr, compute = nullptr;
    std::string err;
    Batch batch[2];
    int cur = -1;
    bool auto_reorder = true;      // accumulate from a series-major copy of a batch that interleaves its series
    bool shadow_valid = false;     // the other batch slot holds that copy of batch[cur]
    // join table
    DevBuf slots, build_keys, build_pods;
    uint32_t n_slots = 0, n_pods = 0;
    bool has_table = false;
    // statistics job
    bool job_open = false, job_final = false;
    kgs_window_spec spec{};
    uint32_t W = 1, n_panes = 0;
    KindState kind[3];
    DevBuf arena, vals, desc, tmp, tmp2, tmp3, tmp4;
    DevBuf sortbuf[4];
    // decoded Prometheus answer (promjson.cuh)
    DevBuf mx_text, mx_cta_trans, mx_cta_counts, mx_chunk_state, mx_chunk_counts, mx_pos, mx_ts, mx_val, mx_off, mx_spans,
        mx_long, mx_out, mx_ctl;
    uint64_t mx_n_series = 0, mx_n_samples = 0, mx_n_long = 0;
    std::vector<cudaEvent_t> mx_arrived;   // one event per piece of text on the copy stream
    void* mx_stage[3] = {nullptr, nullptr, nullptr};          // pinned pieces a pageable answer is staged through
    cudaEvent_t mx_stage_free[3] = {nullptr, nullptr, nullptr};
    size_t mx_stage_bytes = 0;
    int mx_stage_threads = 0;              // host threads copying into a pinned piece (0 = three quarters of the cores, at most 12)
    uint32_t mx_piece_bytes = 32u << 20;
    bool mx_valid = false;
    size_t clear_bytes = 0;
    // The accumulators exist twice and alternate between jobs: the half the NEXT job will use is cleared on a stream of
    // its own while this job's finalize kernels run, so that the 24 MB memset of configs[1] is not on the step's path.
    cudaStream_t prep = nullptr;
    cudaEvent_t acc_idle = nullptr, acc_clean[2] = {nullptr, nullptr};
    struct PreCleared { bool valid = false; void* base = nullptr; size_t bytes = 0, shift = 0; } pre[2];
    bool preclear_on = true, preclear_pending = false;
    size_t acc_shift_other = 0;    // where the other half of the accumulators lies in the arena of the open job
    uint32_t* ctl = nullptr;       // [0] run cursor [1] run-log overflow [2] big cells listed [3] wide cells listed
    uint32_t* rare_list = nullptr; // cells left to k_finalize_rare: wide ones from the front, big ones from the back
    uint32_t rare_cap = 0;
    uint8_t* done = nullptr;       // per window cell: finished by k_finalize_groups in finalize pass done_gen
    uint32_t done_gen = 0;
    size_t done_bytes = 0;
    bool last_was_accum = false;   // the last work queued on the compute stream is k_accum
    bool pdl = true;               // finalize kernels launched as programmatic dependents (not while profiling per kernel)
    bool groups_path = true;       // k_finalize_groups for tumbling-window jobs that include the per-GPU table
    uint32_t thread_cell_max = 256;   // cells above this many samples are counted by a warp, not by a thread
    size_t result_off = 0, result_bytes = 0, exchange_bytes = 0;
    int packed_exchange = 0;       // option: 0 off, 1 packed, 2 compact (kgs_set_packed_exchange)
    bool packed_on = false, packed_sum64 = false, packed_compact = false;   // state of the open job
    uint32_t* esc_hdr = nullptr;   // compact: escape list header + entries inside the packed region
    uint32_t esc_cap = 0;
    size_t packed_off = 0, packed_bytes = 0;
    uint64_t vals_cap = 0, vals_used = 0, accumulated = 0;
    uint32_t desc_cap = 0, desc_pool_base = 0, desc_priv_next = 0;
    uint64_t run_cap_override = 0;
    uint64_t job_seq = 0;
    bool slide = true;             // W > 1 jobs keep a sliding histogram per chunk of windows (finalize_slide.cuh)
    uint32_t slide_chunk = 8;      // measured best on B200 (4: 3.0 ms, 8: 1.9, 16: 2.1, 64: 5.0 for the cfg5-shaped test)
    int accum_waves = 2;           // CTAs per resident slot: > 1 keeps the tail short when another kernel (a collective) shares the SMs
    bool dense_path = true;
    bool dense16_path = false;     // k_accum: two tiles at once (16 samples per lane) where the input is dense; measured slower, KGS_DENSE16=1
    bool accum_tma = false;        // stage tiles through shared memory with cp.async.bulk + mbarrier
    int accum_ctas_per_sm = 4;     // resident CTAs of 8 warps per SM at 64 registers
    // table exchange (exchange.cuh)
    cudaStream_t xchg = nullptr;               // pushes / collectives run here, beside the next job's kernels
    cudaEvent_t fin_done = nullptr;            // the finalize kernels of the job being exchanged
    cudaEvent_t region_free[2] = {nullptr, nullptr};   // result-region half i has been read by its exchange
    bool region_busy[2] = {false, false};
    int region_half = 0;                       // half the open job writes
    int px_n = 0, px_me = 0;                   // peer exchange: ranks, my rank
    char* px_gathered[kMaxRanks] = {nullptr};  // rank r's gathered buffer as mapped on this device
    bool px_self_copy = true;                  // also copy this rank's tables into its own gathered buffer
    char* px_multicast = nullptr;              // the gathered buffers of ALL ranks behind one multicast address (optional)
    PeerFlags px_flags{};
    size_t px_slot = 0;
    uint32_t px_seq = 0;
    bool px_defer = true;                      // queue a push's copies behind the NEXT job's accumulate kernel (see kgs_exchange_push)
    struct { bool on = false; char* src = nullptr; size_t nbytes = 0; int half = 0, region_half = 0; uint32_t seq = 0; } px_pending;
    uint32_t* px_timeout = nullptr;            // device word raised by k_peer_barrier when a peer never arrived
    void* nccl_comm = nullptr;                 // own communicator (kgs_nccl_init)
    // measurement
    uint64_t launches = 0;
    bool prof = false;
    std::vector<TimedLaunch> timed;
    std::vector<cudaEvent_t> event_pool;
    double k_ms[KGS_N_KERNELS] = {0};
    uint64_t k_n[KGS_N_KERNELS] = {0};
    cudaEvent_t mark[2] = {nullptr, nullptr};
};

namespace {

int fail(kgs_handle h, int code, const std::string& msg) {
    if (h) h->err = msg;
    return code;
}

#define KGS_CUDA(h, call)                                                                        \
    do {                                                                                         \
        cudaError_t e_ = (call);                                                                 \
        if (e_ != cudaSuccess)                                                                   \
            return fail(h, KGS_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));    \
    } while (0)

int reserve(kgs_handle h, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return KGS_OK;
    if (b.p) KGS_CUDA(h, cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(h, KGS_ERR_NOMEM, "cudaMalloc of " + std::to_string(want) + " bytes failed: " + cudaGetErrorString(e));
    }
    b.cap = want;
    return KGS_OK;
}

// Grow a device buffer that holds live data: allocate, copy the first `keep` bytes, free the old one.
// The compute stream is drained first (queued kernels hold pointers into the old allocation).
int grow_copy(kgs_handle h, DevBuf& b, size_t bytes, size_t keep) {
    if (bytes <= b.cap) return KGS_OK;
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    DevBuf bigger;
    int rc = reserve(h, bigger, bytes);
    if (rc) return rc;
    if (keep && b.p) KGS_CUDA(h, cudaMemcpy(bigger.p, b.p, keep, cudaMemcpyDeviceToDevice));
    if (b.p) KGS_CUDA(h, cudaFree(b.p));
    b = bigger;
    return KGS_OK;
}

cudaEvent_t pooled_event(kgs_handle h) {
    if (!h->event_pool.empty()) {
        cudaEvent_t e = h->event_pool.back();
        h->event_pool.pop_back();
        return e;
    }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

// Every kernel launch goes through here: counts it and, with profiling on, brackets it with
// events on the launching stream.
template <class F>
void launch(kgs_handle h, int kid, cudaStream_t st, F&& f) {
    if (h->prof) {
        TimedLaunch t{kid, pooled_event(h), pooled_event(h)};
        cudaEventRecord(t.e0, st);
        f();
        cudaEventRecord(t.e1, st);
        h->timed.push_back(t);
    } else {
        f();
    }
    h->launches++;
}

int drain_timed(kgs_handle h) {
    for (auto& t : h->timed) {
        KGS_CUDA(h, cudaEventSynchronize(t.e1));
        float ms = 0.f;
        KGS_CUDA(h, cudaEventElapsedTime(&ms, t.e0, t.e1));
        h->k_ms[t.kid] += ms;
        h->k_n[t.kid] += 1;
        h->event_pool.push_back(t.e0);
        h->event_pool.push_back(t.e1);
    }
    h->timed.clear();
    return KGS_OK;
}

// Launch with programmatic dependent launch (PDL): the kernel may be scheduled while its predecessor on
// the stream is still draining; it waits in cudaGridDependencySynchronize() (first statement of the
// kernels launched this way) until the predecessor has completed and its writes are visible.  The
// predecessors call cudaTriggerProgrammaticLaunchCompletion() when they start, so what is overlapped is
// the successor's launch and ramp-up with the predecessor's tail.
template <class P>
cudaError_t launch_after(void (*kern)(P), unsigned grid, unsigned block, cudaStream_t st, bool pdl, const P& params) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, params);
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// What order is this batch in?  Three windows of up to 512 samples are looked at on the host.
//   0  long runs: fewer than a quarter of the neighbours differ in (gpu, pid) - collector order; decided after a
//      few hundred compares;
//   2  interleaved series: most neighbours differ AND the keys repeat with a period of 2..64 samples - a collector
//      that reports all processes of a GPU (or the GPUs of a small box) second by second;
//   1  disordered otherwise - a time-major feed of many GPUs, a random stream (configs[2]).
// In classes 1 and 2 nearly every sample starts a new run: the kernels stay exact but crawl, and a percentile job
// pays a 16-byte run descriptor per sample and key kind.  kgs_stats_accumulate therefore reads a series-major copy
// of a class-2 batch always, and of a class-1 batch when the job has percentiles (without them the per-CTA
// accumulator cache is the faster way for a random stream).
int input_order_class(const uint32_t* gid, const uint32_t* pid, uint64_t n) {
    if (n < 128) return 0;
    const uint64_t win = n < 512 ? n : 512;
    const uint64_t at[3] = {0, (n - win) / 2, n - win};
    int disordered = 0, periodic = 0;
    for (int w = 0; w < 3; ++w) {
        const uint32_t* g = gid + at[w];
        const uint32_t* p = pid + at[w];
        uint64_t changes = 0;
        for (uint64_t i = 0; i + 1 < win; ++i) changes += (g[i] != g[i + 1]) | (p[i] != p[i + 1]);
        if (changes * 4 < win) continue;
        ++disordered;
        for (uint64_t period = 2; period <= 64 && period * 2 < win; ++period) {
            uint64_t same = 0;
            for (uint64_t i = 0; i + period < win; ++i) same += (g[i] == g[i + period]) & (p[i] == p[i + period]);
            if (same * 4 >= 3 * (win - period)) {
                ++periodic;
                break;
            }
        }
    }
    return periodic >= 2 ? 2 : (disordered >= 2 ? 1 : 0);
}

// A pageable answer (the `bytes` object an HTTP client hands over) on its way to the device.  cudaMemcpyAsync from
// pageable memory goes through the driver's own small staging buffer, synchronously: 10 GB/s.  Here a thread of the
// library copies piece after piece into three pinned buffers with several host threads and queues the pinned piece on
// the copy stream, so the host copy of piece i+1 runs beside the DMA of piece i and the decode of piece i-1.
struct TextStager {
    kgs_ctx* h = nullptr;
    const char* text = nullptr;
    char* dtext = nullptr;
    uint64_t nbytes = 0;
    size_t padded = 0, piece_bytes = 0;
    uint32_t n_pieces = 0;
    int n_threads = 1;
    std::atomic<uint32_t> issued{0};       // pieces whose copy and arrival event have been queued
    std::atomic<int> cuda_error{0};
    std::atomic<bool> stop{false};
    std::thread worker;

    // helper threads live for one answer and share every piece: job counter up = a new piece to copy
    std::vector<std::thread> helpers;
    std::atomic<uint64_t> job{0};
    std::atomic<int> pending{0};
    std::atomic<bool> helpers_stop{false};
    char* job_dst = nullptr;
    const char* job_src = nullptr;
    size_t job_len = 0, job_share = 0;

    void copy_share(int t) {
        const size_t a = std::min(job_len, (size_t)t * job_share), b = std::min(job_len, a + job_share);
        if (b > a) memcpy(job_dst + a, job_src + a, b - a);
    }
    void helper_loop(int t) {
        uint64_t seen = 0;
        for (;;) {
            while (job.load(std::memory_order_acquire) == seen) {
                if (helpers_stop.load(std::memory_order_relaxed)) return;
                __builtin_ia32_pause();
            }
            seen = job.load(std::memory_order_acquire);
            copy_share(t);
            pending.fetch_sub(1, std::memory_order_release);
        }
    }
    void host_copy(char* dst, const char* src, size_t len) {
        if (helpers.empty() || len < (4u << 20)) {
            memcpy(dst, src, len);
            return;
        }
        job_dst = dst; job_src = src; job_len = len;
        job_share = (len / n_threads + 4095) & ~(size_t)4095;
        pending.store((int)helpers.size(), std::memory_order_relaxed);
        job.fetch_add(1, std::memory_order_release);
        copy_share(0);
        while (pending.load(std::memory_order_acquire)) __builtin_ia32_pause();
    }

    void run() {
        cudaError_t e = cudaSetDevice(h->device);
        for (int t = 1; t < n_threads; ++t) helpers.emplace_back([this, t] { helper_loop(t); });
        for (uint32_t i = 0; i < n_pieces && e == cudaSuccess && !stop.load(); ++i) {
            const int slot = (int)(i % 3);
            const size_t off = (size_t)i * piece_bytes;
            const size_t len = off < nbytes ? std::min<size_t>(nbytes - off, piece_bytes) : 0;
            if (i >= 3) e = cudaEventSynchronize(h->mx_stage_free[slot]);      // the DMA that last read this buffer
            if (e == cudaSuccess && len) {
                host_copy(static_cast<char*>(h->mx_stage[slot]), text + off, len);
                e = cudaMemcpyAsync(dtext + off, h->mx_stage[slot], len, cudaMemcpyHostToDevice, h->copy);
            }
            if (e == cudaSuccess && i + 1 == n_pieces) e = cudaMemsetAsync(dtext + nbytes, 0, padded - nbytes, h->copy);
            if (e == cudaSuccess) e = cudaEventRecord(h->mx_arrived[i], h->copy);
            if (e == cudaSuccess) e = cudaEventRecord(h->mx_stage_free[slot], h->copy);
            if (e == cudaSuccess) issued.store(i + 1, std::memory_order_release);
        }
        helpers_stop.store(true);
        for (auto& t : helpers) t.join();
        helpers.clear();
        if (e != cudaSuccess) cuda_error.store((int)e);
    }
    // true once piece k - 1 has been queued (false: the worker failed)
    bool wait_issued(uint32_t k) {
        while (issued.load(std::memory_order_acquire) < k) {
            if (cuda_error.load()) return false;
            std::this_thread::yield();
        }
        return true;
    }
    ~TextStager() {
        stop.store(true);
        if (worker.joinable()) worker.join();
    }
};

// Queue the copies and the barrier of the pending push on the exchange stream, behind whatever the compute stream
// holds at this moment.
int flush_pending_push(kgs_handle h) {
    if (!h->px_pending.on) return KGS_OK;
    h->px_pending.on = false;
    const int n = h->px_n, me = h->px_me, half = h->px_pending.half;
    const uint32_t seq = h->px_pending.seq;
    char* src = h->px_pending.src;
    const size_t nbytes = h->px_pending.nbytes;
    KGS_CUDA(h, cudaEventRecord(h->fin_done, h->compute));
    KGS_CUDA(h, cudaStreamWaitEvent(h->xchg, h->fin_done, 0));
    static const int dbg = getenv("KGS_XCHG_DEBUG") ? atoi(getenv("KGS_XCHG_DEBUG")) : 0;   // experiments only: 1 no self copy, 2 no barrier, 4 no peer copies
    if (h->px_multicast) {   // ONE copy: the NVSwitch replicates the writes into every rank's buffer (this one included)
        char* dst = h->px_multicast + ((size_t)half * n + me) * h->px_slot;
        KGS_CUDA(h, cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDeviceToDevice, h->xchg));
    }
    for (int k = 0; k < n && !h->px_multicast; ++k) {   // the neighbour first: spreads the links
        const int r = (me + 1 + k) % n;
        if ((r == me && ((dbg & 1) || !h->px_self_copy)) || (r != me && (dbg & 4))) continue;
        char* dst = h->px_gathered[r] + ((size_t)half * n + me) * h->px_slot;
        KGS_CUDA(h, cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDeviceToDevice, h->xchg));
    }
    if (!(dbg & 2))
        launch(h, KGS_K_EXCHANGE, h->xchg, [&] {
            k_peer_barrier<<<1, 32, 0, h->xchg>>>(h->px_flags, n, me, half, seq, 10ull * 1000 * 1000 * 1000, h->px_timeout);
        });
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaEventRecord(h->region_free[h->px_pending.region_half], h->xchg));
    h->region_busy[h->px_pending.region_half] = true;
    return KGS_OK;
}

int grid_for(uint64_t work_items, int per_block, int max_blocks) {
    uint64_t b = (work_items + per_block - 1) / per_block;
    if (b < 1) b = 1;
    if (b > (uint64_t)max_blocks) b = max_blocks;
    return (int)b;
}

}  // namespace

extern "C" {

const char* kgs_version(void) { return "libkgs 0.1 (sm_100a)"; }

const char* kgs_last_error(kgs_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int kgs_create(int device, kgs_handle* out) {
    if (!out) return KGS_ERR_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_create_error = std::string("no CUDA device (libkgs has no CPU path): ") +
                         (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        cudaGetLastError();
        return KGS_ERR_CUDA;
    }
    if (device < 0 || device >= count) {
        g_create_error = "device ordinal out of range";
        return KGS_ERR_ARG;
    }
    kgs_ctx* h = new kgs_ctx();
    h->device = device;
    auto bail = [&](cudaError_t ce, const char* what) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(ce);
        delete h;
        return KGS_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&h->copy, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaStreamCreateWithFlags(&h->compute, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    for (int b = 0; b < 2; ++b) {
        if ((e = cudaEventCreateWithFlags(&h->batch[b].ready, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
        if ((e = cudaEventCreateWithFlags(&h->batch[b].consumed, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
        if ((e = cudaEventCreate(&h->mark[b])) != cudaSuccess) return bail(e, "cudaEventCreate");
    }
    if ((e = cudaStreamCreateWithFlags(&h->xchg, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaStreamCreateWithFlags(&h->prep, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaEventCreateWithFlags(&h->acc_idle, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    for (int b = 0; b < 2; ++b)
        if ((e = cudaEventCreateWithFlags(&h->acc_clean[b], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&h->fin_done, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    for (int b = 0; b < 2; ++b)
        if ((e = cudaEventCreateWithFlags(&h->region_free[b], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaMalloc(reinterpret_cast<void**>(&h->px_timeout), 256)) != cudaSuccess) return bail(e, "cudaMalloc");
    if ((e = cudaMemset(h->px_timeout, 0, 256)) != cudaSuccess) return bail(e, "cudaMemset");
    if ((e = cudaFuncSetAttribute(k_sort_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSortSmemBytes)) != cudaSuccess)
        return bail(e, "cudaFuncSetAttribute(k_sort_scatter)");
    if (const char* e = getenv("KGS_SLIDE")) h->slide = atoi(e) != 0;
    if (const char* e = getenv("KGS_SLIDE_CHUNK")) { const int v = atoi(e); if (v >= 1 && v <= 4096) h->slide_chunk = (uint32_t)v; }
    if (const char* e = getenv("KGS_ACCUM_WAVES")) { const int v = atoi(e); if (v >= 1 && v <= 16) h->accum_waves = v; }
    if (const char* e = getenv("KGS_DENSE")) h->dense_path = atoi(e) != 0;
    if (const char* e = getenv("KGS_DENSE16")) h->dense16_path = atoi(e) != 0;
    if (const char* e = getenv("KGS_PARSE_PIECE_KB")) { const int v = atoi(e); if (v >= 16 && v <= (1 << 21)) h->mx_piece_bytes = (uint32_t)v << 10; }
    if (const char* e = getenv("KGS_ACCUM_TMA")) h->accum_tma = atoi(e) != 0;
    if (const char* e = getenv("KGS_FINALIZE_GROUPS")) h->groups_path = atoi(e) != 0;
    if (const char* e = getenv("KGS_AUTO_REORDER")) h->auto_reorder = atoi(e) != 0;
    if (const char* e = getenv("KGS_PDL")) h->pdl = atoi(e) != 0;
    if (const char* e = getenv("KGS_PUSH_DEFER")) h->px_defer = atoi(e) != 0;
    if (const char* e = getenv("KGS_PRECLEAR")) h->preclear_on = atoi(e) != 0;
    if (const char* e = getenv("KGS_STAGE_THREADS")) { const int v = atoi(e); if (v >= 0 && v <= 64) h->mx_stage_threads = v; }
    if (const char* e = getenv("KGS_THREAD_CELL_MAX")) { const int v = atoi(e); if (v >= 1 && v <= 65535) h->thread_cell_max = (uint32_t)v; }
    {   // dynamic shared memory of k_accum: the warps' tile buffers (TMA) or the accumulator cache, beside 21.5 KB of run queues
        const int dyn = (int)std::max<size_t>(kTmaSmemBytes, sizeof(AggTable));
        cudaFuncSetAttribute(k_accum<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
        cudaFuncSetAttribute(k_accum<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
        cudaFuncSetAttribute(k_accum<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
        cudaFuncSetAttribute(k_accum<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
    }
    if (const char* e = getenv("KGS_ACCUM_CTAS_PER_SM")) {   // tuning knob for experiments
        const int v = atoi(e);
        if (v >= 1 && v <= 8) h->accum_ctas_per_sm = v;
    }
    *out = h;
    return KGS_OK;
}

int kgs_destroy(kgs_handle h) {
    if (!h) return KGS_OK;
    cudaSetDevice(h->device);
    h->px_pending.on = false;   // a push nobody waited for is dropped with the handle
    cudaStreamSynchronize(h->copy);
    cudaStreamSynchronize(h->compute);
    if (h->xchg) cudaStreamSynchronize(h->xchg);
    if (h->prep) { cudaStreamSynchronize(h->prep); cudaStreamDestroy(h->prep); }
    if (h->acc_idle) cudaEventDestroy(h->acc_idle);
    for (int b = 0; b < 2; ++b)
        if (h->acc_clean[b]) cudaEventDestroy(h->acc_clean[b]);
    if (h->nccl_comm && nccl_api().ok) nccl_api().comm_destroy(h->nccl_comm);
    if (h->xchg) cudaStreamDestroy(h->xchg);
    if (h->fin_done) cudaEventDestroy(h->fin_done);
    for (int b = 0; b < 2; ++b)
        if (h->region_free[b]) cudaEventDestroy(h->region_free[b]);
    if (h->px_timeout) cudaFree(h->px_timeout);
    for (auto& t : h->timed) {
        cudaEventDestroy(t.e0);
        cudaEventDestroy(t.e1);
    }
    for (auto e : h->event_pool) cudaEventDestroy(e);
    for (int b = 0; b < 2; ++b) {
        for (int c = 0; c < 4; ++c)
            if (h->batch[b].col[c].p) cudaFree(h->batch[b].col[c].p);
        cudaEventDestroy(h->batch[b].ready);
        cudaEventDestroy(h->batch[b].consumed);
        cudaEventDestroy(h->mark[b]);
    }
    DevBuf* bufs[] = {&h->slots, &h->build_keys, &h->build_pods, &h->arena, &h->vals, &h->desc,
                      &h->tmp, &h->tmp2, &h->tmp3, &h->tmp4, &h->mx_text, &h->mx_cta_trans, &h->mx_cta_counts,
                      &h->mx_chunk_state, &h->mx_chunk_counts, &h->mx_pos, &h->mx_ts, &h->mx_val, &h->mx_off,
                      &h->mx_spans, &h->mx_long, &h->mx_out, &h->mx_ctl};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    for (DevBuf& b : h->sortbuf)
        if (b.p) cudaFree(b.p);
    for (auto e : h->mx_arrived) cudaEventDestroy(e);
    for (int i = 0; i < 3; ++i) {
        if (h->mx_stage[i]) cudaFreeHost(h->mx_stage[i]);
        if (h->mx_stage_free[i]) cudaEventDestroy(h->mx_stage_free[i]);
    }
    cudaStreamDestroy(h->copy);
    cudaStreamDestroy(h->compute);
    delete h;
    return KGS_OK;
}

int kgs_sync(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    KGS_CUDA(h, cudaSetDevice(h->device));
    int rc0 = flush_pending_push(h);
    if (rc0) return rc0;
    KGS_CUDA(h, cudaStreamSynchronize(h->copy));
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    KGS_CUDA(h, cudaStreamSynchronize(h->xchg));
    KGS_CUDA(h, cudaStreamSynchronize(h->prep));
    return KGS_OK;
}

int kgs_alloc_pinned(kgs_handle h, size_t nbytes, void** out) {
    if (!h || !out) return KGS_ERR_ARG;
    KGS_CUDA(h, cudaSetDevice(h->device));
    cudaError_t e = cudaHostAlloc(out, nbytes ? nbytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(h, KGS_ERR_NOMEM, std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
    }
    return KGS_OK;
}

int kgs_free_pinned(kgs_handle h, void* p) {
    if (!h) return KGS_ERR_ARG;
    if (p) KGS_CUDA(h, cudaFreeHost(p));
    return KGS_OK;
}

int kgs_ingest(kgs_handle h, const uint32_t* ts, const uint32_t* gpu_gid, const uint32_t* pid,
               const uint32_t* util, uint64_t n) {
    if (!h || (n && (!ts || !gpu_gid || !pid || !util))) return fail(h, KGS_ERR_ARG, "kgs_ingest: null column");
    if (n >= 0xFFFFFC00ull) return fail(h, KGS_ERR_ARG, "kgs_ingest: a batch holds fewer than 2^32 - 1024 samples");
    KGS_CUDA(h, cudaSetDevice(h->device));
    const int b = (h->cur + 1) & 1;
    Batch& B = h->batch[b];
    const uint32_t* src[4] = {ts, gpu_gid, pid, util};
    // the slot may still be read by kernels queued on the compute stream
    KGS_CUDA(h, cudaStreamWaitEvent(h->copy, B.consumed, 0));
    for (int c = 0; c < 4; ++c) {
        if (n * 4 + 64 > B.col[c].cap) {
            // growing a slot frees device memory: make sure nothing is still using it
            KGS_CUDA(h, cudaStreamSynchronize(h->compute));
            KGS_CUDA(h, cudaStreamSynchronize(h->copy));
        }
        int rc = reserve(h, B.col[c], n * 4 + 64);
        if (rc) return rc;
        B.ptr[c] = static_cast<const uint32_t*>(B.col[c].p);
        if (n) KGS_CUDA(h, cudaMemcpyAsync(B.col[c].p, src[c], n * 4, cudaMemcpyHostToDevice, h->copy));
    }
    B.n = n;
    B.order_class = h->auto_reorder ? input_order_class(gpu_gid, pid, n) : 0;
    KGS_CUDA(h, cudaEventRecord(B.ready, h->copy));
    h->cur = b;
    h->shadow_valid = false;
    return KGS_OK;
}

int kgs_batch_reserve(kgs_handle h, uint64_t capacity) {
    if (!h) return KGS_ERR_ARG;
    if (capacity >= 0xFFFFFC00ull) return fail(h, KGS_ERR_ARG, "kgs_batch_reserve: a batch holds fewer than 2^32 - 1024 samples");
    KGS_CUDA(h, cudaSetDevice(h->device));
    const int b = (h->cur + 1) & 1;
    Batch& B = h->batch[b];
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));   // the slot may be freed and reallocated
    KGS_CUDA(h, cudaStreamSynchronize(h->copy));
    for (int c = 0; c < 4; ++c) {
        int rc = reserve(h, B.col[c], capacity * 4 + 64);
        if (rc) return rc;
        B.ptr[c] = static_cast<const uint32_t*>(B.col[c].p);
    }
    B.n = 0;
    B.order_class = 0;
    KGS_CUDA(h, cudaEventRecord(B.ready, h->copy));
    h->cur = b;
    h->shadow_valid = false;
    return KGS_OK;
}

int kgs_batch_append(kgs_handle h, const uint32_t* ts, const uint32_t* gpu_gid, const uint32_t* pid,
                     const uint32_t* util, uint64_t n) {
    if (!h || (n && (!ts || !gpu_gid || !pid || !util))) return fail(h, KGS_ERR_ARG, "kgs_batch_append: null column");
    if (h->cur < 0) return fail(h, KGS_ERR_STATE, "kgs_batch_append: no resident batch (call kgs_batch_reserve)");
    KGS_CUDA(h, cudaSetDevice(h->device));
    Batch& B = h->batch[h->cur];
    const uint32_t* src[4] = {ts, gpu_gid, pid, util};
    for (int c = 0; c < 4; ++c) {
        if (B.col[c].p != B.ptr[c] || (B.n + n) * 4 + 64 > B.col[c].cap)
            return fail(h, KGS_ERR_CAPACITY, "kgs_batch_append: beyond the capacity given to kgs_batch_reserve");
    }
    // kernels queued on the compute stream may be reading the part already there; appending does not touch it
    for (int c = 0; c < 4; ++c)
        if (n) KGS_CUDA(h, cudaMemcpyAsync(static_cast<uint32_t*>(B.col[c].p) + B.n, src[c], n * 4, cudaMemcpyHostToDevice, h->copy));
    if (B.n == 0 && n) B.order_class = h->auto_reorder ? input_order_class(gpu_gid, pid, n) : 0;
    B.n += n;
    h->shadow_valid = false;
    KGS_CUDA(h, cudaEventRecord(B.ready, h->copy));
    return KGS_OK;
}

// Stable radix sort of the resident batch by (gpu_gid, pid) into the OTHER batch slot.  switch_cur: the
// sorted copy becomes the resident batch (kgs_reorder_batch); otherwise the resident batch stays what
// the caller ingested and the copy is kept beside it for kgs_stats_accumulate (shadow_valid).
static int reorder_impl(kgs_handle h, bool switch_cur) {
    KGS_CUDA(h, cudaSetDevice(h->device));
    Batch& src = h->batch[h->cur];
    const uint64_t n = src.n;
    if (n < 2) return KGS_OK;
    cudaStream_t st = h->compute;
    KGS_CUDA(h, cudaStreamWaitEvent(st, src.ready, 0));
    // how many key bits does this batch use?
    int rc;
    KGS_CUDA(h, cudaStreamSynchronize(st));   // tmp buffers may be reallocated
    if ((rc = reserve(h, h->tmp, 64))) return rc;
    KGS_CUDA(h, cudaMemsetAsync(h->tmp.p, 0, 8, st));
    launch(h, KGS_K_SORT, st, [&] {
        k_sort_maxima<<<grid_for(n, 256 * 16, 148 * 8), 256, 0, st>>>(src.ptr[1], src.ptr[2], (uint32_t)n, static_cast<uint32_t*>(h->tmp.p));
    });
    uint32_t mx[2] = {0, 0};
    KGS_CUDA(h, cudaMemcpyAsync(mx, h->tmp.p, 8, cudaMemcpyDeviceToHost, st));
    KGS_CUDA(h, cudaStreamSynchronize(st));
    auto bits = [](uint32_t v) { uint32_t b = 1; while (b < 32 && (v >> b)) ++b; return b; };
    const uint32_t pid_bits = bits(mx[1]), total_bits = bits(mx[0]) + pid_bits;
    const uint32_t passes = (total_bits + 7) / 8;
    // ping-pong between the other batch slot and a scratch set so that the last pass lands in the slot
    Batch& dst = h->batch[(h->cur + 1) & 1];
    KGS_CUDA(h, cudaStreamSynchronize(h->copy));
    uint32_t* slot_cols[4];
    uint32_t* scratch_cols[4];
    for (int c = 0; c < 4; ++c) {
        if ((rc = reserve(h, dst.col[c], n * 4 + 64))) return rc;
        slot_cols[c] = static_cast<uint32_t*>(dst.col[c].p);
        if (passes > 1 && (rc = reserve(h, h->sortbuf[c], n * 4 + 64))) return rc;
        scratch_cols[c] = static_cast<uint32_t*>(h->sortbuf[c].p);
    }
    SortParams P{};
    P.n = (uint32_t)n;
    P.n_ctas = (uint32_t)((n + kSortBlock - 1) / kSortBlock);
    P.pid_bits = pid_bits;
    const uint32_t m = 256 * P.n_ctas, n_part = (m + kScanBlock - 1) / kScanBlock;
    if ((rc = reserve(h, h->tmp2, (size_t)m * 4))) return rc;
    if ((rc = reserve(h, h->tmp3, (size_t)n_part * 4 + 4))) return rc;
    P.counts = static_cast<uint32_t*>(h->tmp2.p);
    uint32_t* partial = static_cast<uint32_t*>(h->tmp3.p);
    const uint32_t* cur_in[4] = {src.ptr[0], src.ptr[1], src.ptr[2], src.ptr[3]};
    for (uint32_t pass = 0; pass < passes; ++pass) {
        const bool to_slot = ((passes - 1 - pass) & 1u) == 0;   // the last pass writes the slot
        for (int c = 0; c < 4; ++c) {
            P.in[c] = cur_in[c];
            P.out[c] = to_slot ? slot_cols[c] : scratch_cols[c];
        }
        P.shift = 8 * pass;
        launch(h, KGS_K_SORT, st, [&] { k_sort_hist<<<P.n_ctas, kSortThreads, 0, st>>>(P); });
        launch(h, KGS_K_SORT, st, [&] { k_scan_reduce<<<n_part, 256, 0, st>>>(P.counts, m, partial); });
        launch(h, KGS_K_SORT, st, [&] { k_scan_top<<<1, 1024, 0, st>>>(partial, n_part); });
        launch(h, KGS_K_SORT, st, [&] { k_scan_apply<<<n_part, 256, 0, st>>>(P.counts, m, partial); });
        launch(h, KGS_K_SORT, st, [&] { k_sort_scatter<<<P.n_ctas, kSortThreads, kSortSmemBytes, st>>>(P); });
        for (int c = 0; c < 4; ++c) cur_in[c] = P.out[c];
    }
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaEventRecord(src.consumed, st));
    for (int c = 0; c < 4; ++c) dst.ptr[c] = slot_cols[c];
    dst.n = n;
    dst.order_class = 0;
    KGS_CUDA(h, cudaEventRecord(dst.ready, st));
    if (switch_cur) {
        h->cur = (h->cur + 1) & 1;
        h->shadow_valid = false;
    } else {
        h->shadow_valid = true;
    }
    return KGS_OK;
}

int kgs_reorder_batch(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    if (h->cur < 0) return fail(h, KGS_ERR_STATE, "kgs_reorder_batch: no resident batch");
    return reorder_impl(h, true);
}

int kgs_input_order_class(const uint32_t* gpu_gid, const uint32_t* pid, uint64_t n) {
    if (n && (!gpu_gid || !pid)) return -1;
    return input_order_class(gpu_gid, pid, n);
}

int kgs_set_auto_reorder(kgs_handle h, int on) {
    if (!h) return KGS_ERR_ARG;
    h->auto_reorder = on != 0;
    return KGS_OK;
}

int kgs_fetch_batch(kgs_handle h, uint32_t* ts, uint32_t* gpu_gid, uint32_t* pid, uint32_t* util) {
    if (!h) return KGS_ERR_ARG;
    if (h->cur < 0) return fail(h, KGS_ERR_STATE, "kgs_fetch_batch: no resident batch");
    KGS_CUDA(h, cudaSetDevice(h->device));
    Batch& B = h->batch[h->cur];
    uint32_t* dst[4] = {ts, gpu_gid, pid, util};
    KGS_CUDA(h, cudaStreamWaitEvent(h->compute, B.ready, 0));
    for (int c = 0; c < 4; ++c)
        if (dst[c] && B.n) KGS_CUDA(h, cudaMemcpyAsync(dst[c], B.ptr[c], B.n * 4, cudaMemcpyDeviceToHost, h->compute));
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    return KGS_OK;
}

int kgs_adopt_device_batch(kgs_handle h, const uint32_t* d_ts, const uint32_t* d_gpu_gid,
                           const uint32_t* d_pid, const uint32_t* d_util, uint64_t n) {
    if (!h || (n && (!d_ts || !d_gpu_gid || !d_pid || !d_util))) return fail(h, KGS_ERR_ARG, "kgs_adopt_device_batch: null column");
    if (n >= 0xFFFFFC00ull) return fail(h, KGS_ERR_ARG, "a batch holds fewer than 2^32 - 1024 samples");
    const uintptr_t all = (uintptr_t)d_ts | (uintptr_t)d_gpu_gid | (uintptr_t)d_pid | (uintptr_t)d_util;
    if (all & 31) return fail(h, KGS_ERR_ARG, "device columns must be 32-byte aligned");
    KGS_CUDA(h, cudaSetDevice(h->device));
    const int b = (h->cur + 1) & 1;
    Batch& B = h->batch[b];
    B.ptr[0] = d_ts; B.ptr[1] = d_gpu_gid; B.ptr[2] = d_pid; B.ptr[3] = d_util;
    B.n = n;
    B.order_class = 0;   // device columns: the host cannot look; the caller may call kgs_reorder_batch
    KGS_CUDA(h, cudaEventRecord(B.ready, h->copy));
    h->cur = b;
    h->shadow_valid = false;
    return KGS_OK;
}

int kgs_build_join(kgs_handle h, const uint64_t* keys, const uint32_t* pod_ids, uint64_t m, uint32_t n_pods) {
    if (!h || (m && (!keys || !pod_ids))) return fail(h, KGS_ERR_ARG, "kgs_build_join: null array");
    if (m >= (1ull << 30)) return fail(h, KGS_ERR_ARG, "kgs_build_join: at most 2^30 - 1 pairs");
    KGS_CUDA(h, cudaSetDevice(h->device));
    for (uint64_t i = 0; i < m; ++i)
        if (keys[i] == kEmptyKey) return fail(h, KGS_ERR_ARG, "kgs_build_join: key 0xFFFFFFFFFFFFFFFF is reserved");
    uint32_t n_slots = 1024;
    while (n_slots < 2 * m) n_slots <<= 1;
    // the old table may still be read by queued kernels
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    int rc;
    if ((rc = reserve(h, h->slots, (size_t)n_slots * sizeof(JoinSlot)))) return rc;
    if ((rc = reserve(h, h->build_keys, m * 8 + 8))) return rc;
    if ((rc = reserve(h, h->build_pods, m * 4 + 4))) return rc;
    cudaStream_t st = h->compute;
    if (m) {
        KGS_CUDA(h, cudaMemcpyAsync(h->build_keys.p, keys, m * 8, cudaMemcpyHostToDevice, st));
        KGS_CUDA(h, cudaMemcpyAsync(h->build_pods.p, pod_ids, m * 4, cudaMemcpyHostToDevice, st));
    }
    JoinSlot* slots = static_cast<JoinSlot*>(h->slots.p);
    launch(h, KGS_K_CLEAR, st, [&] { k_join_clear<<<grid_for(n_slots, 256, 148 * 8), 256, 0, st>>>(slots, n_slots); });
    if (m) {
        launch(h, KGS_K_BUILD, st, [&] {
            k_join_build<<<grid_for(m, 256, 148 * 8), 256, 0, st>>>(
                slots, n_slots - 1, static_cast<const unsigned long long*>(h->build_keys.p), (uint32_t)m);
        });
        launch(h, KGS_K_BUILD_FIN, st, [&] {
            k_join_resolve<<<grid_for(n_slots, 256, 148 * 8), 256, 0, st>>>(
                slots, n_slots, static_cast<const uint32_t*>(h->build_pods.p));
        });
    }
    KGS_CUDA(h, cudaGetLastError());
    // host arrays are pageable in general: do not return before they have been consumed
    KGS_CUDA(h, cudaStreamSynchronize(st));
    h->n_slots = n_slots;
    h->n_pods = n_pods;
    h->has_table = true;
    return KGS_OK;
}

int kgs_probe(kgs_handle h, uint32_t gpus_per_node, uint32_t* out_pod_id) {
    if (!h) return KGS_ERR_ARG;
    if (!h->has_table) return fail(h, KGS_ERR_STATE, "kgs_probe: no join table (call kgs_build_join)");
    if (h->cur < 0) return fail(h, KGS_ERR_STATE, "kgs_probe: no resident batch (call kgs_ingest)");
    if (gpus_per_node == 0) return fail(h, KGS_ERR_ARG, "kgs_probe: gpus_per_node is 0");
    KGS_CUDA(h, cudaSetDevice(h->device));
    Batch& B = h->batch[h->cur];
    if (B.n == 0) return KGS_OK;
    if (!out_pod_id) return fail(h, KGS_ERR_ARG, "kgs_probe: null output");
    int rc;
    if ((rc = reserve(h, h->tmp, B.n * 4 + 16))) return rc;
    cudaStream_t st = h->compute;
    KGS_CUDA(h, cudaStreamWaitEvent(st, B.ready, 0));
    launch(h, KGS_K_PROBE, st, [&] {
        k_join_probe<<<grid_for(B.n / 4 + 1, 256, 148 * 8), 256, 0, st>>>(
            static_cast<const JoinSlot*>(h->slots.p), h->n_slots - 1, B.ptr[1], B.ptr[2], B.n,
            make_fastdiv(gpus_per_node), static_cast<uint32_t*>(h->tmp.p));
    });
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaEventRecord(B.consumed, st));
    KGS_CUDA(h, cudaMemcpyAsync(out_pod_id, h->tmp.p, B.n * 4, cudaMemcpyDeviceToHost, st));
    KGS_CUDA(h, cudaStreamSynchronize(st));
    return KGS_OK;
}

int kgs_join_probe_noalloc(int device, const uint64_t* d_keys, const uint32_t* d_pod_ids, uint64_t m,
                           const uint32_t* d_gpu_gid, const uint32_t* d_pid, uint64_t n, uint32_t gpus_per_node,
                           uint32_t* d_out_pod_id, void* stream) {
    auto bad = [](int code, const std::string& msg) {
        g_create_error = msg;
        return code;
    };
    if (device < 0 || device >= kNoallocMaxDevices) return bad(KGS_ERR_ARG, "kgs_join_probe_noalloc: device ordinal out of range");
    if (m > kNoallocMaxPairs)
        return bad(KGS_ERR_CAPACITY, "kgs_join_probe_noalloc: " + std::to_string(m) + " join pairs exceed the fixed table's limit of " +
                                         std::to_string(kNoallocMaxPairs) + " (KGS_NOALLOC_MAX_PAIRS); use kgs_build_join + kgs_probe");
    if (gpus_per_node == 0) return bad(KGS_ERR_ARG, "kgs_join_probe_noalloc: gpus_per_node is 0");
    if (n >= (1ull << 32)) return bad(KGS_ERR_ARG, "kgs_join_probe_noalloc: fewer than 2^32 samples per call");
    if ((m && (!d_keys || !d_pod_ids)) || (n && (!d_gpu_gid || !d_pid || !d_out_pod_id)))
        return bad(KGS_ERR_ARG, "kgs_join_probe_noalloc: null device pointer");
    if (((uintptr_t)d_gpu_gid | (uintptr_t)d_pid | (uintptr_t)d_out_pod_id) & 15)
        return bad(KGS_ERR_ARG, "kgs_join_probe_noalloc: sample columns and output must be 16-byte aligned");
    // the caller's current device is put back on every way out
    int prev_device = -1;
    cudaError_t e = cudaGetDevice(&prev_device);
    if (e != cudaSuccess) return bad(KGS_ERR_CUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    struct RestoreDevice {
        int dev;
        ~RestoreDevice() { if (dev >= 0) cudaSetDevice(dev); }
    } restore{prev_device == device ? -1 : prev_device};
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return bad(KGS_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    JoinSlot* slots = nullptr;   // the real device address of the module's table (set-up: once per device)
    e = noalloc_table(device, &slots);
    if (e != cudaSuccess) return bad(KGS_ERR_CUDA, std::string("cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint32_t mask = kNoallocSlots - 1;
    k_join_clear<<<grid_for(kNoallocSlots, 256, 148 * 8), 256, 0, st>>>(slots, kNoallocSlots);
    if (m) {
        k_join_build<<<grid_for(m, 256, 148 * 8), 256, 0, st>>>(slots, mask, reinterpret_cast<const unsigned long long*>(d_keys), (uint32_t)m);
        k_join_resolve<<<grid_for(kNoallocSlots, 256, 148 * 8), 256, 0, st>>>(slots, kNoallocSlots, d_pod_ids);
    }
    if (n)
        k_join_probe<<<grid_for(n / 4 + 1, 256, 148 * 8), 256, 0, st>>>(slots, mask, d_gpu_gid, d_pid, n, make_fastdiv(gpus_per_node), d_out_pod_id);
    e = cudaGetLastError();
    if (e != cudaSuccess) return bad(KGS_ERR_CUDA, std::string("kernel launch: ") + cudaGetErrorString(e));
    return KGS_OK;
}

int kgs_set_run_log_capacity(kgs_handle h, uint64_t entries) {
    if (!h) return KGS_ERR_ARG;
    h->run_cap_override = entries;
    return KGS_OK;
}

int kgs_stats_begin(kgs_handle h, const kgs_window_spec* sp) {
    if (!h || !sp) return KGS_ERR_ARG;
    if (sp->key_mask == 0 || sp->key_mask > 7) return fail(h, KGS_ERR_ARG, "key_mask must select 1..3 key kinds");
    if (sp->metric > KGS_METRIC_FB) return fail(h, KGS_ERR_ARG, "unknown metric");
    if (sp->step == 0 || sp->width == 0 || (sp->width > sp->step && sp->width % sp->step))
        return fail(h, KGS_ERR_ARG, "width must be a positive multiple of step, or shorter than step");
    if (sp->n_windows == 0) return fail(h, KGS_ERR_ARG, "n_windows is 0");
    if (sp->gpus_per_node == 0) return fail(h, KGS_ERR_ARG, "gpus_per_node is 0");
    if (sp->n_q > KGS_MAX_Q) return fail(h, KGS_ERR_ARG, "too many percentiles");
    for (uint32_t i = 0; i < sp->n_q; ++i)
        if (!(sp->q[i] >= 0.0 && sp->q[i] <= 1.0)) return fail(h, KGS_ERR_ARG, "percentile outside [0,1]");
    if ((sp->key_mask & 1) && !h->has_table) return fail(h, KGS_ERR_STATE, "per-pod statistics need kgs_build_join first");
    KGS_CUDA(h, cudaSetDevice(h->device));
    const uint32_t W = sp->width < sp->step ? 1u : sp->width / sp->step;
    const uint64_t n_panes = (uint64_t)sp->n_windows + W - 1;
    const uint32_t nk_in[3] = {h->n_pods, sp->n_gpus, sp->n_nodes};
    uint64_t total_hint = sp->total_samples_hint;
    if (!total_hint && h->cur >= 0) total_hint = h->batch[h->cur].n;
    const bool pct = sp->n_q > 0;
    // arena layout: [ctl | accumulators ...] cleared at begin, then results, then the big list
    size_t off = 256, n_kinds = 0;
    struct Lay { size_t cnt, sum, mx, head, oc, os, om, omean, opct; } lay[3];
    for (int k = 0; k < 3; ++k) {
        if (!(sp->key_mask & (1u << k))) continue;
        ++n_kinds;
        const uint64_t pc = (uint64_t)nk_in[k] * n_panes;
        if (pc >= (1ull << 31)) return fail(h, KGS_ERR_ARG, "n_keys * n_panes must stay below 2^31 per key kind");
        lay[k].sum = off; off = align_up(off + pc * 8, 256);
        lay[k].cnt = off; off = align_up(off + pc * 4, 256);
        lay[k].mx = off; off = align_up(off + pc * 4, 256);
        lay[k].head = off; if (pct) off = align_up(off + pc * 4, 256);
    }
    // one byte per window cell: the finalize pass (1..255) in which k_finalize_groups finished the cell
    uint64_t window_cells = 0;
    for (int k = 0; k < 3; ++k)
        if (sp->key_mask & (1u << k)) window_cells += (uint64_t)nk_in[k] * sp->n_windows;
    const size_t done_off = off;
    if (pct) off = align_up(off + window_cells, 256);
    const size_t clear_bytes = off;
    for (int k = 0; k < 3; ++k) {
        if (!(sp->key_mask & (1u << k))) continue;
        const uint64_t wc = (uint64_t)nk_in[k] * sp->n_windows;
        // exchange part (what another shard needs): sum, percentiles, count, max = 24 B/cell at one percentile
        lay[k].os = off; off = align_up(off + wc * 8, 256);
        lay[k].opct = off; off = align_up(off + wc * 8 * sp->n_q, 256);
        lay[k].oc = off; off = align_up(off + wc * 4, 256);
        lay[k].om = off; off = align_up(off + wc * 4, 256);
    }
    const size_t exchange_end = off;
    for (int k = 0; k < 3; ++k) {   // means follow the exchange part (derivable from sum and count, never exchanged)
        if (!(sp->key_mask & (1u << k))) continue;
        lay[k].omean = off; off = align_up(off + (uint64_t)nk_in[k] * sp->n_windows * 8, 256);
    }
    // optional packed exchange tables: sum (u32 if a cell's sum provably fits, else u64), count u32,
    // max u16, and per percentile the two selected values (u8 pair, u16 pair for fb_mib)
    const bool packed = h->packed_exchange != 0;
    const uint64_t metric_max = sp->metric == KGS_METRIC_FB ? 65535u : 255u;
    const bool sum64 = total_hint * metric_max >= (1ull << 32);
    // compact tables (8-bit metrics): count u16, sum 24 bits as u16 + u8, max u8 - 6 + 2*n_q bytes per cell; the
    // few cells above 65 534 samples go on an escape list at the head of the region
    const bool compact = h->packed_exchange == 2 && sp->metric != KGS_METRIC_FB;
    const size_t rank_bytes = sp->metric == KGS_METRIC_FB ? 4 : 2;
    struct PLay { size_t sum, sum_hi, cnt, mx, rank; } play[3];
    const size_t packed_begin = off;
    size_t esc_off = 0;
    uint64_t esc_cap = 0;
    if (packed && compact) {
        esc_cap = window_cells / 64 + 1024;
        esc_off = off;
        off = align_up(off + 16 + esc_cap * 16, 256);
    }
    if (packed) {
        for (int k = 0; k < 3; ++k) {
            if (!(sp->key_mask & (1u << k))) continue;
            const uint64_t wc = (uint64_t)nk_in[k] * sp->n_windows;
            play[k].sum_hi = 0;
            if (compact) {
                play[k].sum = off; off = align_up(off + wc * 2, 256);
                play[k].sum_hi = off; off = align_up(off + wc, 256);
                play[k].cnt = off; off = align_up(off + wc * 2, 256);
                play[k].mx = off; off = align_up(off + wc, 256);
            } else {
                play[k].sum = off; off = align_up(off + wc * (sum64 ? 8 : 4), 256);
                play[k].cnt = off; off = align_up(off + wc * 4, 256);
                play[k].mx = off; off = align_up(off + wc * 2, 256);
            }
            play[k].rank = off; off = align_up(off + wc * rank_bytes * sp->n_q, 256);
        }
    }
    const size_t packed_end = off;
    // two result slots, used by alternate jobs: a caller may still be moving job i's tables (to the
    // host, to other GPUs) while job i+1 is being computed
    const size_t result_area = off - clear_bytes;
    off += result_area;
    const int half = (int)(h->job_seq++ & 1);
    const size_t slot_shift = half ? result_area : 0;
    h->region_half = half;   // (kgs_stats_finalize waits for the exchange of the job two back, which read this half)
    // the rare list has one entry per window cell (a cell is listed at most once), so it cannot overflow
    if (window_cells >= (1ull << 32)) return fail(h, KGS_ERR_ARG, "too many window cells");
    const uint64_t rare_cap = pct ? window_cells : 0;
    const size_t rare_off = off;
    off = align_up(off + (size_t)rare_cap * 8, 256);
    // the second half of the accumulators (jobs alternate; see kgs_ctx::pre).  Not beside an exchange among more than
    // four GPUs: there the window beside the finalize kernels is already full with the exchange copies, and the
    // memset pushes them into the next accumulate kernel (8 GPUs: 0.1899 ms per step with it, 0.1857 without).
    const bool preclear = h->preclear_on && h->px_n <= 4;
    const size_t acc1_off = off;
    if (preclear) off = align_up(off + clear_bytes, 256);
    const size_t acc_shift = (preclear && half) ? acc1_off : 0;
    h->acc_shift_other = (preclear && !half) ? acc1_off : 0;
    // queued kernels may still use the previous job's arena
    if (off > h->arena.cap) {
        KGS_CUDA(h, cudaStreamSynchronize(h->compute));
        KGS_CUDA(h, cudaStreamSynchronize(h->xchg));
        KGS_CUDA(h, cudaStreamSynchronize(h->prep));
    }
    int rc;
    if ((rc = reserve(h, h->arena, off))) return rc;
    char* base = static_cast<char*>(h->arena.p);
    char* acc = base + acc_shift;   // this job's accumulators, control words and done flags
    h->ctl = reinterpret_cast<uint32_t*>(acc);
    for (int k = 0; k < 3; ++k) {
        KindState& K = h->kind[k];
        K = KindState();
        if (!(sp->key_mask & (1u << k))) continue;
        K.n_keys = nk_in[k];
        K.sum = reinterpret_cast<unsigned long long*>(acc + lay[k].sum);
        K.cnt = reinterpret_cast<uint32_t*>(acc + lay[k].cnt);
        K.mx = reinterpret_cast<uint32_t*>(acc + lay[k].mx);
        K.head = pct ? reinterpret_cast<uint32_t*>(acc + lay[k].head) : nullptr;
        K.o_count = reinterpret_cast<uint32_t*>(base + slot_shift + lay[k].oc);
        K.o_sum = reinterpret_cast<unsigned long long*>(base + slot_shift + lay[k].os);
        K.o_mean = reinterpret_cast<double*>(base + slot_shift + lay[k].omean);
        K.o_pct = reinterpret_cast<double*>(base + slot_shift + lay[k].opct);
        K.o_max = reinterpret_cast<uint32_t*>(base + slot_shift + lay[k].om);
        if (packed) {
            K.p_sum = base + slot_shift + play[k].sum;
            K.p_sum_hi = compact ? reinterpret_cast<uint8_t*>(base + slot_shift + play[k].sum_hi) : nullptr;
            K.p_count = base + slot_shift + play[k].cnt;
            K.p_max = base + slot_shift + play[k].mx;
            K.p_rank = sp->n_q ? base + slot_shift + play[k].rank : nullptr;
        }
    }
    h->result_off = clear_bytes + slot_shift;
    h->result_bytes = result_area;
    h->exchange_bytes = exchange_end - clear_bytes;
    h->packed_on = packed;
    h->packed_sum64 = sum64;
    h->packed_compact = packed && compact;
    h->esc_hdr = (packed && compact) ? reinterpret_cast<uint32_t*>(base + slot_shift + esc_off) : nullptr;
    h->esc_cap = (uint32_t)esc_cap;
    h->packed_off = packed_begin + slot_shift;
    h->packed_bytes = packed_end - packed_begin;
    h->rare_list = reinterpret_cast<uint32_t*>(base + rare_off);
    h->rare_cap = (uint32_t)rare_cap;
    h->done = pct ? reinterpret_cast<uint8_t*>(acc + done_off) : nullptr;
    h->done_bytes = pct ? (size_t)window_cells : 0;
    h->done_gen = 0;
    h->clear_bytes = clear_bytes;
    if (pct) {
        const uint64_t vals_cap = total_hint + 16 * 4096 + 64;     // in samples; fb_mib takes 2 B each
        if (vals_cap >= (1ull << 32)) return fail(h, KGS_ERR_ARG, "percentile jobs hold fewer than 2^32 samples per handle; shard the trace");
        // a run cannot outlast its pane: assume runs of step/2 samples, between 2 and 8 (1 Hz sampling, 60 s panes
        // give 60); a third more for the shared pool
        const uint64_t run_guess = sp->step / 2 < 2 ? 2 : (sp->step / 2 > 8 ? 8 : sp->step / 2);
        uint64_t dcap = h->run_cap_override ? h->run_cap_override : (total_hint / run_guess + 65536) * n_kinds * 4 / 3;
        if (dcap > 0xFFFFFFF0ull) dcap = 0xFFFFFFF0ull;
        if (vals_cap > h->vals.cap || dcap * 16 > h->desc.cap) KGS_CUDA(h, cudaStreamSynchronize(h->compute));
        const uint64_t vbytes = sp->metric == KGS_METRIC_FB ? 2 : 1;
        if (vals_cap * vbytes > h->vals.cap) KGS_CUDA(h, cudaStreamSynchronize(h->compute));
        if ((rc = reserve(h, h->vals, vals_cap * vbytes))) return rc;
        if ((rc = reserve(h, h->desc, dcap * 16))) return rc;
        h->vals_cap = vals_cap;
        h->desc_cap = (uint32_t)dcap;
        h->desc_pool_base = (uint32_t)(dcap / 4);          // entries [0, desc_pool_base): shared overflow pool
        h->desc_priv_next = h->desc_pool_base;             // private slabs follow and grow upwards
    } else {
        h->vals_cap = 0;
        h->desc_cap = 0;
    }
    {   // this job's half: cleared beside the previous job's finalize kernels if that job had the same layout, else now
        kgs_ctx::PreCleared& pc = h->pre[half];
        if (preclear && pc.valid && pc.base == h->arena.p && pc.bytes == clear_bytes && pc.shift == acc_shift)
            KGS_CUDA(h, cudaStreamWaitEvent(h->compute, h->acc_clean[half], 0));
        else
            KGS_CUDA(h, cudaMemsetAsync(acc, 0, clear_bytes, h->compute));
        pc.valid = false;
        h->pre[half ^ 1].valid = false;   // cleared for another job's layout, or in use until this job's kernels have passed it
        h->preclear_pending = preclear;
    }
    h->spec = *sp;
    h->W = W;
    h->n_panes = (uint32_t)n_panes;
    h->vals_used = 0;
    h->accumulated = 0;
    h->job_open = true;
    h->job_final = false;
    return KGS_OK;
}

int kgs_stats_accumulate(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    if (!h->job_open) return fail(h, KGS_ERR_STATE, "kgs_stats_accumulate: no open job (call kgs_stats_begin)");
    if (h->cur < 0) return fail(h, KGS_ERR_STATE, "kgs_stats_accumulate: no resident batch (call kgs_ingest)");
    KGS_CUDA(h, cudaSetDevice(h->device));
    if (h->batch[h->cur].n == 0) return KGS_OK;
    // a batch that interleaves its series - or, for a percentile job, any disordered batch - is accumulated from a
    // series-major copy (same tables, long runs)
    const int oc = h->batch[h->cur].order_class;
    const bool want_sorted = h->auto_reorder && (oc == 2 || (oc == 1 && h->spec.n_q > 0));
    if (want_sorted && !h->shadow_valid && h->batch[h->cur].n >= 2) {
        int rc0 = reorder_impl(h, false);
        if (rc0) return rc0;
    }
    const bool from_shadow = h->shadow_valid && want_sorted;
    Batch& B = h->batch[from_shadow ? ((h->cur + 1) & 1) : h->cur];
    const kgs_window_spec& sp = h->spec;
    const bool pct = sp.n_q > 0;
    if (h->accumulated + B.n >= (1ull << 32))
        return fail(h, KGS_ERR_CAPACITY, "a job holds fewer than 2^32 samples per handle (cell counts are 32-bit); shard the trace");
    if (pct && h->vals_used + B.n + 16 > h->vals_cap) {   // total_samples_hint was too small: enlarge the value log
        const uint64_t vbytes = sp.metric == KGS_METRIC_FB ? 2 : 1;
        uint64_t want = 2 * h->vals_cap;
        if (want < h->vals_used + B.n + 65536) want = h->vals_used + B.n + 65536;
        if (want >= (1ull << 32)) return fail(h, KGS_ERR_CAPACITY, "percentile jobs hold fewer than 2^32 samples per handle; shard the trace");
        int rc2 = grow_copy(h, h->vals, want * vbytes, (size_t)h->vals_used * vbytes);
        if (rc2) return rc2;
        h->vals_cap = want;
    }
    AccumParams P{};
    P.ts = B.ptr[0]; P.gid = B.ptr[1]; P.pid = B.ptr[2]; P.util = B.ptr[3];
    P.n = (uint32_t)B.n;
    P.vals_base = h->vals_used;
    P.slots = static_cast<const JoinSlot*>(h->slots.p);
    P.tmask = h->n_slots ? h->n_slots - 1 : 0;
    P.t0 = sp.t0;
    P.dstep = make_fastdiv(sp.step);
    P.n_panes = h->n_panes;
    P.step = sp.step;
    P.win_len = sp.width < sp.step ? sp.width : sp.step;
    P.dgpn = make_fastdiv(sp.gpus_per_node);
    P.mshift = sp.metric == KGS_METRIC_SM ? 0 : sp.metric == KGS_METRIC_MEM ? 8 : 16;
    P.mmask = sp.metric == KGS_METRIC_FB ? 0xFFFFu : 0xFFu;
    P.key_mask = sp.key_mask;
    for (int k = 0; k < 3; ++k) {
        P.n_keys[k] = h->kind[k].n_keys;
        P.cnt[k] = h->kind[k].cnt;
        P.sum[k] = h->kind[k].sum;
        P.mx[k] = h->kind[k].mx;
        P.head[k] = h->kind[k].head;
    }
    P.vals = static_cast<uint8_t*>(h->vals.p);
    P.desc = static_cast<uint4*>(h->desc.p);
    P.desc_cap = h->desc_cap;
    P.ctl = h->ctl;
    P.agg_on = 1;
    P.dense_ok = (sp.step >= 8 && sp.width >= sp.step && h->dense_path) ? 1u : 0u;
    P.dense16_ok = (P.dense_ok && sp.step >= 16 && sp.metric != KGS_METRIC_FB && h->dense16_path) ? 1u : 0u;
    for (int k = 0; k < 3; ++k)
        if ((uint64_t)h->kind[k].n_keys * h->n_panes >= (1ull << 30)) P.agg_on = 0;
    // tile staging through shared memory (cp.async.bulk): forced by KGS_ACCUM_TMA=1; it replaces the accumulator cache
    const bool use_tma = h->accum_tma;
    if (use_tma) P.agg_on = 0;
    cudaStream_t st = h->compute;
    KGS_CUDA(h, cudaStreamWaitEvent(st, B.ready, 0));
    // every warp owns a contiguous chunk of 256-sample tiles; ctas_per_sm resident CTAs of 8 warps
    const uint64_t n_tiles = (B.n + kTileSamples - 1) / kTileSamples;
    const uint64_t max_warps = 148ull * h->accum_ctas_per_sm * kAccumWarps * h->accum_waves;
    const uint64_t tpw = (n_tiles + max_warps - 1) / max_warps;
    P.tiles_per_warp = (uint32_t)(tpw < 2 ? 2 : tpw);
    const int grid = (int)((n_tiles + (uint64_t)P.tiles_per_warp * kAccumWarps - 1) / ((uint64_t)P.tiles_per_warp * kAccumWarps));
    if (pct) {
        // a private run-log slab per warp for this launch (16 entries per tile and key kind is > 3x what
        // collector-ordered input needs); what does not fit goes to the shared pool
        const uint64_t warps = (uint64_t)grid * kAccumWarps;
        const uint64_t run_guess = sp.step / 2 < 2 ? 2 : (sp.step / 2 > 8 ? 8 : sp.step / 2);
        uint64_t per_warp = (uint64_t)P.tiles_per_warp * (kTileSamples / run_guess) * __builtin_popcount(sp.key_mask);
        if (h->run_cap_override) per_warp = (uint64_t)P.tiles_per_warp * kTileSamples * __builtin_popcount(sp.key_mask);   // every sample its own run
        if ((uint64_t)h->desc_priv_next + per_warp * warps > h->desc_cap) {
            // more (or smaller) batches than the sizing assumed: enlarge the log, entries keep their indices
            uint64_t want = 2ull * h->desc_cap;
            if (want < (uint64_t)h->desc_priv_next + per_warp * warps) want = (uint64_t)h->desc_priv_next + per_warp * warps;
            if (want > 0xFFFFFFF0ull) want = 0xFFFFFFF0ull;
            if ((uint64_t)h->desc_priv_next + per_warp * warps > want) per_warp = (want - h->desc_priv_next) / warps;
            int rc2 = grow_copy(h, h->desc, want * 16, (size_t)h->desc_priv_next * 16);
            if (rc2) return rc2;
            h->desc_cap = (uint32_t)want;
            P.desc = static_cast<uint4*>(h->desc.p);
            P.desc_cap = h->desc_cap;
        }
        P.log_base = h->desc_priv_next;
        P.log_per_warp = (uint32_t)per_warp;
        P.pool_end = h->desc_pool_base;
        h->desc_priv_next += (uint32_t)(per_warp * warps);
    }
    launch(h, KGS_K_ACCUM, st, [&] {
        if (use_tma) {
            if (pct) k_accum<true, true><<<grid, 256, kTmaSmemBytes, st>>>(P);
            else k_accum<false, true><<<grid, 256, kTmaSmemBytes, st>>>(P);
        } else {
            if (pct) k_accum<true, false><<<grid, 256, sizeof(AggTable), st>>>(P);
            else k_accum<false, false><<<grid, 256, sizeof(AggTable), st>>>(P);
        }
    });
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaEventRecord(B.consumed, st));
    if (pct) h->vals_used = align_up(h->vals_used + B.n, 16);
    h->accumulated += B.n;
    h->last_was_accum = true;
    {   // the previous job's tables start travelling now, behind this kernel (kgs_exchange_push)
        int rc3 = flush_pending_push(h);
        if (rc3) return rc3;
    }
    if (h->preclear_pending) {   // the other half of the accumulators (the previous job's) is cleared for the next job, behind this kernel
        h->preclear_pending = false;
        const int other = h->region_half ^ 1;
        KGS_CUDA(h, cudaEventRecord(h->acc_idle, st));
        KGS_CUDA(h, cudaStreamWaitEvent(h->prep, h->acc_idle, 0));
        KGS_CUDA(h, cudaMemsetAsync(static_cast<char*>(h->arena.p) + h->acc_shift_other, 0, h->clear_bytes, h->prep));
        KGS_CUDA(h, cudaEventRecord(h->acc_clean[other], h->prep));
        h->pre[other].valid = true;
        h->pre[other].base = h->arena.p;
        h->pre[other].bytes = h->clear_bytes;
        h->pre[other].shift = h->acc_shift_other;
    }
    return KGS_OK;
}

int kgs_stats_finalize(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    if (!h->job_open) return fail(h, KGS_ERR_STATE, "kgs_stats_finalize: no open job");
    KGS_CUDA(h, cudaSetDevice(h->device));
    const kgs_window_spec& sp = h->spec;
    FinParams P{};
    uint32_t slot = 0, cells = 0;
    for (int k = 0; k < 3; ++k) {
        if (!(sp.key_mask & (1u << k))) continue;
        const KindState& K = h->kind[k];
        FinKind& F = P.kind[slot++];
        F.n_keys = K.n_keys;
        F.cell_begin = cells;
        F.cnt = K.cnt; F.sum = K.sum; F.mx = K.mx; F.head = K.head;
        F.o_count = K.o_count; F.o_sum = K.o_sum; F.o_max = K.o_max; F.o_mean = K.o_mean; F.o_pct = K.o_pct;
        F.p_sum = K.p_sum; F.p_sum_hi = K.p_sum_hi; F.p_count = K.p_count; F.p_max = K.p_max; F.p_rank = K.p_rank;
        const uint64_t wc = (uint64_t)K.n_keys * sp.n_windows;
        if (cells + wc >= (1ull << 32)) return fail(h, KGS_ERR_ARG, "too many window cells");
        cells += (uint32_t)wc;
    }
    for (uint32_t s = slot; s < 3; ++s) P.kind[s].cell_begin = 0xFFFFFFFFu;
    P.n_kinds = slot;
    P.twin_pod = P.twin_gpu = -1;
    if (h->W == 1 && (sp.key_mask & 3u) == 3u) {   // pod is slot 0 and gpu slot 1 when both are on
        P.twin_pod = 0;
        P.twin_gpu = 1;
    }
    P.total_cells = cells;
    P.n_windows = sp.n_windows;
    P.n_panes = h->n_panes;
    P.W = h->W;
    P.n_q = sp.n_q;
    for (uint32_t i = 0; i < sp.n_q; ++i) P.q[i] = sp.q[i];
    P.vals = static_cast<const uint8_t*>(h->vals.p);
    P.desc = static_cast<const uint4*>(h->desc.p);
    P.rare_list = h->rare_list;
    P.rare_cap = h->rare_cap;
    P.thread_cell_max = h->thread_cell_max;
    P.ctl = h->ctl;
    P.u16 = sp.metric == KGS_METRIC_FB ? 1u : 0u;
    P.slide_chunk = h->slide_chunk;
    P.packed_sum64 = h->packed_sum64 ? 1u : 0u;
    P.packed_compact = h->packed_compact ? 1u : 0u;
    P.esc_hdr = h->esc_hdr;
    P.esc = h->esc_hdr ? reinterpret_cast<uint4*>(h->esc_hdr + 4) : nullptr;
    P.esc_cap = h->esc_cap;
    if (h->packed_on && !h->packed_compact && !h->packed_sum64) {
        const unsigned long long metric_max = sp.metric == KGS_METRIC_FB ? 65535ull : 255ull;
        if (h->accumulated * metric_max >= (1ull << 32))
            return fail(h, KGS_ERR_CAPACITY, "packed exchange tables hold 32-bit sums, sized from total_samples_hint; this job accumulated " +
                                                 std::to_string(h->accumulated) + " samples: give a larger hint");
    }
    cudaStream_t st = h->compute;
    P.esc_count = h->ctl + 5;
    // A job may be finalized more than once (more batches in between): the lists of cells left to
    // k_finalize_rare start empty each time.
    if (h->job_final) KGS_CUDA(h, cudaMemsetAsync(h->ctl + 2, 0, 20, st));   // ctl[2..6]: rare-list lengths, spare, escape count, CTAs done
    // This job's result region was last read by the exchange of the job two back: that comes first.  (Nothing before
    // the finalize kernels writes the region, so the wait sits here and not at begin: the exchange of job i may run
    // beside the accumulate kernel of job i+2.)
    int rc_push = flush_pending_push(h);
    if (rc_push) return rc_push;
    if (h->region_busy[h->region_half]) {
        KGS_CUDA(h, cudaStreamWaitEvent(h->compute, h->region_free[h->region_half], 0));
        h->region_busy[h->region_half] = false;
    }
    // PDL only where the predecessor on the stream is the library's own kernel and no per-kernel events sit in between
    const bool pdl = h->pdl && !h->prof && h->last_was_accum && !h->job_final;
    h->last_was_accum = false;
    // k_finalize_groups: tumbling (or sampled) windows of an 8-bit metric, per-GPU table part of the job
    const int slot_of[3] = {(sp.key_mask & 1u) ? 0 : -1, (sp.key_mask & 2u) ? (int)__builtin_popcount(sp.key_mask & 1u) : -1,
                            (sp.key_mask & 4u) ? (int)__builtin_popcount(sp.key_mask & 3u) : -1};
    bool groups = h->groups_path && cells && sp.n_q && h->W == 1 && sp.metric != KGS_METRIC_FB && slot_of[1] >= 0;
    if (groups) {
        uint32_t gs_log2 = 0;
        while ((1u << gs_log2) < sp.gpus_per_node && gs_log2 < 5) ++gs_log2;
        const bool node_fits = (1u << gs_log2) >= sp.gpus_per_node;      // else: no node cooperation, one lane per GPU
        P.g_gpu = slot_of[1];
        P.g_pod = slot_of[0];
        P.g_node = node_fits ? slot_of[2] : -1;
        if (!node_fits) gs_log2 = 0;
        P.g_gs_log2 = gs_log2;
        P.g_gpn = node_fits ? sp.gpus_per_node : 1u;
        uint64_t nodes = node_fits ? ((uint64_t)sp.n_gpus + sp.gpus_per_node - 1) / sp.gpus_per_node : sp.n_gpus;
        if (P.g_node >= 0 && sp.n_nodes > nodes) nodes = sp.n_nodes;
        P.g_nodes = (uint32_t)nodes;
        const uint64_t threads = (nodes * sp.n_windows) << gs_log2;
        const uint64_t grid = (threads + kGroupThreads - 1) / kGroupThreads;
        if (grid == 0 || grid > 0x7FFFFFFFull) groups = false;
        if (groups) {
            if (++h->done_gen > 255) {   // the stamp is a byte: start over with clean flags
                KGS_CUDA(h, cudaMemsetAsync(h->done, 0, h->done_bytes, st));
                h->done_gen = 1;
            }
            P.done = h->done;
            P.done_gen = h->done_gen;
            launch(h, KGS_K_FINALIZE_GROUPS, st, [&] {
                switch (gs_log2) {
                    case 0: launch_after(k_finalize_groups<1>, (unsigned)grid, kGroupThreads, st, pdl, P); break;
                    case 1: launch_after(k_finalize_groups<2>, (unsigned)grid, kGroupThreads, st, pdl, P); break;
                    case 2: launch_after(k_finalize_groups<4>, (unsigned)grid, kGroupThreads, st, pdl, P); break;
                    case 3: launch_after(k_finalize_groups<8>, (unsigned)grid, kGroupThreads, st, pdl, P); break;
                    case 4: launch_after(k_finalize_groups<16>, (unsigned)grid, kGroupThreads, st, pdl, P); break;
                    default: launch_after(k_finalize_groups<32>, (unsigned)grid, kGroupThreads, st, pdl, P); break;
                }
            });
        }
    }
    if (!groups) P.done = nullptr;
    bool rare_launched = false;   // "the escape list's length has a kernel that publishes it" (k_finalize_rare, k_finalize_cells)
    if (cells) {
        if (h->W > 1 && h->slide && sp.metric != KGS_METRIC_FB) {
            // sliding windows: per-cell basics, then one thread per chunk of windows keeps the histogram sliding
            launch(h, KGS_K_FINALIZE_BASIC, st, [&] { k_finalize_basic<<<(cells + 255) / 256, 256, 0, st>>>(P); });
            if (sp.n_q) {
                const uint64_t chunks_per_key = (sp.n_windows + h->slide_chunk - 1) / h->slide_chunk;
                uint64_t threads = 0;
                for (uint32_t k = 0; k < P.n_kinds; ++k) threads += (uint64_t)P.kind[k].n_keys * chunks_per_key;
                const unsigned grid = (unsigned)((threads + CellLayout::NT - 1) / CellLayout::NT);
                if (grid) launch(h, KGS_K_FINALIZE, st, [&] { k_finalize_slide<<<grid, CellLayout::NT, 0, st>>>(P); });
            }
        } else {
            rare_launched = !(sp.n_q && sp.metric == KGS_METRIC_FB);   // k_finalize_cells publishes the escape list's length itself
            launch(h, KGS_K_FINALIZE, st, [&] {   // at most one wave of CTAs (7 per SM: 30.7 KB of shared memory each), each walking blocks of cells
                launch_after(k_finalize_cells, (unsigned)grid_for(cells, CellLayout::NT, 148 * 7), CellLayout::NT, st, pdl, P);
            });
        }
        if (sp.n_q && sp.metric == KGS_METRIC_FB) {   // 16-bit values: two-digit radix select, warp per cell
            launch(h, KGS_K_FINALIZE_WIDE, st, [&] { k_finalize_u16<<<148 * 8, 128, 0, st>>>(P); });
        } else if (sp.n_q && h->W > 1 && h->slide) {   // sliding windows: one launch for the windows the sliding histograms could not take
            rare_launched = true;
            const int wide_grid = grid_for(cells, WideLayout::NT, 148 * 6);
            const int big_grid = h->accumulated * h->W > h->thread_cell_max ? 148 * 4 : 0;
            P.rare_wide_ctas = (uint32_t)wide_grid;
            launch(h, KGS_K_FINALIZE_WIDE, st, [&] { launch_after(k_finalize_rare, (unsigned)(wide_grid + big_grid), WideLayout::NT, st, pdl, P); });
        }
    }
    KGS_CUDA(h, cudaGetLastError());
    if (h->packed_compact && !rare_launched)   // k_finalize_rare was not there to do it
        KGS_CUDA(h, cudaMemcpyAsync(h->esc_hdr, h->ctl + 5, 4, cudaMemcpyDeviceToDevice, st));
    h->job_final = true;
    return KGS_OK;
}

int kgs_window_stats(kgs_handle h, const kgs_window_spec* spec) {
    int rc;
    if ((rc = kgs_stats_begin(h, spec))) return rc;
    if ((rc = kgs_stats_accumulate(h))) return rc;
    return kgs_stats_finalize(h);
}

static int check_job_flags(kgs_handle h) {
    uint32_t ctl[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    KGS_CUDA(h, cudaMemcpyAsync(ctl, h->ctl, sizeof(ctl), cudaMemcpyDeviceToHost, h->compute));
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    if (h->packed_compact && h->job_final) {
        const uint32_t n_esc = ctl[5];
        if (n_esc > h->esc_cap)
            return fail(h, KGS_ERR_CAPACITY, "compact exchange tables: " + std::to_string(n_esc) + " cells above 65 534 samples, the escape list holds " +
                                                 std::to_string(h->esc_cap) + " (use kgs_set_packed_exchange(h, 1) for windows this long)");
    }
    if (ctl[1])
        return fail(h, KGS_ERR_CAPACITY,
                    "run log overflow: the shared pool needed " + std::to_string(ctl[0]) + " entries, total capacity " +
                        std::to_string(h->desc_cap) + " (input far from collector order; raise it with kgs_set_run_log_capacity)");
    return KGS_OK;
}

int kgs_job_counter(kgs_handle h, int which, uint64_t* out) {
    if (!h || !out || which < 0 || which >= KGS_N_JOB_COUNTERS) return KGS_ERR_ARG;
    if (!h->job_open) return fail(h, KGS_ERR_STATE, "kgs_job_counter: no job");
    KGS_CUDA(h, cudaSetDevice(h->device));
    uint32_t ctl[8] = {0};
    KGS_CUDA(h, cudaMemcpyAsync(ctl, h->ctl, sizeof(ctl), cudaMemcpyDeviceToHost, h->compute));
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    static const int slot[KGS_N_JOB_COUNTERS] = {0, 4, 2, 3};
    *out = ctl[slot[which]];
    return KGS_OK;
}

int kgs_fetch_stats(kgs_handle h, int key_kind, uint32_t* count, uint64_t* sum, uint32_t* max, double* mean, double* pct) {
    if (!h) return KGS_ERR_ARG;
    if (!h->job_final) return fail(h, KGS_ERR_STATE, "kgs_fetch_stats: no finalized job");
    if (key_kind < 0 || key_kind > 2 || !(h->spec.key_mask & (1u << key_kind))) return fail(h, KGS_ERR_ARG, "key kind not part of this job");
    KGS_CUDA(h, cudaSetDevice(h->device));
    int rc;
    if (h->spec.n_q && (rc = check_job_flags(h))) return rc;
    const KindState& K = h->kind[key_kind];
    const size_t wc = (size_t)K.n_keys * h->spec.n_windows;
    cudaStream_t st = h->compute;
    if (wc) {
        if (count) KGS_CUDA(h, cudaMemcpyAsync(count, K.o_count, wc * 4, cudaMemcpyDeviceToHost, st));
        if (sum) KGS_CUDA(h, cudaMemcpyAsync(sum, K.o_sum, wc * 8, cudaMemcpyDeviceToHost, st));
        if (max) KGS_CUDA(h, cudaMemcpyAsync(max, K.o_max, wc * 4, cudaMemcpyDeviceToHost, st));
        if (mean) KGS_CUDA(h, cudaMemcpyAsync(mean, K.o_mean, wc * 8, cudaMemcpyDeviceToHost, st));
        if (pct && h->spec.n_q) KGS_CUDA(h, cudaMemcpyAsync(pct, K.o_pct, wc * 8 * h->spec.n_q, cudaMemcpyDeviceToHost, st));
    }
    KGS_CUDA(h, cudaStreamSynchronize(st));
    return KGS_OK;
}

int kgs_result_device_ptr(kgs_handle h, int key_kind, int which, void** d_ptr, size_t* nbytes) {
    if (!h || !d_ptr || !nbytes) return KGS_ERR_ARG;
    if (!h->job_final) return fail(h, KGS_ERR_STATE, "kgs_result_device_ptr: no finalized job");
    if (key_kind < 0 || key_kind > 2 || !(h->spec.key_mask & (1u << key_kind))) return fail(h, KGS_ERR_ARG, "key kind not part of this job");
    const KindState& K = h->kind[key_kind];
    const size_t wc = (size_t)K.n_keys * h->spec.n_windows;
    switch (which) {
        case 0: *d_ptr = K.o_count; *nbytes = wc * 4; break;
        case 1: *d_ptr = K.o_sum; *nbytes = wc * 8; break;
        case 2: *d_ptr = K.o_max; *nbytes = wc * 4; break;
        case 3: *d_ptr = K.o_mean; *nbytes = wc * 8; break;
        case 4: *d_ptr = K.o_pct; *nbytes = wc * 8 * h->spec.n_q; break;
        default: return fail(h, KGS_ERR_ARG, "which must be 0..4");
    }
    return KGS_OK;
}

int kgs_result_region(kgs_handle h, void** d_ptr, size_t* exchange_bytes, size_t* total_bytes) {
    if (!h || !d_ptr || !exchange_bytes || !total_bytes) return KGS_ERR_ARG;
    if (!h->job_final) return fail(h, KGS_ERR_STATE, "kgs_result_region: no finalized job");
    *d_ptr = static_cast<char*>(h->arena.p) + h->result_off;
    *exchange_bytes = h->exchange_bytes;
    *total_bytes = h->result_bytes;
    return KGS_OK;
}

int kgs_set_packed_exchange(kgs_handle h, int on) {
    if (!h) return KGS_ERR_ARG;
    if (on < 0 || on > 2) return fail(h, KGS_ERR_ARG, "kgs_set_packed_exchange: 0 (off), 1 (packed) or 2 (compact)");
    h->packed_exchange = on;
    return KGS_OK;
}

int kgs_exchange_region(kgs_handle h, void** d_ptr, size_t* nbytes, int* packed) {
    if (!h || !d_ptr || !nbytes) return KGS_ERR_ARG;
    if (!h->job_final) return fail(h, KGS_ERR_STATE, "kgs_exchange_region: no finalized job");
    if (h->packed_on) {
        *d_ptr = static_cast<char*>(h->arena.p) + h->packed_off;
        *nbytes = h->packed_bytes;
    } else {
        *d_ptr = static_cast<char*>(h->arena.p) + h->result_off;
        *nbytes = h->exchange_bytes;
    }
    if (packed) *packed = h->packed_on ? (h->packed_compact ? 3 : h->packed_sum64 ? 2 : 1) : 0;
    return KGS_OK;
}

int kgs_packed_offset(kgs_handle h, int key_kind, int which, size_t* offset, size_t* nbytes) {
    if (!h || !offset || !nbytes) return KGS_ERR_ARG;
    if (!h->job_final || !h->packed_on) return fail(h, KGS_ERR_STATE, "kgs_packed_offset: no finalized job with packed exchange tables");
    if (key_kind < 0 || key_kind > 2 || !(h->spec.key_mask & (1u << key_kind))) return fail(h, KGS_ERR_ARG, "key kind not part of this job");
    const KindState& K = h->kind[key_kind];
    const size_t wc = (size_t)K.n_keys * h->spec.n_windows;
    const char* base = static_cast<char*>(h->arena.p) + h->packed_off;
    const void* p;
    const bool cp = h->packed_compact;
    switch (which) {
        case 0: p = K.p_count; *nbytes = wc * (cp ? 2 : 4); break;
        case 1: p = K.p_sum; *nbytes = wc * (cp ? 2 : h->packed_sum64 ? 8 : 4); break;
        case 2: p = K.p_max; *nbytes = wc * (cp ? 1 : 2); break;
        case 4: p = K.p_rank ? K.p_rank : K.p_max; *nbytes = K.p_rank ? wc * (h->spec.metric == KGS_METRIC_FB ? 4 : 2) * h->spec.n_q : 0; break;
        case 5: p = cp ? static_cast<const void*>(K.p_sum_hi) : K.p_sum; *nbytes = cp ? wc : 0; break;
        case 6: p = cp ? static_cast<const void*>(h->esc_hdr) : static_cast<const void*>(base); *nbytes = cp ? 16 + (size_t)h->esc_cap * 16 : 0; break;
        default: return fail(h, KGS_ERR_ARG, "which must be 0 (count), 1 (sum), 2 (max), 4 (selected values), 5 (sum, high byte) or 6 (escape list)");
    }
    *offset = (size_t)(static_cast<const char*>(p) - base);
    return KGS_OK;
}

int kgs_result_offset(kgs_handle h, int key_kind, int which, size_t* offset, size_t* nbytes) {
    void* p = nullptr;
    int rc = kgs_result_device_ptr(h, key_kind, which, &p, nbytes);
    if (rc) return rc;
    if (!offset) return KGS_ERR_ARG;
    *offset = (size_t)(static_cast<char*>(p) - (static_cast<char*>(h->arena.p) + h->result_off));
    return KGS_OK;
}

int kgs_fetch_region(kgs_handle h, void* host_dst, size_t nbytes) {
    if (!h || (nbytes && !host_dst)) return KGS_ERR_ARG;
    if (!h->job_final) return fail(h, KGS_ERR_STATE, "kgs_fetch_region: no finalized job");
    if (nbytes > h->result_bytes) return fail(h, KGS_ERR_ARG, "kgs_fetch_region: more bytes than the result region holds");
    KGS_CUDA(h, cudaSetDevice(h->device));
    int rc;
    if (h->spec.n_q && (rc = check_job_flags(h))) return rc;
    if (nbytes)
        KGS_CUDA(h, cudaMemcpyAsync(host_dst, static_cast<char*>(h->arena.p) + h->result_off, nbytes, cudaMemcpyDeviceToHost, h->compute));
    KGS_CUDA(h, cudaStreamSynchronize(h->compute));
    return KGS_OK;
}

// ---- table exchange ----------------------------------------------------------------------------------
static int exchange_source(kgs_handle h, const char* who, char** src, size_t* nbytes) {
    if (!h->job_final) return fail(h, KGS_ERR_STATE, std::string(who) + ": no finalized job");
    if (h->packed_on) {
        *src = static_cast<char*>(h->arena.p) + h->packed_off;
        *nbytes = h->packed_bytes;
    } else {
        *src = static_cast<char*>(h->arena.p) + h->result_off;
        *nbytes = h->exchange_bytes;
    }
    return KGS_OK;
}

int kgs_exchange_setup(kgs_handle h, int n_ranks, int my_rank, void* const* gathered, void* const* flags, size_t slot_bytes) {
    if (!h || !gathered || !flags) return fail(h, KGS_ERR_ARG, "kgs_exchange_setup: null argument");
    if (n_ranks < 1 || n_ranks > kMaxRanks || my_rank < 0 || my_rank >= n_ranks)
        return fail(h, KGS_ERR_ARG, "kgs_exchange_setup: 1 <= n_ranks <= 16 and 0 <= my_rank < n_ranks");
    if (slot_bytes == 0 || (slot_bytes & 15)) return fail(h, KGS_ERR_ARG, "kgs_exchange_setup: slot_bytes must be a positive multiple of 16");
    KGS_CUDA(h, cudaSetDevice(h->device));
    h->px_pending.on = false;
    KGS_CUDA(h, cudaStreamSynchronize(h->xchg));
    for (int r = 0; r < n_ranks; ++r) {
        if (!gathered[r] || !flags[r]) return fail(h, KGS_ERR_ARG, "kgs_exchange_setup: null buffer");
        h->px_gathered[r] = static_cast<char*>(gathered[r]);
        h->px_flags.row[r] = static_cast<uint32_t*>(flags[r]);
    }
    h->px_n = n_ranks;
    h->px_me = my_rank;
    h->px_slot = slot_bytes;
    h->px_seq = 0;
    h->px_multicast = nullptr;
    return KGS_OK;
}

int kgs_exchange_set_self_copy(kgs_handle h, int on) {
    if (!h) return KGS_ERR_ARG;
    h->px_self_copy = on != 0;
    return KGS_OK;
}

int kgs_exchange_set_multicast(kgs_handle h, void* mc_gathered) {
    if (!h) return KGS_ERR_ARG;
    if (h->px_n == 0) return fail(h, KGS_ERR_STATE, "kgs_exchange_set_multicast: call kgs_exchange_setup first");
    KGS_CUDA(h, cudaSetDevice(h->device));
    KGS_CUDA(h, cudaStreamSynchronize(h->xchg));
    h->px_multicast = static_cast<char*>(mc_gathered);
    return KGS_OK;
}

int kgs_exchange_push(kgs_handle h, void** d_gathered, size_t* bytes_per_rank) {
    if (!h) return KGS_ERR_ARG;
    if (h->px_n == 0) return fail(h, KGS_ERR_STATE, "kgs_exchange_push: call kgs_exchange_setup first");
    char* src = nullptr;
    size_t nbytes = 0;
    int rc = exchange_source(h, "kgs_exchange_push", &src, &nbytes);
    if (rc) return rc;
    if (nbytes > h->px_slot)
        return fail(h, KGS_ERR_CAPACITY, "kgs_exchange_push: the job's tables (" + std::to_string(nbytes) + " bytes) exceed slot_bytes (" +
                                             std::to_string(h->px_slot) + ")");
    KGS_CUDA(h, cudaSetDevice(h->device));
    if ((rc = flush_pending_push(h))) return rc;   // an earlier push nobody has started yet
    const int n = h->px_n, me = h->px_me;
    const uint32_t seq = ++h->px_seq;
    const int half = (int)(seq & 1u);
    h->px_pending.on = true;
    h->px_pending.src = src;
    h->px_pending.nbytes = nbytes;
    h->px_pending.half = half;
    h->px_pending.seq = seq;
    h->px_pending.region_half = h->region_half;
    // The copies read this job's tables from HBM and write the peers' tables into it - beside the next job's
    // accumulate kernel, which is the one kernel of a step that is bound by HBM, that costs the most.  So they are
    // queued when that kernel has been launched (kgs_stats_accumulate) and run beside the finalize kernels, which
    // are bound by instruction issue; join / sync / the next push start them at once if no job follows.
    if (!h->px_defer && (rc = flush_pending_push(h))) return rc;
    if (d_gathered) *d_gathered = h->px_gathered[me] + (size_t)half * n * h->px_slot;
    if (bytes_per_rank) *bytes_per_rank = nbytes;
    return KGS_OK;
}

int kgs_exchange_join(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    KGS_CUDA(h, cudaSetDevice(h->device));
    int rc0 = flush_pending_push(h);
    if (rc0) return rc0;
    for (int b = 0; b < 2; ++b)
        if (h->region_busy[b]) KGS_CUDA(h, cudaStreamWaitEvent(h->compute, h->region_free[b], 0));
    return KGS_OK;
}

int kgs_exchange_sync(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    KGS_CUDA(h, cudaSetDevice(h->device));
    int rc0 = flush_pending_push(h);
    if (rc0) return rc0;
    KGS_CUDA(h, cudaStreamSynchronize(h->xchg));
    uint32_t flag = 0;
    KGS_CUDA(h, cudaMemcpy(&flag, h->px_timeout, 4, cudaMemcpyDeviceToHost));
    if (flag) {
        KGS_CUDA(h, cudaMemset(h->px_timeout, 0, 4));
        return fail(h, KGS_ERR_STATE, "kgs_exchange_sync: a peer did not reach the barrier within 10 s");
    }
    return KGS_OK;
}

int kgs_nccl_unique_id(void* id128) {
    if (!id128) return KGS_ERR_ARG;
    const NcclApi& api = nccl_api();
    if (!api.ok) {
        g_create_error = api.load_error ? api.load_error : "NCCL not available";
        return KGS_ERR_STATE;
    }
    const int r = api.get_unique_id(id128);
    if (r) {
        g_create_error = std::string("ncclGetUniqueId: ") + api.get_error_string(r);
        return KGS_ERR_CUDA;
    }
    return KGS_OK;
}

int kgs_nccl_init(kgs_handle h, const void* id128, int n_ranks, int rank) {
    if (!h || !id128 || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(h, KGS_ERR_ARG, "kgs_nccl_init: bad argument");
    const NcclApi& api = nccl_api();
    if (!api.ok) return fail(h, KGS_ERR_STATE, api.load_error ? api.load_error : "NCCL not available");
    KGS_CUDA(h, cudaSetDevice(h->device));
    if (h->nccl_comm) {
        api.comm_destroy(h->nccl_comm);
        h->nccl_comm = nullptr;
    }
    NcclApi::UniqueId id;
    memcpy(id.bytes, id128, sizeof(id.bytes));
    const int r = api.comm_init_rank(&h->nccl_comm, n_ranks, id, rank);
    if (r) return fail(h, KGS_ERR_CUDA, std::string("ncclCommInitRank: ") + api.get_error_string(r));
    return KGS_OK;
}

int kgs_exchange_nccl(kgs_handle h, void* nccl_comm, void* d_recv, size_t recv_bytes, size_t* bytes_per_rank) {
    if (!h || !d_recv) return fail(h, KGS_ERR_ARG, "kgs_exchange_nccl: null receive buffer");
    const NcclApi& api = nccl_api();
    if (!api.ok) return fail(h, KGS_ERR_STATE, api.load_error ? api.load_error : "NCCL not available");
    void* comm = nccl_comm ? nccl_comm : h->nccl_comm;
    if (!comm) return fail(h, KGS_ERR_STATE, "kgs_exchange_nccl: no communicator (pass one, or call kgs_nccl_init)");
    char* src = nullptr;
    size_t nbytes = 0;
    int rc = exchange_source(h, "kgs_exchange_nccl", &src, &nbytes);
    if (rc) return rc;
    if (recv_bytes < nbytes) return fail(h, KGS_ERR_CAPACITY, "kgs_exchange_nccl: receive buffer smaller than one rank's tables");
    KGS_CUDA(h, cudaSetDevice(h->device));
    KGS_CUDA(h, cudaEventRecord(h->fin_done, h->compute));
    KGS_CUDA(h, cudaStreamWaitEvent(h->xchg, h->fin_done, 0));
    const int r = api.all_gather(src, d_recv, nbytes, kNcclUint8, comm, h->xchg);
    if (r) return fail(h, KGS_ERR_CUDA, std::string("ncclAllGather: ") + api.get_error_string(r));
    h->launches++;
    KGS_CUDA(h, cudaEventRecord(h->region_free[h->region_half], h->xchg));
    h->region_busy[h->region_half] = true;
    if (bytes_per_rank) *bytes_per_rank = nbytes;
    return KGS_OK;
}

int kgs_segmented_mean_f64(kgs_handle h, const double* values, const uint64_t* offsets, uint64_t n_series,
                           double mul, double div, double* out) {
    if (!h) return KGS_ERR_ARG;
    if (n_series == 0) return KGS_OK;
    if (!offsets || !out) return fail(h, KGS_ERR_ARG, "kgs_segmented_mean_f64: null array");
    if (n_series >= (1ull << 31)) return fail(h, KGS_ERR_ARG, "too many series");
    const uint64_t nv = offsets[n_series];
    if (nv && !values) return fail(h, KGS_ERR_ARG, "kgs_segmented_mean_f64: null values");
    std::vector<uint32_t> long_ids;
    for (uint64_t s = 0; s < n_series; ++s) {
        if (offsets[s + 1] < offsets[s]) return fail(h, KGS_ERR_ARG, "offsets must be non-decreasing");
        if (offsets[s + 1] - offsets[s] > kSeqMax) long_ids.push_back((uint32_t)s);
    }
    KGS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->compute;
    KGS_CUDA(h, cudaStreamSynchronize(st));   // tmp buffers may be reallocated
    int rc;
    if ((rc = reserve(h, h->tmp, nv * 8 + 8))) return rc;
    if ((rc = reserve(h, h->tmp2, (n_series + 1) * 8))) return rc;
    if ((rc = reserve(h, h->tmp3, n_series * 8))) return rc;
    if ((rc = reserve(h, h->tmp4, long_ids.size() * 4 + 4))) return rc;
    if (nv) KGS_CUDA(h, cudaMemcpyAsync(h->tmp.p, values, nv * 8, cudaMemcpyHostToDevice, st));
    KGS_CUDA(h, cudaMemcpyAsync(h->tmp2.p, offsets, (n_series + 1) * 8, cudaMemcpyHostToDevice, st));
    if (!long_ids.empty())
        KGS_CUDA(h, cudaMemcpyAsync(h->tmp4.p, long_ids.data(), long_ids.size() * 4, cudaMemcpyHostToDevice, st));
    const double* dv = static_cast<const double*>(h->tmp.p);
    const unsigned long long* doff = static_cast<const unsigned long long*>(h->tmp2.p);
    double* dout = static_cast<double*>(h->tmp3.p);
    launch(h, KGS_K_SEGMEAN, st, [&] {
        k_segmean_short<<<(unsigned)((n_series + 127) / 128), 128, 0, st>>>(dv, doff, (uint32_t)n_series, mul, div, dout);
    });
    if (!long_ids.empty()) {
        const uint32_t nl = (uint32_t)long_ids.size();
        launch(h, KGS_K_SEGMEAN, st, [&] {
            k_segmean_long<<<(nl * 32 + 127) / 128, 128, 0, st>>>(dv, doff, static_cast<const uint32_t*>(h->tmp4.p), nl, mul, div, dout);
        });
    }
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaMemcpyAsync(out, dout, n_series * 8, cudaMemcpyDeviceToHost, st));
    KGS_CUDA(h, cudaStreamSynchronize(st));
    return KGS_OK;
}

int kgs_matrix_parse(kgs_handle h, const char* text, uint64_t nbytes, int result_type, uint64_t* n_series,
                     uint64_t* n_samples) {
    if (!h) return KGS_ERR_ARG;
    h->mx_valid = false;
    if (!text && nbytes) return fail(h, KGS_ERR_ARG, "kgs_matrix_parse: null text");
    if (result_type != KGS_RESULT_MATRIX && result_type != KGS_RESULT_VECTOR)
        return fail(h, KGS_ERR_ARG, "kgs_matrix_parse: result_type must be KGS_RESULT_MATRIX or KGS_RESULT_VECTOR");
    if (nbytes >= (1ull << 32) - 2 * kMxTile)
        return fail(h, KGS_ERR_ARG, "kgs_matrix_parse: an answer of 4 GiB or more has to be split by the caller");
    KGS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->compute;
    KGS_CUDA(h, cudaStreamSynchronize(st));   // buffers may be reallocated
    KGS_CUDA(h, cudaStreamSynchronize(h->copy));
    const uint32_t n_ctas = (uint32_t)std::max<uint64_t>(1, (nbytes + kMxTile - 1) / kMxTile);
    const size_t padded = (size_t)n_ctas * kMxTile + kMxChunk;
    const size_t chunks = (size_t)n_ctas * kMxThreads;
    // The text arrives in pieces on the copy stream; a piece is decoded while the next ones are in flight.
    const uint32_t piece_ctas = std::max<uint32_t>(1, h->mx_piece_bytes / kMxTile);
    const uint32_t n_pieces = (n_ctas + piece_ctas - 1) / piece_ctas;
    int rc;
    if ((rc = reserve(h, h->mx_text, padded))) return rc;
    if ((rc = reserve(h, h->mx_cta_trans, n_ctas * sizeof(MxTrans)))) return rc;
    if ((rc = reserve(h, h->mx_cta_counts, n_ctas * sizeof(uint2)))) return rc;
    if ((rc = reserve(h, h->mx_chunk_state, chunks * 4))) return rc;
    if ((rc = reserve(h, h->mx_chunk_counts, chunks * 8))) return rc;
    if ((rc = reserve(h, h->mx_ctl, 64))) return rc;
    while (h->mx_arrived.size() < n_pieces) {
        cudaEvent_t e = nullptr;
        KGS_CUDA(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->mx_arrived.push_back(e);
    }
    char* dtext = static_cast<char*>(h->mx_text.p);
    unsigned long long* ctl = static_cast<unsigned long long*>(h->mx_ctl.p);
    KGS_CUDA(h, cudaMemsetAsync(ctl, 0, 64, st));
    KGS_CUDA(h, cudaMemsetAsync(ctl + 2, 0xFF, 8, st));
    MxParams P{};
    P.text = dtext;
    P.n = nbytes;
    P.n_ctas = n_ctas;
    P.pair_depth = result_type == KGS_RESULT_MATRIX ? 5u : 4u;
    P.carry = reinterpret_cast<MxTrans*>(ctl + 6);
    P.cta_trans = static_cast<MxTrans*>(h->mx_cta_trans.p);
    P.cta_counts = static_cast<uint2*>(h->mx_cta_counts.p);
    P.chunk_state = static_cast<uint32_t*>(h->mx_chunk_state.p);
    P.chunk_counts = static_cast<uint2*>(h->mx_chunk_counts.p);
    P.ctl = ctl;
    auto send_piece = [&](uint32_t i) -> int {
        const size_t off = (size_t)i * piece_ctas * kMxTile;
        const size_t len = off < nbytes ? std::min<size_t>(nbytes - off, (size_t)piece_ctas * kMxTile) : 0;
        if (len) KGS_CUDA(h, cudaMemcpyAsync(dtext + off, text + off, len, cudaMemcpyHostToDevice, h->copy));
        if (i + 1 == n_pieces) KGS_CUDA(h, cudaMemsetAsync(dtext + nbytes, 0, padded - nbytes, h->copy));
        KGS_CUDA(h, cudaEventRecord(h->mx_arrived[i], h->copy));
        return KGS_OK;
    };
    // grow-keeping-contents; `fill` bytes are written behind the kept part (the span table starts as all ones)
    auto ensure = [&](DevBuf& b, size_t need, size_t keep, size_t want, int fill, bool first) -> int {
        const size_t before = b.cap;
        if (need > b.cap) {
            int r = grow_copy(h, b, std::max(need, want), keep);
            if (r) return r;
        }
        if (fill >= 0 && (first || b.cap != before))
            KGS_CUDA(h, cudaMemsetAsync(static_cast<char*>(b.p) + keep, fill, b.cap - keep, st));
        return KGS_OK;
    };
    // whatever way this function is left, no copy may still be reading the caller's buffer
    struct DrainCopies {
        cudaStream_t s;
        ~DrainCopies() { cudaStreamSynchronize(s); }
    } drain{h->copy};
    // pageable text is staged through pinned pieces by a thread of its own (TextStager); pinned text is queued directly
    bool staged = false;
    if (nbytes >= (8u << 20)) {
        cudaPointerAttributes pa{};
        const cudaError_t pe = cudaPointerGetAttributes(&pa, text);
        if (pe != cudaSuccess) cudaGetLastError();
        staged = pe != cudaSuccess || pa.type == cudaMemoryTypeUnregistered;
    }
    TextStager stager;   // declared after `drain`: joined first, then the copy stream is drained
    if (staged) {
        const size_t piece_bytes = (size_t)piece_ctas * kMxTile;
        if (h->mx_stage_bytes < piece_bytes) {
            for (int i = 0; i < 3; ++i) {
                if (h->mx_stage[i]) KGS_CUDA(h, cudaFreeHost(h->mx_stage[i]));
                h->mx_stage[i] = nullptr;
                KGS_CUDA(h, cudaHostAlloc(&h->mx_stage[i], piece_bytes, cudaHostAllocDefault));
                if (!h->mx_stage_free[i]) KGS_CUDA(h, cudaEventCreateWithFlags(&h->mx_stage_free[i], cudaEventDisableTiming));
            }
            h->mx_stage_bytes = piece_bytes;
        }
        int nt = h->mx_stage_threads;
        if (nt <= 0) nt = std::max(1, std::min(12, (int)std::thread::hardware_concurrency() * 3 / 4));
        stager.h = h; stager.text = text; stager.dtext = dtext; stager.nbytes = nbytes; stager.padded = padded;
        stager.piece_bytes = piece_bytes; stager.n_pieces = n_pieces; stager.n_threads = nt;
        stager.worker = std::thread([&stager] { stager.run(); });
    }
    auto piece_queued = [&](uint32_t k) -> int {   // the arrival event of piece k has been recorded (a wait on an unrecorded event is no wait)
        if (staged && !stager.wait_issued(k + 1))
            return fail(h, KGS_ERR_CUDA, std::string("staging the answer: ") + cudaGetErrorString((cudaError_t)stager.cuda_error.load()));
        return KGS_OK;
    };
    unsigned long long head[8];
    uint64_t k_done = 0, s_done = 0;
    uint32_t sent = 0;
    for (uint32_t i = 0; i < n_pieces; ++i) {
        while (!staged && sent < n_pieces && sent <= i + 2)
            if ((rc = send_piece(sent++))) return rc;
        if ((rc = piece_queued(i))) return rc;
        KGS_CUDA(h, cudaStreamWaitEvent(st, h->mx_arrived[i], 0));
        P.cta_first = i * piece_ctas;
        P.cta_count = std::min(piece_ctas, n_ctas - P.cta_first);
        P.last_piece = i + 1 == n_pieces;
        const unsigned grid = P.cta_count;
        launch(h, KGS_K_PARSE, st, [&] { k_mx_classify<<<grid, kMxThreads, 0, st>>>(P); });
        launch(h, KGS_K_PARSE, st, [&] { k_mx_top<<<1, kMxThreads, 0, st>>>(P); });
        launch(h, KGS_K_PARSE, st, [&] { k_mx_count<<<grid, kMxThreads, 0, st>>>(P); });
        launch(h, KGS_K_PARSE, st, [&] { k_mx_top2<<<1, kMxThreads, 0, st>>>(P); });
        KGS_CUDA(h, cudaGetLastError());
        KGS_CUDA(h, cudaMemcpyAsync(head, ctl, 64, cudaMemcpyDeviceToHost, st));
        KGS_CUDA(h, cudaStreamSynchronize(st));
        const uint64_t ns = head[0], nser = head[1];
        if (P.last_piece && head[4] != 0)
            return fail(h, KGS_ERR_PARSE, "kgs_matrix_parse: the text ends inside a string or with unbalanced brackets");
        if (ns >= (1ull << 32) || nser >= (1ull << 31))
            return fail(h, KGS_ERR_ARG, "kgs_matrix_parse: too many samples or series");
        // sizes of the whole answer, projected from the text seen so far
        const double seen = (double)std::min<uint64_t>(nbytes, (uint64_t)(P.cta_first + P.cta_count) * kMxTile);
        const double scale = P.last_piece || seen <= 0 ? 1.0 : 1.1 * (double)nbytes / seen;
        const size_t want_k = (size_t)((double)ns * scale) + 4096, want_s = (size_t)((double)nser * scale) + 256;
        const bool first = i == 0;
        if ((rc = ensure(h->mx_pos, ns * 4 + 4, k_done * 4, want_k * 4, -1, first))) return rc;
        if ((rc = ensure(h->mx_ts, ns * 8 + 8, k_done * 8, want_k * 8, -1, first))) return rc;
        if ((rc = ensure(h->mx_val, ns * 8 + 8, k_done * 8, want_k * 8, -1, first))) return rc;
        if ((rc = ensure(h->mx_off, (nser + 1) * 8, s_done * 8, (want_s + 1) * 8, -1, first))) return rc;
        if ((rc = ensure(h->mx_spans, nser * 32 + 32, s_done * 32, want_s * 32, 0xFF, first))) return rc;
        P.sample_pos = static_cast<uint32_t*>(h->mx_pos.p);
        P.ts = static_cast<double*>(h->mx_ts.p);
        P.val = static_cast<double*>(h->mx_val.p);
        P.offsets = static_cast<unsigned long long*>(h->mx_off.p);
        P.spans = static_cast<unsigned long long*>(h->mx_spans.p);
        P.k_first = (uint32_t)k_done;
        P.k_end = (uint32_t)ns;
        launch(h, KGS_K_PARSE, st, [&] { k_mx_emit<<<grid, kMxThreads, 0, st>>>(P); });
        // a pair that starts in the last bytes of this piece is read on into the next one
        if (!P.last_piece) {
            if ((rc = piece_queued(i + 1))) return rc;
            KGS_CUDA(h, cudaStreamWaitEvent(st, h->mx_arrived[i + 1], 0));
        }
        if (ns > k_done)
            launch(h, KGS_K_PARSE, st, [&] { k_mx_parse<<<(unsigned)((ns - k_done + 127) / 128), 128, 0, st>>>(P); });
        KGS_CUDA(h, cudaGetLastError());
        k_done = ns;
        s_done = nser;
    }
    if ((rc = reserve(h, h->mx_long, s_done * 4 + 4))) return rc;
    if (s_done)
        launch(h, KGS_K_PARSE, st, [&] {
            k_mx_long_series<<<(unsigned)((s_done + 255) / 256), 256, 0, st>>>(P.offsets, (uint32_t)s_done, kSeqMax,
                                                                                static_cast<uint32_t*>(h->mx_long.p), ctl);
        });
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaMemcpyAsync(head, ctl, 64, cudaMemcpyDeviceToHost, st));
    KGS_CUDA(h, cudaStreamSynchronize(st));
    if (head[2] != ~0ull)
        return fail(h, KGS_ERR_PARSE, "kgs_matrix_parse: malformed sample (float() would raise) at byte " + std::to_string(head[2]));
    h->mx_n_samples = k_done;
    h->mx_n_series = s_done;
    h->mx_n_long = head[3];
    h->mx_valid = true;
    if (n_series) *n_series = s_done;
    if (n_samples) *n_samples = k_done;
    return KGS_OK;
}

int kgs_matrix_fetch(kgs_handle h, double* ts, double* values, uint64_t* offsets, uint64_t* spans) {
    if (!h) return KGS_ERR_ARG;
    if (!h->mx_valid) return fail(h, KGS_ERR_STATE, "kgs_matrix_fetch: no decoded answer (kgs_matrix_parse first)");
    KGS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->compute;
    const uint64_t ns = h->mx_n_samples, nser = h->mx_n_series;
    if (ts && ns) KGS_CUDA(h, cudaMemcpyAsync(ts, h->mx_ts.p, ns * 8, cudaMemcpyDeviceToHost, st));
    if (values && ns) KGS_CUDA(h, cudaMemcpyAsync(values, h->mx_val.p, ns * 8, cudaMemcpyDeviceToHost, st));
    if (offsets) KGS_CUDA(h, cudaMemcpyAsync(offsets, h->mx_off.p, (nser + 1) * 8, cudaMemcpyDeviceToHost, st));
    if (spans && nser) KGS_CUDA(h, cudaMemcpyAsync(spans, h->mx_spans.p, nser * 32, cudaMemcpyDeviceToHost, st));
    KGS_CUDA(h, cudaStreamSynchronize(st));
    return KGS_OK;
}

int kgs_matrix_mean(kgs_handle h, double mul, double div, double* out) {
    if (!h) return KGS_ERR_ARG;
    if (!h->mx_valid) return fail(h, KGS_ERR_STATE, "kgs_matrix_mean: no decoded answer (kgs_matrix_parse first)");
    const uint64_t nser = h->mx_n_series;
    if (nser == 0) return KGS_OK;
    if (!out) return fail(h, KGS_ERR_ARG, "kgs_matrix_mean: null out");
    KGS_CUDA(h, cudaSetDevice(h->device));
    cudaStream_t st = h->compute;
    int rc;
    if ((rc = reserve(h, h->mx_out, nser * 8))) return rc;
    const double* dv = static_cast<const double*>(h->mx_val.p);
    const unsigned long long* doff = static_cast<const unsigned long long*>(h->mx_off.p);
    double* dout = static_cast<double*>(h->mx_out.p);
    launch(h, KGS_K_SEGMEAN, st, [&] {
        k_segmean_short<<<(unsigned)((nser + 127) / 128), 128, 0, st>>>(dv, doff, (uint32_t)nser, mul, div, dout);
    });
    if (h->mx_n_long) {
        const uint32_t nl = (uint32_t)h->mx_n_long;
        launch(h, KGS_K_SEGMEAN, st, [&] {
            k_segmean_long<<<(nl * 32 + 127) / 128, 128, 0, st>>>(dv, doff, static_cast<const uint32_t*>(h->mx_long.p), nl, mul, div, dout);
        });
    }
    KGS_CUDA(h, cudaGetLastError());
    KGS_CUDA(h, cudaMemcpyAsync(out, dout, nser * 8, cudaMemcpyDeviceToHost, st));
    KGS_CUDA(h, cudaStreamSynchronize(st));
    return KGS_OK;
}

int kgs_launch_count(kgs_handle h, uint64_t* out) {
    if (!h || !out) return KGS_ERR_ARG;
    *out = h->launches;
    return KGS_OK;
}

int kgs_set_profiling(kgs_handle h, int on) {
    if (!h) return KGS_ERR_ARG;
    int rc = drain_timed(h);
    h->prof = on != 0;
    return rc;
}

int kgs_kernel_time(kgs_handle h, int kernel_id, double* total_ms, uint64_t* launches) {
    if (!h || kernel_id < 0 || kernel_id >= KGS_N_KERNELS) return KGS_ERR_ARG;
    int rc = drain_timed(h);
    if (rc) return rc;
    if (total_ms) *total_ms = h->k_ms[kernel_id];
    if (launches) *launches = h->k_n[kernel_id];
    return KGS_OK;
}

int kgs_reset_kernel_times(kgs_handle h) {
    if (!h) return KGS_ERR_ARG;
    int rc = drain_timed(h);
    for (int i = 0; i < KGS_N_KERNELS; ++i) {
        h->k_ms[i] = 0;
        h->k_n[i] = 0;
    }
    return rc;
}

int kgs_stream(kgs_handle h, int which, void** out_stream) {
    if (!h || !out_stream || which < 0 || which > 2) return KGS_ERR_ARG;
    *out_stream = which == 2 ? (void*)h->xchg : which ? (void*)h->compute : (void*)h->copy;
    return KGS_OK;
}

int kgs_mark(kgs_handle h, int which) {
    if (!h || which < 0 || which > 1) return KGS_ERR_ARG;
    KGS_CUDA(h, cudaSetDevice(h->device));
    KGS_CUDA(h, cudaEventRecord(h->mark[which], h->compute));
    return KGS_OK;
}

int kgs_elapsed_ms(kgs_handle h, double* ms) {
    if (!h || !ms) return KGS_ERR_ARG;
    KGS_CUDA(h, cudaEventSynchronize(h->mark[1]));
    float f = 0.f;
    KGS_CUDA(h, cudaEventElapsedTime(&f, h->mark[0], h->mark[1]));
    *ms = f;
    return KGS_OK;
}

}  // extern "C"

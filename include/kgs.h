/* kgs.h - C ABI of libkgs, the B200 (sm_100a) hot path of kube_gpu_stats.
 *
 * The reference (two Python-2 scripts) exposes no plugin / operator / FFI interface
 * (SURVEY.md section 8b): its boundary is a handful of Python functions.  This header is
 * the boundary a native replacement of those loops binds to.  Each entry point cites the
 * reference lines whose work it takes over ("GUS" = gpu_util_stats/gpu_util_stats.py,
 * "WUG" = who_use_gpu/who_use_gpu.py under /root/reference); entry points for work that
 * BASELINE.json's north_star names but the reference does not contain say so.
 *
 * Conventions
 *   - every call returns an int: KGS_OK or a KGS_ERR_* code; nothing throws across the ABI;
 *     kgs_last_error(h) returns the text of the most recent failure on that handle;
 *   - the caller owns every host buffer it passes; the library owns all device memory it
 *     allocates until kgs_destroy;
 *   - a handle owns three CUDA streams (copy, compute, exchange) on one device; a handle is NOT
 *     thread-safe, distinct handles are independent;
 *   - there is no CPU fallback: without a CUDA device kgs_create fails.
 */
#ifndef KGS_H
#define KGS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct kgs_ctx* kgs_handle;

enum {
    KGS_OK = 0,
    KGS_ERR_ARG = 1,       /* bad argument (null pointer, zero step, width not a multiple of step ...) */
    KGS_ERR_CUDA = 2,      /* a CUDA runtime call failed; see kgs_last_error */
    KGS_ERR_STATE = 3,     /* call out of order (e.g. accumulate before begin, probe before build_join) */
    KGS_ERR_CAPACITY = 4,  /* a fixed-capacity scratch area overflowed (run log, static scratch) */
    KGS_ERR_NOMEM = 5,
    KGS_ERR_PARSE = 6      /* kgs_matrix_parse: the text is not a well-formed answer, or float() would raise */
};

/* key kinds of the windowed statistics (north_star: per-pod / per-GPU / per-node) */
enum { KGS_KEY_POD = 0, KGS_KEY_GPU = 1, KGS_KEY_NODE = 2, KGS_N_KEY_KINDS = 3 };
/* which field of the packed util word is aggregated */
enum { KGS_METRIC_SM = 0, KGS_METRIC_MEM = 1, KGS_METRIC_FB = 2 };

#define KGS_POD_NONE 0xFFFFFFFFu   /* probe result for a (node,pid) with no table entry */
#define KGS_MAX_Q 8

/* One windowed-statistics job.  Windows: pane p = (ts - t0) / step, window j covers panes
 * j .. j + width/step - 1 (width == step: tumbling; width a larger multiple: sliding; width shorter
 * than step: sampled windows [t0 + j*step, t0 + j*step + width) with gaps between them).
 * Rules R1-R6 in oracle/window_stats.py are the specification. */
typedef struct kgs_window_spec {
    uint32_t key_mask;       /* bit k set: produce the table for key kind k */
    uint32_t metric;         /* KGS_METRIC_* */
    uint32_t gpus_per_node;  /* node = gpu_gid / gpus_per_node */
    uint32_t t0;             /* start of window 0, seconds */
    uint32_t step;           /* seconds between window starts (> 0) */
    uint32_t width;          /* window width in seconds: a positive multiple of step, or shorter than step */
    uint32_t n_windows;      /* windows per key */
    uint32_t n_gpus;         /* key domain of KGS_KEY_GPU  (gpu_gid  < n_gpus)  */
    uint32_t n_nodes;        /* key domain of KGS_KEY_NODE (node     < n_nodes) */
    uint32_t n_q;            /* number of percentiles, 0..KGS_MAX_Q */
    double   q[KGS_MAX_Q];   /* each in [0,1] */
    uint64_t total_samples_hint; /* upper bound of samples this job will accumulate (sizes the
                                    value log and run log); 0 = size for the resident batch */
} kgs_window_spec;

/* ---- lifetime ------------------------------------------------------------------------ */
/* No reference counterpart (the reference has no device).  device = CUDA ordinal. */
int kgs_create(int device, kgs_handle* out);
int kgs_destroy(kgs_handle h);
/* Text of the last error on h; with h == NULL the last kgs_create failure of this thread. */
const char* kgs_last_error(kgs_handle h);
/* Wait for everything queued on the handle's streams. */
int kgs_sync(kgs_handle h);
const char* kgs_version(void);

/* ---- ingestion (north_star subsystem 1) ---------------------------------------------- */
/* Replaces the per-sample float() decode + list append of GUS:47-48 / GUS:73-74 as the way
 * samples reach the statistics: the caller parses once into four u32 columns (16 B/sample)
 * in pinned memory and the library copies them with cudaMemcpyAsync on its copy stream. */
int kgs_alloc_pinned(kgs_handle h, size_t nbytes, void** out);
int kgs_free_pinned(kgs_handle h, void* p);
/* Copy one batch of n samples into the next of the handle's two device batch slots and make
 * it the resident batch.  Asynchronous: the host columns must stay valid until kgs_sync or
 * until a later blocking call on the handle returns.  Batches alternate between two slots,
 * so the copy of batch i+1 overlaps the kernels still reading batch i. */
int kgs_ingest(kgs_handle h, const uint32_t* ts, const uint32_t* gpu_gid, const uint32_t* pid,
               const uint32_t* util, uint64_t n);
/* Build one large resident batch from several host pieces (a shard that fits HBM but not one pinned
 * buffer): reserve makes an empty batch of the given capacity resident, append copies n more samples
 * behind those already there (asynchronous, same lifetime rule as kgs_ingest). */
int kgs_batch_reserve(kgs_handle h, uint64_t capacity);
int kgs_batch_append(kgs_handle h, const uint32_t* ts, const uint32_t* gpu_gid, const uint32_t* pid,
                     const uint32_t* util, uint64_t n);
/* Bring the resident batch into series-major collector order (by gpu_gid, then pid, each series in
 * its input order) with a stable radix sort on the GPU; the sorted copy becomes the resident batch.
 * The statistics are the same for any order - this is for speed: input that interleaves the series
 * (a time-major live feed, several processes per GPU delivered second by second) makes every sample
 * its own run.  No reference counterpart (ingestion, north_star subsystem 1).  Queued on the compute
 * stream; needs a second copy of the batch in device memory. */
int kgs_reorder_batch(kgs_handle h);
/* kgs_ingest and the first kgs_batch_append look at a few hundred samples of the host columns.  A batch that
 * interleaves several series with a short period (all processes of a GPU, or the GPUs of a small box, reported
 * second by second) is accumulated from a series-major copy (the same sort, into the other batch slot); so is
 * any other disordered batch (a time-major feed of many GPUs, a random stream) when the job has percentiles -
 * without percentiles the per-CTA accumulator cache handles a random stream faster than a sort would.  The
 * resident batch itself stays as ingested: kgs_probe and kgs_fetch_batch answer in the caller's order.  On by
 * default; 0 switches it off (also KGS_AUTO_REORDER=0).  Columns adopted from device memory are never looked at. */
int kgs_set_auto_reorder(kgs_handle h, int on);
/* The look itself, for callers that want to know (host arrays, no GPU involved): 0 long runs (collector
 * order), 1 disordered, 2 several series interleaved with a period of 2..64 samples; -1 for a null array. */
int kgs_input_order_class(const uint32_t* gpu_gid, const uint32_t* pid, uint64_t n);
/* Copy the resident batch's columns back to host arrays of n entries (any pointer may be NULL). Blocking. */
int kgs_fetch_batch(kgs_handle h, uint32_t* ts, uint32_t* gpu_gid, uint32_t* pid, uint32_t* util);
/* Same, for columns that already live in this device's memory, 32-byte aligned (no copy is made;
 * the caller keeps them alive and unchanged until the work that reads them has been synchronised). */
int kgs_adopt_device_batch(kgs_handle h, const uint32_t* d_ts, const uint32_t* d_gpu_gid,
                           const uint32_t* d_pid, const uint32_t* d_util, uint64_t n);

/* ---- (node,pid) -> pod join (north_star subsystem 2) ---------------------------------- */
/* Build side.  Takes over the dictionaries the reference fills before joining: the
 * server -> pod -> cards table of GUS:139-148 and the pod identity rows of WUG:27-49,
 * re-keyed as (node << 32 | pid) -> pod_id.  Open addressing, linear probing, load <= 0.5.
 * A repeated key keeps the LAST pair, as the reference's dict assignment does (GUS:78, GUS:148).
 * keys/pod_ids are host arrays of m entries; n_pods = key domain of KGS_KEY_POD. */
int kgs_build_join(kgs_handle h, const uint64_t* keys, const uint32_t* pod_ids, uint64_t m,
                   uint32_t n_pods);
/* Probe side for the resident batch: out_pod_id[i] = pod of sample i or KGS_POD_NONE.  Takes
 * over the dict probes of GUS:86-91 (server_pods.get(server, {}).get(pod)).  out_pod_id is a
 * host array of n entries; blocking. */
int kgs_probe(kgs_handle h, uint32_t gpus_per_node, uint32_t* out_pod_id);

/* The same join for callers that cannot allocate GPU memory - an agent sharing its GPUs with the
 * jobs it monitors (a cudaMalloc there can fail, can starve the tenant, and stalls the device), or
 * one living inside a process whose allocator owns all device memory.  This entry point has NO
 * counterpart in the reference; it exists for that deployment constraint only.
 * Everything is a DEVICE pointer supplied by the caller (keys/pod_ids: the m join pairs; gpu_gid,
 * pid: the n samples; out_pod_id: n results, KGS_POD_NONE for a miss; the three sample-sized
 * arrays 16-byte aligned).  The hash table lives in a fixed array inside the loaded module:
 * 131 072 slots x 16 B = 2 MiB per GPU that loads the library, kept until the process exits, hence
 * m <= KGS_NOALLOC_MAX_PAIRS (a larger m returns KGS_ERR_CAPACITY).  No handle is needed; the text
 * of a failure is returned by kgs_last_error(NULL).  The call queues its kernels on `stream` (a
 * cudaStream_t, NULL = default stream) and returns without waiting; it performs no cudaMalloc /
 * cudaFree.  The first call on each device resolves the table's device address - make it before
 * any stream capture.  All calls on a device share the one table and rebuild it from empty, so
 * calls whose kernels may overlap on the same GPU (two streams, two threads) race: serialise them.
 * `stream` must belong to `device`; the caller's current device is the same after the call as before.
 * The keys are read on the GPU only, so the reserved key 0xFFFFFFFFFFFFFFFF (the free-slot marker,
 * which kgs_build_join refuses) cannot be refused here: a pair with that key is left out.  The table's
 * address is looked up once per device and process; it does not survive a cudaDeviceReset. */
#define KGS_NOALLOC_MAX_PAIRS 65536u
int kgs_join_probe_noalloc(int device, const uint64_t* d_keys, const uint32_t* d_pod_ids, uint64_t m,
                           const uint32_t* d_gpu_gid, const uint32_t* d_pid, uint64_t n,
                           uint32_t gpus_per_node, uint32_t* d_out_pod_id, void* stream);

/* ---- windowed statistics (north_star subsystem 3) ------------------------------------- */
/* The mean of GUS:77-80 (sum/len, 0.0 when empty) generalised to windows and to three key
 * kinds; max and percentiles have no reference counterpart.
 * begin: size and clear the accumulators.  accumulate: fold the resident batch in (may be
 * called once per ingested batch).  finalize: per-cell mean / max / percentiles into the
 * device result tables.  kgs_window_stats = begin + accumulate + finalize. */
int kgs_stats_begin(kgs_handle h, const kgs_window_spec* spec);
int kgs_stats_accumulate(kgs_handle h);
int kgs_stats_finalize(kgs_handle h);
int kgs_window_stats(kgs_handle h, const kgs_window_spec* spec);
/* Run-log sizing.  0 (default) = automatic: the log (16 B entries) is sized for runs of
 * min(8, step/2) samples, ample for collector-ordered input at 1 Hz; a quarter of it is a shared
 * overflow pool, the rest is handed out as private slabs per warp and launch, and the log grows
 * when a job brings more batches than total_samples_hint announced.  A non-zero value declares
 * input whose order is unknown or far from collector order (a time-major live feed, a shuffled
 * stream): it is the initial capacity, and every slab is sized for the worst case of one run per
 * sample.  A job that still overflows the pool fails at fetch time with KGS_ERR_CAPACITY. */
int kgs_set_run_log_capacity(kgs_handle h, uint64_t entries);
/* What the kernels of the open (or finalized) job met, for tests and tuning; blocking.
 * KGS_JC_POOL_ENTRIES   run-log entries taken from the shared overflow pool (0 in collector order);
 * KGS_JC_CACHE_CTAS     accumulate CTAs that folded runs in their shared-memory accumulator cache
 *                       first (input far from collector order: the flat random stream of configs[2]);
 * KGS_JC_BIG_CELLS      window cells of the last finalize counted by a whole warp (more than 256 samples, or
 *                       a value >= 104);
 * KGS_JC_WIDE_CELLS     window cells of the last finalize left to the 16-bit select, or (sliding windows) to
 *                       the second kernel.
 * No reference counterpart (the reference has one code path). */
enum { KGS_JC_POOL_ENTRIES = 0, KGS_JC_CACHE_CTAS = 1, KGS_JC_BIG_CELLS = 2, KGS_JC_WIDE_CELLS = 3, KGS_N_JOB_COUNTERS = 4 };
int kgs_job_counter(kgs_handle h, int which, uint64_t* out);
/* Copy one key kind's result table to host arrays of n_keys * n_windows entries, cell index
 * key * n_windows + window (pct: n_q rows of that length).  Any pointer may be NULL. Blocking. */
int kgs_fetch_stats(kgs_handle h, int key_kind, uint32_t* count, uint64_t* sum, uint32_t* max,
                    double* mean, double* pct);
/* Device addresses of the same tables, for a caller that exchanges them GPU-to-GPU (the
 * allgather of per-shard partial tables in north_star).  which: 0 count(u32) 1 sum(u64)
 * 2 max(u32) 3 mean(f64) 4 pct(f64, n_q rows). */
int kgs_result_device_ptr(kgs_handle h, int key_kind, int which, void** d_ptr, size_t* nbytes);
/* All result tables of the finalized job are one contiguous device region (every table starts
 * 256-byte aligned inside it).  Its first exchange_bytes hold what another shard needs - sum, pct,
 * count, max of every kind, 16 + 8*n_q bytes per cell - so ONE collective can move them all; the
 * mean tables (derivable: sum / count) follow, up to total_bytes.  kgs_result_offset gives a
 * table's offset inside the region; kgs_fetch_region copies the first nbytes of the region to
 * the host in one transfer (blocking; use a pinned destination).  Consecutive jobs alternate between
 * two result regions, so job i's region stays intact while job i+1 is computed (not job i+2). */
int kgs_result_region(kgs_handle h, void** d_ptr, size_t* exchange_bytes, size_t* total_bytes);
/* Packed exchange tables, for jobs whose tables travel between GPUs every step.  The option applies
 * from the next kgs_stats_begin; finalize then also writes, per key kind, what another shard needs to
 * rebuild the cell - the receiver recomputes mean = sum/count and each percentile from the two SELECTED
 * values with the sender's own roundings (rule R5), so nothing is lost.
 *   on = 1, packed: sum u32 (u64 when total_samples_hint * metric maximum >= 2^32; a job that accumulates
 *     more than the hint allowed for 32 bits fails at finalize with KGS_ERR_CAPACITY), count u32, max u16,
 *     per percentile u8|u8<<8 (u16|u16<<16 for fb_mib): 10 + 2*n_q bytes per cell instead of 16 + 8*n_q.
 *   on = 2, compact (8-bit metrics; fb_mib jobs fall back to packed): count u16, sum 24 bits (a u16 table
 *     and a u8 table), max u8, per percentile u8|u8<<8: 6 + 2*n_q bytes per cell.  A cell above 65 534
 *     samples has count 0xFFFF in the table and its full count and sum on an escape list at the head of
 *     the region ({u32 entries, 3 x u32 reserved}, then {u32 global cell, u32 count, u64 sum} each; global
 *     cell = cells of the earlier key kinds + key * n_windows + window).  The list has room for 1/64 of
 *     the cells + 1024; beyond that the sender's next fetch fails with KGS_ERR_CAPACITY and the receiver
 *     sees entries > capacity - meant for windows of minutes, not of days.
 * kgs_exchange_region returns the region one collective has to move: the packed tables when they are on
 * (*packed = 1 u32 sums, 2 u64 sums, 3 compact), else the exchange part of kgs_result_region (*packed = 0).
 * kgs_packed_offset: offset and size of one packed table inside that region (which: 0 count, 1 sum -
 * compact: its low 16 bits -, 2 max, 4 selected values, 5 compact: bits 16..23 of the sum, 6 compact: the
 * escape list, key_kind any kind of the job). */
int kgs_set_packed_exchange(kgs_handle h, int on);
int kgs_exchange_region(kgs_handle h, void** d_ptr, size_t* nbytes, int* packed);
int kgs_packed_offset(kgs_handle h, int key_kind, int which, size_t* offset, size_t* nbytes);
int kgs_result_offset(kgs_handle h, int key_kind, int which, size_t* offset, size_t* nbytes);
int kgs_fetch_region(kgs_handle h, void* host_dst, size_t nbytes);

/* ---- table exchange between the GPUs of one box (north_star: allgather of the per-shard tables) ----
 * The reference's tables are keyed by (server, pod) (GUS:70-71) and a pod is bound to one node
 * (WUG:33), so with nodes partitioned over the GPUs every cell lives on exactly one shard: the
 * exchange moves bytes and merges by concatenation.  The reference's only cross-shard arithmetic is
 * WUG:18-20 (GPU totals per type, a few integers), left to the caller.  What travels is the region
 * kgs_exchange_region describes.  Everything here is queued on the handle's third stream
 * (kgs_stream(h, 2, ..)) behind the finalize kernels of the job and runs beside the NEXT job's
 * kernels; a result region is reused two jobs later, and kgs_stats_begin makes that job wait for the
 * exchange that read it.  Not thread-safe per handle, like the rest.
 *
 * Peer pushes (no collective kernel, no SM taken from the statistics kernels):
 *   setup: gathered[r] = rank r's gathered buffer and flags[r] = rank r's flag words, both AS MAPPED
 *     ON THIS DEVICE (cudaIpcOpenMemHandle, cuMemMap, a symmetric-memory allocator ...; for r = my_rank
 *     the local pointers).  A gathered buffer holds 2 halves x n_ranks slots of slot_bytes; the flag
 *     words are 2 x 16 u32, zero before the first push.  n_ranks <= 16.
 *   push: copies this rank's region into slot my_rank of the current half of EVERY rank's buffer
 *     (one cudaMemcpyAsync per rank, copy engines over NVLink), then one small kernel tells every
 *     peer "complete" and waits for every peer's word (10 s limit, reported by kgs_exchange_sync).
 *     The copies are QUEUED when the next job's accumulate kernel has been launched (they then run
 *     beside the finalize kernels, which are not bound by HBM, instead of beside the one kernel that
 *     is); kgs_exchange_join / kgs_exchange_sync / kgs_sync / the next push queue them at once.
 *     KGS_PUSH_DEFER=0: always at once.
 *     Every rank must push the same number of times.  *d_gathered = where rank r's tables of THIS
 *     job will be on this device: (char*)*d_gathered + r * slot_bytes, valid once the exchange stream
 *     has passed the push, until the next-but-one push.  Readers of a half must be ordered (same
 *     stream, or an event) before this rank's next-but-one push.
 *   join: the compute stream waits for every exchange queued so far.  sync: the host does.
 * NCCL: kgs_exchange_nccl queues ONE ncclAllGather of the region into d_recv (n_ranks x the region's
 *   size, rank order) on the exchange stream.  nccl_comm is the caller's ncclComm_t (it must come
 *   from the libnccl.so.2 this process has loaded), or NULL for the handle's own communicator:
 *   kgs_nccl_unique_id (one rank; share the 128 bytes by any means) + kgs_nccl_init (every rank;
 *   collective).  libnccl is looked up with dlopen at the first call; KGS_ERR_STATE if absent. */
int kgs_exchange_setup(kgs_handle h, int n_ranks, int my_rank, void* const* gathered, void* const* flags,
                       size_t slot_bytes);
/* Optional, after setup: mc_gathered = ONE address behind which the NVSwitch replicates every write into
 * the gathered buffers of all ranks (an NVLS multicast mapping of the same allocation, e.g.
 * cuMulticastBindMem or a symmetric-memory allocator's multicast pointer).  push then issues one copy
 * instead of n_ranks: the tables are read from this GPU's memory once and leave it once.  NULL = off. */
int kgs_exchange_set_multicast(kgs_handle h, void* mc_gathered);
/* on = 0: push leaves slot my_rank of this rank's OWN gathered buffer alone - its tables are in its result
 * region anyway (kgs_exchange_region), with the same lifetime; saves a device-local copy per push.  Default 1
 * (the gathered buffer is complete, like the result of an all-gather). */
int kgs_exchange_set_self_copy(kgs_handle h, int on);
int kgs_exchange_push(kgs_handle h, void** d_gathered, size_t* bytes_per_rank);
int kgs_exchange_join(kgs_handle h);
int kgs_exchange_sync(kgs_handle h);
int kgs_nccl_unique_id(void* id128);
int kgs_nccl_init(kgs_handle h, const void* id128, int n_ranks, int rank);
int kgs_exchange_nccl(kgs_handle h, void* nccl_comm, void* d_recv, size_t recv_bytes, size_t* bytes_per_rank);

/* ---- the reference's own statistic, exactly ------------------------------------------- */
/* GUS:77-80 and GUS:51 on a CSR layout: series s owns values[offsets[s] .. offsets[s+1]).
 * out[s] = sum(values)/len (0.0 when empty) when div == 0 (GUS:77-80), otherwise
 * (sum(values) * mul) / div - GUS:51 with mul = step_sec, div = time_range_sec, evaluated in the
 * reference's order.  Series of up to 64 values are summed left to right by one thread, which
 * reproduces the reference's naive fp64 sum bit for bit; longer series use a warp tree (within
 * 1e-12 relative).  values/offsets/out are host arrays; blocking. */
int kgs_segmented_mean_f64(kgs_handle h, const double* values, const uint64_t* offsets,
                           uint64_t n_series, double mul, double div, double* out);

/* ---- value decode of a Prometheus answer on the GPU ---------------------------------------------
 * Replaces the reference's `requests.get(...).json()` + per-sample `float(val[1])` loops
 * (gpu_util_stats.py:25, :47-48, :73-74) for the bulk of an answer: `text` is the response body of
 * /query_range (KGS_RESULT_MATRIX: {"data":{"result":[{"metric":{..},"values":[[ts,"v"],..]},..]}})
 * or /query (KGS_RESULT_VECTOR: "value":[ts,"v"]).  The kernels find every series and every
 * [ts,"v"] pair (strings, escapes and nesting handled on the GPU) and convert both numbers to fp64
 * exactly as Python's float() does (correctly rounded, same grammar).  The result stays in device
 * memory as a CSR: values[offsets[s] .. offsets[s+1]) belong to series s, in text order.
 * `text` is a host pointer, nbytes < 4 GiB; it is copied in 32 MB pieces on the copy stream and a
 * piece is decoded while the next ones are in flight.  Pinned text is queued as it is; pageable text (the
 * bytes object an HTTP client returns) is staged by a thread of the library through three pinned pieces,
 * copied by several host threads (KGS_STAGE_THREADS, default three quarters of the cores, at most 12), so that the host
 * copy, the DMA and the decode of consecutive pieces overlap.
 * Errors: KGS_ERR_PARSE if the text ends inside a string / with open brackets, or a sample is not
 * `[number, "text"]` with text that float() accepts (kgs_last_error names the byte); an escape
 * sequence inside a sample's value string is refused too (Prometheus never writes one).  Blocking;
 * no copy reads `text` any more when the call returns, also on error. */
enum { KGS_RESULT_MATRIX = 0, KGS_RESULT_VECTOR = 1 };
int kgs_matrix_parse(kgs_handle h, const char* text, uint64_t nbytes, int result_type, uint64_t* n_series,
                     uint64_t* n_samples);
/* Copy the decoded answer to host arrays (any pointer may be NULL): ts and values have n_samples
 * entries, offsets n_series + 1, spans 4 per series: byte index of the series' '{', of the '[' and
 * ']' of its values array (UINT64_MAX when it has none) and of its '}' - what the caller needs to
 * decode the small label part (`metric`) of each series itself.  Blocking. */
int kgs_matrix_fetch(kgs_handle h, double* ts, double* values, uint64_t* offsets, uint64_t* spans);
/* kgs_segmented_mean_f64 over the decoded answer where it lies (no copy of the values): out has
 * n_series entries.  mul / div as in kgs_segmented_mean_f64.  Blocking. */
int kgs_matrix_mean(kgs_handle h, double mul, double div, double* out);

/* ---- measurement ---------------------------------------------------------------------- */
/* Kernel launches issued by this handle since creation (bench.py's gpu_launches). */
int kgs_launch_count(kgs_handle h, uint64_t* out);
/* With profiling on, every kernel launch is bracketed by CUDA events on the stream it runs
 * on; kgs_kernel_time returns the accumulated device time and launch count of one kernel
 * (ids: KGS_K_*), kgs_reset_kernel_times clears them. */
enum { KGS_K_BUILD = 0, KGS_K_BUILD_FIN = 1, KGS_K_PROBE = 2, KGS_K_ACCUM = 3, KGS_K_FINALIZE = 4,
       KGS_K_FINALIZE_BIG = 5, KGS_K_SEGMEAN = 6, KGS_K_CLEAR = 7, KGS_K_FINALIZE_WIDE = 8, KGS_K_FINALIZE_BASIC = 9, KGS_K_SORT = 10, KGS_K_PARSE = 11, KGS_K_FINALIZE_GROUPS = 12, KGS_K_EXCHANGE = 13, KGS_N_KERNELS = 14 };
int kgs_set_profiling(kgs_handle h, int on);
int kgs_kernel_time(kgs_handle h, int kernel_id, double* total_ms, uint64_t* launches);
int kgs_reset_kernel_times(kgs_handle h);
/* The handle's CUDA streams as cudaStream_t values (which: 0 copy, 1 compute, 2 exchange), for a caller that
 * wants to order its own device work (a collective, an event) after the library's kernels. */
int kgs_stream(kgs_handle h, int which, void** out_stream);
/* Device time between two marks on the compute stream (bench.py's timed region). */
int kgs_mark(kgs_handle h, int which /* 0 or 1 */);
int kgs_elapsed_ms(kgs_handle h, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* KGS_H */

/*
 * k8s_place_b200 -- C ABI of the B200-native fallback placement path.
 *
 * This is the drop-in boundary for ONE path of AshishGautamX/K8s-LLM-Scheduler: the
 * non-LLM placement decision, reference `HuggingFaceClient._fallback_decision`
 * (scheduler.py:521-559), fed by the node table of `ContextManager.get_node_metrics`
 * (scheduler.py:121-170) and the pod rows of `CustomScheduler._convert_pod_to_spec`
 * (scheduler.py:731-764).  The reference is pure Python and has no FFI; the binding a
 * maintainer would add is a ctypes stub -- see INTEGRATION.md.
 *
 * Conventions
 *   - plain C: pointers + sizes, no C++ or torch types; every call returns a kp_status
 *     and never throws.  kp_last_error() gives a thread-local message for the last
 *     non-OK status.
 *   - "device" entry points take DEVICE pointers, are ordered on the given CUDA stream
 *     (a cudaStream_t passed as void*, NULL = default stream), allocate nothing and
 *     never synchronise: the caller owns every buffer including the workspace.
 *   - "session" entry points take HOST pointers and own their device buffers (allocated
 *     once in kp_session_create); they copy in, run the same kernels, copy out and
 *     synchronise.  This is the form the reference-facing Python adapter binds.
 *   - wire tables: int64 milli-CPU, int64 bytes, int64 pod counts,
 *     fp64 usage percent, uint8 ready flag.  cores = milli/1000.0 and
 *     GiB = bytes/1024/1024/1024 reproduce the floats of scheduler.py:172-187.
 *   - results: int64 node index per pod (-1 = no node, the reference's
 *     selected_node == ""), fp64 score of the chosen node (-1.0 = none, the reference's
 *     best_score sentinel, scheduler.py:529).
 *   - values must satisfy |x| < 2^53 (exact in fp64); n_nodes < 2^31 - 4096.
 *   - threads: the reference calls this path from one thread (scheduler.py:654-729).  The
 *     library itself allows concurrent calls from several host threads as long as they do
 *     not share a workspace, an output buffer or a session; the only shared state is a
 *     per-device pool of helper streams (large node tables), whose use is serialised
 *     inside the library.  Device entry points can be captured in CUDA graphs once the
 *     first (uncaptured) call on that device has run.
 */
#ifndef K8S_PLACE_B200_H
#define K8S_PLACE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KP_ABI_VERSION 1

/* status codes */
#define KP_OK 0
#define KP_ERR_INVALID_ARG 1   /* NULL pointer, negative size, bad enum */
#define KP_ERR_WORKSPACE 2     /* workspace missing, misaligned or too small */
#define KP_ERR_CUDA 3          /* a CUDA runtime call failed (see kp_last_error) */
#define KP_ERR_NO_DEVICE 4     /* no usable sm_100 device */
#define KP_ERR_CAPACITY 5      /* session created for fewer nodes / pods */

/* Strategy ids: the values of config key fallback.strategy (scheduler.py:526,
 * config.yaml:34-36).  Any id other than 0 or 1 takes the round_robin branch, exactly
 * as any unknown string does in the reference (scheduler.py:544). */
#define KP_STRATEGY_RESOURCE_BALANCED 0   /* scheduler.py:537-541 */
#define KP_STRATEGY_LEAST_LOADED 1        /* scheduler.py:542-543 */
#define KP_STRATEGY_ROUND_ROBIN 2         /* scheduler.py:544-545 (picks the MOST loaded node) */

/* Filter modes.
 * KP_FILTER_READY    -- what the reference does: node Ready==True only
 *                       (scheduler.py:531-535); pod requests are not looked at.
 * KP_FILTER_CAPACITY -- north-star extension: additionally pod_count < max_pods,
 *                       cpu_req <= free_cpu and mem_req <= free_mem with
 *                       free = available * (100 - usage%) / 100.0, the only capacity
 *                       formula in the reference (scheduler.py:232-233). */
#define KP_FILTER_READY 0
#define KP_FILTER_CAPACITY 1

/* Node table, structure of arrays, one entry per node in LIST ORDER (= tie-break order,
 * scheduler.py:547).  Mirrors reference type NodeMetrics (scheduler.py:72-84). */
typedef struct kp_node_table {
    int64_t n_nodes;
    const int64_t *alloc_cpu_milli;  /* allocatable cpu, milli-cores   (available_cpu, :139)    */
    const int64_t *alloc_mem_bytes;  /* allocatable memory, bytes      (available_memory, :140) */
    const int64_t *pod_count;        /* pods running on the node       (:147)                   */
    const int64_t *max_pods;         /* allocatable pods               (:141)                   */
    const uint8_t *ready;            /* 1 iff a condition Ready=="True" exists (:532-533)       */
    const double *cpu_usage_pct;     /* cpu_usage_percent              (:150)                   */
    const double *mem_usage_pct;     /* memory_usage_percent           (:151)                   */
} kp_node_table;

/* Pod table, one entry per pending pod in ARRIVAL ORDER.  Mirrors reference type PodSpec
 * (scheduler.py:86-96); only the two request fields matter to this path. */
typedef struct kp_pod_table {
    int64_t n_pods;
    const int64_t *req_cpu_milli;    /* cpu_request, milli-cores (:742-745)           */
    const int64_t *req_mem_bytes;    /* memory_request, bytes    (:747-753)           */
    const int64_t *strategy_ids;     /* optional per-pod strategy id, NULL = uniform  */
} kp_pod_table;

/* ---- library ------------------------------------------------------------------- */
int kp_abi_version(void);
const char *kp_status_string(int status);
const char *kp_last_error(void);
/* Number of SMs etc. of `device`; KP_ERR_NO_DEVICE unless compute capability is 10.x. */
int kp_device_info(int device, int *sm_count, int *cc_major, int *cc_minor);
/* Number of kernels this library has launched in this process (for bench accounting). */
int64_t kp_launch_count(void);
/* Diagnostics for bench.py's roofline: when enabled, kp_snapshot_place records CUDA events
 * on its stream around the node-plan kernels and around the placement kernel.
 * kp_profile_last_ms waits for the last recorded call and returns both durations in ms.
 * Process-global and not thread-safe; off by default (no events are recorded). */
int kp_profile_enable(int enable);
int kp_profile_last_ms(float *plan_ms, float *place_ms);
/* Diagnostic with no counterpart in the reference: the MEASURED issue peak of the snapshot inner
 * loop on the current device.  Runs the three instructions place_kernel executes per pod x node pair
 * (subtract, test with predicate output, predicated update) on a shared-memory tile with nothing
 * around them, in both update mixes the real kernel can use, for at least min_ms milliseconds on
 * `stream`, and returns the best rate in pairs per second.  `scratch` = 256 bytes of device memory.
 * Synchronises the stream.  bench.py reports it as roofline.peak_measured. */
int kp_issue_peak(void *scratch, size_t scratch_bytes, double min_ms, void *stream, double *out_pairs_per_sec);

/* ---- snapshot placement, device pointers ------------------------------------------
 * Replaces: one `_fallback_decision` call per pod (scheduler.py:521-559) over the SAME
 * node table -- P independent decisions.  Fused filter + score + first-best argmax; the
 * P x N matrix is never materialised.
 *
 * `strategy` applies to every pod unless pods->strategy_ids is set.
 * Workspace: kp_snapshot_workspace_bytes(n_nodes, n_pods, per_pod_strategy != 0) bytes,
 * 256-byte aligned device memory; contents are scratch.
 */
size_t kp_snapshot_workspace_bytes(int64_t n_nodes, int64_t n_pods, int per_pod_strategy);
int kp_snapshot_place(const kp_node_table *nodes, const kp_pod_table *pods, int strategy,
                      int filter_mode, int64_t *out_node_idx, double *out_score,
                      void *workspace, size_t workspace_bytes, void *stream);

/* Same decisions for MORE pods against the SAME node table: the workspace still holds the node plan that an
 * earlier kp_snapshot_place call (same node table contents, same strategy, capacity filter, no per-pod strategy
 * ids) left there, and only the placement kernel runs -- the reference's per-pod loop re-reads one cluster table
 * for every pod (scheduler.py:690-729), so does a caller that scores pod batches as they arrive.  The caller
 * vouches for "same table, same workspace"; nothing can check it.  KP_ERR_INVALID_ARG for the Ready-only filter
 * or per-pod strategy ids (their plans are not reusable). */
int kp_snapshot_place_again(const kp_node_table *nodes, const kp_pod_table *pods, int strategy, int filter_mode,
                            int64_t *out_node_idx, double *out_score, void *workspace, size_t workspace_bytes,
                            void *stream);

/* kp_snapshot_place with a first-feasible probe in front of the table scan -- same decisions, bit for bit.
 * A pod's node is the first entry of the best-first table (score descending, list order within equal scores:
 * scheduler.py:547 keeps the earlier of equal scores) that it fits.  The probe tries the first `probe_depth`
 * (at most 64) entries of the pod's strategy table directly and answers the pod on the spot when one fits; only the
 * remaining pods are scanned against the whole table.  With clusters whose best-scoring nodes have room this turns
 * pods x nodes pair tests into a few per pod.  It is an OPTION, not the benchmarked kernel: the pairs evaluated are no
 * longer pods x nodes, so throughput in pairs/s is only "effective" here (kp_snapshot_scanned_pods tells how many
 * pods still took the full scan).  KP_FILTER_READY ignores the probe.  The workspace must be sized with
 * kp_snapshot_workspace_bytes(n_nodes, n_pods, 1) (the grouping pass needs the per-pod permutation) and also
 * serves kp_snapshot_place_again afterwards. */
int kp_snapshot_place_pruned(const kp_node_table *nodes, const kp_pod_table *pods, int strategy, int filter_mode,
                             int probe_depth, int64_t *out_node_idx, double *out_score, void *workspace,
                             size_t workspace_bytes, void *stream);
/* Pods per strategy (3 counts) that the last kp_snapshot_place_pruned call -- or a call with per-pod strategy ids --
 * on this workspace left for the full table scan.  Synchronises `stream`. */
int kp_snapshot_scanned_pods(const void *workspace, void *stream, int64_t *out_scanned);

/* ---- snapshot placement without a workspace, device pointers -------------------------
 * No counterpart in the reference: this entry point exists for hosts that cannot give
 * kp_snapshot_place a workspace because they cannot allocate device memory at the time
 * of the call -- e.g. a scheduler sidecar sharing its GPU with an inference server whose
 * memory pool already owns all of HBM, or a placement loop with a latency budget that
 * must stay clear of cudaMalloc/cudaFree (both take the driver's allocator lock and
 * synchronise the device).  Same inputs, outputs and results as kp_snapshot_place; the
 * scratch is a fixed 2 MiB array inside the library's device image (resident on every
 * GPU that loads the library, until the process exits), so the call allocates nothing.
 *
 * Limits: n_nodes <= 8192; with pods->strategy_ids set, n_pods <= 65536 (no pod limit for
 * a uniform strategy).  Larger calls return KP_ERR_CAPACITY with a message and do no work.
 *
 * The first call on each device looks up the array's device address (the device that is
 * current at the call); make that first call before starting any stream capture.  All
 * calls on a device share the one array: calls that may overlap on the GPU (different
 * streams, different threads) race on it -- serialise them, or use kp_snapshot_place with
 * separate workspaces.  Nothing needs clearing between calls; each call resets what it uses.
 */
int kp_snapshot_place_noalloc(const kp_node_table *nodes, const kp_pod_table *pods, int strategy,
                              int filter_mode, int64_t *out_node_idx, double *out_score,
                              void *stream);

/* ---- snapshot placement with the NODES sharded over several GPUs, device pointers ------------
 * Replaces the same P independent `_fallback_decision` calls (scheduler.py:521-559) when the node list is
 * split over W ranks (one process per GPU): rank r holds the contiguous slice of the list that starts at
 * global index node_offset (offsets grow with the rank), every rank passes the same pods, and the per-pod
 * winner over the whole list -- strict '>' with first-in-list-order ties, scheduler.py:547 -- is reduced
 * across the ranks by this library's kernels over peer-mapped memory (NVLink; CUDA IPC), with no host
 * round trip and no collective library on the data path.  Pods are owned in W contiguous shards
 * (first n_pods % W shards one longer): rank r receives the results of shard r.  Capacity filter only.
 *
 * A kp_exchange owns this rank's share of the exchange memory.  Set-up, once: kp_exchange_create on every
 * rank, kp_exchange_handle -> 64 opaque bytes, gathered from all ranks by the caller (any transport) and
 * given to kp_exchange_connect in rank order.  kp_exchange_connect_local links exchanges that live in ONE
 * process (several GPUs, or several emulated ranks on one GPU) without IPC.
 *
 * kp_snapshot_place_node_sharded = kp_ns_keys + kp_ns_place + kp_ns_reduce, all stream-ordered and
 * asynchronous; every rank must make the same calls in the same order.  (The three steps are separate so
 * that ranks emulated on ONE device can be driven in lock step: a kernel there must never wait for a
 * kernel that has not been launched yet.)  A rank that waits more than 15 s for a peer gives up and sets
 * the error word that kp_exchange_error reads (after synchronising): the results are then invalid.
 * Workspace: kp_snapshot_workspace_bytes(local n_nodes, n_pods, per_pod_strategy).
 */
typedef struct kp_exchange kp_exchange;
int kp_exchange_create(int device, int rank, int world, int64_t max_pods, int64_t max_local_nodes, kp_exchange **out);
int kp_exchange_handle(kp_exchange *ex, unsigned char *out_handle64);
int kp_exchange_connect(kp_exchange *ex, const unsigned char *handles /* world x 64 bytes, rank order */);
int kp_exchange_connect_local(kp_exchange *ex, kp_exchange *const *all /* world exchanges, rank order */);
void kp_exchange_destroy(kp_exchange *ex);
int kp_exchange_error(kp_exchange *ex, int *out_error);
/* First-feasible probe (see kp_snapshot_place_pruned) for this rank's scan of its node slice in kp_ns_place; 0 = off,
 * the default.  The workspace must then be sized with per_pod_strategy = 1.  Ranks may differ; the reduce is unchanged. */
int kp_exchange_set_prune_depth(kp_exchange *ex, int probe_depth);
int kp_ns_keys(kp_exchange *ex, const kp_node_table *local_nodes, int64_t node_offset, const kp_pod_table *pods,
               int strategy, int filter_mode, void *workspace, size_t workspace_bytes, void *stream);
int kp_ns_place(kp_exchange *ex, const kp_node_table *local_nodes, const kp_pod_table *pods, int strategy,
                int filter_mode, void *workspace, size_t workspace_bytes, void *stream);
int kp_ns_reduce(kp_exchange *ex, int64_t *out_idx_shard, double *out_score_shard, void *stream);
int kp_snapshot_place_node_sharded(kp_exchange *ex, const kp_node_table *local_nodes, int64_t node_offset,
                                   const kp_pod_table *pods, int strategy, int filter_mode,
                                   int64_t *out_idx_shard, double *out_score_shard, void *workspace,
                                   size_t workspace_bytes, void *stream);

/* ---- greedy placement with commit, device pointers ---------------------------------
 * Replaces: the sequential loop `_schedule_pod` -> `_fallback_decision` -> bind ->
 * next `get_node_metrics()` (scheduler.py:690-729, :598, :144-151).  Pods are placed in
 * arrival order; each placement does pod_count[chosen] += 1 and re-derives BOTH usage
 * percents of that node as (pod_count / max_pods) * 50 (0 if max_pods == 0); allocatable
 * never changes.  The input node table is not modified: the final state is
 * written to out_pod_count / out_cpu_usage_pct / out_mem_usage_pct (each n_nodes long,
 * may be NULL when not wanted).
 */
size_t kp_greedy_workspace_bytes(int64_t n_nodes, int64_t n_pods);
int kp_greedy_place(const kp_node_table *nodes, const kp_pod_table *pods, int strategy,
                    int filter_mode, int64_t *out_node_idx, double *out_score,
                    int64_t *out_pod_count, double *out_cpu_usage_pct,
                    double *out_mem_usage_pct, void *workspace, size_t workspace_bytes,
                    void *stream);

/* ---- per-node derived values, device pointers --------------------------------------
 * The hoisted per-node part of the path: score under `strategy` (scheduler.py:537-545),
 * free cpu cores / free memory GiB (scheduler.py:232-233) and eligibility under
 * `filter_mode` ignoring pod requests (1 = the node can win for some pod).  Any output
 * may be NULL.  No workspace.
 */
int kp_node_scores(const kp_node_table *nodes, int strategy, int filter_mode,
                   double *out_score, double *out_free_cpu, double *out_free_mem,
                   uint8_t *out_eligible, void *stream);

/* ---- session: host pointers, owns device memory ------------------------------------
 * The reference-facing form: `nodes`, `pods` and the outputs are HOST memory.  Each
 * call brings the tables to the GPU, runs the device path above on the session's stream,
 * returns the results to the host arrays and waits for them.  Pageable memory works
 * (staged copies).  PINNED host memory (cudaHostAlloc / cudaHostRegister, e.g. torch's
 * pin_memory()) is faster: a node table of up to ~18 k rows travels as one packed copy,
 * and in kp_session_snapshot with a single strategy the result arrays are not copied:
 * the placement kernel writes idx/score straight into the caller's arrays; with at most
 * 8192 nodes (every pod row is then read exactly once) the pod columns are not copied
 * either: the kernel reads the requests across the host link itself, paced so that the
 * rows arrive in pod order and scoring overlaps the transfer.
 * Either way the caller's arrays are complete when the call returns.
 * kp_session_last_copy_bytes counts the bytes that crossed the host link in the last
 * call, whichever way they travelled.
 */
typedef struct kp_session kp_session;
int kp_session_create(int device, int64_t max_nodes, int64_t max_pods, kp_session **out);
void kp_session_destroy(kp_session *s);
/* Copy + keep a node table on the device; later kp_session_snapshot/greedy calls with
 * nodes == NULL reuse it (the per-pod loop of the reference re-reads the same cluster). */
int kp_session_set_nodes(kp_session *s, const kp_node_table *nodes);
int kp_session_snapshot(kp_session *s, const kp_node_table *nodes, const kp_pod_table *pods,
                        int strategy, int filter_mode, int64_t *out_node_idx,
                        double *out_score);
int kp_session_greedy(kp_session *s, const kp_node_table *nodes, const kp_pod_table *pods,
                      int strategy, int filter_mode, int64_t *out_node_idx, double *out_score,
                      int64_t *out_pod_count, double *out_cpu_usage_pct,
                      double *out_mem_usage_pct);
/* Probe depth of kp_snapshot_place_pruned for this session's kp_session_snapshot calls (0 = off, the default). */
int kp_session_set_prune_depth(kp_session *s, int probe_depth);
/* ONE decision for ONE pod -- the reference's own call shape: HuggingFaceClient._fallback_decision(node_metrics,
 * reason) is called once per pending pod (scheduler.py:598 -> :521-559) with the node list it was just handed.
 * `nodes` is host memory (required).  Under KP_FILTER_READY the request arguments and the two allocatable columns
 * are not read (pass 0 and any valid arrays).  Tables of up to 2048 rows never leave the host: the columns are
 * gathered into the session's pinned staging buffer, ONE single-CTA kernel reads them across the host link, and the
 * answer comes back through a pinned word the call polls -- no copies, no stream synchronisation; larger tables go
 * through kp_session_snapshot with a one-pod table.  *out_node_idx = -1, *out_score = -1.0: no node can win. */
int kp_session_decide(kp_session *s, const kp_node_table *nodes, int strategy, int filter_mode,
                      int64_t req_cpu_milli, int64_t req_mem_bytes, int64_t *out_node_idx, double *out_score);
/* Bytes moved host->device and device->host by the last session call. */
int kp_session_last_copy_bytes(const kp_session *s, int64_t *h2d, int64_t *d2h);

/* ---- table ingest: Kubernetes quantity strings -> table columns, device pointers ------
 * The step in front of the placement kernels.  String columns are
 * Arrow-style: one byte buffer plus int64 offsets (row i is text[offsets[i] .. offsets[i+1])).
 * The fp64 outputs are the reference's own values bit for bit; the int64 outputs are the wire
 * values of kp_node_table / kp_pod_table that reproduce them.  Per-row status flags:
 */
#define KP_ROW_OK 0
#define KP_ROW_CPU_INEXACT 1 /* the fp64 cpu value is right but no integer milli-core count equals it */
#define KP_ROW_MEM_INEXACT 2 /* the fp64 memory value is right but no integer byte count equals it */
#define KP_ROW_HOST 4        /* a string outside the device grammar ([+-]digits[.digits], < 2^53):
                                the host must redo this row with float()/int() (it may raise, as the
                                reference would); the row's outputs are not meaningful */

/* Pod requests, CustomScheduler._convert_pod_to_spec (scheduler.py:731-764): fp64 sums over a
 * pod's containers in container order; cpu "…m" -> /1000 else float(x or 0) (:742-745); memory
 * only "…Ki" / "…Mi" / "…Gi", anything else contributes 0 (:747-753).  Containers without
 * resources.requests are left out by the caller (:738); a missing key is the string "0" (:739-740).
 */
typedef struct {
    int64_t n_pods;
    const int64_t *container_begin; /* [n_pods + 1] rows of pod p: container_begin[p] .. container_begin[p+1]-1 */
    const uint8_t *cpu_text;        /* requests['cpu'] of every container row */
    const int64_t *cpu_offsets;     /* [n_containers + 1] */
    const uint8_t *mem_text;        /* requests['memory'] */
    const int64_t *mem_offsets;     /* [n_containers + 1] */
} kp_pod_request_strings;

/* out_cpu_request / out_memory_request: PodSpec.cpu_request / memory_request (cores, GiB);
 * out_req_cpu_milli / out_req_mem_bytes: the kp_pod_table columns.  Any output but
 * out_status may be NULL. */
int kp_ingest_pod_requests(const kp_pod_request_strings *in, int64_t *out_req_cpu_milli,
                           int64_t *out_req_mem_bytes, double *out_cpu_request,
                           double *out_memory_request, uint8_t *out_status, void *stream);

/* Node rows, ContextManager.get_node_metrics (scheduler.py:139-151) with _parse_cpu (:172-176)
 * and _parse_memory (:178-187): allocatable cpu / memory / pods strings and the number of pods
 * on the node -> the kp_node_table columns (ready comes from the node conditions and is not a
 * quantity: the caller fills it).  cpu_usage_pct == mem_usage_pct == (pod_count/max_pods)*50,
 * 0 when max_pods <= 0 (:150-151). */
typedef struct {
    int64_t n_nodes;
    const uint8_t *cpu_text;   /* allocatable['cpu'] */
    const int64_t *cpu_offsets;
    const uint8_t *mem_text;   /* allocatable['memory'] */
    const int64_t *mem_offsets;
    const uint8_t *pods_text;  /* allocatable['pods'], int() */
    const int64_t *pods_offsets;
    const int64_t *pod_count;  /* len(pods on the node), scheduler.py:144-147 */
} kp_node_strings;

/* out_available_cpu / out_available_memory: NodeMetrics.available_cpu / available_memory. */
int kp_ingest_nodes(const kp_node_strings *in, int64_t *out_alloc_cpu_milli,
                    int64_t *out_alloc_mem_bytes, int64_t *out_max_pods, double *out_cpu_usage_pct,
                    double *out_mem_usage_pct, double *out_available_cpu,
                    double *out_available_memory, uint8_t *out_status, void *stream);

/* ---- cache keys in bulk, device pointers ---------------------------------------------
 * RequestCache._generate_key (scheduler.py:265-271): md5 over
 *     f"{cpu_request}_{memory_request}_{priority}_" + "_".join(f"{name}_{cpu%:.1f}_{mem%:.1f}" for nodes sorted by name)
 * for every pod of a batch against one node table.  The caller formats
 * the per-pod prefixes (Python float repr) and the node text once; the device hashes
 * prefix_j + nodes_text for all pods in parallel.  prefix_offsets has n_pods + 1 entries;
 * out_digest receives 16 bytes per pod in hexdigest() byte order and must be 4-byte aligned.
 */
int kp_cache_keys(const uint8_t *prefix_text, const int64_t *prefix_offsets, int64_t n_pods,
                  const uint8_t *nodes_text, int64_t nodes_len, uint8_t *out_digest, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* K8S_PLACE_B200_H */

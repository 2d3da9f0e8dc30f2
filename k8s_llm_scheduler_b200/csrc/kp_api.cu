// C ABI of the placement library (declared in include/k8s_place_b200.h).
// Validation, error strings, device queries and the host-pointer "session" that owns its
// device buffers.  No exceptions cross this boundary; every entry point returns a status.
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "kp_common.cuh"

namespace kp {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what)
{
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return KP_ERR_CUDA;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count_cached()
{
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
            v = 148;
        cached[dev] = v;
    }
    return cached[dev];
}

ProfileEvents &profile_events()
{
    static ProfileEvents pe{false, false, nullptr, nullptr, nullptr};
    return pe;
}

static int check_tables(const kp_node_table *nodes, const kp_pod_table *pods)
{
    if (!nodes || !pods) { set_error("nodes / pods table pointer is NULL"); return KP_ERR_INVALID_ARG; }
    if (nodes->n_nodes < 0 || pods->n_pods < 0) { set_error("negative table size"); return KP_ERR_INVALID_ARG; }
    if (nodes->n_nodes >= (int64_t)0x7fffffff - 4096) { set_error("n_nodes too large"); return KP_ERR_INVALID_ARG; }
    if (nodes->n_nodes > 0 &&
        (!nodes->alloc_cpu_milli || !nodes->alloc_mem_bytes || !nodes->pod_count || !nodes->max_pods ||
         !nodes->ready || !nodes->cpu_usage_pct || !nodes->mem_usage_pct)) {
        set_error("node table has a NULL column");
        return KP_ERR_INVALID_ARG;
    }
    if (pods->n_pods > 0 && (!pods->req_cpu_milli || !pods->req_mem_bytes)) {
        set_error("pod table has a NULL column");
        return KP_ERR_INVALID_ARG;
    }
    return KP_OK;
}

static int check_filter(int filter_mode)
{
    if (filter_mode != KP_FILTER_READY && filter_mode != KP_FILTER_CAPACITY) {
        set_error("filter_mode must be KP_FILTER_READY (0) or KP_FILTER_CAPACITY (1), got %d", filter_mode);
        return KP_ERR_INVALID_ARG;
    }
    return KP_OK;
}

}  // namespace kp

using namespace kp;

// ---------------------------------------------------------------------------------------
// session
// ---------------------------------------------------------------------------------------
struct kp_session {
    int device = 0;
    int64_t max_nodes = 0, max_pods = 0;
    cudaStream_t stream = nullptr;
    // device buffers
    unsigned char *d_nodes = nullptr;   // 4 x i64 + 2 x f64 columns + ready bytes
    unsigned char *d_pods = nullptr;    // 3 x i64 columns
    unsigned char *d_out = nullptr;     // idx i64 | score f64
    unsigned char *d_state = nullptr;   // greedy: pc i64 | cu f64 | mu f64
    void *d_ws = nullptr;
    size_t ws_bytes = 0;
    int64_t resident_nodes = -1;        // >= 0: a node table is resident
    int64_t h2d = 0, d2h = 0;
    // pinned staging for small node tables: the seven columns travel as ONE copy
    unsigned *d_flow = nullptr;         // flow-control words of the in-place snapshot (kept zero between calls)
    unsigned char *h_stage = nullptr;
    size_t stage_bytes = 0;
    int direct = 3;                     // bit 0: pinned pod columns read in place, bit 1: pinned results written in place
    // node plan kept in d_ws for the resident table (kp_session_set_nodes + snapshots with nodes == NULL):
    // signature of the last snapshot's plan, 0 = none; cleared whenever the table or the workspace changes hands
    unsigned long long plan_sig = 0;
    // large uploads are pipelined: pod columns go up in chunks on copy_stream while earlier chunks are scored
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t chunk_ready[8] = {nullptr};
    // kp_session_decide: answer slot in pinned host memory (the kernel stores, the host polls) and its call counter
    DecideSlot *h_slot = nullptr;
    unsigned long long decide_seq = 0;
    int prune_depth = 0;                // kp_session_set_prune_depth: first-feasible probe in kp_session_snapshot
};

namespace {

constexpr int kPipeMaxChunks = 8;
constexpr int64_t kPipeChunkMinPodsDefault = 1 << 19;   // below two pieces of this size one copy + one launch is as good
constexpr unsigned kFlowWindowCtas = 64;      // request rows of this many CTAs in flight on the host link
constexpr size_t kStageMaxBytes = 1u << 20;   // node tables up to ~18 k rows travel as one copy

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// Columns are packed for the RESIDENT table size (not max_nodes) so that a small table is one
// contiguous range: 6 x 8-byte columns, then the ready bytes.
kp_node_table session_node_view(const kp_session *s, int64_t n)
{
    const size_t col = align_up(sizeof(int64_t) * (size_t)(n > 0 ? n : 0));
    kp_node_table t{};
    t.n_nodes = n;
    t.alloc_cpu_milli = reinterpret_cast<const int64_t *>(s->d_nodes + 0 * col);
    t.alloc_mem_bytes = reinterpret_cast<const int64_t *>(s->d_nodes + 1 * col);
    t.pod_count = reinterpret_cast<const int64_t *>(s->d_nodes + 2 * col);
    t.max_pods = reinterpret_cast<const int64_t *>(s->d_nodes + 3 * col);
    t.cpu_usage_pct = reinterpret_cast<const double *>(s->d_nodes + 4 * col);
    t.mem_usage_pct = reinterpret_cast<const double *>(s->d_nodes + 5 * col);
    t.ready = reinterpret_cast<const uint8_t *>(s->d_nodes + 6 * col);
    return t;
}

// Device-side alias of p when p lies in pinned host memory the GPU can address, else nullptr.
void *device_alias(const void *p)
{
    cudaPointerAttributes a{};
    if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    return (reinterpret_cast<uintptr_t>(a.devicePointer) & 7u) ? nullptr : a.devicePointer;
}

int upload_nodes(kp_session *s, const kp_node_table *nodes)
{
    const int64_t n = nodes->n_nodes;
    if (n > s->max_nodes) {
        set_error("session holds at most %lld nodes, got %lld", (long long)s->max_nodes, (long long)n);
        return KP_ERR_CAPACITY;
    }
    const kp_node_table d = session_node_view(s, n);
    const size_t b8 = sizeof(int64_t) * (size_t)n;
    const size_t col = align_up(b8);
    const size_t packed = 6 * col + (size_t)n;
    if (n > 0 && s->h_stage && packed <= s->stage_bytes) {
        // gather the columns on the host (tens of KB) and send one copy instead of seven
        const void *src[6] = {nodes->alloc_cpu_milli, nodes->alloc_mem_bytes, nodes->pod_count,
                              nodes->max_pods, nodes->cpu_usage_pct, nodes->mem_usage_pct};
        for (int c = 0; c < 6; ++c) memcpy(s->h_stage + c * col, src[c], b8);
        memcpy(s->h_stage + 6 * col, nodes->ready, (size_t)n);
        KP_CUDA(cudaMemcpyAsync(s->d_nodes, s->h_stage, packed, cudaMemcpyHostToDevice, s->stream));
    } else if (n > 0) {
        KP_CUDA(cudaMemcpyAsync((void *)d.alloc_cpu_milli, nodes->alloc_cpu_milli, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)d.alloc_mem_bytes, nodes->alloc_mem_bytes, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)d.pod_count, nodes->pod_count, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)d.max_pods, nodes->max_pods, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)d.cpu_usage_pct, nodes->cpu_usage_pct, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)d.mem_usage_pct, nodes->mem_usage_pct, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)d.ready, nodes->ready, (size_t)n, cudaMemcpyHostToDevice, s->stream));
    }
    s->h2d += 6 * (int64_t)b8 + n;
    s->resident_nodes = n;
    s->plan_sig = 0;                    // a new table: whatever plan the workspace holds is stale
    return KP_OK;
}

int upload_pods(kp_session *s, const kp_pod_table *pods, kp_pod_table *dview)
{
    const int64_t p = pods->n_pods;
    if (p > s->max_pods) {
        set_error("session holds at most %lld pods, got %lld", (long long)s->max_pods, (long long)p);
        return KP_ERR_CAPACITY;
    }
    const size_t col = align_up(sizeof(int64_t) * (size_t)s->max_pods);
    const size_t b8 = sizeof(int64_t) * (size_t)p;
    dview->n_pods = p;
    dview->req_cpu_milli = reinterpret_cast<const int64_t *>(s->d_pods);
    dview->req_mem_bytes = reinterpret_cast<const int64_t *>(s->d_pods + col);
    dview->strategy_ids = pods->strategy_ids ? reinterpret_cast<const int64_t *>(s->d_pods + 2 * col) : nullptr;
    if (p > 0) {
        KP_CUDA(cudaMemcpyAsync((void *)dview->req_cpu_milli, pods->req_cpu_milli, b8, cudaMemcpyHostToDevice, s->stream));
        KP_CUDA(cudaMemcpyAsync((void *)dview->req_mem_bytes, pods->req_mem_bytes, b8, cudaMemcpyHostToDevice, s->stream));
        s->h2d += 2 * (int64_t)b8;
        if (pods->strategy_ids) {
            KP_CUDA(cudaMemcpyAsync((void *)dview->strategy_ids, pods->strategy_ids, b8, cudaMemcpyHostToDevice, s->stream));
            s->h2d += (int64_t)b8;
        }
    }
    return KP_OK;
}

int check_session_sizes(const kp_session *s, const kp_node_table *nodes, const kp_pod_table *pods)
{
    if (nodes && nodes->n_nodes > s->max_nodes) {
        set_error("session holds at most %lld nodes, got %lld", (long long)s->max_nodes, (long long)nodes->n_nodes);
        return KP_ERR_CAPACITY;
    }
    if (pods->n_pods > s->max_pods) {
        set_error("session holds at most %lld pods, got %lld", (long long)s->max_pods, (long long)pods->n_pods);
        return KP_ERR_CAPACITY;
    }
    return KP_OK;
}

// error return after work was enqueued: drain the stream first (the staging buffer is reused by the next call)
int fail_drained(kp_session *s, int st)
{
    cudaStreamSynchronize(s->stream);
    return st;
}

}  // namespace

extern "C" {

int kp_abi_version(void) { return KP_ABI_VERSION; }

const char *kp_status_string(int status)
{
    switch (status) {
        case KP_OK: return "ok";
        case KP_ERR_INVALID_ARG: return "invalid argument";
        case KP_ERR_WORKSPACE: return "workspace missing, misaligned or too small";
        case KP_ERR_CUDA: return "CUDA runtime error";
        case KP_ERR_NO_DEVICE: return "no usable sm_100 device";
        case KP_ERR_CAPACITY: return "session capacity exceeded";
        default: return "unknown status";
    }
}

const char *kp_last_error(void) { return g_err; }

int64_t kp_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int kp_profile_enable(int enable)
{
    ProfileEvents &pe = profile_events();
    if (enable && !pe.e0) {
        KP_CUDA(cudaEventCreate(&pe.e0));
        KP_CUDA(cudaEventCreate(&pe.e1));
        KP_CUDA(cudaEventCreate(&pe.e2));
    }
    pe.enabled = enable != 0;
    pe.valid = false;
    return KP_OK;
}

int kp_profile_last_ms(float *plan_ms, float *place_ms)
{
    ProfileEvents &pe = profile_events();
    if (!pe.enabled || !pe.valid) { set_error("no profiled kp_snapshot_place call recorded"); return KP_ERR_INVALID_ARG; }
    KP_CUDA(cudaEventSynchronize(pe.e2));
    float a = 0.f, b = 0.f;
    KP_CUDA(cudaEventElapsedTime(&a, pe.e0, pe.e1));
    KP_CUDA(cudaEventElapsedTime(&b, pe.e1, pe.e2));
    if (plan_ms) *plan_ms = a;
    if (place_ms) *place_ms = b;
    return KP_OK;
}

int kp_issue_peak(void *scratch, size_t scratch_bytes, double min_ms, void *stream, double *out_pairs_per_sec)
{
    return issue_peak_device(scratch, scratch_bytes, min_ms, (cudaStream_t)stream, out_pairs_per_sec);
}

int kp_device_info(int device, int *sm_count, int *cc_major, int *cc_minor)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
        set_error("CUDA device %d not available (%d devices visible)", device, count);
        cudaGetLastError();
        return KP_ERR_NO_DEVICE;
    }
    int sms = 0, maj = 0, min = 0;
    KP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    KP_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, device));
    KP_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, device));
    if (sm_count) *sm_count = sms;
    if (cc_major) *cc_major = maj;
    if (cc_minor) *cc_minor = min;
    if (maj != 10) {
        set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, maj, min);
        return KP_ERR_NO_DEVICE;
    }
    return KP_OK;
}

size_t kp_snapshot_workspace_bytes(int64_t n_nodes, int64_t n_pods, int per_pod_strategy)
{
    return snapshot_workspace_bytes(n_nodes, n_pods, per_pod_strategy != 0);
}

int kp_snapshot_place(const kp_node_table *nodes, const kp_pod_table *pods, int strategy,
                      int filter_mode, int64_t *out_node_idx, double *out_score, void *workspace,
                      size_t workspace_bytes, void *stream)
{
    int st = check_tables(nodes, pods);
    if (st != KP_OK) return st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) {
        set_error("output pointer is NULL");
        return KP_ERR_INVALID_ARG;
    }
    return snapshot_place_device(*nodes, *pods, strategy, filter_mode, out_node_idx, out_score,
                                 workspace, workspace_bytes, (cudaStream_t)stream);
}

// kp_snapshot_place with the first-feasible probe in front of the table scan (kp_snapshot.cu, GroupPods): same results,
// far fewer pairs evaluated when most pods fit one of the best-scoring nodes.
int kp_snapshot_place_pruned(const kp_node_table *nodes, const kp_pod_table *pods, int strategy, int filter_mode, int probe_depth,
                             int64_t *out_node_idx, double *out_score, void *workspace, size_t workspace_bytes, void *stream)
{
    int st = check_tables(nodes, pods);
    if (st != KP_OK) return st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) {
        set_error("output pointer is NULL");
        return KP_ERR_INVALID_ARG;
    }
    if (probe_depth < 0) { set_error("probe depth %d is negative", probe_depth); return KP_ERR_INVALID_ARG; }
    PlaceTuning tuning;
    tuning.prune_depth = probe_depth;
    return snapshot_place_device(*nodes, *pods, strategy, filter_mode, out_node_idx, out_score, workspace, workspace_bytes,
                                 (cudaStream_t)stream, tuning);
}

int kp_snapshot_scanned_pods(const void *workspace, void *stream, int64_t *out_scanned)
{
    if (!workspace || !out_scanned) { set_error("workspace / output is NULL"); return KP_ERR_INVALID_ARG; }
    return snapshot_scanned_pods(workspace, (cudaStream_t)stream, out_scanned);
}

int kp_snapshot_place_again(const kp_node_table *nodes, const kp_pod_table *pods, int strategy, int filter_mode,
                            int64_t *out_node_idx, double *out_score, void *workspace, size_t workspace_bytes, void *stream)
{
    int st = check_tables(nodes, pods);
    if (st != KP_OK) return st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) {
        set_error("output pointer is NULL");
        return KP_ERR_INVALID_ARG;
    }
    if (snapshot_plan_signature(nodes->n_nodes, pods->n_pods, pods->strategy_ids != nullptr, strategy, filter_mode) == 0ull) {
        set_error("kp_snapshot_place_again: only the capacity filter with one strategy keeps a reusable node plan");
        return KP_ERR_INVALID_ARG;
    }
    PlaceTuning tuning;
    tuning.reuse_plan = true;
    return snapshot_place_device(*nodes, *pods, strategy, filter_mode, out_node_idx, out_score, workspace, workspace_bytes,
                                 (cudaStream_t)stream, tuning);
}

int kp_snapshot_place_noalloc(const kp_node_table *nodes, const kp_pod_table *pods, int strategy,
                              int filter_mode, int64_t *out_node_idx, double *out_score, void *stream)
{
    int st = check_tables(nodes, pods);
    if (st != KP_OK) return st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) {
        set_error("output pointer is NULL");
        return KP_ERR_INVALID_ARG;
    }
    return snapshot_place_noalloc(*nodes, *pods, strategy, filter_mode, out_node_idx, out_score,
                                  (cudaStream_t)stream);
}

size_t kp_greedy_workspace_bytes(int64_t n_nodes, int64_t n_pods)
{
    return greedy_workspace_bytes(n_nodes, n_pods);
}

int kp_greedy_place(const kp_node_table *nodes, const kp_pod_table *pods, int strategy,
                    int filter_mode, int64_t *out_node_idx, double *out_score,
                    int64_t *out_pod_count, double *out_cpu_usage_pct, double *out_mem_usage_pct,
                    void *workspace, size_t workspace_bytes, void *stream)
{
    int st = check_tables(nodes, pods);
    if (st != KP_OK) return st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) {
        set_error("output pointer is NULL");
        return KP_ERR_INVALID_ARG;
    }
    return greedy_place_device(*nodes, *pods, strategy, filter_mode, out_node_idx, out_score,
                               out_pod_count, out_cpu_usage_pct, out_mem_usage_pct, workspace,
                               workspace_bytes, (cudaStream_t)stream);
}

int kp_node_scores(const kp_node_table *nodes, int strategy, int filter_mode, double *out_score,
                   double *out_free_cpu, double *out_free_mem, uint8_t *out_eligible, void *stream)
{
    kp_pod_table none{};
    int st = check_tables(nodes, &none);
    if (st != KP_OK) return st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    return node_scores_device(*nodes, strategy, filter_mode, out_score, out_free_cpu, out_free_mem,
                              out_eligible, (cudaStream_t)stream);
}

int kp_ingest_pod_requests(const kp_pod_request_strings *in, int64_t *out_req_cpu_milli,
                           int64_t *out_req_mem_bytes, double *out_cpu_request,
                           double *out_memory_request, uint8_t *out_status, void *stream)
{
    if (!in || in->n_pods < 0) { set_error("pod request strings: NULL or negative size"); return KP_ERR_INVALID_ARG; }
    if (in->n_pods > 0 && (!in->container_begin || !in->cpu_offsets || !in->mem_offsets || !out_status)) {
        set_error("pod request strings: container_begin / offsets / out_status is NULL");
        return KP_ERR_INVALID_ARG;
    }
    return ingest_pods_device(*in, out_req_cpu_milli, out_req_mem_bytes, out_cpu_request,
                              out_memory_request, out_status, (cudaStream_t)stream);
}

int kp_ingest_nodes(const kp_node_strings *in, int64_t *out_alloc_cpu_milli, int64_t *out_alloc_mem_bytes,
                    int64_t *out_max_pods, double *out_cpu_usage_pct, double *out_mem_usage_pct,
                    double *out_available_cpu, double *out_available_memory, uint8_t *out_status,
                    void *stream)
{
    if (!in || in->n_nodes < 0) { set_error("node strings: NULL or negative size"); return KP_ERR_INVALID_ARG; }
    if (in->n_nodes > 0 && (!in->cpu_offsets || !in->mem_offsets || !in->pods_offsets || !in->pod_count || !out_status)) {
        set_error("node strings: offsets, pod_count or out_status is NULL");
        return KP_ERR_INVALID_ARG;
    }
    return ingest_nodes_device(*in, out_alloc_cpu_milli, out_alloc_mem_bytes, out_max_pods,
                               out_cpu_usage_pct, out_mem_usage_pct, out_available_cpu,
                               out_available_memory, out_status, (cudaStream_t)stream);
}

int kp_cache_keys(const uint8_t *prefix_text, const int64_t *prefix_offsets, int64_t n_pods, const uint8_t *nodes_text,
                  int64_t nodes_len, uint8_t *out_digest, void *stream)
{
    if (n_pods < 0 || nodes_len < 0) { set_error("cache keys: negative size"); return KP_ERR_INVALID_ARG; }
    if (n_pods > 0 && (!prefix_offsets || !out_digest || (nodes_len > 0 && !nodes_text))) {
        set_error("cache keys: prefix_offsets / nodes_text / out_digest is NULL");
        return KP_ERR_INVALID_ARG;
    }
    if (reinterpret_cast<uintptr_t>(out_digest) & 3u) { set_error("cache keys: out_digest must be 4-byte aligned"); return KP_ERR_INVALID_ARG; }
    return cache_keys_device(prefix_text, prefix_offsets, n_pods, nodes_text, nodes_len, out_digest, (cudaStream_t)stream);
}

int kp_session_create(int device, int64_t max_nodes, int64_t max_pods, kp_session **out)
{
    if (!out || max_nodes < 0 || max_pods < 0) { set_error("bad session arguments"); return KP_ERR_INVALID_ARG; }
    *out = nullptr;
    int st = kp_device_info(device, nullptr, nullptr, nullptr);
    if (st != KP_OK) return st;
    DeviceGuard guard(device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice");
    kp_session *s = new (std::nothrow) kp_session();
    if (!s) { set_error("out of host memory"); return KP_ERR_INVALID_ARG; }
    s->device = device;
    s->max_nodes = max_nodes;
    s->max_pods = max_pods;
    const size_t ncol = align_up(sizeof(int64_t) * (size_t)max_nodes);
    const size_t pcol = align_up(sizeof(int64_t) * (size_t)max_pods);
    const size_t ws_snap = snapshot_workspace_bytes(max_nodes, max_pods, true);
    const size_t ws_greedy = greedy_workspace_bytes(max_nodes, max_pods);
    s->ws_bytes = ws_snap > ws_greedy ? ws_snap : ws_greedy;
    cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->d_nodes, 7 * ncol + kAlign);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->d_pods, 3 * pcol + kAlign);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->d_out, 2 * pcol + kAlign);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->d_state, 3 * ncol + kAlign);
    if (e == cudaSuccess) e = cudaMalloc(&s->d_ws, s->ws_bytes + kAlign);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking);
    for (int c = 0; e == cudaSuccess && c < kPipeMaxChunks; ++c) e = cudaEventCreateWithFlags(&s->chunk_ready[c], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->d_flow, kAlign);
    if (e == cudaSuccess) e = cudaMemset(s->d_flow, 0, kAlign);
    if (e == cudaSuccess) {
        const size_t want = 6 * ncol + (size_t)max_nodes;
        s->stage_bytes = want < kStageMaxBytes ? want : kStageMaxBytes;
        if (s->stage_bytes) e = cudaHostAlloc((void **)&s->h_stage, s->stage_bytes, cudaHostAllocDefault);
    }
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&s->h_slot, sizeof(DecideSlot), cudaHostAllocDefault);
    if (e == cudaSuccess) memset(s->h_slot, 0, sizeof(DecideSlot));
    if (const char *v = getenv("KP_SESSION_DIRECT")) s->direct = atoi(v) & 3;   // experiments only
    if (e != cudaSuccess) {
        kp_session_destroy(s);
        return cuda_fail(e, "kp_session_create allocation");
    }
    *out = s;
    return KP_OK;
}

void kp_session_destroy(kp_session *s)
{
    if (!s) return;
    DeviceGuard guard(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    cudaFree(s->d_nodes);
    cudaFree(s->d_pods);
    cudaFree(s->d_out);
    cudaFree(s->d_state);
    cudaFree(s->d_ws);
    cudaFree(s->d_flow);
    if (s->h_stage) cudaFreeHost(s->h_stage);
    if (s->h_slot) cudaFreeHost(s->h_slot);
    for (int c = 0; c < kPipeMaxChunks; ++c) if (s->chunk_ready[c]) cudaEventDestroy(s->chunk_ready[c]);
    if (s->copy_stream) { cudaStreamSynchronize(s->copy_stream); cudaStreamDestroy(s->copy_stream); }
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

int kp_session_set_nodes(kp_session *s, const kp_node_table *nodes)
{
    if (!s) { set_error("session is NULL"); return KP_ERR_INVALID_ARG; }
    kp_pod_table none{};
    int st = check_tables(nodes, &none);
    if (st != KP_OK) return st;
    DeviceGuard guard(s->device);
    s->h2d = s->d2h = 0;
    if ((st = upload_nodes(s, nodes)) != KP_OK) return st;
    KP_CUDA(cudaStreamSynchronize(s->stream));
    return KP_OK;
}

int kp_session_snapshot(kp_session *s, const kp_node_table *nodes, const kp_pod_table *pods,
                        int strategy, int filter_mode, int64_t *out_node_idx, double *out_score)
{
    if (!s || !pods) { set_error("session / pods is NULL"); return KP_ERR_INVALID_ARG; }
    int st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    DeviceGuard guard(s->device);
    s->h2d = s->d2h = 0;
    // sizes are checked before anything is enqueued: the node upload is an async copy out of the
    // session's one pinned staging buffer, and an early return must not leave it in flight
    if ((st = check_session_sizes(s, nodes, pods)) != KP_OK) return st;
    if (nodes) {
        if ((st = check_tables(nodes, pods)) != KP_OK) return st;
        if ((st = upload_nodes(s, nodes)) != KP_OK) return st;
    } else if (s->resident_nodes < 0) {
        set_error("no resident node table: pass nodes or call kp_session_set_nodes first");
        return KP_ERR_INVALID_ARG;
    }
    const kp_node_table dn = session_node_view(s, s->resident_nodes);
    if ((st = check_tables(&dn, pods)) != KP_OK) return fail_drained(s, st);
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) { set_error("output pointer is NULL"); return fail_drained(s, KP_ERR_INVALID_ARG); }
    const size_t pcol = align_up(sizeof(int64_t) * (size_t)s->max_pods);
    const size_t b8 = sizeof(int64_t) * (size_t)pods->n_pods;
    int64_t *d_idx = reinterpret_cast<int64_t *>(s->d_out);
    double *d_score = reinterpret_cast<double *>(s->d_out + pcol);
    kp_pod_table dp{};
    // Pinned (device-addressable) host buffers are used in place when the kernels touch every
    // pod row once: the requests are read across the host link by the placement kernel itself
    // and the results are written straight into the caller's arrays -- no staging copies, and
    // the transfers overlap the scoring CTA by CTA.
    bool in_place = false, out_in_place = false;
    // (requests: only when every row is read once -- a large table is cut into node segments and
    // each segment would fetch the rows again; results: whenever pods keep their order, every row is
    // written exactly once, by the CTA that finishes the pod chunk)
    const int prune = filter_mode == KP_FILTER_CAPACITY ? s->prune_depth : 0;
    if (s->direct && b8 && pods->strategy_ids == nullptr) {
        // (with the probe the grouping pass reads every request twice: the columns are copied, not read in place)
        const bool once = snapshot_single_pass(s->resident_nodes, false) && prune == 0;
        const void *rc = once ? device_alias(pods->req_cpu_milli) : nullptr;
        const void *rm = once ? device_alias(pods->req_mem_bytes) : nullptr;
        if ((s->direct & 1) && rc && rm && pods->n_pods <= s->max_pods) {
            dp.n_pods = pods->n_pods;
            dp.req_cpu_milli = static_cast<const int64_t *>(rc);
            dp.req_mem_bytes = static_cast<const int64_t *>(rm);
            s->h2d += 2 * (int64_t)b8;
            in_place = true;
        }
        void *oi = device_alias(out_node_idx), *os = device_alias(out_score);
        if ((s->direct & 2) && oi && os) {
            d_idx = static_cast<int64_t *>(oi);
            d_score = static_cast<double *>(os);
            out_in_place = true;
        }
    }
    // Many pods, one strategy: the pod columns go up in TWO pieces on a second stream -- a small first one, so that
    // scoring starts early, and the rest, which travels while the first is scored (against the same node plan).
    // Scoring a pod costs about n_nodes / 3500 times what copying it costs, which sizes the first piece; more,
    // equal pieces would only multiply the placement kernel's start-up and tail.
    int64_t kPipeChunkMinPods = kPipeChunkMinPodsDefault;
    if (const char *v = getenv("KP_PIPE_MIN")) kPipeChunkMinPods = std::max<int64_t>(1024, atoll(v));   // experiments only
    const int64_t P = pods->n_pods;
    const bool pipelined = !in_place && pods->strategy_ids == nullptr && filter_mode == KP_FILTER_CAPACITY && P <= s->max_pods &&
                           P >= 2 * kPipeChunkMinPods;
    int nchunks = pipelined ? 2 : 0;
    int64_t cut[kPipeMaxChunks + 1] = {0, 0, P};
    if (pipelined && prune > 0) {
        // with the probe a pod costs about as much to place as to copy: equal pieces, several of them, so that piece
        // c's answers travel to the host (written in place by the kernels) while piece c + 1 travels to the device
        nchunks = (int)std::min<int64_t>(kPipeMaxChunks, P / kPipeChunkMinPods);
        for (int c = 0; c <= nchunks; ++c) cut[c] = (P * c / nchunks) & ~(int64_t)255;
        cut[nchunks] = P;
    } else if (pipelined) {
        const double f = std::min(0.5, std::max(1.0 / 16, 3500.0 / (3500.0 + (double)dn.n_nodes)));
        cut[1] = std::max<int64_t>(kPipeChunkMinPods, (int64_t)(f * (double)P));
    }
    unsigned long long sig = 0;
    const unsigned long long want_sig = snapshot_plan_signature(dn.n_nodes, P, pods->strategy_ids != nullptr, strategy, filter_mode);
    const bool resident_plan = s->plan_sig != 0 && s->plan_sig == want_sig;
    s->plan_sig = 0;
    if (nchunks >= 2) {
        const int64_t *h_rc = pods->req_cpu_milli, *h_rm = pods->req_mem_bytes;
        int64_t *g_rc = reinterpret_cast<int64_t *>(s->d_pods), *g_rm = reinterpret_cast<int64_t *>(s->d_pods + pcol);
        for (int c = 0; c < nchunks; ++c) {
            const int64_t lo = cut[c], hi = cut[c + 1];
            const size_t bytes = sizeof(int64_t) * (size_t)(hi - lo);
            KP_CUDA(cudaMemcpyAsync(g_rc + lo, h_rc + lo, bytes, cudaMemcpyHostToDevice, s->copy_stream));
            KP_CUDA(cudaMemcpyAsync(g_rm + lo, h_rm + lo, bytes, cudaMemcpyHostToDevice, s->copy_stream));
            KP_CUDA(cudaEventRecord(s->chunk_ready[c], s->copy_stream));
        }
        s->h2d += 2 * (int64_t)b8;
        for (int c = 0; c < nchunks; ++c) {
            const int64_t lo = cut[c], hi = cut[c + 1];
            KP_CUDA(cudaStreamWaitEvent(s->stream, s->chunk_ready[c], 0));
            kp_pod_table part{hi - lo, g_rc + lo, g_rm + lo, nullptr};
            PlaceTuning tuning;
            tuning.prune_depth = prune;
            tuning.plan_sig = &sig;
            tuning.reuse_plan = c > 0 || resident_plan;
            st = snapshot_place_device(dn, part, strategy, filter_mode, d_idx + lo, d_score + lo, s->d_ws, s->ws_bytes, s->stream, tuning);
            if (st != KP_OK) { cudaStreamSynchronize(s->copy_stream); return fail_drained(s, st); }
        }
    } else {
        if (!in_place && (st = upload_pods(s, pods, &dp)) != KP_OK) return fail_drained(s, st);
        PlaceTuning tuning;
        tuning.prune_depth = prune;
        if (in_place) {
            tuning.flow = s->d_flow;
            tuning.flow_window = kFlowWindowCtas;
        }
        // resident table, same strategy / filter as the snapshot before: its node plan is still in d_ws
        tuning.plan_sig = &sig;
        tuning.reuse_plan = resident_plan;
        st = snapshot_place_device(dn, dp, strategy, filter_mode, d_idx, d_score, s->d_ws, s->ws_bytes, s->stream, tuning);
        if (st != KP_OK) return fail_drained(s, st);
    }
    s->plan_sig = sig;
    if (b8) {
        if (!out_in_place) {
            KP_CUDA(cudaMemcpyAsync(out_node_idx, d_idx, b8, cudaMemcpyDeviceToHost, s->stream));
            KP_CUDA(cudaMemcpyAsync(out_score, d_score, b8, cudaMemcpyDeviceToHost, s->stream));
        }
        s->d2h += 2 * (int64_t)b8;
    }
    KP_CUDA(cudaStreamSynchronize(s->stream));
    return KP_OK;
}

int kp_session_set_prune_depth(kp_session *s, int probe_depth)
{
    if (!s || probe_depth < 0) { set_error("session is NULL or probe depth negative"); return KP_ERR_INVALID_ARG; }
    s->prune_depth = probe_depth;
    return KP_OK;
}

// One decision for one pod: the reference's own call shape (scheduler.py:598 -> :521-559).  Small tables are scored
// where they lie (columns gathered into the pinned staging buffer, read across the host link by ONE kernel that also
// writes the answer into pinned memory); the host polls the answer's sequence word instead of synchronising.
int kp_session_decide(kp_session *s, const kp_node_table *nodes, int strategy, int filter_mode, int64_t req_cpu_milli,
                      int64_t req_mem_bytes, int64_t *out_node_idx, double *out_score)
{
    if (!s || !nodes || !out_node_idx || !out_score) { set_error("session / nodes / output is NULL"); return KP_ERR_INVALID_ARG; }
    int st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    const int64_t n = nodes->n_nodes;
    const size_t b8 = sizeof(int64_t) * (size_t)(n > 0 ? n : 0);
    const size_t col = align_up(b8);
    if (n < 0 || n > kDecideMaxNodes || n > s->max_nodes || !s->h_stage || 6 * col + (size_t)n > s->stage_bytes) {
        // larger tables: the batched path with a one-pod table
        const int64_t rc = req_cpu_milli, rm = req_mem_bytes;
        const kp_pod_table one{1, &rc, &rm, nullptr};
        return kp_session_snapshot(s, nodes, &one, strategy, filter_mode, out_node_idx, out_score);
    }
    kp_pod_table none{};
    if ((st = check_tables(nodes, &none)) != KP_OK) return st;
    DeviceGuard guard(s->device);
    s->h2d = s->d2h = 0;
    const void *src[6] = {nodes->alloc_cpu_milli, nodes->alloc_mem_bytes, nodes->pod_count,
                          nodes->max_pods, nodes->cpu_usage_pct, nodes->mem_usage_pct};
    for (int c = 0; c < 6; ++c) memcpy(s->h_stage + c * col, src[c], b8);
    if (n > 0) memcpy(s->h_stage + 6 * col, nodes->ready, (size_t)n);
    kp_node_table t{};
    t.n_nodes = n;
    t.alloc_cpu_milli = reinterpret_cast<const int64_t *>(s->h_stage + 0 * col);
    t.alloc_mem_bytes = reinterpret_cast<const int64_t *>(s->h_stage + 1 * col);
    t.pod_count = reinterpret_cast<const int64_t *>(s->h_stage + 2 * col);
    t.max_pods = reinterpret_cast<const int64_t *>(s->h_stage + 3 * col);
    t.cpu_usage_pct = reinterpret_cast<const double *>(s->h_stage + 4 * col);
    t.mem_usage_pct = reinterpret_cast<const double *>(s->h_stage + 5 * col);
    t.ready = s->h_stage + 6 * col;
    const unsigned long long seq = ++s->decide_seq;
    if ((st = decide_device(t, strategy, filter_mode, req_cpu_milli, req_mem_bytes, s->h_slot, seq, s->stream)) != KP_OK) return st;
    s->h2d += 6 * (int64_t)b8 + n;
    s->d2h += 2 * (int64_t)sizeof(int64_t);
    // the answer arrives as a release store at system scope: poll it (a few microseconds), and only if it stays away
    // fall back on the stream -- which also reports a launch that failed
    volatile unsigned long long *flag = &s->h_slot->seq;
    bool seen = false;
    for (int spin = 0; spin < (1 << 22); ++spin) {
        if (__atomic_load_n(flag, __ATOMIC_ACQUIRE) == seq) { seen = true; break; }
        __builtin_ia32_pause();
    }
    if (!seen) {
        KP_CUDA(cudaStreamSynchronize(s->stream));
        if (__atomic_load_n(flag, __ATOMIC_ACQUIRE) != seq) { set_error("decide: the kernel finished without an answer"); return KP_ERR_CUDA; }
    }
    *out_node_idx = s->h_slot->idx;
    *out_score = s->h_slot->score;
    return KP_OK;
}

int kp_session_greedy(kp_session *s, const kp_node_table *nodes, const kp_pod_table *pods,
                      int strategy, int filter_mode, int64_t *out_node_idx, double *out_score,
                      int64_t *out_pod_count, double *out_cpu_usage_pct, double *out_mem_usage_pct)
{
    if (!s || !pods) { set_error("session / pods is NULL"); return KP_ERR_INVALID_ARG; }
    int st;
    if ((st = check_filter(filter_mode)) != KP_OK) return st;
    DeviceGuard guard(s->device);
    s->h2d = s->d2h = 0;
    // sizes are checked before anything is enqueued: the node upload is an async copy out of the
    // session's one pinned staging buffer, and an early return must not leave it in flight
    if ((st = check_session_sizes(s, nodes, pods)) != KP_OK) return st;
    if (nodes) {
        if ((st = check_tables(nodes, pods)) != KP_OK) return st;
        if ((st = upload_nodes(s, nodes)) != KP_OK) return st;
    } else if (s->resident_nodes < 0) {
        set_error("no resident node table: pass nodes or call kp_session_set_nodes first");
        return KP_ERR_INVALID_ARG;
    }
    const int64_t n = s->resident_nodes;
    const kp_node_table dn = session_node_view(s, n);
    if ((st = check_tables(&dn, pods)) != KP_OK) return fail_drained(s, st);
    if (pods->n_pods > 0 && (!out_node_idx || !out_score)) { set_error("output pointer is NULL"); return fail_drained(s, KP_ERR_INVALID_ARG); }
    kp_pod_table dp{};
    if ((st = upload_pods(s, pods, &dp)) != KP_OK) return fail_drained(s, st);
    const size_t pcol = align_up(sizeof(int64_t) * (size_t)s->max_pods);
    const size_t ncol = align_up(sizeof(int64_t) * (size_t)s->max_nodes);
    int64_t *d_idx = reinterpret_cast<int64_t *>(s->d_out);
    double *d_score = reinterpret_cast<double *>(s->d_out + pcol);
    int64_t *d_pc = reinterpret_cast<int64_t *>(s->d_state);
    double *d_cu = reinterpret_cast<double *>(s->d_state + ncol);
    double *d_mu = reinterpret_cast<double *>(s->d_state + 2 * ncol);
    s->plan_sig = 0;                    // the greedy kernels take over the workspace
    st = greedy_place_device(dn, dp, strategy, filter_mode, d_idx, d_score, d_pc, d_cu, d_mu, s->d_ws,
                             s->ws_bytes, s->stream);
    if (st != KP_OK) return fail_drained(s, st);
    const size_t b8 = sizeof(int64_t) * (size_t)pods->n_pods;
    const size_t n8 = sizeof(int64_t) * (size_t)n;
    if (b8) {
        KP_CUDA(cudaMemcpyAsync(out_node_idx, d_idx, b8, cudaMemcpyDeviceToHost, s->stream));
        KP_CUDA(cudaMemcpyAsync(out_score, d_score, b8, cudaMemcpyDeviceToHost, s->stream));
        s->d2h += 2 * (int64_t)b8;
    }
    if (n8 && out_pod_count) { KP_CUDA(cudaMemcpyAsync(out_pod_count, d_pc, n8, cudaMemcpyDeviceToHost, s->stream)); s->d2h += n8; }
    if (n8 && out_cpu_usage_pct) { KP_CUDA(cudaMemcpyAsync(out_cpu_usage_pct, d_cu, n8, cudaMemcpyDeviceToHost, s->stream)); s->d2h += n8; }
    if (n8 && out_mem_usage_pct) { KP_CUDA(cudaMemcpyAsync(out_mem_usage_pct, d_mu, n8, cudaMemcpyDeviceToHost, s->stream)); s->d2h += n8; }
    KP_CUDA(cudaStreamSynchronize(s->stream));
    return KP_OK;
}

int kp_session_last_copy_bytes(const kp_session *s, int64_t *h2d, int64_t *d2h)
{
    if (!s) { set_error("session is NULL"); return KP_ERR_INVALID_ARG; }
    if (h2d) *h2d = s->h2d;
    if (d2h) *d2h = s->d2h;
    return KP_OK;
}

}  // extern "C"

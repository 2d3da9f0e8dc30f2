// Shared device/host helpers for the placement kernels (sm_100a only).
//
// Exact-arithmetic contract: every floating-point step is written with the IEEE
// round-to-nearest intrinsics (__dadd_rn, __dmul_rn, __ddiv_rn, ...) in the reference's
// own operation order, so that nvcc can never contract a*b+c into an FMA (a contraction
// changes resource_balanced score bits).  The library is additionally
// compiled with -fmad=false.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/k8s_place_b200.h"

namespace kp {

// ---------------------------------------------------------------------------------------
// exact fp64 node arithmetic
// ---------------------------------------------------------------------------------------

// scheduler.py:150-151   (pod_count / max_pods) * 50 if max_pods > 0 else 0
__device__ __forceinline__ double usage_a3(long long pc, long long mp)
{
    if (mp > 0) return __dmul_rn(__ddiv_rn((double)pc, (double)mp), 50.0);
    return 0.0;
}

// scheduler.py:537-541   left-associated: ((cpu*0.35) + (mem*0.35)) + (pod*0.30)
__device__ __forceinline__ double score_resource_balanced(double cu, double mu, long long pc,
                                                          long long mp)
{
    const double cpu_score = __ddiv_rn(__dsub_rn(100.0, cu), 100.0);
    const double mem_score = __ddiv_rn(__dsub_rn(100.0, mu), 100.0);
    const double pod_score = mp > 0 ? __ddiv_rn((double)(mp - pc), (double)mp) : 0.0;
    const double a = __dmul_rn(cpu_score, 0.35);
    const double b = __dmul_rn(mem_score, 0.35);
    const double c = __dmul_rn(pod_score, 0.30);
    return __dadd_rn(__dadd_rn(a, b), c);
}

// scheduler.py:542-543   (100 - cpu%) + (100 - mem%)
__device__ __forceinline__ double score_least_loaded(double cu, double mu)
{
    return __dadd_rn(__dsub_rn(100.0, cu), __dsub_rn(100.0, mu));
}

// scheduler.py:232-233   available * (100 - usage%) / 100.0, with available rebuilt from
// the integer wire format: cores = milli / 1000.0 (scheduler.py:174-175),
// GiB = bytes / 1024 / 1024 / 1024 (scheduler.py:187).
__device__ __forceinline__ double free_cpu_cores(long long alloc_milli, double cu)
{
    const double avail = __ddiv_rn((double)alloc_milli, 1000.0);
    return __ddiv_rn(__dmul_rn(avail, __dsub_rn(100.0, cu)), 100.0);
}
// bytes / 1024 / 1024 / 1024: each division by a power of two is exact for |bytes| < 2^53
// (no underflow: the smallest non-zero magnitude is 2^-30), so the three divisions equal ONE
// multiplication by 2^-30 bit for bit -- one instruction instead of three division sequences.
__device__ __forceinline__ double bytes_to_gib(long long bytes)
{
    return __dmul_rn((double)bytes, 9.31322574615478515625e-10);   // 2^-30
}
// float(text) / 1024 / 1024 / 1024 for a byte count that is already a double (scheduler.py:187)
__device__ __forceinline__ double bytes_to_gib_f64(double bytes)
{
    return __dmul_rn(bytes, 9.31322574615478515625e-10);           // 2^-30, exact scaling
}
__device__ __forceinline__ double free_mem_gib(long long alloc_bytes, double mu)
{
    const double avail = bytes_to_gib(alloc_bytes);
    return __ddiv_rn(__dmul_rn(avail, __dsub_rn(100.0, mu)), 100.0);
}
// pod request in the reference's units (scheduler.py:742-753, single quantity)
__device__ __forceinline__ double req_cores(long long milli)
{
    return __ddiv_rn((double)milli, 1000.0);
}
__device__ __forceinline__ double req_gib(long long bytes)
{
    return bytes_to_gib(bytes);
}

__host__ __device__ __forceinline__ int canon_strategy(long long sid)
{
    return sid == 0 ? 0 : (sid == 1 ? 1 : 2);   // scheduler.py:537/542/544
}

// ---------------------------------------------------------------------------------------
// order-preserving 64-bit keys: a < b (as doubles / int64) <=> key(a) < key(b) (unsigned)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long f64_order_key(double x)
{
    x = __dadd_rn(x, 0.0);                                   // -0.0 -> +0.0 (they compare equal)
    unsigned long long b = (unsigned long long)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ unsigned long long i64_order_key(long long x)
{
    return (unsigned long long)x ^ 0x8000000000000000ull;
}

// Score of node under strategy s and whether it can ever beat the reference's -1 sentinel
// with a strict '>' (scheduler.py:529,547).  NaN never wins (NaN > x is false).
struct NodeScore {
    double score;             // fp64 value reported to the caller
    unsigned long long key;   // order-preserving, larger = better; 0 = cannot win
};
__device__ __forceinline__ NodeScore node_score(int s, double cu, double mu, long long pc,
                                                long long mp)
{
    NodeScore r;
    if (s == 2) {                                   // scheduler.py:544-545: int compare
        r.score = (double)pc;
        r.key = pc > -1 ? i64_order_key(pc) : 0ull;
    } else {
        r.score = s == 0 ? score_resource_balanced(cu, mu, pc, mp) : score_least_loaded(cu, mu);
        r.key = r.score > -1.0 ? f64_order_key(r.score) : 0ull;
    }
    return r;
}

// ---------------------------------------------------------------------------------------
// small PTX wrappers: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "KP_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra KP_DONE_%=;\n"
        "bra KP_WAIT_%=;\n"
        "KP_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// cluster-wide barrier with release / acquire at cluster scope: orders shared-memory and DSMEM
// traffic of the cluster's CTAs without the device-wide fence that cooperative_groups'
// cluster.sync() emits (MEMBAR.ALL.GPU + ERRBAR showed up as the top stall of the kernels using it)
__device__ __forceinline__ void cluster_barrier()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// global -> shared bulk copy; bytes must be a multiple of 16, both addresses 16-B aligned
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gmem_src, uint32_t bytes,
                                          uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------------------------------
// host-side helpers
// ---------------------------------------------------------------------------------------
constexpr int kMaxStrategies = 3;
constexpr int kTileNodes = 4096;            // nodes per shared-memory tile of the snapshot kernel
constexpr size_t kAlign = 256;

inline size_t align_up(size_t x, size_t a = kAlign) { return (x + a - 1) / a * a; }

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);   // records message, returns KP_ERR_CUDA
void count_launch(int n = 1);

#define KP_CUDA(call)                                                     \
    do {                                                                  \
        cudaError_t e__ = (call);                                         \
        if (e__ != cudaSuccess) return ::kp::cuda_fail(e__, #call);       \
    } while (0)

#define KP_LAUNCH_CHECK(name)                                             \
    do {                                                                  \
        ::kp::count_launch();                                             \
        cudaError_t e__ = cudaGetLastError();                             \
        if (e__ != cudaSuccess) return ::kp::cuda_fail(e__, name);        \
    } while (0)

// internal entry points (one per .cu file)
// Launch options of the placement kernel that only the host-buffer session uses.
struct PlaceTuning {
    int resident_cap = 0;          // > 0: at most this many CTAs resident per SM
    unsigned *flow = nullptr;      // three zeroed device words (arrived, finished, spare flag): pace the in-place reads of pinned pod rows
    unsigned flow_window = 0;      // CTAs whose rows may be in flight at once
    // Node-plan reuse (session with a resident node table): `plan_sig` receives a signature of everything the
    // plan in the workspace depends on besides the node table itself (strategies, filter, layout); with
    // `reuse_plan` set the caller vouches that the workspace still holds the plan of an earlier call on the SAME
    // node table with the SAME signature, and the plan kernels are skipped.
    bool reuse_plan = false;
    unsigned long long *plan_sig = nullptr;
    bool reuse_plan_forced = false;   // reuse even with per-pod strategies: the plan was built by a plan_only call just before
    bool plan_only = false;           // build the node plan in the workspace and return (no pods are scored)
    int prune_depth = 0;              // > 0: first-feasible probe of this many best-first nodes before the table scan (kp_snapshot.cu);
                                      //      the workspace must be sized for per-pod strategies
    // Node-sharded form (kp_exchange.cu): besides index and score the epilogue stores, per pod, the chosen node's
    // GLOBAL rank position gpos_by_pos[strategy * gpos_stride + local rank position] (0xffffffff = no node).
    uint32_t *out_gpos = nullptr;
    const uint32_t *gpos_by_pos = nullptr;
    long long gpos_stride = 0;
};
// where the plan of a snapshot workspace keeps what the node-sharded exchange publishes
struct PlanView { const uint32_t *nvalid; const double *score_pos; int npad; };
PlanView snapshot_plan_view(int64_t n_nodes, int64_t n_pods, bool per_pod, void *workspace);
// signature the session compares before it sets reuse_plan (0 = this call's plan cannot be reused)
unsigned long long snapshot_plan_signature(int64_t n_nodes, int64_t n_pods, bool per_pod, int strategy, int filter_mode);
int snapshot_place_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                          int filter_mode, int64_t *out_idx, double *out_score, void *ws,
                          size_t ws_bytes, cudaStream_t stream, const PlaceTuning &tuning = PlaceTuning());
size_t snapshot_workspace_bytes(int64_t n_nodes, int64_t n_pods, bool per_pod);
bool snapshot_single_pass(int64_t n_nodes, bool per_pod);
int snapshot_scanned_pods(const void *workspace, cudaStream_t stream, int64_t *out_scanned);
int snapshot_place_noalloc(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                           int filter_mode, int64_t *out_idx, double *out_score, cudaStream_t stream);
int greedy_place_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                        int filter_mode, int64_t *out_idx, double *out_score,
                        int64_t *out_pod_count, double *out_cu, double *out_mu, void *ws,
                        size_t ws_bytes, cudaStream_t stream);
size_t greedy_workspace_bytes(int64_t n_nodes, int64_t n_pods);
// kp_greedy_fast.cu: integer state-machine greedy; *done_flag (device) tells the fp64 kernel to skip
size_t greedy_fast_workspace_bytes(int64_t n_nodes, int64_t n_pods);
int greedy_fast_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy, int filter_mode,
                       int64_t *out_idx, double *out_score, int64_t *out_pc, double *out_cu, double *out_mu,
                       void *ws, size_t ws_bytes, cudaStream_t stream, const uint32_t **done_flag);
int node_scores_device(const kp_node_table &nodes, int strategy, int filter_mode,
                       double *out_score, double *out_free_cpu, double *out_free_mem,
                       uint8_t *out_eligible, cudaStream_t stream);
// kp_ingest.cu: quantity strings -> table columns
int ingest_pods_device(const kp_pod_request_strings &in, int64_t *out_milli, int64_t *out_bytes,
                       double *out_cpu, double *out_mem, uint8_t *out_status, cudaStream_t stream);
int ingest_nodes_device(const kp_node_strings &in, int64_t *out_milli, int64_t *out_bytes,
                        int64_t *out_max_pods, double *out_cu, double *out_mu, double *out_avail_cpu,
                        double *out_avail_mem, uint8_t *out_status, cudaStream_t stream);
// kp_cachekey.cu: md5 of (per-pod prefix + shared node text) for a batch of pods
int cache_keys_device(const uint8_t *prefix, const int64_t *prefix_off, int64_t n_pods, const uint8_t *nodes_text,
                      int64_t nodes_len, uint8_t *out, cudaStream_t stream);
int sm_count_cached();
// kp_diag.cu: measured issue peak of the snapshot inner loop
// one decision for one pod (kp_decide.cu): the answer slot lives in pinned host memory
constexpr int kDecideMaxNodes = 2048;
struct DecideSlot { long long idx; double score; unsigned long long seq; unsigned long long pad; };
int decide_device(const kp_node_table &nodes, int strategy, int filter_mode, long long req_cpu_milli, long long req_mem_bytes,
                  DecideSlot *slot, unsigned long long seq, cudaStream_t stream);

int issue_peak_device(void *scratch, size_t scratch_bytes, double min_ms, cudaStream_t stream, double *pairs_per_sec);
// Ordinal of the current device, clamped to [0, kMaxDevices): index of per-device once-flags.  Function
// attributes (opt-in shared memory sizes) belong to a device's context, so such flags are kept per device.
constexpr int kMaxDevices = 64;
inline int device_slot()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) dev = 0;
    return dev;
}
// kp_sort.cu: library radix sort / scan for the large-table plan (temp storage from the workspace)
size_t sort_temp_bytes(int64_t n);
int sort_pairs_u64(void *temp, size_t temp_bytes, const unsigned long long *keys_in, unsigned long long *keys_out,
                   const uint32_t *vals_in, uint32_t *vals_out, int n, cudaStream_t stream);
int inclusive_sum_u32(void *temp, size_t temp_bytes, const uint32_t *in, uint32_t *out, int n, cudaStream_t stream);
// kp_ranksort.cu: hand-written sort of several key arrays at once (chunk sort + rank by merge counting), for up to
// kRankSortMax entries per array
constexpr int kRankSortMax = 65536;
struct RankSortArrays { int idx[8]; int count; };
size_t rank_sort_scratch_bytes(int64_t n, int arrays);
int rank_sort_pairs_u64(void *scratch, size_t scratch_bytes, const unsigned long long *keys_in, size_t in_stride, const RankSortArrays &which,
                        int n, unsigned long long *kout, uint32_t *vout, size_t out_stride, cudaStream_t stream);
int exclusive_sum_u32(void *temp, size_t temp_bytes, const uint32_t *in, uint32_t *out, int n, cudaStream_t stream);
int sort_pairs_u32_desc(void *temp, size_t temp_bytes, const uint32_t *keys_in, uint32_t *keys_out, const uint32_t *vals_in,
                        uint32_t *vals_out, int64_t n, cudaStream_t stream);
// diagnostics (kp_profile_enable): events around the node plan and the placement kernel
struct ProfileEvents { bool enabled; bool valid; cudaEvent_t e0, e1, e2; };
ProfileEvents &profile_events();

}  // namespace kp

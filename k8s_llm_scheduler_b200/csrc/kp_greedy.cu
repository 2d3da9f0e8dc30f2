// Greedy placement with commit: pods in arrival order, each placement visible to the next.
//
// Replaces the reference's sequential loop `_schedule_pod` -> `_fallback_decision` -> bind
// -> next `get_node_metrics()` (scheduler.py:690-729, :598, :144-151).  The state
// transition of one bind, as the next decision sees it, is exactly
//     pod_count[chosen] += 1
//     cpu_usage[chosen] = mem_usage[chosen] = (pod_count / max_pods) * 50   (0 if max_pods == 0)
// (SURVEY.md C3); allocatable never changes.  A pod that gets no node changes nothing
// (scheduler.py:727-729).
//
// This is a P-step dependent chain, so it is latency-bound, not throughput-bound.  One
// persistent CTA of 1024 threads owns the whole node table: every thread keeps the "hot"
// state of its nodes in registers (the two capacity thresholds as fp64 and one
// order-preserving 64-bit score key per strategy in play); the "cold" per-node constants a
// commit needs (available cores / GiB, max_pods, pod_count) sit in shared memory; pods are
// staged through shared memory in batches.  Per pod: every thread tests its nodes (2 fp64
// compares + 1 key compare each), a warp-level then block-level (key desc, index asc)
// reduction built from hardware REDUX finds the winner with ONE block barrier, and the owning
// thread commits: it recomputes that node's usage, thresholds and score key with the same
// exact-fp64 operation sequence as the snapshot path (cu == mu after a commit, so the shared
// sub-expression 100 - usage is computed once; the dependent chain is two divisions deep).
// Usage outputs are derived once at the end from the final pod counts.  Tables beyond the
// register budget use the same kernel with the hot state in global memory (slower).  This
// kernel is the fallback: kp_greedy_fast.cu (integer state machine, thread-block clusters for
// large tables) runs first and this one returns at once when that path did the work.
#include <algorithm>

#include "kp_common.cuh"

namespace kp {

constexpr int kGreedyThreads = 1024;
constexpr int kPodBatch = 1024;
constexpr int kGreedyColdBytes = 32;          // per node in shared memory: avail_cpu, avail_mem, max_pods, pod_count
constexpr int kGreedySmemNodes = 6144;        // cold state fits shared memory up to here (192 KB)

struct GreedyParams {
    const long long *alloc_cpu, *alloc_mem, *pod_count0, *max_pods;
    const uint8_t *ready;
    const double *cu0, *mu0;
    int n_nodes;
    const long long *req_cpu, *req_mem, *sid;   // sid NULL = uniform
    long long n_pods;
    int uniform_strategy;
    int capacity;
    // cold state when it does not fit shared memory: [n] each
    double *g_avail_cpu, *g_avail_mem;
    long long *g_pc;
    // spilled hot state (only used by the NPT == 0 variant): [n] each
    double *g_ta, *g_tb;
    unsigned long long *g_key;   // [3][n]
    long long *out_idx;
    double *out_score;
    long long *out_pc;
    double *out_cu, *out_mu;
    const uint32_t *skip;        // device flag: non-zero = the fast path (kp_greedy_fast.cu) already did the work
};

struct Hot {
    double ta, tb;
    unsigned long long key[kMaxStrategies];
};

struct Cold {               // per-node constants + the mutable pod count
    double *avail_cpu, *avail_mem;
    long long *mp, *pc;
};

// hot state from (pc, mp, cu, mu): initial table, arbitrary usage values (scheduler.py:150-151
// need not hold for the caller's table)
template <bool MULTI>
__device__ __forceinline__ Hot hot_initial(const GreedyParams &g, int n, double avail_cpu, double avail_mem,
                                           long long pc, long long mp, double cu, double mu)
{
    Hot h;
    bool ok = g.ready[n] != 0;
    h.ta = h.tb = __longlong_as_double(0x7ff0000000000000ll);
    if (g.capacity) {
        h.ta = __ddiv_rn(__dmul_rn(avail_cpu, __dsub_rn(100.0, cu)), 100.0);     // scheduler.py:232
        h.tb = __ddiv_rn(__dmul_rn(avail_mem, __dsub_rn(100.0, mu)), 100.0);     // scheduler.py:233
        ok = ok && pc < mp;            // NaN thresholds fail the compares on their own
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        h.key[s] = 0ull;
        if (!MULTI && s != g.uniform_strategy) continue;
        if (ok) h.key[s] = node_score(s, cu, mu, pc, mp).key;
    }
    return h;
}

// hot state after a commit: cu == mu == (pc / mp) * 50, so 100 - usage is shared.  Every
// value is produced by the same operation sequence as free_cpu_cores / free_mem_gib /
// score_resource_balanced / score_least_loaded would produce from (cu, mu) = (u, u).
template <bool MULTI>
__device__ __forceinline__ Hot hot_after_commit(const GreedyParams &g, double avail_cpu, double avail_mem,
                                                long long pc, long long mp)
{
    Hot h;
    const double u = usage_a3(pc, mp);
    const double hm = __dsub_rn(100.0, u);
    bool ok = true;                                   // the node was just chosen, so it is Ready
    h.ta = h.tb = __longlong_as_double(0x7ff0000000000000ll);
    if (g.capacity) {
        h.ta = __ddiv_rn(__dmul_rn(avail_cpu, hm), 100.0);
        h.tb = __ddiv_rn(__dmul_rn(avail_mem, hm), 100.0);
        ok = pc < mp;
    }
    h.key[0] = h.key[1] = h.key[2] = 0ull;
    if (ok) {
        if (MULTI || g.uniform_strategy == 0) {
            const double cs = __ddiv_rn(hm, 100.0);
            const double ps = mp > 0 ? __ddiv_rn((double)(mp - pc), (double)mp) : 0.0;
            const double a = __dmul_rn(cs, 0.35);
            const double sc = __dadd_rn(__dadd_rn(a, a), __dmul_rn(ps, 0.30));       // scheduler.py:541
            h.key[0] = sc > -1.0 ? f64_order_key(sc) : 0ull;
        }
        if (MULTI || g.uniform_strategy == 1) {
            const double sc = __dadd_rn(hm, hm);                                       // scheduler.py:543
            h.key[1] = sc > -1.0 ? f64_order_key(sc) : 0ull;
        }
        if (MULTI || g.uniform_strategy == 2) h.key[2] = pc > -1 ? i64_order_key(pc) : 0ull;   // :545
    }
    return h;
}

__device__ __forceinline__ double key_score(int s, unsigned long long k)
{
    if (s == 2) return (double)(long long)(k ^ 0x8000000000000000ull);
    const unsigned long long b = (k >> 63) ? (k ^ 0x8000000000000000ull) : ~k;
    return __longlong_as_double((long long)b);
}

// (key desc, idx asc) argmax over the warp with three hardware reductions
__device__ __forceinline__ void warp_argmax(unsigned long long &key, uint32_t &idx)
{
    const uint32_t hi = (uint32_t)(key >> 32), lo = (uint32_t)key;
    const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    const uint32_t mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? idx : 0xffffffffu);
    key = ((unsigned long long)mh << 32) | ml;
    idx = mi;
}

// NPT > 0: hot state of nodes {k * 1024 + tid} in registers.  NPT == 0: hot state in global memory.
// SMEM_COLD: cold state in dynamic shared memory (n <= kGreedySmemNodes), else in global memory.
template <int NPT, bool MULTI, bool SMEM_COLD>
__global__ void __launch_bounds__(kGreedyThreads, 1)
greedy_kernel(const GreedyParams g)
{
    extern __shared__ __align__(16) unsigned char cold_raw[];
    __shared__ double s_a[kPodBatch], s_b[kPodBatch];
    __shared__ int s_s[kPodBatch];
    __shared__ unsigned long long w_key[2][32];
    __shared__ uint32_t w_idx[2][32];

    if (g.skip && *g.skip) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int KS = MULTI ? kMaxStrategies : 1;
    constexpr int NREG = NPT > 0 ? NPT : 1;
    double ta[NREG], tb[NREG];
    unsigned long long key[KS][NREG];

    Cold c;
    if (SMEM_COLD) {
        c.avail_cpu = reinterpret_cast<double *>(cold_raw);
        c.avail_mem = c.avail_cpu + g.n_nodes;
        c.mp = reinterpret_cast<long long *>(c.avail_mem + g.n_nodes);
        c.pc = c.mp + g.n_nodes;
    } else {
        c.avail_cpu = g.g_avail_cpu;
        c.avail_mem = g.g_avail_mem;
        c.mp = const_cast<long long *>(g.max_pods);
        c.pc = g.g_pc;
    }
    const int us = g.uniform_strategy;

    // ---- initial state ----
    const int per_thread = NPT > 0 ? NPT : (g.n_nodes + kGreedyThreads - 1) / kGreedyThreads;
#pragma unroll
    for (int k = 0; k < (NPT > 0 ? NPT : 1); ++k) {
        for (int kk = k; kk < per_thread; kk += (NPT > 0 ? per_thread : 1)) {
            const int n = kk * kGreedyThreads + tid;
            Hot h;
            h.ta = h.tb = 0.0;
            h.key[0] = h.key[1] = h.key[2] = 0ull;
            if (n < g.n_nodes) {
                const long long pc = g.pod_count0[n], mp = g.max_pods[n];
                const double avail_cpu = __ddiv_rn((double)g.alloc_cpu[n], 1000.0);                       // :174-175
                const double avail_mem = bytes_to_gib(g.alloc_mem[n]);                                  // :187
                c.avail_cpu[n] = avail_cpu;
                c.avail_mem[n] = avail_mem;
                if (SMEM_COLD) c.mp[n] = mp;
                c.pc[n] = pc;
                h = hot_initial<MULTI>(g, n, avail_cpu, avail_mem, pc, mp, g.cu0[n], g.mu0[n]);
            }
            if (NPT > 0) {
                ta[k] = h.ta; tb[k] = h.tb;
                if (MULTI) { key[0][k] = h.key[0]; key[KS > 1 ? 1 : 0][k] = h.key[1]; key[KS > 2 ? 2 : 0][k] = h.key[2]; }
                else key[0][k] = us == 0 ? h.key[0] : (us == 1 ? h.key[1] : h.key[2]);
            } else if (n < g.n_nodes) {
                g.g_ta[n] = h.ta; g.g_tb[n] = h.tb;
#pragma unroll
                for (int s = 0; s < kMaxStrategies; ++s) g.g_key[(size_t)s * g.n_nodes + n] = h.key[s];
            }
        }
    }
    __syncthreads();

    for (long long base = 0; base < g.n_pods; base += kPodBatch) {
        const int cnt = (int)min((long long)kPodBatch, g.n_pods - base);
        __syncthreads();                       // previous batch fully consumed
        if (tid < cnt) {
            s_a[tid] = req_cores(g.req_cpu[base + tid]);
            s_b[tid] = req_gib(g.req_mem[base + tid]);
            s_s[tid] = g.sid ? canon_strategy(g.sid[base + tid]) : us;
        }
        __syncthreads();
        for (int j = 0; j < cnt; ++j) {
            const double a = s_a[j], b = s_b[j];
            const int s = s_s[j];
            unsigned long long bk = 0ull;
            uint32_t bi = 0xffffffffu;
            if (NPT > 0) {
#pragma unroll
                for (int k = 0; k < NREG; ++k) {
                    unsigned long long kk;
                    if (MULTI) kk = s == 0 ? key[0][k] : (s == 1 ? key[KS > 1 ? 1 : 0][k] : key[KS > 2 ? 2 : 0][k]);
                    else kk = key[0][k];
                    const bool fits = a <= ta[k] && b <= tb[k];
                    if (fits && kk > bk) { bk = kk; bi = (uint32_t)(k * kGreedyThreads + tid); }
                }
            } else {
                const unsigned long long *gk = g.g_key + (size_t)s * g.n_nodes;
                for (int n = tid; n < g.n_nodes; n += kGreedyThreads) {
                    const unsigned long long kk = gk[n];
                    const bool fits = a <= g.g_ta[n] && b <= g.g_tb[n];
                    if (fits && kk > bk) { bk = kk; bi = (uint32_t)n; }
                }
            }
            warp_argmax(bk, bi);
            const int par = j & 1;
            if (lane == 0) { w_key[par][warp] = bk; w_idx[par][warp] = bi; }
            __syncthreads();
            bk = w_key[par][lane];
            bi = w_idx[par][lane];
            warp_argmax(bk, bi);
            // every thread now holds the block-wide winner
            if (bk == 0ull) {
                if (tid == 0) { g.out_idx[base + j] = -1; g.out_score[base + j] = -1.0; }
                continue;
            }
            if ((int)(bi % kGreedyThreads) == tid) {
                const int n = (int)bi;
                // commit (scheduler.py:598 -> :144-151)
                const long long pc = c.pc[n] + 1;
                c.pc[n] = pc;
                const Hot h = hot_after_commit<MULTI>(g, c.avail_cpu[n], c.avail_mem[n], pc, c.mp[n]);
                if (NPT > 0) {
                    const int k = n / kGreedyThreads;
#pragma unroll
                    for (int kk = 0; kk < NREG; ++kk) {
                        if (kk == k) {
                            ta[kk] = h.ta; tb[kk] = h.tb;
                            if (MULTI) { key[0][kk] = h.key[0]; key[KS > 1 ? 1 : 0][kk] = h.key[1]; key[KS > 2 ? 2 : 0][kk] = h.key[2]; }
                            else key[0][kk] = us == 0 ? h.key[0] : (us == 1 ? h.key[1] : h.key[2]);
                        }
                    }
                } else {
                    g.g_ta[n] = h.ta; g.g_tb[n] = h.tb;
#pragma unroll
                    for (int q = 0; q < kMaxStrategies; ++q) g.g_key[(size_t)q * g.n_nodes + n] = h.key[q];
                }
                g.out_idx[base + j] = n;
                g.out_score[base + j] = key_score(s, bk);
            }
            if (NPT == 0) __syncthreads();     // global hot state must be visible to the block
        }
    }
    __syncthreads();
    // final state: usage follows scheduler.py:150-151 for every node that received a pod
    for (int n = tid; n < g.n_nodes; n += kGreedyThreads) {
        const long long pc = c.pc[n];
        const bool touched = pc != g.pod_count0[n];
        const double u = usage_a3(pc, g.max_pods[n]);
        if (g.out_pc) g.out_pc[n] = pc;
        if (g.out_cu) g.out_cu[n] = touched ? u : g.cu0[n];
        if (g.out_mu) g.out_mu[n] = touched ? u : g.mu0[n];
    }
}

// workspace: avail_cpu[n] | avail_mem[n] | pc[n] | g_ta[n] | g_tb[n] | g_key[3][n]
struct GreedyLayout { size_t off_acpu, off_amem, off_pc, off_ta, off_tb, off_key, total; };
static GreedyLayout greedy_layout(int64_t n)
{
    GreedyLayout L{};
    size_t o = 0;
    L.off_acpu = o; o = align_up(o + sizeof(double) * n);
    L.off_amem = o; o = align_up(o + sizeof(double) * n);
    L.off_pc = o;   o = align_up(o + sizeof(long long) * n);
    L.off_ta = o;   o = align_up(o + sizeof(double) * n);
    L.off_tb = o;   o = align_up(o + sizeof(double) * n);
    L.off_key = o;  o = align_up(o + sizeof(unsigned long long) * kMaxStrategies * n);
    L.total = std::max<size_t>(o, kAlign);
    return L;
}

size_t greedy_workspace_bytes(int64_t n_nodes, int64_t n_pods)
{
    if (n_nodes < 0 || n_pods < 0) return 0;
    return greedy_layout(n_nodes).total + greedy_fast_workspace_bytes(n_nodes, n_pods);
}

template <int NPT, bool MULTI>
static int launch_greedy(const GreedyParams &g, cudaStream_t stream)
{
    if (g.n_nodes <= kGreedySmemNodes) {
        const size_t smem = (size_t)kGreedyColdBytes * g.n_nodes;
        static bool attr_set[kMaxDevices] = {false};                 // the opt-in is per device
        const int dev = device_slot();
        if (!attr_set[dev]) {
            KP_CUDA(cudaFuncSetAttribute(greedy_kernel<NPT, MULTI, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kGreedyColdBytes * kGreedySmemNodes));
            attr_set[dev] = true;
        }
        greedy_kernel<NPT, MULTI, true><<<1, kGreedyThreads, smem, stream>>>(g);
    } else {
        greedy_kernel<NPT, MULTI, false><<<1, kGreedyThreads, 0, stream>>>(g);
    }
    KP_LAUNCH_CHECK("greedy_kernel");
    return KP_OK;
}

int greedy_place_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                        int filter_mode, int64_t *out_idx, double *out_score,
                        int64_t *out_pod_count, double *out_cu, double *out_mu, void *ws,
                        size_t ws_bytes, cudaStream_t stream)
{
    const int64_t n = nodes.n_nodes;
    const GreedyLayout L = greedy_layout(n);
    const size_t need_bytes = greedy_workspace_bytes(n, pods.n_pods);
    if (!ws || ws_bytes < need_bytes || (reinterpret_cast<uintptr_t>(ws) & (kAlign - 1))) {
        set_error("greedy workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", need_bytes,
                  ws_bytes, ws);
        return KP_ERR_WORKSPACE;
    }
    unsigned char *w = static_cast<unsigned char *>(ws);
    GreedyParams g{};
    g.alloc_cpu = (const long long *)nodes.alloc_cpu_milli;
    g.alloc_mem = (const long long *)nodes.alloc_mem_bytes;
    g.pod_count0 = (const long long *)nodes.pod_count;
    g.max_pods = (const long long *)nodes.max_pods;
    g.ready = nodes.ready;
    g.cu0 = nodes.cpu_usage_pct;
    g.mu0 = nodes.mem_usage_pct;
    g.n_nodes = (int)n;
    g.req_cpu = (const long long *)pods.req_cpu_milli;
    g.req_mem = (const long long *)pods.req_mem_bytes;
    g.sid = (const long long *)pods.strategy_ids;
    g.n_pods = pods.n_pods;
    g.uniform_strategy = canon_strategy(strategy);
    g.capacity = filter_mode == KP_FILTER_CAPACITY;
    g.g_avail_cpu = reinterpret_cast<double *>(w + L.off_acpu);
    g.g_avail_mem = reinterpret_cast<double *>(w + L.off_amem);
    g.g_pc = reinterpret_cast<long long *>(w + L.off_pc);
    g.g_ta = reinterpret_cast<double *>(w + L.off_ta);
    g.g_tb = reinterpret_cast<double *>(w + L.off_tb);
    g.g_key = reinterpret_cast<unsigned long long *>(w + L.off_key);
    g.out_idx = (long long *)out_idx;
    g.out_score = out_score;
    g.out_pc = (long long *)out_pod_count;
    g.out_cu = out_cu;
    g.out_mu = out_mu;
    if (pods.n_pods == 0 && !out_pod_count && !out_cu && !out_mu) return KP_OK;

    // fast path first (integer state machine); the fp64 kernel below runs only if its flag stays 0
    const int fst = greedy_fast_device(nodes, pods, strategy, filter_mode, out_idx, out_score, out_pod_count, out_cu,
                                       out_mu, w + L.total, ws_bytes - L.total, stream, &g.skip);
    if (fst != KP_OK) return fst;

    const bool multi = pods.strategy_ids != nullptr;
    const int npt = (int)((n + kGreedyThreads - 1) / kGreedyThreads);
    if (!multi) {
        if (npt <= 1) return launch_greedy<1, false>(g, stream);
        if (npt <= 2) return launch_greedy<2, false>(g, stream);
        if (npt <= 3) return launch_greedy<3, false>(g, stream);
        if (npt <= 5) return launch_greedy<5, false>(g, stream);
        return launch_greedy<0, false>(g, stream);
    }
    if (npt <= 1) return launch_greedy<1, true>(g, stream);
    if (npt <= 2) return launch_greedy<2, true>(g, stream);
    if (npt <= 3) return launch_greedy<3, true>(g, stream);
    return launch_greedy<0, true>(g, stream);
}

}  // namespace kp

// Greedy placement, fast path: the sequential commit loop as an all-integer state machine.
//
// Same contract as kp_greedy.cu (reference loop scheduler.py:690-729 -> :598 -> :144-151,
// capacity filter of SURVEY.md A13); bit-identical results, different mechanics:
//
//  * After its first commit a node's state depends only on its SHAPE (allocatable cpu,
//    allocatable memory, max_pods) and its pod count.  Nodes of one shape share a TRAJECTORY
//    of states indexed by pod count, so a commit is a shared-memory table lookup instead of a
//    chain of fp64 divisions on the critical path.  (Clusters are built from instance types:
//    a handful of shapes.  More than 128 shapes, or trajectories longer than 4096 states in
//    total, fall back to kp_greedy.cu.)
//  * Every state that can occur -- each node's initial state and every trajectory state --
//    is computed ONCE with the exact fp64 routines, then rank-compressed (radix sort + scan of
//    run starts): thresholds become dense ranks packed in one 32-bit word with guard bits (the
//    snapshot kernel's SWAR test), scores become dense ranks, and a node's key is
//    (score rank << idx bits) | (idx mask - node index): ONE unsigned compare orders nodes by
//    (score desc, index asc) -- the reference's strict '>' with first-in-list-order ties.
//  * A node's hot state is therefore two registers; the per-pod block arg-max is one REDUX.MAX
//    per warp, one barrier, one more REDUX.MAX; the winner's index is read back out of the key.
//  * Tables that exceed one CTA's registers run on a thread-block CLUSTER (2, 4 or 8 CTAs):
//    per pod each CTA sends its best key to every peer with st.async (the word lands in the
//    peer's shared memory and completes a transaction on the peer's mbarrier) and then waits
//    on its own mbarrier only -- no cluster barrier on the per-pod path.
//
//  * EXACT BATCHING of a pod-independent prefix (one strategy, resource_balanced / least_loaded).  While every
//    pod fits every state a node can still reach, a decision does not depend on the pod: the winner is the best
//    (score, index) key among the nodes' current states, and a commit replaces that node's state by the next one
//    of its trajectory, whose score is lower.  The sequence of chosen nodes is therefore the merge of the nodes'
//    state streams by (score desc, index asc) -- ONE descending sort of all reachable (node, state) entries by the
//    same 32-bit key the loop compares -- and pod j of the prefix simply gets entry j (or no node once the entries
//    run out: the cluster is full).  The prefix ends at the first pod that does not fit the tightest reachable
//    state; from there the sequential loop continues on the advanced node states.  Config 3 (every pod fits every
//    node with room) is one such prefix from the first pod to the last: a sort for the whole GPU instead of a
//    million dependent steps on one SM.  Preconditions (monotone streams, every reachable state choosable, entry
//    count within the workspace) are checked on the device; pairs are not evaluated at all on this path, which
//    bench.py reports separately from the pair rate of the scanning kernels.
//
// Whether the preconditions hold is decided on the device (ok flag); the fp64 kernel is always
// launched behind this path and returns at once when the flag says the work is done.
#include <algorithm>
#include <cooperative_groups.h>

#include "kp_common.cuh"

namespace kp {

namespace cg = cooperative_groups;

constexpr int kFastMaxShapes = 128;
constexpr int kFastMaxTraj = 4096;             // trajectory states kept in shared memory
constexpr int kFastPodBatch = 1024;
constexpr uint32_t kFastGuards = 0x80008000u;
constexpr uint32_t kFastMaxDistinct = 32000;
constexpr uint32_t kNoEntry = 0xffffffffu;

struct FastHeader {
    uint32_t ok;               // 1 = fast path valid (set by the prep kernels, read by both greedy kernels)
    uint32_t n_shapes;
    uint32_t traj_total;
    uint32_t distinct[2];
    uint32_t score_ranks[kMaxStrategies];
    uint32_t wmax_lo, wmax_hi;  // largest threshold rank field (rank + 1) of any state that exists: a pod beyond it fits nothing, ever
    uint32_t traj_base[kFastMaxShapes + 1];
    long long shape_cpu[kFastMaxShapes], shape_mem[kFastMaxShapes], shape_mp[kFastMaxShapes];
};

struct FastIn {
    const long long *alloc_cpu, *alloc_mem, *pod_count0, *max_pods;
    const uint8_t *ready;
    const double *cu0, *mu0;
    int n;
    const long long *req_cpu, *req_mem, *sid;
    long long n_pods;
    int uniform_strategy, strat_mask;
};

// ---------------------------------------------------------------------------------------
// 1. shapes: dedupe (alloc_cpu, alloc_mem, max_pods) over the nodes that can ever be chosen
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1)
fast_shapes_kernel(FastIn in, FastHeader *hdr, uint32_t *__restrict__ shape_id)
{
    __shared__ long long s_cpu[kFastMaxShapes], s_mem[kFastMaxShapes], s_mp[kFastMaxShapes];
    __shared__ int s_count, s_pending, s_overflow;
    const int tid = threadIdx.x;
    if (tid == 0) { s_count = 0; s_pending = 0x7fffffff; s_overflow = 0; }
    __syncthreads();
    bool bad_value = false;
    for (int base = 0; base < in.n; base += 1024) {
        const int n = base + tid;
        long long c = 0, m = 0, mp = 0;
        bool want = false;
        if (n < in.n) {
            c = in.alloc_cpu[n]; m = in.alloc_mem[n]; mp = in.max_pods[n];
            const long long pc = in.pod_count0[n];
            want = in.ready[n] != 0 && mp > 0 && pc < mp;            // otherwise never chosen, never committed
            if (want && (pc < 0 || mp >= (1ll << 24))) bad_value = true;
            // choosable under the entries' rule (ready && pc < mp) but without a trajectory: max_pods <= 0 with
            // a negative pod count.  Such a table goes to the fp64 kernel.
            if (in.ready[n] != 0 && pc < mp && mp <= 0) bad_value = true;
        }
        int found = -1;
        for (;;) {
            const int cnt = s_count;
            if (want && found < 0)
                for (int q = 0; q < cnt; ++q)
                    if (s_cpu[q] == c && s_mem[q] == m && s_mp[q] == mp) { found = q; break; }
            if (want && found < 0) atomicMin(&s_pending, n);
            __syncthreads();
            const int who = s_pending;
            if (who == 0x7fffffff) break;                            // block-uniform
            if (n == who) {
                if (cnt < kFastMaxShapes) { s_cpu[cnt] = c; s_mem[cnt] = m; s_mp[cnt] = mp; s_count = cnt + 1; }
                else { s_overflow = 1; }
            }
            __syncthreads();
            if (tid == 0) s_pending = 0x7fffffff;
            if (s_overflow) { want = false; }
            __syncthreads();
        }
        __syncthreads();
        if (n < in.n) shape_id[n] = found >= 0 ? (uint32_t)found : kNoEntry;
    }
    const int any_bad = __syncthreads_or(bad_value ? 1 : 0);
    if (tid == 0) {
        uint32_t total = 0;
        const int cnt = s_count;
        for (int q = 0; q < cnt; ++q) {
            hdr->traj_base[q] = total;
            hdr->shape_cpu[q] = s_cpu[q]; hdr->shape_mem[q] = s_mem[q]; hdr->shape_mp[q] = s_mp[q];
            total += (uint32_t)s_mp[q] + 1u;                          // states pod_count = 0 .. max_pods
            if (total > (1u << 30)) total = 1u << 30;
        }
        hdr->traj_base[cnt] = total;
        hdr->n_shapes = (uint32_t)cnt;
        hdr->traj_total = total;
        hdr->wmax_lo = 0u;
        hdr->wmax_hi = 0u;
        hdr->ok = (!s_overflow && !any_bad && total <= (uint32_t)kFastMaxTraj) ? 1u : 0u;
    }
}

// ---------------------------------------------------------------------------------------
// 2. every state that can occur: entries [0, n) = the nodes' initial states,
//    entries [n, n + kFastMaxTraj) = trajectory states (shape, pod_count)
// ---------------------------------------------------------------------------------------
struct FastEntries {
    unsigned long long *key_t[2];                 // order keys of free cpu / free memory
    unsigned long long *key_s[kMaxStrategies];    // order keys of the score (0 = state cannot be chosen)
    double *score[kMaxStrategies];                // fp64 score of the state
};

__global__ void __launch_bounds__(256)
fast_entries_kernel(FastIn in, const FastHeader *__restrict__ hdr, FastEntries e, uint32_t *__restrict__ iota)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int total = in.n + kFastMaxTraj;
    if (i >= total) return;
    iota[i] = (uint32_t)i;
    double cu = 0, mu = 0, avail_cpu = 0, avail_mem = 0;
    long long pc = 0, mp = 0;
    bool exists = false, ready = true;
    if (i < in.n) {
        exists = true;
        cu = in.cu0[i]; mu = in.mu0[i]; pc = in.pod_count0[i]; mp = in.max_pods[i];
        ready = in.ready[i] != 0;
        avail_cpu = __ddiv_rn((double)in.alloc_cpu[i], 1000.0);
        avail_mem = bytes_to_gib(in.alloc_mem[i]);
    } else if (hdr->ok) {
        const uint32_t t = (uint32_t)(i - in.n);
        if (t < hdr->traj_total) {
            int s = 0;
            const int ns = (int)hdr->n_shapes;
            while (s + 1 < ns && hdr->traj_base[s + 1] <= t) ++s;
            exists = true;
            mp = hdr->shape_mp[s];
            pc = (long long)(t - hdr->traj_base[s]);
            cu = mu = usage_a3(pc, mp);                                // scheduler.py:150-151
            avail_cpu = __ddiv_rn((double)hdr->shape_cpu[s], 1000.0);
            avail_mem = bytes_to_gib(hdr->shape_mem[s]);
        }
    }
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    double ta = inf, tb = inf;
    bool ok = false;
    if (exists) {
        ta = __ddiv_rn(__dmul_rn(avail_cpu, __dsub_rn(100.0, cu)), 100.0);   // scheduler.py:232
        tb = __ddiv_rn(__dmul_rn(avail_mem, __dsub_rn(100.0, mu)), 100.0);   // scheduler.py:233
        ok = ready && pc < mp && ta == ta && tb == tb;
        if (ta != ta) ta = inf;
        if (tb != tb) tb = inf;
    }
    e.key_t[0][i] = f64_order_key(ta);
    e.key_t[1][i] = f64_order_key(tb);
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(in.strat_mask & (1 << s))) continue;
        NodeScore ns;
        ns.key = 0ull; ns.score = -1.0;
        if (exists) ns = node_score(s, cu, mu, pc, mp);
        e.key_s[s][i] = ok ? ns.key : 0ull;
        e.score[s][i] = ns.score;
    }
}

// after a sort of one key array: run starts of the sorted keys
__global__ void __launch_bounds__(256)
fast_flags_kernel(const unsigned long long *__restrict__ kout, int total, uint32_t *__restrict__ flag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < total) flag[i] = (i == 0 || kout[i] != kout[i - 1]) ? 1u : 0u;
}

// dense rank back to the entries; unique values (thresholds only); number of distinct values
__global__ void __launch_bounds__(256)
fast_rank_kernel(const unsigned long long *__restrict__ kout, const uint32_t *__restrict__ vout,
                 const uint32_t *__restrict__ flag, const uint32_t *__restrict__ incl, int total,
                 uint32_t *__restrict__ rank, double *__restrict__ uniq, uint32_t *__restrict__ n_distinct)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const uint32_t d = incl[i] - 1;
    rank[vout[i]] = d;
    if (uniq && flag[i]) {
        const unsigned long long k = kout[i];
        const unsigned long long b = (k >> 63) ? (k ^ 0x8000000000000000ull) : ~k;
        uniq[d] = __longlong_as_double((long long)b);
    }
    if (i == total - 1) *n_distinct = d + 1;
}

// packed threshold word and per-strategy score rank of every entry; final validity check
__global__ void __launch_bounds__(256)
fast_pack_kernel(FastIn in, FastHeader *hdr, const uint32_t *__restrict__ rank_t0, const uint32_t *__restrict__ rank_t1,
                 const uint32_t *__restrict__ rank_s, const unsigned long long *__restrict__ key_s0, int total,
                 int idx_bits, uint32_t *__restrict__ entry_w, uint32_t *__restrict__ entry_key)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        bool ok = hdr->ok != 0 && hdr->distinct[0] <= kFastMaxDistinct && hdr->distinct[1] <= kFastMaxDistinct;
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s)
            if ((in.strat_mask & (1 << s)) && (hdr->score_ranks[s] + 1u) >= (1u << (32 - idx_bits))) ok = false;
        hdr->ok = ok ? 1u : 0u;
    }
    // field-wise maximum over the states that exist (node rows and trajectory states; the unused tail of the
    // trajectory table carries +inf thresholds and must not count)
    const bool exists = i < total && (i < in.n || (uint32_t)(i - in.n) < hdr->traj_total);
    const uint32_t mlo = __reduce_max_sync(0xffffffffu, exists ? rank_t0[i] + 1u : 0u);
    const uint32_t mhi = __reduce_max_sync(0xffffffffu, exists ? rank_t1[i] + 1u : 0u);
    if ((threadIdx.x & 31) == 0) {
        if (mlo) atomicMax(&hdr->wmax_lo, mlo);
        if (mhi) atomicMax(&hdr->wmax_hi, mhi);
    }
    if (i >= total) return;
    entry_w[i] = (((rank_t1[i] + 1u) | 0x8000u) << 16) | ((rank_t0[i] + 1u) | 0x8000u);
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(in.strat_mask & (1 << s))) continue;
        const bool can = key_s0[(size_t)s * total + i] != 0ull;
        entry_key[(size_t)s * total + i] = can ? rank_s[(size_t)s * total + i] + 1u : 0u;     // 0 = cannot be chosen
    }
}

__device__ __forceinline__ int lower_bound_lt_g(const double *arr, int n, double x)
{
    int lo = 0;
    for (int step = 1 << (31 - __clz(n | 1)); step > 0; step >>= 1) {
        const int nxt = lo + step;
        if (nxt <= n && arr[nxt - 1] < x) lo = nxt;
    }
    return lo;
}

// pods -> packed request ranks
__global__ void __launch_bounds__(256)
fast_pods_kernel(FastIn in, const FastHeader *__restrict__ hdr, const double *__restrict__ uniq0,
                 const double *__restrict__ uniq1, uint32_t *__restrict__ xs)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= in.n_pods || !hdr->ok) return;
    const uint32_t ra = (uint32_t)lower_bound_lt_g(uniq0, (int)hdr->distinct[0], req_cores(in.req_cpu[j]));
    const uint32_t rb = (uint32_t)lower_bound_lt_g(uniq1, (int)hdr->distinct[1], req_gib(in.req_mem[j]));
    xs[j] = ((rb + 1u) << 16) | (ra + 1u);
}

// ---------------------------------------------------------------------------------------
// 2b. exact batching of the pod-independent prefix
// ---------------------------------------------------------------------------------------
constexpr long long kBatchMinPods = 16384;      // host: calls with fewer pods skip the attempt
constexpr unsigned long long kBatchMinRun = 4096;   // device: a shorter prefix is left to the loop

struct BatchHeader {
    uint32_t ok;                     // streams monotone, reachable states choosable, entry count within the buffers
    uint32_t wmin;                   // field-wise minimum threshold word over every reachable choosable state
    unsigned long long jstar;        // first pod that does not fit wmin (n_pods if none)
    unsigned long long total;        // number of (node, state) entries
    unsigned long long cap;          // entries the buffers hold
};

// pods decided by the sorted prefix (0: batching not used); the same predicate in every kernel that needs it
__device__ __forceinline__ unsigned long long batch_n_done(const FastHeader *hdr, const BatchHeader *b)
{
    if (!b || !hdr->ok || !b->ok || b->total > b->cap || b->jstar < kBatchMinRun) return 0ull;
    return b->jstar;
}

// One block.  Per choosable node: entry count = max_pods - pod_count; per shape: the lowest pod count any of its
// choosable nodes starts from (every trajectory state above it is reachable).  Then every reachable state is
// checked: it must be choosable, its key must not rise along the stream, and its threshold word joins wmin.
__global__ void __launch_bounds__(1024, 1)
batch_check_kernel(FastIn in, const FastHeader *__restrict__ hdr, const uint32_t *__restrict__ shape_id, const uint32_t *__restrict__ entry_w,
                   const uint32_t *__restrict__ entry_key, int total, uint32_t *__restrict__ cnt, BatchHeader *b, unsigned long long cap)
{
    __shared__ unsigned s_minpc[kFastMaxShapes];
    __shared__ unsigned s_wmin_lo, s_wmin_hi, s_bad;
    __shared__ unsigned long long s_total;
    const int tid = threadIdx.x;
    if (tid < kFastMaxShapes) s_minpc[tid] = 0xffffffffu;
    if (tid == 0) { s_wmin_lo = 0xffffu; s_wmin_hi = 0xffffu; s_bad = 0u; s_total = 0ull; }
    __syncthreads();
    const bool fast_ok = hdr->ok != 0;
    const uint32_t *key = entry_key + (size_t)in.uniform_strategy * total;
    unsigned wlo = 0xffffu, whi = 0xffffu, bad = 0u;
    unsigned long long mine_total = 0ull;
    for (int i = tid; i < in.n; i += 1024) {
        uint32_t c = 0u;
        const uint32_t sh = shape_id[i];
        if (fast_ok && sh != kNoEntry && key[i] != 0u) {                  // choosable now
            const long long pc = in.pod_count0[i], mp = in.max_pods[i];
            c = (uint32_t)(mp - pc);
            atomicMin(&s_minpc[sh & 0x7fu], (unsigned)pc);
            const uint32_t w = entry_w[i];
            wlo = min(wlo, w & 0xffffu); whi = min(whi, w >> 16);
            if (pc + 1 < mp) {                                            // the stream must not rise at its first commit
                const uint32_t nxt = key[in.n + hdr->traj_base[sh & 0x7fu] + (uint32_t)pc + 1u];
                if (nxt > key[i]) bad = 1u;
            }
        }
        cnt[i] = c;
        mine_total += c;
    }
    __syncthreads();                                                       // s_minpc complete
    if (fast_ok) {
        const int ns = (int)hdr->n_shapes;
        for (int t = tid; t < (int)hdr->traj_total; t += 1024) {
            int sh = 0;
            while (sh + 1 < ns && hdr->traj_base[sh + 1] <= (uint32_t)t) ++sh;
            const unsigned pc = (unsigned)t - hdr->traj_base[sh];
            const unsigned mp = (unsigned)hdr->shape_mp[sh];
            if (s_minpc[sh] == 0xffffffffu || pc <= s_minpc[sh] || pc >= mp) continue;   // not reachable, or full
            const uint32_t k = key[in.n + t];
            if (k == 0u) bad = 1u;                                         // a state that can never be chosen ends a stream early
            if (pc > s_minpc[sh] + 1u && k > key[in.n + t - 1]) bad = 1u;  // keys must not rise along the trajectory
            const uint32_t w = entry_w[in.n + t];
            wlo = min(wlo, w & 0xffffu); whi = min(whi, w >> 16);
        }
    }
    atomicMin(&s_wmin_lo, wlo);
    atomicMin(&s_wmin_hi, whi);
    if (bad) s_bad = 1u;
    atomicAdd(&s_total, mine_total);
    __syncthreads();
    if (tid == 0) {
        b->ok = (fast_ok && !s_bad) ? 1u : 0u;
        b->wmin = (s_wmin_hi << 16) | s_wmin_lo;
        b->jstar = (unsigned long long)in.n_pods;
        b->total = s_total;
        b->cap = cap;
    }
}

// first pod whose request does not fit the tightest reachable state
__global__ void __launch_bounds__(256)
batch_violation_kernel(FastIn in, const FastHeader *__restrict__ hdr, const uint32_t *__restrict__ xs, BatchHeader *b)
{
    if (!hdr->ok || !b->ok) return;
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    bool viol = false;
    if (j < in.n_pods) viol = ((~(b->wmin - xs[j])) & kFastGuards) != 0u;
    const unsigned m = __ballot_sync(0xffffffffu, viol);
    if (m && (threadIdx.x & 31) == __ffs(m) - 1) atomicMin(&b->jstar, (unsigned long long)j);
}

// one warp per node: its entries (initial state, then the trajectory states up to max_pods - 1) with the loop's own key
__global__ void __launch_bounds__(256)
batch_entries_kernel(FastIn in, const FastHeader *__restrict__ hdr, const BatchHeader *__restrict__ b, const uint32_t *__restrict__ shape_id,
                     const uint32_t *__restrict__ entry_key, int total, int idx_bits, const uint32_t *__restrict__ cnt,
                     const uint32_t *__restrict__ off, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    if (batch_n_done(hdr, b) == 0ull) return;
    const int node = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (node >= in.n) return;
    const uint32_t c = cnt[node];
    if (c == 0u) return;
    const uint32_t *key = entry_key + (size_t)in.uniform_strategy * total;
    const uint32_t idx_mask = (1u << idx_bits) - 1u;
    const uint32_t low = idx_mask - (uint32_t)node;
    const uint32_t t0 = (uint32_t)in.n + hdr->traj_base[shape_id[node] & 0x7fu] + (uint32_t)in.pod_count0[node];   // entry of state k is t0 + k
    const uint32_t o = off[node];
    for (uint32_t k = lane; k < c; k += 32) {
        const uint32_t e = k == 0u ? (uint32_t)node : t0 + k;
        keys[o + k] = (key[e] << idx_bits) | low;
        vals[o + k] = e;
    }
}

// pod j of the prefix takes sorted entry j; behind the last entry the cluster is full
__global__ void __launch_bounds__(256)
batch_assign_kernel(FastIn in, const FastHeader *__restrict__ hdr, const BatchHeader *__restrict__ b, const uint32_t *__restrict__ keys,
                    const uint32_t *__restrict__ vals, int idx_bits, long long *__restrict__ out_idx, uint32_t *__restrict__ out_entry,
                    uint32_t *__restrict__ commits)
{
    const unsigned long long done = batch_n_done(hdr, b);
    const unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= done) return;
    if (j < b->total) {
        const uint32_t idx_mask = (1u << idx_bits) - 1u;
        const int node = (int)(idx_mask - (keys[j] & idx_mask));
        out_idx[j] = node;
        out_entry[j] = vals[j];
        atomicAdd(&commits[node], 1u);
    } else {
        out_idx[j] = -1;
        out_entry[j] = kNoEntry;
    }
}

// ---------------------------------------------------------------------------------------
// 3. the greedy loop
// ---------------------------------------------------------------------------------------
struct FastRun {
    const FastHeader *hdr;
    const uint32_t *shape_id;      // [n]
    const uint32_t *entry_w;       // [total]
    const uint32_t *entry_key;     // [3][total]
    const uint32_t *xs;            // [P]
    uint32_t *out_entry;           // [P] entry index of the chosen node's state at decision time
    long long *out_idx;
    long long *out_pc;             // final pod counts (may be NULL)
    double *out_cu, *out_mu;
    int total, idx_bits, nodes_per_cta;
    const BatchHeader *batch;      // exact batching of the prefix (NULL = not attempted)
    const uint32_t *commits;       // [n] commits the prefix made to every node
};

// dynamic shared memory: traj_w[kFastMaxTraj] | traj_key[KS][kFastMaxTraj] | c_t[nodes_per_cta]
// 256 threads x NPT nodes per CTA.  Few fat warps beat many thin ones here: the per-pod cost that
// does not scale with the table (reductions, barrier, winner decode -- ~100 instructions) is paid
// per warp, so a CTA is 8 warps holding up to 80 nodes per thread (two registers per node, four
// with per-pod strategies) and larger tables add CTAs through a cluster (CSIZE), not warps.
#ifndef KP_FAST_THREADS
#define KP_FAST_THREADS 256
#endif
constexpr int kFastThreads = KP_FAST_THREADS;

// best key among this thread's nodes that the pod fits: w - x keeps both guard bits iff it fits in
// both dimensions.  Four independent maxima so the dependent-max chain does not bound the scan.
template <int NPT, int KS, int Q>
__device__ __forceinline__ uint32_t fast_scan(const uint32_t (&w)[NPT], const uint32_t (&key)[KS][NPT], uint32_t x)
{
    uint32_t b0 = 0u, b1 = 0u, b2 = 0u, b3 = 0u;
#pragma unroll
    for (int k = 0; k < NPT; ++k) {
        const uint32_t t = w[k] - x;
        const bool fits = ((~t) & kFastGuards) == 0u;
        const uint32_t kk = key[Q][k];
        if ((k & 3) == 0) { if (fits) b0 = max(b0, kk); }
        else if ((k & 3) == 1) { if (fits) b1 = max(b1, kk); }
        else if ((k & 3) == 2) { if (fits) b2 = max(b2, kk); }
        else { if (fits) b3 = max(b3, kk); }
    }
    return max(max(b0, b1), max(b2, b3));
}

// Registers cannot be indexed at run time: the committed node's slot is reached through a binary tree
// of branches (about log2(NPT) compares for the one lane that commits) instead of NPT predicated moves.
template <int NPT, int KS, int LO, int HI>
__device__ __forceinline__ void fast_set_slot(uint32_t (&w)[NPT], uint32_t (&key)[KS][NPT], int slot, uint32_t nw,
                                              const uint32_t (&nk)[KS])
{
    if constexpr (LO == HI) {
        w[LO] = nw;
#pragma unroll
        for (int q = 0; q < KS; ++q) key[q][LO] = nk[q];
    } else {
        constexpr int MID = (LO + HI) / 2;
        if (slot <= MID) fast_set_slot<NPT, KS, LO, MID>(w, key, slot, nw, nk);
        else fast_set_slot<NPT, KS, MID + 1, HI>(w, key, slot, nw, nk);
    }
}

// Per-thread summary of its choosable nodes: field-wise minimum of the threshold words and the best
// key.  A pod that fits the minimum fits every one of them, so the thread's candidate is the best key
// and the scan can be skipped -- the common case when the cluster has headroom (config 3: every pod
// fits every node with room).  The owner of a committed node recomputes its summary.
template <int NPT>
__device__ __forceinline__ void fast_summary(const uint32_t (&w)[NPT], const uint32_t (&key)[NPT], uint32_t &wmin,
                                             uint32_t &kbest)
{
    uint32_t m = 0xffffffffu, b = 0u;
#pragma unroll
    for (int k = 0; k < NPT; ++k) {
        if (key[k] != 0u) m = __vminu2(m, w[k]);
        b = max(b, key[k]);
    }
    wmin = m;
    kbest = b;
}

// one 32-bit word into CTA `target` of this cluster at the address `slot` has in the calling CTA, completing
// 4 bytes of the transaction count of the mbarrier at `bar` in the SAME target CTA
__device__ __forceinline__ void dsmem_send_u32(uint32_t slot, uint32_t bar, uint32_t target, uint32_t value)
{
    uint32_t rslot, rbar;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rslot) : "r"(slot), "r"(target));
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbar) : "r"(bar), "r"(target));
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(rslot), "r"(value),
                 "r"(rbar)
                 : "memory");
}

template <int NPT, bool MULTI, int CSIZE>
__global__ void __launch_bounds__(kFastThreads, 1)
fast_greedy_kernel(FastIn in, FastRun r)
{
    extern __shared__ __align__(16) unsigned char fsm[];
    __shared__ uint32_t s_x[kFastPodBatch + 1];
    __shared__ uint8_t s_sid[kFastPodBatch + 4];
    __shared__ uint32_t w_res[2][32];
    __shared__ uint32_t c_res[2][8];                                   // best key of every CTA of the cluster
    __shared__ __align__(8) uint64_t xbar[2];                          // mbarriers of the per-pod exchange
    __shared__ uint32_t s_tbase[kFastMaxShapes + 1];
    __shared__ uint32_t s_bound[2][32];
    __shared__ uint16_t s_j[kFastPodBatch];                            // position in the batch of every pod that is scanned
    __shared__ uint32_t s_wcnt[32];

    if (!r.hdr->ok) return;                                          // the fp64 kernel does the work instead
    constexpr int KS = MULTI ? kMaxStrategies : 1;
    constexpr int NW = kFastThreads / 32;
    constexpr bool CLUSTER = CSIZE > 1;
    cg::cluster_group cluster = cg::this_cluster();
    const int crank = CLUSTER ? (int)cluster.block_rank() : 0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int total = r.total;
    uint32_t *traj_w = reinterpret_cast<uint32_t *>(fsm);
    uint32_t *traj_key = traj_w + kFastMaxTraj;
    uint32_t *c_t = traj_key + KS * kFastMaxTraj;                      // per node: trajectory index of its state;
                                                                       // bit 31: the node has been committed to
    const int us = in.uniform_strategy;
    const uint32_t idx_mask = (1u << r.idx_bits) - 1u;
    const int cta_base = crank * r.nodes_per_cta;
    const long long n_done = (long long)batch_n_done(r.hdr, r.batch);   // pods the sorted prefix has already decided
    // dead-pod bound: field-wise largest threshold of any node that can still be chosen.  It starts from every state
    // that exists (prep kernels) and is refreshed from the LIVE node states at every pod batch: thresholds only shrink
    // along a trajectory and a full node drops out, so the bound tightens as the cluster fills -- once no node is
    // choosable it is 0 and every remaining pod is answered without a scan, a reduction or an exchange.
    uint32_t wmax_lo = r.hdr->wmax_lo, wmax_hi = r.hdr->wmax_hi;

    if (tid <= kFastMaxShapes) s_tbase[tid] = r.hdr->traj_base[tid];
    for (int i = tid; i < kFastMaxTraj; i += kFastThreads) {
        traj_w[i] = r.entry_w[in.n + i];
#pragma unroll
        for (int q = 0; q < KS; ++q) traj_key[q * kFastMaxTraj + i] = r.entry_key[(size_t)(MULTI ? q : us) * total + in.n + i];
    }
    uint32_t w[NPT], key[KS][NPT];
#pragma unroll
    for (int k = 0; k < NPT; ++k) {
        const int local = k * kFastThreads + tid;
        const int n = cta_base + local;
        w[k] = kFastGuards;                                            // fields 0: nothing fits
#pragma unroll
        for (int q = 0; q < KS; ++q) key[q][k] = 0u;
        if (local < r.nodes_per_cta && n < in.n) {
            // trajectory index of the state with this node's pod count (its NEXT state is one further);
            // nodes that can never be chosen have no shape and are never committed to
            const uint32_t sh = r.shape_id[n];
            uint32_t ct = sh == kNoEntry ? 0u : (r.hdr->traj_base[sh & 0x7fu] + (uint32_t)in.pod_count0[n]) & 0x7fffffffu;
            // the state this node is in when the loop starts: its initial entry, or -- after a batched prefix that
            // committed to it -- the trajectory state that many pods further
            const uint32_t adv = n_done ? r.commits[n] : 0u;
            uint32_t e = (uint32_t)n;
            if (adv) { ct = (ct + adv) | 0x80000000u; e = (uint32_t)in.n + (ct & 0x7fffffffu); }
            c_t[local] = ct;
            w[k] = r.entry_w[e];
#pragma unroll
            for (int q = 0; q < KS; ++q) {
                const uint32_t kr = r.entry_key[(size_t)(MULTI ? q : us) * total + e];
                key[q][k] = kr ? (kr << r.idx_bits) | (idx_mask - (uint32_t)n) : 0u;
            }
        }
    }
    uint32_t wmin = 0u, kbest = 0u;                                    // uniform strategy only (one key per node)
    if (!MULTI) fast_summary<NPT>(w, key[0], wmin, kbest);
    if (CLUSTER && tid == 0) {
        mbar_init(&xbar[0], 1);
        mbar_init(&xbar[1], 1);
        mbar_fence_init();
    }
    uint32_t round = 0;
    __syncthreads();
    if (CLUSTER) cluster_barrier();                                    // every CTA's mbarriers exist before any peer signals them
    // The live refresh of the dead-pod bound needs thresholds that never grow along a trajectory (usage rises with the
    // pod count, so they do not -- unless an allocatable is negative).  Checked here once; otherwise the bound stays
    // the static one.  (Every CTA of a cluster reads the same table and reaches the same verdict.)
    bool mono = true;
    {
        const int n_shapes = (int)r.hdr->n_shapes;
        for (int sh = warp; sh < n_shapes; sh += NW)
            for (uint32_t i = s_tbase[sh] + lane; i + 1 < s_tbase[sh + 1]; i += 32)
                mono = mono && (traj_w[i + 1] & 0x7fffu) <= (traj_w[i] & 0x7fffu) && ((traj_w[i + 1] >> 16) & 0x7fffu) <= ((traj_w[i] >> 16) & 0x7fffu);
        mono = __syncthreads_and(mono) != 0;
    }

    // Maximum of one word over the CTAs of the cluster, through DSMEM without a cluster barrier: every CTA sends
    // its word to every peer with st.async, which writes it into the peer's slot AND signals the peer's
    // mbarrier (complete_tx); each CTA then waits on its OWN mbarrier only.  A cluster barrier
    // here (even release / acquire at cluster scope) cost ~0.5 us per pod in ERRBAR + barrier
    // latency -- 40 % of the kernel; tagging plain stores and spinning was slower still.
    // Slots and mbarriers alternate by round parity: a CTA can be at most one round ahead of a
    // peer, because round r + 1 needs the peer's word of round r + 1.
    auto cluster_max = [&](uint32_t mine) -> uint32_t {
        const int xp = (int)(round & 1u);
        if (tid == 0) mbar_expect_tx(&xbar[xp], 4u * (CSIZE - 1));
        if (tid < CSIZE && tid != crank)
            dsmem_send_u32(smem_u32(&c_res[xp][crank]), smem_u32(&xbar[xp]), (uint32_t)tid, mine);
        mbar_wait(&xbar[xp], (round >> 1) & 1u);
        uint32_t g = mine;
#pragma unroll
        for (int c = 0; c < CSIZE; ++c) if (c != crank) g = max(g, c_res[xp][c]);
        ++round;
        return g;
    };

    for (long long base = n_done; base < in.n_pods; base += kFastPodBatch) {
        const int cnt = (int)min((long long)kFastPodBatch, in.n_pods - base);
        __syncthreads();                                               // the batch before is no longer read (s_x, s_sid, s_bound)
        // the dead-pod bound from the live states (every CTA of a cluster ends up with the same two values): per
        // choosable node the largest thresholds it can still have -- its current state's, and for a node that has not
        // been committed to yet also those of the first trajectory state it would move to (an initial row need not
        // lie on its shape's trajectory: the caller's usage figures are replaced at the first commit)
        if (mono) {
            uint32_t mlo = 0u, mhi = 0u;
#pragma unroll
            for (int k = 0; k < NPT; ++k) {
                bool choosable = false;
#pragma unroll
                for (int q = 0; q < KS; ++q) choosable |= key[q][k] != 0u;
                if (choosable) {
                    mlo = max(mlo, w[k] & 0x7fffu);
                    mhi = max(mhi, (w[k] >> 16) & 0x7fffu);
                    const uint32_t ct = c_t[k * kFastThreads + tid];
                    if (!(ct >> 31)) {
                        const uint32_t nw = traj_w[(ct & 0x7fffffffu) + 1u];
                        mlo = max(mlo, nw & 0x7fffu);
                        mhi = max(mhi, (nw >> 16) & 0x7fffu);
                    }
                }
            }
            mlo = __reduce_max_sync(0xffffffffu, mlo);
            mhi = __reduce_max_sync(0xffffffffu, mhi);
            if (lane == 0) { s_bound[0][warp] = mlo; s_bound[1][warp] = mhi; }
            __syncthreads();
            mlo = __reduce_max_sync(0xffffffffu, lane < NW ? s_bound[0][lane] : 0u);
            mhi = __reduce_max_sync(0xffffffffu, lane < NW ? s_bound[1][lane] : 0u);
            if (CLUSTER) {
                mlo = cluster_max(mlo);
                mhi = cluster_max(mhi);
            }
            wmax_lo = min(wmax_lo, mlo);
            wmax_hi = min(wmax_hi, mhi);
        }
        // Load the batch.  Pods beyond the bound fit nothing: they are answered here, in parallel, and the rest is
        // compacted (in order) so that the sequential loop below only walks pods that need a decision.  Every CTA of
        // a cluster sees the same requests and the same bound, hence the same list.
        int live = 0;
        for (int i0 = 0; i0 < cnt; i0 += kFastThreads) {
            const int i = i0 + tid;
            uint32_t xi = 0u;
            bool keep = false;
            if (i < cnt) {
                xi = r.xs[base + i];
                keep = !((xi & 0xffffu) > wmax_lo || (xi >> 16) > wmax_hi);
                if (!keep && crank == 0) { r.out_idx[base + i] = -1; r.out_entry[base + i] = kNoEntry; }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) s_wcnt[warp] = (uint32_t)__popc(bal);
            __syncthreads();
            uint32_t before = 0u, all = 0u;
#pragma unroll
            for (int q = 0; q < NW; ++q) {
                const uint32_t c = s_wcnt[q];
                before += q < warp ? c : 0u;
                all += c;
            }
            if (keep) {
                const int at = live + (int)before + __popc(bal & ((1u << lane) - 1u));
                s_x[at] = xi;
                s_sid[at] = (uint8_t)(in.sid ? canon_strategy(in.sid[base + i]) : us);
                s_j[at] = (uint16_t)i;
            }
            live += (int)all;
            __syncthreads();                                           // s_wcnt is rewritten by the next chunk
        }
        uint32_t x = s_x[0];
        int s = s_sid[0];
        for (int jj = 0; jj < live; ++jj) {
            const int j = (int)s_j[jj];
            uint32_t best;
            if (!MULTI) {
                const bool all_fit = ((~(wmin - x)) & kFastGuards) == 0u;
                if (__all_sync(0xffffffffu, all_fit)) best = kbest;    // warp-uniform: nobody in the warp needs the scan
                else best = fast_scan<NPT, KS, 0>(w, key, x);
            } else if (s == 0) best = fast_scan<NPT, KS, 0>(w, key, x);         // block-uniform branch per pod
            else if (s == 1) best = fast_scan<NPT, KS, (KS > 1 ? 1 : 0)>(w, key, x);
            else best = fast_scan<NPT, KS, (KS > 2 ? 2 : 0)>(w, key, x);
            const uint32_t x_next = s_x[jj + 1];                       // next pod's request, off the critical path
            const int s_next = s_sid[jj + 1];                          // (one spare element past the batch)
            const int par = jj & 1;
            best = __reduce_max_sync(0xffffffffu, best);
            if (lane == 0) w_res[par][warp] = best;
            __syncthreads();
            best = __reduce_max_sync(0xffffffffu, lane < NW ? w_res[par][lane] : 0u);
            if (CLUSTER) best = cluster_max(best);
            x = x_next;
            s = s_next;
            if (best == 0u) {
                if (crank == 0 && tid == 0) { r.out_idx[base + j] = -1; r.out_entry[base + j] = kNoEntry; }
                continue;
            }
            const int n = (int)(idx_mask - (best & idx_mask));
            const int local = n - cta_base;
            if (local >= 0 && local < r.nodes_per_cta && (local % kFastThreads) == tid) {
                // commit (scheduler.py:598 -> :144-151): one more pod, next state of the shape's trajectory
                const uint32_t ct = c_t[local];
                const uint32_t t2 = (ct & 0x7fffffffu) + 1u;
                c_t[local] = t2 | 0x80000000u;
                const uint32_t nw = traj_w[t2];
                r.out_idx[base + j] = n;
                r.out_entry[base + j] = (ct >> 31) ? (uint32_t)in.n + (ct & 0x7fffffffu) : (uint32_t)n;
                uint32_t nk[KS];
#pragma unroll
                for (int q = 0; q < KS; ++q) {
                    const uint32_t kr = traj_key[q * kFastMaxTraj + t2];
                    nk[q] = kr ? (kr << r.idx_bits) | (idx_mask - (uint32_t)n) : 0u;
                }
                fast_set_slot<NPT, KS, 0, NPT - 1>(w, key, local / kFastThreads, nw, nk);
                if (!MULTI) {
                    // The summary after the commit: the best key exactly (this node held it), the minimum
                    // thresholds only as a lower bound -- folding the node's new word in is exact while
                    // thresholds shrink with every commit (the usual case) and merely conservative
                    // otherwise: a stale-low minimum sends the thread through the full scan, never past it.
                    uint32_t b0 = 0u, b1 = 0u, b2 = 0u, b3 = 0u;          // four chains: this is on the critical path
#pragma unroll
                    for (int k = 0; k < NPT; ++k) {
                        if ((k & 3) == 0) b0 = max(b0, key[0][k]);
                        else if ((k & 3) == 1) b1 = max(b1, key[0][k]);
                        else if ((k & 3) == 2) b2 = max(b2, key[0][k]);
                        else b3 = max(b3, key[0][k]);
                    }
                    kbest = max(max(b0, b1), max(b2, b3));
                    if (nk[0] != 0u) wmin = __vminu2(wmin, nw);
                }
            }
        }
    }
    __syncthreads();
    if (CLUSTER) cluster_barrier();                                    // nobody leaves while a peer may still store into it
    // final state: usage follows scheduler.py:150-151 for every node that received a pod
    for (int local = tid; local < r.nodes_per_cta; local += kFastThreads) {
        const int n = cta_base + local;
        if (n >= in.n) break;
        const uint32_t ct = c_t[local];
        const bool touched = (ct >> 31) != 0;
        long long pc = in.pod_count0[n];
        if (touched) pc = (long long)((ct & 0x7fffffffu) - s_tbase[r.shape_id[n] & 0x7fu]);
        const double u = usage_a3(pc, in.max_pods[n]);
        if (r.out_pc) r.out_pc[n] = pc;
        if (r.out_cu) r.out_cu[n] = touched ? u : in.cu0[n];
        if (r.out_mu) r.out_mu[n] = touched ? u : in.mu0[n];
    }
}

// scores of the chosen states, in parallel over the pods
__global__ void __launch_bounds__(256)
fast_scores_kernel(FastIn in, const FastHeader *__restrict__ hdr, const uint32_t *__restrict__ out_entry,
                   const double *__restrict__ score, int total, double *__restrict__ out_score)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= in.n_pods || !hdr->ok) return;
    const uint32_t e = out_entry[j];
    const int s = in.sid ? canon_strategy(in.sid[j]) : in.uniform_strategy;
    out_score[j] = e == kNoEntry ? -1.0 : score[(size_t)s * total + e];
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct FastLayout {
    int64_t total;
    size_t off_hdr, off_shape, off_iota, off_keys, off_score, off_kout, off_vout, off_flag, off_incl, off_rank,
        off_uniq, off_w, off_key32, off_xs, off_entry, off_tmp, tmp_bytes, bytes;
    // exact batching (only for calls with at least kBatchMinPods pods): entry buffers for up to batch_cap entries
    size_t off_bhdr, off_bcnt, off_boff, off_bcommits, off_bkeys[2], off_bvals[2];
    int64_t batch_cap;
};

static FastLayout fast_layout(int64_t n, int64_t p)
{
    FastLayout L{};
    L.total = n + kFastMaxTraj;
    const size_t T = (size_t)L.total;
    size_t o = 0;
    L.off_hdr = o;   o = align_up(o + sizeof(FastHeader));
    L.off_shape = o; o = align_up(o + sizeof(uint32_t) * n);
    L.off_iota = o;  o = align_up(o + sizeof(uint32_t) * T);
    L.off_keys = o;  o = align_up(o + sizeof(uint64_t) * 5 * T);
    L.off_score = o; o = align_up(o + sizeof(double) * kMaxStrategies * T);
    L.off_kout = o;  o = align_up(o + sizeof(uint64_t) * 5 * T);      // sorted keys / source entries of the five key arrays
    L.off_vout = o;  o = align_up(o + sizeof(uint32_t) * 5 * T);
    L.off_flag = o;  o = align_up(o + sizeof(uint32_t) * T);
    L.off_incl = o;  o = align_up(o + sizeof(uint32_t) * T);
    L.off_rank = o;  o = align_up(o + sizeof(uint32_t) * 5 * T);
    L.off_uniq = o;  o = align_up(o + sizeof(double) * 2 * T);
    L.off_w = o;     o = align_up(o + sizeof(uint32_t) * T);
    L.off_key32 = o; o = align_up(o + sizeof(uint32_t) * kMaxStrategies * T);
    L.off_xs = o;    o = align_up(o + sizeof(uint32_t) * p);
    L.off_entry = o; o = align_up(o + sizeof(uint32_t) * p);
    L.batch_cap = p >= kBatchMinPods ? 2 * p : 0;                     // more reachable states than twice the pods: the loop does it
    L.off_bhdr = o;     o = align_up(o + sizeof(BatchHeader));
    L.off_bcnt = o;     o = align_up(o + sizeof(uint32_t) * (L.batch_cap ? n : 0));
    L.off_boff = o;     o = align_up(o + sizeof(uint32_t) * (L.batch_cap ? n : 0));
    L.off_bcommits = o; o = align_up(o + sizeof(uint32_t) * (L.batch_cap ? n : 0));
    for (int b = 0; b < 2; ++b) {
        L.off_bkeys[b] = o; o = align_up(o + sizeof(uint32_t) * L.batch_cap);
        L.off_bvals[b] = o; o = align_up(o + sizeof(uint32_t) * L.batch_cap);
    }
    L.tmp_bytes = std::max(sort_temp_bytes(L.total), L.batch_cap ? sort_temp_bytes(L.batch_cap) : (size_t)0);
    if (L.total <= kRankSortMax) L.tmp_bytes = std::max(L.tmp_bytes, rank_sort_scratch_bytes(L.total, 5));
    L.off_tmp = o;   o = align_up(o + L.tmp_bytes);
    L.bytes = o;
    return L;
}

constexpr int64_t kFastMinPods = 2048;         // below this the set-up costs more than it saves
constexpr int kFastMaxNptUniform = 80;         // node slots per thread: two registers per node ...
constexpr int kFastMaxNptMulti = 40;           // ... four with per-pod strategies

static int ceil_log2(int64_t v) { int b = 0; while ((1ll << b) < v) ++b; return b; }

struct FastShape { int npt, cluster; };

static size_t fast_smem_bytes(bool multi, int nodes_per_cta)
{
    return sizeof(uint32_t) * kFastMaxTraj * (multi ? 4 : 2) + 4 * (size_t)nodes_per_cta;
}

// Can the fast path be attempted for this call?  (Host-side, sizes only.)  One CTA while the table
// fits its registers (20 480 nodes, 10 240 with per-pod strategies), then clusters of 2, 4, 8 CTAs
// at the largest slot count; beyond that the fp64 kernel does the work.
static bool fast_eligible(int64_t n, int64_t p, int filter_mode, bool multi, FastShape *shape, int *idx_bits)
{
    if (filter_mode != KP_FILTER_CAPACITY || p < kFastMinPods || n < 1) return false;
    if (const char *off = getenv("KP_GREEDY_NO_FAST")) if (off[0] == '1') return false;
    const int ib = std::max(1, ceil_log2(n));
    if (ib + ceil_log2(n + kFastMaxTraj + 2) > 32) return false;      // key = score rank | inverted index
    *idx_bits = ib;
    const int max_npt = multi ? kFastMaxNptMulti : kFastMaxNptUniform;
    int c = 1;
    if (const char *force = getenv("KP_GREEDY_CLUSTER")) c = std::max(1, std::min(8, atoi(force)));   // experiments only
    for (;; c *= 2) {
        if (c > 8) return false;
        const int64_t per_cta = (n + c - 1) / c;
        const int need = (int)((per_cta + kFastThreads - 1) / kFastThreads);
        if (need > max_npt) continue;
        int npt = max_npt;                                             // clusters exist at the largest slot count only
        if (c == 1) for (int cand : {5, 10, 20, 40, 80}) if (need <= cand) { npt = cand; break; }
        if (fast_smem_bytes(multi, npt * kFastThreads) > 200 * 1024) return false;
        *shape = FastShape{npt, c};
        return true;
    }
}

size_t greedy_fast_workspace_bytes(int64_t n, int64_t p)
{
    return fast_layout(n, p).bytes;
}

template <int NPT, bool MULTI, int CSIZE>
static int launch_fast(const FastIn &in, const FastRun &r, size_t smem, cudaStream_t stream)
{
    static bool attr_set[kMaxDevices] = {false};                     // the opt-in is per device
    const int dev = device_slot();
    if (!attr_set[dev]) {
        KP_CUDA(cudaFuncSetAttribute(fast_greedy_kernel<NPT, MULTI, CSIZE>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(CSIZE, 1, 1);
    cfg.blockDim = dim3(kFastThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CSIZE;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    KP_CUDA(cudaLaunchKernelEx(&cfg, fast_greedy_kernel<NPT, MULTI, CSIZE>, in, r));
    count_launch();
    return KP_OK;
}

template <bool MULTI>
static int dispatch_fast(const FastShape &sh, const FastIn &in, const FastRun &r, size_t smem, cudaStream_t stream)
{
    constexpr int MAXN = MULTI ? kFastMaxNptMulti : kFastMaxNptUniform;
    switch (sh.cluster) {
        case 2: return launch_fast<MAXN, MULTI, 2>(in, r, smem, stream);
        case 4: return launch_fast<MAXN, MULTI, 4>(in, r, smem, stream);
        case 8: return launch_fast<MAXN, MULTI, 8>(in, r, smem, stream);
        default: break;
    }
    if (sh.npt <= 5) return launch_fast<5, MULTI, 1>(in, r, smem, stream);
    if (sh.npt <= 10) return launch_fast<10, MULTI, 1>(in, r, smem, stream);
    if (sh.npt <= 20) return launch_fast<20, MULTI, 1>(in, r, smem, stream);
    if (sh.npt <= 40) return launch_fast<40, MULTI, 1>(in, r, smem, stream);
    return launch_fast<MAXN, MULTI, 1>(in, r, smem, stream);
}

// Runs the fast path when the sizes allow it.  *done_flag (device) is 1 afterwards iff the
// results are complete; the caller launches the fp64 kernel with that flag as its skip switch.
int greedy_fast_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy, int filter_mode,
                       int64_t *out_idx, double *out_score, int64_t *out_pc, double *out_cu, double *out_mu,
                       void *ws, size_t ws_bytes, cudaStream_t stream, const uint32_t **done_flag)
{
    *done_flag = nullptr;
    const int64_t n = nodes.n_nodes, p = pods.n_pods;
    const bool multi = pods.strategy_ids != nullptr;
    FastShape shape{};
    int idx_bits = 1;
    if (!fast_eligible(n, p, filter_mode, multi, &shape, &idx_bits)) return KP_OK;
    const FastLayout L = fast_layout(n, p);
    if (!ws || ws_bytes < L.bytes) return KP_OK;                     // not enough scratch: fp64 kernel only
    unsigned char *w = static_cast<unsigned char *>(ws);
    FastHeader *hdr = reinterpret_cast<FastHeader *>(w + L.off_hdr);
    const int total = (int)L.total;

    FastIn in{};
    in.alloc_cpu = (const long long *)nodes.alloc_cpu_milli;
    in.alloc_mem = (const long long *)nodes.alloc_mem_bytes;
    in.pod_count0 = (const long long *)nodes.pod_count;
    in.max_pods = (const long long *)nodes.max_pods;
    in.ready = nodes.ready;
    in.cu0 = nodes.cpu_usage_pct;
    in.mu0 = nodes.mem_usage_pct;
    in.n = (int)n;
    in.req_cpu = (const long long *)pods.req_cpu_milli;
    in.req_mem = (const long long *)pods.req_mem_bytes;
    in.sid = (const long long *)pods.strategy_ids;
    in.n_pods = p;
    in.uniform_strategy = canon_strategy(strategy);
    in.strat_mask = multi ? 7 : (1 << in.uniform_strategy);

    uint32_t *shape_id = reinterpret_cast<uint32_t *>(w + L.off_shape);
    uint32_t *iota = reinterpret_cast<uint32_t *>(w + L.off_iota);
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(w + L.off_keys);
    double *score = reinterpret_cast<double *>(w + L.off_score);
    unsigned long long *kout = reinterpret_cast<unsigned long long *>(w + L.off_kout);
    uint32_t *vout = reinterpret_cast<uint32_t *>(w + L.off_vout);
    uint32_t *flag = reinterpret_cast<uint32_t *>(w + L.off_flag);
    uint32_t *incl = reinterpret_cast<uint32_t *>(w + L.off_incl);
    uint32_t *rank = reinterpret_cast<uint32_t *>(w + L.off_rank);
    double *uniq = reinterpret_cast<double *>(w + L.off_uniq);
    uint32_t *entry_w = reinterpret_cast<uint32_t *>(w + L.off_w);
    uint32_t *key32 = reinterpret_cast<uint32_t *>(w + L.off_key32);
    uint32_t *xs = reinterpret_cast<uint32_t *>(w + L.off_xs);
    uint32_t *out_entry = reinterpret_cast<uint32_t *>(w + L.off_entry);
    void *tmp = w + L.off_tmp;

    fast_shapes_kernel<<<1, 1024, 0, stream>>>(in, hdr, shape_id);
    KP_LAUNCH_CHECK("fast_shapes_kernel");
    FastEntries e{};
    e.key_t[0] = keys + (size_t)3 * total;
    e.key_t[1] = keys + (size_t)4 * total;
    for (int s = 0; s < kMaxStrategies; ++s) { e.key_s[s] = keys + (size_t)s * total; e.score[s] = score + (size_t)s * total; }
    const int gb = (total + 255) / 256;
    fast_entries_kernel<<<gb, 256, 0, stream>>>(in, hdr, e, iota);
    KP_LAUNCH_CHECK("fast_entries_kernel");
    // sorted order of the five key arrays (0..2 score orders, 3..4 thresholds): all of them in two launches of our own
    // for tables up to kRankSortMax entries (kp_ranksort.cu), one library radix sort per array beyond
    if (total <= kRankSortMax) {
        RankSortArrays which{};
        for (int a = 0; a < 5; ++a)
            if (a >= kMaxStrategies || (in.strat_mask & (1 << a))) which.idx[which.count++] = a;
        const int st = rank_sort_pairs_u64(tmp, L.tmp_bytes, keys, (size_t)total, which, total, kout, vout, (size_t)total, stream);
        if (st != KP_OK) return st;
    }
    for (int a = 0; a < 5; ++a) {
        if (a < kMaxStrategies && !(in.strat_mask & (1 << a))) continue;
        unsigned long long *ko = kout + (size_t)a * total;
        uint32_t *vo = vout + (size_t)a * total;
        int st = KP_OK;
        if (total > kRankSortMax) st = sort_pairs_u64(tmp, L.tmp_bytes, keys + (size_t)a * total, ko, iota, vo, total, stream);
        if (st != KP_OK) return st;
        fast_flags_kernel<<<gb, 256, 0, stream>>>(ko, total, flag);
        KP_LAUNCH_CHECK("fast_flags_kernel");
        st = inclusive_sum_u32(tmp, L.tmp_bytes, flag, incl, total, stream);
        if (st != KP_OK) return st;
        fast_rank_kernel<<<gb, 256, 0, stream>>>(ko, vo, flag, incl, total, rank + (size_t)a * total,
                                                 a >= kMaxStrategies ? uniq + (size_t)(a - kMaxStrategies) * total : nullptr,
                                                 a >= kMaxStrategies ? &hdr->distinct[a - kMaxStrategies] : &hdr->score_ranks[a]);
        KP_LAUNCH_CHECK("fast_rank_kernel");
    }
    fast_pack_kernel<<<gb, 256, 0, stream>>>(in, hdr, rank + (size_t)3 * total, rank + (size_t)4 * total, rank, keys, total,
                                             idx_bits, entry_w, key32);
    KP_LAUNCH_CHECK("fast_pack_kernel");
    fast_pods_kernel<<<(int)((p + 255) / 256), 256, 0, stream>>>(in, hdr, uniq, uniq + total, xs);
    KP_LAUNCH_CHECK("fast_pods_kernel");

    // exact batching of the pod-independent prefix (one strategy whose scores fall with every commit)
    BatchHeader *bh = nullptr;
    uint32_t *commits = nullptr;
    if (L.batch_cap > 0 && !multi && in.uniform_strategy != 2 && n < (1 << 24)) {
        bh = reinterpret_cast<BatchHeader *>(w + L.off_bhdr);
        uint32_t *cnt = reinterpret_cast<uint32_t *>(w + L.off_bcnt), *off = reinterpret_cast<uint32_t *>(w + L.off_boff);
        commits = reinterpret_cast<uint32_t *>(w + L.off_bcommits);
        uint32_t *bk0 = reinterpret_cast<uint32_t *>(w + L.off_bkeys[0]), *bk1 = reinterpret_cast<uint32_t *>(w + L.off_bkeys[1]);
        uint32_t *bv0 = reinterpret_cast<uint32_t *>(w + L.off_bvals[0]), *bv1 = reinterpret_cast<uint32_t *>(w + L.off_bvals[1]);
        KP_CUDA(cudaMemsetAsync(commits, 0, sizeof(uint32_t) * (size_t)n, stream));
        KP_CUDA(cudaMemsetAsync(bk0, 0, sizeof(uint32_t) * (size_t)L.batch_cap, stream));          // unused slots sort last
        batch_check_kernel<<<1, 1024, 0, stream>>>(in, hdr, shape_id, entry_w, key32, total, cnt, bh, (unsigned long long)L.batch_cap);
        KP_LAUNCH_CHECK("batch_check_kernel");
        batch_violation_kernel<<<(int)((p + 255) / 256), 256, 0, stream>>>(in, hdr, xs, bh);
        KP_LAUNCH_CHECK("batch_violation_kernel");
        int st = exclusive_sum_u32(tmp, L.tmp_bytes, cnt, off, (int)n, stream);
        if (st != KP_OK) return st;
        batch_entries_kernel<<<(int)((n * 32 + 255) / 256), 256, 0, stream>>>(in, hdr, bh, shape_id, key32, total, idx_bits, cnt, off, bk0, bv0);
        KP_LAUNCH_CHECK("batch_entries_kernel");
        st = sort_pairs_u32_desc(tmp, L.tmp_bytes, bk0, bk1, bv0, bv1, L.batch_cap, stream);
        if (st != KP_OK) return st;
        batch_assign_kernel<<<(int)((p + 255) / 256), 256, 0, stream>>>(in, hdr, bh, bk1, bv1, idx_bits, (long long *)out_idx, out_entry, commits);
        KP_LAUNCH_CHECK("batch_assign_kernel");
    }

    FastRun r{};
    r.hdr = hdr;
    r.batch = bh;
    r.commits = commits;
    r.shape_id = shape_id;
    r.entry_w = entry_w;
    r.entry_key = key32;
    r.xs = xs;
    r.out_entry = out_entry;
    r.out_idx = (long long *)out_idx;
    r.out_pc = (long long *)out_pc;
    r.out_cu = out_cu;
    r.out_mu = out_mu;
    r.total = total;
    r.idx_bits = idx_bits;
    r.nodes_per_cta = shape.npt * kFastThreads;
    const size_t smem = fast_smem_bytes(multi, r.nodes_per_cta);
    const int st = multi ? dispatch_fast<true>(shape, in, r, smem, stream)
                         : dispatch_fast<false>(shape, in, r, smem, stream);
    if (st != KP_OK) return st;
    fast_scores_kernel<<<(int)((p + 255) / 256), 256, 0, stream>>>(in, hdr, out_entry, score, total, out_score);
    KP_LAUNCH_CHECK("fast_scores_kernel");
    *done_flag = &hdr->ok;
    return KP_OK;
}

}  // namespace kp

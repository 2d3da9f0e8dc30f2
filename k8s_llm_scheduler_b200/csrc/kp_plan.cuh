// Node plan of snapshot placement (included by kp_snapshot.cu): everything that depends on
// the node alone, computed once per snapshot.
//
//   * per node: score under each strategy in play (exact fp64, reference op order,
//     scheduler.py:537-545), eligibility (Ready :531-535, score > -1 sentinel :529/:547, and in
//     capacity mode pod_count < max_pods), free-cpu / free-memory thresholds (:232-233).
//   * nodes ordered best-first under (score desc, index asc) -- the reference's "strict >,
//     first in list order wins" (:547) as a total order.
//   * DENSE RANK COMPRESSION of the two thresholds: RA[n] = #{distinct free_cpu values <
//     free_cpu[n]}; for a pod, ra(req) = #{distinct values < req} (binary search).  Then
//       req <= free_cpu[n]   <=>   ra(req) <= RA[n]
//     exactly, for any fp64 inputs, so the per-pair fp64 compares become small-integer
//     compares; with <= 32000 distinct values per dimension both ranks of a node are packed
//     into ONE 32-bit word with a guard bit above each 15-bit field.
//
// Two implementations:
//   small tables       (<= 8192 nodes) plan_count_kernel, one launch: no sort -- every node counts
//                      its own sorted position against the whole table, in parallel over the
//                      whole GPU; plain "number of smaller thresholds" ranks (no dense pass).
//   large tables       (any size) derive -> radix sorts (kp_sort.cu) -> post_sort -> scan of the
//                      run starts (dense ranks) -> tables.
#pragma once

#include <cooperative_groups.h>

#include "kp_common.cuh"

namespace kp {

namespace cg = cooperative_groups;

constexpr int kKeyArrays = kMaxStrategies + 2;   // 3 score orders + cpu threshold + mem threshold
constexpr int kSmallMaxNodes = 8192;             // plan_count_kernel limit (one key array in shared memory)
constexpr uint32_t kSwarMaxDistinct = 32000;     // fields hold rank+1 <= 32001 < 0x7fff

struct NodePtrs {
    const long long *alloc_cpu, *alloc_mem, *pod_count, *max_pods;
    const uint8_t *ready;
    const double *cu, *mu;
    int n;
};

// Everything the plan needs from one node.
//   skey[s]  = ~score_key (ascending sort = best first); ~0 = ineligible (sorts last)
//   tkey[0/1] = order keys of free_cpu / free_mem (ascending); NaN parked at +inf
struct Derived {
    unsigned long long skey[kMaxStrategies];
    double score[kMaxStrategies];
    unsigned long long tkey[2];
};

__device__ __forceinline__ Derived derive_node(const NodePtrs &t, int n, int strat_mask)
{
    Derived d;
    const double cu = t.cu[n], mu = t.mu[n];
    const long long pc = t.pod_count[n], mp = t.max_pods[n];
    const bool rdy = t.ready[n] != 0;
    double ta = free_cpu_cores(t.alloc_cpu[n], cu);
    double tb = free_mem_gib(t.alloc_mem[n], mu);
    const bool nan = (ta != ta) || (tb != tb);
    const bool cap_ok = !nan && pc < mp;
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    if (ta != ta) ta = inf;          // NaN never fits: park it at +inf, the node is ineligible anyway
    if (tb != tb) tb = inf;
    d.tkey[0] = f64_order_key(ta);
    d.tkey[1] = f64_order_key(tb);
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        d.skey[s] = ~0ull;
        d.score[s] = 0.0;
        if (!(strat_mask & (1 << s))) continue;
        const NodeScore ns = node_score(s, cu, mu, pc, mp);
        d.score[s] = ns.score;
        if (rdy && cap_ok && ns.key != 0ull) d.skey[s] = ~ns.key;
    }
    return d;
}

__device__ __forceinline__ double key_to_f64(unsigned long long k)
{
    const unsigned long long b = (k >> 63) ? (k ^ 0x8000000000000000ull) : ~k;
    return __longlong_as_double((long long)b);
}

__device__ __forceinline__ uint32_t pack_ranks(uint32_t ra, uint32_t rb)
{
    return (((rb + 1u) | 0x8000u) << 16) | ((ra + 1u) | 0x8000u);
}

struct PlanOut {
    SnapHeader *hdr;
    double *score;      // [3][n]
    double *uniq;       // [2][npad]
    uint32_t *order;    // [3][npad]
    double *score_pos;  // [3][npad] score of the node at each rank position
    uint2 *ab;          // [3][npad]
    uint32_t *packed;   // [3][npad]
    int npad;
};

// ---------------------------------------------------------------------------------------
// small-table plan (n <= kSmallMaxNodes): plan_count_kernel, ONE launch
// ---------------------------------------------------------------------------------------
// No sort at all: every node finds its own place by counting, fully in parallel over the
// whole GPU.  For a key array K (ascending order wanted, ties broken by node index):
//     pos[n]  = #{m : K[m] < K[n]  or  (K[m] == K[n] and m < n)}      stable sorted position
//     less[n] = #{m : K[m] < K[n]}                                     threshold rank
// Below 32 k nodes the plain count `less` already fits the packed 15-bit rank fields, so no
// dense-rank pass is needed: req <= thr[n]  <=>  #{m : thr[m] < req} <= less[n], with the pod
// side found by binary search in the sorted threshold array (duplicates kept).
//
// A thread-block CLUSTER of 8 CTAs owns 32 nodes (one per lane).  Every CTA first derives the
// keys it needs itself (its eighth of the table and the 32 nodes) into shared memory, so there
// is no separate derive pass and no key array in global memory.  The 64 warps then each count
// over one 64th of the table (every lane reads the same key: broadcast loads) and add their
// partial counts into the leader CTA's shared memory with distributed-shared-memory atomics;
// after a cluster barrier the leader writes the nodes' entries of the sorted threshold arrays
// and of every best-first table.  Cluster 0 also counts the eligible nodes per strategy and
// writes the padding behind them.
constexpr int kCountThreads = 256;
constexpr int kCountNodes = 32;
constexpr int kCountCluster = 8;
constexpr int kCountChunks = kCountCluster * (kCountThreads / 32);

constexpr int kCountMaxChunk = kSmallMaxNodes / kCountCluster + 2;   // table rows one CTA counts over

__device__ __forceinline__ void dsmem_add(cg::cluster_group &cluster, uint32_t *local, uint32_t v)
{
    if (v) atomicAdd(cluster.map_shared_rank(local, 0), v);
}

__global__ void __cluster_dims__(kCountCluster, 1, 1) __launch_bounds__(kCountThreads)
plan_count_kernel(NodePtrs t, int strat_mask, PlanOut o)
{
    __shared__ unsigned long long s_mkeys[kKeyArrays][kCountMaxChunk];   // keys of this CTA's table rows
    __shared__ unsigned long long s_nkeys[kKeyArrays][kCountNodes];      // keys of the cluster's 32 nodes
    __shared__ uint32_t acc_pos[kKeyArrays][kCountNodes];               // accumulated in the leader CTA only
    __shared__ uint32_t acc_less[2][kCountNodes];
    __shared__ uint32_t acc_live[kMaxStrategies];
    __shared__ double s_nscore[kMaxStrategies][kCountNodes];

    cg::cluster_group cluster = cg::this_cluster();
    const int crank = (int)cluster.block_rank();
    const int group = blockIdx.x / kCountCluster;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = kCountThreads / 32;
    const int n = t.n;
    const int g0 = group * kCountNodes;
    const int node = g0 + lane;
    const int chunk = crank * NW + warp;
    const int m0 = (int)(((long long)n * chunk) / kCountChunks), m1 = (int)(((long long)n * (chunk + 1)) / kCountChunks);
    const int cta_m0 = (int)(((long long)n * (crank * NW)) / kCountChunks);
    const int cta_m1 = (int)(((long long)n * (crank * NW + NW)) / kCountChunks);
    const int mcount = cta_m1 - cta_m0;

    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // place_kernel may start its prologue
    for (int i = tid; i < kKeyArrays * kCountNodes; i += kCountThreads) (&acc_pos[0][0])[i] = 0;
    if (tid < 2 * kCountNodes) (&acc_less[0][0])[tid] = 0;
    if (tid < kMaxStrategies) acc_live[tid] = 0;
    if (group == 0 && crank == 0 && tid < kMaxStrategies) {           // header of the snapshot
        o.hdr->group_count[tid] = 0;
        o.hdr->group_cursor[tid] = 0;
        if (!(strat_mask & (1 << tid))) o.hdr->nvalid[tid] = 0;
        if (tid == 0) {
            o.hdr->ticket = 0;
            o.hdr->use_swar = 1;
            o.hdr->reserved = 0;
            o.hdr->distinct[0] = (uint32_t)n;                           // sorted threshold arrays keep duplicates
            o.hdr->distinct[1] = (uint32_t)n;
        }
    }
    // derive exactly the keys this CTA needs: its table rows and the cluster's 32 nodes
    for (int i = tid; i < mcount + kCountNodes; i += kCountThreads) {
        const bool is_row = i < mcount;
        const int id = is_row ? cta_m0 + i : g0 + (i - mcount);
        Derived d;
        d.tkey[0] = d.tkey[1] = 0ull;
        d.skey[0] = d.skey[1] = d.skey[2] = ~0ull;
        if (id < n) {
            d = derive_node(t, id, strat_mask);
            if (!is_row && crank == 0) {
#pragma unroll
                for (int s = 0; s < kMaxStrategies; ++s)
                    if (strat_mask & (1 << s)) { o.score[(size_t)s * n + id] = d.score[s]; s_nscore[s][i - mcount] = d.score[s]; }
            }
        }
#pragma unroll
        for (int j = 0; j < kKeyArrays; ++j) {
            const unsigned long long k = j < kMaxStrategies ? d.skey[j] : d.tkey[j - kMaxStrategies];
            if (is_row) s_mkeys[j][i] = k; else s_nkeys[j][i - mcount] = k;
        }
    }
    __syncthreads();
    cluster_barrier();                                         // every CTA's accumulators are initialised

    for (int j = kKeyArrays - 1; j >= 0; --j) {                // thresholds (3, 4), then strategies
        if (j < kMaxStrategies && !(strat_mask & (1 << j))) continue;
        const unsigned long long *K = s_mkeys[j] - cta_m0;     // indexable by the global row number
        const unsigned long long kn = s_nkeys[j][lane];
        uint32_t c_pos = 0, c_less = 0;
        if (m1 <= g0) {                                        // every m is below the node group: ties count
#pragma unroll 8
            for (int m = m0; m < m1; ++m) {
                const unsigned long long km = K[m];            // same address in every lane
                c_pos += km <= kn;
                if (j >= kMaxStrategies) c_less += km < kn;
            }
        } else if (m0 >= g0 + kCountNodes) {                   // every m is above: ties do not count
#pragma unroll 8
            for (int m = m0; m < m1; ++m) c_pos += K[m] < kn;
            c_less = c_pos;
        } else {
            for (int m = m0; m < m1; ++m) {
                const unsigned long long km = K[m];
                const uint32_t lt = km < kn;
                c_less += lt;
                c_pos += lt | (uint32_t)((km == kn) & (m < node));
            }
        }
        if (node < n) {
            dsmem_add(cluster, &acc_pos[j][lane], c_pos);
            if (j >= kMaxStrategies) dsmem_add(cluster, &acc_less[j - kMaxStrategies][lane], c_less);
        }
        if (j < kMaxStrategies && group == 0) {                // eligible nodes of this strategy table
            uint32_t live = 0;
            for (int m = m0 + lane; m < m1; m += 32) live += K[m] != ~0ull;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) live += __shfl_xor_sync(0xffffffffu, live, d);
            if (lane == 0) dsmem_add(cluster, &acc_live[j], live);
        }
    }
    cluster_barrier();                                         // all partial counts have landed in the leader
    if (crank != 0) return;
    if (group == 0) {
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            if (!(strat_mask & (1 << s))) continue;
            const int nv = (int)acc_live[s];
            if (tid == 0) o.hdr->nvalid[s] = (uint32_t)nv;
            for (int i = nv + tid; i < o.npad; i += kCountThreads) {   // padding behind the eligible nodes
                o.ab[(size_t)s * o.npad + i] = make_uint2(0xffffffffu, 0xffffffffu);
                o.packed[(size_t)s * o.npad + i] = kSwarGuards;
            }
        }
    }
    if (warp == 0 && node < n) {
        const uint2 r = make_uint2(acc_less[0][lane], acc_less[1][lane]);
#pragma unroll
        for (int jt = 0; jt < 2; ++jt)
            o.uniq[(size_t)jt * o.npad + acc_pos[kMaxStrategies + jt][lane]] = key_to_f64(s_nkeys[kMaxStrategies + jt][lane]);
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            if (!(strat_mask & (1 << s))) continue;
            if (s_nkeys[s][lane] == ~0ull) continue;                      // ineligible: not in the table
            const uint32_t p = acc_pos[s][lane];
            o.order[(size_t)s * o.npad + p] = (uint32_t)node;
            o.score_pos[(size_t)s * o.npad + p] = s_nscore[s][lane];
            o.ab[(size_t)s * o.npad + p] = r;
            o.packed[(size_t)s * o.npad + p] = pack_ranks(r.x, r.y);
        }
    }
}

// ---------------------------------------------------------------------------------------
// large-table plan (any n)
// ---------------------------------------------------------------------------------------
// COUNT: accumulate the eligible-node counts with atomics (header zeroed beforehand); the
// small-table path instead initialises the header here and counts in plan_count_kernel.
template <bool COUNT>
__global__ void __launch_bounds__(256)
derive_kernel(NodePtrs t, int strat_mask, unsigned long long *__restrict__ keys,
              double *__restrict__ score, SnapHeader *__restrict__ hdr, uint32_t *__restrict__ iota)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = n < t.n;
    Derived d;
    if (live) {
        if (iota) iota[n] = (uint32_t)n;
        d = derive_node(t, n, strat_mask);
        keys[(size_t)3 * t.n + n] = d.tkey[0];
        keys[(size_t)4 * t.n + n] = d.tkey[1];
    }
    if (!COUNT && n < kMaxStrategies) {
        hdr->group_count[n] = 0;
        hdr->group_cursor[n] = 0;
        if (!(strat_mask & (1 << n))) hdr->nvalid[n] = 0;
        if (n == 0) {
            hdr->ticket = 0;
            hdr->use_swar = 1;
            hdr->reserved = 0;
            hdr->distinct[0] = (uint32_t)t.n;       // sorted threshold arrays keep duplicates
            hdr->distinct[1] = (uint32_t)t.n;
        }
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(strat_mask & (1 << s))) continue;
        bool elig = false;
        if (live) {
            elig = d.skey[s] != ~0ull;
            score[(size_t)s * t.n + n] = d.score[s];
            keys[(size_t)s * t.n + n] = d.skey[s];
        }
        if (COUNT) {
            const unsigned ballot = __ballot_sync(0xffffffffu, elig);
            if ((threadIdx.x & 31) == 0 && ballot) atomicAdd(&hdr->nvalid[s], __popc(ballot));
        }
    }
}

// After the radix sorts (kp_sort.cu): kout[j] / vout[j] = keys of array j in ascending order and
// the node index of every sorted position.  One pass turns that into what the tables need:
//   pos[j][node] = sorted position of the node in array j,
//   sorted[t][i] = threshold value at sorted position i,  flag[t][i] = 1 iff it starts a new run
// (the inclusive scan of flag gives dense rank + 1).
struct SortList { int idx[kKeyArrays]; int count; };

__global__ void __launch_bounds__(256)
post_sort_kernel(const unsigned long long *__restrict__ kout, const uint32_t *__restrict__ vout,
                 int n_nodes, int npad, SortList list, uint32_t *__restrict__ pos,
                 double *__restrict__ sorted, uint32_t *__restrict__ flag)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const int j = list.idx[blockIdx.y];
    const unsigned long long k = kout[(size_t)j * n_nodes + i];
    pos[(size_t)j * n_nodes + vout[(size_t)j * n_nodes + i]] = (uint32_t)i;
    if (j >= kMaxStrategies) {
        const int t = j - kMaxStrategies;
        sorted[(size_t)t * npad + i] = key_to_f64(k);
        flag[(size_t)t * npad + i] = (i == 0 || kout[(size_t)j * n_nodes + i - 1] != k) ? 1u : 0u;
    }
}

// incl[t][i] = inclusive scan of flag[t]: dense rank of sorted position i is incl - 1 and the
// number of distinct thresholds is incl[t][n - 1].
__global__ void __launch_bounds__(256)
scatter_tables_kernel(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ pos,
                      const double *__restrict__ sorted, const uint32_t *__restrict__ flag,
                      const uint32_t *__restrict__ incl, int n_nodes, int strat_mask, PlanOut o)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= o.npad) return;
    const uint32_t d0 = incl[n_nodes - 1], d1 = incl[(size_t)o.npad + n_nodes - 1];
    const bool swar = d0 <= kSwarMaxDistinct && d1 <= kSwarMaxDistinct;
    if (n == 0) {
        o.hdr->distinct[0] = d0;
        o.hdr->distinct[1] = d1;
        o.hdr->use_swar = swar ? 1u : 0u;
    }
    uint2 r = make_uint2(0u, 0u);
    if (n < n_nodes) {
#pragma unroll
        for (int t = 0; t < 2; ++t)                                // unique threshold values, ascending
            if (flag[(size_t)t * o.npad + n]) o.uniq[(size_t)t * o.npad + incl[(size_t)t * o.npad + n] - 1] = sorted[(size_t)t * o.npad + n];
        r.x = incl[pos[(size_t)3 * n_nodes + n]] - 1;
        r.y = incl[(size_t)o.npad + pos[(size_t)4 * n_nodes + n]] - 1;
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(strat_mask & (1 << s))) continue;
        if (n >= (int)o.hdr->nvalid[s]) {                     // padding behind the eligible nodes
            o.ab[(size_t)s * o.npad + n] = make_uint2(0xffffffffu, 0xffffffffu);
            o.packed[(size_t)s * o.npad + n] = kSwarGuards;
        }
        if (n >= n_nodes || keys[(size_t)s * n_nodes + n] == ~0ull) continue;   // ineligible: not in the table
        const uint32_t p = pos[(size_t)s * n_nodes + n];
        o.order[(size_t)s * o.npad + p] = (uint32_t)n;
        o.score_pos[(size_t)s * o.npad + p] = o.score[(size_t)s * n_nodes + n];
        o.ab[(size_t)s * o.npad + p] = r;
        o.packed[(size_t)s * o.npad + p] = swar ? pack_ranks(r.x, r.y) : kSwarGuards;
    }
}

}  // namespace kp

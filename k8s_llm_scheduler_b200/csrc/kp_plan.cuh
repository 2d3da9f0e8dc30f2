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
//   plan_small_kernel  (<= 4096 nodes) ONE launch of a thread-block CLUSTER with one CTA per
//                      independent sort (free-cpu ranks, free-mem ranks, one per strategy);
//                      bitonic sorts held in registers (warp shuffles below distance 32,
//                      shared memory above); rank arrays cross CTAs through global memory
//                      over a cluster barrier.
//   multi-kernel path  (any size) derive -> counting rank -> sorted thresholds -> dense ranks
//                      -> tables.
#pragma once

#include <cooperative_groups.h>

#include "kp_common.cuh"

namespace kp {

namespace cg = cooperative_groups;

constexpr int kKeyArrays = kMaxStrategies + 2;   // 3 score orders + cpu threshold + mem threshold
constexpr int kSmallMaxNodes = 4096;             // plan_small_kernel limit
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
    uint2 *ab;          // [3][npad]
    uint32_t *packed;   // [3][npad]
    uint32_t *rank;     // [2][npad] dense threshold rank per node (small plan: exchanged between CTAs)
    int npad;
};

// block-wide scan helper for 1024 threads: returns the exclusive prefix of `cnt`
// and leaves the block total in *total
__device__ __forceinline__ uint32_t block_exclusive_1024(uint32_t cnt, uint32_t *warp_sums, uint32_t *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    __syncthreads();                       // warp_sums may still be read from a previous call
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warp_sums[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w += v;
        }
        warp_sums[lane] = w;
        if (lane == 31) *total = w;
    }
    __syncthreads();
    return incl - cnt + (warp > 0 ? warp_sums[warp - 1] : 0u);
}

// ---------------------------------------------------------------------------------------
// plan_small_kernel
// ---------------------------------------------------------------------------------------
// Bitonic sort of m = 1024 * E (key, idx) pairs, E in {1, 2, 4}.  Thread t holds elements
// e * 1024 + t in registers.  Compare-exchange partners at distance j are: another element of
// the same thread (j >= 1024), the same element slot of a lane in the same warp (j < 32:
// __shfl_xor), or of another warp (32 <= j < 1024: through shared memory, double-buffered so
// that one barrier per stage suffices).
struct SmallSmem {
    unsigned long long xkey[2][kSmallMaxNodes];   // exchange buffers for the cross-warp stages
    uint16_t xidx[2][kSmallMaxNodes];
    unsigned long long key[kSmallMaxNodes];       // sorted result
    uint16_t idx[kSmallMaxNodes];
    uint32_t warp_sums[32];
    uint32_t total;
};

// compare-exchange between two elements held by the same thread (distance JE * 1024)
template <int E, int JE>
__device__ __forceinline__ void in_thread_stage(unsigned long long (&key)[E], uint32_t (&idx)[E], int k)
{
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const int f = e ^ JE;
        if (f > e && f < E) {
            const int i = e * 1024 + (int)threadIdx.x;
            const bool asc = (i & k) == 0;
            const bool gt = key[e] > key[f] || (key[e] == key[f] && idx[e] > idx[f]);
            if (gt == asc) {
                const unsigned long long tk = key[e]; key[e] = key[f]; key[f] = tk;
                const uint32_t ti = idx[e]; idx[e] = idx[f]; idx[f] = ti;
            }
        }
    }
}

template <int E>
__device__ __forceinline__ void bitonic_sort_regs(unsigned long long (&key)[E], uint32_t (&idx)[E], SmallSmem &sm)
{
    const int tid = threadIdx.x;
    int buf = 0;
    constexpr int m = 1024 * E;
    for (int k = 2; k <= m; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= 1024) {
                if (E >= 4 && j == 2048) in_thread_stage<E, (E >= 4 ? 2 : 1)>(key, idx, k);
                else in_thread_stage<E, 1>(key, idx, k);
                continue;
            }
            if (j >= 32) {
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    sm.xkey[buf][e * 1024 + tid] = key[e];
                    sm.xidx[buf][e * 1024 + tid] = (uint16_t)idx[e];
                }
                __syncthreads();
            }
#pragma unroll
            for (int e = 0; e < E; ++e) {
                const int i = e * 1024 + tid;
                unsigned long long ok;
                uint32_t oi;
                if (j >= 32) {
                    ok = sm.xkey[buf][i ^ j];
                    oi = sm.xidx[buf][i ^ j];
                } else {
                    ok = __shfl_xor_sync(0xffffffffu, key[e], j);
                    oi = __shfl_xor_sync(0xffffffffu, idx[e], j);
                }
                const bool lower = (i & j) == 0;             // lower partner keeps the smaller (if ascending)
                const bool asc = (i & k) == 0;
                const bool mine_gt = key[e] > ok || (key[e] == ok && idx[e] > oi);
                if (mine_gt == (lower == asc)) { key[e] = ok; idx[e] = oi; }   // keep min iff lower == asc
            }
            if (j >= 32) buf ^= 1;
        }
    }
    __syncthreads();
#pragma unroll
    for (int e = 0; e < E; ++e) {
        sm.key[e * 1024 + tid] = key[e];
        sm.idx[e * 1024 + tid] = (uint16_t)idx[e];
    }
    __syncthreads();
}

// One cluster, one CTA per sort.  CTA 0: free-cpu thresholds, CTA 1: free-memory thresholds,
// CTA 2 + i: best-first order of the i-th strategy in strat_mask.
template <int E>
__global__ void __launch_bounds__(1024, 1)
plan_small_kernel(NodePtrs t, int strat_mask, PlanOut o)
{
    extern __shared__ __align__(128) unsigned char plan_smem_raw[];
    SmallSmem &sm = *reinterpret_cast<SmallSmem *>(plan_smem_raw);
    cg::cluster_group cluster = cg::this_cluster();
    const int role = (int)cluster.block_rank();
    const int tid = threadIdx.x;
    const int n = t.n;

    int my_s = -1;                                   // strategy sorted by this CTA (role >= 2)
    if (role >= 2) {
        int seen = 0;
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s)
            if (strat_mask & (1 << s)) { if (seen == role - 2) my_s = s; ++seen; }
    }

    // derive the one key array this CTA sorts; element e * 1024 + tid lives in this thread
    unsigned long long key[E];
    uint32_t idx[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const int i = e * 1024 + tid;
        key[e] = ~0ull;
        idx[e] = (uint32_t)i;
        if (i < n) {
            const Derived d = derive_node(t, i, role >= 2 ? (1 << my_s) : 0);
            if (role < 2) key[e] = d.tkey[role];
            else {
                key[e] = d.skey[my_s];
                o.score[(size_t)my_s * n + i] = d.score[my_s];
            }
        }
    }
    if (role == 0) {
        if (tid < kMaxStrategies) {
            o.hdr->group_count[tid] = 0;
            o.hdr->group_cursor[tid] = 0;
            if (!(strat_mask & (1 << tid))) o.hdr->nvalid[tid] = 0;
        }
        if (tid == 0) { o.hdr->ticket = 0; o.hdr->use_swar = 1; o.hdr->reserved = 0; }
    }
    bitonic_sort_regs<E>(key, idx, sm);

    if (role < 2) {
        // dense ranks + unique values; each thread owns E contiguous sorted positions
        const int i0 = tid * E;
        uint32_t cnt = 0;
#pragma unroll
        for (int q = 0; q < E; ++q) {
            const int i = i0 + q;
            cnt += (i < n && (i == 0 || sm.key[i] != sm.key[i - 1])) ? 1u : 0u;
        }
        uint32_t run = block_exclusive_1024(cnt, sm.warp_sums, &sm.total);   // distinct values before i0
#pragma unroll
        for (int q = 0; q < E; ++q) {
            const int i = i0 + q;
            if (i < n) {
                const bool first = i == 0 || sm.key[i] != sm.key[i - 1];
                if (first) { o.uniq[(size_t)role * o.npad + run] = key_to_f64(sm.key[i]); ++run; }
                o.rank[(size_t)role * o.npad + sm.idx[i]] = run - 1;
            }
        }
        if (tid == 0) o.hdr->distinct[role] = sm.total;
        __threadfence();
    }
    cluster.sync();                                  // rank arrays of CTAs 0 and 1 are now visible
    if (role >= 2) {
        const volatile uint32_t *rank_a = o.rank;
        const volatile uint32_t *rank_b = o.rank + o.npad;
        uint32_t mine = 0;
        for (int i = tid; i < o.npad; i += 1024) {
            const bool live = i < n && sm.key[i] != ~0ull;
            uint2 r = make_uint2(0xffffffffu, 0xffffffffu);
            uint32_t pk = kSwarGuards;                 // pad: fields 0 < every pod field
            if (live) {
                const uint32_t node = sm.idx[i];
                r = make_uint2(rank_a[node], rank_b[node]);
                pk = pack_ranks(r.x, r.y);
                o.order[(size_t)my_s * o.npad + i] = node;
                ++mine;
            }
            o.ab[(size_t)my_s * o.npad + i] = r;
            o.packed[(size_t)my_s * o.npad + i] = pk;
        }
        (void)block_exclusive_1024(mine, sm.warp_sums, &sm.total);
        if (tid == 0) o.hdr->nvalid[my_s] = sm.total;
    }
}

// ---------------------------------------------------------------------------------------
// multi-kernel plan (any n)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
derive_kernel(NodePtrs t, int strat_mask, unsigned long long *__restrict__ keys,
              double *__restrict__ score, SnapHeader *__restrict__ hdr)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = n < t.n;
    Derived d;
    if (live) {
        d = derive_node(t, n, strat_mask);
        keys[(size_t)3 * t.n + n] = d.tkey[0];
        keys[(size_t)4 * t.n + n] = d.tkey[1];
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
        const unsigned ballot = __ballot_sync(0xffffffffu, elig);
        if ((threadIdx.x & 31) == 0 && ballot) atomicAdd(&hdr->nvalid[s], __popc(ballot));
    }
}

// Stable ascending rank by counting, tiled through shared memory.
//   pos[j][n]  = #{m : K[m] < K[n]  or  (K[m] == K[n] and m < n)}
// grid = (ceil(n/256), chunks, active arrays); partial counts are added atomically.
struct RankList { int idx[kKeyArrays]; int count; };
constexpr int kRankTile = 1024;

__global__ void __launch_bounds__(256)
rank_kernel(const unsigned long long *__restrict__ keys, uint32_t *__restrict__ pos, int n_nodes,
            int chunk_len, RankList list)
{
    __shared__ unsigned long long tile[kRankTile];
    const int j = list.idx[blockIdx.z];
    const unsigned long long *K = keys + (size_t)j * n_nodes;
    const int n = blockIdx.x * 256 + threadIdx.x;
    const unsigned long long kn = n < n_nodes ? K[n] : 0ull;
    const int m0 = blockIdx.y * chunk_len;
    const int m1 = min(n_nodes, m0 + chunk_len);
    uint32_t c_pos = 0;
    for (int base = m0; base < m1; base += kRankTile) {
        for (int i = threadIdx.x; i < kRankTile; i += 256) {
            const int m = base + i;
            tile[i] = m < m1 ? K[m] : ~0ull;
        }
        __syncthreads();
        const int cnt = min(kRankTile, m1 - base);
        if (base + cnt <= n) {                 // every m in the tile is < n: ties count
#pragma unroll 8
            for (int i = 0; i < cnt; ++i) c_pos += tile[i] <= kn;
        } else if (base > n) {                 // every m is > n: ties do not count
#pragma unroll 8
            for (int i = 0; i < cnt; ++i) c_pos += tile[i] < kn;
        } else {
            for (int i = 0; i < cnt; ++i) {
                const unsigned long long km = tile[i];
                c_pos += (km < kn) | ((km == kn) & (base + i < n));
            }
        }
        __syncthreads();
    }
    if (n < n_nodes) atomicAdd(&pos[(size_t)j * n_nodes + n], c_pos);
}

__global__ void __launch_bounds__(256)
scatter_sorted_kernel(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ pos,
                      int n_nodes, int npad, double *__restrict__ sorted)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_nodes) return;
#pragma unroll
    for (int j = 0; j < 2; ++j)
        sorted[(size_t)j * npad + pos[(size_t)(3 + j) * n_nodes + n]] = key_to_f64(keys[(size_t)(3 + j) * n_nodes + n]);
}

// one CTA per threshold array: dense rank of every sorted position + unique values
__global__ void __launch_bounds__(1024, 1)
dense_kernel(const double *__restrict__ sorted, int n_nodes, int npad, uint32_t *__restrict__ dense,
             double *__restrict__ uniq, SnapHeader *__restrict__ hdr)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t total;
    const int tid = threadIdx.x;
    const int j = blockIdx.x;
    const int per = (n_nodes + 1023) / 1024;
    const int i0 = min(n_nodes, tid * per), i1 = min(n_nodes, i0 + per);
    const double *S = sorted + (size_t)j * npad;
    uint32_t cnt = 0;
    for (int i = i0; i < i1; ++i) cnt += (i == 0 || S[i] != S[i - 1]) ? 1u : 0u;
    uint32_t run = block_exclusive_1024(cnt, warp_sums, &total);
    for (int i = i0; i < i1; ++i) {
        if (i == 0 || S[i] != S[i - 1]) { uniq[(size_t)j * npad + run] = S[i]; ++run; }
        dense[(size_t)j * npad + i] = run - 1;
    }
    if (tid == 0) hdr->distinct[j] = total;
}

__global__ void __launch_bounds__(256)
scatter_tables_kernel(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ pos,
                      const uint32_t *__restrict__ dense, int n_nodes, int strat_mask, PlanOut o)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= o.npad) return;
    const bool swar = o.hdr->distinct[0] <= kSwarMaxDistinct && o.hdr->distinct[1] <= kSwarMaxDistinct;
    if (n == 0) o.hdr->use_swar = swar ? 1u : 0u;
    uint2 r = make_uint2(0u, 0u);
    if (n < n_nodes) {
        r.x = dense[pos[(size_t)3 * n_nodes + n]];
        r.y = dense[(size_t)o.npad + pos[(size_t)4 * n_nodes + n]];
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
        o.ab[(size_t)s * o.npad + p] = r;
        o.packed[(size_t)s * o.npad + p] = swar ? pack_ranks(r.x, r.y) : kSwarGuards;
    }
}

}  // namespace kp

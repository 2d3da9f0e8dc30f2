// Snapshot placement: P independent fallback decisions against one node table.
//
// Replaces one `_fallback_decision` call per pod (scheduler.py:521-559).  The reference
// recomputes every node's score for every pod; here everything that depends on the node
// alone is hoisted into a per-snapshot "node plan", and the P x N part that remains is a
// pure compare/select loop that never touches HBM.
//
// Node plan (per snapshot):
//   * per node: score under each strategy in play (exact fp64, reference op order),
//     eligibility (Ready, score > -1 sentinel, and in capacity mode pod_count < max_pods),
//     free-cpu / free-memory thresholds (scheduler.py:232-233).
//   * nodes ordered best-first under (score desc, index asc) -- the reference's "strict >,
//     first in list order wins" (scheduler.py:547) as a total order.
//   * DENSE RANK COMPRESSION of the two thresholds: RA[n] = #{distinct free_cpu values <
//     free_cpu[n]}; for a pod, ra(req) = #{distinct values < req} (binary search).  Then
//       req <= free_cpu[n]   <=>   ra(req) <= RA[n]
//     exactly, for any fp64 inputs, so the per-pair fp64 compares become small-integer
//     compares.  When both dimensions have <= 32000 distinct values (always true below 32 k
//     nodes, and in any cluster whose nodes come in a few shapes) the two ranks of a node are
//     packed into ONE 32-bit word with a guard bit above each 15-bit field.
//   Two implementations: plan_small_kernel (<= 4096 nodes: one CTA, bitonic sorts in shared
//   memory, one launch) and the multi-kernel path (counting rank, any size).
//
// place_kernel (the hot loop): one thread owns R pods in registers; tiles of the best-first
// table are staged in shared memory with 1-D bulk async copies (TMA engine, UBLKCP) on a
// 2-deep mbarrier ring; every lane reads the same node (broadcast LDS.128 = 4 packed nodes).
// Per pod x node pair, packed form:  t = node_word - pod_word  (both guard bits survive iff
// neither field borrowed, i.e. the pod fits in both dimensions);  p = (~t & 0x80008000) != 0;
// if !p remember the node.  That is 3 instructions: the subtract issues on the FMA pipe
// (IMAD.IADD), the test on the ALU pipe (LOP3 with predicate output), and the update
// alternates per pod between SEL (ALU) and a guarded IMAD (FMA) so that both 16-lane pipes
// carry 1.5 ops per pair (tools/issue_bench.cu has the measurements behind this).  Scanning
// worst -> best and overwriting on "fits" leaves the best-ranked fitting node, which IS the
// argmax with the reference's tie-break: no per-pair index bookkeeping, no P x N matrix.
//
// KP_FILTER_READY (the reference's own filter) is pod-independent: every pod of a strategy
// gets the same node, so that mode is an argmax over nodes plus a fill.
#include <algorithm>
#include <cstdarg>
#include <cstdio>

#include "kp_common.cuh"

namespace kp {

// ---------------------------------------------------------------------------------------
// workspace layout
// ---------------------------------------------------------------------------------------
struct SnapHeader {
    uint32_t nvalid[kMaxStrategies];        // eligible nodes per strategy
    uint32_t group_count[kMaxStrategies];   // pods per strategy (per-pod strategy mode)
    uint32_t group_cursor[kMaxStrategies];
    uint32_t ticket;
    uint32_t distinct[2];                   // distinct free-cpu / free-mem thresholds
    uint32_t use_swar;                      // 1: packed 15-bit ranks are valid
    uint32_t reserved;
    long long winner[kMaxStrategies];       // ready mode: argmax node per strategy
    double winner_score[kMaxStrategies];
};

constexpr int kKeyArrays = kMaxStrategies + 2;   // 3 score orders + cpu threshold + mem threshold
constexpr int kArgmaxMaxBlocks = 256;
constexpr int kSmallMaxNodes = 4096;             // plan_small_kernel limit
constexpr uint32_t kSwarMaxDistinct = 32000;     // fields hold rank+1 <= 32001 < 0x7fff
constexpr uint32_t kSwarGuards = 0x80008000u;
constexpr uint32_t kSwarDeadPod = 0x7fff7fffu;   // larger than any node field: never fits
constexpr int kTileBytes = 32768;                // per shared-memory tile buffer
constexpr int kTileNodesPacked = kTileBytes / 4;
constexpr int kTileNodesWide = kTileBytes / 8;

struct SnapLayout {
    int64_t n, p;
    int64_t npad;        // n rounded up to a multiple of 4, plus slack for 16-byte copies
    int64_t perm_len;
    size_t off_header, off_pos, zero_bytes;             // [header | pos] zeroed together
    size_t off_keys, off_score, off_sorted, off_uniq, off_dense, off_order, off_ab, off_packed,
        off_partial, off_perm, total;
};

static SnapLayout make_layout(int64_t n, int64_t p, bool per_pod)
{
    SnapLayout L{};
    L.n = n;
    L.p = p;
    L.npad = (n + 3) / 4 * 4 + 4;
    L.perm_len = per_pod ? p + 3 * 2048 : 0;
    size_t o = 0;
    L.off_header = o;  o = align_up(o + sizeof(SnapHeader));
    L.off_pos = o;     o = align_up(o + sizeof(uint32_t) * kKeyArrays * n);
    L.zero_bytes = o;
    L.off_keys = o;    o = align_up(o + sizeof(uint64_t) * kKeyArrays * n);
    L.off_score = o;   o = align_up(o + sizeof(double) * kMaxStrategies * n);
    L.off_sorted = o;  o = align_up(o + sizeof(double) * 2 * L.npad);
    L.off_uniq = o;    o = align_up(o + sizeof(double) * 2 * L.npad);
    L.off_dense = o;   o = align_up(o + sizeof(uint32_t) * 2 * L.npad);
    L.off_order = o;   o = align_up(o + sizeof(uint32_t) * kMaxStrategies * L.npad);
    L.off_ab = o;      o = align_up(o + sizeof(uint2) * kMaxStrategies * L.npad);
    L.off_packed = o;  o = align_up(o + sizeof(uint32_t) * kMaxStrategies * L.npad);
    L.off_partial = o; o = align_up(o + (sizeof(uint64_t) + sizeof(uint32_t) * 2) * kMaxStrategies * kArgmaxMaxBlocks);
    L.off_perm = o;    o = align_up(o + sizeof(uint32_t) * L.perm_len);
    L.total = o;
    return L;
}

size_t snapshot_workspace_bytes(int64_t n_nodes, int64_t n_pods, bool per_pod)
{
    if (n_nodes < 0 || n_pods < 0) return 0;
    return make_layout(n_nodes, n_pods, per_pod).total;
}

struct NodePtrs {
    const long long *alloc_cpu, *alloc_mem, *pod_count, *max_pods;
    const uint8_t *ready;
    const double *cu, *mu;
    int n;
};

static NodePtrs node_ptrs(const kp_node_table &t)
{
    return NodePtrs{(const long long *)t.alloc_cpu_milli, (const long long *)t.alloc_mem_bytes,
                    (const long long *)t.pod_count, (const long long *)t.max_pods, t.ready,
                    t.cpu_usage_pct, t.mem_usage_pct, (int)t.n_nodes};
}

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
    int npad;
};

// ---------------------------------------------------------------------------------------
// plan_small_kernel: whole node plan in ONE CTA (n <= 4096)
// ---------------------------------------------------------------------------------------
struct SmallSmem {
    unsigned long long key[kSmallMaxNodes];
    uint16_t idx[kSmallMaxNodes];
    uint16_t rank[2][kSmallMaxNodes];
    uint32_t warp_sums[32];
    uint32_t total;
};

// ascending bitonic sort of (key, idx) pairs, m = power of two >= 1024, 1024 threads
__device__ __forceinline__ void bitonic_sort(unsigned long long *key, uint16_t *idx, int m)
{
    for (int k = 2; k <= m; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int c = threadIdx.x; c < (m >> 1); c += 1024) {
                const int i = ((c & ~(j - 1)) << 1) | (c & (j - 1));
                const int l = i | j;
                const unsigned long long ka = key[i], kb = key[l];
                const uint16_t ia = idx[i], ib = idx[l];
                const bool gt = ka > kb || (ka == kb && ia > ib);
                const bool asc = (i & k) == 0;
                if (gt == asc) { key[i] = kb; key[l] = ka; idx[i] = ib; idx[l] = ia; }
            }
            __syncthreads();
        }
    }
}

// block-wide inclusive scan helper for 1024 threads: returns the exclusive prefix of `cnt`
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

__global__ void __launch_bounds__(1024, 1)
plan_small_kernel(NodePtrs t, int strat_mask, unsigned long long *__restrict__ keys, PlanOut o)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    SmallSmem &sm = *reinterpret_cast<SmallSmem *>(smem_raw);
    const int tid = threadIdx.x;
    const int n = t.n;
    int m = 1024;
    while (m < n) m <<= 1;

    // derive: keys to global scratch (re-read per sort pass), scores to their final place
    for (int i = tid; i < n; i += 1024) {
        const Derived d = derive_node(t, i, strat_mask);
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            if (!(strat_mask & (1 << s))) continue;
            keys[(size_t)s * n + i] = d.skey[s];
            o.score[(size_t)s * n + i] = d.score[s];
        }
        keys[(size_t)3 * n + i] = d.tkey[0];
        keys[(size_t)4 * n + i] = d.tkey[1];
    }
    if (tid < kMaxStrategies) { o.hdr->group_count[tid] = 0; o.hdr->group_cursor[tid] = 0; }
    if (tid == 0) { o.hdr->ticket = 0; o.hdr->use_swar = 1; o.hdr->reserved = 0; }
    __syncthreads();

    // thresholds: sort, dense ranks, unique values
    for (int j = 0; j < 2; ++j) {
        for (int i = tid; i < m; i += 1024) {
            sm.key[i] = i < n ? keys[(size_t)(3 + j) * n + i] : ~0ull;
            sm.idx[i] = (uint16_t)i;
        }
        __syncthreads();
        bitonic_sort(sm.key, sm.idx, m);
        // each thread owns a contiguous chunk of per = m / 1024 sorted positions
        const int per = m >> 10;
        const int i0 = tid * per;
        uint32_t cnt = 0;
        for (int q = 0; q < per; ++q) {
            const int i = i0 + q;
            cnt += (i < n && (i == 0 || sm.key[i] != sm.key[i - 1])) ? 1u : 0u;
        }
        uint32_t run = block_exclusive_1024(cnt, sm.warp_sums, &sm.total);   // distinct values before i0
        for (int q = 0; q < per; ++q) {
            const int i = i0 + q;
            if (i >= n) break;
            const bool first = i == 0 || sm.key[i] != sm.key[i - 1];
            if (first) { o.uniq[(size_t)j * o.npad + run] = key_to_f64(sm.key[i]); ++run; }
            sm.rank[j][sm.idx[i]] = (uint16_t)(run - 1);
        }
        if (tid == 0) o.hdr->distinct[j] = sm.total;
        __syncthreads();
    }

    // score orders
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(strat_mask & (1 << s))) { if (tid == 0) o.hdr->nvalid[s] = 0; continue; }
        for (int i = tid; i < m; i += 1024) {
            sm.key[i] = i < n ? keys[(size_t)s * n + i] : ~0ull;
            sm.idx[i] = (uint16_t)i;
        }
        __syncthreads();
        bitonic_sort(sm.key, sm.idx, m);
        uint32_t mine = 0;
        for (int i = tid; i < o.npad; i += 1024) {
            const bool live = i < n && sm.key[i] != ~0ull;
            uint2 r = make_uint2(0xffffffffu, 0xffffffffu);
            uint32_t pk = kSwarGuards;                 // pad: fields 0 < every pod field
            if (live) {
                const uint32_t node = sm.idx[i];
                r = make_uint2(sm.rank[0][node], sm.rank[1][node]);
                pk = pack_ranks(r.x, r.y);
                o.order[(size_t)s * o.npad + i] = node;
                ++mine;
            }
            o.ab[(size_t)s * o.npad + i] = r;
            o.packed[(size_t)s * o.npad + i] = pk;
        }
        (void)block_exclusive_1024(mine, sm.warp_sums, &sm.total);   // also the barrier before sm.key is reused
        if (tid == 0) o.hdr->nvalid[s] = sm.total;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// multi-kernel plan (any n): derive -> counting rank -> sorted thresholds -> dense ranks -> tables
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

// one CTA: dense rank of every sorted position + unique values, for both threshold arrays
__global__ void __launch_bounds__(1024, 1)
dense_kernel(const double *__restrict__ sorted, int n_nodes, int npad, uint32_t *__restrict__ dense,
             double *__restrict__ uniq, SnapHeader *__restrict__ hdr)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t total;
    const int tid = threadIdx.x;
    const int per = (n_nodes + 1023) / 1024;
    const int i0 = min(n_nodes, tid * per), i1 = min(n_nodes, i0 + per);
    uint32_t tot[2];
    for (int j = 0; j < 2; ++j) {
        const double *S = sorted + (size_t)j * npad;
        uint32_t cnt = 0;
        for (int i = i0; i < i1; ++i) cnt += (i == 0 || S[i] != S[i - 1]) ? 1u : 0u;
        uint32_t run = block_exclusive_1024(cnt, warp_sums, &total);
        for (int i = i0; i < i1; ++i) {
            if (i == 0 || S[i] != S[i - 1]) { uniq[(size_t)j * npad + run] = S[i]; ++run; }
            dense[(size_t)j * npad + i] = run - 1;
        }
        tot[j] = total;
        __syncthreads();
    }
    if (tid == 0) {
        hdr->distinct[0] = tot[0];
        hdr->distinct[1] = tot[1];
        hdr->use_swar = (tot[0] <= kSwarMaxDistinct && tot[1] <= kSwarMaxDistinct) ? 1u : 0u;
    }
}

__global__ void __launch_bounds__(256)
scatter_tables_kernel(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ pos,
                      const uint32_t *__restrict__ dense, int n_nodes, int strat_mask, PlanOut o)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= o.npad) return;
    const bool swar = o.hdr->use_swar != 0;
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

// ---------------------------------------------------------------------------------------
// per-pod strategy: group pods by strategy so that a CTA scans one best-first table
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
group_count_kernel(const long long *__restrict__ sid, long long n_pods, SnapHeader *hdr)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int g = j < n_pods ? canon_strategy(sid[j]) : -1;
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        const unsigned b = __ballot_sync(0xffffffffu, g == s);
        if ((threadIdx.x & 31) == 0 && b) atomicAdd(&hdr->group_count[s], __popc(b));
    }
}

__device__ __forceinline__ void group_begins(const SnapHeader *hdr, int pods_per_cta,
                                             long long begin[kMaxStrategies])
{
    long long o = 0;
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        begin[s] = o;
        const long long c = hdr->group_count[s];
        o += (c + pods_per_cta - 1) / pods_per_cta * pods_per_cta;
    }
}

__global__ void __launch_bounds__(256)
group_assign_kernel(const long long *__restrict__ sid, long long n_pods, int pods_per_cta,
                    SnapHeader *hdr, uint32_t *__restrict__ perm)
{
    long long begin[kMaxStrategies];
    group_begins(hdr, pods_per_cta, begin);
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int g = j < n_pods ? canon_strategy(sid[j]) : -1;
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        const unsigned b = __ballot_sync(0xffffffffu, g == s);
        if (!b) continue;
        uint32_t base = 0;
        const int leader = __ffs(b) - 1;
        if (lane == leader) base = atomicAdd(&hdr->group_cursor[s], __popc(b));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (g == s) perm[begin[s] + base + __popc(b & ((1u << lane) - 1))] = (uint32_t)j;
    }
}

// ---------------------------------------------------------------------------------------
// place_kernel: the hot loop
// ---------------------------------------------------------------------------------------
struct PlaceParams {
    const long long *req_cpu, *req_mem;
    const uint32_t *perm;              // NULL = identity
    long long n_pods;
    int n_nodes, npad, uniform_strategy;
    int one;                           // always 1; a run-time value so the IMAD update is not folded away
    const double *uniq;                // [2][npad] ascending distinct thresholds
    const uint2 *ab;                   // [3][npad] best-first {RA, RB}
    const uint32_t *packed;            // [3][npad] best-first packed ranks
    const uint32_t *order;             // [3][npad]
    const double *score;               // [3][n]
    const SnapHeader *hdr;
    long long *out_idx;
    double *out_score;
};

// #{i : arr[i] < x} in an ascending array
__device__ __forceinline__ int lower_bound_lt(const double *arr, int n, double x)
{
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (arr[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// t = w - x keeps both guard bits iff neither 15-bit field borrowed  <=>  pod fits in both dimensions
__device__ __forceinline__ void swar_update_alu(int &best, uint32_t x, uint32_t w, int pos)
{
    asm volatile(
        "{ .reg .pred p; .reg .b32 t, d;\n"
        "  sub.u32 t, %1, %2;\n"
        "  lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"      // d = ~t & guards ; p = (d != 0) = does not fit
        "  @!p mov.s32 %0, %3; }"
        : "+r"(best) : "r"(w), "r"(x), "r"(pos));
}
// same test; the update is a guarded multiply-add so that it issues on the FMA pipe:
// best_enc = pos * 1 + x   (decoded later as best_enc - x)
__device__ __forceinline__ void swar_update_fma(int &best_enc, uint32_t x, uint32_t w, int pos, int one)
{
    asm volatile(
        "{ .reg .pred p; .reg .b32 t, d;\n"
        "  sub.u32 t, %1, %2;\n"
        "  lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"
        "  @!p mad.lo.s32 %0, %3, %4, %2; }"
        : "+r"(best_enc) : "r"(w), "r"(x), "r"(pos), "r"(one));
}

template <int THREADS, int R>
__global__ void __launch_bounds__(THREADS)
place_kernel(const PlaceParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned char *tile0 = smem_raw;
    unsigned char *tile1 = smem_raw + kTileBytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + 2 * kTileBytes);

    const int tid = threadIdx.x;
    const long long slot0 = (long long)blockIdx.x * (THREADS * R);
    int s = p.uniform_strategy;
    long long gbegin = 0, gcount = p.n_pods;
    if (s < 0) {
        long long begin[kMaxStrategies];
        group_begins(p.hdr, THREADS * R, begin);
        s = 0;
#pragma unroll
        for (int g = 1; g < kMaxStrategies; ++g) if (slot0 >= begin[g]) s = g;
        gbegin = begin[s];
        gcount = p.hdr->group_count[s];
    }
    if (slot0 - gbegin >= gcount) return;                    // CTA-uniform

    const int nvalid = (int)p.hdr->nvalid[s];
    const bool swar = p.hdr->use_swar != 0;
    const int da = (int)p.hdr->distinct[0], db = (int)p.hdr->distinct[1];
    const int tile_nodes = swar ? kTileNodesPacked : kTileNodesWide;
    const int elem = swar ? 4 : 8;
    const int ntiles = (nvalid + tile_nodes - 1) / tile_nodes;
    const unsigned char *table = swar ? reinterpret_cast<const unsigned char *>(p.packed + (size_t)s * p.npad)
                                      : reinterpret_cast<const unsigned char *>(p.ab + (size_t)s * p.npad);

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }

    // ---- pod prologue: requests -> dense threshold ranks (exact, see file header) ----
    int ra[R], rb[R];
    long long pod[R];
    const double *ua = p.uniq, *ub = p.uniq + p.npad;
    const bool stage = da <= kTileBytes / 8 && db <= kTileBytes / 8;
    if (stage) {        // both unique-threshold arrays fit the (still unused) tile buffers
        double *sa = reinterpret_cast<double *>(tile0);
        double *sb = reinterpret_cast<double *>(tile1);
        for (int i = tid; i < da; i += THREADS) sa[i] = ua[i];
        for (int i = tid; i < db; i += THREADS) sb[i] = ub[i];
        ua = sa;
        ub = sb;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long long slot = slot0 + (long long)r * THREADS + tid;
        const bool live = slot - gbegin < gcount;
        pod[r] = live ? (p.perm ? (long long)p.perm[slot] : slot) : -1;
        ra[r] = rb[r] = 0x7ffe;                              // dead slot: above every node rank
        if (live) {
            ra[r] = lower_bound_lt(ua, da, req_cores(p.req_cpu[pod[r]]));
            rb[r] = lower_bound_lt(ub, db, req_gib(p.req_mem[pod[r]]));
        }
    }
    if (stage) {
        // generic-proxy accesses to the buffers are done; order them before the async-proxy writes
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    auto tile_count = [&](int t) { return min(tile_nodes, ((nvalid + 3) & ~3) - t * tile_nodes); };
    if (tid == 0 && ntiles > 0) {
        const int t = ntiles - 1;
        const uint32_t bytes = (uint32_t)(tile_count(t) * elem);
        mbar_expect_tx(&bars[0], bytes);
        bulk_load(tile0, table + (size_t)t * tile_nodes * elem, bytes, &bars[0]);
    }

    int best[R];
    if (swar) {
        // ---- packed path: 3 instructions per pair, split over the ALU and FMA pipes ----
        uint32_t x[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            x[r] = pod[r] >= 0 ? (((uint32_t)rb[r] + 1u) << 16) | ((uint32_t)ra[r] + 1u) : kSwarDeadPod;
            best[r] = (r & 1) ? (int)x[r] - 1 : -1;          // odd pods hold pos + x (see swar_update_fma)
        }
        int it = 0;
        for (int t = ntiles - 1; t >= 0; --t, ++it) {
            const int buf = it & 1;
            if (tid == 0 && t > 0) {
                const uint32_t bytes = (uint32_t)(tile_count(t - 1) * elem);
                mbar_expect_tx(&bars[buf ^ 1], bytes);
                bulk_load(buf ? tile0 : tile1, table + (size_t)(t - 1) * tile_nodes * elem, bytes, &bars[buf ^ 1]);
            }
            mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
            const uint4 *tp = reinterpret_cast<const uint4 *>(buf ? tile1 : tile0);
            const int base = t * tile_nodes;
            const int nquad = tile_count(t) >> 2;
            for (int i = nquad - 1; i >= 0; --i) {
                const uint4 v = tp[i];                       // four nodes, same address in every lane
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 3; k >= 0; --k) {
                    const int pos = base + 4 * i + k;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        if (r & 1) swar_update_fma(best[r], x[r], w[k], pos, p.one);
                        else swar_update_alu(best[r], x[r], w[k], pos);
                    }
                }
            }
            __syncthreads();                                 // buffer free for the next prefetch
        }
#pragma unroll
        for (int r = 0; r < R; ++r) if (r & 1) best[r] -= (int)x[r];
    } else {
        // ---- wide path (more than 32000 distinct thresholds): two 32-bit compares + select ----
#pragma unroll
        for (int r = 0; r < R; ++r) {
            best[r] = -1;
            if (pod[r] < 0) ra[r] = rb[r] = 0x7fffffff;
        }
        int it = 0;
        for (int t = ntiles - 1; t >= 0; --t, ++it) {
            const int buf = it & 1;
            if (tid == 0 && t > 0) {
                const uint32_t bytes = (uint32_t)(tile_count(t - 1) * elem);
                mbar_expect_tx(&bars[buf ^ 1], bytes);
                bulk_load(buf ? tile0 : tile1, table + (size_t)(t - 1) * tile_nodes * elem, bytes, &bars[buf ^ 1]);
            }
            mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
            const uint4 *tp = reinterpret_cast<const uint4 *>(buf ? tile1 : tile0);
            const int base = t * tile_nodes;
            const int npair = tile_count(t) >> 1;
#pragma unroll 2
            for (int i = npair - 1; i >= 0; --i) {
                const uint4 v = tp[i];                       // two nodes
                const int p1 = base + 2 * i + 1, p0 = base + 2 * i;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (ra[r] <= (int)v.z && rb[r] <= (int)v.w) best[r] = p1;
                    if (ra[r] <= (int)v.x && rb[r] <= (int)v.y) best[r] = p0;
                }
            }
            __syncthreads();
        }
    }

    // ---- epilogue: rank position -> node index and score ----
    const uint32_t *order = p.order + (size_t)s * p.npad;
    const double *score = p.score + (size_t)s * p.n_nodes;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (pod[r] < 0) continue;
        long long idx = -1;
        double sc = -1.0;
        if (best[r] >= 0) {
            idx = order[best[r]];
            sc = score[idx];
        }
        p.out_idx[pod[r]] = idx;
        p.out_score[pod[r]] = sc;
    }
}

// ---------------------------------------------------------------------------------------
// KP_FILTER_READY: argmax over nodes (+ fill).  The filter does not look at the pod
// (scheduler.py:521: _fallback_decision takes no pod), so one argmax serves every pod.
// ---------------------------------------------------------------------------------------
struct Cand { unsigned long long key; uint32_t idx; };
__device__ __forceinline__ bool better(const Cand &a, const Cand &b)
{
    return a.key > b.key || (a.key == b.key && a.idx < b.idx);
}
__device__ __forceinline__ Cand warp_best(Cand c)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Cand d;
        d.key = __shfl_xor_sync(0xffffffffu, c.key, o);
        d.idx = __shfl_xor_sync(0xffffffffu, c.idx, o);
        if (better(d, c)) c = d;
    }
    return c;
}

struct FillParams {
    const long long *sid;      // NULL = uniform
    int uniform_strategy;
    long long n_pods;
    long long *out_idx;
    double *out_score;
};

__device__ __forceinline__ void fill_range(const FillParams &f, const SnapHeader *hdr,
                                           long long j0, long long stride)
{
    for (long long j = j0; j < f.n_pods; j += stride) {
        const int s = f.sid ? canon_strategy(f.sid[j]) : f.uniform_strategy;
        f.out_idx[j] = hdr->winner[s];
        f.out_score[j] = hdr->winner_score[s];
    }
}

constexpr long long kInlineFillMax = 8192;

__global__ void __launch_bounds__(256)
ready_argmax_kernel(NodePtrs t, int strat_mask, SnapHeader *hdr, unsigned long long *part_key,
                    uint32_t *part_idx, FillParams fill)
{
    __shared__ Cand sm[kMaxStrategies][8];
    __shared__ bool is_last;
    Cand c[kMaxStrategies];
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) { c[s].key = 0ull; c[s].idx = 0xffffffffu; }
    for (int n = blockIdx.x * 256 + threadIdx.x; n < t.n; n += gridDim.x * 256) {
        if (!t.ready[n]) continue;
        const double cu = t.cu[n], mu = t.mu[n];
        const long long pc = t.pod_count[n], mp = t.max_pods[n];
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            if (!(strat_mask & (1 << s))) continue;
            Cand d;
            d.key = node_score(s, cu, mu, pc, mp).key;
            d.idx = (uint32_t)n;
            if (d.key != 0ull && better(d, c[s])) c[s] = d;
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        c[s] = warp_best(c[s]);
        if (lane == 0) sm[s][warp] = c[s];
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            Cand d = lane < 8 ? sm[s][lane] : Cand{0ull, 0xffffffffu};
            d = warp_best(d);
            if (lane == 0) {
                part_key[s * kArgmaxMaxBlocks + blockIdx.x] = d.key;
                part_idx[s * kArgmaxMaxBlocks + blockIdx.x] = d.idx;
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&hdr->ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (warp == 0) {
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            Cand d{0ull, 0xffffffffu};
            for (int b = lane; b < (int)gridDim.x; b += 32) {
                Cand e{part_key[s * kArgmaxMaxBlocks + b], part_idx[s * kArgmaxMaxBlocks + b]};
                if (better(e, d)) d = e;
            }
            d = warp_best(d);
            if (lane == 0) {
                const bool ok = (strat_mask & (1 << s)) && d.key != 0ull;
                hdr->winner[s] = ok ? (long long)d.idx : -1;
                hdr->winner_score[s] =
                    ok ? node_score(s, t.cu[d.idx], t.mu[d.idx], t.pod_count[d.idx], t.max_pods[d.idx]).score
                       : -1.0;
            }
        }
    }
    __syncthreads();
    if (fill.n_pods <= kInlineFillMax) fill_range(fill, hdr, threadIdx.x, 256);
}

__global__ void __launch_bounds__(256)
fill_kernel(FillParams fill, const SnapHeader *hdr)
{
    fill_range(fill, hdr, (long long)blockIdx.x * 256 + threadIdx.x, (long long)gridDim.x * 256);
}

// ---------------------------------------------------------------------------------------
// per-node derived values (kp_node_scores)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
node_scores_kernel(NodePtrs t, int strategy, int capacity, double *out_score, double *out_fc,
                   double *out_fm, uint8_t *out_elig)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= t.n) return;
    const double cu = t.cu[n], mu = t.mu[n];
    const long long pc = t.pod_count[n], mp = t.max_pods[n];
    const NodeScore ns = node_score(strategy, cu, mu, pc, mp);
    const double fc = free_cpu_cores(t.alloc_cpu[n], cu);
    const double fm = free_mem_gib(t.alloc_mem[n], mu);
    bool elig = t.ready[n] != 0 && ns.key != 0ull;
    if (capacity) elig = elig && pc < mp && fc == fc && fm == fm;
    if (out_score) out_score[n] = ns.score;
    if (out_fc) out_fc[n] = fc;
    if (out_fm) out_fm[n] = fm;
    if (out_elig) out_elig[n] = elig ? 1 : 0;
}

int node_scores_device(const kp_node_table &nodes, int strategy, int filter_mode,
                       double *out_score, double *out_free_cpu, double *out_free_mem,
                       uint8_t *out_eligible, cudaStream_t stream)
{
    if (nodes.n_nodes == 0) return KP_OK;
    const NodePtrs t = node_ptrs(nodes);
    node_scores_kernel<<<(t.n + 255) / 256, 256, 0, stream>>>(
        t, canon_strategy(strategy), filter_mode == KP_FILTER_CAPACITY, out_score, out_free_cpu,
        out_free_mem, out_eligible);
    KP_LAUNCH_CHECK("node_scores_kernel");
    return KP_OK;
}

// ---------------------------------------------------------------------------------------
// orchestration
// ---------------------------------------------------------------------------------------
template <int THREADS, int R>
static int launch_place(const PlaceParams &pp, long long grid, cudaStream_t stream)
{
    constexpr size_t smem = 2 * kTileBytes + 2 * sizeof(uint64_t);
    static bool attr_set = false;
    if (!attr_set) {
        KP_CUDA(cudaFuncSetAttribute(place_kernel<THREADS, R>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    place_kernel<THREADS, R><<<(unsigned)grid, THREADS, smem, stream>>>(pp);
    KP_LAUNCH_CHECK("place_kernel");
    return KP_OK;
}

int snapshot_place_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                          int filter_mode, int64_t *out_idx, double *out_score, void *ws,
                          size_t ws_bytes, cudaStream_t stream)
{
    const int64_t n = nodes.n_nodes, P = pods.n_pods;
    if (P == 0) return KP_OK;
    const bool per_pod = pods.strategy_ids != nullptr;
    const SnapLayout L = make_layout(n, P, per_pod);
    if (!ws || ws_bytes < L.total || (reinterpret_cast<uintptr_t>(ws) & (kAlign - 1))) {
        set_error("snapshot workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", L.total,
                  ws_bytes, ws);
        return KP_ERR_WORKSPACE;
    }
    unsigned char *w = static_cast<unsigned char *>(ws);
    SnapHeader *hdr = reinterpret_cast<SnapHeader *>(w + L.off_header);
    const int uniform = per_pod ? -1 : canon_strategy(strategy);
    const int strat_mask = per_pod ? 7 : (1 << uniform);
    const bool capacity = filter_mode == KP_FILTER_CAPACITY;
    const NodePtrs t = node_ptrs(nodes);

    if (!capacity || n == 0) {
        // pod-independent filter: one argmax per strategy + fill  (n == 0: winners stay -1)
        KP_CUDA(cudaMemsetAsync(hdr, 0, sizeof(SnapHeader), stream));
        FillParams fill{(const long long *)pods.strategy_ids, uniform < 0 ? 0 : uniform, P,
                        (long long *)out_idx, out_score};
        const int blocks = (int)std::min<int64_t>(kArgmaxMaxBlocks, std::max<int64_t>(1, (n + 1023) / 1024));
        ready_argmax_kernel<<<blocks, 256, 0, stream>>>(
            t, strat_mask, hdr, reinterpret_cast<unsigned long long *>(w + L.off_partial),
            reinterpret_cast<uint32_t *>(w + L.off_partial + sizeof(uint64_t) * kMaxStrategies * kArgmaxMaxBlocks),
            fill);
        KP_LAUNCH_CHECK("ready_argmax_kernel");
        if (P > kInlineFillMax) {
            const int fb = (int)std::min<int64_t>((P + 255) / 256, (int64_t)sm_count_cached() * 8);
            fill_kernel<<<fb, 256, 0, stream>>>(fill, hdr);
            KP_LAUNCH_CHECK("fill_kernel");
        }
        return KP_OK;
    }

    // ---- node plan ----
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(w + L.off_keys);
    uint32_t *pos = reinterpret_cast<uint32_t *>(w + L.off_pos);
    double *sorted = reinterpret_cast<double *>(w + L.off_sorted);
    uint32_t *dense = reinterpret_cast<uint32_t *>(w + L.off_dense);
    uint32_t *perm = per_pod ? reinterpret_cast<uint32_t *>(w + L.off_perm) : nullptr;
    PlanOut po{};
    po.hdr = hdr;
    po.score = reinterpret_cast<double *>(w + L.off_score);
    po.uniq = reinterpret_cast<double *>(w + L.off_uniq);
    po.order = reinterpret_cast<uint32_t *>(w + L.off_order);
    po.ab = reinterpret_cast<uint2 *>(w + L.off_ab);
    po.packed = reinterpret_cast<uint32_t *>(w + L.off_packed);
    po.npad = (int)L.npad;

    ProfileEvents &pe = profile_events();
    if (pe.enabled) KP_CUDA(cudaEventRecord(pe.e0, stream));
    if (n <= kSmallMaxNodes) {
        static bool attr_set = false;
        if (!attr_set) {
            KP_CUDA(cudaFuncSetAttribute(plan_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(SmallSmem)));
            attr_set = true;
        }
        plan_small_kernel<<<1, 1024, sizeof(SmallSmem), stream>>>(t, strat_mask, keys, po);
        KP_LAUNCH_CHECK("plan_small_kernel");
    } else {
        KP_CUDA(cudaMemsetAsync(w, 0, L.zero_bytes, stream));
        const int nb = (int)((n + 255) / 256);
        derive_kernel<<<nb, 256, 0, stream>>>(t, strat_mask, keys, po.score, hdr);
        KP_LAUNCH_CHECK("derive_kernel");
        RankList list{};
        for (int s = 0; s < kMaxStrategies; ++s) if (strat_mask & (1 << s)) list.idx[list.count++] = s;
        list.idx[list.count++] = 3;
        list.idx[list.count++] = 4;
        {
            // enough chunks to give every SM a few CTAs, but at least one rank tile each
            const int sms = sm_count_cached();
            int chunks = std::max(1, (4 * sms) / std::max(1, nb * list.count));
            chunks = (int)std::min<int64_t>(chunks, (n + kRankTile - 1) / kRankTile);
            const int chunk_len = (int)(((n + chunks - 1) / chunks + kRankTile - 1) / kRankTile * kRankTile);
            chunks = (int)((n + chunk_len - 1) / chunk_len);
            rank_kernel<<<dim3(nb, chunks, list.count), 256, 0, stream>>>(keys, pos, (int)n, chunk_len, list);
            KP_LAUNCH_CHECK("rank_kernel");
        }
        scatter_sorted_kernel<<<nb, 256, 0, stream>>>(keys, pos, (int)n, (int)L.npad, sorted);
        KP_LAUNCH_CHECK("scatter_sorted_kernel");
        dense_kernel<<<1, 1024, 0, stream>>>(sorted, (int)n, (int)L.npad, dense, po.uniq, hdr);
        KP_LAUNCH_CHECK("dense_kernel");
        scatter_tables_kernel<<<(int)((L.npad + 255) / 256), 256, 0, stream>>>(keys, pos, dense, (int)n,
                                                                               strat_mask, po);
        KP_LAUNCH_CHECK("scatter_tables_kernel");
    }

    // ---- pods ----
    const int sms = sm_count_cached();
    int cfg;   // pods per CTA: 2048 (256x8), 1024 (128x8), 512 (128x4), 256 (64x4)
    if (P >= (int64_t)2048 * 2 * sms) cfg = 2048;
    else if (P >= (int64_t)1024 * 2 * sms) cfg = 1024;
    else if (P >= (int64_t)512 * 2 * sms) cfg = 512;
    else cfg = 256;
    if (per_pod) {
        const int gb = (int)((P + 255) / 256);
        group_count_kernel<<<gb, 256, 0, stream>>>((const long long *)pods.strategy_ids, P, hdr);
        KP_LAUNCH_CHECK("group_count_kernel");
        group_assign_kernel<<<gb, 256, 0, stream>>>((const long long *)pods.strategy_ids, P, cfg,
                                                    hdr, perm);
        KP_LAUNCH_CHECK("group_assign_kernel");
    }
    PlaceParams pp{};
    pp.req_cpu = (const long long *)pods.req_cpu_milli;
    pp.req_mem = (const long long *)pods.req_mem_bytes;
    pp.perm = perm;
    pp.n_pods = P;
    pp.n_nodes = (int)n;
    pp.npad = (int)L.npad;
    pp.uniform_strategy = uniform;
    pp.one = 1;
    pp.uniq = po.uniq;
    pp.ab = po.ab;
    pp.packed = po.packed;
    pp.order = po.order;
    pp.score = po.score;
    pp.hdr = hdr;
    pp.out_idx = (long long *)out_idx;
    pp.out_score = out_score;
    const long long grid = (P + cfg - 1) / cfg + (per_pod ? kMaxStrategies - 1 : 0);
    if (pe.enabled) KP_CUDA(cudaEventRecord(pe.e1, stream));
    int st;
    switch (cfg) {
        case 2048: st = launch_place<256, 8>(pp, grid, stream); break;
        case 1024: st = launch_place<128, 8>(pp, grid, stream); break;
        case 512: st = launch_place<128, 4>(pp, grid, stream); break;
        default: st = launch_place<64, 4>(pp, grid, stream); break;
    }
    if (st == KP_OK && pe.enabled) {
        KP_CUDA(cudaEventRecord(pe.e2, stream));
        pe.valid = true;
    }
    return st;
}

// ---------------------------------------------------------------------------------------
// kp_snapshot_place_noalloc: the snapshot path for callers that cannot provide scratch
// ---------------------------------------------------------------------------------------
// Some hosts of this library cannot hand us a workspace: a scheduler sidecar that shares
// its GPU with an inference server whose allocator already owns all of HBM has nothing to
// cudaMalloc from, and a placement loop with a latency budget must not call cudaMalloc /
// cudaFree at all (both take the driver's allocator lock and synchronise the device).
// For them the node plan lives in this fixed array instead, which is part of the loaded
// module: every GPU that loads the library carries these 2 MiB until the process exits.
//
// The array's name is only a host-side handle in host code.  The device address is
// looked up with cudaGetSymbolAddress, once per device, while that device is current.
constexpr int64_t kNoAllocMaxNodes = 8192;
constexpr int64_t kNoAllocMaxPodsPerPodStrategy = 65536;
constexpr size_t kNoAllocScratchBytes = 2u << 20;          // 2 MiB >= layout(8192 nodes, 65536 pods, per-pod)
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

int snapshot_place_noalloc(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                           int filter_mode, int64_t *out_idx, double *out_score, cudaStream_t stream)
{
    const bool per_pod = pods.strategy_ids != nullptr;
    if (nodes.n_nodes > kNoAllocMaxNodes) {
        set_error("kp_snapshot_place_noalloc: %lld nodes exceed the built-in scratch limit of %lld "
                  "(use kp_snapshot_place with a caller-provided workspace)",
                  (long long)nodes.n_nodes, (long long)kNoAllocMaxNodes);
        return KP_ERR_CAPACITY;
    }
    if (per_pod && pods.n_pods > kNoAllocMaxPodsPerPodStrategy) {
        set_error("kp_snapshot_place_noalloc: %lld pods with per-pod strategy ids exceed the built-in "
                  "scratch limit of %lld (uniform-strategy calls have no pod limit)",
                  (long long)pods.n_pods, (long long)kNoAllocMaxPodsPerPodStrategy);
        return KP_ERR_CAPACITY;
    }
    if (make_layout(nodes.n_nodes, pods.n_pods, per_pod).total > kNoAllocScratchBytes) {
        set_error("kp_snapshot_place_noalloc: internal scratch too small for this call");
        return KP_ERR_CAPACITY;
    }
    // per-device set-up on the first call: resolve the array's device address
    static void *scratch_addr[64] = {nullptr};
    int dev = -1;
    KP_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("kp_snapshot_place_noalloc: device ordinal %d unsupported", dev); return KP_ERR_NO_DEVICE; }
    if (!scratch_addr[dev]) {
        void *ptr = nullptr;
        KP_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        scratch_addr[dev] = ptr;
    }
    // the array keeps the previous call's contents; snapshot_place_device rewrites every
    // counter, rank accumulator and table entry (including padding) it later reads
    return snapshot_place_device(nodes, pods, strategy, filter_mode, out_idx, out_score,
                                 scratch_addr[dev], kNoAllocScratchBytes, stream);
}

}  // namespace kp

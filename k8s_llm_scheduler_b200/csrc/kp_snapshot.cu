// Snapshot placement: P independent fallback decisions against one node table.
//
// Replaces one `_fallback_decision` call per pod (scheduler.py:521-559).  The reference
// recomputes every node's score for every pod; here everything that depends on the node
// alone is hoisted into a per-snapshot "node plan", and the P x N part that remains is a
// pure compare/select loop that never touches HBM:
//
//   K_derive   per node: score under each strategy in play (exact fp64, reference op
//              order), eligibility (Ready, score > -1 sentinel, and in capacity mode
//              pod_count < max_pods), free cpu / free memory thresholds.
//   K_rank     stable rank of every node under (score desc, index asc) -- the reference's
//              "strict >, first in list order wins" (scheduler.py:547) as a total order --
//              and rank-compression of the two thresholds: RA[n] = #{m : free_cpu[m] <
//              free_cpu[n]}, so that  req <= free_cpu[n]  <=>  ra(req) <= RA[n]  with
//              ra(req) = #{m : free_cpu[m] < req}.  Exact for any fp64 inputs; turns the
//              per-pair fp64 compares into 32-bit integer compares.
//   K_scatter  node ranks written in best-first order: ab[s][pos] = {RA, RB}.
//   K_place    the hot loop.  One thread owns R pods in registers; tiles of the best-first
//              {RA, RB} table are staged in shared memory with 1-D bulk async copies
//              (TMA engine) on a 2-deep mbarrier ring; every lane reads the same node
//              (broadcast LDS.128 = 2 nodes); per pod x node pair: 2 integer compares and
//              one select.  Scanning worst -> best and overwriting on "fits" leaves the
//              best-ranked fitting node, whose index IS the argmax with the reference's
//              tie-break; no per-pair index bookkeeping, no P x N matrix.
//
// KP_FILTER_READY (the reference's own filter) is pod-independent: every pod of a
// strategy gets the same node, so that mode is an argmax over nodes plus a fill.
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
    long long winner[kMaxStrategies];       // ready mode: argmax node per strategy
    double winner_score[kMaxStrategies];
};

constexpr int kKeyArrays = kMaxStrategies + 2;   // 3 score orders + cpu threshold + mem threshold
constexpr int kArgmaxMaxBlocks = 256;

struct SnapLayout {
    int64_t n, p;
    int64_t npad;        // n rounded up to a multiple of 4, plus slack for 16-byte copies
    int64_t perm_len;
    size_t off_header, off_pos, off_less, zero_bytes;   // [header | pos | less] zeroed together
    size_t off_keys, off_score, off_sorted_ta, off_sorted_tb, off_order, off_ab, off_partial,
        off_perm, total;
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
    L.off_less = o;    o = align_up(o + sizeof(uint32_t) * 2 * n);
    L.zero_bytes = o;
    L.off_keys = o;    o = align_up(o + sizeof(uint64_t) * kKeyArrays * n);
    L.off_score = o;   o = align_up(o + sizeof(double) * kMaxStrategies * n);
    L.off_sorted_ta = o; o = align_up(o + sizeof(double) * L.npad);
    L.off_sorted_tb = o; o = align_up(o + sizeof(double) * L.npad);
    L.off_order = o;   o = align_up(o + sizeof(uint32_t) * kMaxStrategies * L.npad);
    L.off_ab = o;      o = align_up(o + sizeof(uint2) * kMaxStrategies * L.npad);
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

// ---------------------------------------------------------------------------------------
// K_derive
// ---------------------------------------------------------------------------------------
// keys[s][n]  = ~score_key (ascending sort = best first); ~0 = ineligible (sorts last)
// keys[3][n]  = order key of free_cpu, keys[4][n] = order key of free_mem (ascending)
__global__ void __launch_bounds__(256)
derive_kernel(NodePtrs t, int strat_mask, int capacity, unsigned long long *__restrict__ keys,
              double *__restrict__ score, SnapHeader *__restrict__ hdr)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = n < t.n;
    double cu = 0, mu = 0;
    long long pc = 0, mp = 0;
    bool rdy = false;
    if (live) {
        cu = t.cu[n]; mu = t.mu[n]; pc = t.pod_count[n]; mp = t.max_pods[n];
        rdy = t.ready[n] != 0;
    }
    bool cap_ok = true;
    if (capacity) {
        double ta = 0, tb = 0;
        if (live) {
            ta = free_cpu_cores(t.alloc_cpu[n], cu);
            tb = free_mem_gib(t.alloc_mem[n], mu);
        }
        const bool nan = (ta != ta) || (tb != tb);
        cap_ok = !nan && pc < mp;
        if (ta != ta) ta = __longlong_as_double(0x7ff0000000000000ll);   // NaN never fits: park at +inf
        if (tb != tb) tb = __longlong_as_double(0x7ff0000000000000ll);
        if (live) {
            keys[(size_t)3 * t.n + n] = f64_order_key(ta);
            keys[(size_t)4 * t.n + n] = f64_order_key(tb);
        }
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(strat_mask & (1 << s))) continue;
        bool elig = false;
        if (live) {
            const NodeScore ns = node_score(s, cu, mu, pc, mp);
            elig = rdy && cap_ok && ns.key != 0ull;
            score[(size_t)s * t.n + n] = ns.score;
            keys[(size_t)s * t.n + n] = elig ? ~ns.key : ~0ull;
        }
        const unsigned ballot = __ballot_sync(0xffffffffu, elig);
        if ((threadIdx.x & 31) == 0 && ballot) atomicAdd(&hdr->nvalid[s], __popc(ballot));
    }
}

// ---------------------------------------------------------------------------------------
// K_rank: stable ascending rank by counting, tiled through shared memory.
//   pos[j][n]  = #{m : K[m] < K[n]  or  (K[m] == K[n] and m < n)}
//   less[j][n] = #{m : K[m] < K[n]}                      (threshold arrays only)
// grid = (ceil(n/256), chunks, active arrays); partial counts are added atomically.
// ---------------------------------------------------------------------------------------
struct RankList { int idx[kKeyArrays]; int count; };
constexpr int kRankTile = 1024;

__global__ void __launch_bounds__(256)
rank_kernel(const unsigned long long *__restrict__ keys, uint32_t *__restrict__ pos,
            uint32_t *__restrict__ less, int n_nodes, int chunk_len, RankList list)
{
    __shared__ unsigned long long tile[kRankTile];
    const int j = list.idx[blockIdx.z];
    const unsigned long long *K = keys + (size_t)j * n_nodes;
    const int n = blockIdx.x * 256 + threadIdx.x;
    const unsigned long long kn = n < n_nodes ? K[n] : 0ull;
    const int m0 = blockIdx.y * chunk_len;
    const int m1 = min(n_nodes, m0 + chunk_len);
    uint32_t c_pos = 0, c_less = 0;
    for (int base = m0; base < m1; base += kRankTile) {
        for (int i = threadIdx.x; i < kRankTile; i += 256) {
            const int m = base + i;
            tile[i] = m < m1 ? K[m] : ~0ull;
        }
        __syncthreads();
        const int cnt = min(kRankTile, m1 - base);
        if (base + cnt <= n) {                 // every m in the tile is < n: ties count
#pragma unroll 8
            for (int i = 0; i < cnt; ++i) {
                const unsigned long long km = tile[i];
                c_less += km < kn;
                c_pos += km <= kn;
            }
        } else if (base > n) {                 // every m is > n: ties do not count
#pragma unroll 8
            for (int i = 0; i < cnt; ++i) {
                const unsigned long long km = tile[i];
                const uint32_t lt = km < kn;
                c_less += lt;
                c_pos += lt;
            }
        } else {
            for (int i = 0; i < cnt; ++i) {
                const unsigned long long km = tile[i];
                const uint32_t lt = km < kn;
                c_less += lt;
                c_pos += lt | (uint32_t)((km == kn) & (base + i < n));
            }
        }
        __syncthreads();
    }
    if (n < n_nodes) {
        atomicAdd(&pos[(size_t)j * n_nodes + n], c_pos);
        if (j >= kMaxStrategies) atomicAdd(&less[(size_t)(j - kMaxStrategies) * n_nodes + n], c_less);
    }
}

// ---------------------------------------------------------------------------------------
// K_scatter
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double key_to_f64(unsigned long long k)
{
    const unsigned long long b = (k >> 63) ? (k ^ 0x8000000000000000ull) : ~k;
    return __longlong_as_double((long long)b);
}

__global__ void __launch_bounds__(256)
scatter_kernel(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ pos,
               const uint32_t *__restrict__ less, int n_nodes, int npad, int strat_mask,
               int capacity, double *__restrict__ sorted_ta, double *__restrict__ sorted_tb,
               uint32_t *__restrict__ order, uint2 *__restrict__ ab)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_nodes) return;
    uint2 r = make_uint2(0u, 0u);
    if (capacity) {
        sorted_ta[pos[(size_t)3 * n_nodes + n]] = key_to_f64(keys[(size_t)3 * n_nodes + n]);
        sorted_tb[pos[(size_t)4 * n_nodes + n]] = key_to_f64(keys[(size_t)4 * n_nodes + n]);
        r = make_uint2(less[n], less[(size_t)n_nodes + n]);
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        if (!(strat_mask & (1 << s))) continue;
        if (keys[(size_t)s * n_nodes + n] == ~0ull) continue;    // ineligible: not in the table
        const uint32_t p = pos[(size_t)s * n_nodes + n];
        order[(size_t)s * npad + p] = (uint32_t)n;
        ab[(size_t)s * npad + p] = r;
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
// K_place: the hot loop
// ---------------------------------------------------------------------------------------
struct PlaceParams {
    const long long *req_cpu, *req_mem;
    const uint32_t *perm;              // NULL = identity
    long long n_pods;
    int n_nodes, npad, capacity, uniform_strategy, pods_per_cta;
    const double *sorted_ta, *sorted_tb;
    const uint2 *ab;                   // [3][npad], best-first
    const uint32_t *order;             // [3][npad]
    const double *score;               // [3][n]
    const SnapHeader *hdr;
    long long *out_idx;
    double *out_score;
};

// #{i : arr[i] < x} in an ascending array
template <typename Ptr>
__device__ __forceinline__ int lower_bound_lt(Ptr arr, int n, double x)
{
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (arr[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

template <int THREADS, int R>
__global__ void __launch_bounds__(THREADS)
place_kernel(const PlaceParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint2 *tile0 = reinterpret_cast<uint2 *>(smem_raw);
    uint2 *tile1 = tile0 + kTileNodes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(tile1 + kTileNodes);

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
    const int ntiles = (nvalid + kTileNodes - 1) / kTileNodes;
    const uint2 *ab = p.ab + (size_t)s * p.npad;

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }

    // ---- pod prologue: requests -> threshold ranks (exact, see file header) ----
    int ra[R], rb[R], best[R];
    long long pod[R];
    const bool small_n = p.capacity && p.n_nodes <= kTileNodes;
    if (small_n) {      // stage both sorted threshold arrays in the (still unused) tile buffers
        double *sa = reinterpret_cast<double *>(tile0);
        double *sb = reinterpret_cast<double *>(tile1);
        for (int i = tid; i < p.n_nodes; i += THREADS) { sa[i] = p.sorted_ta[i]; sb[i] = p.sorted_tb[i]; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long long slot = slot0 + (long long)r * THREADS + tid;
        const bool live = slot - gbegin < gcount;
        pod[r] = live ? (p.perm ? (long long)p.perm[slot] : slot) : -1;
        best[r] = -1;
        ra[r] = rb[r] = live ? 0 : 0x7fffffff;
        if (live && p.capacity) {
            const double a = req_cores(p.req_cpu[pod[r]]);
            const double b = req_gib(p.req_mem[pod[r]]);
            if (small_n) {
                ra[r] = lower_bound_lt(reinterpret_cast<const double *>(tile0), p.n_nodes, a);
                rb[r] = lower_bound_lt(reinterpret_cast<const double *>(tile1), p.n_nodes, b);
            } else {
                ra[r] = lower_bound_lt(p.sorted_ta, p.n_nodes, a);
                rb[r] = lower_bound_lt(p.sorted_tb, p.n_nodes, b);
            }
        }
    }
    if (small_n) {
        // generic-proxy accesses to the buffers are done; order them before the async-proxy writes
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    // ---- main loop: tiles worst -> best, nodes within a tile worst -> best ----
    auto tile_count = [&](int t) { return min(kTileNodes, ((nvalid + 3) & ~3) - t * kTileNodes); };
    if (tid == 0 && ntiles > 0) {
        const int t = ntiles - 1;
        const uint32_t bytes = (uint32_t)tile_count(t) * sizeof(uint2);
        mbar_expect_tx(&bars[0], bytes);
        bulk_load(tile0, ab + (size_t)t * kTileNodes, bytes, &bars[0]);
    }
    int it = 0;
    for (int t = ntiles - 1; t >= 0; --t, ++it) {
        const int buf = it & 1;
        if (tid == 0 && t > 0) {
            const uint32_t bytes = (uint32_t)tile_count(t - 1) * sizeof(uint2);
            mbar_expect_tx(&bars[buf ^ 1], bytes);
            bulk_load(buf ? tile0 : tile1, ab + (size_t)(t - 1) * kTileNodes, bytes, &bars[buf ^ 1]);
        }
        mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
        const uint4 *tp = reinterpret_cast<const uint4 *>(buf ? tile1 : tile0);
        const int base = t * kTileNodes;
        const int npair = tile_count(t) >> 1;
#pragma unroll 2
        for (int i = npair - 1; i >= 0; --i) {
            const uint4 v = tp[i];                       // two nodes, same address in every lane
            const int p1 = base + 2 * i + 1, p0 = base + 2 * i;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (ra[r] <= (int)v.z && rb[r] <= (int)v.w) best[r] = p1;
                if (ra[r] <= (int)v.x && rb[r] <= (int)v.y) best[r] = p0;
            }
        }
        __syncthreads();                                 // buffer free for the next prefetch
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
    constexpr size_t smem = 2 * kTileNodes * sizeof(uint2) + 2 * sizeof(uint64_t);
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
    uint32_t *less = reinterpret_cast<uint32_t *>(w + L.off_less);
    double *score = reinterpret_cast<double *>(w + L.off_score);
    double *sorted_ta = reinterpret_cast<double *>(w + L.off_sorted_ta);
    double *sorted_tb = reinterpret_cast<double *>(w + L.off_sorted_tb);
    uint32_t *order = reinterpret_cast<uint32_t *>(w + L.off_order);
    uint2 *ab = reinterpret_cast<uint2 *>(w + L.off_ab);
    uint32_t *perm = per_pod ? reinterpret_cast<uint32_t *>(w + L.off_perm) : nullptr;

    ProfileEvents &pe = profile_events();
    if (pe.enabled) KP_CUDA(cudaEventRecord(pe.e0, stream));
    KP_CUDA(cudaMemsetAsync(w, 0, L.zero_bytes, stream));
    KP_CUDA(cudaMemsetAsync(ab, 0xff, sizeof(uint2) * kMaxStrategies * L.npad, stream));
    const int nb = (int)((n + 255) / 256);
    derive_kernel<<<nb, 256, 0, stream>>>(t, strat_mask, 1, keys, score, hdr);
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
        rank_kernel<<<dim3(nb, chunks, list.count), 256, 0, stream>>>(keys, pos, less, (int)n,
                                                                        chunk_len, list);
        KP_LAUNCH_CHECK("rank_kernel");
    }
    scatter_kernel<<<nb, 256, 0, stream>>>(keys, pos, less, (int)n, (int)L.npad, strat_mask, 1,
                                           sorted_ta, sorted_tb, order, ab);
    KP_LAUNCH_CHECK("scatter_kernel");

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
    pp.capacity = 1;
    pp.uniform_strategy = uniform;
    pp.pods_per_cta = cfg;
    pp.sorted_ta = sorted_ta;
    pp.sorted_tb = sorted_tb;
    pp.ab = ab;
    pp.order = order;
    pp.score = score;
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
// module: every GPU that loads the library carries these 1.5 MiB until the process exits.
//
// The array's name is only a host-side handle in host code.  The device address is
// looked up with cudaGetSymbolAddress, once per device, while that device is current.
constexpr int64_t kNoAllocMaxNodes = 8192;
constexpr int64_t kNoAllocMaxPodsPerPodStrategy = 65536;
constexpr size_t kNoAllocScratchBytes = 3u << 19;          // 1.5 MiB >= layout(8192 nodes, 65536 pods, per-pod)
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
    // the array keeps the previous call's contents; snapshot_place_device re-zeroes the
    // counters / rank accumulators and re-fills the padding of the rank table on every call
    return snapshot_place_device(nodes, pods, strategy, filter_mode, out_idx, out_score,
                                 scratch_addr[dev], kNoAllocScratchBytes, stream);
}

}  // namespace kp

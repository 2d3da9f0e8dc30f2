// Snapshot placement: P independent fallback decisions against one node table.
//
// Replaces one `_fallback_decision` call per pod (scheduler.py:521-559).  The reference
// recomputes every node's score for every pod; here everything that depends on the node
// alone is hoisted into a per-snapshot "node plan" (kp_plan.cuh), and the P x N part that
// remains is a pure compare/select loop over shared memory (kp_place.cuh) that never
// touches HBM and never materialises the P x N matrix.
//
// KP_FILTER_READY (the reference's own filter) is pod-independent: every pod of a strategy
// gets the same node, so that mode is an argmax over nodes plus a fill (this file).
#include <algorithm>
#include <mutex>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "kp_common.cuh"

namespace kp {

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

}  // namespace kp

#include "kp_place.cuh"
#include "kp_plan.cuh"

namespace kp {

// ---------------------------------------------------------------------------------------
// workspace layout
// ---------------------------------------------------------------------------------------
constexpr int kArgmaxMaxBlocks = 256;
constexpr int kSmallPad = 1536;
constexpr int64_t kSmallMaxPods = 1 << 19;
constexpr int kMaxChunkPods = 2048;
constexpr int kProbeMax = 64;                 // first-feasible probe: at most this many best-first nodes per strategy
constexpr size_t kProbeIntBytes = sizeof(long long) * (2 + 2 * 3 * kProbeMax);     // struct ProbeInt (defined with the grouping pass)

struct SnapLayout {
    int64_t n, p;
    int64_t npad;        // n rounded up to a multiple of 4, plus slack for 16-byte copies
    int64_t slots;       // pod slots incl. per-strategy chunk padding
    int64_t chunks;      // pod chunks at the smallest chunk size (32)
    size_t off_header, off_chunk_done, chunk_done_bytes;      // both zeroed by the large-table plan; the per-pod arrays
                                                              // (chunk_done, perm, best) sit BEHIND the plan, whose offsets
                                                              // therefore depend on the node count alone
    size_t off_keys, off_score, off_score_pos, off_uniq, off_order, off_ab, off_packed, off_partial, off_probe, off_probe_int, off_perm;
    size_t off_pos, off_kout, off_vout, off_iota, off_sorted, off_flag, off_incl, off_sort_tmp, off_best;   // large tables
    size_t sort_tmp_bytes, total;
};

static SnapLayout make_layout(int64_t n, int64_t p, bool per_pod)
{
    SnapLayout L{};
    L.n = n;
    L.p = p;
    // small tables: place_small_kernel loads whole register slices without bounds checks, up to 256 entries
    // past the table per slice (kSmallPad covers five slices); the plan fills the padding with entries that fit nothing
    L.npad = (n + 3) / 4 * 4 + 4 + (n > kSmallMaxNodes ? 0 : kSmallPad);
    L.slots = p + 3 * kMaxChunkPods;
    L.chunks = L.slots / 32 + 1;
    const bool big = n > kSmallMaxNodes;                 // multi-kernel plan, node segments possible
    size_t o = 0;
    L.off_header = o;     o = align_up(o + sizeof(SnapHeader));
    L.off_keys = o;    o = align_up(o + sizeof(uint64_t) * kKeyArrays * n);
    L.off_score = o;   o = align_up(o + sizeof(double) * kMaxStrategies * n);
    L.off_score_pos = o; o = align_up(o + sizeof(double) * kMaxStrategies * L.npad);
    L.off_uniq = o;    o = align_up(o + sizeof(double) * 2 * L.npad);
    L.off_order = o;   o = align_up(o + sizeof(uint32_t) * kMaxStrategies * L.npad);
    L.off_ab = o;      o = align_up(o + sizeof(uint2) * kMaxStrategies * L.npad);
    L.off_packed = o;  o = align_up(o + sizeof(uint32_t) * kMaxStrategies * L.npad);
    L.off_partial = o; o = align_up(o + (sizeof(uint64_t) + sizeof(uint32_t) * 2) * kMaxStrategies * kArgmaxMaxBlocks);
    L.off_probe = o;   o = align_up(o + sizeof(double2) * (kMaxStrategies * kProbeMax + 1));      // + the largest thresholds
    L.off_probe_int = o; o = align_up(o + kProbeIntBytes);
    const int64_t nb = big ? n : 0, npb = big ? L.npad : 0;
    L.sort_tmp_bytes = big ? align_up(sort_temp_bytes(n)) : 0;            // per sorted array (they run concurrently)
    L.off_pos = o;      o = align_up(o + sizeof(uint32_t) * kKeyArrays * nb);
    L.off_kout = o;     o = align_up(o + sizeof(uint64_t) * kKeyArrays * nb);
    L.off_vout = o;     o = align_up(o + sizeof(uint32_t) * kKeyArrays * nb);
    L.off_iota = o;     o = align_up(o + sizeof(uint32_t) * nb);
    L.off_sorted = o;   o = align_up(o + sizeof(double) * 2 * npb);
    L.off_flag = o;     o = align_up(o + sizeof(uint32_t) * 2 * npb);
    L.off_incl = o;     o = align_up(o + sizeof(uint32_t) * 2 * npb);
    L.off_sort_tmp = o; o = align_up(o + L.sort_tmp_bytes * (kMaxStrategies + 2));
    L.off_chunk_done = o; L.chunk_done_bytes = sizeof(uint32_t) * (big ? L.chunks : 0); o = align_up(o + L.chunk_done_bytes);
    L.off_perm = o;     o = align_up(o + sizeof(uint32_t) * (per_pod ? L.slots : 0));
    L.off_best = o;     o = align_up(o + sizeof(uint32_t) * (big ? L.slots : 0));
    L.total = o;
    return L;
}

size_t snapshot_workspace_bytes(int64_t n_nodes, int64_t n_pods, bool per_pod)
{
    if (n_nodes < 0 || n_pods < 0) return 0;
    return make_layout(n_nodes, n_pods, per_pod).total;
}

// True when the snapshot reads every pod row exactly once and writes every result row exactly
// once, in pod order (one node segment, no strategy grouping): the host session may then hand
// the kernels pinned host memory directly instead of staging copies.
bool snapshot_single_pass(int64_t n_nodes, bool per_pod)
{
    return n_nodes <= kSmallMaxNodes && !per_pod;
}

// What a node plan depends on besides the node table: which strategy table was built and the filter (the plan's
// place in the workspace depends on the node count alone).  Plans with per-pod strategies are not reused: their
// header carries per-call pod group counters.
unsigned long long snapshot_plan_signature(int64_t n_nodes, int64_t n_pods, bool per_pod, int strategy, int filter_mode)
{
    if (per_pod || filter_mode != KP_FILTER_CAPACITY || n_nodes <= 0) return 0ull;
    const SnapLayout L = make_layout(n_nodes, n_pods, false);
    return 0x8000000000000000ull | ((unsigned long long)canon_strategy(strategy) << 60) |
           ((unsigned long long)(L.off_keys & 0xfffffffffffull) << 16) | (unsigned long long)(n_nodes & 0xffff);
}

PlanView snapshot_plan_view(int64_t n_nodes, int64_t n_pods, bool per_pod, void *workspace)
{
    const SnapLayout L = make_layout(n_nodes, n_pods, per_pod);
    unsigned char *w = static_cast<unsigned char *>(workspace);
    return PlanView{reinterpret_cast<const SnapHeader *>(w + L.off_header)->nvalid, reinterpret_cast<const double *>(w + L.off_score_pos),
                    (int)L.npad};
}

// pods per strategy that the last grouped call (per-pod strategies, or the first-feasible probe) left for the table scan
int snapshot_scanned_pods(const void *workspace, cudaStream_t stream, int64_t *out_scanned)
{
    SnapHeader h;
    KP_CUDA(cudaMemcpyAsync(&h, workspace, sizeof(SnapHeader), cudaMemcpyDeviceToHost, stream));
    KP_CUDA(cudaStreamSynchronize(stream));
    for (int s = 0; s < kMaxStrategies; ++s) out_scanned[s] = (int64_t)h.group_count[s];
    return KP_OK;
}

static NodePtrs node_ptrs(const kp_node_table &t)
{
    return NodePtrs{(const long long *)t.alloc_cpu_milli, (const long long *)t.alloc_mem_bytes,
                    (const long long *)t.pod_count, (const long long *)t.max_pods, t.ready,
                    t.cpu_usage_pct, t.mem_usage_pct, (int)t.n_nodes};
}

// ---------------------------------------------------------------------------------------
// per-pod strategy: group pods by strategy so that a CTA scans one best-first table
// ---------------------------------------------------------------------------------------
// Pods grouped by strategy.  A block owns kGroupItems consecutive pods, aggregates in shared memory
// and touches the three global counters once: with one global atomic per warp the two kernels took
// 120 us each for 2 M pods (serialised on three addresses); the data itself is 16 MB = 3 us of HBM.
constexpr int kGroupPerThread = 8;
constexpr int kGroupItems = 256 * kGroupPerThread;

// Pods that exceed the LARGEST free-cpu or free-memory threshold of the table fit no node whatever the strategy
// (their threshold rank is past every node's): they are answered here ("no node") and never enter a strategy
// group, so the placement kernel does not scan the table for them.  Exact: the test is the kernel's own
// request > threshold compare against the last entry of the sorted threshold arrays (which hold every node's
// thresholds, NaN parked at +inf -- a table with such a node simply rejects nobody early).
//
// First-feasible probe (PlaceTuning::prune_depth > 0).  The best-first table is sorted by (score desc, node index asc),
// so the node a pod gets is the FIRST table entry it fits (scheduler.py:547: strict '>' keeps the earlier of equal
// scores).  The grouping pass therefore tries the first prune_depth entries of the pod's strategy table directly --
// the thresholds of those few nodes, prepared by probe_table_kernel -- and a pod that fits one of them is answered on
// the spot; only the rest enter a strategy group and are scanned against the whole table by the placement kernel.
// The result is the same bit for bit; the number of pairs EVALUATED drops from pods x nodes to roughly pods x (depth
// of the first fit), which is why this is an option and not the measured roofline kernel (SURVEY.md section 7,
// "algorithmic honesty": evaluated and effective pairs are reported separately wherever it is used).
//
// Integer images of the fp64 thresholds the grouping pass compares against.  A request is an int64 (milli-cores,
// bytes) and the reference compares fl(milli / 1000.0) resp. bytes * 2^-30 with a node's free capacity T; both maps are
// monotone, so "request fits" is "request <= M" for M = the LARGEST integer whose image is <= T.  M is found once per
// threshold (limit_cpu / limit_mem below, exact: the candidates around floor(T * 1000) are tested with the very
// division the pod side would do), and the per-pod work of the pass is integer compares: no fp64 division, no
// conversions.  The images cover requests with |x| < 2^53 (where (double)x is exact -- the ABI's stated range); a
// request outside takes the fp64 compares.
constexpr long long kExactMax = (1ll << 53) - 1;
struct ProbeInt {
    long long dead_cpu, dead_mem;                       // a request above either fits no node of the table
    long long cpu[kMaxStrategies][kProbeMax], mem[kMaxStrategies][kProbeMax];
};

static_assert(sizeof(ProbeInt) == kProbeIntBytes, "workspace layout reserves exactly one ProbeInt");

// largest m in [-kExactMax, kExactMax] with fl(m / 1000.0) <= T  (-kExactMax - 1: none).  No NaN reaches this.
__device__ long long limit_cpu(double T)
{
    if (req_cores(kExactMax) <= T) return kExactMax;
    if (!(req_cores(-kExactMax) <= T)) return -kExactMax - 1;
    // m / 1000 <= T  <=>  m <= 1000 T in real arithmetic; the rounded division admits at most one more m, the rounded
    // product moves the floor by at most one: the answer is within [g - 1, g + 2]
    const double g = floor(__dmul_rn(T, 1000.0));
    long long c = (long long)g + 2;
    c = c > kExactMax ? kExactMax : c;
    while (c > -kExactMax && req_cores(c) > T) --c;
    return c;
}
// largest b in [-kExactMax, kExactMax] with b * 2^-30 <= T: the scaling is exact, so b <= floor(T * 2^30)
__device__ long long limit_mem(double T)
{
    const double g = floor(__dmul_rn(T, 1073741824.0));            // exact unless it overflows to +-inf
    if (g >= (double)kExactMax) return kExactMax;
    if (g < -(double)kExactMax) return -kExactMax - 1;
    return (long long)g;
}

struct GroupPods {
    const long long *sid, *req_cpu, *req_mem;      // sid == NULL: every pod uses `uniform`
    int uniform;
    long long n_pods;
    const double *uniq;            // [2][npad] sorted thresholds
    int npad, small_n;             // small_n > 0: arrays keep duplicates and hold small_n entries; else hdr->distinct
    long long *out_idx;
    double *out_score;
    uint32_t *out_gpos;            // node-sharded form (may be NULL)
    const double2 *probe;          // [3][kProbeMax] (free cpu, free mem) of the best-first nodes
    const ProbeInt *lim;           // integer images of the probe table and of the largest thresholds
    int probe_depth;               // 0 = no probe
    const uint32_t *order;         // [3][npad] node index by best-first position
    const double *score_pos;       // [3][npad]
    const uint32_t *gpos_by_pos;   // node-sharded form: global position by local position
    size_t gpos_stride;
};
constexpr int kGroupDead = kMaxStrategies;
constexpr int kGroupDone = kMaxStrategies + 1;       // answered by the probe: `found` holds the table position

__device__ __forceinline__ void group_limits(const GroupPods &g, const SnapHeader *hdr, double &ta_max, double &tb_max)
{
    const int da = g.small_n > 0 ? g.small_n : (int)hdr->distinct[0], db = g.small_n > 0 ? g.small_n : (int)hdr->distinct[1];
    const double ninf = __longlong_as_double((long long)0xfff0000000000000ull);
    ta_max = da > 0 ? g.uniq[da - 1] : ninf;
    tb_max = db > 0 ? g.uniq[(size_t)g.npad + db - 1] : ninf;
}
// what every pod of a thread is compared with first, loaded once per thread: the table's largest thresholds and the
// first probed node of every strategy (most pods are settled by these; deeper probe entries are read on demand)
struct GroupHot {
    long long dead_cpu, dead_mem;
    long long cpu0[kMaxStrategies], mem0[kMaxStrategies];
};
__device__ __forceinline__ GroupHot group_hot(const GroupPods &g)
{
    GroupHot h;
    h.dead_cpu = __ldg(&g.lim->dead_cpu);
    h.dead_mem = __ldg(&g.lim->dead_mem);
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        // no probe: an entry nothing fits
        h.cpu0[s] = g.probe_depth > 0 ? __ldg(&g.lim->cpu[s][0]) : -kExactMax - 1;
        h.mem0[s] = g.probe_depth > 0 ? __ldg(&g.lim->mem[s][0]) : -kExactMax - 1;
    }
    return h;
}
// qc / qm: the pod's requests as they lie in the table; s: its canonical strategy
__device__ __forceinline__ int group_of(const GroupPods &g, const GroupHot &h, long long qc, long long qm, int s, int &found)
{
    const ProbeInt *lim = g.lim;
    const bool exact = (unsigned long long)qc + (unsigned long long)kExactMax <= 2ull * kExactMax &&
                       (unsigned long long)qm + (unsigned long long)kExactMax <= 2ull * kExactMax;
    if (exact) {
        if (qc > h.dead_cpu || qm > h.dead_mem) return kGroupDead;
        const long long c0 = s == 0 ? h.cpu0[0] : (s == 1 ? h.cpu0[1] : h.cpu0[2]);
        const long long m0 = s == 0 ? h.mem0[0] : (s == 1 ? h.mem0[1] : h.mem0[2]);
        if (qc <= c0 && qm <= m0) { found = 0; return kGroupDone; }
        for (int k = 1; k < g.probe_depth; ++k)                    // entries past the table fit nothing
            if (qc <= __ldg(&lim->cpu[s][k]) && qm <= __ldg(&lim->mem[s][k])) { found = k; return kGroupDone; }
        return s;
    }
    // a request beyond 2^53: (double)x rounds, the integer images do not apply -- the fp64 compares themselves
    double ta_max, tb_max;
    {
        const double2 d = __ldg(g.probe + kMaxStrategies * kProbeMax);
        ta_max = d.x;
        tb_max = d.y;
    }
    const double rc = req_cores(qc), rm = req_gib(qm);
    if (rc > ta_max || rm > tb_max) return kGroupDead;
    const double2 *pt = g.probe + s * kProbeMax;
    for (int k = 0; k < g.probe_depth; ++k) {
        const double2 th = __ldg(pt + k);
        if (rc <= th.x && rm <= th.y) { found = k; return kGroupDone; }
    }
    return s;
}

// After the node plan, one CTA: (free cpu, free mem) of the first kProbeMax best-first nodes of every planned strategy,
// the largest thresholds of the table (entry [3][0] of the fp64 table), and the integer images of all of them.
__global__ void __launch_bounds__(kMaxStrategies * kProbeMax)
probe_table_kernel(NodePtrs t, int strat_mask, const SnapHeader *hdr, GroupPods g, double2 *__restrict__ probe, ProbeInt *__restrict__ lim)
{
    const int s = threadIdx.x / kProbeMax, k = threadIdx.x % kProbeMax;
    const double ninf = __longlong_as_double((long long)0xfff0000000000000ull);
    double2 v = make_double2(ninf, ninf);
    if ((strat_mask & (1 << s)) && k < g.probe_depth && k < (int)hdr->nvalid[s]) {
        const uint32_t n = g.order[(size_t)s * g.npad + k];
        v.x = free_cpu_cores(t.alloc_cpu[n], t.cu[n]);
        v.y = free_mem_gib(t.alloc_mem[n], t.mu[n]);
    }
    probe[threadIdx.x] = v;
    lim->cpu[s][k] = limit_cpu(v.x);
    lim->mem[s][k] = limit_mem(v.y);
    if (threadIdx.x == 0) {
        double ta_max, tb_max;
        group_limits(g, hdr, ta_max, tb_max);
        probe[kMaxStrategies * kProbeMax] = make_double2(ta_max, tb_max);
        lim->dead_cpu = limit_cpu(ta_max);
        lim->dead_mem = limit_mem(tb_max);
    }
}

// one block's pods, loaded before anything is computed (sixteen to twenty-four independent loads per thread in flight)
struct GroupRows { long long qc[kGroupPerThread], qm[kGroupPerThread]; int s[kGroupPerThread]; };
__device__ __forceinline__ void group_load(const GroupPods &gp, long long base, GroupRows &r)
{
#pragma unroll
    for (int i = 0; i < kGroupPerThread; ++i) {
        const long long j = base + i * 256 + threadIdx.x;
        const bool in = j < gp.n_pods;
        r.qc[i] = in ? gp.req_cpu[j] : 0;
        r.qm[i] = in ? gp.req_mem[j] : 0;
        r.s[i] = in ? (gp.sid ? canon_strategy(gp.sid[j]) : gp.uniform) : -1;
    }
}

__global__ void __launch_bounds__(256)
group_count_kernel(GroupPods gp, SnapHeader *hdr)
{
    __shared__ unsigned s_cnt[kMaxStrategies];
    if (threadIdx.x < kMaxStrategies) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    GroupRows r;
    group_load(gp, (long long)blockIdx.x * kGroupItems, r);
    const GroupHot hot = group_hot(gp);
    unsigned c[kMaxStrategies] = {0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < kGroupPerThread; ++i) {
        if (r.s[i] >= 0) {
            int found;
            const int g = group_of(gp, hot, r.qc[i], r.qm[i], r.s[i], found);
#pragma unroll
            for (int s = 0; s < kMaxStrategies; ++s) c[s] += g == s;
        }
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        const unsigned w = __reduce_add_sync(0xffffffffu, c[s]);
        if ((threadIdx.x & 31) == 0 && w) atomicAdd(&s_cnt[s], w);
    }
    __syncthreads();
    if (threadIdx.x < kMaxStrategies && s_cnt[threadIdx.x]) atomicAdd(&hdr->group_count[threadIdx.x], s_cnt[threadIdx.x]);
}
__global__ void __launch_bounds__(256)
group_assign_kernel(GroupPods gp, int chunk_pods, SnapHeader *hdr, uint32_t *__restrict__ perm)
{
    __shared__ unsigned s_cnt[kMaxStrategies], s_cur[kMaxStrategies];
    __shared__ long long s_base[kMaxStrategies];
    if (threadIdx.x < kMaxStrategies) { s_cnt[threadIdx.x] = 0u; s_cur[threadIdx.x] = 0u; }
    __syncthreads();
    const long long base = (long long)blockIdx.x * kGroupItems;
    const int lane = threadIdx.x & 31;
    GroupRows r;
    group_load(gp, base, r);
    const GroupHot hot = group_hot(gp);
    int g[kGroupPerThread];
    unsigned c[kMaxStrategies] = {0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < kGroupPerThread; ++i) {
        const long long j = base + i * 256 + threadIdx.x;
        int found = 0;
        g[i] = r.s[i] >= 0 ? group_of(gp, hot, r.qc[i], r.qm[i], r.s[i], found) : -1;
        if (g[i] == kGroupDead) {                                 // fits nothing: answered here, never scanned
            gp.out_idx[j] = -1;
            gp.out_score[j] = -1.0;
            if (gp.out_gpos) gp.out_gpos[j] = 0xffffffffu;
        } else if (g[i] == kGroupDone) {                          // fits one of the first best-first nodes: that node is the answer
            const int s = r.s[i];
            const size_t at = (size_t)s * gp.npad + found;
            gp.out_idx[j] = (long long)gp.order[at];
            gp.out_score[j] = gp.score_pos[at];
            if (gp.out_gpos) gp.out_gpos[j] = gp.gpos_by_pos[(size_t)s * gp.gpos_stride + found];
        }
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) c[s] += g[i] == s;
    }
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        const unsigned w = __reduce_add_sync(0xffffffffu, c[s]);
        if (lane == 0 && w) atomicAdd(&s_cnt[s], w);
    }
    __syncthreads();
    if (threadIdx.x == 0) {                                   // this block's range inside every group
        long long begin[kMaxStrategies];
        group_begins(hdr, chunk_pods, begin);
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s)
            s_base[s] = begin[s] + (s_cnt[s] ? (long long)atomicAdd(&hdr->group_cursor[s], s_cnt[s]) : 0ll);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kGroupPerThread; ++i) {
        const long long j = base + i * 256 + threadIdx.x;
#pragma unroll
        for (int s = 0; s < kMaxStrategies; ++s) {
            const unsigned b = __ballot_sync(0xffffffffu, g[i] == s);
            if (!b) continue;
            unsigned off = 0;
            const int leader = __ffs(b) - 1;
            if (lane == leader) off = atomicAdd(&s_cur[s], (unsigned)__popc(b));
            off = __shfl_sync(0xffffffffu, off, leader);
            if (g[i] == s) perm[s_base[s] + off + __popc(b & ((1u << lane) - 1))] = (uint32_t)j;
        }
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
// place_kernel configurations, largest pods-per-CTA first
struct PlaceCfg { int threads, r, nsplit, chunk; };
// (index, threads, pods per lane, node splits); pods per CTA = 32 * R * warps / NSPLIT
#define KP_PLACE_CFG_LIST(X) \
    X(0, 128, 16, 1) X(1, 128, 8, 1) X(2, 128, 4, 1) X(3, 256, 4, 4) X(4, 256, 4, 8) X(5, 256, 2, 8) \
    X(6, 256, 8, 8) /* not in the size ladder: few pods, one wave of 8-warp CTAs with 8 pods per lane */
// Measured on c2 and not kept (profiles/r01_place_cfg_sweep.txt): 128-thread CTAs with one or two node
// splits (fewer, longer warps: less prologue per pair) lose more to latency than they save.
#define KP_CFG_ROW(i, t, r, ns) {t, r, ns, 32 * r * (t / 32 / ns)},
constexpr PlaceCfg kPlaceCfgs[] = {KP_PLACE_CFG_LIST(KP_CFG_ROW)};
#undef KP_CFG_ROW
constexpr int kNumPlaceCfgs = sizeof(kPlaceCfgs) / sizeof(kPlaceCfgs[0]);
constexpr int kLadderCfgs = 6;           // configurations 0..5 form the size ladder, largest pods-per-CTA first
constexpr int kFallbackCfg = 4;          // few pods: 128 pods per CTA, 8 warps splitting the nodes

template <int THREADS, int R, int NSPLIT>
static int place_occupancy()
{
    static int occ = 0;                                             // same on every device this library runs on (sm_100)
    if (occ == 0) {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, place_kernel<THREADS, R, NSPLIT>, THREADS, 0) != cudaSuccess || v <= 0)
            v = 1;
        occ = v;
    }
    return occ;
}

static int cfg_occupancy(int c)
{
    switch (c) {
#define KP_CFG_CASE(i, t, r, ns) case i: return place_occupancy<t, r, ns>();
        KP_PLACE_CFG_LIST(KP_CFG_CASE)
#undef KP_CFG_CASE
        default: return 1;
    }
}

// Programmatic dependent launch: the kernel may start while its predecessor in the stream is
// still draining; place_kernel does its plan-independent prologue first and then executes
// griddepcontrol.wait before touching anything the plan wrote.
// resident_cap > 0 limits the CTAs resident per SM (by reserving unused dynamic shared memory):
// used when the pods arrive over the host link, where several short waves pipeline the
// transfers with the scoring instead of every CTA waiting for its rows at the same time.
template <int THREADS, int R, int NSPLIT>
static cudaError_t launch_place_pdl(const PlaceParams &pp, dim3 grid, cudaStream_t stream, int resident_cap)
{
    static size_t static_smem = 0;                                  // a property of the kernel image
    static size_t max_dyn_set_dev[kMaxDevices] = {0};               // the opt-in is per device
    size_t &max_dyn_set = max_dyn_set_dev[device_slot()];
    size_t dyn = 0;
    if (resident_cap > 0) {
        if (static_smem == 0) {
            cudaFuncAttributes fa{};
            cudaError_t e = cudaFuncGetAttributes(&fa, place_kernel<THREADS, R, NSPLIT>);
            if (e != cudaSuccess) return e;
            static_smem = fa.sharedSizeBytes ? fa.sharedSizeBytes : 1;
        }
        const size_t per_sm = 228u * 1024u, reserve = 1024u, max_cta = 227u * 1024u;
        size_t want = per_sm / (size_t)resident_cap - reserve;         // static + dynamic per CTA
        if (want > max_cta) want = max_cta;
        if (want > static_smem) dyn = (want - static_smem) & ~(size_t)127;
        if (dyn > max_dyn_set) {
            cudaError_t e = cudaFuncSetAttribute(place_kernel<THREADS, R, NSPLIT>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
            if (e != cudaSuccess) return e;
            max_dyn_set = dyn;
        }
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(THREADS, 1, 1);
    cfg.dynamicSmemBytes = dyn;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, place_kernel<THREADS, R, NSPLIT>, pp);
}

static int launch_place_cfg(int c, const PlaceParams &pp, dim3 grid, cudaStream_t stream, int resident_cap)
{
    cudaError_t e = cudaErrorInvalidValue;
    switch (c) {
#define KP_CFG_CASE(i, t, r, ns) case i: e = launch_place_pdl<t, r, ns>(pp, grid, stream, resident_cap); break;
        KP_PLACE_CFG_LIST(KP_CFG_CASE)
#undef KP_CFG_CASE
        default: break;
    }
    if (e != cudaSuccess) return cuda_fail(e, "place_kernel launch");
    count_launch();
    return KP_OK;
}

// place_small_kernel (tables of the one-launch plan): NPT = node words per lane, a multiple of 8 up to 64
template <int NPT>
static cudaError_t launch_small_pdl(const PlaceParams &pp, int ctas, cudaStream_t stream)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)ctas, 1, 1);
    cfg.blockDim = dim3(kSmallThreads, 1, 1);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, place_small_kernel<NPT>, pp);
}

static int launch_place_small(const PlaceParams &pp, int64_t n, int64_t P, bool per_pod, cudaStream_t stream)
{
    const int nslices = (int)((n + 1791) / 1792);                       // at most 56 words per lane and slice
    int npt = (int)((n + 32 * nslices - 1) / (32 * nslices));
    npt = std::max(8, (npt + 7) / 8 * 8);
    const int occ = small_occupancy(npt);
    const int64_t max_ctas = (int64_t)sm_count_cached() * occ;
    // enough pods per warp to pay for loading its table slice (16), otherwise every resident warp gets a share
    const int64_t want = per_pod ? ((P + 31) / 32 + kMaxStrategies - 1 + kSmallWarps - 1) / kSmallWarps
                                 : (P + 16 * kSmallWarps - 1) / (16 * kSmallWarps);
    int ctas = (int)std::max<int64_t>(1, std::min(want, max_ctas));
    if (const char *v = getenv("KP_SMALL_CTAS")) ctas = std::max(1, atoi(v));       // experiments only
    PlaceParams pq = pp;
    const int64_t nwarps = (int64_t)ctas * kSmallWarps;
    pq.share_q = P / nwarps;
    pq.share_r = (int)(P % nwarps);
    cudaError_t e = cudaErrorInvalidValue;
    switch (npt) {
        case 8: e = launch_small_pdl<8>(pq, ctas, stream); break;
        case 16: e = launch_small_pdl<16>(pq, ctas, stream); break;
        case 24: e = launch_small_pdl<24>(pq, ctas, stream); break;
        case 32: e = launch_small_pdl<32>(pq, ctas, stream); break;
        case 40: e = launch_small_pdl<40>(pq, ctas, stream); break;
        case 48: e = launch_small_pdl<48>(pq, ctas, stream); break;
        case 56: e = launch_small_pdl<56>(pq, ctas, stream); break;
        default: break;
    }
    if (e != cudaSuccess) return cuda_fail(e, "place_small_kernel launch");
    count_launch();
    return KP_OK;
}

// Side streams for the independent radix sorts of the large-table plan (one pool per device, made on
// first use and kept for the life of the process).  NULL = could not be created: the sorts then run one
// after the other on the caller's stream.
constexpr int kSideStreams = 4;                 // up to 5 sorted arrays: 3 strategies + 2 thresholds
struct SideStreams {
    cudaStream_t s[kSideStreams];
    cudaEvent_t fork, join[kSideStreams];
    std::mutex mu;                              // the events are shared: one fork/join sequence is enqueued at a time
};
static SideStreams *side_streams()
{
    static SideStreams pool[64];
    static int state[64] = {0};                 // 0 = not tried, 1 = ready, -1 = unavailable
    static std::mutex once;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(once);                // first use may come from several host threads
    if (state[dev] == 0) {
        bool ok = getenv("KP_NO_SIDE_STREAMS") == nullptr;
        SideStreams &p = pool[dev];
        if (ok) ok = cudaEventCreateWithFlags(&p.fork, cudaEventDisableTiming) == cudaSuccess;
        for (int i = 0; ok && i < kSideStreams; ++i)
            ok = cudaStreamCreateWithFlags(&p.s[i], cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&p.join[i], cudaEventDisableTiming) == cudaSuccess;
        if (!ok) cudaGetLastError();
        state[dev] = ok ? 1 : -1;
    }
    return state[dev] == 1 ? &pool[dev] : nullptr;
}

int snapshot_place_device(const kp_node_table &nodes, const kp_pod_table &pods, int strategy,
                          int filter_mode, int64_t *out_idx, double *out_score, void *ws,
                          size_t ws_bytes, cudaStream_t stream, const PlaceTuning &tuning)
{
    const int64_t n = nodes.n_nodes, P = pods.n_pods;
    if (P == 0) return KP_OK;
    int resident_cap = tuning.resident_cap;
    unsigned flow_window = tuning.flow_window;
    if (const char *v = getenv("KP_PLACE_RESIDENT")) resident_cap = atoi(v);        // experiments only
    if (const char *v = getenv("KP_FLOW_WINDOW")) flow_window = (unsigned)atoi(v);  // experiments only
    const bool by_sid = pods.strategy_ids != nullptr;
    // first-feasible probe: pods go through the grouping pass (and its permutation) even with one strategy
    const int prune_depth = filter_mode == KP_FILTER_CAPACITY ? std::max(0, std::min(tuning.prune_depth, kProbeMax)) : 0;
    const bool per_pod = by_sid || prune_depth > 0;                // "grouped": the placement kernel reads pods through perm
    const SnapLayout L = make_layout(n, P, per_pod);
    if (!ws || ws_bytes < L.total || (reinterpret_cast<uintptr_t>(ws) & (kAlign - 1))) {
        set_error("snapshot workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", L.total,
                  ws_bytes, ws);
        return KP_ERR_WORKSPACE;
    }
    unsigned char *w = static_cast<unsigned char *>(ws);
    SnapHeader *hdr = reinterpret_cast<SnapHeader *>(w + L.off_header);
    const int uniform = per_pod ? -1 : canon_strategy(strategy);
    const int strat_mask = by_sid ? 7 : (1 << canon_strategy(strategy));
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

    // ---- work decomposition for place_kernel ----
    // Enough CTAs that the hardware scheduler can balance them (>= ~8 per resident slot): first
    // by cutting the node table into segments (grid.y; needs the zeroed merge buffers of the
    // multi-kernel plan), then by carrying fewer pods per CTA.
    const int sms = sm_count_cached();
    const bool big = n > kSmallMaxNodes;
    const int ntiles_wide = (int)((L.npad + kTileNodesWide - 1) / kTileNodesWide);
    const int max_nseg = big ? ntiles_wide : 1;
    int cfg = kFallbackCfg;
    // The top of the ladder carries 16 pods per lane: the per-iteration overhead of the inner loop (one LDS.128,
    // four position adds, loop control: 11 instructions) is then spread over 64 pairs instead of 32, and 16
    // independent compare chains per lane need only 3 warps per scheduler (profiles/r01_place_cfg_sweep.txt).
    for (int c = 0; c < kLadderCfgs; ++c) {
        const int64_t chunks = (P + kPlaceCfgs[c].chunk - 1) / kPlaceCfgs[c].chunk;
        if (chunks * max_nseg >= (int64_t)4 * sms * cfg_occupancy(c)) { cfg = c; break; }
    }
    if (const char *force = getenv("KP_FORCE_PLACE_CFG")) {      // experiments only
        const int c = atoi(force);
        if (c >= 0 && c < kNumPlaceCfgs) cfg = c;
    }
    // small tables: the nodes-in-registers kernel (kp_place.cuh); its per-pod strategy groups are 32 slots wide
    // (pinned host pod rows read in place keep the paced place_kernel: the order in which the rows cross the
    // host link matters more there than the instruction count)
    // and large pod counts / per-pod strategies keep it too: with thousands of pods per warp its fixed costs vanish
    // and its tile loop is a little tighter (profiles/r02_slice_probe.txt)
    bool use_small = !big && !per_pod && P <= kSmallMaxPods && !(tuning.flow && flow_window > 0);
    if (const char *v = getenv("KP_SMALL_OLD")) if (v[0] == '1') use_small = false;   // experiments only
    const int chunk_pods = use_small ? 32 : kPlaceCfgs[cfg].chunk;
    const int64_t n_chunks = (P + chunk_pods - 1) / chunk_pods + (per_pod ? kMaxStrategies - 1 : 0);
    int nseg = 1, tiles_per_seg = ntiles_wide;
    if (max_nseg > 1) {
        const int64_t slots = (int64_t)sms * cfg_occupancy(cfg);
        int64_t want = std::max<int64_t>(1, (8 * slots + n_chunks - 1) / n_chunks);
        want = std::min<int64_t>(want, max_nseg);
        tiles_per_seg = (int)((ntiles_wide + want - 1) / want);
        tiles_per_seg += tiles_per_seg & 1;                      // even: packed tiles are two wide tiles
        nseg = (ntiles_wide + tiles_per_seg - 1) / tiles_per_seg;
    }
    if (n_chunks > 0x7fffffff || nseg > 65535) {
        set_error("snapshot too large for one launch (%lld pod chunks, %d node segments)", (long long)n_chunks, nseg);
        return KP_ERR_INVALID_ARG;
    }

    // ---- node plan ----
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(w + L.off_keys);
    uint32_t *pos = reinterpret_cast<uint32_t *>(w + L.off_pos);
    double *sorted = reinterpret_cast<double *>(w + L.off_sorted);
    uint32_t *perm = per_pod ? reinterpret_cast<uint32_t *>(w + L.off_perm) : nullptr;
    uint32_t *best_pos = reinterpret_cast<uint32_t *>(w + L.off_best);
    PlanOut po{};
    po.hdr = hdr;
    po.score = reinterpret_cast<double *>(w + L.off_score);
    po.uniq = reinterpret_cast<double *>(w + L.off_uniq);
    po.order = reinterpret_cast<uint32_t *>(w + L.off_order);
    po.score_pos = reinterpret_cast<double *>(w + L.off_score_pos);
    po.ab = reinterpret_cast<uint2 *>(w + L.off_ab);
    po.packed = reinterpret_cast<uint32_t *>(w + L.off_packed);
    po.npad = (int)L.npad;

    ProfileEvents &pe = profile_events();
    if (pe.enabled) KP_CUDA(cudaEventRecord(pe.e0, stream));
    const unsigned long long sig = snapshot_plan_signature(n, P, by_sid, strategy, filter_mode);
    if (tuning.plan_sig) *tuning.plan_sig = sig;
    if (tuning.reuse_plan && (sig != 0ull || tuning.reuse_plan_forced)) {
        // the workspace holds this table's plan: only the per-call merge counters of the segmented form are reset
        if (L.chunk_done_bytes) KP_CUDA(cudaMemsetAsync(w + L.off_chunk_done, 0, L.chunk_done_bytes, stream));
        // (and the group counters of the probe's grouping pass, which a fresh plan leaves zeroed)
        if (per_pod) KP_CUDA(cudaMemsetAsync(hdr->group_count, 0, sizeof(hdr->group_count) + sizeof(hdr->group_cursor), stream));
    } else if (!big) {
        // no sort, no memset, one launch: every node counts its own sorted positions
        const int groups = (int)((n + kCountNodes - 1) / kCountNodes);
        plan_count_kernel<<<groups * kCountCluster, kCountThreads, 0, stream>>>(t, strat_mask, po);
        KP_LAUNCH_CHECK("plan_count_kernel");
    } else {
        KP_CUDA(cudaMemsetAsync(hdr, 0, sizeof(SnapHeader), stream));
        if (L.chunk_done_bytes) KP_CUDA(cudaMemsetAsync(w + L.off_chunk_done, 0, L.chunk_done_bytes, stream));
        const int nb = (int)((n + 255) / 256);
        uint32_t *iota = reinterpret_cast<uint32_t *>(w + L.off_iota);
        unsigned long long *kout = reinterpret_cast<unsigned long long *>(w + L.off_kout);
        uint32_t *vout = reinterpret_cast<uint32_t *>(w + L.off_vout);
        uint32_t *flag = reinterpret_cast<uint32_t *>(w + L.off_flag);
        uint32_t *incl = reinterpret_cast<uint32_t *>(w + L.off_incl);
        void *tmp = w + L.off_sort_tmp;
        derive_kernel<true><<<nb, 256, 0, stream>>>(t, strat_mask, keys, po.score, hdr, iota);
        KP_LAUNCH_CHECK("derive_kernel");
        SortList list{};
        for (int s = 0; s < kMaxStrategies; ++s) if (strat_mask & (1 << s)) list.idx[list.count++] = s;
        list.idx[list.count++] = 3;
        list.idx[list.count++] = 4;
        if (n <= kRankSortMax) {
            // every key array in two launches of our own (kp_ranksort.cu); the radix sort of the library needs ten
            // latency-bound launches per array at these sizes
            RankSortArrays which{};
            for (int q = 0; q < list.count; ++q) which.idx[which.count++] = list.idx[q];
            const int st = rank_sort_pairs_u64(tmp, L.sort_tmp_bytes * (kMaxStrategies + 2), keys, (size_t)n, which, (int)n, kout, vout, (size_t)n,
                                               stream);
            if (st != KP_OK) return st;
        } else {
        // the sorts are independent and latency-bound (tens of thousands of keys, ~9 launches each):
        // all but the first run on side streams, forked from and joined back into the caller's stream
        SideStreams *side = side_streams();
        std::unique_lock<std::mutex> side_lock;
        if (side) {
            side_lock = std::unique_lock<std::mutex>(side->mu);
            KP_CUDA(cudaEventRecord(side->fork, stream));
        }
        for (int q = 0; q < list.count; ++q) {
            const int j = list.idx[q];
            cudaStream_t sq = (side && q > 0) ? side->s[q - 1] : stream;
            if (sq != stream) KP_CUDA(cudaStreamWaitEvent(sq, side->fork, 0));
            const int st = sort_pairs_u64(static_cast<unsigned char *>(tmp) + (size_t)q * L.sort_tmp_bytes, L.sort_tmp_bytes,
                                          keys + (size_t)j * n, kout + (size_t)j * n, iota, vout + (size_t)j * n, (int)n, sq);
            if (st != KP_OK) return st;
            if (sq != stream) {
                KP_CUDA(cudaEventRecord(side->join[q - 1], sq));
                KP_CUDA(cudaStreamWaitEvent(stream, side->join[q - 1], 0));
            }
        }
        if (side_lock.owns_lock()) side_lock.unlock();
        }
        post_sort_kernel<<<dim3(nb, list.count), 256, 0, stream>>>(kout, vout, (int)n, (int)L.npad, list, pos, sorted, flag);
        KP_LAUNCH_CHECK("post_sort_kernel");
        for (int tq = 0; tq < 2; ++tq) {
            const int st = inclusive_sum_u32(tmp, L.sort_tmp_bytes, flag + (size_t)tq * L.npad, incl + (size_t)tq * L.npad,
                                             (int)n, stream);
            if (st != KP_OK) return st;
        }
        scatter_tables_kernel<<<(int)((L.npad + 255) / 256), 256, 0, stream>>>(keys, pos, sorted, flag, incl, (int)n,
                                                                               strat_mask, po);
        KP_LAUNCH_CHECK("scatter_tables_kernel");
    }

    if (tuning.plan_only) return KP_OK;

    // ---- pods ----
    if (per_pod) {
        const int gb = (int)((P + kGroupItems - 1) / kGroupItems);
        // thresholds the pass compares against (largest of the table; the probed best-first nodes) and their integer images
        double2 *probe = reinterpret_cast<double2 *>(w + L.off_probe);
        ProbeInt *lim = reinterpret_cast<ProbeInt *>(w + L.off_probe_int);
        GroupPods gp{(const long long *)pods.strategy_ids, (const long long *)pods.req_cpu_milli, (const long long *)pods.req_mem_bytes,
                     canon_strategy(strategy), P, po.uniq, (int)L.npad, big ? 0 : (int)n, (long long *)out_idx, out_score, tuning.out_gpos,
                     probe, lim, prune_depth, po.order, po.score_pos, tuning.gpos_by_pos, (size_t)tuning.gpos_stride};
        probe_table_kernel<<<1, kMaxStrategies * kProbeMax, 0, stream>>>(t, strat_mask, hdr, gp, probe, lim);
        KP_LAUNCH_CHECK("probe_table_kernel");
        group_count_kernel<<<gb, 256, 0, stream>>>(gp, hdr);
        KP_LAUNCH_CHECK("group_count_kernel");
        group_assign_kernel<<<gb, 256, 0, stream>>>(gp, chunk_pods, hdr, perm);
        KP_LAUNCH_CHECK("group_assign_kernel");
    }
    if (nseg > 1) KP_CUDA(cudaMemsetAsync(best_pos, 0xff, sizeof(uint32_t) * n_chunks * chunk_pods, stream));
    PlaceParams pp{};
    pp.req_cpu = (const long long *)pods.req_cpu_milli;
    pp.req_mem = (const long long *)pods.req_mem_bytes;
    pp.perm = perm;
    pp.n_pods = P;
    pp.n_nodes = (int)n;
    pp.npad = (int)L.npad;
    pp.uniform_strategy = uniform;
    pp.one = 1;
    pp.tiles_per_seg_wide = tiles_per_seg;
    pp.uniq = po.uniq;
    pp.ab = po.ab;
    pp.packed = po.packed;
    pp.order = po.order;
    pp.score = po.score;
    pp.score_by_pos = po.score_pos;
    pp.small_table_len = use_small ? (int)n : ((!big && !per_pod) ? (int)((n + 3) / 4 * 4) : 0);
    pp.hdr = hdr;
    pp.chunk_done = reinterpret_cast<uint32_t *>(w + L.off_chunk_done);
    pp.best_pos = best_pos;
    pp.out_idx = (long long *)out_idx;
    pp.out_score = out_score;
    pp.out_gpos = tuning.out_gpos;
    pp.gpos_by_pos = tuning.gpos_by_pos;
    pp.gpos_stride = tuning.gpos_stride;
    // pacing applies to the kernels that fetch a CTA's requests up front (node splits, one strategy)
    const bool paced = tuning.flow && flow_window > 0 && !per_pod && !big && !use_small && kPlaceCfgs[cfg].nsplit > 1;
    pp.flow = paced ? tuning.flow : nullptr;
    pp.flow_window = flow_window;
    if (pe.enabled) KP_CUDA(cudaEventRecord(pe.e1, stream));
    const int st = use_small ? launch_place_small(pp, n, P, per_pod, stream)
                             : launch_place_cfg(cfg, pp, dim3((unsigned)n_chunks, (unsigned)nseg, 1), stream, resident_cap);
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
constexpr int64_t kNoAllocMaxPods = 65536;
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
    if (per_pod && pods.n_pods > kNoAllocMaxPods) {
        set_error("kp_snapshot_place_noalloc: %lld pods with per-pod strategy ids exceed the built-in "
                  "scratch limit of %lld (uniform-strategy calls have no pod limit)",
                  (long long)pods.n_pods, (long long)kNoAllocMaxPods);
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
    // counter, merge buffer and table entry (including padding) it later reads
    return snapshot_place_device(nodes, pods, strategy, filter_mode, out_idx, out_score,
                                 scratch_addr[dev], kNoAllocScratchBytes, stream);
}

}  // namespace kp

// decide_kernel: ONE decision for ONE pod -- the call shape of the reference itself.
//
// The reference asks for a node once per pending pod (scheduler.py:598 -> :521-559: filter, score, first-best argmax
// over the node list it was just handed).  Through the batched snapshot path that is a staging copy, a header memset,
// a kernel and a stream synchronisation for a 3-node table -- five driver calls and their latencies for a few hundred
// bytes.  Here the whole decision is one launch of one CTA:
//
//   * the node columns stay in the session's PINNED HOST staging buffer; every thread loads the rows it owns straight
//     across the host link (all loads of a thread are issued before the first use: one round trip, not seven);
//   * score and filter are the same exact-fp64 functions the plan uses (kp_common.cuh), the argmax is a warp / CTA
//     reduction on (order key, node index) -- larger key first, then the FIRST node in list order (scheduler.py:547);
//   * the answer goes back as two words in pinned host memory followed by a sequence number stored with release
//     semantics at system scope; the host polls that word instead of synchronising the stream.
//
// Tables above kDecideMaxNodes rows take the batched path (one SM cannot keep the host link busy).
#include <algorithm>

#include "kp_common.cuh"

namespace kp {

constexpr int kDecideThreads = 1024;
constexpr int kDecidePerThread = 2;

struct DecideNodes {
    const long long *alloc_cpu, *alloc_mem, *pod_count, *max_pods;
    const uint8_t *ready;
    const double *cu, *mu;
    int n;
};

struct Pick { unsigned long long key; uint32_t idx; double score; };

__device__ __forceinline__ bool pick_better(const Pick &a, const Pick &b)
{
    return a.key > b.key || (a.key == b.key && a.idx < b.idx);
}
__device__ __forceinline__ Pick pick_warp_best(Pick c)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Pick d;
        d.key = __shfl_xor_sync(0xffffffffu, c.key, o);
        d.idx = __shfl_xor_sync(0xffffffffu, c.idx, o);
        d.score = __shfl_xor_sync(0xffffffffu, c.score, o);
        if (pick_better(d, c)) c = d;
    }
    return c;
}

__global__ void __launch_bounds__(kDecideThreads)
decide_kernel(DecideNodes t, int s, int capacity, long long req_cpu_milli, long long req_mem_bytes, DecideSlot *slot, unsigned long long seq)
{
    __shared__ Pick sm[kDecideThreads / 32];
    // every load first (predicated, independent), then the arithmetic
    long long pc[kDecidePerThread], mp[kDecidePerThread], ac[kDecidePerThread], am[kDecidePerThread];
    double cu[kDecidePerThread], mu[kDecidePerThread];
    uint8_t rdy[kDecidePerThread];
#pragma unroll
    for (int k = 0; k < kDecidePerThread; ++k) {
        const int n = (int)threadIdx.x + k * (int)blockDim.x;
        rdy[k] = 0;
        pc[k] = mp[k] = ac[k] = am[k] = 0;
        cu[k] = mu[k] = 0.0;
        if (n < t.n) {
            rdy[k] = t.ready[n];
            pc[k] = t.pod_count[n];
            mp[k] = t.max_pods[n];
            cu[k] = t.cu[n];
            mu[k] = t.mu[n];
            if (capacity) {
                ac[k] = t.alloc_cpu[n];
                am[k] = t.alloc_mem[n];
            }
        }
    }
    const double rc = req_cores(req_cpu_milli), rm = req_gib(req_mem_bytes);
    Pick best{0ull, 0xffffffffu, -1.0};
#pragma unroll
    for (int k = 0; k < kDecidePerThread; ++k) {
        const int n = (int)threadIdx.x + k * (int)blockDim.x;
        if (n >= t.n || !rdy[k]) continue;                                     // scheduler.py:531-533
        if (capacity) {
            // pc < max_pods, request <= free capacity in both dimensions (NaN never fits: the comparisons are false)
            if (!(pc[k] < mp[k])) continue;
            if (!(rc <= free_cpu_cores(ac[k], cu[k]) && rm <= free_mem_gib(am[k], mu[k]))) continue;
        }
        const NodeScore ns = node_score(s, cu[k], mu[k], pc[k], mp[k]);
        const Pick c{ns.key, (uint32_t)n, ns.score};
        if (ns.key != 0ull && pick_better(c, best)) best = c;               // key 0: cannot beat the -1 sentinel
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = (blockDim.x + 31) >> 5;
    best = pick_warp_best(best);
    if (nwarps > 1) {
        if (lane == 0) sm[warp] = best;
        __syncthreads();
        if (warp != 0) return;
        best = lane < nwarps ? sm[lane] : Pick{0ull, 0xffffffffu, -1.0};
        best = pick_warp_best(best);
    }
    if (threadIdx.x == 0) {
        const bool ok = best.key != 0ull;
        slot->idx = ok ? (long long)best.idx : -1;
        slot->score = ok ? best.score : -1.0;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&slot->seq), "l"(seq) : "memory");
    }
}

// t: columns the device can address (pinned host memory or device memory), t.n_nodes <= kDecideMaxNodes.
int decide_device(const kp_node_table &nodes, int strategy, int filter_mode, long long req_cpu_milli, long long req_mem_bytes,
                  DecideSlot *slot, unsigned long long seq, cudaStream_t stream)
{
    const int64_t n = nodes.n_nodes;
    if (n < 0 || n > kDecideMaxNodes) { set_error("decide: %lld nodes (at most %d)", (long long)n, kDecideMaxNodes); return KP_ERR_INVALID_ARG; }
    DecideNodes t{(const long long *)nodes.alloc_cpu_milli, (const long long *)nodes.alloc_mem_bytes, (const long long *)nodes.pod_count,
               (const long long *)nodes.max_pods, nodes.ready, nodes.cpu_usage_pct, nodes.mem_usage_pct, (int)n};
    int threads = (int)((n + kDecidePerThread - 1) / kDecidePerThread);
    threads = std::max(32, std::min(kDecideThreads, (threads + 31) / 32 * 32));
    decide_kernel<<<1, threads, 0, stream>>>(t, canon_strategy(strategy), filter_mode == KP_FILTER_CAPACITY, req_cpu_milli, req_mem_bytes, slot, seq);
    KP_LAUNCH_CHECK("decide_kernel");
    return KP_OK;
}

}  // namespace kp

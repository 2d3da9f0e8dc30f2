// Node-sharded snapshot placement: the per-pod arg-max reduce over NVLink peer memory.
//
// BASELINE config 5 / SURVEY.md §8(e): every rank (one process, one GPU) holds a contiguous slice of the node
// list and scores ALL pods against its slice; the per-pod winner over the whole list is the arg-max under
// (score desc, node index asc) -- scheduler.py:547's strict '>' with first-in-list-order ties -- over the
// ranks' local winners.  The reduce is done by our own kernels over peer-mapped memory (CUDA IPC), not by a
// collective library and not through the host:
//
//   1. keys     every rank builds the node plan of its slice (kp_plan.cuh) and publishes, in its exchange buffer,
//               the order keys of its eligible nodes' scores in best-first order; it raises flag 0 in every
//               peer's buffer (st.release.sys).
//   2. place    every rank pulls all ranks' sorted keys (plain peer loads) and finds, by binary search in each
//               peer's list, how many nodes of the whole cluster sort before each of ITS nodes -- the node's
//               GLOBAL rank position, a 32-bit number that is unique per node and orders exactly like
//               (score desc, index asc).  Then the ordinary placement kernel scans the slice; its epilogue
//               additionally stores the winner's global position per pod.  Flag 1 is raised.
//   3. reduce   pods are owned in contiguous shards; the owner pulls the W candidate positions of each of its
//               pods from the peers' buffers, takes the minimum, and fetches index and score from the one rank
//               that holds the winner.  4 W + 16 bytes cross NVLink per pod, coalesced; nothing is reduced twice.
//
// Flags carry an epoch that only grows; per-pod buffers alternate by epoch parity, so a rank may start the next
// snapshot while a peer still reads the last one.  Waits are bounded (a peer that never arrives sets an error
// word instead of hanging the GPU).  All tables a kernel stages with bulk copies live in this GPU's own memory;
// peer memory is only ever touched by ordinary loads and stores.
#include <algorithm>
#include <cstring>
#include <new>

#include "kp_common.cuh"

namespace kp {

constexpr int kMaxWorld = 16;
constexpr unsigned long long kWaitNs = 15000000000ull;         // 15 s: give up on a peer (a rank may still be loading its CUDA modules)

struct ExHeader {                       // start of every rank's buffer
    uint32_t flag[2][kMaxWorld];        // flag[phase][source rank] = epoch reached (written by that rank)
    uint32_t error;                     // set by a kernel of this rank that gave up waiting
    uint32_t nvalid[kMaxStrategies];    // published: eligible nodes of this rank's slice per strategy
    long long n_local;                  // this rank's slice
    long long node_offset;
};
struct ExMeta {                         // pulled copy of every rank's header fields
    uint32_t nvalid[kMaxWorld][kMaxStrategies];
    long long node_offset[kMaxWorld];
};

struct ExLayout {
    size_t off_keys, off_gathered, off_gbase, off_gpos, off_pod[2], pod_stride_idx, pod_stride_score, total;
};

static ExLayout ex_layout(int world, int64_t max_pods, int64_t max_local)
{
    ExLayout L{};
    size_t o = align_up(sizeof(ExHeader));
    L.off_keys = o;     o = align_up(o + sizeof(uint64_t) * kMaxStrategies * max_local);            // published: sorted keys
    L.off_gathered = o; o = align_up(o + sizeof(uint64_t) * kMaxStrategies * max_local * world);    // pulled copy: [s][rank][pos]
    L.off_gbase = o;    o = align_up(o + sizeof(ExMeta));
    L.off_gpos = o;     o = align_up(o + sizeof(uint32_t) * kMaxStrategies * max_local);            // global position per local position
    for (int b = 0; b < 2; ++b) {
        L.off_pod[b] = o;
        o = align_up(o + sizeof(uint32_t) * max_pods);      // candidate global position per pod
        L.pod_stride_idx = o - L.off_pod[b];
        o = align_up(o + sizeof(long long) * max_pods);     // local node index
        L.pod_stride_score = o - L.off_pod[b];
        o = align_up(o + sizeof(double) * max_pods);        // score
    }
    L.total = o;
    return L;
}

}  // namespace kp

using namespace kp;

struct kp_exchange {
    int device = 0, rank = 0, world = 1;
    int64_t max_pods = 0, max_local = 0;
    unsigned char *base = nullptr;
    unsigned char *peer[kMaxWorld] = {nullptr};
    bool ipc_opened[kMaxWorld] = {false};
    bool connected = false;
    uint32_t epoch = 0;
    int64_t n_pods = 0;                 // of the call in flight
    int prune_depth = 0;                // kp_exchange_set_prune_depth: first-feasible probe in kp_ns_place
    ExLayout L{};
};

namespace kp {

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys_u64(const void *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ld_sys_u32(const void *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

struct ExPeers { unsigned char *p[kMaxWorld]; };

// Block-wide: wait until every rank has raised flag[phase] to `epoch` in THIS rank's buffer.  Returns false
// (and records the error) when a peer does not arrive in time.
__device__ __forceinline__ bool wait_all_ranks(ExHeader *mine, int phase, int world, uint32_t epoch)
{
    __shared__ int s_ok;
    if (threadIdx.x == 0) {
        int ok = 1;
        const unsigned long long t0 = global_ns();
        for (int r = 0; r < world && ok; ++r) {
            while ((int)(ld_acquire_sys(&mine->flag[phase][r]) - epoch) < 0) {
                __nanosleep(200);
                if (global_ns() - t0 > kWaitNs) { ok = 0; mine->error = 1u + (uint32_t)phase; break; }
            }
        }
        s_ok = ok;
    }
    __syncthreads();
    return s_ok != 0;
}

// ---- phase 1: publish this slice's score keys in best-first order -------------------------------------------
// key = ~order_key(score): ascending = best first, the order the plan's tables are in.  The plan keeps the fp64
// score of the node at every rank position; the key is rebuilt from it (round_robin scores are pod counts and
// order as integers, like the reference's int compare, scheduler.py:544-549).
__device__ __forceinline__ unsigned long long sort_key_of_score(int s, double score)
{
    return s == 2 ? ~i64_order_key((long long)score) : ~f64_order_key(score);
}

// one block: writes the keys, then raises flag 0 in every rank's buffer (including its own)
__global__ void __launch_bounds__(1024)
ns_publish_kernel(ExPeers peers, int rank, int world, uint32_t epoch, int strat_mask, const uint32_t *__restrict__ plan_nvalid,
                  const double *__restrict__ score_pos, int npad, long long n_local, long long node_offset, size_t off_keys, long long max_local)
{
    unsigned char *base = peers.p[rank];
    ExHeader *h = reinterpret_cast<ExHeader *>(base);
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(base + off_keys);
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        const int nv = (strat_mask & (1 << s)) && n_local > 0 ? (int)plan_nvalid[s] : 0;
        if (threadIdx.x == 0) h->nvalid[s] = (uint32_t)nv;
        for (int i = threadIdx.x; i < nv; i += blockDim.x) keys[(size_t)s * max_local + i] = sort_key_of_score(s, score_pos[(size_t)s * npad + i]);
    }
    if (threadIdx.x == 0) { h->n_local = n_local; h->node_offset = node_offset; }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) st_release_sys(&reinterpret_cast<ExHeader *>(peers.p[threadIdx.x])->flag[0][rank], epoch);
}

// raise flag[phase][my rank] = epoch in every rank's buffer (one thread per rank)
__global__ void ns_signal_kernel(ExPeers peers, int rank, int world, int phase, uint32_t epoch)
{
    const int r = threadIdx.x;
    if (r >= world) return;
    __threadfence_system();
    st_release_sys(&reinterpret_cast<ExHeader *>(peers.p[r])->flag[phase][rank], epoch);
}

// ---- phase 2a: pull every rank's sorted keys into local memory ---------------------------------------------
// grid (x, world): block row r copies rank r's lists to gathered[s][r][pos]
__global__ void __launch_bounds__(256)
ns_gather_keys_kernel(ExPeers peers, int rank, int world, uint32_t epoch, size_t off_keys, size_t off_gathered, size_t off_meta,
                      long long max_local)
{
    unsigned char *base = peers.p[rank];
    if (!wait_all_ranks(reinterpret_cast<ExHeader *>(base), 0, world, epoch)) return;
    const int r = blockIdx.y;
    const ExHeader *h = reinterpret_cast<const ExHeader *>(peers.p[r]);
    ExMeta *meta = reinterpret_cast<ExMeta *>(base + off_meta);
    unsigned long long *gathered = reinterpret_cast<unsigned long long *>(base + off_gathered);
    const unsigned long long *src = reinterpret_cast<const unsigned long long *>(peers.p[r] + off_keys);
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        const uint32_t nv = ld_sys_u32(&h->nvalid[s]);
        if (blockIdx.x == 0 && threadIdx.x == 0) meta->nvalid[r][s] = nv;
        unsigned long long *dst = gathered + ((size_t)s * world + r) * max_local;
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += gridDim.x * blockDim.x)
            dst[i] = ld_sys_u64(src + (size_t)s * max_local + i);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) meta->node_offset[r] = (long long)ld_sys_u64(&h->node_offset);
}

// ---- phase 2b: global rank position of every local rank position -------------------------------------------
// gpos[s][pos] = pos + sum over the other ranks r of #{keys of r that sort before mine}: keys equal to mine count
// for ranks below me (their nodes come first in list order) and not for ranks above.  One binary search per peer
// list, all peers side by side (independent load chains).
__global__ void __launch_bounds__(256)
ns_gpos_kernel(const unsigned char *base, int rank, int world, int strat_mask, size_t off_gathered, size_t off_meta, size_t off_gpos,
               long long max_local)
{
    const int s = blockIdx.y;
    if (!(strat_mask & (1 << s))) return;
    const ExMeta *meta = reinterpret_cast<const ExMeta *>(base + off_meta);
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= (int)meta->nvalid[rank][s]) return;
    const unsigned long long *K = reinterpret_cast<const unsigned long long *>(base + off_gathered) + (size_t)s * world * max_local;
    const unsigned long long kn = K[(size_t)rank * max_local + pos];
    int lo[kMaxWorld], nv[kMaxWorld], top = 1;
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
        lo[r] = 0;
        nv[r] = (r < world && r != rank) ? (int)meta->nvalid[r][s] : 0;
        top = max(top, nv[r]);
    }
    for (int step = 1 << (31 - __clz(top)); step > 0; step >>= 1) {
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r) {
            const int nxt = lo[r] + step;
            if (nxt <= nv[r]) {                                   // (nv == 0 for ranks that do not take part)
                const unsigned long long km = K[(size_t)r * max_local + nxt - 1];
                if (r < rank ? km <= kn : km < kn) lo[r] = nxt;
            }
        }
    }
    uint32_t g = (uint32_t)pos;
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) g += (uint32_t)lo[r];
    reinterpret_cast<uint32_t *>(const_cast<unsigned char *>(base) + off_gpos)[(size_t)s * max_local + pos] = g;
}

// ---- phase 3: owner-side reduce over the peers' candidates -------------------------------------------------
__global__ void __launch_bounds__(256)
ns_reduce_kernel(ExPeers peers, int rank, int world, uint32_t epoch, size_t off_pod, size_t stride_idx, size_t stride_score, size_t off_meta,
                 long long pod_lo, long long pod_hi, long long *__restrict__ out_idx, double *__restrict__ out_score)
{
    unsigned char *base = peers.p[rank];
    if (!wait_all_ranks(reinterpret_cast<ExHeader *>(base), 1, world, epoch)) return;
    const ExMeta *meta = reinterpret_cast<const ExMeta *>(base + off_meta);
    for (long long j = pod_lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; j < pod_hi; j += (long long)gridDim.x * blockDim.x) {
        uint32_t g[kMaxWorld];
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)                         // all candidates requested before any is looked at
            g[r] = r < world ? ld_sys_u32(peers.p[r] + off_pod + sizeof(uint32_t) * j) : 0xffffffffu;
        uint32_t best = 0xffffffffu;
        int who = -1;
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)
            if (g[r] < best) { best = g[r]; who = r; }
        long long idx = -1;
        double sc = -1.0;
        if (who >= 0) {
            const unsigned long long li = ld_sys_u64(peers.p[who] + off_pod + stride_idx + sizeof(long long) * j);
            const unsigned long long ls = ld_sys_u64(peers.p[who] + off_pod + stride_score + sizeof(double) * j);
            idx = (long long)li + meta->node_offset[who];
            sc = __longlong_as_double((long long)ls);
        }
        out_idx[j - pod_lo] = idx;
        out_score[j - pod_lo] = sc;
    }
}

static void shard_bounds(int64_t n, int rank, int world, int64_t *lo, int64_t *hi)
{
    const int64_t base = n / world, rem = n % world;
    *lo = rank * base + (rank < rem ? rank : rem);
    *hi = *lo + base + (rank < rem ? 1 : 0);
}

static ExPeers peers_of(const kp_exchange *ex)
{
    ExPeers p{};
    for (int r = 0; r < ex->world; ++r) p.p[r] = ex->peer[r];
    return p;
}

}  // namespace kp

extern "C" {

int kp_exchange_create(int device, int rank, int world, int64_t max_pods, int64_t max_local_nodes, kp_exchange **out)
{
    if (!out || world < 1 || world > kMaxWorld || rank < 0 || rank >= world || max_pods < 0 || max_local_nodes < 0) {
        set_error("kp_exchange_create: bad arguments (world 1..%d)", kMaxWorld);
        return KP_ERR_INVALID_ARG;
    }
    *out = nullptr;
    int st = kp_device_info(device, nullptr, nullptr, nullptr);
    if (st != KP_OK) return st;
    int prev = -1;
    cudaGetDevice(&prev);
    KP_CUDA(cudaSetDevice(device));
    kp_exchange *ex = new (std::nothrow) kp_exchange();
    if (!ex) { set_error("out of host memory"); return KP_ERR_INVALID_ARG; }
    ex->device = device; ex->rank = rank; ex->world = world;
    ex->max_pods = max_pods; ex->max_local = max_local_nodes;
    ex->L = ex_layout(world, max_pods, max_local_nodes);
    cudaError_t e = cudaMalloc((void **)&ex->base, ex->L.total);
    if (e == cudaSuccess) e = cudaMemset(ex->base, 0, ex->L.total);
    if (prev >= 0) cudaSetDevice(prev);
    if (e != cudaSuccess) { delete ex; return cuda_fail(e, "kp_exchange_create allocation"); }
    ex->peer[rank] = ex->base;
    ex->connected = world == 1;
    *out = ex;
    return KP_OK;
}

int kp_exchange_handle(kp_exchange *ex, unsigned char *out_handle64)
{
    if (!ex || !out_handle64) { set_error("kp_exchange_handle: NULL argument"); return KP_ERR_INVALID_ARG; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    int prev = -1;
    cudaGetDevice(&prev);
    KP_CUDA(cudaSetDevice(ex->device));
    const cudaError_t e = cudaIpcGetMemHandle(&h, ex->base);
    if (prev >= 0) cudaSetDevice(prev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaIpcGetMemHandle");
    memcpy(out_handle64, &h, 64);
    return KP_OK;
}

int kp_exchange_connect(kp_exchange *ex, const unsigned char *handles)
{
    if (!ex || !handles) { set_error("kp_exchange_connect: NULL argument"); return KP_ERR_INVALID_ARG; }
    int prev = -1;
    cudaGetDevice(&prev);
    KP_CUDA(cudaSetDevice(ex->device));
    for (int r = 0; r < ex->world; ++r) {
        if (r == ex->rank || ex->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + 64 * (size_t)r, 64);
        void *p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { if (prev >= 0) cudaSetDevice(prev); return cuda_fail(e, "cudaIpcOpenMemHandle (peer access between the GPUs is required)"); }
        ex->peer[r] = static_cast<unsigned char *>(p);
        ex->ipc_opened[r] = true;
    }
    if (prev >= 0) cudaSetDevice(prev);
    ex->connected = true;
    return KP_OK;
}

int kp_exchange_connect_local(kp_exchange *ex, kp_exchange *const *all)
{
    if (!ex || !all) { set_error("kp_exchange_connect_local: NULL argument"); return KP_ERR_INVALID_ARG; }
    for (int r = 0; r < ex->world; ++r) {
        if (!all[r] || all[r]->world != ex->world || all[r]->rank != r || all[r]->L.total != ex->L.total) {
            set_error("kp_exchange_connect_local: exchange %d does not belong to this group", r);
            return KP_ERR_INVALID_ARG;
        }
        ex->peer[r] = all[r]->base;
    }
    // exchanges on other devices of this process: plain peer access instead of IPC
    int prev = -1;
    cudaGetDevice(&prev);
    KP_CUDA(cudaSetDevice(ex->device));
    for (int r = 0; r < ex->world; ++r) {
        if (all[r]->device == ex->device) continue;
        const cudaError_t e = cudaDeviceEnablePeerAccess(all[r]->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
            if (prev >= 0) cudaSetDevice(prev);
            return cuda_fail(e, "cudaDeviceEnablePeerAccess (peer access between the GPUs is required)");
        }
        cudaGetLastError();
    }
    if (prev >= 0) cudaSetDevice(prev);
    ex->connected = true;
    return KP_OK;
}

void kp_exchange_destroy(kp_exchange *ex)
{
    if (!ex) return;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(ex->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < ex->world; ++r)
        if (ex->ipc_opened[r]) cudaIpcCloseMemHandle(ex->peer[r]);
    cudaFree(ex->base);
    if (prev >= 0) cudaSetDevice(prev);
    delete ex;
}

int kp_exchange_set_prune_depth(kp_exchange *ex, int probe_depth)
{
    if (!ex || probe_depth < 0) { set_error("kp_exchange_set_prune_depth: NULL exchange or negative depth"); return KP_ERR_INVALID_ARG; }
    ex->prune_depth = probe_depth;
    return KP_OK;
}

int kp_exchange_error(kp_exchange *ex, int *out_error)
{
    if (!ex || !out_error) { set_error("kp_exchange_error: NULL argument"); return KP_ERR_INVALID_ARG; }
    uint32_t v = 0;
    KP_CUDA(cudaMemcpy(&v, ex->base + offsetof(ExHeader, error), sizeof(v), cudaMemcpyDeviceToHost));
    *out_error = (int)v;
    if (v) set_error("node-sharded exchange: a peer did not reach phase %u within %llu ms", v - 1u, kWaitNs / 1000000ull);
    return KP_OK;
}

static int ns_check(const kp_exchange *ex, const char *what)
{
    if (!ex) { set_error("%s: exchange is NULL", what); return KP_ERR_INVALID_ARG; }
    if (!ex->connected) { set_error("%s: exchange is not connected to its peers", what); return KP_ERR_INVALID_ARG; }
    return KP_OK;
}

static int ns_check_tables(const kp_exchange *ex, const kp_node_table *nodes, const kp_pod_table *pods, int filter_mode, const char *what)
{
    if (!nodes || !pods || pods->n_pods < 0 || pods->n_pods > ex->max_pods || nodes->n_nodes < 0 || nodes->n_nodes > ex->max_local) {
        set_error("%s: table sizes exceed the exchange (%lld pods, %lld nodes per rank)", what, (long long)ex->max_pods, (long long)ex->max_local);
        return KP_ERR_CAPACITY;
    }
    if (filter_mode != KP_FILTER_CAPACITY) {
        set_error("node-sharded placement implements the capacity filter (the Ready-only filter is pod-independent: one arg-max)");
        return KP_ERR_INVALID_ARG;
    }
    return KP_OK;
}

int kp_ns_keys(kp_exchange *ex, const kp_node_table *nodes, int64_t node_offset, const kp_pod_table *pods, int strategy, int filter_mode,
               void *workspace, size_t workspace_bytes, void *stream)
{
    int st = ns_check(ex, "kp_ns_keys");
    if (st != KP_OK) return st;
    if ((st = ns_check_tables(ex, nodes, pods, filter_mode, "kp_ns_keys")) != KP_OK) return st;
    cudaStream_t sq = (cudaStream_t)stream;
    ex->epoch += 1;
    const bool per_pod = pods->strategy_ids != nullptr;
    const int strat_mask = per_pod ? 7 : (1 << canon_strategy(strategy));
    const int64_t n = nodes->n_nodes;
    PlanView pv{nullptr, nullptr, 0};
    if (n > 0 && pods->n_pods > 0) {
        // the node plan of this slice, and nothing else yet
        PlaceTuning tuning;
        tuning.plan_only = true;
        st = snapshot_place_device(*nodes, *pods, strategy, filter_mode, nullptr, nullptr, workspace, workspace_bytes, sq, tuning);
        if (st != KP_OK) return st;
        pv = snapshot_plan_view(n, pods->n_pods, per_pod, workspace);
    }
    ns_publish_kernel<<<1, 1024, 0, sq>>>(peers_of(ex), ex->rank, ex->world, ex->epoch, strat_mask, pv.nvalid, pv.score_pos, pv.npad,
                                          pv.nvalid ? n : 0, node_offset, ex->L.off_keys, ex->max_local);
    KP_LAUNCH_CHECK("ns_publish_kernel");
    return KP_OK;
}

int kp_ns_place(kp_exchange *ex, const kp_node_table *nodes, const kp_pod_table *pods, int strategy, int filter_mode,
                void *workspace, size_t workspace_bytes, void *stream)
{
    int st = ns_check(ex, "kp_ns_place");
    if (st != KP_OK) return st;
    if ((st = ns_check_tables(ex, nodes, pods, filter_mode, "kp_ns_place")) != KP_OK) return st;
    cudaStream_t sq = (cudaStream_t)stream;
    const bool per_pod = pods->strategy_ids != nullptr;
    const int strat_mask = per_pod ? 7 : (1 << canon_strategy(strategy));
    const ExPeers peers = peers_of(ex);
    ns_gather_keys_kernel<<<dim3(4, ex->world), 256, 0, sq>>>(peers, ex->rank, ex->world, ex->epoch, ex->L.off_keys, ex->L.off_gathered,
                                                             ex->L.off_gbase, ex->max_local);
    KP_LAUNCH_CHECK("ns_gather_keys_kernel");
    const int n = (int)nodes->n_nodes;
    if (n > 0) {
        ns_gpos_kernel<<<dim3((n + 255) / 256, kMaxStrategies), 256, 0, sq>>>(ex->base, ex->rank, ex->world, strat_mask, ex->L.off_gathered,
                                                                              ex->L.off_gbase, ex->L.off_gpos, ex->max_local);
        KP_LAUNCH_CHECK("ns_gpos_kernel");
    }
    ex->n_pods = pods->n_pods;
    unsigned char *pb = ex->base + ex->L.off_pod[ex->epoch & 1u];
    PlaceTuning tuning;
    tuning.out_gpos = reinterpret_cast<uint32_t *>(pb);
    tuning.gpos_by_pos = reinterpret_cast<const uint32_t *>(ex->base + ex->L.off_gpos);
    tuning.gpos_stride = ex->max_local;
    tuning.reuse_plan = tuning.reuse_plan_forced = true;      // kp_ns_keys built it in this workspace a moment ago
    tuning.prune_depth = ex->prune_depth;                     // (optional) local first fit by probe; the reduce is unchanged
    if (pods->n_pods > 0 && n == 0)                      // an empty slice offers no candidate
        KP_CUDA(cudaMemsetAsync(tuning.out_gpos, 0xff, sizeof(uint32_t) * (size_t)pods->n_pods, sq));
    if (pods->n_pods > 0 && n > 0) {
        st = snapshot_place_device(*nodes, *pods, strategy, filter_mode, reinterpret_cast<int64_t *>(pb + ex->L.pod_stride_idx),
                                   reinterpret_cast<double *>(pb + ex->L.pod_stride_score), workspace, workspace_bytes, sq, tuning);
        if (st != KP_OK) return st;
    }
    ns_signal_kernel<<<1, kMaxWorld, 0, sq>>>(peers, ex->rank, ex->world, 1, ex->epoch);
    KP_LAUNCH_CHECK("ns_signal_kernel");
    return KP_OK;
}

int kp_ns_reduce(kp_exchange *ex, int64_t *out_idx_shard, double *out_score_shard, void *stream)
{
    int st = ns_check(ex, "kp_ns_reduce");
    if (st != KP_OK) return st;
    int64_t lo = 0, hi = 0;
    shard_bounds(ex->n_pods, ex->rank, ex->world, &lo, &hi);
    if (hi > lo && (!out_idx_shard || !out_score_shard)) { set_error("kp_ns_reduce: output pointer is NULL"); return KP_ERR_INVALID_ARG; }
    const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((hi - lo + 255) / 256, (int64_t)sm_count_cached() * 8));
    ns_reduce_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(peers_of(ex), ex->rank, ex->world, ex->epoch, ex->L.off_pod[ex->epoch & 1u],
                                                              ex->L.pod_stride_idx, ex->L.pod_stride_score, ex->L.off_gbase, lo, hi,
                                                              (long long *)out_idx_shard, out_score_shard);
    KP_LAUNCH_CHECK("ns_reduce_kernel");
    return KP_OK;
}

int kp_snapshot_place_node_sharded(kp_exchange *ex, const kp_node_table *local_nodes, int64_t node_offset, const kp_pod_table *pods,
                                   int strategy, int filter_mode, int64_t *out_idx_shard, double *out_score_shard, void *workspace,
                                   size_t workspace_bytes, void *stream)
{
    int st = kp_ns_keys(ex, local_nodes, node_offset, pods, strategy, filter_mode, workspace, workspace_bytes, stream);
    if (st != KP_OK) return st;
    if ((st = kp_ns_place(ex, local_nodes, pods, strategy, filter_mode, workspace, workspace_bytes, stream)) != KP_OK) return st;
    return kp_ns_reduce(ex, out_idx_shard, out_score_shard, stream);
}

}  // extern "C"

// place_kernel: the hot loop of snapshot placement (included by kp_snapshot.cu).
//
// Work decomposition.  grid.x = pod chunks, grid.y = node segments (a whole number of
// shared-memory tiles each).  Inside a CTA the warps form a PODW x NSPLIT arrangement: warp
// (podw, split) owns 32*R pods (R per lane, in registers) and scans slice `split` of every
// tile.  Small inputs use NSPLIT > 1 so that a CTA carries few pods but many warps (fills the
// machine when there are few pods); large tables use grid.y > 1 so that there are many more
// CTAs than resident slots (no wave-quantisation tail).
//
// Tiles of the best-first node table are staged in shared memory with 1-D bulk async copies
// (TMA engine, SASS UBLKCP) on a 2-deep mbarrier ring; every lane reads the same nodes
// (one broadcast LDS.128 = 4 packed nodes).  Scanning worst -> best and overwriting the
// candidate whenever the pod fits leaves the best-ranked fitting node of the range; ranges
// are merged by taking the minimum rank position: across the warps of a CTA through shared
// memory, across node segments with atomicMin in global memory, after which the CTA that
// finishes a pod chunk last maps position -> node index / fp64 score and stores the result.
//
// Per pod x node pair, packed form (use_swar):
//     t = node_word - pod_word           IMAD.IADD   (FMA pipe)
//     p = (~t & 0x80008000) != 0         LOP3 -> P   (ALU pipe)   both guard bits survive iff
//     @!p best = pos * 1 + pod_word      IMAD        (FMA pipe)   neither 15-bit field borrowed
// tools/issue_bench.cu measured this mix (variant F) at 11.0 Tpairs/s on B200 = 89 % of the
// 3-instruction issue bound, against 6.0 for two integer compares + select (ALU pipe only).
// The wide form (more than 32000 distinct thresholds) uses two 32-bit compares + select.
#pragma once

#include "kp_common.cuh"

namespace kp {

#ifndef KP_TILE_BYTES
#define KP_TILE_BYTES 4096
#endif
#ifndef KP_UPDATE_MODE
#define KP_UPDATE_MODE 1      // candidate update: 0 = SEL (ALU pipe), 1 = alternate per pod, 2 = guarded IMAD (FMA pipe)
#endif                        // 3 = IMAD that reads the candidate (best * 0 + pos: one vector source), 4 = alternate SEL / that IMAD
constexpr int kTileBytes = KP_TILE_BYTES;              // per shared-memory tile buffer
constexpr int kTileNodesPacked = kTileBytes / 4;       // 1024
constexpr int kTileNodesWide = kTileBytes / 8;         // 512
constexpr uint32_t kSwarGuards = 0x80008000u;
constexpr uint32_t kSwarDeadPod = 0x7fff7fffu;         // larger than any node field: never fits
constexpr uint32_t kNoPos = 0xffffffffu;
constexpr long long kFlowNever = (long long)0x8000000000000001ull;   // sentinel for the arrival vote
constexpr int kUniqStage = 1024;                       // sorted thresholds staged in shared memory up to this many

struct PlaceParams {
    const long long *req_cpu, *req_mem;
    const uint32_t *perm;              // slot -> pod (per-pod strategies), NULL = identity
    long long n_pods;
    int n_nodes, npad, uniform_strategy;
    int one;                           // always 1; run-time value so the IMAD update is not folded away
    int tiles_per_seg_wide;            // tiles (of 512 nodes) per node segment; the packed form uses half
    int small_table_len;               // > 0: small-table plan -- packed form, thresholds sorted WITH duplicates
                                       // (n_nodes of them), table length known on the host: no header reads
    uint32_t *out_gpos;                // node-sharded form: global rank position of the chosen node per pod (NULL = off)
    const uint32_t *gpos_by_pos;       // [3][gpos_stride] global rank position of every local rank position
    long long gpos_stride;
    long long share_q;                 // place_small_kernel, one strategy: warp g owns share_q pods, the first share_r
    int share_r;                       // warps one more (the split is made on the host: no 64-bit division per warp)
    const double *uniq;                // [2][npad] ascending distinct thresholds
    const uint2 *ab;                   // [3][npad] best-first {RA, RB}
    const uint32_t *packed;            // [3][npad] best-first packed ranks
    const uint32_t *order;             // [3][npad] rank position -> node index
    const double *score;               // [3][n]
    const double *score_by_pos;        // [3][npad] score of the node at each rank position
    const SnapHeader *hdr;
    uint32_t *chunk_done;              // [chunks] finished node segments per pod chunk (zeroed; grid.y > 1)
    uint32_t *best_pos;                // [slots] merged best rank position (0xffffffff; grid.y > 1)
    long long *out_idx;
    double *out_score;
    // Host-link flow control (pods read in place from pinned host memory, see kp_api.cu): CTA c
    // asks for its rows only once the rows of CTA c - flow_window have arrived, so the requests
    // reach the link in pod order and early CTAs score while later rows are still in flight.
    // flow[0] = CTAs whose rows have arrived, flow[1] = CTAs finished (the last one resets both).
    unsigned *flow;                    // NULL = off
    unsigned flow_window;
};

// #{i : arr[i] < x} in an ascending array: branch-free, fixed depth (same trip count in every lane)
__device__ __forceinline__ int lower_bound_lt(const double *arr, int n, double x)
{
    int lo = 0;
    for (int step = 1 << (31 - __clz(n | 1)); step > 0; step >>= 1) {
        const int nxt = lo + step;
        if (nxt <= n && arr[nxt - 1] < x) lo = nxt;
    }
    return lo;
}

__device__ __forceinline__ void group_begins(const SnapHeader *hdr, int chunk_pods,
                                             long long begin[kMaxStrategies])
{
    long long o = 0;
#pragma unroll
    for (int s = 0; s < kMaxStrategies; ++s) {
        begin[s] = o;
        const long long c = hdr->group_count[s];
        o += (c + chunk_pods - 1) / chunk_pods * chunk_pods;
    }
}

// packed test + update on the FMA pipe: best_enc = pos * 1 + x   (decoded later as best_enc - x)
__device__ __forceinline__ void swar_update(int &best_enc, uint32_t x, uint32_t w, int pos, int one)
{
    asm volatile(
        "{ .reg .pred p; .reg .b32 t, d;\n"
        "  sub.u32 t, %1, %2;\n"
        "  lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"      // d = ~t & guards ; p = (d != 0) = does not fit
        "  @!p mad.lo.s32 %0, %3, %4, %2; }"
        : "+r"(best_enc) : "r"(w), "r"(x), "r"(pos), "r"(one));
}

// packed test + update on the ALU pipe (SEL): best = pos
__device__ __forceinline__ void swar_update_alu(int &best, uint32_t x, uint32_t w, int pos)
{
    asm volatile(
        "{ .reg .pred p; .reg .b32 t, d;\n"
        "  sub.u32 t, %1, %2;\n"
        "  lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"
        "  @!p mov.s32 %0, %3; }"
        : "+r"(best) : "r"(w), "r"(x), "r"(pos));
}
// packed test + update on the FMA pipe, second form: best = best * zero + pos (zero and pos are warp-uniform,
// so the instruction has one vector-register source; nothing to decode afterwards)
__device__ __forceinline__ void swar_update_self(int &best, uint32_t x, uint32_t w, int pos, int zero)
{
    asm volatile(
        "{ .reg .pred p; .reg .b32 t, d;\n"
        "  sub.u32 t, %1, %2;\n"
        "  lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"
        "  @!p mad.lo.s32 %0, %0, %4, %3; }"
        : "+r"(best) : "r"(w), "r"(x), "r"(pos), "r"(zero));
}
__device__ __forceinline__ constexpr bool update_on_fma(int r)             // the x-encoded form (decoded after the scan)
{
    return KP_UPDATE_MODE == 2 || (KP_UPDATE_MODE == 1 && (r & 1));
}
__device__ __forceinline__ constexpr bool update_self(int r)
{
    return KP_UPDATE_MODE == 3 || (KP_UPDATE_MODE == 4 && (r & 1));
}
__device__ __forceinline__ void swar_step(int &best, uint32_t x, uint32_t w, int pos, int one, int r)
{
    if (update_on_fma(r)) swar_update(best, x, w, pos, one);
    else if (update_self(r)) swar_update_self(best, x, w, pos, one - 1);
    else swar_update_alu(best, x, w, pos);
}

template <int THREADS, int R, int NSPLIT>
__global__ void __launch_bounds__(THREADS, (NSPLIT > 1 && R < 8) ? 6 : 3)
place_kernel(const PlaceParams p)
{
    constexpr int W = THREADS / 32;
    constexpr int PODW = W / NSPLIT;                         // pod-warps per CTA
    constexpr int CHUNK = 32 * R * PODW;                     // pods per CTA
    static_assert(PODW * NSPLIT == W, "warps must factor into pod-warps x node splits");

    __shared__ __align__(128) unsigned char tiles[2][kTileBytes];
    __shared__ uint64_t bars[2];
    __shared__ int comb[NSPLIT > 1 ? NSPLIT : 1][NSPLIT > 1 ? CHUNK : 1];
    __shared__ int s_last;
    __shared__ __align__(16) double s_uniq[2][kUniqStage];
    __shared__ uint64_t ubar;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int podw = warp / NSPLIT, split = warp % NSPLIT;
    const long long slot0 = (long long)blockIdx.x * CHUNK;

    // ---- before the node plan is needed: barrier set-up and the pods' requests ----
    // (launched with programmatic dependent launch, this part overlaps the tail of the plan)
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&ubar, 1);
        mbar_fence_init();
    }
    constexpr int ITEMS = NSPLIT > 1 ? (2 * CHUNK + THREADS - 1) / THREADS : 1;
    double pre_req[ITEMS];
    const bool prefetched = NSPLIT > 1 && p.uniform_strategy >= 0;       // slot == pod: requests need no plan data
    if (prefetched) {
        const bool paced = p.flow != nullptr;
        if (paced) {
            if (tid == 0) {
                const unsigned need = blockIdx.x > p.flow_window ? blockIdx.x - p.flow_window : 0u;
                // pacing is only a hint: give up after ~1 ms rather than depend on CTA dispatch order
                for (int spin = 0; spin < 16384 && *reinterpret_cast<volatile unsigned *>(p.flow) < need; ++spin)
                    __nanosleep(64);
            }
            __syncthreads();
        }
        bool arrived = true;
#pragma unroll
        for (int e = 0; e < ITEMS; ++e) {
            const int item = tid + e * THREADS;
            const int q = item % CHUNK, dim = item / CHUNK;
            const long long pod = slot0 + q;
            pre_req[e] = 0.0;
            if (item < 2 * CHUNK && pod < p.n_pods) {
                const long long raw = dim == 0 ? p.req_cpu[pod] : p.req_mem[pod];
                pre_req[e] = dim == 0 ? req_cores(raw) : req_gib(raw);
                arrived = arrived && raw != kFlowNever;
            }
        }
        if (paced) {
            // the vote consumes every loaded value, so passing the barrier means the rows are here
            if (!__syncthreads_and(arrived)) p.flow[2] = 1u;      // never true for int64 requests that mean anything
            if (tid == 0) atomicAdd(p.flow, 1u);
        }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");                   // the plan kernels have completed

    int s = p.uniform_strategy;
    long long gbegin = 0, gcount = p.n_pods;
    if (s < 0) {
        long long begin[kMaxStrategies];
        group_begins(p.hdr, CHUNK, begin);
        s = 0;
#pragma unroll
        for (int g = 1; g < kMaxStrategies; ++g) if (slot0 >= begin[g]) s = g;
        gbegin = begin[s];
        gcount = p.hdr->group_count[s];
    }
    if (slot0 - gbegin >= gcount) return;                    // CTA-uniform

    const bool small = p.small_table_len > 0;
    const int nvalid4 = small ? p.small_table_len : ((int)p.hdr->nvalid[s] + 3) & ~3;
    const bool swar = small || p.hdr->use_swar != 0;
    const int da = small ? p.n_nodes : (int)p.hdr->distinct[0];
    const int db = small ? p.n_nodes : (int)p.hdr->distinct[1];
    const int tile_nodes = swar ? kTileNodesPacked : kTileNodesWide;
    const int elem = swar ? 4 : 8;
    const int tiles_per_seg = swar ? (p.tiles_per_seg_wide + 1) / 2 : p.tiles_per_seg_wide;
    const int tile_shift = swar ? 10 : 9;                            // log2(tile_nodes): a shift, not a run-time division
    static_assert(kTileNodesPacked == 1 << 10 && kTileNodesWide == 1 << 9, "tile sizes are powers of two");
    const int ntiles_all = (nvalid4 + tile_nodes - 1) >> tile_shift;
    const int t_lo = min(ntiles_all, (int)blockIdx.y * tiles_per_seg);          // this CTA's tiles [t_lo, t_hi)
    const int t_hi = min(ntiles_all, t_lo + tiles_per_seg);
    const unsigned char *table = swar ? reinterpret_cast<const unsigned char *>(p.packed + (size_t)s * p.npad)
                                      : reinterpret_cast<const unsigned char *>(p.ab + (size_t)s * p.npad);

    // sorted thresholds (when both arrays fit shared memory) and the first tile on their way
    // as bulk copies while the threads get to the ranking step
    auto tile_count = [&](int t) { return min(tile_nodes, nvalid4 - t * tile_nodes); };
    const double *ua = p.uniq, *ub = p.uniq + p.npad;
    const uint32_t ua_bytes = (uint32_t)((da + 1) & ~1) * 8u, ub_bytes = (uint32_t)((db + 1) & ~1) * 8u;
    const bool stage_uniq = da <= kUniqStage && db <= kUniqStage && ua_bytes + ub_bytes > 0;
    __syncthreads();                                         // barrier objects initialised
    if (tid == 0) {
        if (stage_uniq) {
            mbar_expect_tx(&ubar, ua_bytes + ub_bytes);
            if (ua_bytes) bulk_load(s_uniq[0], ua, ua_bytes, &ubar);
            if (ub_bytes) bulk_load(s_uniq[1], ub, ub_bytes, &ubar);
        }
        if (t_hi > t_lo) {
            const int t = t_hi - 1;
            const uint32_t bytes = (uint32_t)(tile_count(t) * elem);
            mbar_expect_tx(&bars[0], bytes);
            bulk_load(tiles[0], table + (size_t)t * tile_nodes * elem, bytes, &bars[0]);
        }
    }

    // ---- pod prologue: requests -> threshold ranks (exact, see kp_plan.cuh) ----
    if (stage_uniq) {
        mbar_wait(&ubar, 0);
        ua = s_uniq[0];
        ub = s_uniq[1];
    }
    uint32_t ra[R], rb[R];
    bool live[R];
    if (NSPLIT > 1) {
        // the NSPLIT warps that share a pod would each repeat its division + binary searches:
        // compute every (pod, dimension) rank once per CTA and hand it over in shared memory
#pragma unroll
        for (int e = 0; e < ITEMS; ++e) {
            const int item = tid + e * THREADS;
            if (item >= 2 * CHUNK) break;
            const int q = item % CHUNK, dim = item / CHUNK;
            const long long slot = slot0 + q;
            uint32_t rk = 0;
            if (slot - gbegin < gcount) {
                double req = pre_req[e];
                if (!prefetched) {
                    const long long pod = p.perm ? (long long)p.perm[slot] : slot;
                    req = dim == 0 ? req_cores(p.req_cpu[pod]) : req_gib(p.req_mem[pod]);
                }
                rk = dim == 0 ? (uint32_t)lower_bound_lt(ua, da, req) : (uint32_t)lower_bound_lt(ub, db, req);
            }
            comb[dim][q] = (int)rk;                          // comb is free until the merge step
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int q = (podw * R + r) * 32 + lane;
        const long long slot = slot0 + q;
        live[r] = slot - gbegin < gcount;
        ra[r] = rb[r] = 0;
        if (NSPLIT > 1) {
            ra[r] = (uint32_t)comb[0][q];
            rb[r] = (uint32_t)comb[1][q];
        } else if (live[r]) {
            const long long pod = p.perm ? (long long)p.perm[slot] : slot;
            ra[r] = (uint32_t)lower_bound_lt(ua, da, req_cores(p.req_cpu[pod]));
            rb[r] = (uint32_t)lower_bound_lt(ub, db, req_gib(p.req_mem[pod]));
        }
    }
    if (NSPLIT > 1) __syncthreads();                         // comb fully read before the merge step reuses it

    int best[R];
    if (swar) {
        uint32_t x[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            x[r] = live[r] ? ((rb[r] + 1u) << 16) | (ra[r] + 1u) : kSwarDeadPod;
            best[r] = update_on_fma(r) ? (int)x[r] - 1 : -1;  // FMA form holds pos + x; pos = -1
        }
        int it = 0;
        for (int t = t_hi - 1; t >= t_lo; --t, ++it) {
            const int buf = it & 1;
            if (tid == 0 && t > t_lo) {
                const uint32_t bytes = (uint32_t)(tile_count(t - 1) * elem);
                mbar_expect_tx(&bars[buf ^ 1], bytes);
                bulk_load(tiles[buf ^ 1], table + (size_t)(t - 1) * tile_nodes * elem, bytes, &bars[buf ^ 1]);
            }
            mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
            const uint4 *tp = reinterpret_cast<const uint4 *>(tiles[buf]);
            const int base = t * tile_nodes;
            const int nquad = tile_count(t) >> 2;
            const int q0 = NSPLIT > 1 ? (nquad * split) / NSPLIT : 0;
            const int q1 = NSPLIT > 1 ? (nquad * (split + 1)) / NSPLIT : nquad;
#pragma unroll(R >= 8 ? 1 : 2)
            for (int i = q1 - 1; i >= q0; --i) {
                const uint4 v = tp[i];                       // four nodes, same address in every lane
                const int pos = base + 4 * i;
#pragma unroll
                for (int r = 0; r < R; ++r) swar_step(best[r], x[r], v.w, pos + 3, p.one, r);
#pragma unroll
                for (int r = 0; r < R; ++r) swar_step(best[r], x[r], v.z, pos + 2, p.one, r);
#pragma unroll
                for (int r = 0; r < R; ++r) swar_step(best[r], x[r], v.y, pos + 1, p.one, r);
#pragma unroll
                for (int r = 0; r < R; ++r) swar_step(best[r], x[r], v.x, pos, p.one, r);
            }
            __syncthreads();                                 // buffer free for the next prefetch
        }
#pragma unroll
        for (int r = 0; r < R; ++r) if (update_on_fma(r)) best[r] -= (int)x[r];
    } else {
        int a[R], b[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            a[r] = live[r] ? (int)ra[r] : 0x7fffffff;
            b[r] = live[r] ? (int)rb[r] : 0x7fffffff;
            best[r] = -1;
        }
        int it = 0;
        for (int t = t_hi - 1; t >= t_lo; --t, ++it) {
            const int buf = it & 1;
            if (tid == 0 && t > t_lo) {
                const uint32_t bytes = (uint32_t)(tile_count(t - 1) * elem);
                mbar_expect_tx(&bars[buf ^ 1], bytes);
                bulk_load(tiles[buf ^ 1], table + (size_t)(t - 1) * tile_nodes * elem, bytes, &bars[buf ^ 1]);
            }
            mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
            const uint4 *tp = reinterpret_cast<const uint4 *>(tiles[buf]);
            const int base = t * tile_nodes;
            const int npair = tile_count(t) >> 1;
            const int q0 = NSPLIT > 1 ? (npair * split) / NSPLIT : 0;
            const int q1 = NSPLIT > 1 ? (npair * (split + 1)) / NSPLIT : npair;
#pragma unroll 2
            for (int i = q1 - 1; i >= q0; --i) {
                const uint4 v = tp[i];                       // two nodes
                const int pos = base + 2 * i;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (a[r] <= (int)v.z && b[r] <= (int)v.w) best[r] = pos + 1;
                    if (a[r] <= (int)v.x && b[r] <= (int)v.y) best[r] = pos;
                }
            }
            __syncthreads();
        }
    }

    // ---- merge the node splits of this CTA (minimum rank position wins) ----
    if (NSPLIT > 1) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            comb[split][(podw * R + r) * 32 + lane] = best[r] >= 0 ? best[r] : 0x7fffffff;
        __syncthreads();
        if (split == 0) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                int m = 0x7fffffff;
#pragma unroll
                for (int q = 0; q < NSPLIT; ++q) m = min(m, comb[q][(podw * R + r) * 32 + lane]);
                best[r] = m == 0x7fffffff ? -1 : m;
            }
        }
    }

    // ---- merge the node segments (grid.y) through global memory ----
    if (gridDim.y > 1) {
        if (split == 0) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (live[r] && best[r] >= 0)
                    atomicMin(p.best_pos + slot0 + (podw * R + r) * 32 + lane, (uint32_t)best[r]);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = atomicAdd(p.chunk_done + blockIdx.x, 1u) == gridDim.y - 1;
        __syncthreads();
        if (!s_last) return;                                 // another segment's CTA will finish this chunk
        __threadfence();
        if (split == 0) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint32_t bp = *reinterpret_cast<volatile uint32_t *>(p.best_pos + slot0 + (podw * R + r) * 32 + lane);
                best[r] = bp == kNoPos ? -1 : (int)bp;
            }
        }
    }

    // ---- epilogue: rank position -> node index and score ----
    if (split != 0) return;
    const uint32_t *order = p.order + (size_t)s * p.npad;
    const double *score = p.score_by_pos + (size_t)s * p.npad;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (!live[r]) continue;
        const long long slot = slot0 + (podw * R + r) * 32 + lane;
        const long long pod = p.perm ? (long long)p.perm[slot] : slot;
        long long idx = -1;
        double sc = -1.0;
        if (best[r] >= 0) {
            idx = order[best[r]];
            sc = score[best[r]];
        }
        p.out_idx[pod] = idx;
        p.out_score[pod] = sc;
        if (p.out_gpos) p.out_gpos[pod] = best[r] >= 0 ? p.gpos_by_pos[(size_t)s * p.gpos_stride + best[r]] : 0xffffffffu;
    }
    if (p.flow && tid == 0) {                                // flow control off again for the next launch
        __threadfence();
        if (atomicAdd(p.flow + 1, 1u) == gridDim.x - 1) {
            p.flow[0] = 0u;
            p.flow[1] = 0u;
        }
    }
}


// ---------------------------------------------------------------------------------------
// place_small_kernel: small tables (the one-launch plan, <= 8192 nodes) -- NODES in registers
// ---------------------------------------------------------------------------------------
// place_kernel keeps pods in registers and broadcasts nodes from shared memory; its fixed costs
// (a CTA-wide prologue, tile staging, barriers, 11 loop instructions per four nodes spread over
// only R = 4 pods per lane) are what a 1 000-node table cannot amortise: 14.1 M warp instructions
// for 9.4 M useful on config 2.  Here the roles are swapped.  A warp holds a slice of the
// best-first table in registers -- NPT packed words per lane, four consecutive table positions per
// 16-byte load -- and its pods are broadcast one after the other: per pod x node pair the same
// three instructions (subtract, LOP3 with predicate output, predicated update; the update
// alternates between the ALU and the FMA pipe from pod to pod), the position is an immediate
// operand (no index arithmetic, no loop), and one REDUX.MIN over the lanes per pod yields the
// best-ranked fitting node of the slice.  Per pod: 3 NPT + ~8 instructions.
//
// Warps are autonomous: each takes at most 32 pods at a time (lane = pod for the request ->
// rank prologue and for the position -> node / score epilogue, coalesced both ways), so any number
// of pods splits evenly over all resident warps, no barrier follows the start-up, and warps in
// their prologue or epilogue overlap with warps that scan.  Tables longer than 32 NPT are
// scanned slice by slice (re-loaded from L1/L2 per slice), the running minimum kept per pod.
#ifndef KP_SMALL_UPDATE
#define KP_SMALL_UPDATE 1         // candidate update: 0 = alternate VIMNMX / IMAD per pod, 1 = IMAD for every pod, 2 = VIMNMX for every pod
#endif
constexpr int kSmallWarps = 8;
constexpr int kSmallThreads = 32 * kSmallWarps;
constexpr int kSmallUniqStage = 2048;                  // sorted thresholds staged in shared memory up to this many nodes
constexpr uint32_t kSmallNone = 0x7fffff00u;           // "no node": above every position, survives + lane

// The update writes the position (an immediate) over `best`.  Both forms read `best` on purpose: a plain
// predicated move of a constant gets its constant hoisted into a register by ptxas (one register per node
// slot), and `one * imm` is hoisted the same way.  FMA == true: best = best * zero + imm (IMAD, FMA pipe);
// FMA == false: best = min(best, imm) (VIMNMX, ALU pipe; positions only fall during the scan); `zero` is a run-time 0.
// The subtraction is written as x * (-1) + w with a run-time -1 so that it issues as IMAD on the FMA pipe:
// per pair the FMA pipe then carries 1.5 instructions (subtract, every other update) and the ALU pipe 1.5
// (test, every other update) -- the two 16-lane pipes are the issue bound.
template <bool FMA, int IMM>
__device__ __forceinline__ void small_step(uint32_t &best, uint32_t x, uint32_t w, uint32_t zero, uint32_t minus1)
{
    if (FMA)
        asm volatile("{ .reg .pred p; .reg .b32 t, d; mad.lo.u32 t, %2, %5, %1; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"
                     "  @!p mad.lo.u32 %0, %0, %3, %4; }"
                     : "+r"(best) : "r"(w), "r"(x), "r"(zero), "n"(IMM), "r"(minus1));
    else
        asm volatile("{ .reg .pred p; .reg .b32 t, d; mad.lo.u32 t, %2, %4, %1; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0;\n"
                     "  @!p min.u32 %0, %0, %3; }"
                     : "+r"(best) : "r"(w), "r"(x), "n"(IMM), "r"(minus1));
}

template <int NPT, int K, int NP>
struct SmallScan {
    // nodes K-1 .. 0 of the slice against NP pods: worst -> best, "fits" overwrites
    static __device__ __forceinline__ void run(uint32_t (&b)[NP], const uint32_t (&x)[NP], const uint32_t (&w)[NPT], uint32_t zero,
                                               uint32_t minus1)
    {
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            if (KP_SMALL_UPDATE == 1 || (KP_SMALL_UPDATE == 0 && (q & 1)))
                small_step<true, 128 * ((K - 1) / 4) + (K - 1) % 4>(b[q], x[q], w[K - 1], zero, minus1);
            else
                small_step<false, 128 * ((K - 1) / 4) + (K - 1) % 4>(b[q], x[q], w[K - 1], zero, minus1);
        }
        SmallScan<NPT, K - 1, NP>::run(b, x, w, zero, minus1);
    }
};
template <int NPT, int NP>
struct SmallScan<NPT, 0, NP> {
    static __device__ __forceinline__ void run(uint32_t (&)[NP], const uint32_t (&)[NP], const uint32_t (&)[NPT], uint32_t, uint32_t) {}
};

// NP pods (lanes first .. first + NP - 1 of the pass hold their requests and running minima) against the slice in w
template <int NPT, int NP>
__device__ __forceinline__ void small_group(const uint32_t (&w)[NPT], uint32_t xmine, uint32_t &mine, int first, int lane,
                                            uint32_t slice_pos, uint32_t zero, uint32_t minus1)
{
    uint32_t x[NP], b[NP];
#pragma unroll
    for (int q = 0; q < NP; ++q) {
        x[q] = __shfl_sync(0xffffffffu, xmine, first + q);
        b[q] = kSmallNone;
    }
    SmallScan<NPT, NPT, NP>::run(b, x, w, zero, minus1);
#pragma unroll
    for (int q = 0; q < NP; ++q) {
        const uint32_t r = __reduce_min_sync(0xffffffffu, b[q] + slice_pos);
        if (lane == first + q) mine = min(mine, r);
    }
}

// two searches of the same depth side by side (independent load chains): #{a[i] < xa} and #{b[i] < xb}
__device__ __forceinline__ void lower_bound_lt_pair(const double *a, const double *b, int n, double xa, double xb,
                                                    uint32_t &ra, uint32_t &rb)
{
    int la = 0, lb = 0;
    for (int step = 1 << (31 - __clz(n | 1)); step > 0; step >>= 1) {
        const int na = la + step, nb = lb + step;
        const double va = a[min(na, n) - 1], vb = b[min(nb, n) - 1];
        if (na <= n && va < xa) la = na;
        if (nb <= n && vb < xb) lb = nb;
    }
    ra = (uint32_t)la;
    rb = (uint32_t)lb;
}

// w[4 j .. 4 j + 3] = table[slice + 128 j + 4 lane .. + 3]: one 16-byte load per four nodes.  The table is
// padded (with entries that fit nothing) far enough that no bounds check is needed.
template <int NPT>
__device__ __forceinline__ void small_load_slice(uint32_t (&w)[NPT], const uint32_t *table, int slice, int lane)
{
    static_assert(NPT % 4 == 0, "node words are loaded four at a time");
    const uint4 *src = reinterpret_cast<const uint4 *>(table + (size_t)slice * 32 * NPT) + lane;
#pragma unroll
    for (int j = 0; j < NPT / 4; ++j) {
        const uint4 v = src[32 * j];
        w[4 * j] = v.x; w[4 * j + 1] = v.y; w[4 * j + 2] = v.z; w[4 * j + 3] = v.w;
    }
}

#ifndef KP_SMALL_GROUP
#define KP_SMALL_GROUP 4          // pods scanned together (4 or 8; measured equal on config 2): the per-group overhead is shared by this many pods
#endif
#ifndef KP_SMALL_OCC32
#define KP_SMALL_OCC32 3          // resident CTAs per SM asked of the compiler for NPT <= 32
#endif
constexpr int small_occupancy(int npt) { return npt <= 32 ? KP_SMALL_OCC32 : (npt <= 40 ? 3 : 2); }

template <int NPT>
__global__ void __launch_bounds__(kSmallThreads, small_occupancy(NPT))
place_small_kernel(const PlaceParams p)
{
    __shared__ __align__(16) double s_uniq[2][kSmallUniqStage];
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned gwarp = blockIdx.x * kSmallWarps + (tid >> 5), nwarps = gridDim.x * kSmallWarps;
    const int n = p.n_nodes;                                  // thresholds: n sorted values per dimension (duplicates kept)
    const int nslices = (p.small_table_len + 32 * NPT - 1) / (32 * NPT);   // best-first table entries to scan, in slices
    const bool uniform = p.uniform_strategy >= 0;

    // this warp's pods.  One strategy: a contiguous, balanced share of all pods.  Per-pod strategies: 32-slot
    // batches of the strategy-grouped slot space, strided over the warps.
    long long lo, hi;
    if (uniform) {
        lo = p.share_q * gwarp + min(gwarp, (unsigned)p.share_r);
        hi = lo + p.share_q + (gwarp < (unsigned)p.share_r ? 1 : 0);
    } else {
        lo = (long long)gwarp * 32;
        hi = ((p.n_pods + 31) / 32 + (kMaxStrategies - 1)) * 32;
    }
    const long long stride = uniform ? 32 : (long long)nwarps * 32;

    // before the node plan is needed (programmatic dependent launch): the first requests
    long long raw_c = 0, raw_m = 0;
    bool pre = uniform && lo + lane < hi;
    if (pre) { raw_c = p.req_cpu[lo + lane]; raw_m = p.req_mem[lo + lane]; }
    asm volatile("griddepcontrol.wait;" ::: "memory");        // the plan kernel has completed

    uint32_t w[NPT];
    int loaded = -1;                                          // strategy * 8 + slice held in w
    if (uniform && nslices == 1) {                            // the whole table: on its way before anything else
        small_load_slice<NPT>(w, p.packed + (size_t)p.uniform_strategy * p.npad, 0, lane);
        loaded = p.uniform_strategy * 8;
    }
    // sorted thresholds: in shared memory when they fit (searched at shared-memory latency; every warp of the
    // kernel is in its first prologue at the same time, so that latency is exposed), else from L1 / L2
    const bool staged = n <= kSmallUniqStage;
    if (staged) {
        for (int i = tid; i < n; i += kSmallThreads) {
            s_uniq[0][i] = p.uniq[i];
            s_uniq[1][i] = p.uniq[p.npad + i];
        }
        __syncthreads();                                      // the only block-wide barrier
    }
    const double *ua = staged ? s_uniq[0] : p.uniq, *ub = staged ? s_uniq[1] : p.uniq + p.npad;
    const uint32_t zero = (uint32_t)p.one - 1u;               // 0, but not to the compiler (see small_step)
    const uint32_t minus1 = zero - (uint32_t)p.one;           // -1

    for (long long s0 = lo; s0 < hi; s0 += stride) {
        int s = p.uniform_strategy, count;
        if (uniform) {
            count = (int)min((long long)32, hi - s0);
        } else {
            long long gb[kMaxStrategies];
            group_begins(p.hdr, 32, gb);
            s = 0;
#pragma unroll
            for (int g = 1; g < kMaxStrategies; ++g) if (s0 >= gb[g]) s = g;
            const long long left = (long long)p.hdr->group_count[s] - (s0 - gb[s]);
            if (left <= 0) continue;                          // warp-uniform
            count = (int)min((long long)32, left);
        }
        // ---- prologue: lane = pod; request -> packed threshold ranks ----
        uint32_t xmine = kSwarDeadPod;
        if (lane < count) {
            if (!pre) {
                const long long pod = p.perm ? (long long)p.perm[s0 + lane] : s0 + lane;
                raw_c = p.req_cpu[pod];
                raw_m = p.req_mem[pod];
            }
            uint32_t ra, rb;
            lower_bound_lt_pair(ua, ub, n, req_cores(raw_c), req_gib(raw_m), ra, rb);
            xmine = ((rb + 1u) << 16) | (ra + 1u);
        }
        pre = false;
        uint32_t mine = kSmallNone;
        const uint32_t *table = p.packed + (size_t)s * p.npad;
        for (int sl = nslices - 1; sl >= 0; --sl) {
            if (loaded != s * 8 + sl) {                       // one slice: loaded once per strategy, not per pass
                small_load_slice<NPT>(w, table, sl, lane);
                loaded = s * 8 + sl;
            }
            const uint32_t slice_pos = (uint32_t)(sl * 32 * NPT + 4 * lane);
#pragma unroll 1
            for (int g = 0; g + KP_SMALL_GROUP <= count; g += KP_SMALL_GROUP)
                small_group<NPT, KP_SMALL_GROUP>(w, xmine, mine, g, lane, slice_pos, zero, minus1);
            if (KP_SMALL_GROUP == 8 && (count & 4)) small_group<NPT, 4>(w, xmine, mine, count & ~7, lane, slice_pos, zero, minus1);
            if (count & 2) small_group<NPT, 2>(w, xmine, mine, count & ~3, lane, slice_pos, zero, minus1);
            if (count & 1) small_group<NPT, 1>(w, xmine, mine, count & ~1, lane, slice_pos, zero, minus1);
        }
        // ---- epilogue: lane = pod; rank position -> node index and score ----
        if (lane < count) {
            const long long pod = p.perm ? (long long)p.perm[s0 + lane] : s0 + lane;
            long long idx = -1;
            double sc = -1.0;
            if (mine < kSmallNone) {
                idx = p.order[(size_t)s * p.npad + mine];
                sc = p.score_by_pos[(size_t)s * p.npad + mine];
            }
            p.out_idx[pod] = idx;
            p.out_score[pod] = sc;
            if (p.out_gpos) p.out_gpos[pod] = mine < kSmallNone ? p.gpos_by_pos[(size_t)s * p.gpos_stride + mine] : 0xffffffffu;
        }
    }
}

}  // namespace kp

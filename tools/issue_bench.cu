// Micro-benchmark: which instruction mix for the pod x node inner loop issues fastest on
// sm_100a?  Every variant does the same logical work per pair (does pod fit node in both
// dimensions? if so remember the node) over a 4096-node shared-memory tile, R pods per
// thread in registers, nodes broadcast to all lanes -- exactly the shape of place_kernel.
//
//   A   2 x ISETP (32-bit rank compares) + SEL                      -> 3 ALU-pipe ops
//   C   SWAR: packed 15-bit ranks with guard bits: IMAD.IADD (FMA pipe) + LOP3->P + SEL
//   D   2 x DSETP (fp64 thresholds, no rank compression) + SEL      -> 2 FP64-pipe + 1 ALU
//   E   DSETP + ISETP + SEL
//   M   half the pods per thread use C, half use D (spreads over ALU / FMA / FP64 pipes)
//   N   half A, half D
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/issue_bench tools/issue_bench.cu
// Run:   tools/issue_bench            (prints pairs/s and pairs/clk/SM per variant)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int TILE = 4096;
constexpr int THREADS = 256;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

struct Tile {
    uint32_t ra[TILE], rb[TILE];   // 32-bit ranks
    uint32_t packed[TILE];         // (rb | guard) << 16 | (ra | guard), 15-bit ranks
    double ta[TILE], tb[TILE];     // fp64 thresholds
};

__device__ __forceinline__ void upd_int(int &best, int ra, int rb, int A, int B, int pos)
{
    if (ra <= A && rb <= B) best = pos;
}
__device__ __forceinline__ void upd_swar(int &best, uint32_t x, uint32_t w, int pos)
{
    // t = w - x; guard bits (15, 31) survive iff neither field borrowed, i.e. both ranks fit
    asm volatile("{ .reg .pred p; .reg .b32 t, d; sub.u32 t, %1, %2; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0; @!p mov.s32 %0, %3; }"
                 : "+r"(best) : "r"(w), "r"(x), "r"(pos));
}
// same test, but the update is a guarded IMAD (FMA pipe): best_enc = pos * one + x; decode pos = best_enc - x
__device__ __forceinline__ void upd_swar_fma(int &best_enc, uint32_t x, uint32_t w, int pos, int one)
{
    asm volatile("{ .reg .pred p; .reg .b32 t, d; sub.u32 t, %1, %2; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0; @!p mad.lo.s32 %0, %3, %4, %2; }"
                 : "+r"(best_enc) : "r"(w), "r"(x), "r"(pos), "r"(one));
}
__device__ __forceinline__ void upd_f64(int &best, double a, double b, double A, double B, int pos)
{
    if (a <= A && b <= B) best = pos;
}
__device__ __forceinline__ void upd_mixed(int &best, double a, int rb, double A, int B, int pos)
{
    if (a <= A && rb <= B) best = pos;
}

template <char V, int R>
__global__ void __launch_bounds__(THREADS) bench_kernel(const Tile *__restrict__ g, int iters, int *__restrict__ out,
                                                        unsigned long long *__restrict__ clk, int one)
{
    extern __shared__ __align__(16) unsigned char smem[];
    // layout depends on variant
    uint2 *s_ab = reinterpret_cast<uint2 *>(smem);                       // A: {ra, rb}
    uint32_t *s_pk = reinterpret_cast<uint32_t *>(smem);                 // C: packed
    double2 *s_t = reinterpret_cast<double2 *>(smem);                    // D: {ta, tb}
    // E: {ta (double), rb (int), pad}
    for (int i = threadIdx.x; i < TILE; i += THREADS) {
        if (V == 'A') s_ab[i] = make_uint2(g->ra[i], g->rb[i]);
        if (V == 'C' || V == 'B' || V == 'F') s_pk[i] = g->packed[i];
        if (V == 'D') s_t[i] = make_double2(g->ta[i], g->tb[i]);
        if (V == 'E') { s_t[i].x = g->ta[i]; reinterpret_cast<int *>(&s_t[i].y)[0] = (int)g->rb[i]; }
        if (V == 'M') { s_t[i] = make_double2(g->ta[i], g->tb[i]); reinterpret_cast<uint32_t *>(smem + TILE * 16)[i] = g->packed[i]; }
        if (V == 'N') { s_t[i] = make_double2(g->ta[i], g->tb[i]); reinterpret_cast<uint2 *>(smem + TILE * 16)[i] = make_uint2(g->ra[i], g->rb[i]); }
    }
    __syncthreads();
    const int gtid = blockIdx.x * THREADS + threadIdx.x;
    int best[R];
    int ra[R], rb[R];
    uint32_t x[R];
    double a[R], b[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const uint32_t h = (uint32_t)(gtid * R + r) * 2654435761u;
        ra[r] = (h >> 3) % 3000; rb[r] = (h >> 17) % 3000;
        x[r] = ((uint32_t)rb[r] << 16) | (uint32_t)ra[r];
        a[r] = (double)ra[r] + 0.5; b[r] = (double)rb[r] + 0.5;
        best[r] = -1;
    }
    const unsigned long long c0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (V == 'A') {
            const uint4 *tp = reinterpret_cast<const uint4 *>(s_ab);
#pragma unroll 2
            for (int i = TILE / 2 - 1; i >= 0; --i) {
                const uint4 v = tp[i];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    upd_int(best[r], ra[r], rb[r], (int)v.z, (int)v.w, 2 * i + 1);
                    upd_int(best[r], ra[r], rb[r], (int)v.x, (int)v.y, 2 * i);
                }
            }
        } else if (V == 'C') {
            const uint4 *tp = reinterpret_cast<const uint4 *>(s_pk);
            for (int i = TILE / 4 - 1; i >= 0; --i) {
                const uint4 v = tp[i];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    upd_swar(best[r], x[r], v.w, 4 * i + 3);
                    upd_swar(best[r], x[r], v.z, 4 * i + 2);
                    upd_swar(best[r], x[r], v.y, 4 * i + 1);
                    upd_swar(best[r], x[r], v.x, 4 * i);
                }
            }
        } else if (V == 'B' || V == 'F') {
            // B: half the pods update with SEL (ALU), half with a guarded IMAD (FMA); F: all IMAD
            const uint4 *tp = reinterpret_cast<const uint4 *>(s_pk);
            for (int i = TILE / 4 - 1; i >= 0; --i) {
                const uint4 v = tp[i];
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 3; k >= 0; --k) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        if (V == 'F' || (r & 1)) upd_swar_fma(best[r], x[r], w[k], 4 * i + k, one);
                        else upd_swar(best[r], x[r], w[k], 4 * i + k);
                    }
                }
            }
        } else if (V == 'D') {
#pragma unroll 2
            for (int i = TILE - 1; i >= 0; --i) {
                const double2 v = s_t[i];
#pragma unroll
                for (int r = 0; r < R; ++r) upd_f64(best[r], a[r], b[r], v.x, v.y, i);
            }
        } else if (V == 'E') {
#pragma unroll 2
            for (int i = TILE - 1; i >= 0; --i) {
                const double2 v = s_t[i];
                const int B = reinterpret_cast<const int *>(&v.y)[0];
#pragma unroll
                for (int r = 0; r < R; ++r) upd_mixed(best[r], a[r], rb[r], v.x, B, i);
            }
        } else if (V == 'M') {
            const uint32_t *pk = reinterpret_cast<const uint32_t *>(smem + TILE * 16);
#pragma unroll 2
            for (int i = TILE - 1; i >= 0; --i) {
                const double2 v = s_t[i];
                const uint32_t w = pk[i];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (r & 1) upd_f64(best[r], a[r], b[r], v.x, v.y, i);
                    else upd_swar(best[r], x[r], w, i);
                }
            }
        } else if (V == 'N') {
            const uint2 *ab = reinterpret_cast<const uint2 *>(smem + TILE * 16);
#pragma unroll 2
            for (int i = TILE - 1; i >= 0; --i) {
                const double2 v = s_t[i];
                const uint2 w = ab[i];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (r & 1) upd_f64(best[r], a[r], b[r], v.x, v.y, i);
                    else upd_int(best[r], ra[r], rb[r], (int)w.x, (int)w.y, i);
                }
            }
        }
        // keep iterations from being merged: perturb the pods a little
#pragma unroll
        for (int r = 0; r < R; ++r) { ra[r] ^= (best[r] & 1); x[r] ^= (best[r] & 1); a[r] += (double)(best[r] & 1); }
    }
    const unsigned long long c1 = clock64();
    int acc = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) acc += best[r];
    out[gtid] = acc;
    if (threadIdx.x == 0) clk[blockIdx.x] = c1 - c0;
}

template <char V, int R>
static void run(const char *what, const Tile *d_tile, int sms, int ctas_per_sm, int iters)
{
    const size_t smem = (V == 'M' || V == 'N') ? TILE * 16 + TILE * 8 : TILE * 16;
    CK(cudaFuncSetAttribute(bench_kernel<V, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = sms * ctas_per_sm;
    int *out;
    unsigned long long *clk;
    CK(cudaMalloc(&out, sizeof(int) * grid * THREADS));
    CK(cudaMalloc(&clk, sizeof(unsigned long long) * grid));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    bench_kernel<V, R><<<grid, THREADS, smem>>>(d_tile, 2, out, clk, 1);
    CK(cudaDeviceSynchronize());
    float best_ms = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0));
        bench_kernel<V, R><<<grid, THREADS, smem>>>(d_tile, iters, out, clk, 1);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best_ms) best_ms = ms;
    }
    std::vector<unsigned long long> h(grid);
    CK(cudaMemcpy(h.data(), clk, sizeof(unsigned long long) * grid, cudaMemcpyDeviceToHost));
    double cyc = 0;
    for (auto c : h) cyc += (double)c;
    cyc /= grid;
    const double pairs = (double)grid * THREADS * R * TILE * iters;
    const double pairs_per_clk_sm = (double)ctas_per_sm * THREADS * R * TILE * iters / cyc;
    printf("%-34s R=%d ctas/SM=%d  %8.3f ms  %7.3f Tpairs/s  %6.2f pairs/clk/SM  (%.0f MHz eff)\n", what, R,
           ctas_per_sm, best_ms, pairs / best_ms * 1e-9, pairs_per_clk_sm, cyc / (best_ms * 1e3));
    CK(cudaFree(out));
    CK(cudaFree(clk));
}

int main()
{
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    Tile *h = new Tile;
    for (int i = 0; i < TILE; ++i) {
        const uint32_t v = (uint32_t)i * 2246822519u;
        h->ra[i] = (v >> 5) % 3000; h->rb[i] = (v >> 18) % 3000;
        h->packed[i] = ((h->rb[i] | 0x8000u) << 16) | (h->ra[i] | 0x8000u);
        h->ta[i] = (double)h->ra[i] + 0.25; h->tb[i] = (double)h->rb[i] + 0.25;
    }
    Tile *d;
    CK(cudaMalloc(&d, sizeof(Tile)));
    CK(cudaMemcpy(d, h, sizeof(Tile), cudaMemcpyHostToDevice));
    printf("SMs: %d   (issue bound at 3 instr/pair = 42.67 pairs/clk/SM; single 16-lane pipe at 3 = 21.33)\n", sms);
    const int iters = 24;
    for (int occ : {1, 2, 4}) {
        run<'A', 8>("A  ISETP+ISETP+SEL", d, sms, occ, iters);
        run<'C', 8>("C  SWAR IMAD.IADD+LOP3.P+SEL", d, sms, occ, iters);
        run<'B', 8>("B  SWAR, half SEL half @P IMAD", d, sms, occ, iters);
        run<'F', 8>("F  SWAR, all @P IMAD", d, sms, occ, iters);
        run<'D', 8>("D  DSETP+DSETP+SEL", d, sms, occ, iters);
        run<'E', 8>("E  DSETP+ISETP+SEL", d, sms, occ, iters);
        run<'M', 8>("M  half SWAR, half DSETPx2", d, sms, occ, iters);
        run<'N', 8>("N  half ISETPx2, half DSETPx2", d, sms, occ, iters);
    }
    run<'A', 4>("A  ISETP+ISETP+SEL", d, sms, 4, iters);
    run<'A', 16>("A  ISETP+ISETP+SEL", d, sms, 2, iters);
    run<'C', 16>("C  SWAR", d, sms, 2, iters);
    run<'B', 16>("B  SWAR half/half", d, sms, 2, iters);
    run<'B', 4>("B  SWAR half/half", d, sms, 4, iters);
    run<'B', 6>("B  SWAR half/half", d, sms, 4, iters);
    run<'M', 16>("M  half SWAR, half DSETPx2", d, sms, 2, iters);
    return 0;
}

// Micro-benchmark, second round: can the pod x node test be done in TWO instructions per pair?
// Ranks held as fp16 pairs (exact for ranks <= 2048): HSETP2.GE yields one predicate per dimension,
// and the update is a SEL guarded by one predicate with the other as its select operand.
//   F   (reference) SWAR: IMAD.IADD + LOP3->P + @!P IMAD            3 instructions per pair
//   H   HSETP2.GE P,Q + @P SEL best, pos, best, Q                     2 instructions per pair
//   I   HSETP2.GE P,Q + PLOP3 + @R IMAD                               3 instructions per pair
//   J   HSET2.GE mask + ISETP + @P IMAD                               3 instructions per pair
// Same shape as place_kernel: R pods per lane in registers, nodes broadcast from a shared-memory tile.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/issue_bench2 tools/issue_bench2.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

constexpr int TILE = 4096;
constexpr int THREADS = 256;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void upd_f(int &best_enc, uint32_t x, uint32_t w, int pos, int one)
{
    asm volatile("{ .reg .pred p; .reg .b32 t, d; sub.u32 t, %1, %2; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0; @!p mad.lo.s32 %0, %3, %4, %2; }"
                 : "+r"(best_enc) : "r"(w), "r"(x), "r"(pos), "r"(one));
}
// G: like F, but the update reads `best` (best = best * zero + pos): pos and zero are warp-uniform, so the
// instruction has ONE vector-register source
__device__ __forceinline__ void upd_g(int &best, uint32_t x, uint32_t w, int pos, int zero)
{
    asm volatile("{ .reg .pred p; .reg .b32 t, d; sub.u32 t, %1, %2; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0; @!p mad.lo.s32 %0, %0, %4, %3; }"
                 : "+r"(best) : "r"(w), "r"(x), "r"(pos), "r"(zero));
}
__device__ __forceinline__ void upd_h(int &best, uint32_t x, uint32_t w, int pos)
{
    asm volatile("{ .reg .pred p, q; setp.ge.f16x2 p|q, %1, %2; @p selp.s32 %0, %3, %0, q; }"
                 : "+r"(best) : "r"(w), "r"(x), "r"(pos));
}
__device__ __forceinline__ void upd_i(int &best, uint32_t x, uint32_t w, int pos, int zero)
{
    asm volatile("{ .reg .pred p, q, r; setp.ge.f16x2 p|q, %1, %2; and.pred r, p, q; @r mad.lo.s32 %0, %0, %4, %3; }"
                 : "+r"(best) : "r"(w), "r"(x), "r"(pos), "r"(zero));
}
__device__ __forceinline__ void upd_j(int &best, uint32_t x, uint32_t w, int pos, int zero)
{
    const __half2 a = *reinterpret_cast<const __half2 *>(&w), b = *reinterpret_cast<const __half2 *>(&x);
    if (__hbge2(a, b)) best = best * zero + pos;
}

template <char V, int R>
__global__ void __launch_bounds__(THREADS) bench_kernel(int iters, int *__restrict__ out, int one)
{
    __shared__ __align__(16) uint32_t s_pk[TILE];
    for (int i = threadIdx.x; i < TILE; i += THREADS) {
        const uint32_t v = (uint32_t)i * 2246822519u;
        const uint32_t ra = (v >> 5) % 2000u, rb = (v >> 18) % 2000u;
        if (V == 'F' || V == 'G') s_pk[i] = ((rb | 0x8000u) << 16) | (ra | 0x8000u);
        else {
            const __half2 h = __halves2half2(__ushort2half_rn((unsigned short)ra), __ushort2half_rn((unsigned short)rb));
            s_pk[i] = *reinterpret_cast<const uint32_t *>(&h);
        }
    }
    __syncthreads();
    const int gtid = blockIdx.x * THREADS + threadIdx.x;
    int best[R];
    uint32_t x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const uint32_t h = (uint32_t)(gtid * R + r) * 2654435761u;
        const uint32_t ra = (h >> 3) % 2000u, rb = (h >> 17) % 2000u;
        if (V == 'F' || V == 'G') x[r] = (rb << 16) | ra;
        else {
            const __half2 hh = __halves2half2(__ushort2half_rn((unsigned short)ra), __ushort2half_rn((unsigned short)rb));
            x[r] = *reinterpret_cast<const uint32_t *>(&hh);
        }
        best[r] = -1;
    }
    const int zero = one - 1;
    const uint4 *tp = reinterpret_cast<const uint4 *>(s_pk);
    for (int it = 0; it < iters; ++it) {
        for (int i = TILE / 4 - 1; i >= 0; --i) {
            const uint4 v = tp[i];
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 3; k >= 0; --k) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (V == 'F') upd_f(best[r], x[r], w[k], 4 * i + k, one);
                    if (V == 'G') upd_g(best[r], x[r], w[k], 4 * i + k, zero);
                    if (V == 'H') upd_h(best[r], x[r], w[k], 4 * i + k);
                    if (V == 'I') upd_i(best[r], x[r], w[k], 4 * i + k, zero);
                    if (V == 'J') upd_j(best[r], x[r], w[k], 4 * i + k, zero);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) x[r] ^= (uint32_t)(best[r] & 1) << 10;
    }
    int acc = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) acc += best[r];
    out[gtid] = acc;
}

template <char V, int R>
static void run(const char *what, int sms, int ctas_per_sm, int iters)
{
    const int grid = sms * ctas_per_sm;
    int *out;
    CK(cudaMalloc(&out, sizeof(int) * grid * THREADS));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    bench_kernel<V, R><<<grid, THREADS>>>(2, out, 1);
    CK(cudaDeviceSynchronize());
    float best_ms = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0));
        bench_kernel<V, R><<<grid, THREADS>>>(iters, out, 1);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best_ms) best_ms = ms;
    }
    std::vector<int> h(grid * THREADS);
    CK(cudaMemcpy(h.data(), out, sizeof(int) * grid * THREADS, cudaMemcpyDeviceToHost));
    long long sum = 0;
    for (int v : h) sum += v;
    const double pairs = (double)grid * THREADS * R * TILE * iters;
    printf("%-44s R=%-2d ctas/SM=%d  %8.3f ms  %7.3f Tpairs/s   checksum %lld\n", what, R, ctas_per_sm, best_ms,
           pairs / best_ms * 1e-9, sum);
    CK(cudaFree(out));
}

int main()
{
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    printf("SMs: %d\n", sms);
    const int iters = 24;
    for (int occ : {2, 4}) {
        run<'F', 8>("F  SWAR IMAD.IADD + LOP3.P + @P IMAD (3/pair)", sms, occ, iters);
        run<'G', 8>("G  SWAR, update best = best*zero + pos (1 vector source)", sms, occ, iters);
        run<'H', 8>("H  HSETP2 P,Q + @P SEL ..,Q (2/pair)", sms, occ, iters);
        run<'I', 8>("I  HSETP2 + PLOP3 + @P IMAD (3/pair)", sms, occ, iters);
        run<'J', 8>("J  HSET2 mask + ISETP + @P IMAD (3/pair)", sms, occ, iters);
    }
    run<'G', 16>("G  SWAR, update best = best*zero + pos", sms, 2, iters);
    run<'G', 16>("G  SWAR, update best = best*zero + pos", sms, 3, iters);
    run<'F', 16>("F  SWAR IMAD.IADD + LOP3.P + @P IMAD", sms, 3, iters);
    run<'H', 4>("H  HSETP2 P,Q + @P SEL ..,Q (2/pair)", sms, 4, iters);
    run<'H', 16>("H  HSETP2 P,Q + @P SEL ..,Q (2/pair)", sms, 2, iters);
    return 0;
}

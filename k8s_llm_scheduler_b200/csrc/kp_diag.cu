// Diagnostics: the MEASURED instruction-issue peak of the snapshot inner loop (kp_issue_peak).
//
// place_kernel's bound is instruction issue (SURVEY.md §8(d) regime X): per pod x node pair it
// executes three lane-ops out of shared memory and touches no HBM.  The theoretical issue rate
// (SMs x 4 schedulers x 32 lanes x clock) is not reachable by any real instruction mix; this
// kernel runs the same three instructions per pair as place_kernel (kp_place.cuh: subtract on
// the FMA pipe, LOP3 with predicate output on the ALU pipe, predicated update) on a
// shared-memory tile with nothing else around them -- no plan, no pod prologue, no tile
// staging, no epilogue -- in the two update mixes the real kernel can use (tools/issue_bench.cu
// variants B and F, profiles/r01_issue_bench.txt).  The best rate it reaches is what bench.py
// reports as roofline.peak_measured.  It has no counterpart in the reference.
#include "kp_common.cuh"

namespace kp {

constexpr int kPeakTile = 4096;          // packed nodes in shared memory (16 KB)
constexpr int kPeakThreads = 256;
constexpr int kPeakR = 8;                // pods per lane

__device__ __forceinline__ void peak_upd_alu(int &best, uint32_t x, uint32_t w, int pos)
{
    asm volatile("{ .reg .pred p; .reg .b32 t, d; sub.u32 t, %1, %2; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0; @!p mov.s32 %0, %3; }"
                 : "+r"(best) : "r"(w), "r"(x), "r"(pos));
}
__device__ __forceinline__ void peak_upd_fma(int &best_enc, uint32_t x, uint32_t w, int pos, int one)
{
    asm volatile("{ .reg .pred p; .reg .b32 t, d; sub.u32 t, %1, %2; lop3.b32.or d|p, t, 0x80008000, 0, 0x0c, 0; @!p mad.lo.s32 %0, %3, %4, %2; }"
                 : "+r"(best_enc) : "r"(w), "r"(x), "r"(pos), "r"(one));
}

// ALL_FMA: every pod updates with the guarded IMAD; otherwise pods alternate SEL / IMAD (place_kernel's mode 1)
template <bool ALL_FMA>
__global__ void __launch_bounds__(kPeakThreads)
issue_peak_kernel(int iters, int one, int *__restrict__ sink)
{
    __shared__ __align__(16) uint32_t s_pk[kPeakTile];
    for (int i = threadIdx.x; i < kPeakTile; i += kPeakThreads) {
        const uint32_t v = (uint32_t)i * 2246822519u;
        const uint32_t ra = (v >> 5) % 3000u, rb = (v >> 18) % 3000u;
        s_pk[i] = ((rb | 0x8000u) << 16) | (ra | 0x8000u);
    }
    __syncthreads();
    const int gtid = blockIdx.x * kPeakThreads + threadIdx.x;
    int best[kPeakR];
    uint32_t x[kPeakR];
#pragma unroll
    for (int r = 0; r < kPeakR; ++r) {
        const uint32_t h = (uint32_t)(gtid * kPeakR + r) * 2654435761u;
        x[r] = (((h >> 17) % 3000u) << 16) | ((h >> 3) % 3000u);
        best[r] = -1;
    }
    const uint4 *tp = reinterpret_cast<const uint4 *>(s_pk);
    for (int it = 0; it < iters; ++it) {
        for (int i = kPeakTile / 4 - 1; i >= 0; --i) {
            const uint4 v = tp[i];                           // four nodes, same address in every lane
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 3; k >= 0; --k) {
#pragma unroll
                for (int r = 0; r < kPeakR; ++r) {
                    if (ALL_FMA || (r & 1)) peak_upd_fma(best[r], x[r], w[k], 4 * i + k, one);
                    else peak_upd_alu(best[r], x[r], w[k], 4 * i + k);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < kPeakR; ++r) x[r] ^= (uint32_t)(best[r] & 1);      // keep the passes apart
    }
    int acc = 0;
#pragma unroll
    for (int r = 0; r < kPeakR; ++r) acc += best[r];
    if (acc == 0x7ffffff1) *sink = acc;                      // never true; keeps the work alive
}

template <bool ALL_FMA>
static int time_mix(int grid, int iters, double min_ms, int *sink, cudaStream_t stream, cudaEvent_t e0, cudaEvent_t e1,
                    double *pairs_per_sec)
{
    issue_peak_kernel<ALL_FMA><<<grid, kPeakThreads, 0, stream>>>(2, 1, sink);       // warm-up
    KP_LAUNCH_CHECK("issue_peak_kernel");
    double best = 0.0, spent = 0.0;
    for (int rep = 0; rep < 64 && (rep < 3 || spent < min_ms); ++rep) {
        KP_CUDA(cudaEventRecord(e0, stream));
        issue_peak_kernel<ALL_FMA><<<grid, kPeakThreads, 0, stream>>>(iters, 1, sink);
        KP_LAUNCH_CHECK("issue_peak_kernel");
        KP_CUDA(cudaEventRecord(e1, stream));
        KP_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        KP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        spent += ms;
        const double rate = (double)grid * kPeakThreads * kPeakR * kPeakTile * iters / (ms * 1e-3);
        if (rate > best) best = rate;
    }
    *pairs_per_sec = best;
    return KP_OK;
}

int issue_peak_device(void *scratch, size_t scratch_bytes, double min_ms, cudaStream_t stream, double *pairs_per_sec)
{
    if (!scratch || scratch_bytes < 256 || !pairs_per_sec) {
        set_error("kp_issue_peak: needs 256 bytes of device scratch and an output pointer");
        return KP_ERR_INVALID_ARG;
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    KP_CUDA(cudaEventCreate(&e0));
    if (cudaEventCreate(&e1) != cudaSuccess) { cudaEventDestroy(e0); return cuda_fail(cudaGetLastError(), "cudaEventCreate"); }
    const int grid = sm_count_cached() * 4;                 // 4 resident CTAs per SM: the best point of the sweep
    const int iters = 24;                                   // ~10 ms per launch
    double alt = 0.0, fma = 0.0;
    int st = time_mix<false>(grid, iters, min_ms * 0.5, static_cast<int *>(scratch), stream, e0, e1, &alt);
    if (st == KP_OK) st = time_mix<true>(grid, iters, min_ms * 0.5, static_cast<int *>(scratch), stream, e0, e1, &fma);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (st != KP_OK) return st;
    *pairs_per_sec = alt > fma ? alt : fma;
    return KP_OK;
}

}  // namespace kp

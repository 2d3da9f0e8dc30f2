// Sorting the node plan's key arrays without a library: chunk sort + rank by merge counting.
//
// The plans sort a handful of 64-bit key arrays of n entries (n = nodes, or nodes + trajectory states) -- tens of
// thousands of keys, several arrays at once.  A radix sort of 64-bit keys is ten latency-bound launches per array
// (about 0.1 ms each however few keys there are; profiles/r02_c3_batched_launches.txt), which made the sorts the
// largest part of the plan.  Here every array is sorted by two launches that handle ALL arrays together:
//
//   chunk_sort_kernel   a CTA sorts 2048 (key, index) pairs in shared memory (bitonic network, ties by index, so the
//                       order is the stable order the reference's "first in list order" needs);
//   chunk_rank_kernel   every element finds its final position = its position in its own chunk + the number of
//                       elements of every OTHER chunk that sort before it: one binary search per chunk, eight chunks
//                       side by side (independent load chains), and scatters (key, index) to that position.
//
// Work grows with n^2 / 2048, so this form is for tables up to kRankSortMax entries; larger ones keep the radix sort
// (kp_sort.cu).  Scratch: 12 bytes per padded entry and array.
#include "kp_common.cuh"

namespace kp {

constexpr int kChunk = 2048;                  // entries a CTA sorts in shared memory (two per thread)
constexpr int kChunkThreads = kChunk / 2;

__device__ __forceinline__ bool pair_less(unsigned long long ka, uint32_t ia, unsigned long long kb, uint32_t ib)
{
    return ka < kb || (ka == kb && ia < ib);
}

// grid (chunks, arrays).  keys_in[which[a]][n]; chunk a, c -> ckey / cidx[(a * chunks + c) * kChunk ..)
__global__ void __launch_bounds__(kChunkThreads)
chunk_sort_kernel(const unsigned long long *__restrict__ keys_in, size_t in_stride, RankSortArrays which, int n,
                  unsigned long long *__restrict__ ckey, uint32_t *__restrict__ cidx)
{
    __shared__ unsigned long long sk[kChunk];
    __shared__ uint32_t si[kChunk];
    const int a = blockIdx.y, c = blockIdx.x, t = threadIdx.x;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int l = t + h * kChunkThreads, g = c * kChunk + l;
        sk[l] = g < n ? keys_in[(size_t)which.idx[a] * in_stride + g] : ~0ull;   // padding sorts last (index above every real one)
        si[l] = g < n ? (uint32_t)g : 0xffffffffu;
    }
    __syncthreads();
    // bitonic network: in every step thread t owns one compare-exchange (lower element i, partner i + j)
    for (int k = 2; k <= kChunk; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int i = 2 * j * (t / j) + (t % j);
            const bool up = (i & k) == 0;
            const unsigned long long k0 = sk[i], k1 = sk[i + j];
            const uint32_t i0 = si[i], i1 = si[i + j];
            if (pair_less(k1, i1, k0, i0) == up) { sk[i] = k1; si[i] = i1; sk[i + j] = k0; si[i + j] = i0; }
            __syncthreads();
        }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int l = t + h * kChunkThreads;
        const size_t o = ((size_t)a * gridDim.x + c) * kChunk + l;
        ckey[o] = sk[l];
        cidx[o] = si[l];
    }
}

// grid (ceil(chunks * kChunk / 256), arrays).  Final position of every real element, scattered to kout / vout[a][n].
constexpr int kRankWays = 8;

__global__ void __launch_bounds__(256)
chunk_rank_kernel(const unsigned long long *__restrict__ ckey, const uint32_t *__restrict__ cidx, int chunks, int n, RankSortArrays which,
                  unsigned long long *__restrict__ kout, uint32_t *__restrict__ vout, size_t out_stride)
{
    const int a = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;                   // element of the chunked array
    if (e >= chunks * kChunk) return;
    const unsigned long long *K = ckey + (size_t)a * chunks * kChunk;
    const uint32_t *I = cidx + (size_t)a * chunks * kChunk;
    const unsigned long long ke = K[e];
    const uint32_t ie = I[e];
    if (ie == 0xffffffffu) return;                                          // padding
    const int mine = e / kChunk;
    uint32_t pos = (uint32_t)(e - mine * kChunk);
    for (int c0 = 0; c0 < chunks; c0 += kRankWays) {
        int lo[kRankWays];
#pragma unroll
        for (int w = 0; w < kRankWays; ++w) lo[w] = 0;
        for (int step = kChunk >> 1; step > 0; step >>= 1) {               // #{x in chunk : x < mine} by halving steps
#pragma unroll
            for (int w = 0; w < kRankWays; ++w) {
                const int c = c0 + w;
                if (c < chunks && c != mine) {
                    const size_t at = (size_t)c * kChunk + lo[w] + step - 1;
                    const unsigned long long km = K[at];
                    if (km < ke || (km == ke && I[at] < ie)) lo[w] += step;        // the index is read only on equal keys
                }
            }
        }
#pragma unroll
        for (int w = 0; w < kRankWays; ++w) {
            const int c = c0 + w;
            if (c < chunks && c != mine) {
                // the halving search covers all positions but the last; the last element of the chunk is checked on its own
                const size_t last = (size_t)c * kChunk + kChunk - 1;
                if (lo[w] == kChunk - 1 && (K[last] < ke || (K[last] == ke && I[last] < ie))) lo[w] = kChunk;
                pos += (uint32_t)lo[w];
            }
        }
    }
    kout[(size_t)which.idx[a] * out_stride + pos] = ke;
    vout[(size_t)which.idx[a] * out_stride + pos] = ie;
}

size_t rank_sort_scratch_bytes(int64_t n, int arrays)
{
    const int64_t chunks = (n + kChunk - 1) / kChunk;
    return align_up((size_t)arrays * (size_t)chunks * kChunk * (sizeof(unsigned long long) + sizeof(uint32_t)) + kAlign);
}

// keys_in[which.idx[a] * in_stride + i], a < which.count, i < n  ->  kout / vout[which.idx[a] * out_stride + rank]: keys
// ascending (ties by index) and the index every sorted position came from.  All arrays in two launches.
int rank_sort_pairs_u64(void *scratch, size_t scratch_bytes, const unsigned long long *keys_in, size_t in_stride, const RankSortArrays &which,
                        int n, unsigned long long *kout, uint32_t *vout, size_t out_stride, cudaStream_t stream)
{
    const int arrays = which.count;
    if (n <= 0 || arrays <= 0) return KP_OK;
    if (n > kRankSortMax || scratch_bytes < rank_sort_scratch_bytes(n, arrays)) {
        set_error("rank sort: %d keys x %d arrays does not fit (limit %d keys, scratch %zu bytes)", n, arrays, kRankSortMax, scratch_bytes);
        return KP_ERR_WORKSPACE;
    }
    const int chunks = (n + kChunk - 1) / kChunk;
    unsigned long long *ckey = static_cast<unsigned long long *>(scratch);
    uint32_t *cidx = reinterpret_cast<uint32_t *>(ckey + (size_t)arrays * chunks * kChunk);
    chunk_sort_kernel<<<dim3(chunks, arrays), kChunkThreads, 0, stream>>>(keys_in, in_stride, which, n, ckey, cidx);
    KP_LAUNCH_CHECK("chunk_sort_kernel");
    chunk_rank_kernel<<<dim3((chunks * kChunk + 255) / 256, arrays), 256, 0, stream>>>(ckey, cidx, chunks, n, which, kout, vout, out_stride);
    KP_LAUNCH_CHECK("chunk_rank_kernel");
    return KP_OK;
}

}  // namespace kp

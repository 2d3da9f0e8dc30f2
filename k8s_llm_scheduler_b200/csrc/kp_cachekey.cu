// Cache keys in bulk: RequestCache._generate_key (scheduler.py:265-271)
// is  md5(f"{pod_data}_{nodes_data}")  where nodes_data -- every node's name and usage, sorted by
// name -- is the same for all pods of a batch and grows with the cluster (~20 bytes per node), so the
// reference hashes the whole cluster description again for every pod.  MD5 chains block to block and
// the pod part comes FIRST, so the shared suffix cannot be hashed once; what can be done is to hash
// all pods of a batch in parallel: one thread per pod, the shared node text read through the cache
// by every thread at (almost) the same addresses.
//
// Message of pod j = prefix_j (host-formatted "cpu_mem_priority_", at most 255 bytes) followed by
// the node text.  Output: the 16 digest bytes per pod (hexdigest() order).  RFC 1321, little-endian.
#include "kp_common.cuh"

namespace kp {
namespace {

__constant__ uint32_t c_md5_k[64] = {
    0xd76aa478u, 0xe8c7b756u, 0x242070dbu, 0xc1bdceeeu, 0xf57c0fafu, 0x4787c62au, 0xa8304613u, 0xfd469501u,
    0x698098d8u, 0x8b44f7afu, 0xffff5bb1u, 0x895cd7beu, 0x6b901122u, 0xfd987193u, 0xa679438eu, 0x49b40821u,
    0xf61e2562u, 0xc040b340u, 0x265e5a51u, 0xe9b6c7aau, 0xd62f105du, 0x02441453u, 0xd8a1e681u, 0xe7d3fbc8u,
    0x21e1cde6u, 0xc33707d6u, 0xf4d50d87u, 0x455a14edu, 0xa9e3e905u, 0xfcefa3f8u, 0x676f02d9u, 0x8d2a4c8au,
    0xfffa3942u, 0x8771f681u, 0x6d9d6122u, 0xfde5380cu, 0xa4beea44u, 0x4bdecfa9u, 0xf6bb4b60u, 0xbebfbc70u,
    0x289b7ec6u, 0xeaa127fau, 0xd4ef3085u, 0x04881d05u, 0xd9d4d039u, 0xe6db99e5u, 0x1fa27cf8u, 0xc4ac5665u,
    0xf4292244u, 0x432aff97u, 0xab9423a7u, 0xfc93a039u, 0x655b59c3u, 0x8f0ccc92u, 0xffeff47du, 0x85845dd1u,
    0x6fa87e4fu, 0xfe2ce6e0u, 0xa3014314u, 0x4e0811a1u, 0xf7537e82u, 0xbd3af235u, 0x2ad7d2bbu, 0xeb86d391u};

__device__ __forceinline__ uint32_t rotl(uint32_t x, int s) { return __funnelshift_l(x, x, s); }

// one 64-byte block, message words in w[16]
__device__ __forceinline__ void md5_block(uint32_t (&h)[4], const uint32_t (&w)[16])
{
    constexpr int S[4][4] = {{7, 12, 17, 22}, {5, 9, 14, 20}, {4, 11, 16, 23}, {6, 10, 15, 21}};
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        uint32_t f;
        int g;
        if (i < 16) { f = (b & c) | (~b & d); g = i; }
        else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
        else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
        else { f = c ^ (b | ~d); g = (7 * i) & 15; }
        const uint32_t t = a + f + c_md5_k[i] + w[g];
        a = d;
        d = c;
        c = b;
        b = b + rotl(t, S[i >> 4][i & 3]);
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

__global__ void __launch_bounds__(128)
cache_keys_kernel(const uint8_t *__restrict__ prefix, const long long *__restrict__ prefix_off, long long n_pods,
                  const uint8_t *__restrict__ nodes_text, long long nodes_len, uint8_t *__restrict__ out, int text_aligned)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_pods) return;
    const long long p0 = prefix_off[j];
    const long long lp = prefix_off[j + 1] - p0;
    const long long total = lp + nodes_len;                       // message length in bytes
    const long long n_blocks = (total + 8) / 64 + 1;              // + 0x80 + 64-bit length
    auto byte_at = [&](long long i) -> uint32_t {                 // i-th byte of the padded message
        if (i < lp) return prefix[p0 + i];
        if (i < total) return nodes_text[i - lp];
        if (i == total) return 0x80u;
        const long long tail = n_blocks * 64 - 8;
        if (i >= tail) return (uint32_t)(((unsigned long long)total * 8ull) >> (8 * (i - tail))) & 0xffu;
        return 0u;
    };
    uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    const uint32_t *nodes32 = reinterpret_cast<const uint32_t *>(nodes_text);   // used only when 4-byte aligned
    for (long long b = 0; b < n_blocks; ++b) {
        uint32_t w[16];
        const long long first = b * 64, last = first + 64;
        if (text_aligned && first >= lp && last + 4 <= total) {           // (+4: the shifted form reads one word ahead)
            // the block lies inside the shared node text: aligned 32-bit loads + a funnel shift by this
            // pod's fixed misalignment
            const long long o = first - lp;                       // offset into the node text
            const int sh = (int)(o & 3) * 8;
            const long long q = o >> 2;
            uint32_t lo = nodes32[q];
#pragma unroll
            for (int t = 0; t < 16; ++t) {
                // the word after the last needed one may lie past the text: only touch it when the bytes are used
                const uint32_t hi = (sh != 0) ? nodes32[q + t + 1] : 0u;
                w[t] = sh ? __funnelshift_r(lo, hi, sh) : lo;
                lo = sh ? hi : ((t < 15) ? nodes32[q + t + 1] : 0u);
            }
        } else {
#pragma unroll
            for (int t = 0; t < 16; ++t) {
                const long long i = first + 4 * t;
                w[t] = byte_at(i) | (byte_at(i + 1) << 8) | (byte_at(i + 2) << 16) | (byte_at(i + 3) << 24);
            }
        }
        md5_block(h, w);
    }
    uint32_t *o32 = reinterpret_cast<uint32_t *>(out + 16 * j);
    o32[0] = h[0]; o32[1] = h[1]; o32[2] = h[2]; o32[3] = h[3];   // little-endian words = digest byte order
}

}  // namespace

int cache_keys_device(const uint8_t *prefix, const int64_t *prefix_off, int64_t n_pods, const uint8_t *nodes_text,
                      int64_t nodes_len, uint8_t *out, cudaStream_t stream)
{
    if (n_pods == 0) return KP_OK;
    const int blocks = (int)((n_pods + 127) / 128);
    const int text_aligned = (reinterpret_cast<uintptr_t>(nodes_text) & 3u) == 0;
    cache_keys_kernel<<<blocks, 128, 0, stream>>>(prefix, (const long long *)prefix_off, n_pods, nodes_text, nodes_len, out,
                                                  text_aligned);
    KP_LAUNCH_CHECK("cache_keys_kernel");
    return KP_OK;
}

}  // namespace kp

// What ONE per-pod exchange costs a node-sharded GREEDY placement: a persistent kernel per
// GPU; every round each rank stores one packed 64-bit (round tag | best key) word into every peer's buffer over
// NVLink (plain peer stores) and polls its own buffer until the words of all peers for that round have arrived,
// then reduces them (max) -- exactly the exchange a node-sharded greedy loop would need once per pod, with nothing
// else in the loop.  The round tag travels in the same 8-byte word as the payload (one store is one message; no
// fence), slots alternate by round parity (a rank can be at most one round ahead of a peer).  Waits are bounded.
//
// One process drives all GPUs (peer access enabled between them); one kernel per GPU, launched back to back, so
// kernels that wait on each other never share a device.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/p2p_exchange_bench tools/p2p_exchange_bench.cu
// Run:   tools/p2p_exchange_bench [rounds]        (uses 2, 4 and 8 GPUs as far as the box has them)
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int kMaxW = 8;
struct Peers { unsigned long long *slot[kMaxW]; };     // slot[g] = base of GPU g's buffer: [2 parities][kMaxW sources]

__device__ __forceinline__ void st_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__global__ void exchange_kernel(Peers peers, int me, int world, int rounds, unsigned long long *out, int *failed)
{
    const int t = threadIdx.x;                          // one thread per peer, all in one warp
    unsigned long long acc = 0;
    for (int r = 1; r <= rounds; ++r) {
        const int par = r & 1;
        const unsigned long long mine = ((unsigned long long)r << 32) | (unsigned)((me * 2654435761u) ^ (r * 40503u));
        unsigned long long got = mine;
        if (t < world && t != me) {
            st_sys(peers.slot[t] + par * kMaxW + me, mine);
            const unsigned long long *src = peers.slot[me] + par * kMaxW + t;
            long long spins = 0;
            while (((got = ld_sys(src)) >> 32) != (unsigned long long)r) {
                if (++spins > 200000000ll) { *failed = r; break; }
            }
        }
        // the arg-max over the ranks' words, as every rank of a greedy loop would compute it
        unsigned key = (unsigned)got;
        if (t >= world) key = 0;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) key = max(key, __shfl_xor_sync(0xffu, key, o, 8));
        acc += key;
        if (*(volatile int *)failed) break;
    }
    if (t == 0) *out = acc;
}

static double run(int world, int rounds)
{
    std::vector<unsigned long long *> buf(world), out(world);
    std::vector<int *> failed(world);
    std::vector<cudaStream_t> st(world);
    std::vector<cudaEvent_t> e0(world), e1(world);
    for (int g = 0; g < world; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaMalloc(&buf[g], sizeof(unsigned long long) * 2 * kMaxW));
        CK(cudaMemset(buf[g], 0, sizeof(unsigned long long) * 2 * kMaxW));
        CK(cudaMalloc(&out[g], 8));
        CK(cudaMalloc(&failed[g], 4));
        CK(cudaMemset(failed[g], 0, 4));
        CK(cudaStreamCreate(&st[g]));
        CK(cudaEventCreate(&e0[g]));
        CK(cudaEventCreate(&e1[g]));
    }
    Peers p{};
    for (int g = 0; g < world; ++g) p.slot[g] = buf[g];
    for (int g = 0; g < world; ++g) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
    const auto t0 = std::chrono::steady_clock::now();
    for (int g = 0; g < world; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaEventRecord(e0[g], st[g]));
        exchange_kernel<<<1, 8, 0, st[g]>>>(p, g, world, rounds, out[g], failed[g]);
        CK(cudaEventRecord(e1[g], st[g]));
    }
    float worst = 0.f;
    int bad = 0;
    for (int g = 0; g < world; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaStreamSynchronize(st[g]));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0[g], e1[g]));
        worst = ms > worst ? ms : worst;
        int f = 0;
        CK(cudaMemcpy(&f, failed[g], 4, cudaMemcpyDeviceToHost));
        bad |= f;
    }
    const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("world %d: %d rounds, slowest rank %.3f ms on the device (%.3f ms wall) = %.3f us per round%s\n", world, rounds, worst,
           wall * 1e3, worst * 1e3 / rounds, bad ? "  ** a wait timed out **" : "");
    for (int g = 0; g < world; ++g) {
        CK(cudaSetDevice(g));
        cudaFree(buf[g]); cudaFree(out[g]); cudaFree(failed[g]);
        cudaStreamDestroy(st[g]); cudaEventDestroy(e0[g]); cudaEventDestroy(e1[g]);
    }
    return worst * 1e3 / rounds;
}

int main(int argc, char **argv)
{
    const int rounds = argc > 1 ? atoi(argv[1]) : 100000;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    printf("GPUs visible: %d\n", n);
    for (int a = 0; a < n && a < kMaxW; ++a)
        for (int b = 0; b < n && b < kMaxW; ++b) {
            if (a == b) continue;
            int can = 0;
            CK(cudaDeviceCanAccessPeer(&can, a, b));
            if (!can) { printf("no peer access %d -> %d\n", a, b); return 1; }
            CK(cudaSetDevice(a));
            const cudaError_t pe = cudaDeviceEnablePeerAccess(b, 0);
            if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) { printf("peer access %d -> %d: %s\n", a, b, cudaGetErrorString(pe)); return 1; }
            cudaGetLastError();
        }
    run(1, rounds);                                      // the loop alone: no peer, no wait
    for (int w : {2, 4, 8})
        if (w <= n) { run(w, 1000); run(w, rounds); }
    printf("for comparison, one GPU places a pod in 0.22 us (config 3) to 0.82 us (config 5, greedy form) -- bench.py extras\n");
    return 0;
}

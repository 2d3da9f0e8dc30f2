// How fast can a kernel read / write PINNED HOST memory directly (no copy engine)?
// Compares plain coalesced loads/stores with bulk async copies (TMA engine) of various piece
// sizes, for the sizes the session API moves (1.6 MB = c2's pod table) and a large one.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/hostlink tools/hostlink_bench.cu && /tmp/hostlink
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// each CTA handles `piece` bytes; grid covers the buffer
__global__ void read_ldg64(const unsigned long long *src, size_t n, unsigned long long *sink)
{
    unsigned long long acc = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc ^= src[i];
    if (acc == 0x123456789abcdefull) *sink = acc;
}
__global__ void read_ldg128(const ulonglong2 *src, size_t n, unsigned long long *sink)
{
    unsigned long long acc = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        ulonglong2 v = src[i];
        acc ^= v.x ^ v.y;
    }
    if (acc == 0x123456789abcdefull) *sink = acc;
}
__global__ void read_bulk(const unsigned char *src, size_t bytes, uint32_t piece, unsigned long long *sink)
{
    extern __shared__ __align__(128) unsigned char buf[];
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;
    unsigned long long acc = 0;
    for (size_t off = (size_t)blockIdx.x * piece; off < bytes; off += (size_t)gridDim.x * piece) {
        const uint32_t b = (uint32_t)(bytes - off < piece ? bytes - off : piece);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(b) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(buf)), "l"(src + off), "r"(b), "r"(smem_u32(&bar)) : "memory");
        }
        asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(smem_u32(&bar)), "r"(phase) : "memory");
        phase ^= 1;
        for (uint32_t i = threadIdx.x * 8; i < b; i += blockDim.x * 8) acc ^= *reinterpret_cast<unsigned long long *>(buf + i);
        __syncthreads();
    }
    if (acc == 0x123456789abcdefull) *sink = acc;
}
__global__ void write_stg64(unsigned long long *dst, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = i;
}
__global__ void write_bulk(unsigned char *dst, size_t bytes, uint32_t piece)
{
    extern __shared__ __align__(128) unsigned char buf[];
    for (size_t off = (size_t)blockIdx.x * piece; off < bytes; off += (size_t)gridDim.x * piece) {
        const uint32_t b = (uint32_t)(bytes - off < piece ? bytes - off : piece);
        for (uint32_t i = threadIdx.x * 8; i < b; i += blockDim.x * 8) *reinterpret_cast<unsigned long long *>(buf + i) = off + i;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + off), "r"(smem_u32(buf)), "r"(b) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// both directions at once: read one pinned buffer, write another
__global__ void copy_h2h(const unsigned long long *src, unsigned long long *dst, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i] + 1;
}

template <class F>
static float time_us(F launch, int iters)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i) launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < iters; ++i) launch();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, a, b));
    return ms * 1000.f / iters;
}

int main()
{
    unsigned long long *sink;
    CK(cudaMalloc(&sink, 8));
    const size_t sizes[2] = {1600000, 64u << 20};
    for (size_t bytes : sizes) {
        unsigned char *h, *d;
        CK(cudaHostAlloc(&h, bytes, cudaHostAllocDefault));
        CK(cudaMalloc(&d, bytes));
        for (size_t i = 0; i < bytes; ++i) h[i] = (unsigned char)(i * 131u);
        const int iters = bytes > (8u << 20) ? 10 : 100;
        auto report = [&](const char *name, float us) { printf("%9zu B  %-34s %9.1f us  %7.2f GB/s\n", bytes, name, us, bytes / us * 1e-3); };
        report("copy engine H2D", time_us([&] { cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, 0); }, iters));
        report("copy engine D2H", time_us([&] { cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, 0); }, iters));
        const int grids[3] = {148, 592, 1776};
        for (int g : grids) {
            char nm[96];
            snprintf(nm, sizeof nm, "kernel read  LDG.64  grid %d x256", g);
            report(nm, time_us([&] { read_ldg64<<<g, 256>>>((const unsigned long long *)h, bytes / 8, sink); }, iters));
            snprintf(nm, sizeof nm, "kernel read  LDG.128 grid %d x256", g);
            report(nm, time_us([&] { read_ldg128<<<g, 256>>>((const ulonglong2 *)h, bytes / 16, sink); }, iters));
            snprintf(nm, sizeof nm, "kernel write STG.64  grid %d x256", g);
            report(nm, time_us([&] { write_stg64<<<g, 256>>>((unsigned long long *)h, bytes / 8); }, iters));
        }
        {
            unsigned char *h2, *d2;
            CK(cudaHostAlloc(&h2, bytes, cudaHostAllocDefault));
            CK(cudaMalloc(&d2, bytes));
            cudaStream_t sa, sb;
            CK(cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking));
            CK(cudaStreamCreateWithFlags(&sb, cudaStreamNonBlocking));
            for (int g : grids) {
                char nm[96];
                snprintf(nm, sizeof nm, "kernel read+write (both ways) grid %d", g);
                report(nm, time_us([&] { copy_h2h<<<g, 256>>>((const unsigned long long *)h, (unsigned long long *)h2, bytes / 8); }, iters));
            }
            // copy engines, both directions at once on two streams (events on the legacy stream fence both)
            cudaEvent_t ea, eb, e0;
            CK(cudaEventCreateWithFlags(&ea, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&eb, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&e0, cudaEventDisableTiming));
            report("copy engines H2D || D2H", time_us([&] {
                cudaEventRecord(e0, 0);
                cudaStreamWaitEvent(sa, e0, 0);
                cudaStreamWaitEvent(sb, e0, 0);
                cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, sa);
                cudaMemcpyAsync(h2, d2, bytes, cudaMemcpyDeviceToHost, sb);
                cudaEventRecord(ea, sa);
                cudaEventRecord(eb, sb);
                cudaStreamWaitEvent(0, ea, 0);
                cudaStreamWaitEvent(0, eb, 0);
            }, iters));
            CK(cudaDeviceSynchronize());
            CK(cudaFreeHost(h2));
            CK(cudaFree(d2));
        }
        const uint32_t pieces[4] = {1024, 2048, 4096, 16384};
        for (uint32_t pc : pieces) {
            char nm[96];
            const int g = (int)((bytes + pc - 1) / pc < 1776 ? (bytes + pc - 1) / pc : 1776);
            snprintf(nm, sizeof nm, "kernel read  bulk %5u B grid %d", pc, g);
            report(nm, time_us([&] { read_bulk<<<g, 128, pc>>>(h, bytes, pc, sink); }, iters));
            snprintf(nm, sizeof nm, "kernel write bulk %5u B grid %d", pc, g);
            report(nm, time_us([&] { write_bulk<<<g, 128, pc>>>(h, bytes, pc); }, iters));
        }
        CK(cudaDeviceSynchronize());
        CK(cudaFreeHost(h));
        CK(cudaFree(d));
    }
    printf("ok\n");
    return 0;
}

// Library sorts / scans of the plans: prefix scans, the 32-bit entry sort of the batched greedy prefix, and the
// key-array sorts of tables beyond kRankSortMax entries (up to that size kp_ranksort.cu sorts them with our own
// two kernels).
//
// The plan of a large table orders up to five 64-bit key arrays of n nodes (three strategy
// orders, two threshold orders).  That is a plain key-value radix sort, not part of the hot
// P x N loop, so it uses CUB (ships with the CUDA toolkit): cub::DeviceRadixSort (stable, so
// equal keys keep ascending node index = the reference's "first in list order wins") and
// cub::DeviceScan.  All temporary storage comes from the caller's workspace; nothing here
// allocates.  Kept in its own translation unit because the CUB templates dominate build time.
#include <algorithm>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "kp_common.cuh"

namespace kp {

size_t sort_temp_bytes(int64_t n)
{
    if (n <= 0) return kAlign;
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (const unsigned long long *)nullptr, (unsigned long long *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 64, (cudaStream_t)0);
    cub::DeviceScan::InclusiveSum(nullptr, b, (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, (cudaStream_t)0);
    // the queries depend on the device CUB sees; keep a floor so that a size computed without a
    // device (0 on failure) can never under-provision a later call
    return align_up(std::max({a, b, (size_t)n * 16 + ((size_t)1 << 20)}) + kAlign);
}

int sort_pairs_u64(void *temp, size_t temp_bytes, const unsigned long long *keys_in, unsigned long long *keys_out,
                   const uint32_t *vals_in, uint32_t *vals_out, int n, cudaStream_t stream)
{
    KP_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, 64, stream));
    count_launch(3);        // histogram + scan + onesweep passes (several kernels; counted coarsely)
    return KP_OK;
}

// descending 32-bit keys with 32-bit values (the batched greedy prefix: state entries by (score rank, node index))
int sort_pairs_u32_desc(void *temp, size_t temp_bytes, const uint32_t *keys_in, uint32_t *keys_out, const uint32_t *vals_in,
                        uint32_t *vals_out, int64_t n, cudaStream_t stream)
{
    if (n <= 0) return KP_OK;
    KP_CUDA(cub::DeviceRadixSort::SortPairsDescending(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, 32, stream));
    count_launch(3);
    return KP_OK;
}

int exclusive_sum_u32(void *temp, size_t temp_bytes, const uint32_t *in, uint32_t *out, int n, cudaStream_t stream)
{
    KP_CUDA(cub::DeviceScan::ExclusiveSum(temp, temp_bytes, in, out, n, stream));
    count_launch(2);
    return KP_OK;
}

int inclusive_sum_u32(void *temp, size_t temp_bytes, const uint32_t *in, uint32_t *out, int n, cudaStream_t stream)
{
    KP_CUDA(cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, n, stream));
    count_launch(2);
    return KP_OK;
}

}  // namespace kp

// 32-bit key/value radix sort used by the kNN Morton ordering (call site it replaces:
// cub::DeviceRadixSort::SortPairs in simple_knn.cu:213-214).  distCUDA2 runs once per training; this is
// still the toolkit's CUB onesweep (a header library compiled into this .so) -- the rasterizer's
// (tile | depth) ordering no longer uses it (binning.cu).
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace gsm {

size_t sort32_temp_bytes(size_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n);
  return align_up(bytes + 256, 128);
}

cudaError_t launch_sort_pairs32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in,
                                uint32_t* vals_out, size_t n, int end_bit, void* temp,
                                size_t temp_bytes, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  return cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out,
                                         (int64_t)n, 0, end_bit, stream);
}

}  // namespace gsm

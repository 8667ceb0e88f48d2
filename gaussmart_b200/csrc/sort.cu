// (tile|depth) key sort: stable LSD radix sort of u64 keys with u32 payloads over bits [0,end_bit).
// Call site it replaces: cub::DeviceRadixSort::SortPairs in rasterizer_impl.cu:304-309 and
// simple_knn.cu:213-214.
//
// v1: thin wrapper over CUB's onesweep (toolkit header library) so that the end-to-end path is
// parity-green first; the hand-written onesweep that replaces it is tracked in DESIGN.md.
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace gsm {

size_t sort_temp_bytes(size_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr,
                                  (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n);
  return align_up(bytes + 256, 128);
}

cudaError_t launch_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in,
                              uint32_t* vals_out, size_t n, int end_bit, void* temp,
                              size_t temp_bytes, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  return cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out,
                                         (int64_t)n, 0, end_bit, stream);
}

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

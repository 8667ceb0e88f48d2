// 32-bit key/value LSD radix sort for the kNN Morton ordering, hand-written for sm_100a.
// Call site it replaces: cub::DeviceRadixSort::SortPairs in simple_knn.cu:213-214 (the rasterizer's
// (tile | depth) ordering has its own per-tile sort in binning.cu).
//
// Stable, 8-bit digits, ceil(end_bit / 8) passes (4 for the 30-bit Morton codes).  The grid is
// persistent: at most SORT_MAX_BLOCKS CTAs, each owning one contiguous slice of the input, so the
// (digit, block) count table stays small enough for a single-CTA scan at any n.  Per pass:
//   1. count   : per-block digit histogram of its slice           -> counts[digit][block]
//   2. scan    : exclusive scan of the flat table (digit-major)   -> first output slot of every
//                (digit, block) pair; digit-major order makes the result globally stable
//   3. scatter : the block walks its slice tile by tile; each warp ranks its 32 keys with ballot
//                peer masks (no shared-memory atomics), a (digit, warp) scan inside the tile gives
//                every warp its slot, and the block's running per-digit base advances by the tile's
//                digit counts.  Equal digits keep (block, tile, warp, lane) = input order.
// HBM traffic per pass: 4 B (count) + 8 B read + 8 B written per element.
#include "common.cuh"

namespace gsm {

constexpr int SORT_THREADS = 512;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_TILE = 4096;          // elements per tile (8 per thread)
constexpr int SORT_MAX_BLOCKS = 148 * 4; // persistent grid bound
constexpr int SORT_RADIX = 256;

struct Sort32Plan {
  int blocks;
  size_t per_block;  // elements per block slice (multiple of SORT_TILE)
};

static Sort32Plan sort32_plan(size_t n) {
  Sort32Plan pl{};
  const size_t tiles = (n + SORT_TILE - 1) / SORT_TILE;
  const size_t tiles_per_block = (tiles + SORT_MAX_BLOCKS - 1) / SORT_MAX_BLOCKS;
  pl.per_block = (tiles_per_block ? tiles_per_block : 1) * SORT_TILE;
  pl.blocks = (int)((n + pl.per_block - 1) / pl.per_block);
  if (pl.blocks < 1) pl.blocks = 1;
  return pl;
}

size_t sort32_temp_bytes(size_t n) {
  const Sort32Plan pl = sort32_plan(n);
  // counts table + ping-pong buffers for keys and values (odd pass counts end in the temp buffers)
  return align_up((size_t)SORT_RADIX * pl.blocks * 4, 256) + 2 * align_up(n * 4, 256) + 256;
}

// lanes of the warp holding the same 8-bit digit (restricted to valid lanes), from 8 ballots
__device__ __forceinline__ unsigned peers8(uint32_t d, bool valid) {
  unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const bool bit = (d >> k) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, valid && bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

__global__ void __launch_bounds__(SORT_THREADS)
sort32_count_kernel(const uint32_t* __restrict__ keys, size_t n, size_t per_block, int shift, uint32_t dmask,
                    uint32_t* __restrict__ counts, int blocks) {
  __shared__ uint32_t whist[SORT_WARPS][SORT_RADIX];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int i = t; i < SORT_WARPS * SORT_RADIX; i += SORT_THREADS) (&whist[0][0])[i] = 0u;
  __syncthreads();
  const size_t beg = (size_t)blockIdx.x * per_block;
  const size_t end = beg + per_block < n ? beg + per_block : n;
  // warp-strided over the slice: the count does not depend on the order
  for (size_t it = beg + (size_t)warp * 32; it < end; it += SORT_THREADS) {
    const size_t i = it + lane;
    const bool valid = i < end;
    const uint32_t d = valid ? ((__ldg(keys + i) >> shift) & dmask) : 0u;
    const unsigned peers = peers8(d, valid);
    if (valid && lane == __ffs(peers) - 1) whist[warp][d] += __popc(peers);
    __syncwarp();
  }
  __syncthreads();
  for (int d = t; d < SORT_RADIX; d += SORT_THREADS) {
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) s += whist[w][d];
    counts[(size_t)d * blocks + blockIdx.x] = s;
  }
}

// exclusive scan of counts[0 .. m) in place, one CTA (m <= 256 * SORT_MAX_BLOCKS)
__global__ void __launch_bounds__(1024) sort32_scan_kernel(uint32_t* __restrict__ counts, int m) {
  __shared__ uint32_t warp_sums[32];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int per = (m + 1023) / 1024;
  const int beg = min(t * per, m), end = min(beg + per, m);
  uint32_t local = 0;
  for (int i = beg; i < end; i++) local += counts[i];
  uint32_t inc = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  uint32_t run = inc - local;
  for (int w = 0; w < warp; w++) run += warp_sums[w];
  for (int i = beg; i < end; i++) { const uint32_t c = counts[i]; counts[i] = run; run += c; }
}

__global__ void __launch_bounds__(SORT_THREADS)
sort32_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                      uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, size_t n,
                      size_t per_block, int shift, uint32_t dmask, const uint32_t* __restrict__ offsets,
                      int blocks) {
  __shared__ uint32_t whist[SORT_WARPS][SORT_RADIX];
  __shared__ uint32_t base[SORT_RADIX];   // next output slot of digit d for this block
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int d = t; d < SORT_RADIX; d += SORT_THREADS) base[d] = offsets[(size_t)d * blocks + blockIdx.x];
  const size_t beg = (size_t)blockIdx.x * per_block;
  const size_t end = beg + per_block < n ? beg + per_block : n;
  constexpr int PER_WARP = SORT_TILE / SORT_WARPS;  // contiguous elements per warp per tile
  for (size_t tile = beg; tile < end; tile += SORT_TILE) {
    for (int i = t; i < SORT_WARPS * SORT_RADIX; i += SORT_THREADS) (&whist[0][0])[i] = 0u;
    __syncthreads();
    const size_t wbeg = tile + (size_t)warp * PER_WARP;
    uint32_t k[PER_WARP / 32], v[PER_WARP / 32], rank[PER_WARP / 32];
    // sweep 1: load, per-warp digit counts, rank of every element among the warp's equal digits so far
#pragma unroll
    for (int r = 0; r < PER_WARP / 32; r++) {
      const size_t i = wbeg + (size_t)r * 32 + lane;
      const bool valid = i < end;
      k[r] = valid ? __ldg(keys_in + i) : 0xffffffffu;
      v[r] = valid ? __ldg(vals_in + i) : 0u;
      const uint32_t d = (k[r] >> shift) & dmask;
      const unsigned peers = peers8(d, valid);
      const int leader = __ffs(peers) - 1;
      uint32_t old = 0;
      if (valid && lane == leader) { old = whist[warp][d]; whist[warp][d] = old + __popc(peers); }
      old = __shfl_sync(0xffffffffu, old, valid ? leader : lane);
      rank[r] = old + __popc(peers & ((1u << lane) - 1u));
      __syncwarp();
    }
    __syncthreads();
    // (digit, warp) exclusive scan within the tile on top of the block's running base
    for (int d = t; d < SORT_RADIX; d += SORT_THREADS) {
      uint32_t run = base[d];
#pragma unroll
      for (int w = 0; w < SORT_WARPS; w++) { const uint32_t c = whist[w][d]; whist[w][d] = run; run += c; }
      base[d] = run;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < PER_WARP / 32; r++) {
      const size_t i = wbeg + (size_t)r * 32 + lane;
      if (i < end) {
        const uint32_t slot = whist[warp][(k[r] >> shift) & dmask] + rank[r];
        keys_out[slot] = k[r];
        vals_out[slot] = v[r];
      }
    }
    __syncthreads();
  }
}

cudaError_t launch_sort_pairs32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in,
                                uint32_t* vals_out, size_t n, int end_bit, void* temp,
                                size_t temp_bytes, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  if (n > 0xffffffffull || temp_bytes < sort32_temp_bytes(n)) return cudaErrorInvalidValue;
  const Sort32Plan pl = sort32_plan(n);
  char* tp = (char*)temp;
  uint32_t* counts = (uint32_t*)tp;
  tp += align_up((size_t)SORT_RADIX * pl.blocks * 4, 256);
  uint32_t* tkeys = (uint32_t*)tp;
  tp += align_up(n * 4, 256);
  uint32_t* tvals = (uint32_t*)tp;
  const int passes = end_bit <= 0 ? 0 : (end_bit + 7) / 8;
  if (passes == 0) {
    cudaMemcpyAsync(keys_out, keys_in, n * 4, cudaMemcpyDeviceToDevice, stream);
    return cudaMemcpyAsync(vals_out, vals_in, n * 4, cudaMemcpyDeviceToDevice, stream);
  }
  // ping-pong so that the last pass lands in (keys_out, vals_out)
  const uint32_t* src_k = keys_in;
  const uint32_t* src_v = vals_in;
  for (int p = 0; p < passes; p++) {
    const bool to_out = ((passes - 1 - p) & 1) == 0;
    uint32_t* dst_k = to_out ? keys_out : tkeys;
    uint32_t* dst_v = to_out ? vals_out : tvals;
    const int width = end_bit - 8 * p < 8 ? end_bit - 8 * p : 8;  // key bits >= end_bit do not take part
    const uint32_t dmask = (1u << width) - 1u;
    sort32_count_kernel<<<pl.blocks, SORT_THREADS, 0, stream>>>(src_k, n, pl.per_block, 8 * p, dmask, counts, pl.blocks);
    sort32_scan_kernel<<<1, 1024, 0, stream>>>(counts, SORT_RADIX * pl.blocks);
    sort32_scatter_kernel<<<pl.blocks, SORT_THREADS, 0, stream>>>(src_k, src_v, dst_k, dst_v, n, pl.per_block, 8 * p,
                                                                  dmask, counts, pl.blocks);
    src_k = dst_k;
    src_v = dst_v;
  }
  return cudaGetLastError();
}

}  // namespace gsm

extern "C" size_t gsm_sort_pairs32_temp_bytes(size_t n) { return gsm::sort32_temp_bytes(n); }

extern "C" int gsm_sort_pairs32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* values_in,
                                uint32_t* values_out, size_t n, int end_bit, void* temp, size_t temp_bytes,
                                void* stream) {
  if (end_bit < 0 || end_bit > 32) { gsm::set_error("gsm_sort_pairs32: end_bit %d out of range", end_bit); return GSM_ERR_INVALID_ARGUMENT; }
  if (n > 0 && (temp == nullptr || temp_bytes < gsm::sort32_temp_bytes(n))) {
    gsm::set_error("gsm_sort_pairs32: temp storage too small"); return GSM_ERR_WORKSPACE_TOO_SMALL;
  }
  GSM_CUDA(gsm::launch_sort_pairs32(keys_in, keys_out, values_in, values_out, n, end_bit, temp, temp_bytes,
                                    (cudaStream_t)stream));
  return GSM_OK;
}

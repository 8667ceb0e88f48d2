// Tile binning for sm_100a: inclusive scan of tile counts, duplicate-with-keys, tile ranges.
//
// Restates rasterizer_impl.cu:70-138 (duplicateWithKeys, identifyTileRanges) and the
// cub::DeviceScan::InclusiveSum call site rasterizer_impl.cu:278.  All integer work: results are
// bit-exact by construction (key = tile_id << 32 | bits(depth), emitted y-major / x-minor in
// surfel order, so a stable sort reproduces the reference's permutation).
#include "common.cuh"

namespace gsm {

// ---------------------------------------------------------------------------------------------
// Inclusive scan: single-pass decoupled look-back (one kernel, 2 x 4 bytes/elem of HBM traffic).
// Tile state word: bits 31..30 = status (0 invalid, 1 aggregate, 2 inclusive prefix), 29..0 value
// is not enough for R up to 2^32, so state is a 64-bit word: high 32 = status, low 32 = value.
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;  // 4 x uint4 per thread
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

size_t scan_temp_bytes(int P) {
  const size_t tiles = ((size_t)P + SCAN_TILE - 1) / SCAN_TILE;
  return align_up(sizeof(unsigned long long) * (tiles + 1) + 16, 128);
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int P,
            unsigned long long* __restrict__ state, unsigned int* __restrict__ ticket) {
  __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
  __shared__ uint32_t s_tile, s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);  // dynamic tile id => forward progress
  __syncthreads();
  const uint32_t tile = s_tile;
  const size_t base = (size_t)tile * SCAN_TILE;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  // blocked arrangement within a warp-striped layout: thread t owns items [t*16, t*16+16)
  uint32_t v[SCAN_ITEMS];
  const size_t tbase = base + (size_t)threadIdx.x * SCAN_ITEMS;
  if (tbase + SCAN_ITEMS <= (size_t)P) {
    const uint4* src = reinterpret_cast<const uint4*>(in + tbase);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; i++) {
      const uint4 q = __ldg(src + i);
      v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (tbase + i < (size_t)P) ? __ldg(in + tbase + i) : 0u;
  }
  uint32_t tsum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; i++) { tsum += v[i]; v[i] = tsum; }

  // warp inclusive scan of thread sums
  uint32_t inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  uint32_t warp_off = 0, block_sum = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; w++) {
    const uint32_t s = warp_sums[w];
    if (w < warp) warp_off += s;
    block_sum += s;
  }
  const uint32_t thread_excl = warp_off + inc - tsum;

  // publish aggregate, look back for the exclusive prefix (warp 0)
  if (warp == 0) {
    if (lane == 0) {
      const unsigned long long st = ((tile == 0 ? 2ull : 1ull) << 32) | block_sum;
      atomicExch(&state[tile], st);
    }
    uint32_t prefix = 0;
    if (tile > 0) {
      int look = (int)tile - 1;
      while (true) {
        const int j = look - lane;
        unsigned long long st = 0;
        if (j >= 0) {
          do { st = *reinterpret_cast<volatile unsigned long long*>(&state[j]); } while ((st >> 32) == 0);
        } else {
          st = 2ull << 32;  // virtual tile before the first: inclusive prefix 0
        }
        const uint32_t status = (uint32_t)(st >> 32), val = (uint32_t)st;
        const unsigned full = __ballot_sync(0xffffffffu, status == 2);
        // sum values of lanes up to and including the first lane with an inclusive prefix
        const int first = full ? (__ffs(full) - 1) : 32;
        uint32_t contrib = (lane <= first) ? val : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
        prefix += contrib;
        if (full) break;
        look -= 32;
      }
      if (lane == 0) {
        atomicExch(&state[tile], (2ull << 32) | (unsigned long long)(uint32_t)(prefix + block_sum));
      }
    }
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  const uint32_t off = s_prefix + thread_excl;
  if (tbase + SCAN_ITEMS <= (size_t)P) {
    uint4* dst = reinterpret_cast<uint4*>(out + tbase);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; i++)
      dst[i] = make_uint4(v[4 * i] + off, v[4 * i + 1] + off, v[4 * i + 2] + off, v[4 * i + 3] + off);
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
      if (tbase + i < (size_t)P) out[tbase + i] = v[i] + off;
  }
}

cudaError_t launch_inclusive_scan(const uint32_t* in, uint32_t* out, int P, void* temp,
                                  size_t temp_bytes, cudaStream_t stream) {
  if (P <= 0) return cudaSuccess;
  const size_t need = scan_temp_bytes(P);
  if (temp_bytes < need) return cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(temp, 0, need, stream);
  if (e != cudaSuccess) return e;
  const int tiles = (P + SCAN_TILE - 1) / SCAN_TILE;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(temp);
  unsigned long long* state = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(temp) + 16);
  scan_kernel<<<tiles, SCAN_THREADS, 0, stream>>>(in, out, P, state, ticket);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// duplicate-with-keys (rasterizer_impl.cu:70-111).  One warp per 32 surfels; each surfel's tile
// rect is re-derived from the record's AABB centre and its integer radius exactly as getRect
// (auxiliary.h:67-77) does, then the warp cooperatively emits the instances of its 32 surfels so
// the 12-byte pair stores of one surfel are issued by adjacent lanes.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
duplicate_kernel(int P, const float4* __restrict__ records, const uint32_t* __restrict__ offsets,
                 const int* __restrict__ radii, uint64_t* __restrict__ keys,
                 uint32_t* __restrict__ values, int grid_x, int grid_y) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  int r = 0;
  if (idx < P) r = __ldg(radii + idx);
  uint32_t off = 0, depth_bits = 0;
  int minx = 0, miny = 0, w = 0, n = 0;
  if (r > 0) {
    off = (idx == 0) ? 0u : __ldg(offsets + idx - 1);
    const float4* rec = records + (size_t)idx * REC_F4;
    const float cx = __ldg(&rec[0].w), cy = __ldg(&rec[1].w);
    depth_bits = __float_as_uint(__ldg(&rec[3].w));
    const float rf = (float)r;
    minx = min(grid_x, max(0, (int)__fmul_rn(__fsub_rn(cx, rf), 0.0625f)));
    miny = min(grid_y, max(0, (int)__fmul_rn(__fsub_rn(cy, rf), 0.0625f)));
    const int maxx = min(grid_x, max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(cx, rf), 16.0f), -1.0f), 0.0625f)));
    const int maxy = min(grid_y, max(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(cy, rf), 16.0f), -1.0f), 0.0625f)));
    w = maxx - minx;
    n = w * (maxy - miny);
  }
  // small rects (the common case, <= 4 tiles) are written by the owning lane; large ones are
  // spread over the warp.
  const int BIG = 16;
  if (n > 0 && n <= BIG) {
    for (int t = 0; t < n; t++) {
      const int y = miny + t / w, x = minx + t % w;
      keys[off + t] = ((uint64_t)(uint32_t)(y * grid_x + x) << 32) | depth_bits;
      values[off + t] = (uint32_t)idx;
    }
  }
  unsigned big = __ballot_sync(0xffffffffu, n > BIG);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const int bn = __shfl_sync(0xffffffffu, n, src);
    const int bw = __shfl_sync(0xffffffffu, w, src);
    const int bminx = __shfl_sync(0xffffffffu, minx, src);
    const int bminy = __shfl_sync(0xffffffffu, miny, src);
    const uint32_t boff = __shfl_sync(0xffffffffu, off, src);
    const uint32_t bdepth = __shfl_sync(0xffffffffu, depth_bits, src);
    const uint32_t bidx = (uint32_t)(idx - lane + src);
    for (int t = lane; t < bn; t += 32) {
      const int y = bminy + t / bw, x = bminx + t % bw;
      keys[boff + t] = ((uint64_t)(uint32_t)(y * grid_x + x) << 32) | bdepth;
      values[boff + t] = bidx;
    }
  }
}

void launch_duplicate_with_keys(int P, const float4* records, const uint32_t* offsets,
                                const int* radii, uint64_t* keys, uint32_t* values, int grid_x,
                                int grid_y, cudaStream_t stream) {
  if (P <= 0) return;
  duplicate_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, records, offsets, radii, keys, values,
                                                         grid_x, grid_y);
}

// ---------------------------------------------------------------------------------------------
// tile ranges (rasterizer_impl.cu:116-138): ranges[t] = [first, last+1) of tile t in the sorted
// list; untouched tiles stay (0,0) (the memset of rasterizer_impl.cu:311 is issued by the launcher
// on the same stream first).  Fused with it, because it streams the same sorted list: the per-
// instance STRIP MASK -- bit s set iff the surfel's conservative contribution box (common.cuh,
// CULL_RHO) meets rows [2s, 2s+1] of this tile -- which lets each compositing warp skip the
// entries that cannot reach any of its 32 pixels without looking at the surfel.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tile_ranges_kernel(size_t R, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ point_list,
                   const short4* __restrict__ bbox, uint2* __restrict__ ranges,
                   uint8_t* __restrict__ strip_mask, int grid_x, int no_cull) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= R) return;
  const uint32_t cur = (uint32_t)(__ldg(keys + idx) >> 32);
  if (idx == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(__ldg(keys + idx - 1) >> 32);
    if (cur != prev) {
      ranges[prev].y = (uint32_t)idx;
      ranges[cur].x = (uint32_t)idx;
    }
  }
  if (idx == R - 1) ranges[cur].y = (uint32_t)R;

  uint32_t mask = 0xffu;
  if (!no_cull) {
    const short4 bx = __ldg(bbox + __ldg(point_list + idx));
    const int tx = (int)(cur % (uint32_t)grid_x) * TILE_X, ty = (int)(cur / (uint32_t)grid_x) * TILE_Y;
    mask = 0;
    if (bx.x <= tx + TILE_X - 1 && bx.z >= tx) {
      const int r0 = max((int)bx.y - ty, 0), r1 = min((int)bx.w - ty, TILE_Y - 1);
      if (r0 <= r1) {
        const int s0 = r0 / STRIP_ROWS, s1 = r1 / STRIP_ROWS;
        mask = ((2u << s1) - 1u) & ~((1u << s0) - 1u);
      }
    }
  }
  strip_mask[idx] = (uint8_t)mask;
}

void launch_tile_ranges(size_t R, const uint64_t* keys_sorted, const uint32_t* point_list,
                        const short4* bbox, uint2* ranges, uint8_t* strip_mask, int grid_x,
                        int num_tiles, int no_cull, cudaStream_t stream) {
  cudaMemsetAsync(ranges, 0, sizeof(uint2) * (size_t)num_tiles, stream);
  if (R == 0) return;
  tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, stream>>>(R, keys_sorted, point_list, bbox, ranges,
                                                                       strip_mask, grid_x, no_cull);
}

}  // namespace gsm

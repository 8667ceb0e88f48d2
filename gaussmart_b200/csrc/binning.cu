// Tile binning for sm_100a: (tile | depth) ordering of the (surfel, tile) instances, tile ranges,
// strip masks.
//
// WHAT it produces is exactly what the reference produces with duplicateWithKeys +
// cub::DeviceRadixSort::SortPairs + identifyTileRanges (rasterizer_impl.cu:70-138,278-319):
//   keys_sorted[R]  = (tile_id << 32 | bits(view depth)) ascending,
//   point_list[R]   = surfel index of every instance, ties (equal tile AND equal depth bits) in
//                     ascending surfel index (what a stable sort of the reference's emission order gives),
//   ranges[t]       = [first, last+1) of tile t, (0,0) for untouched tiles.
// All integer work: bit-exact by construction.
//
// HOW is different.  The reference sorts R 96-bit pairs globally with a 6-pass LSD radix sort over
// 44 bits (152 B of HBM traffic per instance).  The key has structure the global sort ignores: its
// high part is a small dense integer (the tile) that the compositing kernels only ever need as a
// PARTITION.  So:
//   1. the forward preprocess counts instances per tile (one red.global.add per instance);
//   2. tile_scan_kernel turns the counts into ranges (exclusive scan over the Nt tiles, not over
//      the P surfels) -- this replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:278);
//   3. scatter_kernel reads every visible surfel's tile rect (computed by the forward preprocess
//      exactly as getRect does, auxiliary.h:67-77) and appends (depth bits, surfel) to its tiles' segments through a
//      per-tile cursor (order inside a segment is arbitrary at this point);
//   4. tile_sort_kernel sorts each segment by depth bits with a block-wide LSD radix sort held
//      entirely in shared memory (only as many 8-bit passes as the segment's depth bits actually
//      differ in, typically 4), fixes the order of equal-depth runs to ascending surfel index,
//      and writes point_list, keys_sorted and the strip mask of every instance.
// HBM traffic: 8 B written + 8 B read + 13 B written per instance (29 B vs 152 B).
#include "common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace gsm {

// ---------------------------------------------------------------------------------------------
// 2. exclusive scan of the per-tile instance counts -> ranges, total R; resets the cursors
// ---------------------------------------------------------------------------------------------
constexpr int TSCAN_THREADS = 1024;
constexpr int TSCAN_CLUSTER = 8;   // CTAs of one thread-block cluster share the tiles; totals cross through DSMEM

// tile_count / tile_cursor are [Nt][TILE_SUBS]: a tile's instances are counted and appended through
// TILE_SUBS sub-cursors selected by the surfel index, because same-address atomics WITH return
// serialise in L2 (~27 cycles each): one cursor per tile makes the hottest tile the critical path.
//
// One cluster of TSCAN_CLUSTER CTAs (round 1 and 2a: one CTA, 0.10 ms of a 3840x2160 frame's 32,400 tiles).  Every CTA
// owns a contiguous range of whole chunks: it sums its range, the CTAs exchange the sums through distributed shared
// memory (no global scratch, nothing to zero between frames), then each scans its range from its own offset.
__global__ void __cluster_dims__(TSCAN_CLUSTER, 1, 1) __launch_bounds__(TSCAN_THREADS)
tile_scan_kernel(const uint32_t* __restrict__ tile_count, uint32_t* __restrict__ tile_cursor,
                 uint2* __restrict__ ranges, int num_tiles, unsigned long long capacity,
                 const uint32_t* __restrict__ flags, ForwardStatus* __restrict__ status) {
  // a CTA walks its tiles in chunks of TSCAN_THREADS, one tile (TILE_SUBS counters = 32 B) per thread and chunk, so
  // every access is coalesced; the next chunk's counters are requested before the current chunk is scanned
  __shared__ unsigned long long warp_sums[2][TSCAN_THREADS / 32];
  __shared__ unsigned long long cta_total;      // read by the other CTAs of the cluster
  __shared__ unsigned long long cta_offset[2];  // sum over the lower ranks, sum over all ranks
  static_assert(TILE_SUBS == 8, "a tile's counters are read as two uint4");
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const uint4* cnt4 = reinterpret_cast<const uint4*>(tile_count);
  uint4* cur4 = reinterpret_cast<uint4*>(tile_cursor);
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
  const int chunks = (num_tiles + TSCAN_THREADS - 1) / TSCAN_THREADS;
  const int per = ((chunks + TSCAN_CLUSTER - 1) / TSCAN_CLUSTER) * TSCAN_THREADS;
  const int begin = min(rank * per, num_tiles), end = min(begin + per, num_tiles);

  // ---- pass 1: the range's total
  unsigned long long mine = 0;
  for (int i = begin + t; i < end; i += TSCAN_THREADS) {
    const uint4 c0 = __ldg(cnt4 + 2 * (size_t)i), c1 = __ldg(cnt4 + 2 * (size_t)i + 1);
    mine += (unsigned long long)c0.x + c0.y + c0.z + c0.w + c1.x + c1.y + c1.z + c1.w;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if (lane == 0) warp_sums[0][warp] = mine;
  __syncthreads();
  if (warp == 0) {
    unsigned long long v = warp_sums[0][lane];
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) cta_total = v;
  }
  cluster.sync();
  if (warp == 0) {
    unsigned long long v = 0;
    if (lane < TSCAN_CLUSTER) v = *cluster.map_shared_rank(&cta_total, lane);
    unsigned long long below = lane < rank ? v : 0ull, all = v;
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      below += __shfl_xor_sync(0xffffffffu, below, o);
      all += __shfl_xor_sync(0xffffffffu, all, o);
    }
    if (lane == 0) { cta_offset[0] = below; cta_offset[1] = all; }
  }
  cluster.sync();  // (also: no CTA leaves while another still reads its cta_total)

  // ---- pass 2: scan of the range (the counters come from L2 this time)
  unsigned long long carry = cta_offset[0];
  uint4 n0 = zero, n1 = zero;
  if (begin + t < end) { n0 = __ldg(cnt4 + 2 * (size_t)(begin + t)); n1 = __ldg(cnt4 + 2 * (size_t)(begin + t) + 1); }
  int buf = 0;
  for (int base = begin; base < end; base += TSCAN_THREADS, buf ^= 1) {
    const int i = base + t;
    const uint4 c0 = n0, c1 = n1;
    const int nxt = i + TSCAN_THREADS;
    n0 = zero; n1 = zero;
    if (nxt < end) { n0 = __ldg(cnt4 + 2 * (size_t)nxt); n1 = __ldg(cnt4 + 2 * (size_t)nxt + 1); }
    const unsigned long long local = (unsigned long long)c0.x + c0.y + c0.z + c0.w + c1.x + c1.y + c1.z + c1.w;
    unsigned long long inc = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane == 31) warp_sums[buf][warp] = inc;
    __syncthreads();  // (double-buffered sums: one barrier per chunk)
    unsigned long long warp_off = 0, total = 0;
    for (int w = 0; w < TSCAN_THREADS / 32; w++) {
      const unsigned long long s = warp_sums[buf][w];
      if (w < warp) warp_off += s;
      total += s;
    }
    if (i < end) {
      unsigned long long run = carry + warp_off + inc - local;
      const unsigned long long first = run;
      uint4 o;  // absolute slot of each sub-segment's next pair
      o.x = (uint32_t)run; run += c0.x; o.y = (uint32_t)run; run += c0.y;
      o.z = (uint32_t)run; run += c0.z; o.w = (uint32_t)run; run += c0.w;
      cur4[2 * (size_t)i] = o;
      o.x = (uint32_t)run; run += c1.x; o.y = (uint32_t)run; run += c1.y;
      o.z = (uint32_t)run; run += c1.z; o.w = (uint32_t)run; run += c1.w;
      cur4[2 * (size_t)i + 1] = o;
      ranges[i] = run > first ? make_uint2((uint32_t)first, (uint32_t)run) : make_uint2(0u, 0u);
    }
    carry += total;
  }
  if (rank == 0 && t == 0) {
    // R, and whether it fits the binning workspace the host sized from its high-water mark: if not, every kernel
    // behind this one treats the frame as empty (see ForwardStatus) and the host renders it again with a larger one
    const unsigned long long R = cta_offset[1];
    ForwardStatus st;
    st.num_rendered = R;
    st.overflow = R > capacity ? 1u : 0u;
    st.prefiltered_violation = flags[0];
    *status = st;
  }
}

void launch_tile_scan(const uint32_t* tile_count, uint32_t* tile_cursor, uint2* ranges, int num_tiles,
                      unsigned long long capacity, const uint32_t* flags, ForwardStatus* status, cudaStream_t stream) {
  tile_scan_kernel<<<TSCAN_CLUSTER, TSCAN_THREADS, 0, stream>>>(tile_count, tile_cursor, ranges, num_tiles, capacity, flags, status);
}

// ---------------------------------------------------------------------------------------------
// Block mask of one (surfel, tile) instance: which of the tile's 8 blocks (8x4 pixels) the surfel's conservative
// contribution region -- box AND diagonal slabs = an octagon, common.cuh -- can reach.  Computed by the scatter, where
// the surfel's box / slabs are loaded ONCE per surfel as a coalesced stream, and carried through the sort in the top
// byte of the pair's value (surfel index < 2^24); round 1 gathered box and slabs per INSTANCE at the end of the sort
// (24 B in 32-B sectors from random surfels: 2.5x the algorithmic DRAM traffic of that kernel).
// ---------------------------------------------------------------------------------------------
constexpr uint32_t ID_BITS = 24, ID_MASK = (1u << ID_BITS) - 1u;

__device__ __forceinline__ uint32_t instance_block_mask(const short4 bx, const float4 dg, int tx, int ty) {
  uint32_t mask = 0;
  const int c0 = max((int)bx.x - tx, 0), c1 = min((int)bx.z - tx, TILE_X - 1);
  const int r0 = max((int)bx.y - ty, 0), r1 = min((int)bx.w - ty, TILE_Y - 1);
  if (c0 <= c1 && r0 <= r1) {
    mask = block_mask(c0, c1, r0, r1);
    // keep only the blocks that meet the surfel's diagonal slabs (cull record: bounds of x+y and x-y over
    // its contribution region, slack included): with the box this bounds the region by an octagon
#if GSM_BLOCK_W == 8
    // 2 x 4 blocks of 8 x 4 pixels: block (bx, by) sits on the diagonals u = 2bx + by (x+y = s + 4u) and
    // v = 2bx - by (x-y = d + 4v), so each slab bound is a range of u or v and the four tests
    //   x1+y1 = s+4u+10 >= dg.x,  x0+y0 = s+4u <= dg.y,  x1-y0 = d+4v+7 >= dg.z,  x0-y1 = d+4v-3 <= dg.w
    // become four table lookups (one byte per threshold: the blocks on the right side of it) -- no loop over blocks
    const float s = (float)(tx + ty), d = (float)(tx - ty);
    const int ulo = (int)fminf(fmaxf(ceilf((dg.x - 10.0f - s) * 0.25f), 0.0f), 6.0f);
    const int uhi = (int)fminf(fmaxf(floorf((dg.y - s) * 0.25f), -1.0f), 5.0f);
    const int vlo = (int)fminf(fmaxf(ceilf((dg.z - 7.0f - d) * 0.25f), -3.0f), 3.0f);
    const int vhi = (int)fminf(fmaxf(floorf((dg.w + 3.0f - d) * 0.25f), -4.0f), 2.0f);
    mask &= (uint32_t)((0x80a0e8fafeffull >> (8 * ulo)) & (0xff7f5f17050100ull >> (8 * (uhi + 1))) &
                       (0x20a2bafbfffull >> (8 * (vlo + 3))) & (0xfffdf5d4504000ull >> (8 * (vhi + 4)))) & 0xffu;
#else
    uint32_t keep = 0;
    for (uint32_t rest = mask; rest; rest &= rest - 1u) {
      const int b = __ffs(rest) - 1;
      const float x0 = (float)(tx + (b % BLOCKS_X) * BLOCK_W), y0 = (float)(ty + (b / BLOCKS_X) * BLOCK_H);
      const float x1 = x0 + (float)(BLOCK_W - 1), y1 = y0 + (float)(BLOCK_H - 1);
      if (x1 + y1 >= dg.x && x0 + y0 <= dg.y && x1 - y0 >= dg.z && x0 - y1 <= dg.w) keep |= 1u << b;
    }
    mask = keep;
#endif
  }
  return mask;
}

// ---------------------------------------------------------------------------------------------
// 3. scatter (the reference's duplicateWithKeys, rasterizer_impl.cu:70-111).  One thread LOADS one surfel, but the
//    (surfel, tile) instances of a warp's 32 surfels are flattened over its lanes: an inclusive scan of the tile
//    counts, then lane l takes instances l, l + 32, ... of the warp's list, finds the owning surfel by a 5-step search
//    over the scanned counts (shuffles) and reads that surfel's rect / box / slabs from shared memory.  (One thread per
//    surfel with a loop over its tiles ran at 9 of 32 lanes: most surfels touch 1-4 tiles, a few up to 16.)
//    Rects of more than 16 tiles are spread over the warp one surfel at a time, as before.
// ---------------------------------------------------------------------------------------------
// PACKED: P <= 2^24, the pair's value is (block mask << 24 | surfel); otherwise the plain surfel index (the sort's
// epilogue then gathers box and slabs itself, as in round 1).
#ifndef GSM_SCATTER_UNROLL
#define GSM_SCATTER_UNROLL 2
#endif
constexpr int SCATTER_THREADS = 256;
constexpr int SCATTER_U = GSM_SCATTER_UNROLL;   // instances per lane and round (returning atomics in flight)
template <bool PACKED>
__global__ void __launch_bounds__(SCATTER_THREADS)
scatter_kernel(int P, const uint4* __restrict__ rects, const int* __restrict__ radii,
               uint32_t* __restrict__ tile_cursor, uint2* __restrict__ pairs, int grid_x,
               const short4* __restrict__ bbox, const float4* __restrict__ cull, int no_cull,
               const ForwardStatus* __restrict__ status) {
  if (status->overflow) return;  // R exceeds the pair buffer: nothing may be appended
  __shared__ uint4 s_rect[SCATTER_THREADS];   // minx | miny << 16, width, ceil(2^16 / width), depth bits
  __shared__ short4 s_box[PACKED ? SCATTER_THREADS : 1];
  __shared__ float4 s_slab[PACKED ? SCATTER_THREADS : 1];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int wbase = threadIdx.x - lane;
  int r = 0;
  if (idx < P) r = __ldg(radii + idx);
  uint32_t depth_bits = 0;
  int minx = 0, miny = 0, w = 0, n = 0;
  if (r > 0) {
    // tile rect (getRect, auxiliary.h:67-77) and depth bits as left by the forward preprocess
    const uint4 q = __ldg(rects + idx);
    minx = (int)(q.x & 0xffffu); miny = (int)(q.x >> 16);
    w = (int)(q.y & 0xffffu) - minx;
    n = w * ((int)(q.y >> 16) - miny);
    depth_bits = q.z;
  }
  auto value_of = [&](uint32_t surfel, const short4 b, const float4 g, uint32_t tile_x, uint32_t tile_y) -> uint32_t {
    if (!PACKED) return surfel;
    const uint32_t m = no_cull ? 0xffu : instance_block_mask(b, g, (int)tile_x * TILE_X, (int)tile_y * TILE_Y);
    return surfel | (m << ID_BITS);
  };
  const int BIG = 16;
  const int n_small = (n > 0 && n <= BIG) ? n : 0;
  int incl = n_small;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  if (n_small) {
    // t / w for t < 16, w <= 16 as (t * ceil(2^16 / w)) >> 16: exact (error term t * (m w - 2^16) < 2^16)
    s_rect[threadIdx.x] = make_uint4((uint32_t)minx | ((uint32_t)miny << 16), (uint32_t)w,
                                     (65535u / (uint32_t)w) + 1u, depth_bits);
    if (PACKED && !no_cull) { s_box[threadIdx.x] = __ldg(bbox + idx); s_slab[threadIdx.x] = __ldg(cull + idx); }
  }
  __syncwarp();
  // two instances per lane and round: both cursor atomics (they return a value) are issued before either result is used
  for (int j0 = 0; j0 < total; j0 += 32 * SCATTER_U) {
    uint32_t slot[SCATTER_U], val[SCATTER_U], dep[SCATTER_U];
#pragma unroll
    for (int u = 0; u < SCATTER_U; u++) {
      const int j = j0 + 32 * u + lane;
      int s = 0;  // owner: the first lane whose inclusive count exceeds j
#pragma unroll
      for (int step = 16; step; step >>= 1) {
        const int v = __shfl_sync(0xffffffffu, incl, s + step - 1);
        if (v <= j) s += step;
      }
      s = min(s, 31);
      const int first = __shfl_sync(0xffffffffu, incl - n_small, s);
      slot[u] = 0xffffffffu;
      if (j < total) {
        const uint4 q = s_rect[wbase + s];
        const uint32_t t = (uint32_t)(j - first);
        const uint32_t row = (t * q.z) >> 16;
        const uint32_t tx = (q.x & 0xffffu) + (t - row * q.y), ty = (q.x >> 16) + row;
        const uint32_t surfel = (uint32_t)(idx - lane + s);
        const uint32_t tile = ty * (uint32_t)grid_x + tx;
        slot[u] = atomicAdd(tile_cursor + tile * TILE_SUBS + (surfel & (TILE_SUBS - 1)), 1u);
        short4 b = make_short4(0, 0, 0, 0);
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (PACKED && !no_cull) { b = s_box[wbase + s]; g = s_slab[wbase + s]; }
        val[u] = value_of(surfel, b, g, tx, ty);
        dep[u] = q.w;
      }
    }
#pragma unroll
    for (int u = 0; u < SCATTER_U; u++)
      if (slot[u] != 0xffffffffu) pairs[slot[u]] = make_uint2(dep[u], val[u]);
  }
  unsigned big = __ballot_sync(0xffffffffu, n > BIG);
  while (big) {
    const int src = __ffs(big) - 1;
    big &= big - 1;
    const int bn = __shfl_sync(0xffffffffu, n, src);
    const int bw = __shfl_sync(0xffffffffu, w, src);
    const int bminx = __shfl_sync(0xffffffffu, minx, src);
    const int bminy = __shfl_sync(0xffffffffu, miny, src);
    const uint32_t bdepth = __shfl_sync(0xffffffffu, depth_bits, src);
    const uint32_t bidx = (uint32_t)(idx - lane + src);
    short4 bbx = make_short4(0, 0, 0, 0);
    float4 bdg = make_float4(0.f, 0.f, 0.f, 0.f);
    if (PACKED && !no_cull) { bbx = __ldg(bbox + bidx); bdg = __ldg(cull + bidx); }   // (same address in every lane: one request)
    for (int t = lane; t < bn; t += 32) {
      const uint32_t ty = (uint32_t)(bminy + t / bw), tx = (uint32_t)(bminx + t % bw);
      const uint32_t tile = ty * (uint32_t)grid_x + tx;
      const uint32_t slot = atomicAdd(tile_cursor + tile * TILE_SUBS + (bidx & (TILE_SUBS - 1)), 1u);
      pairs[slot] = make_uint2(bdepth, value_of(bidx, bbx, bdg, tx, ty));
    }
  }
}

void launch_scatter(int P, const uint4* rects, const int* radii, uint32_t* tile_cursor, uint2* pairs,
                    int grid_x, const short4* bbox, const float4* cull, int no_cull, int packed,
                    const ForwardStatus* status, cudaStream_t stream) {
  if (P <= 0) return;
  if (packed)
    scatter_kernel<true><<<(P + 255) / 256, 256, 0, stream>>>(P, rects, radii, tile_cursor, pairs, grid_x, bbox, cull, no_cull, status);
  else
    scatter_kernel<false><<<(P + 255) / 256, 256, 0, stream>>>(P, rects, radii, tile_cursor, pairs, grid_x, bbox, cull, no_cull, status);
}

// ---------------------------------------------------------------------------------------------
// 4. per-tile sort by depth bits: block-wide LSD radix sort (8-bit digits) of (depth, surfel)
//    pairs.  Segments of up to TS_CAP_SMALL / TS_CAP_LARGE pairs are sorted in shared memory
//    (two launches with different dynamic shared memory sizes so small tiles keep 3 CTAs per SM);
//    longer ones ping-pong between their own slices of the pair buffer and of keys_sorted in
//    global memory (same code, generic pointers).  Each launch covers all tiles; a CTA whose
//    segment length is outside the launch's class returns at once.
// ---------------------------------------------------------------------------------------------
#ifndef GSM_TS_CAP_SMALL
#define GSM_TS_CAP_SMALL 4096
#endif
#ifndef GSM_TS_CAP_LARGE
#define GSM_TS_CAP_LARGE 9216
#endif
constexpr uint32_t TS_CAP_SMALL = GSM_TS_CAP_SMALL;    // pairs per shared-memory buffer of the small class (2 x 32 KB + 32 KB histograms: 2 CTAs/SM)
constexpr uint32_t TS_CAP_LARGE = GSM_TS_CAP_LARGE;    // large class (2 x 72 KB), 1024 threads
constexpr int TS_SMALL_THREADS = 512, TS_SMALL_BITS = 9;   // <= 9-bit digits: 26 differing bits -> 3 passes
constexpr int TS_LARGE_THREADS = 1024, TS_LARGE_BITS = 8;   // (9-bit digits measured slower here: 355 vs 275 us)

struct TileSortParams {
  const uint2* ranges;
  uint2* pairs;            // in: unsorted (depth bits, surfel) per tile segment; scratch afterwards
  uint64_t* keys_sorted;   // out (and scratch of the global-memory class)
  uint32_t* point_list;    // out
  uint8_t* strip_mask;     // out
  const short4* bbox;
  const float4* cull;
  int grid_x;
  int no_cull;
  uint32_t n_lo, n_hi;     // this launch handles segments with n_lo < n <= n_hi
  uint32_t cap;            // pairs per shared-memory buffer (0: global-memory class)
  int packed;              // the pairs' values carry the block mask in their top byte (scatter_kernel<true>)
  const ForwardStatus* status;
};

// One stable LSD pass over `bits` bits at `shift`: A -> B.  Warp w owns the contiguous range
// [w*chunk, (w+1)*chunk) of A; whist[w][d] first counts, then (after the scan over (digit, warp))
// holds the next output slot of digit d for warp w.
// lanes holding the same digit as this lane (match.any semantics, restricted to `valid` lanes), built
// from `bits` ballots: fixed latency, whereas MATCH.ANY iterates over the distinct values in the warp
// (typically 32 distinct digits here) and dominated the pass.
__device__ __forceinline__ unsigned digit_peers(uint32_t d, bool valid, int bits) {
  unsigned peers = __ballot_sync(0xffffffffu, valid);
  for (int k = 0; k < bits; k++) {
    const bool bit = (d >> k) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, valid && bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

// ITEMS > 0 and `cached`: the segment fits ITEMS elements per thread (shared-memory classes), so the elements and
// their peer masks found in sweep 1 stay in registers for sweep 2 (no second read, no second round of ballots).
// BY_ID: digits are taken from the surfel index (e.y) instead of the depth bits (e.x): the tie-break passes.
template <int THREADS, int MAXBITS, int ITEMS, bool BY_ID>
__device__ __forceinline__ void radix_pass(const uint2* __restrict__ A, uint2* __restrict__ B, uint32_t n,
                                           uint32_t kmin, int shift, int bits, uint32_t (*whist)[1 << MAXBITS],
                                           uint32_t* warp_tot, int t, int lane, int warp, bool cached) {
  constexpr int WARPS = THREADS / 32, MAXRADIX = 1 << MAXBITS;
  const uint32_t radix = 1u << bits, dmask = radix - 1u;
  const uint32_t chunk = ((n + THREADS - 1) / THREADS) * 32;
  const uint32_t wbeg = min(warp * chunk, n), wend = min(wbeg + chunk, n);
  for (int i = t; i < WARPS * MAXRADIX; i += THREADS) (&whist[0][0])[i] = 0u;
  __syncthreads();
  uint2 ce[ITEMS];
  unsigned cp[ITEMS];
  // sweep 1: per-warp digit histograms; lanes with equal digits are aggregated so that one lane per
  // distinct digit touches shared memory (no shared-memory atomics: 2 cycles per lane)
  if (cached) {
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
      const uint32_t it = wbeg + 32u * k;
      cp[k] = 0u;
      if (it < wend) {  // warp-uniform
        const uint32_t i = it + lane;
        const bool valid = i < wend;
        uint2 e = make_uint2(0u, 0u);
        if (valid) e = A[i];
        const uint32_t d = valid ? ((((BY_ID ? (e.y & ID_MASK) : e.x) - kmin) >> shift) & dmask) : 0u;
        const unsigned peers = digit_peers(d, valid, bits);
        if (valid && lane == __ffs(peers) - 1) whist[warp][d] += __popc(peers);
        ce[k] = e;
        cp[k] = valid ? peers : 0u;
        __syncwarp();
      }
    }
  } else {
    for (uint32_t it = wbeg; it < wend; it += 32) {
      const uint32_t i = it + lane;
      const bool valid = i < wend;
      const uint32_t d = valid ? ((((BY_ID ? (A[i].y & ID_MASK) : A[i].x) - kmin) >> shift) & dmask) : 0u;
      const unsigned peers = digit_peers(d, valid, bits);
      if (valid && lane == __ffs(peers) - 1) whist[warp][d] += __popc(peers);
      __syncwarp();
    }
  }
  __syncthreads();
  // exclusive scan over (digit, warp): each thread owns DPT consecutive digits
  {
    constexpr int DPT = (MAXRADIX + THREADS - 1) / THREADS;
    uint32_t tot[DPT];
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < DPT; k++) {
      const uint32_t d = t * DPT + k;
      uint32_t run = 0;
      if (d < radix) {
        for (int w = 0; w < WARPS; w++) { const uint32_t c = whist[w][d]; whist[w][d] = run; run += c; }
      }
      tot[k] = run;
      mine += run;
    }
    uint32_t inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t base = inc - mine;
    for (int w = 0; w < warp; w++) base += warp_tot[w];
#pragma unroll
    for (int k = 0; k < DPT; k++) {
      const uint32_t d = t * DPT + k;
      if (d < radix) {
        for (int w = 0; w < WARPS; w++) whist[w][d] += base;
      }
      base += tot[k];
    }
  }
  __syncthreads();
  // sweep 2: stable scatter
  if (cached) {
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
      if (wbeg + 32u * k < wend) {  // warp-uniform
        const unsigned peers = cp[k];
        const bool valid = peers != 0u;
        const uint2 e = ce[k];
        const uint32_t d = (((BY_ID ? (e.y & ID_MASK) : e.x) - kmin) >> shift) & dmask;
        const int leader = valid ? __ffs(peers) - 1 : lane;
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        uint32_t old = 0;
        if (valid && lane == leader) { old = whist[warp][d]; whist[warp][d] = old + __popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader);
        if (valid) B[old + rank] = e;
        __syncwarp();
      }
    }
    __syncthreads();
    return;
  }
  for (uint32_t it = wbeg; it < wend; it += 32) {
    const uint32_t i = it + lane;
    const bool valid = i < wend;
    uint2 e = make_uint2(0u, 0u);
    if (valid) e = A[i];
    const uint32_t d = valid ? ((((BY_ID ? (e.y & ID_MASK) : e.x) - kmin) >> shift) & dmask) : 0u;
    const unsigned peers = digit_peers(d, valid, bits);
    const int leader = valid ? __ffs(peers) - 1 : lane;
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    uint32_t old = 0;
    if (valid && lane == leader) { old = whist[warp][d]; whist[warp][d] = old + __popc(peers); }
    old = __shfl_sync(0xffffffffu, old, leader);
    if (valid) B[old + rank] = e;
    __syncwarp();
  }
  __syncthreads();
}

template <int THREADS, int MAXBITS, int ITEMS>
__global__ void __launch_bounds__(THREADS) tile_sort_kernel(const TileSortParams p) {
  constexpr int WARPS = THREADS / 32;
  extern __shared__ __align__(16) uint2 ts_dyn[];
  // dynamic shared memory: [pair buffer A | pair buffer B] (cap pairs each; none in the global-memory class),
  // then the per-warp digit histograms
  uint32_t (*whist)[1 << MAXBITS] = reinterpret_cast<uint32_t (*)[1 << MAXBITS]>(ts_dyn + 2 * (size_t)p.cap);
  __shared__ uint32_t warp_tot[WARPS];
  const uint32_t tile = blockIdx.x;
  const uint2 range = __ldg(p.ranges + tile);
  const uint32_t n = range.y - range.x;
  if (n <= p.n_lo || n > p.n_hi || p.status->overflow) return;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const bool in_smem = p.cap != 0;
  uint2* A = in_smem ? ts_dyn : p.pairs + range.x;
  uint2* B = in_smem ? ts_dyn + p.cap : reinterpret_cast<uint2*>(p.keys_sorted + range.x);
  if (in_smem) {
    const uint2* src = p.pairs + range.x;
    for (uint32_t i = t; i < n; i += THREADS) A[i] = __ldg(src + i);
  }
  __syncthreads();
  // span of the segment's depth keys: digits are taken from key - min (order-preserving), so the passes cover
  // log2(max - min) bits -- 26 for view depths between 0.2 and 8 -- instead of every bit position in which two
  // keys differ (31: the exponent's top bit flips at depth 2.0)
  uint32_t kmin = 0xffffffffu, kmax = 0u;
  for (uint32_t i = t; i < n; i += THREADS) { const uint32_t k = A[i].x; kmin = min(kmin, k); kmax = max(kmax, k); }
  kmin = __reduce_min_sync(0xffffffffu, kmin);
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  __shared__ uint32_t warp_min[WARPS];
  if (lane == 0) { warp_tot[warp] = kmax; warp_min[warp] = kmin; }
  __syncthreads();
  for (int w = 0; w < WARPS; w++) { kmax = max(kmax, warp_tot[w]); kmin = min(kmin, warp_min[w]); }
  __syncthreads();
  const uint32_t span = kmax - kmin;
  const int nbits = span ? 32 - __clz(span) : 0;
  const int passes = (nbits + MAXBITS - 1) / MAXBITS;
  const int bits = passes ? (nbits + passes - 1) / passes : 0;  // equal digit widths, <= MAXBITS
  for (int shift = 0; shift < nbits; shift += bits) {
    radix_pass<THREADS, MAXBITS, ITEMS, false>(A, B, n, kmin, shift, bits, whist, warp_tot, t, lane, warp, in_smem);
    uint2* tmp = A; A = B; B = tmp;
  }
  // Equal depth bits must come out in ascending surfel index (the reference's STABLE sort keeps its emission order,
  // rasterizer_impl.cu:70-111,304-309), and the scatter appended in arbitrary order.  Ties are rare (cloned surfels:
  // runs of 2), so short runs are fixed in place; a segment with a run longer than TIE_RUN (a fronto-parallel planar
  // cloud: thousands of equal view depths in one tile) is instead re-sorted as a whole on (depth, surfel): LSD passes
  // over the surfel-index span first, then the depth passes again -- O(n) per pass, no O(run^2) cliff.
  constexpr uint32_t TIE_RUN = 4;
  bool long_run = false;
  for (uint32_t i = t; i + TIE_RUN < n; i += THREADS) long_run |= A[i].x == A[i + TIE_RUN].x;
  if (__syncthreads_or(long_run)) {
    uint32_t imin = 0xffffffffu, imax = 0u;
    for (uint32_t i = t; i < n; i += THREADS) { const uint32_t k = A[i].y & ID_MASK; imin = min(imin, k); imax = max(imax, k); }
    imin = __reduce_min_sync(0xffffffffu, imin);
    imax = __reduce_max_sync(0xffffffffu, imax);
    if (lane == 0) { warp_tot[warp] = imax; warp_min[warp] = imin; }
    __syncthreads();
    for (int w = 0; w < WARPS; w++) { imax = max(imax, warp_tot[w]); imin = min(imin, warp_min[w]); }
    __syncthreads();
    const uint32_t ispan = imax - imin;
    const int ibits_all = ispan ? 32 - __clz(ispan) : 0;
    const int ipasses = (ibits_all + MAXBITS - 1) / MAXBITS;
    const int ibits = ipasses ? (ibits_all + ipasses - 1) / ipasses : 0;
    for (int shift = 0; shift < ibits_all; shift += ibits) {
      radix_pass<THREADS, MAXBITS, ITEMS, true>(A, B, n, imin, shift, ibits, whist, warp_tot, t, lane, warp, in_smem);
      uint2* tmp = A; A = B; B = tmp;
    }
    for (int shift = 0; shift < nbits; shift += bits) {
      radix_pass<THREADS, MAXBITS, ITEMS, false>(A, B, n, kmin, shift, bits, whist, warp_tot, t, lane, warp, in_smem);
      uint2* tmp = A; A = B; B = tmp;
    }
  } else {
    for (uint32_t i = t; i + 1 < n; i += THREADS) {
      const uint32_t k = A[i].x;
      if (A[i + 1].x == k && (i == 0 || A[i - 1].x != k)) {
        uint32_t j = i + 1;
        while (j + 1 < n && A[j + 1].x == k) j++;
        for (uint32_t a = i + 1; a <= j; a++) {  // insertion sort of the run's <= TIE_RUN surfel indices
          const uint32_t v = A[a].y;
          uint32_t b = a;
          while (b > i && (A[b - 1].y & ID_MASK) > (v & ID_MASK)) { A[b].y = A[b - 1].y; b--; }
          A[b].y = v;
        }
      }
    }
  }
  __syncthreads();
  const int tx = (int)(tile % (uint32_t)p.grid_x) * TILE_X, ty = (int)(tile / (uint32_t)p.grid_x) * TILE_Y;
  for (uint32_t i = t; i < n; i += THREADS) {
    const uint2 e = A[i];  // (in the global class A may alias keys_sorted[range.x + i]: read before write, same thread)
    const uint32_t surfel = p.packed ? (e.y & ID_MASK) : e.y;
    p.point_list[range.x + i] = surfel;
    p.keys_sorted[range.x + i] = ((uint64_t)tile << 32) | e.x;
    uint32_t mask = 0xffu;
    if (p.packed) mask = e.y >> ID_BITS;
    else if (!p.no_cull) mask = instance_block_mask(__ldg(p.bbox + surfel), __ldg(p.cull + surfel), tx, ty);
    p.strip_mask[range.x + i] = (uint8_t)mask;
  }
}

void launch_tile_sort(const uint2* ranges, uint2* pairs, uint64_t* keys_sorted, uint32_t* point_list,
                      uint8_t* strip_mask, const short4* bbox, const float4* cull, int grid_x, int num_tiles,
                      int no_cull, int packed, const ForwardStatus* status, cudaStream_t stream) {
  auto small = tile_sort_kernel<TS_SMALL_THREADS, TS_SMALL_BITS, (TS_CAP_SMALL + TS_SMALL_THREADS - 1) / TS_SMALL_THREADS>;
  auto large = tile_sort_kernel<TS_LARGE_THREADS, TS_LARGE_BITS, (TS_CAP_LARGE + TS_LARGE_THREADS - 1) / TS_LARGE_THREADS>;
  constexpr size_t hist_small = (size_t)(TS_SMALL_THREADS / 32) * (1u << TS_SMALL_BITS) * 4;
  constexpr size_t hist_large = (size_t)(TS_LARGE_THREADS / 32) * (1u << TS_LARGE_BITS) * 4;
  static PerDeviceLevel configured;  // the dynamic shared-memory limit is a per-device attribute
  configured.ensure(1, [&] {
    cudaFuncSetAttribute(small, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * TS_CAP_SMALL * sizeof(uint2) + hist_small));
    cudaFuncSetAttribute(large, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * TS_CAP_LARGE * sizeof(uint2) + hist_large));
  });
  TileSortParams p{};
  p.ranges = ranges; p.pairs = pairs; p.keys_sorted = keys_sorted; p.point_list = point_list;
  p.strip_mask = strip_mask; p.bbox = bbox; p.cull = cull; p.grid_x = grid_x; p.no_cull = no_cull; p.packed = packed; p.status = status;
  p.n_lo = 0; p.n_hi = TS_CAP_SMALL; p.cap = TS_CAP_SMALL;
  small<<<num_tiles, TS_SMALL_THREADS, 2 * TS_CAP_SMALL * sizeof(uint2) + hist_small, stream>>>(p);
  p.n_lo = TS_CAP_SMALL; p.n_hi = TS_CAP_LARGE; p.cap = TS_CAP_LARGE;
  large<<<num_tiles, TS_LARGE_THREADS, 2 * TS_CAP_LARGE * sizeof(uint2) + hist_large, stream>>>(p);
  p.n_lo = TS_CAP_LARGE; p.n_hi = 0xffffffffu; p.cap = 0;
  large<<<num_tiles, TS_LARGE_THREADS, hist_large, stream>>>(p);
}

}  // namespace gsm

// Forward per-tile compositing (K4) for sm_100a.
//
// Restates cuda_rasterizer/forward.cu:256-448 (renderCUDA): depth-ordered alpha compositing of
// colour, depth, normal, distortion moments and median depth, one thread per pixel.
// Differences in HOW (not in WHAT):
//   * surfel data arrive as one packed 80-byte record (5 x 128-bit loads per staged surfel
//     instead of 16 scalar loads from 4 arrays + 3 global loads per contributing pair);
//   * batches are double-buffered in shared memory with cp.async so the gather of batch i+1
//     overlaps the blending of batch i;
//   * termination is voted per warp (__all_sync) inside a batch and per CTA between batches.
// The arithmetic up to alpha and the transmittance test uses explicit rn intrinsics in the
// reference build's association so that every discrete decision (skip / contribute / stop /
// median) matches the reference bit for bit.
#include "common.cuh"
#include <cuda_pipeline.h>

namespace gsm {

#define MUL(a, b) __fmul_rn((a), (b))
#define ADD(a, b) __fadd_rn((a), (b))
#define SUB(a, b) __fsub_rn((a), (b))
#define FMA(a, b, c) __fmaf_rn((a), (b), (c))

// Warp-independent design: a CTA holds FWD_WARPS warps, each of which composites one 8x4-pixel block of
// a tile ENTIRELY on its own -- it scans the tile's strip masks (1 byte per list entry, streamed),
// compacts the entries that can reach its strip into a small position queue, gathers their packed
// records into its private double-buffered shared-memory stage with cp.async, and blends.  There is
// no __syncthreads anywhere: a warp with a short (heavily culled or early-terminated) list retires
// at once and the hardware block scheduler backfills, instead of waiting at a per-batch barrier
// for the slowest strip of the tile.
// GSM_FWD_WALK (see the chunk loop): 2 = entry-parallel pre-test + per-pixel walk of the candidates (production);
// 1 = pixel-parallel pre-test + per-pixel walk; 0 = one entry at a time for the whole warp (round 1).  Measured at C4
// (5 M surfels, 1237x822), bit-identical images in all three: 1.19 ms (0), 1.00 ms (1), 0.95 ms (2); 3840x2160: 4.17 /
// 4.24 / 4.04 ms.  GSM_FWD_MIN_CTAS = 8 caps the kernel at 64 registers: without it the build of variant 2 takes 72
// and runs at 1.7 ms.
#ifndef GSM_FWD_WALK
#define GSM_FWD_WALK 2
#endif
#ifndef GSM_FWD_WARPS
#define GSM_FWD_WARPS 4
#endif
constexpr int FWD_WARPS = GSM_FWD_WARPS;
constexpr int FWD_CHUNK = 32;   // entries blended per stage
constexpr int FWD_QUEUE = 256;  // position queue (power of two, >= FWD_CHUNK - 1 + 128: a scan round queues up to 128)

struct FwdWarpSmem {
  float4 q[2][REC_F4][FWD_CHUNK];  // [stage][field][entry]  5 KB
  uint32_t pos[2][FWD_CHUNK];      // 1-based position of the staged entry in the tile list
  uint32_t queue[FWD_QUEUE];
  uint32_t hit[FWD_CHUNK];         // per staged entry: the pixels of the block that blended it
  float cut[FWD_CHUNK];            // per staged entry: upper bound of rho for alpha >= 1/255 (pre-test below)
};

// HANDOVER: collect what the backward needs (per-entry pixel sets, dead block bits); false for inference renders
#ifndef GSM_FWD_MIN_CTAS
#define GSM_FWD_MIN_CTAS 8
#endif
template <bool STATS, bool HANDOVER>
__global__ void __launch_bounds__(FWD_WARPS * 32, GSM_FWD_MIN_CTAS) render_fwd_kernel(const FwdRenderParams p) {
  __shared__ __align__(16) FwdWarpSmem smem_all[FWD_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  FwdWarpSmem& sm = smem_all[warp];

  const uint32_t strip_global = blockIdx.x * FWD_WARPS + warp;
  const uint32_t tile = strip_global / STRIPS_PER_TILE, strip = strip_global % STRIPS_PER_TILE;
  if (tile >= (uint32_t)(p.grid_x * p.grid_y)) return;
  const uint32_t tile_x = tile % (uint32_t)p.grid_x, tile_y = tile / (uint32_t)p.grid_x;
  const uint32_t pixx = tile_x * TILE_X + (uint32_t)block_px((int)strip, lane);
  const uint32_t pixy = tile_y * TILE_Y + (uint32_t)block_py((int)strip, lane);
  const bool inside = pixx < (uint32_t)p.W && pixy < (uint32_t)p.H;
  const float pxf = (float)pixx, pyf = (float)pixy;
  uint2 range = __ldg(p.ranges + tile);
  if (p.status->overflow) range = make_uint2(0u, 0u);  // R exceeded the binning capacity: the frame is rendered empty
  const uint32_t total = range.y - range.x;
  const uint32_t strip_bit = 1u << strip;

  bool done = !inside;
  float T = 1.0f;
  uint32_t contributor = 0, last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, N0 = 0.f, N1 = 0.f, N2 = 0.f;
  float D = 0.f, M1 = 0.f, M2 = 0.f, distortion = 0.f, median_depth = 0.f;
  uint32_t median_contributor = 0;  // the reference converts a float -1 to u32 when never set: undefined
                                    // (garbage in its build); only read where last_contributor > 0
  uint32_t n_contributing = 0, n_evaluated = 0;  // STATS only
  bool stopped = false;                          // STATS only: terminated by the transmittance test
  const float far_ratio = __uint_as_float(GSM_FAR_RATIO_BITS);

  // ---- scan / queue / stage machinery (all warp-uniform state) ----
  uint32_t scan = 0;           // next list offset (0-based within the tile list) to scan
  uint32_t qhead = 0, qcount = 0;
  auto fill = [&]() {          // scan strip masks until a full chunk is queued or the list ends
    // 128 entries per round: every lane reads one aligned 32-bit word of the byte-per-entry masks (a quarter of the
    // dependent loads of a byte per lane; this loop held 12 % of the kernel's stall samples), four ballots keep the
    // list order.  (Plain loads: other blocks' warps clear THEIR bits of the same words meanwhile.)
    const uint32_t lt = (1u << lane) - 1u;
    while (qcount < FWD_CHUNK && scan < total) {
      const uint32_t first = range.x + scan;          // absolute position of the next entry to scan
      const uint32_t wpos = (first & ~3u) + 4u * (uint32_t)lane;
      uint32_t wv = 0u;
      if (wpos < range.y) wv = *reinterpret_cast<const volatile uint32_t*>(p.strip_mask + wpos);
      bool hit[4];
      unsigned m[4];
#pragma unroll
      for (int b = 0; b < 4; b++) {
        const uint32_t pos = wpos + (uint32_t)b;
        hit[b] = pos >= first && pos < range.y && ((wv >> (8 * b)) & strip_bit);
        m[b] = __ballot_sync(0xffffffffu, hit[b]);
      }
      uint32_t at = qhead + qcount + __popc(m[0] & lt) + __popc(m[1] & lt) + __popc(m[2] & lt) + __popc(m[3] & lt);
#pragma unroll
      for (int b = 0; b < 4; b++)
        if (hit[b]) sm.queue[(at++) & (FWD_QUEUE - 1)] = wpos + (uint32_t)b - range.x + 1u;
      qcount += __popc(m[0]) + __popc(m[1]) + __popc(m[2]) + __popc(m[3]);
      scan = (first & ~3u) + 128u - range.x;
    }
    __syncwarp();
  };
  auto issue = [&](int s) -> int {  // gather up to FWD_CHUNK queued entries into stage s
    const int n = (int)min(qcount, (uint32_t)FWD_CHUNK);
    if (lane < n) {
      const uint32_t pos1 = sm.queue[(qhead + lane) & (FWD_QUEUE - 1)];
      sm.pos[s][lane] = pos1;
      const uint32_t id = __ldg(p.point_list + range.x + pos1 - 1u);
      const float4* src = p.records + (size_t)id * REC_F4;
#pragma unroll
      for (int f = 0; f < REC_F4; f++) __pipeline_memcpy_async(&sm.q[s][f][lane], src + f, sizeof(float4));
    }
    __pipeline_commit();
    qhead += n;
    qcount -= n;
    return n;
  };

  fill();
  int n_cur = issue(0);
  int s = 0;
  while (n_cur > 0) {
    fill();
    const int n_next = issue(s ^ 1);  // prefetch the next chunk while this one is blended
    __pipeline_wait_prior(1);
    __syncwarp();
    const float4* __restrict__ q0 = sm.q[s][0];
    const float4* __restrict__ q1 = sm.q[s][1];
    const float4* __restrict__ q2 = sm.q[s][2];
    const float4* __restrict__ q3 = sm.q[s][3];
    const float4* __restrict__ q4 = sm.q[s][4];
    const uint32_t* __restrict__ qpos = sm.pos[s];
    bool all_done = false;
    if (HANDOVER) sm.hit[lane] = 0u;
    // alpha = min(0.99, o exp(-rho/2)) >= 1/255 needs rho <= 2 ln(255 o); 0.1 % + 1e-3 covers every rounding of
    // the exact evaluation below (expf, the products, the two divisions: a few ulp).  Negative for o < 1/255.
    if (lane < n_cur) sm.cut[lane] = p.no_cull ? 3.0e38f : FMA(2.0f * __logf(255.0f * q2[lane].w), 1.001f, 1.0e-3f);
    __syncwarp();
#if GSM_FWD_WALK
    // Two phases per chunk.  (1) EVERY lane runs the cheap geometric pre-test of its pixel against every staged entry,
    // at full lane occupancy, and keeps the entries that can reach alpha >= 1/255 as a bit set.  (2) Each lane then
    // walks ITS OWN candidates in list order through the exact evaluation: the warp's trip count is the busiest
    // pixel's number of candidates in the chunk (~5 of 32 entries), not the number of entries some pixel of the block
    // passes (~17 of 32), which is what one-entry-at-a-time costs with 5-6 of 32 lanes active in the expensive part.
    uint32_t mine = 0;
#if GSM_FWD_WALK == 2
    // (1) with LANE = ENTRY: lane j finds the pixels of the block that can reach alpha >= 1/255 for staged entry j:
    // the same pre-test, with the same expressions (k = px Tw - Tu and l = py Tw - Tv by one FMA each, p = k x l), as
    // the pixel-parallel form below -- but k is shared by the 4 pixels of a column and l by the 8 pixels of a row, the
    // record is read once per entry instead of once per (entry, pixel), and nothing diverges: ~19 warp-instructions per
    // entry instead of 37.  (An expansion of rho3d <= cut into a quadratic per row is cheaper still, ~7, but cancels
    // catastrophically for needle-thin splats: it failed the needle scene of test_conservative_culling_changes_nothing.)
    {
      uint32_t pm = 0u;
      if (lane < n_cur) {
        const float4 a = q0[lane], b = q1[lane], c = q2[lane];
        const float cut = sm.cut[lane];
        const float bxf = pxf - (float)(lane % BLOCK_W), byf = pyf - (float)(lane / BLOCK_W);   // block origin
        float lx[BLOCK_H], ly[BLOCK_H], lz[BLOCK_H], dyv[BLOCK_H];
#pragma unroll
        for (int r = 0; r < BLOCK_H; r++) {
          const float y = byf + (float)r;
          lx[r] = FMA(y, c.x, -b.x); ly[r] = FMA(y, c.y, -b.y); lz[r] = FMA(y, c.z, -b.z);
          dyv[r] = SUB(b.w, y);
        }
#pragma unroll
        for (int col = 0; col < BLOCK_W; col++) {
          const float x = bxf + (float)col;
          const float kx = FMA(x, c.x, -a.x), ky = FMA(x, c.y, -a.y), kz = FMA(x, c.z, -a.z);
          const float dxv = SUB(a.w, x), dx2 = MUL(dxv, dxv);
#pragma unroll
          for (int r = 0; r < BLOCK_H; r++) {
            const float ppx = FMA(ky, lz[r], -MUL(kz, ly[r]));
            const float ppy = FMA(kz, lx[r], -MUL(kx, lz[r]));
            const float ppz = FMA(kx, ly[r], -MUL(ky, lx[r]));
            const float r2 = FMA(dyv[r], dyv[r], dx2);
            const float rho2d = ADD(r2, r2);
            const bool fail = (ppx * ppx + ppy * ppy) > cut * (ppz * ppz) && rho2d > cut;   // (NaNs pass)
            pm |= (fail ? 0u : 1u) << (r * BLOCK_W + col);
          }
        }
      }
      // (2) transpose (five shuffles): every lane collects the entries ITS pixel passed
      mine = warp_transpose_bits(pm, lane);
    }
#else
    for (int j = 0; j < n_cur; j++) {
      const float4 a = q0[j], b = q1[j], c = q2[j];
      const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
      const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
      const float ppx = FMA(ky, lz, -MUL(kz, ly));
      const float ppy = FMA(kz, lx, -MUL(kx, lz));
      const float ppz = FMA(kx, ly, -MUL(ky, lx));
      const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
      const float r2 = FMA(dy, dy, MUL(dx, dx));
      const float rho2d = ADD(r2, r2);
      const float cut = sm.cut[j];
      // (the same conservative pre-test as below: NaNs pass)
      const bool fail = (ppx * ppx + ppy * ppy) > cut * (ppz * ppz) && rho2d > cut;
      mine |= (fail ? 0u : 1u) << j;
    }
#endif
    if (STATS && !done) n_evaluated += (uint32_t)n_cur;
    if (done) mine = 0u;
    while (__any_sync(0xffffffffu, mine != 0u)) {
      if (mine == 0u) continue;
      const int j = __ffs(mine) - 1;
      mine &= mine - 1u;
      contributor = qpos[j];  // position in the tile list, 1-based (forward.cu:350)
      const float4 a = q0[j], b = q1[j], c = q2[j];
      const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
      const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
      const float ppx = FMA(ky, lz, -MUL(kz, ly));
      const float ppy = FMA(kz, lx, -MUL(kx, lz));
      const float ppz = FMA(kx, ly, -MUL(ky, lx));
      const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
      const float r2 = FMA(dy, dy, MUL(dx, dx));
      const float rho2d = ADD(r2, r2);
      if (ppz == 0.0f) continue;
      const float sx = __fdiv_rn(ppx, ppz), sy = __fdiv_rn(ppy, ppz);
      const float rho3d = FMA(sx, sx, MUL(sy, sy));
      const float rho = fminf(rho3d, rho2d);
      const float depth = ADD(c.z, FMA(c.x, sx, MUL(c.y, sy)));
      if (depth < NEAR_N) continue;
      const float power = MUL(rho, -0.5f);
      if (power > 0.0f) continue;
      const float alpha = fminf(MUL(c.w, expf(power)), ALPHA_MAX);
      if (alpha < ALPHA_MIN) continue;
      const float test_T = MUL(T, SUB(1.0f, alpha));
      if (test_T < T_EPS) { done = true; mine = 0u; if (STATS) stopped = true; continue; }

      const float w = MUL(T, alpha);
      const float A = SUB(1.0f, T);
      const float m = MUL(ADD(__fdiv_rn(-NEAR_N, depth), 1.0f), far_ratio);
      const float mm = MUL(m, m);
      distortion = FMA(w, FMA(-M1, ADD(m, m), FMA(A, mm, M2)), distortion);
      D = FMA(depth, w, D);
      M1 = FMA(m, w, M1);
      M2 = FMA(mm, w, M2);
      if (T > 0.5f) { median_depth = depth; median_contributor = contributor; }
      const float4 nrm = q3[j], col = q4[j];
      N0 = FMA(w, nrm.x, N0); N1 = FMA(w, nrm.y, N1); N2 = FMA(w, nrm.z, N2);
      C0 = FMA(w, col.x, C0); C1 = FMA(w, col.y, C1); C2 = FMA(w, col.z, C2);
      T = test_T;
      last_contributor = contributor;
      if (HANDOVER) atomicOr(&sm.hit[j], 1u << lane);
      if (STATS) n_contributing++;
    }
    if (__all_sync(0xffffffffu, done)) all_done = true;
#else
    for (int j = 0; j < n_cur; j++) {
      if (__all_sync(0xffffffffu, done)) { all_done = true; break; }
      if (done) continue;
      contributor = qpos[j];  // position in the tile list, 1-based (forward.cu:350)
      if (STATS) n_evaluated++;
      const float4 a = q0[j], b = q1[j], c = q2[j];
      // k = px*Tw - Tu ; l = py*Tw - Tv ; p = k x l   (forward.cu:355-358), fused as the
      // reference's SASS fuses them (ptxas contracts the un-suffixed PTX mul/sub pairs)
      const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
      const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
      const float ppx = FMA(ky, lz, -MUL(kz, ly));
      const float ppy = FMA(kz, lx, -MUL(kx, lz));
      const float ppz = FMA(kx, ly, -MUL(ky, lx));
      const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
      const float r2 = FMA(dy, dy, MUL(dx, dx));
      const float rho2d = ADD(r2, r2);
      {
        // pre-test before the two IEEE divisions and the exponential: can this pixel reach alpha >= 1/255 at all?
        // rho3d = (ppx^2 + ppy^2) / ppz^2, compared without dividing (NaNs pass).  More than half of the queued
        // entries fail it for every pixel of the block (their box meets the block, their footprint does not), and
        // then the warp skips the rest of the iteration.
        const float cut = sm.cut[j];
        if ((ppx * ppx + ppy * ppy) > cut * (ppz * ppz) && rho2d > cut) continue;
      }
      if (ppz == 0.0f) continue;
      const float sx = __fdiv_rn(ppx, ppz), sy = __fdiv_rn(ppy, ppz);
      const float rho3d = FMA(sx, sx, MUL(sy, sy));
      const float rho = fminf(rho3d, rho2d);
      const float depth = ADD(c.z, FMA(c.x, sx, MUL(c.y, sy)));
      if (depth < NEAR_N) continue;
      const float power = MUL(rho, -0.5f);
      if (power > 0.0f) continue;
      const float alpha = fminf(MUL(c.w, expf(power)), ALPHA_MAX);
      if (alpha < ALPHA_MIN) continue;
      const float test_T = MUL(T, SUB(1.0f, alpha));
      if (test_T < T_EPS) { done = true; if (STATS) stopped = true; continue; }

      const float w = MUL(T, alpha);
      const float A = SUB(1.0f, T);
      const float m = MUL(ADD(__fdiv_rn(-NEAR_N, depth), 1.0f), far_ratio);
      const float mm = MUL(m, m);
      distortion = FMA(w, FMA(-M1, ADD(m, m), FMA(A, mm, M2)), distortion);
      D = FMA(depth, w, D);
      M1 = FMA(m, w, M1);
      M2 = FMA(mm, w, M2);
      if (T > 0.5f) { median_depth = depth; median_contributor = contributor; }
      const float4 nrm = q3[j], col = q4[j];
      N0 = FMA(w, nrm.x, N0); N1 = FMA(w, nrm.y, N1); N2 = FMA(w, nrm.z, N2);
      C0 = FMA(w, col.x, C0); C1 = FMA(w, col.y, C1); C2 = FMA(w, col.z, C2);
      T = test_T;
      last_contributor = contributor;
      if (HANDOVER) atomicOr(&sm.hit[j], 1u << lane);
      if (STATS) n_contributing++;
    }
#endif
    // Hand the backward pass what this pass found out.  The backward's activity test of a pair is this kernel's
    // contribution test, bit for bit (pixels that stopped earlier are excluded by their last contributor), so
    //  * an entry that no pixel of this block blended is a no-op there too: this block's bit of the entry's mask
    //    is cleared and the backward never queues it (57 % of the queued (block, entry) pairs at C4);
    //  * for the others the 32-bit set of blending pixels is stored (pix_mask[block][entry]): the backward walks,
    //    per pixel, exactly its own contributors.
    // Only this warp reads this bit / writes this word, and it is past these entries.
    if (HANDOVER) {
      __syncwarp();
      const uint32_t my_mask = sm.hit[lane];  // lane j: the pixels that blended entry j
      if (lane < n_cur) {
        const size_t at = (size_t)range.x + qpos[lane] - 1u;
        if (my_mask) p.pix_mask[(size_t)strip * p.pix_plane + at] = my_mask;
        else atomicAnd(reinterpret_cast<unsigned int*>(p.strip_mask + (at & ~(size_t)3)), ~(strip_bit << (8u * (unsigned)(at & 3))));
      }
    }
    if (all_done) break;
    __syncwarp();  // every lane is done reading stage s before it is refilled two issues from now
    n_cur = n_next;
    s ^= 1;
  }
  __pipeline_wait_prior(0);

  if (STATS) {  // work counters for the roofline model (bench only)
    // [0] pairs the reference's loop evaluates (list position at which the pixel stops, forward.cu:344-424)
    // [1] contributing pairs  [2] records staged  [3] pairs this kernel evaluated after culling
    unsigned long long e = inside ? (stopped ? contributor : (uint32_t)total) : 0u, c = n_contributing, v = n_evaluated;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      e += __shfl_xor_sync(0xffffffffu, e, o);
      c += __shfl_xor_sync(0xffffffffu, c, o);
      v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    if (lane == 0) { atomicAdd(p.stats + 0, e); atomicAdd(p.stats + 1, c); atomicAdd(p.stats + 3, v); }
    if (lane == 0 && strip == 0) atomicAdd(p.stats + 2, (unsigned long long)total);
  }

  if (inside) {
    const size_t N = (size_t)p.W * p.H;
    const size_t pix = (size_t)pixy * p.W + pixx;
    p.final_T[pix] = T;
    p.final_T[pix + N] = M1;
    p.final_T[pix + 2 * N] = M2;
    p.n_contrib[pix] = last_contributor;
    p.n_contrib[pix + N] = median_contributor;
    p.out_color[pix] = FMA(T, __ldg(p.background + 0), C0);
    p.out_color[pix + N] = FMA(T, __ldg(p.background + 1), C1);
    p.out_color[pix + 2 * N] = FMA(T, __ldg(p.background + 2), C2);
    p.out_others[pix] = D;
    p.out_others[pix + N] = SUB(1.0f, T);
    p.out_others[pix + 2 * N] = N0;
    p.out_others[pix + 3 * N] = N1;
    p.out_others[pix + 4 * N] = N2;
    p.out_others[pix + 5 * N] = median_depth;
    p.out_others[pix + 6 * N] = distortion;
  }
}

void launch_render_fwd(const FwdRenderParams& p, cudaStream_t stream) {
  const unsigned strips = (unsigned)(p.grid_x * p.grid_y) * STRIPS_PER_TILE;
  const unsigned grid = (strips + FWD_WARPS - 1) / FWD_WARPS;
  if (p.stats && p.pix_mask != nullptr) render_fwd_kernel<true, true><<<grid, FWD_WARPS * 32, 0, stream>>>(p);
  else if (p.stats) render_fwd_kernel<true, false><<<grid, FWD_WARPS * 32, 0, stream>>>(p);
  else if (p.pix_mask == nullptr) render_fwd_kernel<false, false><<<grid, FWD_WARPS * 32, 0, stream>>>(p);
  else render_fwd_kernel<false, true><<<grid, FWD_WARPS * 32, 0, stream>>>(p);
}

}  // namespace gsm

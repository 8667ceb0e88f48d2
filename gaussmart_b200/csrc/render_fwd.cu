// Forward per-tile compositing (K4) for sm_100a.
//
// Restates cuda_rasterizer/forward.cu:256-448 (renderCUDA): one CTA per 16x16 tile, one thread
// per pixel, depth-ordered alpha compositing of colour, depth, normal, distortion moments and
// median depth.  Differences in HOW (not in WHAT):
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

constexpr int FWD_BATCH = 256;

struct FwdSmem {
  float4 q[2][REC_F4][FWD_BATCH];       // [stage][field][entry]   2*5*256*16 = 40 KB
  short4 bbox[2][FWD_BATCH];            // conservative contribution boxes of the staged entries
  uint8_t list[TILE_PIX / 32][FWD_BATCH];  // per-warp compacted entry indices (ascending = depth order)
};

__device__ __forceinline__ void stage_batch(float4 (*q)[FWD_BATCH], short4* sbox,
                                            const float4* __restrict__ records, const short4* __restrict__ bbox,
                                            const uint32_t* __restrict__ point_list, uint32_t begin,
                                            uint32_t end, int t) {
  const uint32_t i = begin + t;
  if (i < end) {
    const uint32_t id = __ldg(point_list + i);
    const float4* src = records + (size_t)id * REC_F4;
#pragma unroll
    for (int f = 0; f < REC_F4; f++) __pipeline_memcpy_async(&q[f][t], src + f, sizeof(float4));
    __pipeline_memcpy_async(&sbox[t], bbox + id, sizeof(short4));
  }
}

template <bool STATS>
__global__ void __launch_bounds__(TILE_PIX) render_fwd_kernel(const FwdRenderParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FwdSmem& sm = *reinterpret_cast<FwdSmem*>(smem_raw);

  const int t = threadIdx.x;
  const uint32_t pixx = blockIdx.x * TILE_X + (t & 15);
  const uint32_t pixy = blockIdx.y * TILE_Y + (t >> 4);
  const bool inside = pixx < (uint32_t)p.W && pixy < (uint32_t)p.H;
  const float pxf = (float)pixx, pyf = (float)pixy;
  const uint2 range = __ldg(p.ranges + blockIdx.y * p.grid_x + blockIdx.x);
  const int total = (int)(range.y - range.x);
  const int rounds = (total + FWD_BATCH - 1) / FWD_BATCH;

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

  // this warp's pixel strip (16 x 2), inclusive integer pixel bounds
  const int warp = t >> 5, lane = t & 31;
  const int sx_lo = blockIdx.x * TILE_X, sx_hi = sx_lo + TILE_X - 1;
  const int sy_lo = blockIdx.y * TILE_Y + 2 * warp, sy_hi = sy_lo + 1;
  uint8_t* __restrict__ mylist = sm.list[warp];

  if (rounds > 0) stage_batch(sm.q[0], sm.bbox[0], p.records, p.bbox, p.point_list, range.x, range.y, t);
  __pipeline_commit();

  int toDo = total;
  for (int i = 0; i < rounds; i++, toDo -= FWD_BATCH) {
    const int s = i & 1;
    // batch i has landed for this thread's own copies; the CTA-wide vote (forward.cu:325-327)
    // doubles as the barrier that publishes stage s and proves stage s^1 (read in round i-1) free
    __pipeline_wait_prior(0);
    if (__syncthreads_count(done) == TILE_PIX) break;
    if (i + 1 < rounds)
      stage_batch(sm.q[s ^ 1], sm.bbox[s ^ 1], p.records, p.bbox, p.point_list,
                  range.x + (uint32_t)(i + 1) * FWD_BATCH, range.y, t);
    __pipeline_commit();

    const int n = min(FWD_BATCH, toDo);
    const float4* __restrict__ q0 = sm.q[s][0];
    const float4* __restrict__ q1 = sm.q[s][1];
    const float4* __restrict__ q2 = sm.q[s][2];
    const float4* __restrict__ q3 = sm.q[s][3];
    const float4* __restrict__ q4 = sm.q[s][4];
    if (!__all_sync(0xffffffffu, done)) {
      // compact, per warp, the entries whose contribution box meets this warp's strip; entries
      // outside cannot pass the alpha test for any of its pixels, and a skipped entry has no side
      // effect in the reference either (every `continue` of forward.cu:355-392 precedes all updates)
      int cnt = 0;
#pragma unroll
      for (int r = 0; r < FWD_BATCH / 32; r++) {
        const int j = r * 32 + lane;
        bool hit = false;
        if (j < n) {
          const short4 bx = sm.bbox[s][j];
          hit = p.no_cull || (bx.x <= sx_hi && bx.z >= sx_lo && bx.y <= sy_hi && bx.w >= sy_lo);
        }
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (hit) mylist[cnt + __popc(m & ((1u << lane) - 1u))] = (uint8_t)j;
        cnt += __popc(m);
      }
      __syncwarp();
      const uint32_t base = (uint32_t)i * FWD_BATCH + 1u;
      for (int e = 0; e < cnt; e++) {
        if (__all_sync(0xffffffffu, done)) break;
        if (done) continue;
        const int j = mylist[e];
        contributor = base + (uint32_t)j;  // position in the tile list, 1-based (forward.cu:350)
        if (STATS) n_evaluated++;
        const float4 a = q0[j], b = q1[j], c = q2[j];
        // k = px*Tw - Tu ; l = py*Tw - Tv ; p = k x l   (forward.cu:355-358), fused as the
        // reference's SASS fuses them (ptxas contracts the un-suffixed PTX mul/sub pairs)
        const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
        const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
        const float ppx = FMA(ky, lz, -MUL(kz, ly));
        const float ppy = FMA(kz, lx, -MUL(kx, lz));
        const float ppz = FMA(kx, ly, -MUL(ky, lx));
        if (ppz == 0.0f) continue;
        const float sx = __fdiv_rn(ppx, ppz), sy = __fdiv_rn(ppy, ppz);
        const float rho3d = FMA(sx, sx, MUL(sy, sy));
        const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
        const float r2 = FMA(dy, dy, MUL(dx, dx));
        const float rho2d = ADD(r2, r2);
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
        if (STATS) n_contributing++;
      }
    }
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
    if ((t & 31) == 0) { atomicAdd(p.stats + 0, e); atomicAdd(p.stats + 1, c); atomicAdd(p.stats + 3, v); }
    if (t == 0) atomicAdd(p.stats + 2, (unsigned long long)total);
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
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(render_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FwdSmem));
    cudaFuncSetAttribute(render_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FwdSmem));
    configured = true;
  }
  dim3 grid(p.grid_x, p.grid_y, 1);
  if (p.stats) render_fwd_kernel<true><<<grid, TILE_PIX, sizeof(FwdSmem), stream>>>(p);
  else render_fwd_kernel<false><<<grid, TILE_PIX, sizeof(FwdSmem), stream>>>(p);
}

}  // namespace gsm

// Backward per-tile compositing (K5) for sm_100a.
//
// Restates cuda_rasterizer/backward.cu:143-446 (renderCUDA): the tile's depth-sorted list is
// replayed back to front, T is reconstructed by division (backward.cu:319) and the per-pair
// gradients w.r.t. the homography rows Tu/Tv/Tw, the AABB centre, colour, normal and opacity are
// produced.  Differences in HOW:
//   * the replay starts at the tile's deepest contributor (max n_contrib over the tile, and per
//     warp over the warp's 32 pixels) instead of the end of the list;
//   * the 18 per-pair partial gradients are reduced across the warp with a recursive-halving
//     shuffle network (21 shuffles) and leave the SM as ONE red.global.add per value per
//     warp-surfel instead of one atomic per value per pixel-surfel (backward.cu:339-443);
//   * they land in a compact 80-byte per-surfel gradient record that the backward preprocess
//     expands, so only 80*P bytes are zero-filled per backward instead of 76 floats * P.
#include "common.cuh"
#include <cuda_pipeline.h>

namespace gsm {

#define MUL(a, b) __fmul_rn((a), (b))
#define ADD(a, b) __fadd_rn((a), (b))
#define SUB(a, b) __fsub_rn((a), (b))
#define FMA(a, b, c) __fmaf_rn((a), (b), (c))

// Warp-independent design (see render_fwd.cu): each warp replays one 16x2 strip back to front on its
// own -- scans the strip masks downwards from its deepest contributor, queues the entries that can
// reach the strip, gathers their records into a private double-buffered stage, and reduces its 32
// pixels' partial gradients with shuffles before one red.global.add per value.  No __syncthreads.
#ifndef GSM_BWD_WARPS
#define GSM_BWD_WARPS 4
#endif
constexpr int BWD_WARPS = GSM_BWD_WARPS;
constexpr int BWD_CHUNK = 32;
constexpr int BWD_QUEUE = 128;

struct BwdWarpSmem {
  float4 q[2][REC_F4][BWD_CHUNK];
  uint32_t id[2][BWD_CHUNK];
  uint32_t pos[2][BWD_CHUNK];  // 0-based list position k of the staged entry (the reference's `contributor`)
  uint32_t queue[BWD_QUEUE];
};

// Recursive-halving reduce-scatter of 16 values over 32 lanes + 2 extra values.
// On return lane L holds in v[0] the warp sum of slot (L>>1)&15 (both lanes of a pair hold it),
// e0 holds the warp sum of extra slot (L>>4) (16 or 17).
__device__ __forceinline__ void warp_reduce_scatter18(float (&v)[16], float& e0, float e1, int lane) {
  const unsigned FULL = 0xffffffffu;
  {
    const bool up = lane & 16;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const float send = up ? v[i] : v[i + 8];
      const float keep = up ? v[i + 8] : v[i];
      v[i] = keep + __shfl_xor_sync(FULL, send, 16);
    }
    const float send = up ? e0 : e1;
    const float keep = up ? e1 : e0;
    e0 = keep + __shfl_xor_sync(FULL, send, 16);
  }
  {
    const bool up = lane & 8;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float send = up ? v[i] : v[i + 4];
      const float keep = up ? v[i + 4] : v[i];
      v[i] = keep + __shfl_xor_sync(FULL, send, 8);
    }
    e0 += __shfl_xor_sync(FULL, e0, 8);
  }
  {
    const bool up = lane & 4;
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const float send = up ? v[i] : v[i + 2];
      const float keep = up ? v[i + 2] : v[i];
      v[i] = keep + __shfl_xor_sync(FULL, send, 4);
    }
    e0 += __shfl_xor_sync(FULL, e0, 4);
  }
  {
    const bool up = lane & 2;
    const float send = up ? v[0] : v[1];
    const float keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(FULL, send, 2);
    e0 += __shfl_xor_sync(FULL, e0, 2);
  }
  v[0] += __shfl_xor_sync(FULL, v[0], 1);
  e0 += __shfl_xor_sync(FULL, e0, 1);
}

#ifndef GSM_BWD_MIN_CTAS
#define GSM_BWD_MIN_CTAS 5
#endif
__global__ void __launch_bounds__(BWD_WARPS * 32, GSM_BWD_MIN_CTAS) render_bwd_kernel(const BwdRenderParams p) {
  __shared__ __align__(16) BwdWarpSmem smem_all[BWD_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  BwdWarpSmem& sm = smem_all[warp];

  const uint32_t strip_global = blockIdx.x * BWD_WARPS + warp;
  const uint32_t tile = strip_global / STRIPS_PER_TILE, strip = strip_global % STRIPS_PER_TILE;
  if (tile >= (uint32_t)(p.grid_x * p.grid_y)) return;
  const uint32_t tile_x = tile % (uint32_t)p.grid_x, tile_y = tile / (uint32_t)p.grid_x;
  const uint32_t pixx = tile_x * TILE_X + (uint32_t)block_px((int)strip, lane);
  const uint32_t pixy = tile_y * TILE_Y + (uint32_t)block_py((int)strip, lane);
  const bool inside = pixx < (uint32_t)p.W && pixy < (uint32_t)p.H;
  const float pxf = (float)pixx, pyf = (float)pixy;
  const size_t N = (size_t)p.W * p.H;
  const size_t pix = (size_t)pixy * p.W + pixx;
  const uint2 range = __ldg(p.ranges + tile);
  const uint32_t strip_bit = 1u << strip;

  const float T_final = inside ? __ldg(p.final_T + pix) : 0.0f;
  const float final_D = inside ? __ldg(p.final_T + pix + N) : 0.0f;
  const float final_D2 = inside ? __ldg(p.final_T + pix + 2 * N) : 0.0f;
  const float final_A = 1.0f - T_final;
  const int last_contributor = inside ? (int)__ldg(p.n_contrib + pix) : 0;
  const int median_contributor = inside ? (int)__ldg(p.n_contrib + pix + N) : 0;

  float dL_dpixel[3] = {0.f, 0.f, 0.f}, dL_dnormal2D[3] = {0.f, 0.f, 0.f};
  float dL_ddepth = 0.f, dL_daccum = 0.f, dL_dreg = 0.f, dL_dmedian_depth = 0.f;
  if (inside) {
#pragma unroll
    for (int ch = 0; ch < 3; ch++) dL_dpixel[ch] = __ldg(p.dL_dpix + ch * N + pix);
    dL_ddepth = __ldg(p.dL_dothers + 0 * N + pix);
    dL_daccum = __ldg(p.dL_dothers + 1 * N + pix);
#pragma unroll
    for (int ch = 0; ch < 3; ch++) dL_dnormal2D[ch] = __ldg(p.dL_dothers + (2 + ch) * N + pix);
    dL_dmedian_depth = __ldg(p.dL_dothers + 5 * N + pix);
    dL_dreg = __ldg(p.dL_dothers + 6 * N + pix);
  }
  float bg_dot_dpixel = 0.f;
#pragma unroll
  for (int ch = 0; ch < 3; ch++) bg_dot_dpixel += __ldg(p.background + ch) * dL_dpixel[ch];

  // deepest contributor of the warp: entries at or beyond it are behind every pixel of the strip
  const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);

  float T = T_final;
  float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
  float last_alpha = 0.f, last_depth = 0.f, last_normal[3] = {0.f, 0.f, 0.f};
  float accum_depth_rec = 0.f, accum_alpha_rec = 0.f, accum_normal_rec[3] = {0.f, 0.f, 0.f};
  float last_dL_dT = 0.f;
  const float far_ratio = __uint_as_float(GSM_FAR_RATIO_BITS);

  // ---- scan / queue / stage machinery (warp-uniform state); list offsets run DOWN from warp_last-1 ----
  int scan = warp_last - 1;  // next 0-based list offset to scan (descending); < 0 => exhausted
  uint32_t qhead = 0, qcount = 0;
  auto fill = [&]() {
    while (qcount < BWD_CHUNK && scan >= 0) {
      const int o = scan - lane;
      const bool hit = o >= 0 && (__ldg(p.strip_mask + range.x + o) & strip_bit);
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (hit) sm.queue[(qhead + qcount + __popc(m & ((1u << lane) - 1u))) & (BWD_QUEUE - 1)] = (uint32_t)o;
      qcount += __popc(m);
      scan -= 32;
    }
    __syncwarp();
  };
  auto issue = [&](int s) -> int {
    const int n = (int)min(qcount, (uint32_t)BWD_CHUNK);
    if (lane < n) {
      const uint32_t o = sm.queue[(qhead + lane) & (BWD_QUEUE - 1)];
      sm.pos[s][lane] = o;
      const uint32_t id = __ldg(p.point_list + range.x + o);
      sm.id[s][lane] = id;
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
    const int n_next = issue(s ^ 1);
    __pipeline_wait_prior(1);
    __syncwarp();
    for (int j = 0; j < n_cur; j++) {
      const int k = (int)sm.pos[s][j];  // == the reference's `contributor` after its decrement
      bool active = k < last_contributor;
      const float4 a = sm.q[s][0][j], b = sm.q[s][1][j], c = sm.q[s][2][j];
      const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
      const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
      const float ppx = FMA(ky, lz, -MUL(kz, ly));
      const float ppy = FMA(kz, lx, -MUL(kx, lz));
      const float ppz = FMA(kx, ly, -MUL(ky, lx));
      active = active && (ppz != 0.0f);
      const float sx = __fdiv_rn(ppx, ppz), sy = __fdiv_rn(ppy, ppz);
      const float rho3d = FMA(sx, sx, MUL(sy, sy));
      const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
      const float r2 = FMA(dy, dy, MUL(dx, dx));
      const float rho2d = ADD(r2, r2);
      const float rho = fminf(rho3d, rho2d);
      const float c_d = ADD(c.z, FMA(c.x, sx, MUL(c.y, sy)));
      active = active && !(c_d < NEAR_N);
      const float power = MUL(rho, -0.5f);
      active = active && !(power > 0.0f);
      const float G = expf(power);
      const float alpha = fminf(MUL(c.w, G), ALPHA_MAX);
      active = active && !(alpha < ALPHA_MIN);
      if (!__any_sync(0xffffffffu, active)) continue;

      float v[16];
      float e0 = 0.f, e1 = 0.f;
#pragma unroll
      for (int z = 0; z < 16; z++) v[z] = 0.f;

      if (active) {
        const float4 nrm = sm.q[s][3][j], col = sm.q[s][4][j];
        const float normal[3] = {nrm.x, nrm.y, nrm.z};
        const float color[3] = {col.x, col.y, col.z};
        // T is reconstructed by IEEE division exactly as the reference does (backward.cu:319): its error
        // compounds along the list.  The other quotients below feed single gradient terms (bar: 1e-4 rel)
        // and use the 2-ulp MUFU reciprocal instead of the reference's 8 div.rn per pair.
        const float one_m_alpha = SUB(1.0f, alpha);
        T = __fdiv_rn(T, one_m_alpha);
        const float w = alpha * T;
        float dL_dalpha = 0.0f;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = color[ch];
          dL_dalpha += (color[ch] - accum_rec[ch]) * dL_dpixel[ch];
          v[11 + ch] = w * dL_dpixel[ch];
        }
        float dL_dz = 0.0f;
        const float near_over_d = __fdiv_rn(NEAR_N, c_d);  // exact: m_d enters a cancelling sum below
        const float m_d = far_ratio * (1.0f - near_over_d);
        const float dmd_dd = (FAR_N / ((FAR_N - NEAR_N) * NEAR_N)) * near_over_d * near_over_d;
        if (k == median_contributor - 1) dL_dz += dL_dmedian_depth;
        const float dL_dweight = (final_D2 + m_d * m_d * final_A - 2.0f * m_d * final_D) * dL_dreg;
        dL_dalpha += dL_dweight - last_dL_dT;
        last_dL_dT = dL_dweight * alpha + (1.0f - alpha) * last_dL_dT;
        const float dL_dmd = 2.0f * (T * alpha) * (m_d * final_A - final_D) * dL_dreg;
        dL_dz += dL_dmd * dmd_dd;

        accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
        last_depth = c_d;
        dL_dalpha += (c_d - accum_depth_rec) * dL_ddepth;
        accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
        dL_dalpha += (1.0f - accum_alpha_rec) * dL_daccum;
        float dn[3];
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
          last_normal[ch] = normal[ch];
          dL_dalpha += (normal[ch] - accum_normal_rec[ch]) * dL_dnormal2D[ch];
          dn[ch] = alpha * T * dL_dnormal2D[ch];
        }
        v[14] = dn[0]; v[15] = dn[1]; e0 = dn[2];

        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final * __fdividef(1.0f, one_m_alpha)) * bg_dot_dpixel;

        const float dL_dG = c.w * dL_dalpha;
        dL_dz += alpha * T * dL_ddepth;

        if (rho3d <= rho2d) {
          const float dL_dsx = dL_dG * -G * sx + dL_dz * c.x;
          const float dL_dsy = dL_dG * -G * sy + dL_dz * c.y;
          const float rpz = __fdividef(1.0f, ppz);
          const float dsx_pz = dL_dsx * rpz, dsy_pz = dL_dsy * rpz;
          const float dpx = dsx_pz, dpy = dsy_pz, dpz = -(dsx_pz * sx + dsy_pz * sy);
          // dL_dk = l x dL_dp ; dL_dl = dL_dp x k
          const float dkx = ly * dpz - lz * dpy, dky = lz * dpx - lx * dpz, dkz = lx * dpy - ly * dpx;
          const float dlx = dpy * kz - dpz * ky, dly = dpz * kx - dpx * kz, dlz = dpx * ky - dpy * kx;
          v[0] = -dkx; v[1] = -dky; v[2] = -dkz;
          v[3] = -dlx; v[4] = -dly; v[5] = -dlz;
          v[6] = pxf * dkx + pyf * dlx + dL_dz * sx;
          v[7] = pxf * dky + pyf * dly + dL_dz * sy;
          v[8] = pxf * dkz + pyf * dlz + dL_dz;
        } else {
          const float dG_ddelx = -G * FILTER_INV_SQUARE * dx;
          const float dG_ddely = -G * FILTER_INV_SQUARE * dy;
          v[9] = dL_dG * dG_ddelx;
          v[10] = dL_dG * dG_ddely;
          v[6] = sx * dL_dz;
          v[7] = sy * dL_dz;
          v[8] = dL_dz;
        }
        e1 = G * dL_dalpha;
      }

      warp_reduce_scatter18(v, e0, e1, lane);
      float* g = p.grad_records + (size_t)sm.id[s][j] * GREC;
      if ((lane & 1) == 0) atomicAdd(g + (lane >> 1), v[0]);
      else if (lane == 1) atomicAdd(g + 16, e0);
      else if (lane == 17) atomicAdd(g + 17, e0);
    }
    __syncwarp();
    n_cur = n_next;
    s ^= 1;
  }
  __pipeline_wait_prior(0);
}

void launch_render_bwd(const BwdRenderParams& p, cudaStream_t stream) {
  const unsigned strips = (unsigned)(p.grid_x * p.grid_y) * STRIPS_PER_TILE;
  const unsigned grid = (strips + BWD_WARPS - 1) / BWD_WARPS;
  render_bwd_kernel<<<grid, BWD_WARPS * 32, 0, stream>>>(p);
}

}  // namespace gsm

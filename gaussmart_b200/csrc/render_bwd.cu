// Backward per-tile compositing (K5) for sm_100a.
//
// Restates cuda_rasterizer/backward.cu:143-446 (renderCUDA): the tile's depth-sorted list is
// replayed back to front, T is reconstructed by division (backward.cu:319) and the per-pair
// gradients w.r.t. the homography rows Tu/Tv/Tw, the AABB centre, colour, normal and opacity are
// produced.  Differences in HOW:
//   * each warp owns one 8x4-pixel block of a tile and replays it on its own (no __syncthreads),
//     starting at the block's deepest contributor instead of the end of the list;
//   * the per-pixel SEQUENTIAL part of a pair (transmittance, running colour/depth/normal sums,
//     dL/dalpha) runs in the pixel's lane as the entries go by, but the pair's GEOMETRIC part (the
//     chain through the ray-splat intersection to Tu/Tv/Tw, ~2/3 of the arithmetic) is deferred:
//     active pairs -- typically 4-5 of a block's 32 pixels per entry -- are pushed as 16-byte
//     records into a small shared-memory ring, one entry per quarter-warp slot, and processed
//     four entries at a time with every quarter-warp working on its own entry;
//   * the 18 per-pair partial gradients are reduced inside the quarter-warp with a recursive-halving
//     shuffle network (17 shuffles for 4 entries) and leave the SM as one red.global.add.v2.f32 per
//     lane into a compact 80-byte per-surfel gradient record, instead of one atomic per value per
//     pixel-surfel pair (backward.cu:339-443); the backward preprocess expands the record, so only
//     80*P bytes are zero-filled per backward instead of 76 floats * P.
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
#ifndef GSM_BWD_CHUNK
#define GSM_BWD_CHUNK 32
#endif
constexpr int BWD_CHUNK = GSM_BWD_CHUNK;  // entries per stage (<= 32)
constexpr int BWD_QUEUE = 64;             // position queue (power of two, >= BWD_CHUNK + 32)
static_assert(BWD_CHUNK <= 32 && BWD_CHUNK + 32 <= BWD_QUEUE, "queue too small");

#ifndef GSM_BWD_SLOTS
#define GSM_BWD_SLOTS 32
#endif
constexpr int SLOT_W = 4;                      // lanes (pairs) per slot: an entry takes ceil(pairs / 4) slots
constexpr int SLOTS_PER_PASS = 32 / SLOT_W;    // 8 entries-slots retired per warp pass
constexpr int BWD_SLOTS = GSM_BWD_SLOTS;       // slots of deferred pairs per batch of entries (one entry takes <= 8)
static_assert(BWD_SLOTS >= 32 / SLOT_W && BWD_SLOTS % SLOTS_PER_PASS == 0, "a batch must hold at least one entry");

struct BwdWarpSmem {
  float4 q[2][REC_F4][BWD_CHUNK];
  uint32_t id[2][BWD_CHUNK];
  uint32_t pos[2][BWD_CHUNK];    // 0-based list position k of the staged entry (the reference's `contributor`)
  uint32_t pmask[2][BWD_CHUNK];  // pixels of this block that blended the entry (from the forward pass)
  uint32_t queue[BWD_QUEUE];
  float4 pend[BWD_SLOTS][SLOT_W];  // deferred pairs: (G*dL_dalpha, dL_dz, alpha*T, bits: pixel | 3d-branch << 5)
  uint32_t qinfo[BWD_SLOTS];     // per slot: stage entry j | pairs in the slot << 8
  uint32_t sbase[BWD_CHUNK];     // first slot of entry j (chunk-wide numbering)
  float4 pixc[2][32];            // per pixel: (dL_dpixel.rgb, dL_dnormal.x), (dL_dnormal.yz, -, -)
};

// red.global.add.v2.f32: two adjacent floats in one L2 reduction request (8-byte aligned)
__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}
// red.global.add.v4.f32: four adjacent floats (16-byte aligned)
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Deferred geometric part of up to 8 entries at once: lanes 4q..4q+3 work on slot first+q (< nslots valid), lane i of
// the four on the slot's i-th pair.  Recomputes the ray-splat intersection of the pair (MUFU
// reciprocal: these values feed gradient terms only), forms the 18 partial gradients, reduces them
// over the slot's 4 lanes and adds them to the entry's gradient record (one red.global.add.v4.f32 per lane).
__device__ __forceinline__ void geometry_pass(BwdWarpSmem& sm, int s, int first, int nslots, int lane, int bx0, int by0,
                                              float* __restrict__ grad_records) {
  const unsigned FULL = 0xffffffffu;
  const int qtr = lane / SLOT_W, i = lane % SLOT_W;
  const int slot = min(first + qtr, BWD_SLOTS - 1);
  const uint32_t qf = sm.qinfo[slot];
  const int j = (int)(qf & 31u);
  const bool slot_valid = qtr < nslots;
  const bool valid = slot_valid && i < (int)(qf >> 8);
  float v[16];
  float e0 = 0.f, e1 = 0.f;
#pragma unroll
  for (int z = 0; z < 16; z++) v[z] = 0.f;
  if (valid) {
    const float4 pr = sm.pend[slot][i];
    const uint32_t bits = __float_as_uint(pr.w);
    const int pix = (int)(bits & 31u);
    const float pxf = (float)(bx0 + pix % BLOCK_W), pyf = (float)(by0 + pix / BLOCK_W);
    const float4 a = sm.q[s][0][j], b = sm.q[s][1][j], c = sm.q[s][2][j];
    const float4 pc0 = sm.pixc[0][pix], pc1 = sm.pixc[1][pix];
    const float dL_dz = pr.y, w = pr.z;
    e1 = pr.x;
    const float cw_e1 = c.w * e1;  // dL_dG * G
    const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
    const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
    const float ppx = FMA(ky, lz, -MUL(kz, ly));
    const float ppy = FMA(kz, lx, -MUL(kx, lz));
    const float ppz = FMA(kx, ly, -MUL(ky, lx));
    const float rpz = __fdividef(1.0f, ppz);
    const float sx = ppx * rpz, sy = ppy * rpz;
    v[11] = w * pc0.x; v[12] = w * pc0.y; v[13] = w * pc0.z;
    v[14] = w * pc0.w; v[15] = w * pc1.x; e0 = w * pc1.y;
    if (bits & 32u) {  // rho3d <= rho2d: gradient through the ray-splat intersection
      const float dL_dsx = dL_dz * c.x - cw_e1 * sx;
      const float dL_dsy = dL_dz * c.y - cw_e1 * sy;
      const float dpx = dL_dsx * rpz, dpy = dL_dsy * rpz, dpz = -(dpx * sx + dpy * sy);
      // dL_dk = l x dL_dp ; dL_dl = dL_dp x k
      const float dkx = ly * dpz - lz * dpy, dky = lz * dpx - lx * dpz, dkz = lx * dpy - ly * dpx;
      const float dlx = dpy * kz - dpz * ky, dly = dpz * kx - dpx * kz, dlz = dpx * ky - dpy * kx;
      v[0] = -dkx; v[1] = -dky; v[2] = -dkz;
      v[3] = -dlx; v[4] = -dly; v[5] = -dlz;
      v[6] = pxf * dkx + pyf * dlx + dL_dz * sx;
      v[7] = pxf * dky + pyf * dly + dL_dz * sy;
      v[8] = pxf * dkz + pyf * dlz + dL_dz;
    } else {  // low-pass branch: gradient to the AABB centre
      const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
      v[9] = cw_e1 * (-FILTER_INV_SQUARE * dx);
      v[10] = cw_e1 * (-FILTER_INV_SQUARE * dy);
      v[6] = sx * dL_dz;
      v[7] = sy * dL_dz;
      v[8] = dL_dz;
    }
  }
  // recursive halving over the slot's 4 lanes: lane i ends with the sums of values 4i .. 4i+3
  {
    const bool up = lane & 2;
#pragma unroll
    for (int z = 0; z < 8; z++) {
      const float send = up ? v[z] : v[z + 8];
      const float keep = up ? v[z + 8] : v[z];
      v[z] = keep + __shfl_xor_sync(FULL, send, 2);
    }
    const float send = up ? e0 : e1;
    const float keep = up ? e1 : e0;
    e0 = keep + __shfl_xor_sync(FULL, send, 2);  // lanes 0-1 of the slot: value 16, lanes 2-3: value 17
  }
  {
    const bool up = lane & 1;
#pragma unroll
    for (int z = 0; z < 4; z++) {
      const float send = up ? v[z] : v[z + 4];
      const float keep = up ? v[z + 4] : v[z];
      v[z] = keep + __shfl_xor_sync(FULL, send, 1);
    }
    e0 += __shfl_xor_sync(FULL, e0, 1);
  }
  const float e17 = __shfl_down_sync(FULL, e0, 2);  // lane 0 of the slot fetches value 17 from lane 2
  if (slot_valid) {
    float* g = grad_records + (size_t)sm.id[s][j] * GREC;
    red_add_v4(g + 4 * i, v[0], v[1], v[2], v[3]);
    if (i == 0) red_add_v2(g + 16, e0, e17);
  }
}

#ifndef GSM_BWD_MIN_CTAS
#define GSM_BWD_MIN_CTAS 6
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

  sm.pixc[0][lane] = make_float4(dL_dpixel[0], dL_dpixel[1], dL_dpixel[2], dL_dnormal2D[0]);
  sm.pixc[1][lane] = make_float4(dL_dnormal2D[1], dL_dnormal2D[2], 0.f, 0.f);
  const int bx0 = (int)pixx - (lane % BLOCK_W), by0 = (int)pixy - (lane / BLOCK_W);  // block origin
  // deepest contributor of the warp: entries at or beyond it are behind every pixel of the block
  const int warp_last = __reduce_max_sync(0xffffffffu, last_contributor);

  float T = T_final;
  float accum_rec[3] = {0.f, 0.f, 0.f}, last_color[3] = {0.f, 0.f, 0.f};
  float last_alpha = 0.f, last_depth = 0.f, last_normal[3] = {0.f, 0.f, 0.f};
  float accum_depth_rec = 0.f, accum_alpha_rec = 0.f, accum_normal_rec[3] = {0.f, 0.f, 0.f};
  float last_dL_dT = 0.f;
  const float far_ratio = __uint_as_float(GSM_FAR_RATIO_BITS);
  const unsigned FULL = 0xffffffffu;
  const uint32_t* __restrict__ pix_plane = p.pix_mask + (size_t)strip * p.pix_plane;

  // ---- scan / queue / stage machinery (warp-uniform state); list offsets run DOWN from warp_last-1 ----
  int scan = warp_last - 1;  // next 0-based list offset to scan (descending); < 0 => exhausted
  uint32_t qhead = 0, qcount = 0;
  auto fill = [&]() {
    while (qcount < BWD_CHUNK && scan >= 0) {
      const int o = scan - lane;
      const bool hit = o >= 0 && (__ldg(p.strip_mask + range.x + o) & strip_bit);
      const unsigned m = __ballot_sync(FULL, hit);
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
      sm.pmask[s][lane] = __ldg(pix_plane + range.x + o);
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

    // ---- the chunk's pixel masks: lane j holds entry j's, every lane collects the entries of ITS pixel ----
    const uint32_t emask = lane < n_cur ? sm.pmask[s][lane] : 0u;   // pixels that blended entry `lane`
    const uint32_t mine = warp_transpose_bits(emask, lane);         // entries of the chunk that my pixel blended
    // slots: entry j takes ceil(popc/8) quarter-warp slots; exclusive prefix over the entries
    const int cnt = __popc(emask), nq = (cnt + SLOT_W - 1) / SLOT_W;
    int incl = nq;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(FULL, incl, o);
      if (lane >= o) incl += v;
    }
    const int excl = incl - nq;
    if (lane < n_cur) sm.sbase[lane] = (uint32_t)excl;
    __syncwarp();

    // ---- batches of consecutive entries whose slots fit the pending buffer ----
    int jb = 0;
    while (jb < n_cur) {
      const int base = __shfl_sync(FULL, excl, jb);
      const unsigned fits = __ballot_sync(FULL, lane >= jb && lane < n_cur && incl - base <= BWD_SLOTS);
      // entries jb .. je-1 (contiguous run of `fits` from jb; at least one, an entry takes <= 4 slots)
      const unsigned run = ~(fits >> jb);
      const int je = jb + (run ? __ffs(run) - 1 : 32 - jb);
      const int nslots = __shfl_sync(FULL, incl, je - 1) - base;
      if (lane >= jb && lane < je) {
        for (int t = 0; t < nq; t++) sm.qinfo[excl - base + t] = (uint32_t)lane | ((uint32_t)min(SLOT_W, cnt - SLOT_W * t) << 8);
      }

      // ---- sequential part: every lane walks its own pixel's entries of the batch, back to front ----
      uint32_t todo = mine & (je >= 32 ? FULL : ((1u << je) - 1u)) & ~((1u << jb) - 1u);
      while (todo) {
        const int j = __ffs(todo) - 1;
        todo &= todo - 1u;
        const int k = (int)sm.pos[s][j];  // == the reference's `contributor` after its decrement
        const float4 a = sm.q[s][0][j], b = sm.q[s][1][j], c = sm.q[s][2][j];
        const float kx = FMA(pxf, c.x, -a.x), ky = FMA(pxf, c.y, -a.y), kz = FMA(pxf, c.z, -a.z);
        const float lx = FMA(pyf, c.x, -b.x), ly = FMA(pyf, c.y, -b.y), lz = FMA(pyf, c.z, -b.z);
        const float ppx = FMA(ky, lz, -MUL(kz, ly));
        const float ppy = FMA(kz, lx, -MUL(kx, lz));
        const float ppz = FMA(kx, ly, -MUL(ky, lx));
        const float sx = __fdiv_rn(ppx, ppz), sy = __fdiv_rn(ppy, ppz);
        const float rho3d = FMA(sx, sx, MUL(sy, sy));
        const float dx = SUB(a.w, pxf), dy = SUB(b.w, pyf);
        const float r2 = FMA(dy, dy, MUL(dx, dx));
        const float rho2d = ADD(r2, r2);
        const float rho = fminf(rho3d, rho2d);
        const float c_d = ADD(c.z, FMA(c.x, sx, MUL(c.y, sy)));
        const float power = MUL(rho, -0.5f);
        const float G = expf(power);
        const float alpha = fminf(MUL(c.w, G), ALPHA_MAX);
        // (the forward blended this pair, so every test of backward.cu:283-316 passes: same expressions, same bits)

        const float4 nrm = sm.q[s][3][j], col = sm.q[s][4][j];
        const float normal[3] = {nrm.x, nrm.y, nrm.z};
        const float color[3] = {col.x, col.y, col.z};
        // T is reconstructed by IEEE division exactly as the reference does (backward.cu:319): its error
        // compounds along the list.  The other quotients below feed single gradient terms (bar: 1e-4 rel)
        // and use the 2-ulp MUFU reciprocal instead of the reference's 8 div.rn per pair.
        const float one_m_alpha = SUB(1.0f, alpha);
        T = __fdiv_rn(T, one_m_alpha);
        const float w = alpha * T;
        float dL_dalpha = 0.0f, dL_dz = 0.0f;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = color[ch];
          dL_dalpha += (color[ch] - accum_rec[ch]) * dL_dpixel[ch];
        }
        const float near_over_d = __fdiv_rn(NEAR_N, c_d);  // exact: m_d enters a cancelling sum below
        const float m_d = far_ratio * (1.0f - near_over_d);
        const float dmd_dd = (FAR_N / ((FAR_N - NEAR_N) * NEAR_N)) * near_over_d * near_over_d;
        if (k == median_contributor - 1) dL_dz += dL_dmedian_depth;
        const float dL_dweight = (final_D2 + m_d * m_d * final_A - 2.0f * m_d * final_D) * dL_dreg;
        dL_dalpha += dL_dweight - last_dL_dT;
        last_dL_dT = dL_dweight * alpha + (1.0f - alpha) * last_dL_dT;
        const float dL_dmd = 2.0f * w * (m_d * final_A - final_D) * dL_dreg;
        dL_dz += dL_dmd * dmd_dd;

        accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
        last_depth = c_d;
        dL_dalpha += (c_d - accum_depth_rec) * dL_ddepth;
        accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
        dL_dalpha += (1.0f - accum_alpha_rec) * dL_daccum;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
          last_normal[ch] = normal[ch];
          dL_dalpha += (normal[ch] - accum_normal_rec[ch]) * dL_dnormal2D[ch];
        }
        dL_dalpha *= T;
        last_alpha = alpha;
        dL_dalpha += (-T_final * __fdividef(1.0f, one_m_alpha)) * bg_dot_dpixel;
        dL_dz += w * dL_ddepth;
        const float e1 = G * dL_dalpha;  // dL_dopacity of the pair; dL_dG * G = opacity * e1

        // my place among the entry's pairs: slot of the entry + rank of my pixel in its mask
        const uint32_t em = sm.pmask[s][j];
        const int rank = __popc(em & ((1u << lane) - 1u));
        const int slot = (int)sm.sbase[j] - base + rank / SLOT_W;
        sm.pend[slot][rank % SLOT_W] =
            make_float4(e1, dL_dz, w, __uint_as_float((uint32_t)lane | (rho3d <= rho2d ? 32u : 0u)));
      }
      __syncwarp();
      for (int first = 0; first < nslots; first += SLOTS_PER_PASS)
        geometry_pass(sm, s, first, min(SLOTS_PER_PASS, nslots - first), lane, bx0, by0, p.grad_records);
      __syncwarp();  // the slots are refilled by the next batch
      jb = je;
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

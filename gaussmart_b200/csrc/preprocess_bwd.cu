// Backward per-surfel preprocess (K6) for sm_100a.
//
// Restates cuda_rasterizer/backward.cu:20-139 (SH backward), :449-579 (compute_transmat_aabb) and
// :581-636 (preprocessCUDA), plus auxiliary.h:128-138 (dnormvdv) and :238-282 (quaternion VJP),
// including the reference's deliberate deviations from autograd: scale modifier
// forced to 1, rsqrt normaliser treated as constant, W/H re-derived from focal*tan in fp32, and
// dL_dmean2D overwritten by the densification heuristic.
// It reads the compact gradient record written by the backward compositing kernel and writes
// EVERY public gradient row (zeros for culled surfels), so no output needs a prior memset.
#include "common.cuh"
#include <cstdlib>

namespace gsm {

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 mk(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

__device__ __forceinline__ void store3(float* dst, f3 v, bool acc = false) {
  if (acc) { dst[0] += v.x; dst[1] += v.y; dst[2] += v.z; }
  else { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }
}

// SH basis coefficients of sh_backward (dL_dsh[k] = coef[k] * dL_dRGB), same expressions, for the
// normalised direction (x, y, z); coefficients beyond degree D are zero.
__device__ __forceinline__ void sh_coefficients(int D, float x, float y, float z, float (&coef)[16]) {
#pragma unroll
  for (int k = 0; k < 16; k++) coef[k] = 0.f;
  const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
  coef[0] = C0;
  if (D > 0) {
    coef[1] = -C1 * y; coef[2] = C1 * z; coef[3] = -C1 * x;
    if (D > 1) {
      const float C20 = 1.0925484305920792f, C21 = -1.0925484305920792f, C22 = 0.31539156525252005f,
                  C23 = -1.0925484305920792f, C24 = 0.5462742152960396f;
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      coef[4] = C20 * xy; coef[5] = C21 * yz; coef[6] = C22 * (2.f * zz - xx - yy);
      coef[7] = C23 * xz; coef[8] = C24 * (xx - yy);
      if (D > 2) {
        const float C30 = -0.5900435899266435f, C31 = 2.890611442640554f, C32 = -0.4570457994644658f,
                    C33 = 0.3731763325901154f, C34 = -0.4570457994644658f, C35 = 1.445305721320277f,
                    C36 = -0.5900435899266435f;
        coef[9] = C30 * y * (3.f * xx - yy); coef[10] = C31 * xy * z; coef[11] = C32 * y * (4.f * zz - xx - yy);
        coef[12] = C33 * z * (2.f * zz - 3.f * xx - 3.f * yy); coef[13] = C34 * x * (4.f * zz - xx - yy);
        coef[14] = C35 * z * (xx - yy); coef[15] = C36 * x * (xx - 3.f * yy);
      }
    }
  }
}

// SH backward (backward.cu:20-139), in place on a shared-memory row: on entry `row` holds this
// surfel's SH coefficients (3M floats), on exit dL_dsh (zeros beyond the active degree).  Every
// coefficient is read before its slot is overwritten.
__device__ void sh_backward(int D, int M, float* row, f3 dir_orig, f3 dL_dRGB, f3& dL_dmean_add) {
  float* dL_dsh = row;
  const float* sh = row;
  const float len = __fsqrt_rn(__fmaf_rn(dir_orig.z, dir_orig.z, __fmaf_rn(dir_orig.x, dir_orig.x, __fmul_rn(dir_orig.y, dir_orig.y))));
  const float x = __fdiv_rn(dir_orig.x, len), y = __fdiv_rn(dir_orig.y, len), z = __fdiv_rn(dir_orig.z, len);
  const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
  f3 dx = mk(0, 0, 0), dy = mk(0, 0, 0), dz = mk(0, 0, 0);
#define SH(k) mk(sh[3 * (k)], sh[3 * (k) + 1], sh[3 * (k) + 2])
  int written = 1;
  if (D > 0) {
    dx = (-C1) * SH(3);
    dy = (-C1) * SH(1);
    dz = C1 * SH(2);
    written = 4;
    if (D > 1) {
      const float C20 = 1.0925484305920792f, C21 = -1.0925484305920792f, C22 = 0.31539156525252005f,
                  C23 = -1.0925484305920792f, C24 = 0.5462742152960396f;
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      const f3 s4 = SH(4), s5 = SH(5), s6 = SH(6), s7 = SH(7), s8 = SH(8);
      dx = dx + (C20 * y) * s4 + (C22 * 2.f * -x) * s6 + (C23 * z) * s7 + (C24 * 2.f * x) * s8;
      dy = dy + (C20 * x) * s4 + (C21 * z) * s5 + (C22 * 2.f * -y) * s6 + (C24 * 2.f * -y) * s8;
      dz = dz + (C21 * y) * s5 + (C22 * 2.f * 2.f * z) * s6 + (C23 * x) * s7;
      written = 9;
      if (D > 2) {
        const float C30 = -0.5900435899266435f, C31 = 2.890611442640554f, C32 = -0.4570457994644658f,
                    C33 = 0.3731763325901154f, C34 = -0.4570457994644658f, C35 = 1.445305721320277f,
                    C36 = -0.5900435899266435f;
        const f3 s9 = SH(9), s10 = SH(10), s11 = SH(11), s12 = SH(12), s13 = SH(13), s14 = SH(14), s15 = SH(15);
        dx = dx + ((C30 * 3.f * 2.f * xy) * s9 + (C31 * yz) * s10 + (C32 * -2.f * xy) * s11 +
                   (C33 * -3.f * 2.f * xz) * s12 + (C34 * (-3.f * xx + 4.f * zz - yy)) * s13 +
                   (C35 * 2.f * xz) * s14 + (C36 * 3.f * (xx - yy)) * s15);
        dy = dy + ((C30 * 3.f * (xx - yy)) * s9 + (C31 * xz) * s10 + (C32 * (-3.f * yy + 4.f * zz - xx)) * s11 +
                   (C33 * -3.f * 2.f * yz) * s12 + (C34 * -2.f * xy) * s13 + (C35 * -2.f * yz) * s14 +
                   (C36 * -3.f * 2.f * xy) * s15);
        dz = dz + ((C31 * xy) * s10 + (C32 * 4.f * 2.f * yz) * s11 + (C33 * 3.f * (2.f * zz - xx - yy)) * s12 +
                   (C34 * 4.f * 2.f * xz) * s13 + (C35 * (xx - yy)) * s14);
        written = 16;
      }
    }
  }
#undef SH
  // every coefficient above has been read: the row now becomes dL_dsh = coef (x) dL_dRGB (zeros beyond degree D)
  {
    float coef[16];
    sh_coefficients(D, x, y, z, coef);
#pragma unroll
    for (int k = 0; k < 16; k++)
      if (k < written) store3(dL_dsh + 3 * k, coef[k] * dL_dRGB);
  }
  for (int k = written * 3; k < M * 3; k++) dL_dsh[k] = 0.0f;

  const f3 dL_ddir = mk(dot(dx, dL_dRGB), dot(dy, dL_dRGB), dot(dz, dL_dRGB));
  // dnormvdv (auxiliary.h:128-138)
  const f3 v = dir_orig, dv = dL_ddir;
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  dL_dmean_add.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  dL_dmean_add.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  dL_dmean_add.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
}

// sh_backward over a 16-byte aligned shared-memory row accessed with 128-bit loads / stores (TMA bulk mode: the row
// arrived by cp.async.bulk and leaves the same way).  Same expressions as sh_backward; the coefficients are read degree
// by degree (at most 6 float4 live), every read precedes the first write.
__device__ __forceinline__ void sh_backward_row4(int D, int M, float4* row4, f3 dir_orig, f3 dL_dRGB, f3& dL_dmean_add) {
  const float len = __fsqrt_rn(__fmaf_rn(dir_orig.z, dir_orig.z, __fmaf_rn(dir_orig.x, dir_orig.x, __fmul_rn(dir_orig.y, dir_orig.y))));
  const float x = __fdiv_rn(dir_orig.x, len), y = __fdiv_rn(dir_orig.y, len), z = __fdiv_rn(dir_orig.z, len);
  const float C1 = 0.4886025119029199f;
  f3 dx = mk(0, 0, 0), dy = mk(0, 0, 0), dz = mk(0, 0, 0);
  int written = 1;
  if (D > 0) {
    {
      const float4 v0 = row4[0], v1 = row4[1], v2 = row4[2];   // floats 0..11: SH(1) = 3..5, SH(2) = 6..8, SH(3) = 9..11
      const f3 s1 = mk(v0.w, v1.x, v1.y), s2 = mk(v1.z, v1.w, v2.x), s3 = mk(v2.y, v2.z, v2.w);
      dx = (-C1) * s3;
      dy = (-C1) * s1;
      dz = C1 * s2;
    }
    written = 4;
    if (D > 1) {
      const float C20 = 1.0925484305920792f, C21 = -1.0925484305920792f, C22 = 0.31539156525252005f,
                  C23 = -1.0925484305920792f, C24 = 0.5462742152960396f;
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      {
        const float4 v3 = row4[3], v4 = row4[4], v5 = row4[5], v6 = row4[6];   // floats 12..27: SH(4) = 12..14 ... SH(8) = 24..26
        const f3 s4 = mk(v3.x, v3.y, v3.z), s5 = mk(v3.w, v4.x, v4.y), s6 = mk(v4.z, v4.w, v5.x), s7 = mk(v5.y, v5.z, v5.w),
                 s8 = mk(v6.x, v6.y, v6.z);
        dx = dx + (C20 * y) * s4 + (C22 * 2.f * -x) * s6 + (C23 * z) * s7 + (C24 * 2.f * x) * s8;
        dy = dy + (C20 * x) * s4 + (C21 * z) * s5 + (C22 * 2.f * -y) * s6 + (C24 * 2.f * -y) * s8;
        dz = dz + (C21 * y) * s5 + (C22 * 2.f * 2.f * z) * s6 + (C23 * x) * s7;
      }
      written = 9;
      if (D > 2) {
        const float C30 = -0.5900435899266435f, C31 = 2.890611442640554f, C32 = -0.4570457994644658f,
                    C33 = 0.3731763325901154f, C34 = -0.4570457994644658f, C35 = 1.445305721320277f,
                    C36 = -0.5900435899266435f;
        const float4 v6 = row4[6], v7 = row4[7], v8 = row4[8], v9 = row4[9], v10 = row4[10], v11 = row4[11];  // floats 24..47
        const f3 s9 = mk(v6.w, v7.x, v7.y), s10 = mk(v7.z, v7.w, v8.x), s11 = mk(v8.y, v8.z, v8.w), s12 = mk(v9.x, v9.y, v9.z),
                 s13 = mk(v9.w, v10.x, v10.y), s14 = mk(v10.z, v10.w, v11.x), s15 = mk(v11.y, v11.z, v11.w);
        dx = dx + ((C30 * 3.f * 2.f * xy) * s9 + (C31 * yz) * s10 + (C32 * -2.f * xy) * s11 +
                   (C33 * -3.f * 2.f * xz) * s12 + (C34 * (-3.f * xx + 4.f * zz - yy)) * s13 +
                   (C35 * 2.f * xz) * s14 + (C36 * 3.f * (xx - yy)) * s15);
        dy = dy + ((C30 * 3.f * (xx - yy)) * s9 + (C31 * xz) * s10 + (C32 * (-3.f * yy + 4.f * zz - xx)) * s11 +
                   (C33 * -3.f * 2.f * yz) * s12 + (C34 * -2.f * xy) * s13 + (C35 * -2.f * yz) * s14 +
                   (C36 * -3.f * 2.f * xy) * s15);
        dz = dz + ((C31 * xy) * s10 + (C32 * 4.f * 2.f * yz) * s11 + (C33 * 3.f * (2.f * zz - xx - yy)) * s12 +
                   (C34 * 4.f * 2.f * xz) * s13 + (C35 * (xx - yy)) * s14);
        written = 16;
      }
    }
  }
  // every coefficient has been read: the row becomes dL_dsh = coef (x) dL_dRGB (zeros beyond degree D)
  {
    float coef[16];
    sh_coefficients(D, x, y, z, coef);
    const float g[3] = {dL_dRGB.x, dL_dRGB.y, dL_dRGB.z};
    const int n4 = 3 * M / 4;
#pragma unroll
    for (int j = 0; j < 12; j++) {
      if (j < n4) {
        float o[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
          const int idx = 4 * j + e, k = idx / 3, c = idx % 3;
          o[e] = k < written ? coef[k] * g[c] : 0.0f;
        }
        row4[j] = make_float4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  const f3 dL_ddir = mk(dot(dx, dL_dRGB), dot(dy, dL_dRGB), dot(dz, dL_dRGB));
  // dnormvdv (auxiliary.h:128-138)
  const f3 v = dir_orig, dv = dL_ddir;
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  dL_dmean_add.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  dL_dmean_add.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  dL_dmean_add.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
}

constexpr int PBWD_THREADS = 128;

// SH rows travel through shared memory so that the 12*M-byte-per-surfel streams (shs in, dL_dsh out)
// are read and written as fully coalesced 128-bit accesses by the warp instead of 3M scalar accesses
// per thread at a 12*M-byte stride.  Row r of the warp sits at smem[r * (3M + 1)] (odd stride: the
// per-lane accesses of the SH polynomial are bank-conflict free).
// (rl floats per source row, landing at column col0 of the shared-memory row of stride rs: the full [P,M,3] tensor is
// rl = 3M, col0 = 0; split storage (features_dc [P,1,3] + features_rest [P,M-1,3], gaussian_model.py:115-118 without the
// torch.cat) is two calls, rl = 3 at column 0 and rl = 3M-3 at column 3)
__device__ __forceinline__ void sh_rows_load(float* wsm, const float* __restrict__ g, int rl, int rs, int col0, int rows,
                                             unsigned vis_mask, int lane) {
  const int total = rows * rl;
  const unsigned rmagic = 0xffffffffu / (unsigned)rl + 1u;  // idx / rl == umulhi(idx, rmagic) for idx < 2^16 (one division per thread instead of one per element)
  if (rows == 32 && (reinterpret_cast<size_t>(g) & 15) == 0) {
    const float4* g4 = reinterpret_cast<const float4*>(g);
    // batches of 4 independent 128-bit loads per lane are issued before any of them is consumed: the row stream is
    // 54 % of the kernel's read bytes and one load per loop iteration left the lanes waiting on HBM one at a time
#ifndef GSM_ROW_LOADS
#define GSM_ROW_LOADS 4
#endif
    constexpr int LOADS_IN_FLIGHT = GSM_ROW_LOADS;
    for (int f0 = lane; f0 < total / 4; f0 += 32 * LOADS_IN_FLIGHT) {
      float4 v[LOADS_IN_FLIGHT];
      bool take[LOADS_IN_FLIGHT];
#pragma unroll
      for (int u = 0; u < LOADS_IN_FLIGHT; u++) {
        const int f = f0 + 32 * u, e0 = 4 * f;
        const int r0 = (int)__umulhi((unsigned)e0, rmagic), r3 = (int)__umulhi((unsigned)(e0 + 3), rmagic);
        take[u] = f < total / 4 && (((vis_mask >> r0) & 1u) || ((vis_mask >> r3) & 1u));
        v[u] = take[u] ? __ldg(g4 + f) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < LOADS_IN_FLIGHT; u++) {
        if (!take[u]) continue;
        const int e0 = 4 * (f0 + 32 * u);
        const float vv[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int e = 0; e < 4; e++) {
          const int idx = e0 + e, r = (int)__umulhi((unsigned)idx, rmagic);
          wsm[r * rs + col0 + (idx - r * rl)] = vv[e];
        }
      }
    }
  } else {
    for (int idx = lane; idx < total; idx += 32) {
      const int r = (int)__umulhi((unsigned)idx, rmagic);
      if ((vis_mask >> r) & 1u) wsm[r * rs + col0 + (idx - r * rl)] = __ldg(g + idx);
    }
  }
}

__device__ __forceinline__ void sh_rows_store(const float* wsm, float* __restrict__ g, int rl, int rs, int col0, int rows,
                                              unsigned vis_mask, bool acc, int lane) {
  const int total = rows * rl;
  const unsigned rmagic = 0xffffffffu / (unsigned)rl + 1u;  // idx / rl == umulhi(idx, rmagic) for idx < 2^16 (one division per thread instead of one per element)
  if (rows == 32 && (reinterpret_cast<size_t>(g) & 15) == 0) {
    float4* g4 = reinterpret_cast<float4*>(g);
    for (int f = lane; f < total / 4; f += 32) {
      const int e0 = 4 * f;
      float vv[4];
      bool any = false;
#pragma unroll
      for (int e = 0; e < 4; e++) {
        const int idx = e0 + e, r = (int)__umulhi((unsigned)idx, rmagic);
        const bool vis = (vis_mask >> r) & 1u;
        vv[e] = vis ? wsm[r * rs + col0 + (idx - r * rl)] : 0.0f;
        any |= vis;
      }
      if (acc) {
        if (!any) continue;
        const float4 o = g4[f];
        g4[f] = make_float4(o.x + vv[0], o.y + vv[1], o.z + vv[2], o.w + vv[3]);
      } else {
        g4[f] = make_float4(vv[0], vv[1], vv[2], vv[3]);
      }
    }
  } else {
    for (int idx = lane; idx < total; idx += 32) {
      const int r = (int)__umulhi((unsigned)idx, rmagic);
      const bool vis = (vis_mask >> r) & 1u;
      const float v = vis ? wsm[r * rs + col0 + (idx - r * rl)] : 0.0f;
      if (acc) { if (vis) g[idx] += v; }
      else g[idx] = v;
    }
  }
}

#ifndef GSM_PBWD_MIN_CTAS
#define GSM_PBWD_MIN_CTAS 6
#endif
// RAW: the raw-parameter entry (activation chain, split SH storage); the standard entry compiles without it.
// BULK: SH rows in (and dL_dsh rows out) by per-row TMA bulk copies (standard entry, M % 4 == 0, aligned tensors): one
// instruction per row each way instead of the cooperative staging loops, the load in flight under the geometry math.
template <bool RAW, bool BULK>
__global__ void __launch_bounds__(PBWD_THREADS, GSM_PBWD_MIN_CTAS) preprocess_bwd_kernel(const BwdPreprocessParams p) {
  extern __shared__ __align__(16) float sh_smem[];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warp_first = idx - lane;
  if (warp_first >= p.P) return;  // whole warp out of range
  const bool valid = idx < p.P;
  const size_t i = (size_t)(valid ? idx : p.P - 1);
  const bool visible = valid && __ldg(p.radii + idx) > 0;
  const bool acc = p.accumulate != 0;  // outputs += (view-parallel accumulation over local views)
  const unsigned vis_mask = __ballot_sync(0xffffffffu, visible);
  const int rows = min(32, p.P - warp_first);
  const int row_stride = BULK ? SH_BULK_STRIDE : 3 * p.M + 1;
  float* wsm = sh_smem + (size_t)warp * 32 * row_stride;
  float* my_row = wsm + lane * row_stride;
  // factored mode (view-parallel): dL_dRGB is written instead of the 3M-float dL_dsh row, which the
  // receiver rebuilds as basis(dir) (x) dL_dRGB (sh_grad_expand_kernel); the SH term of dL_dmean3D stays here
  const bool factored = p.dL_drgb != nullptr || p.n_rgb_peers > 0;
  const bool has_sh = p.shs != nullptr && (p.dL_dsh != nullptr || factored);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sh_smem + (size_t)(PBWD_THREADS / 32) * 32 * SH_BULK_STRIDE) + warp;
  float4* zero4 = reinterpret_cast<float4*>(sh_smem + (size_t)(PBWD_THREADS / 32) * 32 * SH_BULK_STRIDE + 2 * (PBWD_THREADS / 32)) + warp * 12;
  if (BULK) {   // (launched only with has_sh)
    if (lane == 0) mbar_init(bar, 1);
    if (lane < 12) zero4[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    if (visible) bulk_load(my_row, p.shs + i * (size_t)(3 * p.M), (uint32_t)(12 * p.M), bar);
    if (lane == 0) mbar_expect_tx(bar, (uint32_t)(12 * p.M) * (uint32_t)__popc(vis_mask));
  }
  // the thread's own operands are requested BEFORE the cooperative row staging, so both streams are in flight together
  float4 g0 = make_float4(0, 0, 0, 0), g1 = g0, g2 = g0, g3 = g0, g4 = g0;
  f3 pos = mk(0, 0, 0);
  if (visible) {
    const float4* g = reinterpret_cast<const float4*>(p.grad_records + i * GREC);
    g0 = __ldg(g + 0); g1 = __ldg(g + 1); g2 = __ldg(g + 2); g3 = __ldg(g + 3); g4 = __ldg(g + 4);
    pos = mk(__ldg(p.means3D + 3 * i), __ldg(p.means3D + 3 * i + 1), __ldg(p.means3D + 3 * i + 2));
  }
  if (has_sh && !BULK) {
    if (!RAW || p.shs_rest == nullptr) {
      sh_rows_load(wsm, p.shs + (size_t)warp_first * p.M * 3, 3 * p.M, 3 * p.M + 1, 0, rows, vis_mask, lane);
    } else {
      sh_rows_load(wsm, p.shs + (size_t)warp_first * 3, 3, 3 * p.M + 1, 0, rows, vis_mask, lane);
      if (p.M > 1) sh_rows_load(wsm, p.shs_rest + (size_t)warp_first * (p.M - 1) * 3, 3 * p.M - 3, 3 * p.M + 1, 3, rows, vis_mask, lane);
    }
    __syncwarp();
  }

  f3 dmean3D = mk(0, 0, 0);
  float dmean2D_x = 0.f, dmean2D_y = 0.f, dopacity = 0.f;
  f3 dcolor = mk(0, 0, 0), drgb_masked = mk(0, 0, 0);
  float dscale_x = 0.f, dscale_y = 0.f;
  float4 drot = make_float4(0, 0, 0, 0);
  float dT[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};

  if (visible) {
    dT[0] = g0.x; dT[1] = g0.y; dT[2] = g0.z; dT[3] = g0.w; dT[4] = g1.x; dT[5] = g1.y;
    dT[6] = g1.z; dT[7] = g1.w; dT[8] = g2.x;
    const float gcx = g2.y, gcy = g2.z;
    dcolor = mk(g2.w, g3.x, g3.y);
    const f3 dnormal = mk(g3.z, g3.w, g4.x);
    dopacity = g4.y;
    const float raw_dTu_z = dT[2], raw_dTv_z = dT[5];

    const float* vm = p.viewmatrix;
    const float* pm = p.projmatrix;
    // transMats[idx*9+8] of the forward (Tw.z): it depends on the position and the projection matrix only, so it
    // is recomputed in the forward's association (preprocess.cu) instead of gathered from the 80-byte record
    const float depth = p.scales == nullptr
                            ? __ldg(p.transMat_precomp + 9 * i + 8)
                            : __fadd_rn(pm[15], __fmaf_rn(pos.z, pm[11], __fmaf_rn(pos.x, pm[3], __fmul_rn(pos.y, pm[7]))));

    const bool precomp = (p.scales == nullptr);
    f3 T0, T1, T2;            // rows Tu, Tv, Tw
    f3 R0, R1, R2;            // rotation columns
    f3 normal = mk(0, 0, 0);
    float Pm[3][4];           // P[k][r] = (world2ndc * ndc2pix)[k][r]
    float2 sc = make_float2(0, 0);
    float4 q = make_float4(0, 0, 0, 0);
    float raw_norm = 1.0f;
    if (precomp) {
      const float* T = p.transMat_precomp + 9 * i;
      T0 = mk(__ldg(T + 0), __ldg(T + 1), __ldg(T + 2));
      T1 = mk(__ldg(T + 3), __ldg(T + 4), __ldg(T + 5));
      T2 = mk(__ldg(T + 6), __ldg(T + 7), __ldg(T + 8));
    } else {
      sc = __ldg(reinterpret_cast<const float2*>(p.scales) + idx);
      q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
      if (RAW && p.raw) {  // the forward's activations (common.cuh), re-evaluated
        sc = activate_scaling(sc);
        const ActivatedRotation ar = activate_rotation(q);
        q = ar.q;
        raw_norm = ar.norm;
      }
      const float s = rsqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
      const float w = q.x * s, x = q.y * s, y = q.z * s, z = q.w * s;
      R0 = mk(1.f - 2.f * (y * y + z * z), 2.f * (x * y + w * z), 2.f * (x * z - w * y));
      R1 = mk(2.f * (x * y - w * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z + w * x));
      R2 = mk(2.f * (x * z + w * y), 2.f * (y * z - w * x), 1.f - 2.f * (x * x + y * y));
      const f3 L0 = sc.x * R0, L1 = sc.y * R1;  // scale modifier forced to 1 (backward.cu:487)
      const float hW = float(p.W) / 2.0f, hWm1 = float(p.W - 1) / 2.0f;
      const float hH = float(p.H) / 2.0f, hHm1 = float(p.H - 1) / 2.0f;
#pragma unroll
      for (int r = 0; r < 4; r++) {
        Pm[0][r] = pm[4 * r + 0] * hW + pm[4 * r + 3] * hWm1;
        Pm[1][r] = pm[4 * r + 1] * hH + pm[4 * r + 3] * hHm1;
        Pm[2][r] = pm[4 * r + 3];
      }
      // T[k][c] = <M[c], P[k]>, M[0]=(L0,0), M[1]=(L1,0), M[2]=(pos,1)
      float Tm[3][3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        Tm[k][0] = L0.x * Pm[k][0] + L0.y * Pm[k][1] + L0.z * Pm[k][2];
        Tm[k][1] = L1.x * Pm[k][0] + L1.y * Pm[k][1] + L1.z * Pm[k][2];
        Tm[k][2] = pos.x * Pm[k][0] + pos.y * Pm[k][1] + pos.z * Pm[k][2] + Pm[k][3];
      }
      T0 = mk(Tm[0][0], Tm[0][1], Tm[0][2]);
      T1 = mk(Tm[1][0], Tm[1][1], Tm[1][2]);
      T2 = mk(Tm[2][0], Tm[2][1], Tm[2][2]);
      normal = mk(vm[0] * R2.x + vm[4] * R2.y + vm[8] * R2.z, vm[1] * R2.x + vm[5] * R2.y + vm[9] * R2.z,
                  vm[2] * R2.x + vm[6] * R2.y + vm[10] * R2.z);
    }

    f3 dT0 = mk(dT[0], dT[1], dT[2]), dT1 = mk(dT[3], dT[4], dT[5]), dT2 = mk(dT[6], dT[7], dT[8]);
    bool chained = false;
    if (gcx != 0.f || gcy != 0.f) {  // AABB-centre chain (backward.cu:521-535)
      const f3 tv = mk(9.0f, 9.0f, -1.0f);
      const float d = dot(tv, T2 * T2);
      const f3 fv = (1.0f / d) * tv;
      const f3 a0 = gcx * (fv * T2);
      const f3 a1 = gcy * (fv * T2);
      f3 a3 = gcx * (fv * T0) + gcy * (fv * T1);
      const f3 dL_df = gcx * (T0 * T2) + gcy * (T1 * T2);
      const float dL_dd = dot(dL_df, fv) * (-1.0f / d);
      const f3 dd_dT3 = 2.0f * (tv * T2);
      a3 = a3 + dL_dd * dd_dT3;
      dT0 = dT0 + a0; dT1 = dT1 + a1; dT2 = dT2 + a3;
      chained = true;
    }
    if (precomp) {
      if (chained) {
        dT[0] = dT0.x; dT[1] = dT0.y; dT[2] = dT0.z; dT[3] = dT1.x; dT[4] = dT1.y; dT[5] = dT1.z;
        dT[6] = dT2.x; dT[7] = dT2.y; dT[8] = dT2.z;
      }
    } else {
      // dL_dM[c] = sum_k P[k] * dL_dT[k][c]   (backward.cu:554)
      const float dTk[3][3] = {{dT0.x, dT0.y, dT0.z}, {dT1.x, dT1.y, dT1.z}, {dT2.x, dT2.y, dT2.z}};
      f3 dM[3];
#pragma unroll
      for (int c = 0; c < 3; c++) {
        dM[c].x = Pm[0][0] * dTk[0][c] + Pm[1][0] * dTk[1][c] + Pm[2][0] * dTk[2][c];
        dM[c].y = Pm[0][1] * dTk[0][c] + Pm[1][1] * dTk[1][c] + Pm[2][1] * dTk[2][c];
        dM[c].z = Pm[0][2] * dTk[0][c] + Pm[1][2] * dTk[1][c] + Pm[2][2] * dTk[2][c];
      }
      // dL_dtn = V[:3,:3] * dL_dnormal (auxiliary.h:110-118), dual-visible sign (backward.cu:556-561)
      f3 dtn = mk(vm[0] * dnormal.x + vm[1] * dnormal.y + vm[2] * dnormal.z,
                  vm[4] * dnormal.x + vm[5] * dnormal.y + vm[6] * dnormal.z,
                  vm[8] * dnormal.x + vm[9] * dnormal.y + vm[10] * dnormal.z);
      const f3 pv = mk(vm[0] * pos.x + vm[4] * pos.y + vm[8] * pos.z + vm[12],
                       vm[1] * pos.x + vm[5] * pos.y + vm[9] * pos.z + vm[13],
                       vm[2] * pos.x + vm[6] * pos.y + vm[10] * pos.z + vm[14]);
      const float cosv = -(pv.x * normal.x + pv.y * normal.y + pv.z * normal.z);
      const float mult = cosv > 0 ? 1.f : -1.f;
      dtn = mult * dtn;
      const f3 dRS0 = dM[0], dRS1 = dM[1], dRS2 = dtn;
      // v_R columns
      const f3 vR0 = sc.x * dRS0, vR1 = sc.y * dRS1, vR2 = dRS2;
      // quaternion VJP with the normaliser treated as constant (auxiliary.h:238-282)
      const float s = rsqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
      const float w = q.x * s, x = q.y * s, y = q.z * s, z = q.w * s;
      // v_R[c][r]: column c, row r
      drot.x = 2.f * (x * (vR1.z - vR2.y) + y * (vR2.x - vR0.z) + z * (vR0.y - vR1.x));
      drot.y = 2.f * (-2.f * x * (vR1.y + vR2.z) + y * (vR0.y + vR1.x) + z * (vR0.z + vR2.x) + w * (vR1.z - vR2.y));
      drot.z = 2.f * (x * (vR0.y + vR1.x) - 2.f * y * (vR0.x + vR2.z) + z * (vR1.z + vR2.y) + w * (vR2.x - vR0.z));
      drot.w = 2.f * (x * (vR0.z + vR2.x) + y * (vR1.z + vR2.y) - 2.f * z * (vR0.x + vR1.y) + w * (vR0.y - vR1.x));
      dscale_x = dot(dRS0, R0);
      dscale_y = dot(dRS1, R1);
      dmean3D = dM[2];
      if (RAW && p.raw) {
        // chain through the activations: exp' = exp; normalize: (g - q (q.g)) / max(||raw||, eps) (the clamp's own
        // derivative is zero); sigmoid' = o (1 - o)
        dscale_x *= sc.x; dscale_y *= sc.y;
        const float qg = q.x * drot.x + q.y * drot.y + q.z * drot.z + q.w * drot.w;
        const float inv = 1.0f / raw_norm;
        if (raw_norm > 1e-12f) drot = make_float4((drot.x - q.x * qg) * inv, (drot.y - q.y * qg) * inv, (drot.z - q.z * qg) * inv, (drot.w - q.w * qg) * inv);
        else drot = make_float4(drot.x * inv, drot.y * inv, drot.z * inv, drot.w * inv);
      }
    }
    if (RAW && p.raw) {
      const float o = activate_opacity(__ldg(p.opacities + i));
      dopacity *= o * (1.0f - o);
    }

    if (has_sh) {
      const int clamp_bits = (int)__ldg(p.clamp + i);
      f3 dRGB = dcolor;
      if (clamp_bits & 1) dRGB.x = 0.f;
      if (clamp_bits & 2) dRGB.y = 0.f;
      if (clamp_bits & 4) dRGB.z = 0.f;
      const f3 dir_orig = mk(pos.x - p.campos[0], pos.y - p.campos[1], pos.z - p.campos[2]);
      f3 add;
      if (BULK) {
        mbar_wait(bar, 0);
        sh_backward_row4(p.D, p.M, reinterpret_cast<float4*>(my_row), dir_orig, dRGB, add);
      } else {
        sh_backward(p.D, p.M, my_row, dir_orig, dRGB, add);
      }
      dmean3D = dmean3D + add;
      drgb_masked = dRGB;
    }

    // densification heuristic (backward.cu:632-635); in the precomputed-T case the reference reads
    // the (possibly rewritten) dL_dtransMat, otherwise the raw accumulation.
    const float a2 = precomp ? dT[2] : raw_dTu_z;
    const float a5 = precomp ? dT[5] : raw_dTv_z;
    dmean2D_x = a2 * depth * 0.5f * float(p.W);
    dmean2D_y = a5 * depth * 0.5f * float(p.H);
  }

  if (has_sh && factored && p.n_rgb_peers > 0) {
    // the all-gather of the factored exchange, fused: the warp's 32 rows (384 contiguous bytes) go to EVERY rank's
    // receive slot as 24 coalesced 128-bit stores over NVLink (the SH rows in shared memory are spent by now)
    __syncwarp();
    wsm[3 * lane] = drgb_masked.x; wsm[3 * lane + 1] = drgb_masked.y; wsm[3 * lane + 2] = drgb_masked.z;
    __syncwarp();
    if (rows == 32) {
      if (lane < 24) {
        const float4 v = reinterpret_cast<const float4*>(wsm)[lane];
        for (int k = 0; k < p.n_rgb_peers; k++) reinterpret_cast<float4*>(p.rgb_peers[k] + 3 * (size_t)warp_first)[lane] = v;
      }
    } else {
      for (int e = lane; e < 3 * rows; e += 32) {
        const float v = wsm[e];
        for (int k = 0; k < p.n_rgb_peers; k++) p.rgb_peers[k][3 * (size_t)warp_first + e] = v;
      }
    }
  } else if (has_sh && factored) {
    if (valid) store3(p.dL_drgb + 3 * i, drgb_masked);  // per view: never accumulated
  } else if (has_sh && BULK) {
    // each lane hands its own row (dL_dsh, or the zero row for a culled surfel) to the copy engine
    fence_proxy_async();
    __syncwarp();
    if (valid) {
      float* dst = p.dL_dsh + i * (size_t)(3 * p.M);
      if (visible) {
        if (acc) bulk_reduce_add_f32(dst, my_row, (uint32_t)(12 * p.M));
        else bulk_store(dst, my_row, (uint32_t)(12 * p.M));
      } else if (!acc) {
        bulk_store(dst, zero4, (uint32_t)(12 * p.M));
      }
    }
    bulk_commit_and_wait_read();
  } else if (has_sh) {
    __syncwarp();
    if (!RAW || p.dL_dsh_rest == nullptr) {
      sh_rows_store(wsm, p.dL_dsh + (size_t)warp_first * p.M * 3, 3 * p.M, 3 * p.M + 1, 0, rows, vis_mask, acc, lane);
    } else {  // gradients of features_dc and features_rest, no concatenated gradient to slice afterwards
      sh_rows_store(wsm, p.dL_dsh + (size_t)warp_first * 3, 3, 3 * p.M + 1, 0, rows, vis_mask, acc, lane);
      if (p.M > 1) sh_rows_store(wsm, p.dL_dsh_rest + (size_t)warp_first * (p.M - 1) * 3, 3 * p.M - 3, 3 * p.M + 1, 3, rows, vis_mask, acc, lane);
    }
  } else if (p.dL_dsh != nullptr && valid && !acc) {
    float* o = p.dL_dsh + i * p.M * 3;
    for (int k = 0; k < p.M * 3; k++) o[k] = 0.f;
  }
  if (!valid) return;
  // the view-space gradient is per view (gm.py:551-553 takes its norm per view): never accumulated
  p.dL_dmeans2D[3 * i] = dmean2D_x; p.dL_dmeans2D[3 * i + 1] = dmean2D_y; p.dL_dmeans2D[3 * i + 2] = 0.f;
  if (visible) {  // densification statistics of the view (gaussian_model.py:551-553), fused
    if (p.stat_norm) p.stat_norm[i] += sqrtf(dmean2D_x * dmean2D_x + dmean2D_y * dmean2D_y);
    if (p.stat_vis) p.stat_vis[i] += 1.0f;
    if (p.max_radii) p.max_radii[i] = max(p.max_radii[i], __ldg(p.radii + idx));
  }
  if (acc) {
    if (!visible) return;  // adds nothing
    p.dL_dopacity[i] += dopacity;
    store3(p.dL_dmeans3D + 3 * i, dmean3D, true);
    if (p.dL_dcolors) store3(p.dL_dcolors + 3 * i, dcolor, true);
    if (p.dL_dtransMat) {
#pragma unroll
      for (int k = 0; k < 9; k++) p.dL_dtransMat[9 * i + k] += dT[k];
    }
    if (p.dL_dscales) { p.dL_dscales[2 * i] += dscale_x; p.dL_dscales[2 * i + 1] += dscale_y; }
    if (p.dL_drotations) {
      float4* d = reinterpret_cast<float4*>(p.dL_drotations) + i;
      float4 o = *d;
      o.x += drot.x; o.y += drot.y; o.z += drot.z; o.w += drot.w;
      *d = o;
    }
    return;
  }
  p.dL_dopacity[i] = dopacity;
  store3(p.dL_dmeans3D + 3 * i, dmean3D);
  if (p.dL_dcolors) store3(p.dL_dcolors + 3 * i, dcolor);
  if (p.dL_dtransMat) {
#pragma unroll
    for (int k = 0; k < 9; k++) p.dL_dtransMat[9 * i + k] = dT[k];
  }
  if (p.dL_dscales) { p.dL_dscales[2 * i] = dscale_x; p.dL_dscales[2 * i + 1] = dscale_y; }
  if (p.dL_drotations) reinterpret_cast<float4*>(p.dL_drotations)[i] = drot;
}

// View-parallel exchange, receiving side: dL_dsh[p] (=|+=) sum over views v of basis(dir_v(p)) (x) dL_dRGB_v[p].
// dL_dsh of ONE view is a rank-1 product of a view-dependent basis (a function of the surfel position and
// the view's camera centre, both known on every rank) and 3 floats, so ranks exchange the 3 floats
// (12 B/surfel/view) instead of the 12M-byte row (192 B at M = 16).  One thread per surfel; the rows
// leave through shared memory as coalesced 128-bit stores (sh_rows_store).
// BULK (M % 4 == 0, aligned rows): every lane parks its 48 sums in its own 16-byte aligned shared-memory row with 12
// 128-bit stores and hands the row to the copy engine (cp.async.bulk / cp.reduce.async.bulk, one instruction per row;
// rows without a gradient are written from a shared zero row) -- instead of 48 scalar stores and a cooperative copy
// loop with multiply-high row/column indexing (~300 instructions per lane at M = 16).
template <bool BULK>
__global__ void __launch_bounds__(PBWD_THREADS)
sh_grad_expand_kernel(int P, int D, int M, int n_views, const float* __restrict__ means3D,
                      const float* __restrict__ campos_all, const float* __restrict__ drgb_all, size_t view_stride,
                      float* __restrict__ dL_dsh, int accumulate) {
  extern __shared__ float sh_smem[];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warp_first = idx - lane;
  if (warp_first >= P) return;
  const bool valid = idx < P;
  const size_t i = (size_t)(valid ? idx : P - 1);
  const int rows = min(32, P - warp_first);
  float* wsm = sh_smem + (size_t)warp * 32 * (3 * M + 1);
  float* my_row = wsm + lane * (3 * M + 1);
  const f3 pos = mk(__ldg(means3D + 3 * i), __ldg(means3D + 3 * i + 1), __ldg(means3D + 3 * i + 2));
  const int K = min(M, (D + 1) * (D + 1));
  float acc[48];
#pragma unroll
  for (int k = 0; k < 48; k++) acc[k] = 0.f;
  bool any = false;
  // the rows of up to 8 views are requested together: one dependent HBM round trip per view (round 1) left the kernel
  // waiting on memory 8 times in a row at 8 GPUs -- 0.70 ms for 1.5 GB
  constexpr int VB = 8;
  for (int v0 = 0; v0 < n_views; v0 += VB) {
    f3 d[VB];
#pragma unroll
    for (int u = 0; u < VB; u++) {
      d[u] = mk(0.f, 0.f, 0.f);
      if (valid && v0 + u < n_views) {
        const float* g = drgb_all + (size_t)(v0 + u) * view_stride + i * 3;
        d[u] = mk(__ldg(g), __ldg(g + 1), __ldg(g + 2));
      }
    }
#pragma unroll
    for (int u = 0; u < VB; u++) {
      if (d[u].x == 0.f && d[u].y == 0.f && d[u].z == 0.f) continue;  // culled in this view (or no gradient)
      any = true;
      const int v = v0 + u;
      const f3 dir = mk(pos.x - __ldg(campos_all + 3 * v), pos.y - __ldg(campos_all + 3 * v + 1),
                        pos.z - __ldg(campos_all + 3 * v + 2));
      const float len = __fsqrt_rn(__fmaf_rn(dir.z, dir.z, __fmaf_rn(dir.x, dir.x, __fmul_rn(dir.y, dir.y))));
      float coef[16];
      sh_coefficients(D, __fdiv_rn(dir.x, len), __fdiv_rn(dir.y, len), __fdiv_rn(dir.z, len), coef);
#pragma unroll
      for (int k = 0; k < 16; k++) {
        acc[3 * k] += coef[k] * d[u].x; acc[3 * k + 1] += coef[k] * d[u].y; acc[3 * k + 2] += coef[k] * d[u].z;
      }
    }
  }
  if (BULK) {
    float* rows_base = sh_smem + (size_t)warp * 32 * SH_BULK_STRIDE;
    float4* row4 = reinterpret_cast<float4*>(rows_base + lane * SH_BULK_STRIDE);
    float4* zero4 = reinterpret_cast<float4*>(sh_smem + (size_t)(PBWD_THREADS / 32) * 32 * SH_BULK_STRIDE) + warp * 12;
    if (!accumulate && lane < 12) zero4[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int n4 = 3 * M / 4;
    if (any) {
#pragma unroll
      for (int k = 0; k < 12; k++)   // (sums beyond the active degree are exact zeros: their coefficients are)
        if (k < n4) row4[k] = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
    }
    fence_proxy_async();   // (writers fence, then the warp meets, then rows are handed over: the zero row is shared)
    __syncwarp();
    if (valid) {
      float* dst = dL_dsh + i * (size_t)(3 * M);
      if (any) {
        if (accumulate) bulk_reduce_add_f32(dst, row4, (uint32_t)(12 * M));
        else bulk_store(dst, row4, (uint32_t)(12 * M));
      } else if (!accumulate) {
        bulk_store(dst, zero4, (uint32_t)(12 * M));
      }
    }
    bulk_commit_and_wait_read();
    return;
  }
  const unsigned vis_mask = __ballot_sync(0xffffffffu, any);
  if (any) {
#pragma unroll
    for (int k = 0; k < 16; k++) {
      if (k < M) {
        const bool on = k < K;
        my_row[3 * k] = on ? acc[3 * k] : 0.f; my_row[3 * k + 1] = on ? acc[3 * k + 1] : 0.f;
        my_row[3 * k + 2] = on ? acc[3 * k + 2] : 0.f;
      }
    }
    for (int k = 16; k < M; k++) { my_row[3 * k] = 0.f; my_row[3 * k + 1] = 0.f; my_row[3 * k + 2] = 0.f; }
  }
  __syncwarp();
  sh_rows_store(wsm, dL_dsh + (size_t)warp_first * M * 3, 3 * M, 3 * M + 1, 0, rows, vis_mask, accumulate != 0, lane);
}

void launch_sh_grad_expand(int P, int D, int M, int n_views, const float* means3D, const float* campos_all,
                           const float* drgb_all, size_t view_stride, float* dL_dsh, int accumulate, cudaStream_t stream) {
  if (P <= 0 || M <= 0) return;
  const size_t smem = (size_t)(PBWD_THREADS / 32) * 32 * (3 * M + 1) * sizeof(float);
  static PerDeviceLevel configured;
  if (smem > 48 * 1024) configured.ensure(smem, [&] {
    cudaFuncSetAttribute(sh_grad_expand_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  });
  const bool bulk = M % 4 == 0 && M <= 16 && (reinterpret_cast<size_t>(dL_dsh) & 15) == 0 && getenv("GSM_NO_BULK") == nullptr;
  if (bulk) {
    const size_t bsmem = (size_t)(PBWD_THREADS / 32) * 32 * SH_BULK_STRIDE * sizeof(float) + (PBWD_THREADS / 32) * 192;
    sh_grad_expand_kernel<true><<<(P + PBWD_THREADS - 1) / PBWD_THREADS, PBWD_THREADS, bsmem, stream>>>(
        P, D, M, n_views, means3D, campos_all, drgb_all, view_stride, dL_dsh, accumulate);
    return;
  }
  sh_grad_expand_kernel<false><<<(P + PBWD_THREADS - 1) / PBWD_THREADS, PBWD_THREADS, smem, stream>>>(
      P, D, M, n_views, means3D, campos_all, drgb_all, view_stride, dL_dsh, accumulate);
}


void launch_preprocess_bwd(const BwdPreprocessParams& p, cudaStream_t stream) {
  if (p.P <= 0) return;
  const bool raw = p.raw || p.shs_rest != nullptr;
  const bool factored = p.dL_drgb != nullptr || p.n_rgb_peers > 0;
  const bool has_sh = p.shs != nullptr && (p.dL_dsh != nullptr || factored);
  // per-row TMA bulk copies: the concatenated [P,M,3] tensors with 16-byte rows at 16-byte addresses
  const bool bulk = !raw && has_sh && p.M > 0 && p.M <= 16 && p.M % 4 == 0 && (reinterpret_cast<size_t>(p.shs) & 15) == 0 &&
                    (p.dL_dsh == nullptr || (reinterpret_cast<size_t>(p.dL_dsh) & 15) == 0) && getenv("GSM_NO_BULK") == nullptr;
  const size_t smem = bulk ? (size_t)(PBWD_THREADS / 32) * 32 * SH_BULK_STRIDE * sizeof(float) + (PBWD_THREADS / 32) * (8 + 192)
                           : (size_t)(PBWD_THREADS / 32) * 32 * (3 * p.M + 1) * sizeof(float);
  static PerDeviceLevel configured;
  if (smem > 48 * 1024) configured.ensure(smem, [&] {
    cudaFuncSetAttribute(preprocess_bwd_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(preprocess_bwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  });
  const int grid = (p.P + PBWD_THREADS - 1) / PBWD_THREADS;
  if (bulk) preprocess_bwd_kernel<false, true><<<grid, PBWD_THREADS, smem, stream>>>(p);
  else if (raw) preprocess_bwd_kernel<true, false><<<grid, PBWD_THREADS, smem, stream>>>(p);
  else preprocess_bwd_kernel<false, false><<<grid, PBWD_THREADS, smem, stream>>>(p);
}

}  // namespace gsm

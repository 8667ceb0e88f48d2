// Forward per-surfel preprocess (K1) and markVisible (K7) for sm_100a.
//
// Restates, with its own data layout, what the reference does in
//   cuda_rasterizer/forward.cu:20-71   (SH -> RGB, clamp flags)
//   cuda_rasterizer/forward.cu:75-115  (splat->pixel homography T, view-space normal)
//   cuda_rasterizer/forward.cu:119-145 (AABB centre / extent)
//   cuda_rasterizer/forward.cu:175-251 (cull, radius, tile rect)
//   cuda_rasterizer/auxiliary.h:67-77,185-235 (rect, frustum test, quaternion)
//
// radii, the tile rect and the depth key must be BIT-EXACT against the reference build, so every
// expression feeding them is written with explicit round-to-nearest intrinsics in the operation
// order the reference's sm_100a build performs (documented in DESIGN.md, "FP association").
// Output is one packed 80-byte record per visible surfel instead of the reference's six SoA arrays.
#include "common.cuh"
#include <cstdlib>

namespace gsm {

#define MUL(a, b) __fmul_rn((a), (b))
#define ADD(a, b) __fadd_rn((a), (b))
#define SUB(a, b) __fsub_rn((a), (b))
#define FMA(a, b, c) __fmaf_rn((a), (b), (c))

// SH basis constants as fp32 (auxiliary.h:43-60)
__device__ __forceinline__ float3 sh_to_rgb(int D, const float* sh, float3 dir) {
  // sh: this surfel's 3M coefficients, staged in shared memory by the warp (coalesced 128-bit loads)
  const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
  float3 r;
  r.x = MUL(C0, sh[0]); r.y = MUL(C0, sh[1]); r.z = MUL(C0, sh[2]);
  if (D > 0) {
    const float x = dir.x, y = dir.y, z = dir.z;
    float t = MUL(y, C1);
    r.x = FMA(-t, sh[3], r.x); r.y = FMA(-t, sh[4], r.y); r.z = FMA(-t, sh[5], r.z);
    t = MUL(z, C1);
    r.x = FMA(t, sh[6], r.x); r.y = FMA(t, sh[7], r.y); r.z = FMA(t, sh[8], r.z);
    t = MUL(x, C1);
    r.x = FMA(-t, sh[9], r.x); r.y = FMA(-t, sh[10], r.y); r.z = FMA(-t, sh[11], r.z);
    if (D > 1) {
      const float xx = MUL(x, x), yy = MUL(y, y), zz = MUL(z, z);
      const float xy = MUL(x, y), yz = MUL(y, z), xz = MUL(x, z);
      const float C20 = 1.0925484305920792f, C21 = -1.0925484305920792f, C22 = 0.31539156525252005f,
                  C23 = -1.0925484305920792f, C24 = 0.5462742152960396f;
      t = MUL(xy, C20);
      r.x = FMA(t, sh[12], r.x); r.y = FMA(t, sh[13], r.y); r.z = FMA(t, sh[14], r.z);
      t = MUL(yz, C21);
      r.x = FMA(t, sh[15], r.x); r.y = FMA(t, sh[16], r.y); r.z = FMA(t, sh[17], r.z);
      const float zz2 = ADD(zz, zz);
      t = MUL(SUB(SUB(zz2, xx), yy), C22);
      r.x = FMA(t, sh[18], r.x); r.y = FMA(t, sh[19], r.y); r.z = FMA(t, sh[20], r.z);
      t = MUL(xz, C23);
      r.x = FMA(t, sh[21], r.x); r.y = FMA(t, sh[22], r.y); r.z = FMA(t, sh[23], r.z);
      const float xx_yy = SUB(xx, yy);
      t = MUL(xx_yy, C24);
      r.x = FMA(t, sh[24], r.x); r.y = FMA(t, sh[25], r.y); r.z = FMA(t, sh[26], r.z);
      if (D > 2) {
        const float C30 = -0.5900435899266435f, C31 = 2.890611442640554f, C32 = -0.4570457994644658f,
                    C33 = 0.3731763325901154f, C34 = -0.4570457994644658f, C35 = 1.445305721320277f,
                    C36 = -0.5900435899266435f;
                const float zz4_xx_yy = SUB(FMA(zz, 4.0f, -xx), yy);
        t = MUL(MUL(y, C30), FMA(xx, 3.0f, -yy));
        r.x = FMA(t, sh[27], r.x); r.y = FMA(t, sh[28], r.y); r.z = FMA(t, sh[29], r.z);
        t = MUL(z, MUL(xy, C31));
        r.x = FMA(t, sh[30], r.x); r.y = FMA(t, sh[31], r.y); r.z = FMA(t, sh[32], r.z);
        t = MUL(MUL(y, C32), zz4_xx_yy);
        r.x = FMA(t, sh[33], r.x); r.y = FMA(t, sh[34], r.y); r.z = FMA(t, sh[35], r.z);
        t = MUL(MUL(z, C33), FMA(yy, -3.0f, FMA(xx, -3.0f, zz2)));
        r.x = FMA(t, sh[36], r.x); r.y = FMA(t, sh[37], r.y); r.z = FMA(t, sh[38], r.z);
        t = MUL(MUL(x, C34), zz4_xx_yy);
        r.x = FMA(t, sh[39], r.x); r.y = FMA(t, sh[40], r.y); r.z = FMA(t, sh[41], r.z);
        t = MUL(MUL(z, C35), xx_yy);
        r.x = FMA(t, sh[42], r.x); r.y = FMA(t, sh[43], r.y); r.z = FMA(t, sh[44], r.z);
        t = MUL(MUL(x, C36), FMA(yy, -3.0f, xx));
        r.x = FMA(t, sh[45], r.x); r.y = FMA(t, sh[46], r.y); r.z = FMA(t, sh[47], r.z);
      }
    }
  }
  r.x = ADD(r.x, 0.5f); r.y = ADD(r.y, 0.5f); r.z = ADD(r.z, 0.5f);
  return r;
}

// Quaternion (w,x,y,z) -> rotation columns, scaled: L0 = sx*R[:,0], L1 = sy*R[:,1], L2 = R[:,2]
// (auxiliary.h:213-235, forward.cu:88-90).  rsqrtf is the approximate PTX rsqrt, as the reference.
__device__ __forceinline__ void build_L(float4 q, float sx, float sy, float3& L0, float3& L1, float3& L2) {
  float n2 = MUL(q.w, q.w);
  n2 = FMA(q.x, q.x, n2);
  n2 = FMA(q.y, q.y, n2);
  n2 = FMA(q.z, q.z, n2);
  const float s = rsqrtf(n2);
  const float w = MUL(q.x, s), x = MUL(q.y, s), y = MUL(q.z, s), z = MUL(q.w, s);
  const float yy = MUL(y, y), zz = MUL(z, z);
  const float wz = MUL(w, z), wy = MUL(w, y), wx = MUL(w, x);
  const float xx_zz = FMA(x, x, zz), xx_yy = FMA(x, x, yy), yy_zz = ADD(yy, zz);
  // ptxas fuses the products of the off-diagonal terms into the add/sub (reference SASS)
  const float a = FMA(x, y, wz), b = FMA(x, y, -wz), c = FMA(x, z, -wy), d = FMA(x, z, wy), e = FMA(y, z, wx), f = FMA(y, z, -wx);
  const float R00 = SUB(1.0f, ADD(yy_zz, yy_zz)), R01 = ADD(a, a), R02 = ADD(c, c);
  const float R10 = ADD(b, b), R11 = SUB(1.0f, ADD(xx_zz, xx_zz)), R12 = ADD(e, e);
  const float R20 = ADD(d, d), R21 = ADD(f, f), R22 = SUB(1.0f, ADD(xx_yy, xx_yy));
  L0 = make_float3(MUL(sx, R00), MUL(sx, R01), MUL(sx, R02));
  L1 = make_float3(MUL(sy, R10), MUL(sy, R11), MUL(sy, R12));
  L2 = make_float3(R20, R21, R22);
}

__device__ __forceinline__ float dot3_ref(float a0, float a1, float a2, float b0, float b1, float b2) {
  // reference association of a0*b0 + a1*b1 + a2*b2:  fma(a2,b2, fma(a0,b0, a1*b1))
  return FMA(a2, b2, FMA(a0, b0, MUL(a1, b1)));
}

constexpr int PFWD_THREADS = 128;
#ifndef GSM_PFWD_MIN_CTAS
#define GSM_PFWD_MIN_CTAS 8
#endif

// ---- SH rows by TMA bulk copy (M a multiple of 4: 12 M bytes is a multiple of 16) ---------------------------------------
// Every VISIBLE lane fetches its own row with one cp.async.bulk (192 B at M = 16) that completes on the warp's mbarrier,
// issued as soon as the cull decision is known: the copy runs under the rest of the geometry (contribution box, slabs),
// and replaces the cooperative staging loop (per float4: one LDG.128, multiply-high row/column indexing, four STS.32 --
// ~300 instructions per lane at M = 16) by one instruction per row.  Rows sit 208 B apart in shared memory: 16-byte
// aligned for the copy, and conflict-free for the 128-bit reads of the evaluation (52 floats = 13 x 16 B, odd).

// sh_to_rgb over a 16-byte aligned row read with 128-bit loads: same arithmetic, same order as sh_to_rgb above; the row
// is fetched degree by degree (3 / 4 / 5 float4s live at a time) so that the evaluation stays inside 64 registers
__device__ __forceinline__ float3 sh_to_rgb_row4(int D, const float4* row4, float3 dir) {
  const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
  float3 r;
  const float x = dir.x, y = dir.y, z = dir.z;
  {
    const float4 v0 = row4[0];
    r.x = MUL(C0, v0.x); r.y = MUL(C0, v0.y); r.z = MUL(C0, v0.z);
    if (D > 0) {
      const float4 v1 = row4[1], v2 = row4[2];
      const float a[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};   // floats 0..11
      float t = MUL(y, C1);
      r.x = FMA(-t, a[3], r.x); r.y = FMA(-t, a[4], r.y); r.z = FMA(-t, a[5], r.z);
      t = MUL(z, C1);
      r.x = FMA(t, a[6], r.x); r.y = FMA(t, a[7], r.y); r.z = FMA(t, a[8], r.z);
      t = MUL(x, C1);
      r.x = FMA(-t, a[9], r.x); r.y = FMA(-t, a[10], r.y); r.z = FMA(-t, a[11], r.z);
    }
  }
  if (D > 1) {
    const float xx = MUL(x, x), yy = MUL(y, y), zz = MUL(z, z);
    const float xy = MUL(x, y), yz = MUL(y, z), xz = MUL(x, z);
    const float zz2 = ADD(zz, zz);
    const float xx_yy = SUB(xx, yy);
    {
      const float4 v3 = row4[3], v4 = row4[4], v5 = row4[5], v6 = row4[6];
      const float b[16] = {v3.x, v3.y, v3.z, v3.w, v4.x, v4.y, v4.z, v4.w, v5.x, v5.y, v5.z, v5.w, v6.x, v6.y, v6.z, v6.w};  // floats 12..27
      const float C20 = 1.0925484305920792f, C21 = -1.0925484305920792f, C22 = 0.31539156525252005f,
                  C23 = -1.0925484305920792f, C24 = 0.5462742152960396f;
      float t = MUL(xy, C20);
      r.x = FMA(t, b[0], r.x); r.y = FMA(t, b[1], r.y); r.z = FMA(t, b[2], r.z);
      t = MUL(yz, C21);
      r.x = FMA(t, b[3], r.x); r.y = FMA(t, b[4], r.y); r.z = FMA(t, b[5], r.z);
      t = MUL(SUB(SUB(zz2, xx), yy), C22);
      r.x = FMA(t, b[6], r.x); r.y = FMA(t, b[7], r.y); r.z = FMA(t, b[8], r.z);
      t = MUL(xz, C23);
      r.x = FMA(t, b[9], r.x); r.y = FMA(t, b[10], r.y); r.z = FMA(t, b[11], r.z);
      t = MUL(xx_yy, C24);
      r.x = FMA(t, b[12], r.x); r.y = FMA(t, b[13], r.y); r.z = FMA(t, b[14], r.z);
    }
    if (D > 2) {
      const float4 v6 = row4[6], v7 = row4[7], v8 = row4[8], v9 = row4[9], v10 = row4[10], v11 = row4[11];
      // floats 27..47: e[i] = float 27 + i
      const float e[21] = {v6.w, v7.x, v7.y, v7.z, v7.w, v8.x, v8.y, v8.z, v8.w, v9.x, v9.y, v9.z, v9.w,
                           v10.x, v10.y, v10.z, v10.w, v11.x, v11.y, v11.z, v11.w};
      const float C30 = -0.5900435899266435f, C31 = 2.890611442640554f, C32 = -0.4570457994644658f,
                  C33 = 0.3731763325901154f, C34 = -0.4570457994644658f, C35 = 1.445305721320277f,
                  C36 = -0.5900435899266435f;
      const float zz4_xx_yy = SUB(FMA(zz, 4.0f, -xx), yy);
      float t = MUL(MUL(y, C30), FMA(xx, 3.0f, -yy));
      r.x = FMA(t, e[0], r.x); r.y = FMA(t, e[1], r.y); r.z = FMA(t, e[2], r.z);
      t = MUL(z, MUL(xy, C31));
      r.x = FMA(t, e[3], r.x); r.y = FMA(t, e[4], r.y); r.z = FMA(t, e[5], r.z);
      t = MUL(MUL(y, C32), zz4_xx_yy);
      r.x = FMA(t, e[6], r.x); r.y = FMA(t, e[7], r.y); r.z = FMA(t, e[8], r.z);
      t = MUL(MUL(z, C33), FMA(yy, -3.0f, FMA(xx, -3.0f, zz2)));
      r.x = FMA(t, e[9], r.x); r.y = FMA(t, e[10], r.y); r.z = FMA(t, e[11], r.z);
      t = MUL(MUL(x, C34), zz4_xx_yy);
      r.x = FMA(t, e[12], r.x); r.y = FMA(t, e[13], r.y); r.z = FMA(t, e[14], r.z);
      t = MUL(MUL(z, C35), xx_yy);
      r.x = FMA(t, e[15], r.x); r.y = FMA(t, e[16], r.y); r.z = FMA(t, e[17], r.z);
      t = MUL(MUL(x, C36), FMA(yy, -3.0f, xx));
      r.x = FMA(t, e[18], r.x); r.y = FMA(t, e[19], r.y); r.z = FMA(t, e[20], r.z);
    }
  }
  r.x = ADD(r.x, 0.5f); r.y = ADD(r.y, 0.5f); r.z = ADD(r.z, 0.5f);
  return r;
}

// Cooperative, coalesced load of the SH rows of the warp's visible surfels into shared memory
// (row r at wsm[r * (3M + 1)]: odd stride, conflict-free per-lane reads).
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
          const int i = e0 + e, r = (int)__umulhi((unsigned)i, rmagic);
          wsm[r * rs + col0 + (i - r * rl)] = vv[e];
        }
      }
    }
  } else {
    for (int i = lane; i < total; i += 32) {
      const int r = (int)__umulhi((unsigned)i, rmagic);
      if ((vis_mask >> r) & 1u) wsm[r * rs + col0 + (i - r * rl)] = __ldg(g + i);
    }
  }
}

// upper bound of sqrt(x), x >= 0, through the fp32 unit (the fp64 square root is a long software sequence;
// the extents it feeds only need to err upwards)
__device__ __forceinline__ double sqrt_up(double x) { return (double)__fsqrt_ru(__double2float_ru(x)); }

// RAW: the raw-parameter entry (in-kernel activations and / or split SH storage); the standard entry compiles without it
// BULK: the SH rows arrive by per-row TMA bulk copies (standard entry, M a multiple of 4; see above)
template <bool RAW, bool BULK>
__global__ void __launch_bounds__(PFWD_THREADS, GSM_PFWD_MIN_CTAS) preprocess_fwd_kernel(const FwdPreprocessParams p) {
  extern __shared__ __align__(16) float sh_smem[];
  const int gidx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int warp_first = gidx - lane;
  if (warp_first >= p.P) return;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sh_smem + (size_t)(PFWD_THREADS / 32) * 32 * SH_BULK_STRIDE) + warp;
  if (BULK) {
    if (lane == 0) mbar_init(bar, 1);
    __syncwarp();
  }
  const bool valid = gidx < p.P;
  const int idx = valid ? gidx : p.P - 1;  // out-of-range lanes shadow the last surfel and write nothing

  int radius_out = 0;
  uint32_t tiles_out = 0;
  float Tu0 = 0, Tu1 = 0, Tu2 = 0, Tv0 = 0, Tv1 = 0, Tv2 = 0, Tw0 = 0, Tw1 = 0, Tw2 = 0;
  float nx = 0, ny = 0, nz = 0, cx = 0, cy = 0, opac = 0;
  uint32_t rect_min = 0, rect_max = 0;
  short4 box = make_short4(-32768, -32768, 32767, 32767);
  float4 diag = make_float4(-INFINITY, INFINITY, -INFINITY, INFINITY);  // unbounded diagonal slabs

  // camera matrices are 16 floats each: uniform, served by the constant/L1 path
  const float* __restrict__ vm = p.viewmatrix;
  const float* __restrict__ pm = p.projmatrix;

  const float px = __ldg(p.means3D + 3 * (size_t)idx + 0);
  const float py = __ldg(p.means3D + 3 * (size_t)idx + 1);
  const float pz = __ldg(p.means3D + 3 * (size_t)idx + 2);
  // issued before the cull test so that they are in flight together with the position (28 B that a culled
  // surfel would not have needed, against one dependent round trip to HBM less for the visible ones)
  float2 sc_in = make_float2(0.f, 0.f);
  float4 q_in = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.transMat_precomp == nullptr) {
    sc_in = __ldg(reinterpret_cast<const float2*>(p.scales) + idx);
    q_in = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
  }
  float opac_in = __ldg(p.opacities + idx);
  if (RAW && p.raw) {
    // raw parameters (gaussian_model.py:104-113): scaling_activation = exp, rotation_activation = F.normalize,
    // opacity_activation = sigmoid, evaluated as torch's CUDA kernels evaluate them
    sc_in = activate_scaling(sc_in);
    q_in = activate_rotation(q_in).q;
    opac_in = activate_opacity(opac_in);
  }

  // p_view = [p 1] * V   (auxiliary.h:79-87), association: m12 + fma(z,m8, fma(x,m0, y*m4))
  const float vx = ADD(vm[12], FMA(pz, vm[8], FMA(px, vm[0], MUL(py, vm[4]))));
  const float vy = ADD(vm[13], FMA(pz, vm[9], FMA(px, vm[1], MUL(py, vm[5]))));
  const float vz = ADD(vm[14], FMA(pz, vm[10], FMA(px, vm[2], MUL(py, vm[6]))));

  do {
    if (!valid) break;
    if (vz <= NEAR_N) {  // auxiliary.h:200
      if (p.prefiltered) atomicOr(p.flags, 1u);
      break;
    }

    if (p.transMat_precomp == nullptr) {
      const float2 sc = sc_in;
      const float4 q = q_in;
      float3 L0, L1, L2;
      build_L(q, MUL(p.scale_modifier, sc.x), MUL(p.scale_modifier, sc.y), L0, L1, L2);
      // A = M^T * world2ndc, rows (L0,0), (L1,0), (p,1); columns c = 0..3 of projmatrix
      float A0[4], A1[4], A2[4];
#pragma unroll
      for (int c = 0; c < 4; c++) {
        A0[c] = FMA(L0.z, pm[8 + c], FMA(L0.x, pm[c], MUL(L0.y, pm[4 + c])));
        A1[c] = FMA(L1.z, pm[8 + c], FMA(L1.x, pm[c], MUL(L1.y, pm[4 + c])));
        A2[c] = ADD(pm[12 + c], FMA(pz, pm[8 + c], FMA(px, pm[c], MUL(py, pm[4 + c]))));
      }
      const float hW = MUL((float)p.W, 0.5f), hWm1 = MUL((float)(p.W - 1), 0.5f);
      const float hH = MUL((float)p.H, 0.5f), hHm1 = MUL((float)(p.H - 1), 0.5f);
      Tu0 = FMA(hWm1, A0[3], MUL(hW, A0[0])); Tu1 = FMA(hWm1, A1[3], MUL(hW, A1[0])); Tu2 = FMA(hWm1, A2[3], MUL(hW, A2[0]));
      Tv0 = FMA(hHm1, A0[3], MUL(hH, A0[1])); Tv1 = FMA(hHm1, A1[3], MUL(hH, A1[1])); Tv2 = FMA(hHm1, A2[3], MUL(hH, A2[1]));
      Tw0 = A0[3]; Tw1 = A1[3]; Tw2 = A2[3];
      // normal = L2 * V[:3,:3]  (auxiliary.h:100-108)
      nx = FMA(L2.z, vm[8], FMA(L2.x, vm[0], MUL(L2.y, vm[4])));
      ny = FMA(L2.z, vm[9], FMA(L2.x, vm[1], MUL(L2.y, vm[5])));
      nz = FMA(L2.z, vm[10], FMA(L2.x, vm[2], MUL(L2.y, vm[6])));
    } else {
      const float* T = p.transMat_precomp + 9 * (size_t)idx;
      Tu0 = __ldg(T + 0); Tu1 = __ldg(T + 1); Tu2 = __ldg(T + 2);
      Tv0 = __ldg(T + 3); Tv1 = __ldg(T + 4); Tv2 = __ldg(T + 5);
      Tw0 = __ldg(T + 6); Tw1 = __ldg(T + 7); Tw2 = __ldg(T + 8);
      nx = 0.0f; ny = 0.0f; nz = 1.0f;
    }

    // dual-visible flip (forward.cu:209-214): cos = -<p_view, n>
    const float pn = FMA(vz, nz, FMA(vx, nx, MUL(vy, ny)));
    if (pn == 0.0f) break;
    const float mult = (pn < 0.0f) ? 1.0f : -1.0f;
    nx = MUL(nx, mult); ny = MUL(ny, mult); nz = MUL(nz, mult);

    // AABB (forward.cu:119-145), cutoff 3 -> t = (9, 9, -1)
    const float d = FMA(-Tw2, Tw2, FMA(MUL(Tw0, Tw0), 9.0f, MUL(MUL(Tw1, Tw1), 9.0f)));
    if (d == 0.0f) break;
    const float inv = __frcp_rn(d);
    const float f9 = MUL(inv, 9.0f);
    cx = FMA(MUL(Tw2, Tu2), -inv, FMA(f9, MUL(Tu0, Tw0), MUL(f9, MUL(Tu1, Tw1))));
    cy = FMA(MUL(Tw2, Tv2), -inv, FMA(f9, MUL(Tv0, Tw0), MUL(f9, MUL(Tv1, Tw1))));
    const float su = FMA(MUL(Tu2, Tu2), inv, -FMA(f9, MUL(Tu0, Tu0), MUL(f9, MUL(Tu1, Tu1))));
    const float sv = FMA(MUL(Tv2, Tv2), inv, -FMA(f9, MUL(Tv0, Tv0), MUL(f9, MUL(Tv1, Tv1))));
    const float hx = __fsqrt_rn(fmaxf(FMA(cx, cx, su), 1e-4f));
    const float hy = __fsqrt_rn(fmaxf(FMA(cy, cy, sv), 1e-4f));
    const float radius = ceilf(fmaxf(fmaxf(hx, hy), __uint_as_float(GSM_MIN_RADIUS_BITS)));
    const int ri = (int)radius;
    const float rf = (float)ri;

    // tile rect (auxiliary.h:67-77); /16 is exact as *0.0625
    const int minx = min(p.grid_x, max(0, (int)MUL(SUB(cx, rf), 0.0625f)));
    const int miny = min(p.grid_y, max(0, (int)MUL(SUB(cy, rf), 0.0625f)));
    const int maxx = min(p.grid_x, max(0, (int)MUL(ADD(ADD(ADD(cx, rf), 16.0f), -1.0f), 0.0625f)));
    const int maxy = min(p.grid_y, max(0, (int)MUL(ADD(ADD(ADD(cy, rf), 16.0f), -1.0f), 0.0625f)));
    const uint32_t area = (uint32_t)((maxx - minx) * (maxy - miny));
    if (area == 0) break;
    rect_min = (uint32_t)minx | ((uint32_t)miny << 16);
    rect_max = (uint32_t)maxx | ((uint32_t)maxy << 16);
    // per-tile instance counts (binning.cu step 1): one red.global.add per (surfel, tile) instance
    for (int ty = miny; ty < maxy; ty++)
      for (int tx = minx; tx < maxx; tx++)
        atomicAdd(p.tile_count + (ty * p.grid_x + tx) * TILE_SUBS + (idx & (TILE_SUBS - 1)), 1u);

    if (BULK)   // this surfel is visible from here on: its SH row is requested now and lands under the box math below
      bulk_load(sh_smem + ((size_t)warp * 32 + lane) * SH_BULK_STRIDE, p.shs + (size_t)idx * p.M * 3, (uint32_t)(12 * p.M), bar);

    // conservative contribution box (common.cuh, CULL_RHO); fp64 so that the cancellation in
    // centre^2 - <f,T^2> cannot shrink it
    opac = opac_in;
    if (opac < ALPHA_MIN) {
      // alpha = opacity * exp(power) with power <= 0 never reaches 1/255 (forward.cu:385-392): empty box
      box = make_short4(32767, 32767, -32768, -32768);
    } else if (opac >= ALPHA_MIN) {  // (false for NaN: keep the unbounded box)
      // alpha >= 1/255  <=>  rho <= 2 ln(255 opacity); 3% + 0.2 margin on rho, capped at CULL_RHO (opacity <= 1)
      const double c2 = fmin((double)CULL_RHO * fmax(1.0, (double)opac * (double)opac),
                             2.0 * (double)logf(255.0f * opac) * 1.03 + 0.2);  // logf: 1 ulp, inside the 0.2 margin
      const double w0 = Tw0, w1 = Tw1, w2 = Tw2;
      const double dd = c2 * (w0 * w0 + w1 * w1) - w2 * w2;
      if (dd < -1e-6 * w2 * w2 && w2 > 0.0) {  // the whole disc is in front of the camera plane
        const double f0 = c2 / dd, f2 = -1.0 / dd;
        const double ccx = f0 * ((double)Tu0 * w0 + (double)Tu1 * w1) + f2 * (double)Tu2 * w2;
        const double ccy = f0 * ((double)Tv0 * w0 + (double)Tv1 * w1) + f2 * (double)Tv2 * w2;
        const double ex2 = ccx * ccx - (f0 * ((double)Tu0 * Tu0 + (double)Tu1 * Tu1) + f2 * (double)Tu2 * Tu2);
        const double ey2 = ccy * ccy - (f0 * ((double)Tv0 * Tv0 + (double)Tv1 * Tv1) + f2 * (double)Tv2 * Tv2);
        const double ex = sqrt_up(fmax(ex2, 0.0)) * 1.01 + (double)CULL_SLACK_PX;
        const double ey = sqrt_up(fmax(ey2, 0.0)) * 1.01 + (double)CULL_SLACK_PX;
        const double r2d = sqrt_up(0.5 * c2) + (double)CULL_SLACK_PX;  // low-pass disc around the reference centre
        const double x0 = fmin(ccx - ex, (double)cx - r2d), x1 = fmax(ccx + ex, (double)cx + r2d);
        const double y0 = fmin(ccy - ey, (double)cy - r2d), y1 = fmax(ccy + ey, (double)cy + r2d);
        if (x0 == x0 && x1 == x1 && y0 == y0 && y1 == y1) {  // not NaN
          box.x = (short)fmax(-32768.0, fmin(32767.0, floor(x0)));
          box.y = (short)fmax(-32768.0, fmin(32767.0, floor(y0)));
          box.z = (short)fmax(-32768.0, fmin(32767.0, ceil(x1)));
          box.w = (short)fmax(-32768.0, fmin(32767.0, ceil(y1)));
          // diagonal slabs of the same region (common.cuh, "cull record"): the extent of the conic along
          // (1,1) and (1,-1) is sqrt(S00 + S11 +- 2 S01) with S the conic's "covariance" (ex2, ey2 are its
          // diagonal, sxy its off-diagonal term), stretched and slackened like the box and united with the disc
          const double sxy = ccx * ccy - (f0 * ((double)Tu0 * Tv0 + (double)Tu1 * Tv1) + f2 * (double)Tu2 * Tv2);
          const double s00 = fmax(ex2, 0.0), s11 = fmax(ey2, 0.0);
          const double rt2 = 1.4142135623730951;
          const double ep = sqrt_up(fmax(s00 + s11 + 2.0 * sxy, 0.0)) * 1.01 + rt2 * (double)CULL_SLACK_PX + 0.01;
          const double em = sqrt_up(fmax(s00 + s11 - 2.0 * sxy, 0.0)) * 1.01 + rt2 * (double)CULL_SLACK_PX + 0.01;
          const double dp = (double)cx + (double)cy, dm = (double)cx - (double)cy, dr = rt2 * r2d + 0.01;
          const double plo = fmin(ccx + ccy - ep, dp - dr), phi = fmax(ccx + ccy + ep, dp + dr);
          const double mlo = fmin(ccx - ccy - em, dm - dr), mhi = fmax(ccx - ccy + em, dm + dr);
          if (plo == plo && phi == phi && mlo == mlo && mhi == mhi)
            diag = make_float4(__double2float_rd(plo), __double2float_ru(phi), __double2float_rd(mlo), __double2float_ru(mhi));
        }
      }
    }
    radius_out = ri;
    tiles_out = area;
  } while (false);

  // ---- colour (forward.cu:20-71): SH rows of the warp's visible surfels go through shared memory ----
  const bool visible = tiles_out != 0;
  const unsigned vis_mask = __ballot_sync(0xffffffffu, visible);
  float3 rgb = make_float3(0.f, 0.f, 0.f);
  int clamp_bits = 0;
  if (BULK) {
    // every lane of the warp is past its cull decision: arm the barrier with the bytes in flight, then wait for them
    if (lane == 0) mbar_expect_tx(bar, (uint32_t)(12 * p.M) * (uint32_t)__popc(vis_mask));
    if (visible) {
      const float* cp = p.campos;
      const float dx = SUB(px, cp[0]), dy = SUB(py, cp[1]), dz = SUB(pz, cp[2]);
      const float len = __fsqrt_rn(FMA(dz, dz, FMA(dx, dx, MUL(dy, dy))));
      const float3 dir = make_float3(__fdiv_rn(dx, len), __fdiv_rn(dy, len), __fdiv_rn(dz, len));
      mbar_wait(bar, 0);
      rgb = sh_to_rgb_row4(p.D, reinterpret_cast<const float4*>(sh_smem + ((size_t)warp * 32 + lane) * SH_BULK_STRIDE), dir);
      if (rgb.x < 0.0f) { clamp_bits |= 1; rgb.x = 0.0f; }
      if (rgb.y < 0.0f) { clamp_bits |= 2; rgb.y = 0.0f; }
      if (rgb.z < 0.0f) { clamp_bits |= 4; rgb.z = 0.0f; }
    }
  } else if (p.colors_precomp == nullptr) {
    float* wsm = sh_smem + (size_t)warp * 32 * (3 * p.M + 1);
    if (vis_mask != 0) {
      const int rows = min(32, p.P - warp_first), rs = 3 * p.M + 1;
      if (!RAW || p.shs_rest == nullptr) {
        sh_rows_load(wsm, p.shs + (size_t)warp_first * p.M * 3, 3 * p.M, rs, 0, rows, vis_mask, lane);
      } else {  // features_dc + features_rest, no concatenated copy
        sh_rows_load(wsm, p.shs + (size_t)warp_first * 3, 3, rs, 0, rows, vis_mask, lane);
        if (p.M > 1) sh_rows_load(wsm, p.shs_rest + (size_t)warp_first * (p.M - 1) * 3, 3 * p.M - 3, rs, 3, rows, vis_mask, lane);
      }
      __syncwarp();
    }
    if (visible) {
      const float* cp = p.campos;
      const float dx = SUB(px, cp[0]), dy = SUB(py, cp[1]), dz = SUB(pz, cp[2]);
      const float len = __fsqrt_rn(FMA(dz, dz, FMA(dx, dx, MUL(dy, dy))));
      const float3 dir = make_float3(__fdiv_rn(dx, len), __fdiv_rn(dy, len), __fdiv_rn(dz, len));
      rgb = sh_to_rgb(p.D, wsm + lane * (3 * p.M + 1), dir);
      if (rgb.x < 0.0f) { clamp_bits |= 1; rgb.x = 0.0f; }
      if (rgb.y < 0.0f) { clamp_bits |= 2; rgb.y = 0.0f; }
      if (rgb.z < 0.0f) { clamp_bits |= 4; rgb.z = 0.0f; }
    }
  } else if (visible) {
    rgb.x = __ldg(p.colors_precomp + 3 * (size_t)idx + 0);
    rgb.y = __ldg(p.colors_precomp + 3 * (size_t)idx + 1);
    rgb.z = __ldg(p.colors_precomp + 3 * (size_t)idx + 2);
  }
  if (!valid) return;
  if (visible) {
    p.bbox[idx] = box;
    p.clamp[idx] = (uint8_t)clamp_bits;
    p.cull[idx] = diag;
    p.rects[idx] = make_uint4(rect_min, rect_max, __float_as_uint(vz), 0u);
    float4* rec = p.records + (size_t)idx * REC_F4;
    rec[0] = make_float4(Tu0, Tu1, Tu2, cx);
    rec[1] = make_float4(Tv0, Tv1, Tv2, cy);
    rec[2] = make_float4(Tw0, Tw1, Tw2, opac);
    rec[3] = make_float4(nx, ny, nz, vz);
    rec[4] = make_float4(rgb.x, rgb.y, rgb.z, __int_as_float(clamp_bits));
  }
  p.radii[idx] = radius_out;
}

void launch_preprocess_fwd(const FwdPreprocessParams& p, cudaStream_t stream) {
  if (p.P <= 0) return;
  const bool raw = p.raw || p.shs_rest != nullptr;
  // per-row TMA bulk copies need 16-byte rows at 16-byte addresses: the concatenated [P,M,3] tensor with M % 4 == 0
  const bool bulk = !raw && p.colors_precomp == nullptr && p.M > 0 && p.M <= 16 && p.M % 4 == 0 &&
                    (reinterpret_cast<size_t>(p.shs) & 15) == 0 && getenv("GSM_NO_BULK") == nullptr;
  const size_t smem = bulk ? (size_t)(PFWD_THREADS / 32) * 32 * SH_BULK_STRIDE * sizeof(float) + (PFWD_THREADS / 32) * 8
                           : (p.colors_precomp == nullptr ? (size_t)(PFWD_THREADS / 32) * 32 * (3 * p.M + 1) * sizeof(float) : 0);
  static PerDeviceLevel configured;
  if (smem > 48 * 1024) configured.ensure(smem, [&] {
    cudaFuncSetAttribute(preprocess_fwd_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(preprocess_fwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  });
  const int grid = (p.P + PFWD_THREADS - 1) / PFWD_THREADS;
  if (bulk) preprocess_fwd_kernel<false, true><<<grid, PFWD_THREADS, smem, stream>>>(p);
  else if (raw) preprocess_fwd_kernel<true, false><<<grid, PFWD_THREADS, smem, stream>>>(p);
  else preprocess_fwd_kernel<false, false><<<grid, PFWD_THREADS, smem, stream>>>(p);
}

// markVisible (rasterizer_impl.cu:54-66): present = view_z > 0.2
__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means3D,
                                                           const float* __restrict__ vm,
                                                           uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float px = __ldg(means3D + 3 * (size_t)idx + 0);
  const float py = __ldg(means3D + 3 * (size_t)idx + 1);
  const float pz = __ldg(means3D + 3 * (size_t)idx + 2);
  const float vz = ADD(vm[14], FMA(pz, vm[10], FMA(px, vm[2], MUL(py, vm[6]))));
  present[idx] = (vz <= NEAR_N) ? 0 : 1;
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present,
                         cudaStream_t stream) {
  if (P <= 0) return;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, viewmatrix, present);
}

}  // namespace gsm

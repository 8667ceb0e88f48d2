// Fused surface regularisers for sm_100a: everything the training iteration does with the
// rasterizer's allmap between render() and total_loss.backward() --
//   gaussian_renderer/__init__.py:117-147  world-space normal, median / expected depth (nan_to_num), surf_depth,
//                                          surf_normal = depth_to_normal(view, surf_depth) * alpha.detach()
//   utils/point_utils.py:9-37              depths_to_points (pixel rays), central differences, cross, normalize
//   train.py:131-140                       normal_loss = l_n mean(1 - <rend_normal, surf_normal>), dist_loss = l_d mean(rend_dist)
// -- ~35 torch kernels forward with full-image temporaries and their autograd mirror, as two kernels that produce the
// two loss values AND dL/dallmap[7,H,W].
//
// Kernel 1 (per pixel): surf_depth at the pixel and its four neighbours -> points -> normal; the loss terms; the
// gradients that stay on the pixel (normal channels, distortion channel) and the gradient w.r.t. the two finite
// differences (6 floats to scratch).  Kernel 2 (per pixel): adjoint of the central differences (gather from the four
// neighbours' scratch), chain through point = depth * ray + origin, the depth mix, nan_to_num and expected = D/alpha.
// The per-camera ray matrix and origin are host-side constants (the reference rebuilds them with two 4x4 inverses on
// the GPU every iteration); HBM traffic: 7 maps read twice + 6 scratch floats written and gathered + 7 written.
#include "common.cuh"
#include <cfloat>

namespace gsm {

constexpr int REG_THREADS = 256;

struct RegParams {
  const float* allmap;  // [7,H,W]: expected-depth numerator, alpha, normal xyz (view space), median depth, distortion
  int H, W;
  float ray[9];      // row-major 3x3: ray_dir(x, y) = [x y 1] * ray
  float origin[3];   // camera centre
  float view[9];     // viewmatrix[:3,:3] as stored (row-major): rend_normal_k = sum_j n_j * view[k][j]
  float depth_ratio, one_minus_ratio;
  float k_normal;    // grad_scale * lambda_normal / (H W)
  float k_dist;      // grad_scale * lambda_dist / (H W)
  float lambda_normal, lambda_dist;
  float* scratch;    // [6,H,W]: dL/d(dx), dL/d(dy)
  double* sums;      // [2]: sum of normal_error, sum of rend_dist
  float* dL_dallmap; // [7,H,W]
  float* rend_normal;  // optional [3,H,W] (world space)
  float* surf_normal;  // optional [3,H,W]
  float* surf_depth;   // optional [H,W]
  float* loss_out;     // optional [3]: normal_loss + dist_loss, normal_loss, dist_loss
};

// torch.nan_to_num(x, 0, 0): nan -> 0, +inf -> 0, -inf -> lowest float
__device__ __forceinline__ float nan_to_num00(float x) {
  if (isnan(x)) return 0.0f;
  if (isinf(x)) return x > 0.0f ? 0.0f : -FLT_MAX;
  return x;
}

__device__ __forceinline__ float surf_depth_at(const RegParams& p, size_t plane, int y, int x) {
  const size_t at = (size_t)y * p.W + x;
  const float alpha = __ldg(p.allmap + plane + at);
  const float expected = nan_to_num00(__fdiv_rn(__ldg(p.allmap + at), alpha));
  const float median = nan_to_num00(__ldg(p.allmap + 5 * plane + at));
  return __fadd_rn(__fmul_rn(expected, p.one_minus_ratio), __fmul_rn(p.depth_ratio, median));
}

__device__ __forceinline__ float3 ray_at(const RegParams& p, int y, int x) {
  const float fx = (float)x, fy = (float)y;
  return make_float3(fx * p.ray[0] + fy * p.ray[3] + p.ray[6], fx * p.ray[1] + fy * p.ray[4] + p.ray[7],
                     fx * p.ray[2] + fy * p.ray[5] + p.ray[8]);
}

__device__ __forceinline__ float3 point_at(const RegParams& p, size_t plane, int y, int x) {
  const float s = surf_depth_at(p, plane, y, x);
  const float3 r = ray_at(p, y, x);
  // depth * rays_d + rays_o as two torch kernels: product rounded, then the sum
  return make_float3(__fadd_rn(__fmul_rn(s, r.x), p.origin[0]), __fadd_rn(__fmul_rn(s, r.y), p.origin[1]),
                     __fadd_rn(__fmul_rn(s, r.z), p.origin[2]));
}

__device__ __forceinline__ float3 cross3(float3 a, float3 b) {
  return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

__global__ void __launch_bounds__(REG_THREADS) regularizer_forward_kernel(const RegParams p) {
  __shared__ double red[2][REG_THREADS / 32];
  const size_t plane = (size_t)p.H * p.W;
  const size_t i = (size_t)blockIdx.x * REG_THREADS + threadIdx.x;
  double e_sum = 0.0, d_sum = 0.0;
  if (i < plane) {
    const int y = (int)(i / p.W), x = (int)(i - (size_t)y * p.W);
    const float alpha = __ldg(p.allmap + plane + i);
    const float nv0 = __ldg(p.allmap + 2 * plane + i), nv1 = __ldg(p.allmap + 3 * plane + i), nv2 = __ldg(p.allmap + 4 * plane + i);
    const float dist = __ldg(p.allmap + 6 * plane + i);
    float rn[3];
#pragma unroll
    for (int k = 0; k < 3; k++) rn[k] = nv0 * p.view[3 * k + 0] + nv1 * p.view[3 * k + 1] + nv2 * p.view[3 * k + 2];
    float3 n = make_float3(0.f, 0.f, 0.f), gdx = n, gdy = n;
    const bool interior = y >= 1 && y <= p.H - 2 && x >= 1 && x <= p.W - 2;
    if (interior) {
      const float3 pu = point_at(p, plane, y - 1, x), pd = point_at(p, plane, y + 1, x);
      const float3 pl = point_at(p, plane, y, x - 1), pr = point_at(p, plane, y, x + 1);
      const float3 dx = make_float3(pd.x - pu.x, pd.y - pu.y, pd.z - pu.z);   // points[2:] - points[:-2] along rows
      const float3 dy = make_float3(pr.x - pl.x, pr.y - pl.y, pr.z - pl.z);   // along columns
      const float3 c = cross3(dx, dy);
      const float len = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
      const float nrm = fmaxf(len, 1e-12f);  // F.normalize eps
      n = make_float3(c.x / nrm, c.y / nrm, c.z / nrm);
      // d(loss)/d(n) = -k_normal * alpha * rend_normal; through normalize, then the cross product
      const float3 gn = make_float3(-p.k_normal * alpha * rn[0], -p.k_normal * alpha * rn[1], -p.k_normal * alpha * rn[2]);
      float3 gc;
      if (len > 1e-12f) {
        const float dot = n.x * gn.x + n.y * gn.y + n.z * gn.z;
        gc = make_float3((gn.x - n.x * dot) / nrm, (gn.y - n.y * dot) / nrm, (gn.z - n.z * dot) / nrm);
      } else {
        gc = make_float3(gn.x / nrm, gn.y / nrm, gn.z / nrm);
      }
      gdx = cross3(dy, gc);
      gdy = cross3(gc, dx);
    }
    const float3 sn = make_float3(n.x * alpha, n.y * alpha, n.z * alpha);
    e_sum = (double)(1.0f - (rn[0] * sn.x + rn[1] * sn.y + rn[2] * sn.z));
    d_sum = (double)dist;
    p.scratch[i] = gdx.x; p.scratch[plane + i] = gdx.y; p.scratch[2 * plane + i] = gdx.z;
    p.scratch[3 * plane + i] = gdy.x; p.scratch[4 * plane + i] = gdy.y; p.scratch[5 * plane + i] = gdy.z;
    // d(loss)/d(rend_normal_k) = -k_normal * surf_normal_k; back to view space through view[k][j]
#pragma unroll
    for (int j = 0; j < 3; j++)
      p.dL_dallmap[(2 + j) * plane + i] = -p.k_normal * (sn.x * p.view[j] + sn.y * p.view[3 + j] + sn.z * p.view[6 + j]);
    p.dL_dallmap[6 * plane + i] = p.k_dist;
    if (p.rend_normal) { p.rend_normal[i] = rn[0]; p.rend_normal[plane + i] = rn[1]; p.rend_normal[2 * plane + i] = rn[2]; }
    if (p.surf_normal) { p.surf_normal[i] = sn.x; p.surf_normal[plane + i] = sn.y; p.surf_normal[2 * plane + i] = sn.z; }
    if (p.surf_depth) p.surf_depth[i] = surf_depth_at(p, plane, y, x);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    e_sum += __shfl_xor_sync(0xffffffffu, e_sum, o);
    d_sum += __shfl_xor_sync(0xffffffffu, d_sum, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = e_sum; red[1][warp] = d_sum; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int w = 0; w < REG_THREADS / 32; w++) { a += red[0][w]; b += red[1][w]; }
    atomicAdd(p.sums + 0, a);
    atomicAdd(p.sums + 1, b);
  }
}

__global__ void __launch_bounds__(REG_THREADS) regularizer_backward_kernel(const RegParams p) {
  const size_t plane = (size_t)p.H * p.W;
  const size_t i = (size_t)blockIdx.x * REG_THREADS + threadIdx.x;
  if (i == 0 && p.loss_out != nullptr) {
    const double nl = (double)p.lambda_normal * p.sums[0] / (double)plane, dl = (double)p.lambda_dist * p.sums[1] / (double)plane;
    p.loss_out[0] = (float)(nl + dl);
    p.loss_out[1] = (float)nl;
    p.loss_out[2] = (float)dl;
  }
  if (i >= plane) return;
  const int y = (int)(i / p.W), x = (int)(i - (size_t)y * p.W);
  // adjoint of dx(y', x) = P(y'+1, x) - P(y'-1, x) and dy(y, x') = P(y, x'+1) - P(y, x'-1); the scratch of
  // non-interior centres is zero
  float3 g = make_float3(0.f, 0.f, 0.f);
  if (y >= 1)       { const size_t a = i - p.W; g.x += p.scratch[a]; g.y += p.scratch[plane + a]; g.z += p.scratch[2 * plane + a]; }
  if (y <= p.H - 2) { const size_t a = i + p.W; g.x -= p.scratch[a]; g.y -= p.scratch[plane + a]; g.z -= p.scratch[2 * plane + a]; }
  if (x >= 1)       { const size_t a = i - 1; g.x += p.scratch[3 * plane + a]; g.y += p.scratch[4 * plane + a]; g.z += p.scratch[5 * plane + a]; }
  if (x <= p.W - 2) { const size_t a = i + 1; g.x -= p.scratch[3 * plane + a]; g.y -= p.scratch[4 * plane + a]; g.z -= p.scratch[5 * plane + a]; }
  const float3 r = ray_at(p, y, x);
  const float g_s = g.x * r.x + g.y * r.y + g.z * r.z;
  const float alpha = __ldg(p.allmap + plane + i), num = __ldg(p.allmap + i), med = __ldg(p.allmap + 5 * plane + i);
  const float q = __fdiv_rn(num, alpha);
  // nan_to_num passes the gradient where its input is finite; expected = num / alpha
  const float g_q = (isnan(q) || isinf(q)) ? 0.0f : g_s * p.one_minus_ratio;
  const bool live = alpha != 0.0f;   // torch yields 0/0 = NaN here (pixels nothing was blended into); written as 0
  p.dL_dallmap[i] = live ? g_q / alpha : 0.0f;
  p.dL_dallmap[plane + i] = live ? -g_q * (q / alpha) : 0.0f;
  p.dL_dallmap[5 * plane + i] = (isnan(med) || isinf(med)) ? 0.0f : g_s * p.depth_ratio;
}

}  // namespace gsm

using namespace gsm;

extern "C" size_t gsm_surface_regularizer_workspace_bytes(int H, int W) {
  if (H <= 0 || W <= 0) return 256;
  return 256 + align_up((size_t)6 * H * W * sizeof(float), 256);
}

extern "C" int gsm_surface_regularizer(const gsm_regularizer_args* a, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a == nullptr) { set_error("gsm_surface_regularizer: null args"); return GSM_ERR_INVALID_ARGUMENT; }
  if (a->H <= 0 || a->W <= 0) { set_error("gsm_surface_regularizer: bad sizes H=%d W=%d", a->H, a->W); return GSM_ERR_INVALID_ARGUMENT; }
  if (a->allmap == nullptr || a->dL_dallmap == nullptr) { set_error("gsm_surface_regularizer: null allmap / gradient"); return GSM_ERR_INVALID_ARGUMENT; }
  if (a->workspace == nullptr || a->workspace_bytes < gsm_surface_regularizer_workspace_bytes(a->H, a->W)) {
    set_error("gsm_surface_regularizer: workspace too small"); return GSM_ERR_WORKSPACE_TOO_SMALL;
  }
  RegParams p{};
  p.allmap = a->allmap; p.H = a->H; p.W = a->W;
  for (int k = 0; k < 9; k++) { p.ray[k] = a->ray_matrix[k]; p.view[k] = a->view_rotation[k]; }
  for (int k = 0; k < 3; k++) p.origin[k] = a->ray_origin[k];
  const double n = (double)a->H * a->W;
  p.depth_ratio = a->depth_ratio;
  p.one_minus_ratio = (float)(1.0 - (double)a->depth_ratio);
  p.lambda_normal = a->lambda_normal; p.lambda_dist = a->lambda_dist;
  p.k_normal = (float)((double)a->grad_scale * a->lambda_normal / n);
  p.k_dist = (float)((double)a->grad_scale * a->lambda_dist / n);
  char* ws = (char*)a->workspace;
  p.sums = (double*)ws;
  p.scratch = (float*)(ws + 256);
  p.dL_dallmap = a->dL_dallmap;
  p.rend_normal = a->rend_normal; p.surf_normal = a->surf_normal; p.surf_depth = a->surf_depth; p.loss_out = a->loss_out;
  GSM_CUDA(cudaMemsetAsync(p.sums, 0, 16, stream));
  const unsigned blocks = (unsigned)(((size_t)a->H * a->W + REG_THREADS - 1) / REG_THREADS);
  {
    StageMark m(GSM_STAGE_REGULARIZER, stream);
    regularizer_forward_kernel<<<blocks, REG_THREADS, 0, stream>>>(p);
    regularizer_backward_kernel<<<blocks, REG_THREADS, 0, stream>>>(p);
  }
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

// Fused multi-tensor Adam step for sm_100a: the optimizer step that follows the rasterizer's
// backward on every training iteration (train.py:212-213 `gaussians.optimizer.step()`; the optimizer is
// torch.optim.Adam(l, lr=0.0, eps=1e-15) over six parameter groups, scene/gaussian_model.py:283-292).
//
// WHAT (torch/optim/adam.py, the default non-amsgrad, weight_decay = 0, maximize = False branch; every element of every
// parameter is updated on every step, visible this iteration or not):
//     m  = lerp(m, g, 1 - b1)                     v  = v * b2 + (1 - b2) * g * g
//     p += -(lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// with the scalar factors formed in double on the host exactly as the Python code forms them and rounded to fp32 once,
// and the device expressions in the association torch's own elementwise kernels compile to (fma where nvcc contracts),
// so the result tracks torch's foreach path to the last bits.
// HOW: one launch for all tensors; a flat grid of 4096-element chunks, each thread holding four independent 128-bit
// loads of each of the four streams in flight.  Pure HBM streaming: 16 B read + 12 B written per element.
#include "common.cuh"
#include <cmath>

namespace gsm {

constexpr int ADAM_MAX_TENSORS = 8;
constexpr int ADAM_THREADS = 256;
constexpr int ADAM_VEC_PER_THREAD = 4;
constexpr size_t ADAM_CHUNK = (size_t)ADAM_THREADS * ADAM_VEC_PER_THREAD * 4;  // elements per CTA

struct AdamTensor {
  float* p;
  const float* g;
  float* m;
  float* v;
  size_t n;
  float w1;         // 1 - beta1
  float beta2;
  float w2;         // 1 - beta2
  float bc2_sqrt;   // sqrt(1 - beta2^t)
  float eps;
  float neg_step;   // -(lr / (1 - beta1^t))
};
struct AdamBatch {
  AdamTensor t[ADAM_MAX_TENSORS];
  unsigned first_chunk[ADAM_MAX_TENSORS + 1];  // prefix of chunk counts
  int count;
};

// IEEE division and square root, same results as __fdiv_rn / __fsqrt_rn, without their slow paths on the operands
// this kernel meets all the time: surfels that no view has touched yet have grad = m = v = 0 and barely visible ones
// have denormal second moments; the library routines branch to a subroutine for zero / denormal operands, and one
// such lane stalls its warp (measured: 2.2 ms instead of 1.24 ms per step with 30 % zero rows).  Zero is answered
// directly, tiny operands are scaled by an exact power of two into the normal range and the result scaled back.
__device__ __forceinline__ float sqrt_guarded(float v) {
  const bool zero = v == 0.0f, tiny = v < 5.4210109e-20f;                     // 2^-64
  const float in = zero ? 1.0f : (tiny ? v * 1.8446744e19f : v);              // * 2^64, exact
  const float r = __fsqrt_rn(in);
  return zero ? 0.0f : (tiny ? r * 2.3283064e-10f : r);                       // * 2^-32, exact
}
__device__ __forceinline__ float div_guarded(float a, float b) {
  const float ab = fabsf(b), aa = fabsf(a);
  if (!(ab >= 1.0e-30f && ab <= 1.0e30f && aa <= 1.0e30f)) return __fdiv_rn(a, b);  // exotic operands: library path
  const bool zero = aa == 0.0f, tiny = aa < 5.4210109e-20f;
  const float in = zero ? 1.0f : (tiny ? a * 1.8446744e19f : a);
  const float q = __fdiv_rn(in, b);
  return zero ? 0.0f : (tiny ? q * 5.4210109e-20f : q);  // (a quotient below 2^-126 is rounded twice: beyond fp32 += reach)
}

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const AdamTensor& t) {
  // lerp with |w| < 0.5: m + w (g - m); with w >= 0.5: g - (g - m)(1 - w)        (ATen/native/Lerp.h)
  const float diff = __fsub_rn(g, m);
  m = (fabsf(t.w1) < 0.5f) ? __fmaf_rn(t.w1, diff, m) : __fsub_rn(g, __fmul_rn(diff, __fsub_rn(1.0f, t.w1)));
  // _foreach_mul_(beta2) then _foreach_addcmul_(g, g, 1 - beta2): v*b2 + w2*(g*g)   (ForeachPointwiseOp.cu)
  v = __fmaf_rn(t.w2, __fmul_rn(g, g), __fmul_rn(v, t.beta2));
  const float denom = __fadd_rn(div_guarded(sqrt_guarded(v), t.bc2_sqrt), t.eps);
  // addcdiv_(m, denom, value = neg_step): p + neg_step * (m / denom)
  p = __fmaf_rn(t.neg_step, div_guarded(m, denom), p);
}

__global__ void __launch_bounds__(ADAM_THREADS) adam_step_kernel(const __grid_constant__ AdamBatch b) {
  int ti = 0;
#pragma unroll
  for (int k = 1; k < ADAM_MAX_TENSORS; k++) ti += (k < b.count && blockIdx.x >= b.first_chunk[k]) ? 1 : 0;
  const AdamTensor& t = b.t[ti];
  const size_t base = (size_t)(blockIdx.x - b.first_chunk[ti]) * ADAM_CHUNK;
  const size_t n = t.n;
  if (base + ADAM_CHUNK <= n) {
    float4 p[ADAM_VEC_PER_THREAD], g[ADAM_VEC_PER_THREAD], m[ADAM_VEC_PER_THREAD], v[ADAM_VEC_PER_THREAD];
#pragma unroll
    for (int k = 0; k < ADAM_VEC_PER_THREAD; k++) {
      const size_t at = base / 4 + (size_t)k * ADAM_THREADS + threadIdx.x;
      g[k] = __ldcs(reinterpret_cast<const float4*>(t.g) + at);
      p[k] = __ldcs(reinterpret_cast<const float4*>(t.p) + at);
      m[k] = __ldcs(reinterpret_cast<const float4*>(t.m) + at);
      v[k] = __ldcs(reinterpret_cast<const float4*>(t.v) + at);
    }
#pragma unroll
    for (int k = 0; k < ADAM_VEC_PER_THREAD; k++) {
      adam_update(p[k].x, g[k].x, m[k].x, v[k].x, t);
      adam_update(p[k].y, g[k].y, m[k].y, v[k].y, t);
      adam_update(p[k].z, g[k].z, m[k].z, v[k].z, t);
      adam_update(p[k].w, g[k].w, m[k].w, v[k].w, t);
      const size_t at = base / 4 + (size_t)k * ADAM_THREADS + threadIdx.x;
      reinterpret_cast<float4*>(t.p)[at] = p[k];
      __stcs(reinterpret_cast<float4*>(t.m) + at, m[k]);
      __stcs(reinterpret_cast<float4*>(t.v) + at, v[k]);
    }
  } else {  // the tensor's last, ragged chunk
    for (size_t i = base + threadIdx.x; i < n; i += ADAM_THREADS) {
      float p = t.p[i], m = t.m[i], v = t.v[i];
      adam_update(p, t.g[i], m, v, t);
      t.p[i] = p; t.m[i] = m; t.v[i] = v;
    }
  }
}


// ---- densification bookkeeping of the same iteration (train.py:199-203 and
// scene/gaussian_model.py:551-553): for every surfel the filter selects (radii > 0 when no filter is passed, which is
// render()'s visibility_filter),
//     max_radii2D = max(max_radii2D, (float)radii);  xyz_gradient_accum += ||viewspace_grad[i, :]||_2;  denom += 1
// Six indexed torch ops with boolean-mask gathers in the reference; here one pass, four surfels per thread so that every
// stream is a 128-bit access ([P,3] gradient rows: 3 x float4 per four surfels).  The norm is sqrt((x*x + z*z) + y*y)
// with separately rounded products (no fma contraction): the association of torch's reduction over a contiguous dimension
// of three (two lanes per output: lane 0 sums elements 0 and 2, lane 1 holds element 1, then one shuffle-add).
constexpr int DENS_THREADS = 256;

__device__ __forceinline__ void dens_update(bool on, int r, float gx, float gy, float gz, float& mr, float& acc, float& den) {
  if (!on) return;
  mr = fmaxf(mr, (float)r);
  const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gz, gz)), __fmul_rn(gy, gy));
  acc = __fadd_rn(acc, sqrt_guarded(n2));
  den = __fadd_rn(den, 1.0f);
}

__global__ void __launch_bounds__(DENS_THREADS)
densification_stats_kernel(const int* __restrict__ radii, const float* __restrict__ grad, const unsigned char* __restrict__ filter,
                           float* __restrict__ max_radii, float* __restrict__ accum, float* __restrict__ denom, size_t P) {
  const size_t q = (size_t)blockIdx.x * DENS_THREADS + threadIdx.x;  // group of four surfels
  const size_t i0 = q * 4;
  if (i0 >= P) return;
  if (i0 + 4 <= P) {
    const int4 r = __ldg(reinterpret_cast<const int4*>(radii) + q);
    bool on[4];
    if (filter != nullptr) {
      const uchar4 f = __ldg(reinterpret_cast<const uchar4*>(filter) + q);
      on[0] = f.x != 0; on[1] = f.y != 0; on[2] = f.z != 0; on[3] = f.w != 0;
    } else {
      on[0] = r.x > 0; on[1] = r.y > 0; on[2] = r.z > 0; on[3] = r.w > 0;
    }
    if (!(on[0] | on[1] | on[2] | on[3])) return;
    const float4* g4 = reinterpret_cast<const float4*>(grad) + q * 3;
    const float4 a = __ldg(g4), b = __ldg(g4 + 1), c = __ldg(g4 + 2);
    float4 mr = reinterpret_cast<float4*>(max_radii)[q];
    float4 ac = reinterpret_cast<float4*>(accum)[q];
    float4 dn = reinterpret_cast<float4*>(denom)[q];
    dens_update(on[0], r.x, a.x, a.y, a.z, mr.x, ac.x, dn.x);
    dens_update(on[1], r.y, a.w, b.x, b.y, mr.y, ac.y, dn.y);
    dens_update(on[2], r.z, b.z, b.w, c.x, mr.z, ac.z, dn.z);
    dens_update(on[3], r.w, c.y, c.z, c.w, mr.w, ac.w, dn.w);
    reinterpret_cast<float4*>(max_radii)[q] = mr;
    reinterpret_cast<float4*>(accum)[q] = ac;
    reinterpret_cast<float4*>(denom)[q] = dn;
  } else {  // ragged tail (P not a multiple of four)
    for (size_t i = i0; i < P; i++) {
      const int r = radii[i];
      const bool on = filter != nullptr ? filter[i] != 0 : r > 0;
      dens_update(on, r, grad[3 * i], grad[3 * i + 1], grad[3 * i + 2], max_radii[i], accum[i], denom[i]);
    }
  }
}

}  // namespace gsm

using namespace gsm;

extern "C" int gsm_adam_step(const gsm_adam_tensor* tensors, int n_tensors, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n_tensors < 0 || (n_tensors > 0 && tensors == nullptr)) { set_error("gsm_adam_step: bad tensor list"); return GSM_ERR_INVALID_ARGUMENT; }
  for (int i = 0; i < n_tensors; i++) {
    const gsm_adam_tensor& a = tensors[i];
    if (a.n == 0) continue;
    if (a.param == nullptr || a.grad == nullptr || a.exp_avg == nullptr || a.exp_avg_sq == nullptr) {
      set_error("gsm_adam_step: tensor %d has a null pointer", i); return GSM_ERR_INVALID_ARGUMENT;
    }
    if (a.step < 1) { set_error("gsm_adam_step: tensor %d: step must be >= 1 (the count AFTER this step)", i); return GSM_ERR_INVALID_ARGUMENT; }
    if ((((uintptr_t)a.param | (uintptr_t)a.grad | (uintptr_t)a.exp_avg | (uintptr_t)a.exp_avg_sq) & 15u) != 0) {
      set_error("gsm_adam_step: tensor %d is not 16-byte aligned", i); return GSM_ERR_INVALID_ARGUMENT;
    }
    if ((a.n + ADAM_CHUNK - 1) / ADAM_CHUNK > 0x7fffffffull / ADAM_MAX_TENSORS) { set_error("gsm_adam_step: tensor %d too large", i); return GSM_ERR_INVALID_ARGUMENT; }
  }
  StageMark mark(GSM_STAGE_ADAM, stream);
  int i = 0;
  while (i < n_tensors) {
    AdamBatch b{};
    unsigned chunks = 0;
    while (i < n_tensors && b.count < ADAM_MAX_TENSORS) {
      const gsm_adam_tensor& a = tensors[i++];
      if (a.n == 0) continue;
      AdamTensor& t = b.t[b.count];
      t.p = a.param; t.g = a.grad; t.m = a.exp_avg; t.v = a.exp_avg_sq; t.n = a.n;
      // torch/optim/adam.py (_multi_tensor_adam / _single_tensor_adam), Python floats = doubles:
      const double bc1 = 1.0 - std::pow(a.beta1, (double)a.step);
      const double bc2 = 1.0 - std::pow(a.beta2, (double)a.step);
      t.w1 = (float)(1.0 - a.beta1);
      t.beta2 = (float)a.beta2;
      t.w2 = (float)(1.0 - a.beta2);
      t.bc2_sqrt = (float)std::pow(bc2, 0.5);  // `bc ** 0.5`
      t.eps = (float)a.eps;
      t.neg_step = (float)(-(a.lr / bc1));
      b.first_chunk[b.count] = chunks;
      chunks += (unsigned)((a.n + ADAM_CHUNK - 1) / ADAM_CHUNK);
      b.count++;
    }
    for (int k = b.count; k <= ADAM_MAX_TENSORS; k++) b.first_chunk[k] = chunks;
    if (chunks) adam_step_kernel<<<chunks, ADAM_THREADS, 0, stream>>>(b);
  }
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

extern "C" int gsm_densification_stats(const int* radii, const float* viewspace_grad, const unsigned char* update_filter,
                                       float* max_radii2D, float* xyz_gradient_accum, float* denom, size_t P, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P == 0) return GSM_OK;
  if (radii == nullptr || viewspace_grad == nullptr || max_radii2D == nullptr || xyz_gradient_accum == nullptr || denom == nullptr) {
    set_error("gsm_densification_stats: null pointer"); return GSM_ERR_INVALID_ARGUMENT;
  }
  if ((((uintptr_t)radii | (uintptr_t)viewspace_grad | (uintptr_t)max_radii2D | (uintptr_t)xyz_gradient_accum | (uintptr_t)denom) & 15u) != 0 ||
      ((uintptr_t)update_filter & 3u) != 0) {
    set_error("gsm_densification_stats: arrays must be 16-byte aligned (the filter 4-byte)"); return GSM_ERR_INVALID_ARGUMENT;
  }
  const size_t groups = (P + 3) / 4, ctas = (groups + DENS_THREADS - 1) / DENS_THREADS;
  if (ctas > 0x7fffffffull) { set_error("gsm_densification_stats: too many surfels"); return GSM_ERR_INVALID_ARGUMENT; }
  densification_stats_kernel<<<(unsigned)ctas, DENS_THREADS, 0, stream>>>(radii, viewspace_grad, update_filter, max_radii2D,
                                                                        xyz_gradient_accum, denom, P);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

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

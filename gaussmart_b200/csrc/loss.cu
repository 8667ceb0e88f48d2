// Fused photometric loss for sm_100a: (1 - lambda) * L1 + lambda * (1 - SSIM), value AND gradient
// w.r.t. the rendered image in two kernels.  SURVEY.md 8(f) row 1: the caller of the rasterizer on the
// training path (train.py:113-114) evaluates it on every iteration through utils/loss_utils.py:16-17
// (l1_loss) and :32-62 (ssim / _ssim: five grouped 11x11 conv2d with a sigma = 1.5 Gaussian window and
// zero padding, C1 = 0.01^2, C2 = 0.03^2, plain mean), and autograd replays them backwards.
//
// WHAT: loss = (1-l) mean|x - y| + l (1 - mean S),  S = (2 m1 m2 + C1)(2 s12 + C2) / ((m1^2 + m2^2 + C1)(s1 + s2 + C2)),
//       m1 = W*x, m2 = W*y, s1 = W*x^2 - m1^2, s2 = W*y^2 - m2^2, s12 = W*xy - m1 m2   (W: 11x11 Gaussian, zero padded)
//       dL/dx = (1-l)/n sign(x - y) - l/n [ W*a + 2 x (W*b) + y (W*c) ],   n = C H W,
//       a = dS/dm1 (total, through s1 and s12 too), b = dS/d(W*x^2) = -S / B2, c = dS/d(W*xy) = 2 A1 / (B1 B2).
// HOW: the window is separable (outer product of the 1-D Gaussian), so each 32x32 tile (+ halo 5) is filtered in
// shared memory with 11 + 11 taps per map instead of 121: kernel 1 produces the SSIM sums and the three partial
// maps, kernel 2 filters those and assembles the gradient.  HBM traffic: 2 images + 3 maps written, 3 maps + 2
// images read, 1 gradient written = 11 floats per pixel-channel; the library path moves > 100.
#include "common.cuh"
#include <cmath>

namespace gsm {

constexpr int LW = 11, LH = 5;           // window size, half width
constexpr int LT = 32;                   // output tile edge
constexpr int LTH = LT + 2 * LH;         // tile + halo = 42
constexpr int LOSS_THREADS = 256;

__constant__ float c_gauss[LW];

struct LossLayout {
  size_t sums;   // double[2]: sum |x-y|, sum S
  size_t maps;   // float[3][C*H*W]: a, b, c
  size_t total;
};
static LossLayout loss_layout(int C, int H, int W) {
  LossLayout L{};
  const size_t n = (size_t)C * H * W;
  L.sums = 0;
  L.maps = 256;
  L.total = 256 + align_up(3 * n * sizeof(float), 256);
  return L;
}

// tile (with halo) of a [C,H,W] image into shared memory, zeros outside the image (conv2d zero padding)
__device__ __forceinline__ void load_tile(float (*dst)[LTH], const float* __restrict__ img, int H, int W, int x0, int y0) {
  for (int i = threadIdx.x; i < LTH * LTH; i += LOSS_THREADS) {
    const int r = i / LTH, c = i - r * LTH;
    const int y = y0 + r - LH, x = x0 + c - LH;
    dst[r][c] = (y >= 0 && y < H && x >= 0 && x < W) ? __ldg(img + (size_t)y * W + x) : 0.0f;
  }
}

__global__ void __launch_bounds__(LOSS_THREADS)
ssim_forward_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int H, int W,
                    float* __restrict__ map_a, float* __restrict__ map_b, float* __restrict__ map_c,
                    double* __restrict__ sums) {
  __shared__ float t1[LTH][LTH], t2[LTH][LTH];
  __shared__ float hz[5][LTH][LT];  // horizontally filtered x, y, x^2, y^2, xy
  __shared__ double red[2][LOSS_THREADS / 32];
  const int ch = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
  const size_t plane = (size_t)ch * H * W;
  load_tile(t1, img1 + plane, H, W, x0, y0);
  load_tile(t2, img2 + plane, H, W, x0, y0);
  __syncthreads();
  // horizontal pass, 4 adjacent outputs per thread: 14 + 14 shared-memory reads feed 4 x 5 sums (was 22 per output)
  for (int i = threadIdx.x; i < LTH * (LT / 4); i += LOSS_THREADS) {
    const int r = i / (LT / 4), c0 = (i - r * (LT / 4)) * 4;
    float a[LW + 3], b[LW + 3];
#pragma unroll
    for (int k = 0; k < LW + 3; k++) { a[k] = t1[r][c0 + k]; b[k] = t2[r][c0 + k]; }
#pragma unroll
    for (int o = 0; o < 4; o++) {
      float s1 = 0.f, s2 = 0.f, s11 = 0.f, s22 = 0.f, s12 = 0.f;
#pragma unroll
      for (int k = 0; k < LW; k++) {
        const float w = c_gauss[k], x = a[o + k], y = b[o + k];
        s1 += w * x; s2 += w * y; s11 += w * x * x; s22 += w * y * y; s12 += w * x * y;
      }
      hz[0][r][c0 + o] = s1; hz[1][r][c0 + o] = s2; hz[2][r][c0 + o] = s11; hz[3][r][c0 + o] = s22; hz[4][r][c0 + o] = s12;
    }
  }
  __syncthreads();
  const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
  double l1_sum = 0.0, s_sum = 0.0;
  // vertical pass, 4 stacked outputs per thread (one thread per column and group of 4 rows: exactly one round)
  for (int i = threadIdx.x; i < (LT / 4) * LT; i += LOSS_THREADS) {
    const int r0 = (i / LT) * 4, c = i % LT;
    const int x = x0 + c;
    if (x >= W || y0 + r0 >= H) continue;
    float m1[4] = {0.f, 0.f, 0.f, 0.f}, m2[4] = {0.f, 0.f, 0.f, 0.f}, e11[4] = {0.f, 0.f, 0.f, 0.f},
          e22[4] = {0.f, 0.f, 0.f, 0.f}, e12[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < LW + 3; k++) {
      const float h0 = hz[0][r0 + k][c], h1 = hz[1][r0 + k][c], h2 = hz[2][r0 + k][c], h3 = hz[3][r0 + k][c], h4 = hz[4][r0 + k][c];
#pragma unroll
      for (int o = 0; o < 4; o++) {
        if (k - o >= 0 && k - o < LW) {
          const float w = c_gauss[k - o];
          m1[o] += w * h0; m2[o] += w * h1; e11[o] += w * h2; e22[o] += w * h3; e12[o] += w * h4;
        }
      }
    }
#pragma unroll
    for (int o = 0; o < 4; o++) {
      const int r = r0 + o, y = y0 + r;
      if (y >= H) break;
      const float s1 = e11[o] - m1[o] * m1[o], s2 = e22[o] - m2[o] * m2[o], s12 = e12[o] - m1[o] * m2[o];
      const float A1 = 2.f * m1[o] * m2[o] + C1, A2 = 2.f * s12 + C2;
      const float B1 = m1[o] * m1[o] + m2[o] * m2[o] + C1, B2 = s1 + s2 + C2;
      const float iB1 = 1.0f / B1, iB2 = 1.0f / B2;
      const float S = A1 * A2 * iB1 * iB2;
      const size_t at = plane + (size_t)y * W + x;
      map_a[at] = 2.f * m2[o] * (A2 - A1) * iB1 * iB2 + 2.f * m1[o] * S * (iB2 - iB1);
      map_b[at] = -S * iB2;
      map_c[at] = 2.f * A1 * iB1 * iB2;
      s_sum += (double)S;
      l1_sum += (double)fabsf(t1[r + LH][c + LH] - t2[r + LH][c + LH]);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    l1_sum += __shfl_xor_sync(0xffffffffu, l1_sum, o);
    s_sum += __shfl_xor_sync(0xffffffffu, s_sum, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = l1_sum; red[1][warp] = s_sum; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int w = 0; w < LOSS_THREADS / 32; w++) { a += red[0][w]; b += red[1][w]; }
    atomicAdd(sums + 0, a);
    atomicAdd(sums + 1, b);
  }
}

__global__ void __launch_bounds__(LOSS_THREADS)
ssim_backward_kernel(const float* __restrict__ img1, const float* __restrict__ img2, int C, int H, int W,
                     const float* __restrict__ map_a, const float* __restrict__ map_b, const float* __restrict__ map_c,
                     const double* __restrict__ sums, float lambda, float grad_scale, float* __restrict__ loss_out,
                     float* __restrict__ dL_dimg) {
  __shared__ float t[3][LTH][LTH];
  __shared__ float hz[3][LTH][LT];
  const int ch = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
  const size_t plane = (size_t)ch * H * W;
  const double n = (double)C * H * W;
  if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0 && loss_out != nullptr) {
    const double l1 = sums[0] / n, ss = sums[1] / n;
    loss_out[0] = (float)((1.0 - (double)lambda) * l1 + (double)lambda * (1.0 - ss));
    loss_out[1] = (float)l1;
    loss_out[2] = (float)ss;
  }
  load_tile(t[0], map_a + plane, H, W, x0, y0);
  load_tile(t[1], map_b + plane, H, W, x0, y0);
  load_tile(t[2], map_c + plane, H, W, x0, y0);
  __syncthreads();
  for (int i = threadIdx.x; i < LTH * (LT / 4); i += LOSS_THREADS) {
    const int r = i / (LT / 4), c0 = (i - r * (LT / 4)) * 4;
#pragma unroll
    for (int m = 0; m < 3; m++) {
      float a[LW + 3];
#pragma unroll
      for (int k = 0; k < LW + 3; k++) a[k] = t[m][r][c0 + k];
#pragma unroll
      for (int o = 0; o < 4; o++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < LW; k++) acc += c_gauss[k] * a[o + k];
        hz[m][r][c0 + o] = acc;
      }
    }
  }
  __syncthreads();
  const float k_l1 = grad_scale * (1.0f - lambda) / (float)n, k_ss = -grad_scale * lambda / (float)n;
  for (int i = threadIdx.x; i < (LT / 4) * LT; i += LOSS_THREADS) {
    const int r0 = (i / LT) * 4, c = i % LT;
    const int x = x0 + c;
    if (x >= W || y0 + r0 >= H) continue;
    float fa[4] = {0.f, 0.f, 0.f, 0.f}, fb[4] = {0.f, 0.f, 0.f, 0.f}, fc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < LW + 3; k++) {
      const float h0 = hz[0][r0 + k][c], h1 = hz[1][r0 + k][c], h2 = hz[2][r0 + k][c];
#pragma unroll
      for (int o = 0; o < 4; o++) {
        if (k - o >= 0 && k - o < LW) {
          const float w = c_gauss[k - o];
          fa[o] += w * h0; fb[o] += w * h1; fc[o] += w * h2;
        }
      }
    }
#pragma unroll
    for (int o = 0; o < 4; o++) {
      const int y = y0 + r0 + o;
      if (y >= H) break;
      const size_t at = plane + (size_t)y * W + x;
      const float xv = __ldg(img1 + at), yv = __ldg(img2 + at);
      const float d = xv - yv;
      const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);  // torch.abs backward: sign, 0 at 0
      dL_dimg[at] = k_l1 * sgn + k_ss * (fa[o] + 2.f * xv * fb[o] + yv * fc[o]);
    }
  }
}

static void upload_window() {
  // utils/loss_utils.py:22-24: torch.Tensor([exp(-(x - 5)^2 / (2 sigma^2))]) (doubles rounded to fp32), / its fp32 sum
  float g[LW];
  float sum = 0.f;
  for (int k = 0; k < LW; k++) { g[k] = (float)std::exp(-(double)((k - LH) * (k - LH)) / (2.0 * 1.5 * 1.5)); }
  for (int k = 0; k < LW; k++) sum += g[k];
  for (int k = 0; k < LW; k++) g[k] = g[k] / sum;
  cudaMemcpyToSymbol(c_gauss, g, sizeof(g));
}

}  // namespace gsm

using namespace gsm;

extern "C" size_t gsm_photometric_loss_workspace_bytes(int C, int H, int W) {
  if (C <= 0 || H <= 0 || W <= 0) return 256;
  return loss_layout(C, H, W).total;
}

extern "C" int gsm_photometric_loss(const float* image, const float* gt, int C, int H, int W, float lambda_dssim,
                                    float grad_scale, float* loss_out, float* dL_dimage, void* workspace,
                                    size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (C <= 0 || H <= 0 || W <= 0) { set_error("gsm_photometric_loss: bad sizes C=%d H=%d W=%d", C, H, W); return GSM_ERR_INVALID_ARGUMENT; }
  if (image == nullptr || gt == nullptr || dL_dimage == nullptr) { set_error("gsm_photometric_loss: null image / gt / gradient"); return GSM_ERR_INVALID_ARGUMENT; }
  const LossLayout L = loss_layout(C, H, W);
  if (workspace == nullptr || workspace_bytes < L.total) { set_error("gsm_photometric_loss: workspace too small"); return GSM_ERR_WORKSPACE_TOO_SMALL; }
  static PerDeviceLevel window_ready;  // c_gauss is a per-device symbol
  window_ready.ensure(1, [] { upload_window(); });
  char* ws = (char*)workspace;
  double* sums = (double*)(ws + L.sums);
  const size_t n = (size_t)C * H * W;
  float* ma = (float*)(ws + L.maps);
  float* mb = ma + n;
  float* mc = mb + n;
  GSM_CUDA(cudaMemsetAsync(sums, 0, 16, stream));
  const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C);
  {
    StageMark m(GSM_STAGE_LOSS, stream);
    ssim_forward_kernel<<<grid, LOSS_THREADS, 0, stream>>>(image, gt, H, W, ma, mb, mc, sums);
    ssim_backward_kernel<<<grid, LOSS_THREADS, 0, stream>>>(image, gt, C, H, W, ma, mb, mc, sums, lambda_dssim, grad_scale,
                                                            loss_out, dL_dimage);
  }
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

// ---- allocation-free variant ------------------------------------------------------------------------------------
// gsm_photometric_loss_static: the same two kernels with their scratch (double[2] sums + the three partial maps) in a
// file-scope device array instead of a caller-owned workspace, for callers that cannot allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Limit: C*H*W <= GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS (4 Mi pixel-channels: 48 MiB + 256 B
// of device memory held by every GPU that loads this module, until exit).  The array is shared by every call on a
// device and keeps the previous call's contents: the sums are cleared at the start of each call (stream-ordered); the
// maps are fully overwritten by kernel 1 before kernel 2 reads them.  Calls that overlap on one GPU (two streams, two
// host threads) RACE on it -- callers serialise them.
namespace gsm {

constexpr size_t LOSS_STATIC_MAX = (size_t)GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS;
constexpr size_t LOSS_STATIC_BYTES = 256 + 3 * LOSS_STATIC_MAX * sizeof(float);
static __device__ __align__(256) unsigned char g_loss_static_scratch[LOSS_STATIC_BYTES];

// In host code `g_loss_static_scratch` is only a handle of the symbol, not a device address: the address of each
// device's copy is asked for once per device, with that device current.
constexpr int LOSS_STATIC_MAX_DEVICES = 64;
static std::mutex g_loss_static_mu;
static unsigned char* g_loss_static_ptr[LOSS_STATIC_MAX_DEVICES] = {};

static int loss_static_setup(unsigned char** out) {
  int dev = -1;
  GSM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= LOSS_STATIC_MAX_DEVICES) {
    set_error("gsm_photometric_loss_static: device ordinal %d outside [0, %d)", dev, LOSS_STATIC_MAX_DEVICES);
    return GSM_ERR_INVALID_ARGUMENT;
  }
  std::lock_guard<std::mutex> lk(g_loss_static_mu);
  if (g_loss_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    GSM_CUDA(cudaGetSymbolAddress(&p, g_loss_static_scratch));  // current device's copy of the array
    if (p == nullptr || ((uintptr_t)p & 255) != 0) { set_error("gsm_photometric_loss_static: bad scratch address %p", p); return (int)cudaErrorInvalidSymbol; }
    upload_window();                                            // c_gauss is a per-device symbol as well
    GSM_CUDA(cudaGetLastError());
    g_loss_static_ptr[dev] = (unsigned char*)p;
  }
  *out = g_loss_static_ptr[dev];
  return GSM_OK;
}

}  // namespace gsm

extern "C" size_t gsm_photometric_loss_static_max_elements(void) { return LOSS_STATIC_MAX; }

extern "C" int gsm_photometric_loss_static(const float* image, const float* gt, int C, int H, int W, float lambda_dssim,
                                           float grad_scale, float* loss_out, float* dL_dimage, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (C <= 0 || H <= 0 || W <= 0) { set_error("gsm_photometric_loss_static: bad sizes C=%d H=%d W=%d", C, H, W); return GSM_ERR_INVALID_ARGUMENT; }
  if (image == nullptr || gt == nullptr || dL_dimage == nullptr) { set_error("gsm_photometric_loss_static: null image / gt / gradient"); return GSM_ERR_INVALID_ARGUMENT; }
  const size_t n = (size_t)C * H * W;
  if (n > LOSS_STATIC_MAX) {
    set_error("gsm_photometric_loss_static: C*H*W = %zu exceeds the static scratch limit of %zu pixel-channels "
              "(GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS); use gsm_photometric_loss with a caller-owned workspace", n, LOSS_STATIC_MAX);
    return GSM_ERR_WORKSPACE_TOO_SMALL;
  }
  unsigned char* ws = nullptr;
  const int rc = loss_static_setup(&ws);
  if (rc != GSM_OK) return rc;
  double* sums = (double*)ws;
  float* ma = (float*)(ws + 256);
  float* mb = ma + n;
  float* mc = mb + n;
  GSM_CUDA(cudaMemsetAsync(sums, 0, 16, stream));   // the only part that must start at zero
  const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C);
  ssim_forward_kernel<<<grid, LOSS_THREADS, 0, stream>>>(image, gt, H, W, ma, mb, mc, sums);
  ssim_backward_kernel<<<grid, LOSS_THREADS, 0, stream>>>(image, gt, C, H, W, ma, mb, mc, sums, lambda_dssim, grad_scale,
                                                          loss_out, dL_dimage);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

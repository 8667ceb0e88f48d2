// distCUDA2 for sm_100a: mean squared distance of every point to its 3 nearest OTHER points.
//
// Replaces SimpleKNN::knn (simple_knn.cu:186-222) and its kernels coord2Morton / boxMinMax /
// boxMeanDist (simple_knn.cu:46-184).  The reference result is the exact 3-NN mean
// (b0+b1+b2)/3 with squared distances evaluated as fma(dz,dz, fma(dx,dx, dy*dy)); its Morton boxes
// only prune.  Any exact search therefore reproduces it bit for bit.  This implementation keeps
// the Morton ordering but replaces the flat list of 1024-point boxes, which every query scans in
// full (O(P^2/1024) box tests), by an implicit 8-ary bounding-box hierarchy over the Morton order
// traversed with pruning (O(P log P)).
#include "common.cuh"
#include <float.h>

namespace gsm {

size_t sort32_temp_bytes(size_t n);
cudaError_t launch_sort_pairs32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in,
                                uint32_t* vals_out, size_t n, int end_bit, void* temp,
                                size_t temp_bytes, cudaStream_t stream);

constexpr int KNN_FAN = 8;
constexpr int KNN_MAX_LEVELS = 12;

struct KnnLayout {
  size_t minmax;        // int[6] ordered-int encoded min xyz, max xyz
  size_t codes, idx, codes_sorted, idx_sorted;  // u32[P] each
  size_t pts;           // float4[P] Morton-sorted points (w = original index bits)
  size_t boxes;         // float4[2 * total_boxes]
  size_t sort_temp, sort_temp_bytes;
  size_t total;
  int levels;
  size_t level_off[KNN_MAX_LEVELS];
  int level_cnt[KNN_MAX_LEVELS];
};

static KnnLayout knn_layout(int P) {
  KnnLayout L{};
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  const size_t n = (size_t)(P > 0 ? P : 1);
  L.minmax = take(64);
  L.codes = take(4 * n); L.idx = take(4 * n); L.codes_sorted = take(4 * n); L.idx_sorted = take(4 * n);
  L.pts = take(16 * n);
  size_t total_boxes = 0;
  int cnt = (int)((n + KNN_FAN - 1) / KNN_FAN);
  L.levels = 0;
  while (true) {
    L.level_off[L.levels] = total_boxes;
    L.level_cnt[L.levels] = cnt;
    total_boxes += (size_t)cnt;
    L.levels++;
    if (cnt <= KNN_FAN || L.levels >= KNN_MAX_LEVELS) break;
    cnt = (cnt + KNN_FAN - 1) / KNN_FAN;
  }
  L.boxes = take(32 * total_boxes);
  L.sort_temp_bytes = sort32_temp_bytes(n);
  L.sort_temp = take(L.sort_temp_bytes);
  L.total = off;
  return L;
}

// order-preserving float <-> int mapping for atomicMin/Max
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void knn_init_minmax(int* mm) {
  if (threadIdx.x < 3) mm[threadIdx.x] = f2ord(FLT_MAX);
  else if (threadIdx.x < 6) mm[threadIdx.x] = f2ord(-FLT_MAX);
}

__global__ void __launch_bounds__(256) knn_minmax_kernel(int P, const float* __restrict__ pts, int* mm) {
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const float v = __ldg(pts + 3 * (size_t)i + c);
      mn[c] = fminf(mn[c], v); mx[c] = fmaxf(mx[c], v);
    }
  }
#pragma unroll
  for (int c = 0; c < 3; c++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 3; c++) { atomicMin(mm + c, f2ord(mn[c])); atomicMax(mm + 3 + c, f2ord(mx[c])); }
  }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FF;
  x = (x | (x << 8)) & 0x0300F00F;
  x = (x | (x << 4)) & 0x030C30C3;
  x = (x | (x << 2)) & 0x09249249;
  return x;
}

__global__ void __launch_bounds__(256) knn_morton_kernel(int P, const float* __restrict__ pts,
                                                         const int* __restrict__ mm,
                                                         uint32_t* __restrict__ codes, uint32_t* __restrict__ idx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint32_t code = 0;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float lo = ord2f(mm[c]), hi = ord2f(mm[3 + c]);
    const float ext = hi - lo;
    float u = ext > 0.f ? (__ldg(pts + 3 * (size_t)i + c) - lo) / ext : 0.f;
    u = fminf(fmaxf(u, 0.f), 1.f);
    code |= spread10((uint32_t)(u * 1023.0f)) << c;
  }
  codes[i] = code;
  idx[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) knn_gather_kernel(int P, const float* __restrict__ pts,
                                                         const uint32_t* __restrict__ idx_sorted,
                                                         float4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const uint32_t s = __ldg(idx_sorted + i);
  out[i] = make_float4(__ldg(pts + 3 * (size_t)s), __ldg(pts + 3 * (size_t)s + 1), __ldg(pts + 3 * (size_t)s + 2),
                       __uint_as_float(s));
}

// level 0: boxes over 8 consecutive sorted points; level l>0: boxes over 8 child boxes
__global__ void __launch_bounds__(256) knn_boxes_kernel(int level, int count, int child_count,
                                                        const float4* __restrict__ pts,
                                                        const float4* __restrict__ child, float4* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= count) return;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int c = 0; c < KNN_FAN; c++) {
    const int j = b * KNN_FAN + c;
    if (j >= child_count) break;
    if (level == 0) {
      const float4 q = __ldg(pts + j);
      mn[0] = fminf(mn[0], q.x); mn[1] = fminf(mn[1], q.y); mn[2] = fminf(mn[2], q.z);
      mx[0] = fmaxf(mx[0], q.x); mx[1] = fmaxf(mx[1], q.y); mx[2] = fmaxf(mx[2], q.z);
    } else {
      const float4 lo = __ldg(child + 2 * (size_t)j), hi = __ldg(child + 2 * (size_t)j + 1);
      mn[0] = fminf(mn[0], lo.x); mn[1] = fminf(mn[1], lo.y); mn[2] = fminf(mn[2], lo.z);
      mx[0] = fmaxf(mx[0], hi.x); mx[1] = fmaxf(mx[1], hi.y); mx[2] = fmaxf(mx[2], hi.z);
    }
  }
  out[2 * (size_t)b] = make_float4(mn[0], mn[1], mn[2], 0.f);
  out[2 * (size_t)b + 1] = make_float4(mx[0], mx[1], mx[2], 0.f);
}

// squared distance exactly as the reference evaluates it (simple_knn.cu:135-136 after contraction)
__device__ __forceinline__ float dist2_ref(float dx, float dy, float dz) {
  return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

__device__ __forceinline__ void update3(float d, float& b0, float& b1, float& b2) {
  // simple_knn.cu:137-145 (insertion into the ascending triple; strict comparisons)
  if (b0 > d) { const float t = b0; b0 = d; d = t; }
  if (b1 > d) { const float t = b1; b1 = d; d = t; }
  if (b2 > d) { b2 = d; }
}

// lower bound of dist2_ref over any point inside the box (monotone in |dx|,|dy|,|dz|)
__device__ __forceinline__ float box_dist2(float4 lo, float4 hi, float4 p) {
  const float dx = fmaxf(fmaxf(__fsub_rn(lo.x, p.x), __fsub_rn(p.x, hi.x)), 0.f);
  const float dy = fmaxf(fmaxf(__fsub_rn(lo.y, p.y), __fsub_rn(p.y, hi.y)), 0.f);
  const float dz = fmaxf(fmaxf(__fsub_rn(lo.z, p.z), __fsub_rn(p.z, hi.z)), 0.f);
  return dist2_ref(dx, dy, dz);
}

struct KnnLevels {
  int levels;
  int cnt[KNN_MAX_LEVELS];
  unsigned long long off[KNN_MAX_LEVELS];
};

__global__ void __launch_bounds__(128) knn_search_kernel(int P, const float4* __restrict__ pts,
                                                         const float4* __restrict__ boxes, KnnLevels lv,
                                                         float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float4 me = __ldg(pts + i);

  // seed bound from the +-3 Morton neighbours (simple_knn.cu:157-164)
  float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
  for (int j = max(0, i - 3); j <= min(P - 1, i + 3); j++) {
    if (j == i) continue;
    const float4 q = __ldg(pts + j);
    update3(dist2_ref(__fsub_rn(q.x, me.x), __fsub_rn(q.y, me.y), __fsub_rn(q.z, me.z)), b0, b1, b2);
  }
  float bound = b2;
  b0 = FLT_MAX; b1 = FLT_MAX; b2 = FLT_MAX;

  int stack[KNN_FAN * KNN_MAX_LEVELS];
  int sp = 0;
  const int top = lv.levels - 1;
  for (int c = lv.cnt[top] - 1; c >= 0; c--) stack[sp++] = (top << 27) | c;
  while (sp > 0) {
    const int e = stack[--sp];
    const int level = e >> 27, node = e & ((1 << 27) - 1);
    const float4* bx = boxes + 2 * (lv.off[level] + (size_t)node);
    if (box_dist2(__ldg(bx), __ldg(bx + 1), me) > bound) continue;
    if (level == 0) {
      const int beg = node * KNN_FAN, end = min(P, beg + KNN_FAN);
      for (int j = beg; j < end; j++) {
        if (j == i) continue;
        const float4 q = __ldg(pts + j);
        update3(dist2_ref(__fsub_rn(q.x, me.x), __fsub_rn(q.y, me.y), __fsub_rn(q.z, me.z)), b0, b1, b2);
      }
      bound = fminf(bound, b2);
    } else {
      const int beg = node * KNN_FAN, end = min(lv.cnt[level - 1], beg + KNN_FAN);
      // push far-to-near in Morton order relative to the query's own branch is not known cheaply;
      // push all, the bound from the Morton neighbours already prunes almost everything
      for (int c = end - 1; c >= beg; c--) stack[sp++] = ((level - 1) << 27) | c;
    }
  }
  out[__float_as_uint(me.w)] = __fdiv_rn(__fadd_rn(__fadd_rn(b0, b1), b2), 3.0f);
}

}  // namespace gsm

using namespace gsm;

extern "C" size_t gsm_knn_workspace_bytes(int P) { return knn_layout(P).total; }

extern "C" int gsm_knn_dist2(int P, const float* points, float* mean_dist2, void* workspace,
                             size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (P < 0) { set_error("gsm_knn_dist2: negative P"); return GSM_ERR_INVALID_ARGUMENT; }
  if (P == 0) return GSM_OK;
  if (P > (1 << 30)) { set_error("gsm_knn_dist2: P too large"); return GSM_ERR_INVALID_ARGUMENT; }
  const KnnLayout L = knn_layout(P);
  if (workspace == nullptr || workspace_bytes < L.total) {
    set_error("gsm_knn_dist2: workspace too small (%zu < %zu)", workspace_bytes, L.total);
    return GSM_ERR_WORKSPACE_TOO_SMALL;
  }
  char* ws = (char*)workspace;
  int* mm = (int*)(ws + L.minmax);
  uint32_t* codes = (uint32_t*)(ws + L.codes);
  uint32_t* idx = (uint32_t*)(ws + L.idx);
  uint32_t* codes_sorted = (uint32_t*)(ws + L.codes_sorted);
  uint32_t* idx_sorted = (uint32_t*)(ws + L.idx_sorted);
  float4* pts = (float4*)(ws + L.pts);
  float4* boxes = (float4*)(ws + L.boxes);

  StageMark mark(GSM_STAGE_KNN, stream);
  knn_init_minmax<<<1, 32, 0, stream>>>(mm);
  const int nb = (P + 255) / 256;
  knn_minmax_kernel<<<min(nb, 148 * 8), 256, 0, stream>>>(P, points, mm);
  knn_morton_kernel<<<nb, 256, 0, stream>>>(P, points, mm, codes, idx);
  GSM_CUDA(launch_sort_pairs32(codes, codes_sorted, idx, idx_sorted, (size_t)P, 30, ws + L.sort_temp,
                               L.sort_temp_bytes, stream));
  knn_gather_kernel<<<nb, 256, 0, stream>>>(P, points, idx_sorted, pts);
  KnnLevels lv{};
  lv.levels = L.levels;
  for (int l = 0; l < L.levels; l++) {
    lv.cnt[l] = L.level_cnt[l];
    lv.off[l] = L.level_off[l];
    const int child_count = (l == 0) ? P : L.level_cnt[l - 1];
    const float4* child = (l == 0) ? nullptr : boxes + 2 * L.level_off[l - 1];
    knn_boxes_kernel<<<(L.level_cnt[l] + 255) / 256, 256, 0, stream>>>(l, L.level_cnt[l], child_count, pts, child,
                                                                          boxes + 2 * L.level_off[l]);
  }
  knn_search_kernel<<<(P + 127) / 128, 128, 0, stream>>>(P, pts, boxes, lv, mean_dist2);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}

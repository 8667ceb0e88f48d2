// densify_and_prune as stream-compaction kernels for sm_100a (last part).
//
// Replaces GaussianModel.densify_and_prune (scene/gaussian_model.py:535-549) with everything it calls -- densify_and_clone
// (:519-533), densify_and_split (:492-517), densification_postfix (:468-490), cat_tensors_to_optimizer (:446-466),
// prune_points / _prune_optimizer (:413-444) -- run by train.py:205-208 every `densification_interval` iterations.
// The reference does it with ~60 torch ops (boolean-mask gathers with a `nonzero` host sync each, six torch.cat per
// stage, the optimizer state rebuilt three times): every parameter and both Adam moments are copied three times.
//
// WHAT comes out (derived from the reference's statements; `i` runs over the P surfels before the call):
//   g_i      = xyz_gradient_accum_i / denom_i, NaN -> 0                                  (:536-537)
//   big_i    = max(exp(scaling_i)) > percent_dense * extent
//   clone_i  = |g_i| >= max_grad  and not big_i                                          (:521-523)
//   split_i  = g_i >= max_grad  and  big_i     (clones are appended with grad 0: never split, :495-499)
//   pruneA_i = sigmoid(opacity_i) < min_opacity  or  (max_screen_size and max(exp(scaling_i)) > 0.1 extent)
//              -- max_radii2D plays no part: densification_postfix has reset it to zero by then (:489)
//   pruneS_i = the same test on a split child: same opacity, scaling log(exp(scaling_i) / (0.8 N))
//   surfels afterwards, in this order:
//     A  = { i : not split_i and not pruneA_i }           parameters and Adam moments kept
//     B  = { i : clone_i and not pruneA_i }               parameters copied, moments zero
//     S1 = S2 = { i : split_i and not pruneS_i }          N = 2 children each: xyz = R(q_i) (z * (exp(scaling_i), 0)) + xyz_i
//                                                         with z ~ N(0,1) row (k * #split + rank_i) of the caller's normal
//                                                         draw, scaling as above, everything else copied, moments zero
//   xyz_gradient_accum, denom, max_radii2D: zeros (:487-489).
// HOW: flags + block sums -> one-CTA scan of the block sums -> (host reads the four totals to size the new tensors) ->
// source map of the new surfels -> ONE gather launch over all tensors (6 parameters, 12 moments, segments), 128-bit stores.
#include "common.cuh"
#include <cmath>

namespace gsm {

constexpr int DFY_THREADS = 256;
constexpr int DFY_ITEMS = 4;                       // surfels per thread in the flag / map kernels
constexpr int DFY_BLOCK = DFY_THREADS * DFY_ITEMS;
constexpr uint32_t KIND_SHIFT = 30, SRC_MASK = (1u << KIND_SHIFT) - 1u;
enum { KIND_A = 0, KIND_B = 1, KIND_S1 = 2, KIND_S2 = 3 };
enum { FLAG_A = 1, FLAG_B = 2, FLAG_S = 4, FLAG_SPLIT = 8 };

struct DensifyPlanLayout {
  size_t flags;        // u8[P]
  size_t block_sums;   // u64[nblocks][2]: (A << 32 | B), (S << 32 | split)   exclusive after the scan
  size_t totals;       // u64[4]: nA, nB, nS, n_split
  size_t total;
};
static DensifyPlanLayout densify_plan_layout(size_t P) {
  DensifyPlanLayout L{};
  const size_t nb = (P + DFY_BLOCK - 1) / DFY_BLOCK;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 128); return o; };
  L.flags = take(P);
  L.block_sums = take(nb * 16);
  L.totals = take(32);
  L.total = off + 128;
  return L;
}

struct DensifyTests {
  float grad_threshold, dense_limit, min_opacity, world_limit, split_inv;
  int has_size_limit;
};

__device__ __forceinline__ uint32_t densify_flags_of(float accum, float denom, float2 sc_raw, float op_raw, const DensifyTests t) {
  float g = __fdiv_rn(accum, denom);
  if (g != g) g = 0.0f;                                                   // grads[grads.isnan()] = 0.0
  const float2 sc = activate_scaling(sc_raw);                             // get_scaling = exp
  const float smax = fmaxf(sc.x, sc.y);
  const bool big = smax > t.dense_limit;
  const bool clone = fabsf(g) >= t.grad_threshold && !big;                // torch.norm(grads, dim=-1) over one column
  const bool split = g >= t.grad_threshold && big;
  const bool low_opacity = activate_opacity(op_raw) < t.min_opacity;      // get_opacity = sigmoid
  const bool pruneA = low_opacity || (t.has_size_limit && smax > t.world_limit);
  // a split child's scaling: log(get_scaling / (0.8 N)) (torch divides by a Python scalar as a multiplication by
  // its fp32 reciprocal), read back through exp by the prune test
  const float cx = expf(logf(__fmul_rn(sc.x, t.split_inv))), cy = expf(logf(__fmul_rn(sc.y, t.split_inv)));
  const bool pruneS = low_opacity || (t.has_size_limit && fmaxf(cx, cy) > t.world_limit);
  uint32_t f = 0;
  if (!split && !pruneA) f |= FLAG_A;
  if (clone && !pruneA) f |= FLAG_B;
  if (split && !pruneS) f |= FLAG_S;
  if (split) f |= FLAG_SPLIT;
  return f;
}

__device__ __forceinline__ void pack_counts(uint32_t f, unsigned long long& ab, unsigned long long& st) {
  ab = ((unsigned long long)((f & FLAG_A) ? 1u : 0u) << 32) | ((f & FLAG_B) ? 1u : 0u);
  st = ((unsigned long long)((f & FLAG_S) ? 1u : 0u) << 32) | ((f & FLAG_SPLIT) ? 1u : 0u);
}

// block-wide inclusive scan of two packed counters; returns the block totals in (tot_ab, tot_st)
__device__ __forceinline__ void block_scan2(unsigned long long& ab, unsigned long long& st, unsigned long long& tot_ab,
                                            unsigned long long& tot_st, unsigned long long (*wsum)[2]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long a = __shfl_up_sync(0xffffffffu, ab, o), s = __shfl_up_sync(0xffffffffu, st, o);
    if (lane >= o) { ab += a; st += s; }
  }
  if (lane == 31) { wsum[warp][0] = ab; wsum[warp][1] = st; }
  __syncthreads();
  unsigned long long oa = 0, os = 0;
  tot_ab = 0; tot_st = 0;
  for (int w = 0; w < DFY_THREADS / 32; w++) {
    if (w < warp) { oa += wsum[w][0]; os += wsum[w][1]; }
    tot_ab += wsum[w][0]; tot_st += wsum[w][1];
  }
  ab += oa; st += os;
  __syncthreads();
}

__global__ void __launch_bounds__(DFY_THREADS)
densify_flags_kernel(size_t P, const float* __restrict__ accum, const float* __restrict__ denom,
                     const float2* __restrict__ scaling, const float* __restrict__ opacity, const DensifyTests t,
                     uint8_t* __restrict__ flags, unsigned long long* __restrict__ block_sums) {
  __shared__ unsigned long long wsum[DFY_THREADS / 32][2];
  const size_t base = (size_t)blockIdx.x * DFY_BLOCK + (size_t)threadIdx.x * DFY_ITEMS;
  unsigned long long ab = 0, st = 0;
  uint32_t packed = 0;
#pragma unroll
  for (int k = 0; k < DFY_ITEMS; k++) {
    const size_t i = base + k;
    uint32_t f = 0;
    if (i < P) f = densify_flags_of(__ldg(accum + i), __ldg(denom + i), __ldg(scaling + i), __ldg(opacity + i), t);
    packed |= f << (8 * k);
    unsigned long long a, s;
    pack_counts(f, a, s);
    ab += a; st += s;
  }
  if (base + DFY_ITEMS <= P && (reinterpret_cast<size_t>(flags) & 3) == 0) *reinterpret_cast<uint32_t*>(flags + base) = packed;
  else for (int k = 0; k < DFY_ITEMS; k++) if (base + k < P) flags[base + k] = (uint8_t)(packed >> (8 * k));
  unsigned long long tab, tst;
  block_scan2(ab, st, tab, tst, wsum);
  if (threadIdx.x == 0) { block_sums[2 * (size_t)blockIdx.x] = tab; block_sums[2 * (size_t)blockIdx.x + 1] = tst; }
}

// one CTA: exclusive scan of the block sums in place, totals out
__global__ void __launch_bounds__(DFY_THREADS)
densify_scan_kernel(size_t nblocks, unsigned long long* __restrict__ block_sums, unsigned long long* __restrict__ totals) {
  __shared__ unsigned long long wsum[DFY_THREADS / 32][2];
  unsigned long long carry_ab = 0, carry_st = 0;
  for (size_t b0 = 0; b0 < nblocks; b0 += DFY_THREADS) {
    const size_t b = b0 + threadIdx.x;
    unsigned long long ab = 0, st = 0;
    if (b < nblocks) { ab = block_sums[2 * b]; st = block_sums[2 * b + 1]; }
    const unsigned long long mine_ab = ab, mine_st = st;
    unsigned long long tab, tst;
    block_scan2(ab, st, tab, tst, wsum);
    if (b < nblocks) { block_sums[2 * b] = carry_ab + ab - mine_ab; block_sums[2 * b + 1] = carry_st + st - mine_st; }
    carry_ab += tab; carry_st += tst;
  }
  if (threadIdx.x == 0) {
    totals[0] = carry_ab >> 32; totals[1] = carry_ab & 0xffffffffull;
    totals[2] = carry_st >> 32; totals[3] = carry_st & 0xffffffffull;
  }
}

// source map of the new surfels: src_map[dst] = source surfel | kind << 30, zrow[dst] = row of the normal draw (children)
__global__ void __launch_bounds__(DFY_THREADS)
densify_map_kernel(size_t P, const uint8_t* __restrict__ flags, const unsigned long long* __restrict__ block_sums,
                   const unsigned long long* __restrict__ totals, uint32_t* __restrict__ src_map, uint32_t* __restrict__ zrow) {
  __shared__ unsigned long long wsum[DFY_THREADS / 32][2];
  const size_t base = (size_t)blockIdx.x * DFY_BLOCK + (size_t)threadIdx.x * DFY_ITEMS;
  uint32_t f[DFY_ITEMS];
  unsigned long long ab = 0, st = 0;
#pragma unroll
  for (int k = 0; k < DFY_ITEMS; k++) {
    f[k] = (base + k < P) ? flags[base + k] : 0u;
    unsigned long long a, s;
    pack_counts(f[k], a, s);
    ab += a; st += s;
  }
  const unsigned long long mine_ab = ab, mine_st = st;
  unsigned long long tab, tst;
  block_scan2(ab, st, tab, tst, wsum);
  unsigned long long run_ab = block_sums[2 * (size_t)blockIdx.x] + ab - mine_ab;
  unsigned long long run_st = block_sums[2 * (size_t)blockIdx.x + 1] + st - mine_st;
  const uint32_t nA = (uint32_t)totals[0], nB = (uint32_t)totals[1], nS = (uint32_t)totals[2], nT = (uint32_t)totals[3];
#pragma unroll
  for (int k = 0; k < DFY_ITEMS; k++) {
    const uint32_t i = (uint32_t)(base + k);
    const uint32_t rA = (uint32_t)(run_ab >> 32), rB = (uint32_t)run_ab, rS = (uint32_t)(run_st >> 32), rT = (uint32_t)run_st;
    if (f[k] & FLAG_A) src_map[rA] = i | ((uint32_t)KIND_A << KIND_SHIFT);
    if (f[k] & FLAG_B) src_map[nA + rB] = i | ((uint32_t)KIND_B << KIND_SHIFT);
    if (f[k] & FLAG_S) {
      src_map[nA + nB + rS] = i | ((uint32_t)KIND_S1 << KIND_SHIFT);
      src_map[nA + nB + nS + rS] = i | ((uint32_t)KIND_S2 << KIND_SHIFT);
      zrow[nA + nB + rS] = rT;               // stds = get_scaling[mask].repeat(N, 1): copy k of the r-th selected is row k * n + r
      zrow[nA + nB + nS + rS] = nT + rT;
    }
    unsigned long long a, s;
    pack_counts(f[k], a, s);
    run_ab += a; run_st += s;
  }
}

// ---- the gather: every tensor of the model in one launch -------------------------------------------------------
constexpr int DFY_MAX_TENSORS = 20;
enum { ROLE_COPY = 0, ROLE_MOMENT = 1, ROLE_XYZ = 2, ROLE_SCALING = 3 };
struct MoveTensor {
  const uint32_t* in;   // rows of `words` 4-byte words (fp32 everywhere; int64 segments as 2 words); NULL: zeros
  uint32_t* out;
  int words;
  int role;
};
struct MoveBatch {
  MoveTensor t[DFY_MAX_TENSORS];
  unsigned first_chunk[DFY_MAX_TENSORS + 1];
  int count;
  uint32_t P_new;
  const float* xyz; const float2* scaling; const float4* rotation;   // sources of the split children
  const float* normal_draw;                                          // [N * n_split, 3]
  float split_inv;
};
constexpr int MOVE_THREADS = 256;
constexpr size_t MOVE_CHUNK = (size_t)MOVE_THREADS * 4;  // output words per CTA

// build_rotation (utils/general_utils.py:78-99) as torch evaluates it: every elementwise op rounded on its own
__device__ __forceinline__ void split_child_value(const MoveBatch& b, uint32_t src, uint32_t zr, int c, float& xyz_out) {
  const float4 r = __ldg(b.rotation + src);
  const float n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r.x, r.x), __fmul_rn(r.y, r.y)), __fmul_rn(r.z, r.z)), __fmul_rn(r.w, r.w)));
  const float qr = __fdiv_rn(r.x, n), qx = __fdiv_rn(r.y, n), qy = __fdiv_rn(r.z, n), qz = __fdiv_rn(r.w, n);
  float R0, R1, R2;  // row c of the rotation matrix
  if (c == 0) {
    R0 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(qy, qy), __fmul_rn(qz, qz))));
    R1 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(qx, qy), __fmul_rn(qr, qz)));
    R2 = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(qx, qz), __fmul_rn(qr, qy)));
  } else if (c == 1) {
    R0 = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(qx, qy), __fmul_rn(qr, qz)));
    R1 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(qx, qx), __fmul_rn(qz, qz))));
    R2 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(qy, qz), __fmul_rn(qr, qx)));
  } else {
    R0 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(qx, qz), __fmul_rn(qr, qy)));
    R1 = __fmul_rn(2.0f, __fadd_rn(__fmul_rn(qy, qz), __fmul_rn(qr, qx)));
    R2 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(qx, qx), __fmul_rn(qy, qy))));
  }
  const float2 std = activate_scaling(__ldg(b.scaling + src));            // stds = (get_scaling, 0)
  const float* z = b.normal_draw + 3 * (size_t)zr;
  const float s0 = __fmul_rn(__ldg(z + 0), std.x), s1 = __fmul_rn(__ldg(z + 1), std.y), s2 = __fmul_rn(__ldg(z + 2), 0.0f);
  // torch.bmm(rots, samples): a 3-term dot product (cuBLAS; its summation order is not specified -- the parity test
  // holds the children's positions to 1e-6 relative, everything else in this file to the bit)
  const float dot = __fmaf_rn(R2, s2, __fmaf_rn(R1, s1, __fmul_rn(R0, s0)));
  xyz_out = __fadd_rn(dot, __ldg(b.xyz + 3 * (size_t)src + c));
}

__global__ void __launch_bounds__(MOVE_THREADS)
densify_move_kernel(const __grid_constant__ MoveBatch b, const uint32_t* __restrict__ src_map, const uint32_t* __restrict__ zrow) {
  int ti = 0;
#pragma unroll
  for (int k = 1; k < DFY_MAX_TENSORS; k++) ti += (k < b.count && blockIdx.x >= b.first_chunk[k]) ? 1 : 0;
  const MoveTensor& t = b.t[ti];
  const size_t n = (size_t)b.P_new * t.words;
  const size_t e0 = ((size_t)(blockIdx.x - b.first_chunk[ti]) * MOVE_THREADS + threadIdx.x) * 4;
  if (e0 >= n) return;
  uint32_t v[4];
#pragma unroll
  for (int u = 0; u < 4; u++) {
    const size_t e = e0 + u;
    v[u] = 0u;
    if (e >= n) continue;
    const uint32_t row = (uint32_t)(e / (unsigned)t.words), c = (uint32_t)(e - (size_t)row * t.words);
    const uint32_t sm = __ldg(src_map + row), kind = sm >> KIND_SHIFT, src = sm & SRC_MASK;
    if (t.role == ROLE_MOMENT) {
      if (kind == KIND_A && t.in != nullptr) v[u] = __ldg(t.in + (size_t)src * t.words + c);   // clones and children start from zero moments
    } else if (t.role == ROLE_XYZ && kind >= KIND_S1) {
      float x;
      split_child_value(b, src, __ldg(zrow + row), (int)c, x);
      v[u] = __float_as_uint(x);
    } else if (t.role == ROLE_SCALING && kind >= KIND_S1) {
      v[u] = __float_as_uint(logf(__fmul_rn(expf(__uint_as_float(__ldg(t.in + (size_t)src * t.words + c))), b.split_inv)));
    } else {
      v[u] = __ldg(t.in + (size_t)src * t.words + c);
    }
  }
  if (e0 + 4 <= n) *reinterpret_cast<uint4*>(t.out + e0) = make_uint4(v[0], v[1], v[2], v[3]);
  else for (int u = 0; u < 4 && e0 + u < n; u++) t.out[e0 + u] = v[u];
}

}  // namespace gsm

using namespace gsm;

extern "C" size_t gsm_densify_workspace_bytes(size_t P) { return densify_plan_layout(P).total; }

extern "C" int gsm_densify_plan(const gsm_densify_args* a, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a == nullptr || workspace == nullptr) { set_error("gsm_densify_plan: null argument"); return GSM_ERR_INVALID_ARGUMENT; }
  if (a->P == 0 || a->P >= (size_t)SRC_MASK) { set_error("gsm_densify_plan: P = %zu out of range", a->P); return GSM_ERR_INVALID_ARGUMENT; }
  if (a->split_count != 2) { set_error("gsm_densify_plan: the reference splits into N = 2 (gaussian_model.py:492)"); return GSM_ERR_INVALID_ARGUMENT; }
  const DensifyPlanLayout L = densify_plan_layout(a->P);
  if (workspace_bytes < L.total) { set_error("gsm_densify_plan: workspace too small"); return GSM_ERR_WORKSPACE_TOO_SMALL; }
  char* ws = (char*)workspace;
  DensifyTests t{a->grad_threshold, a->dense_limit, a->min_opacity, a->world_limit, a->split_inv, a->has_size_limit};
  const size_t nb = (a->P + DFY_BLOCK - 1) / DFY_BLOCK;
  densify_flags_kernel<<<(unsigned)nb, DFY_THREADS, 0, stream>>>(a->P, a->xyz_gradient_accum, a->denom, (const float2*)a->scaling,
                                                                a->opacity, t, (uint8_t*)(ws + L.flags),
                                                                (unsigned long long*)(ws + L.block_sums));
  densify_scan_kernel<<<1, DFY_THREADS, 0, stream>>>(nb, (unsigned long long*)(ws + L.block_sums), (unsigned long long*)(ws + L.totals));
  GSM_CUDA(cudaGetLastError());
  if (a->totals_host != nullptr)   // pinned or pageable HOST u64[4]; valid once the stream has reached this point
    GSM_CUDA(cudaMemcpyAsync(a->totals_host, ws + L.totals, 32, cudaMemcpyDeviceToHost, stream));
  return GSM_OK;
}

extern "C" int gsm_densify_apply(const gsm_densify_args* a, const void* workspace, const gsm_densify_tensor* tensors, int n_tensors,
                                 size_t P_new, uint32_t* src_map, uint32_t* zrow, const float* normal_draw, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a == nullptr || workspace == nullptr || tensors == nullptr || n_tensors <= 0 || n_tensors > DFY_MAX_TENSORS) {
    set_error("gsm_densify_apply: bad arguments (at most %d tensors)", DFY_MAX_TENSORS); return GSM_ERR_INVALID_ARGUMENT;
  }
  if (P_new == 0) return GSM_OK;
  if (src_map == nullptr || zrow == nullptr) { set_error("gsm_densify_apply: null map"); return GSM_ERR_INVALID_ARGUMENT; }
  const DensifyPlanLayout L = densify_plan_layout(a->P);
  const char* ws = (const char*)workspace;
  const size_t nb = (a->P + DFY_BLOCK - 1) / DFY_BLOCK;
  densify_map_kernel<<<(unsigned)nb, DFY_THREADS, 0, stream>>>(a->P, (const uint8_t*)(ws + L.flags),
                                                              (const unsigned long long*)(ws + L.block_sums),
                                                              (const unsigned long long*)(ws + L.totals), src_map, zrow);
  MoveBatch b{};
  b.count = n_tensors; b.P_new = (uint32_t)P_new;
  b.xyz = a->xyz; b.scaling = (const float2*)a->scaling; b.rotation = (const float4*)a->rotation;
  b.normal_draw = normal_draw; b.split_inv = a->split_inv;
  unsigned chunks = 0;
  for (int i = 0; i < n_tensors; i++) {
    if (tensors[i].out == nullptr || tensors[i].words <= 0 || (reinterpret_cast<size_t>(tensors[i].out) & 15) != 0) {
      set_error("gsm_densify_apply: tensor %d: output must be 16-byte aligned, words > 0", i); return GSM_ERR_INVALID_ARGUMENT;
    }
    if (tensors[i].in == nullptr && tensors[i].role != ROLE_MOMENT) { set_error("gsm_densify_apply: tensor %d has no input", i); return GSM_ERR_INVALID_ARGUMENT; }
    b.t[i].in = (const uint32_t*)tensors[i].in; b.t[i].out = (uint32_t*)tensors[i].out;
    b.t[i].words = tensors[i].words; b.t[i].role = tensors[i].role;
    b.first_chunk[i] = chunks;
    chunks += (unsigned)((P_new * (size_t)tensors[i].words + MOVE_CHUNK - 1) / MOVE_CHUNK);
  }
  b.first_chunk[n_tensors] = chunks;
  if (chunks) densify_move_kernel<<<chunks, MOVE_THREADS, 0, stream>>>(b, src_map, zrow);
  GSM_CUDA(cudaGetLastError());
  return GSM_OK;
}
